"""Group the SASS of an ncu source-page CSV into runs of equal execution count (development aid).
usage: ncu -i X.ncu-rep --page source --csv > src.csv ; python tools/sass_segments.py src.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
seq = []
for r in rows[2:]:
    if len(r) < len(hdr): continue
    seq.append((r[ix['Source']].strip(), int(r[ix['Instructions Executed']]), float(r[ix['Avg. Threads Executed']]), int(r[ix['# Samples']])))
tot = sum(x[1] for x in seq); allsm = sum(x[3] for x in seq)
seg = []; cur = None
for i, (s, ie, at, sm) in enumerate(seq):
    if cur and abs(ie - cur['ie']) <= 0.02 * max(ie, cur['ie']):
        cur['n'] += 1; cur['tot'] += ie; cur['sm'] += sm; cur['end'] = i
    else:
        if cur: seg.append(cur)
        cur = dict(start=i, end=i, ie=ie, n=1, tot=ie, sm=sm, at=at)
seg.append(cur)
print("total warp instructions", tot, "samples", allsm)
for g in sorted(seg, key=lambda g: -g['tot'])[:top]:
    print("%5d-%5d n=%4d exec/inst=%10d tot=%5.1f%% samples=%5.1f%% thr=%4.1f  %s" % (g['start'], g['end'], g['n'], g['ie'], 100 * g['tot'] / tot, 100 * g['sm'] / allsm, g['at'], seq[g['start']][0][:60]))
