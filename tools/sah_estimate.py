"""Expected node visits (sum of internal-node surface areas / root area) of the LBVH the product builds, with per-axis and with
cubic Morton scaling, against a full-sweep SAH build, for the 10k-sphere workload.  CPU only; development aid.
Result: per-axis 65.2, cubic 31.2, SAH 28.7 -- the cubic LBVH is within 8 % of a SAH tree on this scene class."""
import importlib, sys, numpy as np
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
scenes = importlib.import_module("rust-raytrace_b200.scenes")
d = scenes.spheres_10k()
data = d["objects"][1]["data"]
c = data[:, :3]; r = data[:, 3]
lo = c - r[:, None]; hi = c + r[:, None]
def sa(l, h):
    e = h - l; return 2 * (e[0]*e[1] + e[1]*e[2] + e[0]*e[2])
root_sa = sa(lo.min(0), hi.max(0))
# LBVH with cubic morton
clo, chi = c.min(0), c.max(0); ext = (chi - clo).max()
q = np.clip(((c - clo) / ext * 2097152).astype(np.int64), 0, 2097151)
def spread(v):
    v = v & 0x1FFFFF
    v = (v | (v << 32)) & 0x001F00000000FFFF
    v = (v | (v << 16)) & 0x001F0000FF0000FF
    v = (v | (v << 8)) & 0x100F00F00F00F00F
    v = (v | (v << 4)) & 0x10C30C30C30C30C3
    v = (v | (v << 2)) & 0x1249249249249249
    return v
keys = (spread(q[:,0]) << 2) | (spread(q[:,1]) << 1) | spread(q[:,2])
order = np.argsort(keys, kind='stable'); keys = keys[order]; L = lo[order]; H = hi[order]
sys.setrecursionlimit(100000)
def lbvh(a, b, bit):  # range [a,b), returns (sum internal SA, box)
    if b - a == 1: return 0.0, L[a], H[a]
    while bit >= 0:
        m = (keys[a:b] >> bit) & 1
        if m[0] != m[-1]: break
        bit -= 1
    if bit < 0: split = (a + b) // 2
    else: split = a + int(np.searchsorted(m, 1))
    s1, l1, h1 = lbvh(a, split, bit - 1); s2, l2, h2 = lbvh(split, b, bit - 1)
    l = np.minimum(l1, l2); h = np.maximum(h1, h2)
    return s1 + s2 + sa(l, h), l, h
s_lbvh, _, _ = lbvh(0, len(keys), 62)
# per-axis morton (old)
inv = 1.0 / (chi - clo)
q2 = np.clip(((c - clo) * inv * 2097152).astype(np.int64), 0, 2097151)
keys2 = (spread(q2[:,0]) << 2) | (spread(q2[:,1]) << 1) | spread(q2[:,2])
o2 = np.argsort(keys2, kind='stable')
keys_b, L_b, H_b = keys, L, H
keys, L, H = keys2[o2], lo[o2], hi[o2]
s_axis, _, _ = lbvh(0, len(keys), 62)
# SAH binned top-down
def sah(idx):
    if len(idx) == 1: return 0.0, lo[idx[0]], hi[idx[0]]
    l = lo[idx].min(0); h = hi[idx].max(0)
    best = None
    cc = c[idx]
    for ax in range(3):
        o = np.argsort(cc[:, ax]); ii = idx[o]
        n = len(ii)
        # sweep SAH (exact, full sweep)
        lmin = np.minimum.accumulate(lo[ii], 0); lmax = np.maximum.accumulate(hi[ii], 0)
        rmin = np.minimum.accumulate(lo[ii][::-1], 0)[::-1]; rmax = np.maximum.accumulate(hi[ii][::-1], 0)[::-1]
        el = lmax[:-1] - lmin[:-1]; er = rmax[1:] - rmin[1:]
        sl = 2*(el[:,0]*el[:,1]+el[:,1]*el[:,2]+el[:,0]*el[:,2]); sr = 2*(er[:,0]*er[:,1]+er[:,1]*er[:,2]+er[:,0]*er[:,2])
        k = np.arange(1, n)
        cost = sl * k + sr * (n - k)
        j = int(np.argmin(cost))
        if best is None or cost[j] < best[0]: best = (cost[j], ii[:j+1], ii[j+1:])
    s1, _, _ = sah(best[1]); s2, _, _ = sah(best[2])
    return s1 + s2 + sa(l, h), l, h
s_sah, _, _ = sah(np.arange(len(c)))
print("sum internal SA / root SA: per-axis morton %.1f, cubic morton %.1f, full-sweep SAH %.1f" % (s_axis/root_sa, s_lbvh/root_sa, s_sah/root_sa))
