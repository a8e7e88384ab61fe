"""profiles/<prefix>_ncu_<kernel>.txt extracts and profiles/ncu_traffic.json from the .ncu-rep files of one tools/gpu_check.sh visit.
usage: python tools/ncu_traffic.py GPURUN_TAG PROFILE_PREFIX   (e.g. s16h r01_v18)"""
import csv, io, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag, prefix = sys.argv[1], sys.argv[2]
skip = {"k_shadow_bvh": 2, "k_closest_bvh": 12, "k_shade": 12}
out = {"workload": "spheres-10k-4k-d8", "kernels": {}}
for k in ("k_shadow_bvh", "k_closest_bvh", "k_shade"):
    rep = os.path.join(ROOT, "gpurun_out", "%s_full_%s.ncu-rep" % (tag, k))
    txt = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), rep], capture_output=True, text=True).stdout
    name = "%s_ncu_%s.txt" % (prefix, k)
    with open(os.path.join(ROOT, "profiles", name), "w") as f:
        f.write(txt)
    rows = list(csv.reader(io.StringIO(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    ix = {h: i for i, h in enumerate(hdr)}

    def val(m, scale_units=True):
        v = float(vals[ix[m]].replace(",", ""))
        u = units[ix[m]]
        if scale_units:
            v *= {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "us": 1e-3, "ms": 1.0, "ns": 1e-6, "usecond": 1e-3, "msecond": 1.0, "nsecond": 1e-6}.get(u, 1.0)
        return v
    out["kernels"][k] = {
        "dram_bytes_read": val("dram__bytes_read.sum"), "dram_bytes_write": val("dram__bytes_write.sum"),
        "launch_ms_under_ncu": val("gpu__time_duration.sum"),
        "source": "profiles/%s (ncu --set full, one launch of the spheres-10k-4k-d8 frame: -k regex:%s -s %d -c 1)" % (name, k, skip[k]),
        "issue_slots_pct": val("sm__inst_issued.avg.pct_of_peak_sustained_active", False),
        "lanes_per_instruction": val("smsp__thread_inst_executed_per_inst_executed.ratio", False),
        "l1_lsu_wavefronts_pct": val("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", False),
        "warps_active_pct": val("sm__warps_active.avg.pct_of_peak_sustained_active", False),
    }
with open(os.path.join(ROOT, "profiles", "ncu_traffic.json"), "w") as f:
    json.dump(out, f, indent=1)
print(json.dumps(out, indent=1))
