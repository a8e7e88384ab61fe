"""Prints the handful of ncu metrics we track from a .ncu-rep (raw page) and the SASS hot spots (source page).
usage: python tools/ncu_summary.py report.ncu-rep [--sass]"""
import csv, io, subprocess, sys

WANT = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__inst_issued.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__average_warp_latency_per_inst_issued.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio", "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "l1tex__m_l1tex2xbar_write_bytes.sum", "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum"]


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    rows = page(rep, "raw")
    hdr, units, vals = rows[0], rows[1], rows[2]
    ix = {h: i for i, h in enumerate(hdr)}
    print("kernel:", vals[ix["Kernel Name"]][:100])
    for w in WANT:
        if w in ix:
            print("  %-86s %s %s" % (w, vals[ix[w]], units[ix[w]]))
    if "--sass" in sys.argv:
        rows = page(rep, "source")
        h = rows[1]
        ix = {x: i for i, x in enumerate(h)}
        data = rows[2:]
        tot = sum(int(r[ix["Instructions Executed"]]) for r in data)
        tots = sum(int(r[ix["# Samples"]]) for r in data)
        thr = float(sys.argv[sys.argv.index("--sass") + 1]) if len(sys.argv) > sys.argv.index("--sass") + 1 else 0.004
        print("total warp instructions %d, samples %d" % (tot, tots))
        for n, r in enumerate(data):
            ie = int(r[ix["Instructions Executed"]]); s = int(r[ix["# Samples"]])
            if ie > tot * thr or s > tots * thr * 1.5:
                print("%5d %6.2f%% thr %5s smp %5.2f%% | %s" % (n, 100 * ie / tot, r[ix["Avg. Threads Executed"]][:5], 100 * s / max(1, tots), r[ix["Source"]][:100]))


if __name__ == "__main__":
    main()
