"""Summarise an .ncu-rep (raw page) into the handful of metrics the roofline discussion needs.
usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep [extra metric substrings...]"""
import csv, subprocess, sys
rep = sys.argv[1]
extra = sys.argv[2:]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct", "lts__t_bytes.sum", "lts__throughput.avg.pct",
        "smsp__inst_executed.sum", "sm__cycles_elapsed.max", "launch__shared_mem_per_block_dynamic", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "smsp__average_warp", "smsp__warps_issue_stalled", "sm__inst_executed_pipe_fp64", "sm__pipe_fp64_cycles_active", "l1tex__throughput.avg.pct"] + extra
idx = [i for i, h in enumerate(hdr) if any(w in h for w in want) and "per_second" not in h and ".peak_sustained" not in h.replace("pct_of_peak_sustained", "")]
for r in rows[2:]:
    print("=" * 100)
    for i in idx:
        if r[i] not in ("", "0", "n/a"):
            print(f"{hdr[i]:95s} {units[i]:12s} {r[i]}")
