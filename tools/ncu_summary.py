"""Print the key metrics of every kernel in an .ncu-rep (run where ncu is installed, no GPU needed)."""
import csv, subprocess, sys
rep = sys.argv[1]
pat = sys.argv[2:] or None
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.sum",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__cycles_active.avg", "smsp__cycles_active.avg"]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")]
    print("==", name[:100])
    for k in KEYS:
        if k in hdr:
            print("   %-70s %s %s" % (k, r[hdr.index(k)], units[hdr.index(k)]))
    st = [(float(r[i] or 0), hdr[i]) for i in range(len(hdr)) if "smsp__average_warp" in hdr[i] and "issue_stalled" in hdr[i] and hdr[i].endswith("_per_issue_active.ratio") or
          ("smsp__average_warps_issue_stalled" in hdr[i] and hdr[i].endswith("per_issue_active.ratio"))]
    for v, k in sorted(st, reverse=True)[:8]:
        print("   stall %-64s %.2f" % (k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v))
