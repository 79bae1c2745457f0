import csv, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
want = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "sm__warps_active.avg.per_cycle_active",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "smsp__warps_eligible.avg.per_cycle_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "smsp__average_warp_latency_per_inst_issued.ratio"]
for r in rows[2:]:
  print(r[hdr.index("Kernel Name")][:80])
  for i, h in enumerate(hdr):
    if h in want: print("  %-70s %s %s" % (h, r[i], units[i]))
    if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio") and float(r[i] or 0) > 0.1:
      print("  stall %-50s %s" % (h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")], r[i]))
