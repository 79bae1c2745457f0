#!/usr/bin/env python
"""Turns the raw ncu artefacts of one gpurun call (gpurun_out/, scratch) into the small,
tracked summaries under profiles/.

  python profiles/summarize_ncu.py <tag> [--launches gpurun_out/<x>.csv] [--rep gpurun_out/<y>.ncu-rep]

writes profiles/<tag>_launches.txt (per-kernel launch count / total / mean device time from the
`--metrics gpu__time_duration.sum` pass) and profiles/<tag>_kernels.csv (one row per profiled
launch of the `--set full` capture with the metrics DESIGN.md and bench.py quote).
Runs here (no GPU needed): `ncu -i` only reads the report.
"""
import argparse
import collections
import csv
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))

METRICS = [
    "gpu__time_duration.sum",
    "launch__grid_size",
    "launch__block_size",
    "launch__registers_per_thread",
    "launch__occupancy_limit_registers",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum",
    "smsp__thread_inst_executed.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__warps_eligible.avg.per_cycle_active",
    "smsp__average_branch_targets_threads_uniform.pct",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct",
    "l1tex__t_sector_hit_rate.pct",
]


def to_ms(value, unit):
  v = float(value.replace(",", ""))
  return {"ns": v / 1e6, "us": v / 1e3, "ms": v, "s": v * 1e3}.get(unit, v)


def launches(path, out):
  rows = [r for r in csv.reader(open(path)) if len(r) > 5]
  hdr = rows[0]
  ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
  agg = collections.OrderedDict()
  for r in rows[1:]:
    agg.setdefault(r[ki], []).append(to_ms(r[vi], r[ui]))
  total = sum(sum(v) for v in agg.values())
  with open(out, "w") as f:
    f.write("# source: %s (ncu --metrics gpu__time_duration.sum --clock-control none)\n" % path)
    f.write("# per-launch times are cold-cache and serialised: compare shares, not absolutes\n")
    f.write("%6s %10s %10s %7s  %s\n" % ("count", "total_ms", "mean_ms", "share", "kernel"))
    for k, v in agg.items():
      f.write("%6d %10.3f %10.4f %6.1f%%  %s\n" % (len(v), sum(v), sum(v) / len(v),
                                                  100 * sum(v) / total, k[:110]))
  print("wrote", out)


def kernels(rep, out):
  raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True,
                       text=True, check=True).stdout
  rows = list(csv.reader(raw.splitlines()))
  hdr, units = rows[0], rows[1]
  cols = [(m, hdr.index(m)) for m in METRICS if m in hdr]
  ki = hdr.index("Kernel Name")
  with open(out, "w", newline="") as f:
    w = csv.writer(f)
    w.writerow(["kernel"] + ["%s [%s]" % (m, units[i]) for m, i in cols])
    for r in rows[2:]:
      w.writerow([r[ki]] + [r[i] for _, i in cols])
  print("wrote", out)


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("tag")
  ap.add_argument("--launches")
  ap.add_argument("--rep")
  a = ap.parse_args()
  if a.launches:
    launches(a.launches, os.path.join(HERE, a.tag + "_launches.txt"))
  if a.rep:
    kernels(a.rep, os.path.join(HERE, a.tag + "_kernels.csv"))
  return 0


if __name__ == "__main__":
  sys.exit(main())
