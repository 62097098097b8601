#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page raw --csv` (one kernel launch) as a small JSON for profiles/: the metrics the
roofline discussion in DESIGN.md cites, each with its unit.  usage: ncu_summary.py raw.csv [launch_index] > out.json"""
import csv, json, sys
KEYS = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct"]
rows = list(csv.reader(open(sys.argv[1])))
h, units = rows[0], rows[1]
r = rows[2 + (int(sys.argv[2]) if len(sys.argv) > 2 else 0)]
out = {}
for k in KEYS:
    if k in h:
        i = h.index(k)
        out[k] = [r[i], units[i]]
stalls = {k.split("smsp__average_warps_issue_stalled_")[1].split("_per_issue_active")[0]: float(r[i]) for i, k in enumerate(h)
          if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio") and r[i]}
out["top_stalls_warps_per_issue"] = dict(sorted(stalls.items(), key=lambda kv: -kv[1])[:6])
print(json.dumps(out, indent=1))
