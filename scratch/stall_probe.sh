#!/bin/bash
# usage: stall_probe.sh "<env list>" [team]: per-issue stall reasons and icache hit rate of k_run_par at several batch sizes
M=smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio,smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio,smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio,smsp__average_warps_issue_stalled_imc_miss_per_issue_active.ratio,sm__icc_request_hit_rate.pct,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,gpu__time_duration.sum,sm__warps_active.avg.pct_of_peak_sustained_active
for e in $1; do
  ncu --metrics $M --clock-control none -k regex:k_run_par -s 2 -c 1 --csv --log-file gpurun_out/probe_e${e}.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --envs $e --team ${2:-0} --gym-steps 0 > gpurun_out/probe_e${e}.log 2>&1
  python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/probe_e${e}.csv")) if len(r)>5]
h=rows[0]; i_n=h.index("Metric Name"); i_v=h.index("Metric Value")
print("envs", $e, " ".join("%s=%s"%(r[i_n].replace("smsp__average_warps_issue_stalled_","").replace("_per_issue_active.ratio",""), r[i_v]) for r in rows[1:]))
PY
done
