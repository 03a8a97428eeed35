#!/usr/bin/env bash
# Round 2, GPU call Z: why the Cornell box renders 8 % fewer rays/s at s128 than at s1024 -- resident warps, instructions per ray and lanes per instruction at s32 / s128 / s512
set -u
mkdir -p gpurun_out
M="sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,gpu__time_duration.sum,sm__cycles_active.avg,sm__cycles_elapsed.max"
for s in 32 128 512; do
  timeout 300 ncu --metrics $M --clock-control none -k regex:render_kernel -s 1 -c 1 --csv --log-file gpurun_out/r2z_c4_s$s.csv python tools/profile_target.py C4 $s 2 > gpurun_out/r2z_c4_s$s.txt 2>&1
  echo "s$s rc=$?"; tail -1 gpurun_out/r2z_c4_s$s.txt
done
