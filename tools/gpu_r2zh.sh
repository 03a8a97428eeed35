#!/usr/bin/env bash
# Round 2, GPU call ZH: what the flag-bit dispatch changed on C4 (instructions, lanes per instruction, issue slots) -- and on C1
set -u
mkdir -p gpurun_out
M="sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,gpu__time_duration.sum"
timeout 300 ncu --metrics $M --clock-control none -k regex:render_kernel -s 1 -c 1 --csv --log-file gpurun_out/r2zh_c4_s256.csv python tools/profile_target.py C4 256 2 > gpurun_out/r2zh_c4.txt 2>&1; echo "c4 rc=$?"
timeout 300 ncu --metrics $M --clock-control none -k regex:render_kernel -s 1 -c 1 --csv --log-file gpurun_out/r2zh_c1_s16.csv python tools/profile_target.py C1 16 2 > gpurun_out/r2zh_c1.txt 2>&1; echo "c1 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 1 -c 1 -o gpurun_out/r2zh_c4_s256 -f python tools/profile_target.py C4 256 2 > gpurun_out/r2zh_ncu_c4.log 2>&1; echo "c4 full rc=$?"
