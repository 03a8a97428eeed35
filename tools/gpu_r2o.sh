#!/usr/bin/env bash
# Round 2, GPU call O: fresh full captures (with source) of the render kernel on C4 and C2 for the per-line breakdown
set -u
mkdir -p gpurun_out
timeout 120 python tools/profile_target.py C4 256 2 > gpurun_out/r2o_c4.txt 2>&1; tail -1 gpurun_out/r2o_c4.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 1 -c 1 -o gpurun_out/r2o_c4_s256 -f python tools/profile_target.py C4 256 2 > gpurun_out/r2o_ncu_c4.log 2>&1; echo "c4 capture rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 1 -c 1 -o gpurun_out/r2o_c2_s32 -f python tools/profile_target.py C2 32 2 > gpurun_out/r2o_ncu_c2.log 2>&1; echo "c2 capture rc=$?"
ls -la gpurun_out/*.ncu-rep
