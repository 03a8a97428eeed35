#!/usr/bin/env bash
# Round 2, GPU call S: full captures (with source) after the axis-aligned-quad and fused-wrapper changes: C4 s256, C5 800x800 s16
set -u
mkdir -p gpurun_out
timeout 120 python tools/profile_target.py C4 256 2 > gpurun_out/r2s_c4.txt 2>&1; tail -1 gpurun_out/r2s_c4.txt
timeout 120 python tools/profile_target.py C5 16 2 > gpurun_out/r2s_c5.txt 2>&1; tail -1 gpurun_out/r2s_c5.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 1 -c 1 -o gpurun_out/r2s_c4_s256 -f python tools/profile_target.py C4 256 2 > gpurun_out/r2s_ncu_c4.log 2>&1; echo "c4 capture rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 1 -c 1 -o gpurun_out/r2s_c5_s16 -f python tools/profile_target.py C5 16 2 > gpurun_out/r2s_ncu_c5.log 2>&1; echo "c5 capture rc=$?"
cp build/rt_kernels.o gpurun_out/r2s_rt_kernels.o
