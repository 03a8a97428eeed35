#!/usr/bin/env bash
# Round 2, GPU call ZC: full captures (with source) of the other walks with the current build: C2 (mode 1), C3b (mode 5), C5 800x800 (mode 2)
# (gpurun brings back at most 64 MiB: three reports of ~14 MB)
set -u
mkdir -p gpurun_out
cap() { timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 1 -c 1 -o gpurun_out/r2zc_$1 -f python tools/profile_target.py $2 $3 2 > gpurun_out/r2zc_ncu_$1.log 2>&1; echo "$1 rc=$?"; }
cap c2_s32 C2 32
cap c3b_s64 C3b 64
cap c5_s16 C5 16
ls -la gpurun_out
