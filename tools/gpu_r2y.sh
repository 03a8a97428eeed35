#!/usr/bin/env bash
# Round 2, GPU call Y: kernel time against spp on the Cornell box (t = 0.27 ms x spp + 3.4 ms from s64 up: where does the fixed part come from?)
set -u
mkdir -p gpurun_out
timeout 300 python tools/gpu_spp_sweep.py 1 50 1 2 4 8 16 32 64 128 > gpurun_out/r2y_spp.jsonl 2>&1
RT_CHUNK=1000000 timeout 300 python tools/gpu_spp_sweep.py 1 50 1 2 4 8 16 32 64 >> gpurun_out/r2y_spp.jsonl 2>&1
RT_SCATTER=1 timeout 300 python tools/gpu_spp_sweep.py 1 50 8 32 128 >> gpurun_out/r2y_spp.jsonl 2>&1
cat gpurun_out/r2y_spp.jsonl | cut -c1-260
