#!/usr/bin/env bash
# Round 2, GPU call ZK: C5 exactly as BASELINE.json names it (./raytracing 3 s4096 b50 at 3840 wide) on ONE GPU with the final kernels,
# statistical mode (t2), measured directly instead of extrapolated from s512 / s1024
set -u
mkdir -p gpurun_out
cd gpurun_out
( time timeout 200 ../bin/raytracing 3 s4096 b50 w3840 t2 q ) > r2zk_c5_4k_s4096_1gpu.txt 2>&1; echo "rc=$?"
rm -f *.ppm
cat r2zk_c5_4k_s4096_1gpu.txt | cut -c1-400
