#!/usr/bin/env bash
# BASELINE config 5 (4K render of the heaviest scene, sample-sharded over the GPUs of the box) at reduced spp:
# ours on 1 and N GPUs (both shardings), and the reference's own binary at 4K for its paths/s.  usage: gpu_c5.sh N [spp]
set -u
N="${1:-2}"; SPP="${2:-64}"
mkdir -p gpurun_out
run() { echo "== $*"; "$@"; echo "rc=$?"; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
run timeout 300 python -m cuda_raytracing_b200.render 3 s$SPP b50 w3840 t2 --out=gpurun_out/c5_fast_n1.ppm
run timeout 300 $TR -m cuda_raytracing_b200.render 3 s$SPP b50 w3840 t2 --out=gpurun_out/c5_fast_samples_n$N.ppm
run timeout 300 $TR -m cuda_raytracing_b200.render 3 s$SPP b50 w3840 t2 --shard=tiles --out=gpurun_out/c5_fast_tiles_n$N.ppm
cmp gpurun_out/c5_fast_n1.ppm gpurun_out/c5_fast_tiles_n$N.ppm && echo "tiles x$N image == 1-GPU image (byte for byte)"
run timeout 300 python -m cuda_raytracing_b200.render 3 s4 b50 w3840 t0 --out=gpurun_out/c5_exact_n1.ppm
( cd gpurun_out && RT_WIDTH=3840 RT_NO_PPM=1 RT_DUMP_F32=/dev/null timeout 600 ../oracle/_ref/raytracing_ref_ext 3 s4 b50 ) | sed 's/^/reference 4K s4: /'
rm -f gpurun_out/*.ppm
