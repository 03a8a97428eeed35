#!/usr/bin/env bash
# Scaling record on an N-GPU box: bench.py at 1 and N GPUs, BASELINE config 5 (4K, scene 3, s4096 b50) sample-sharded
# over N GPUs in the statistical mode, and the exact mode tile-sharded at reduced spp.  usage: gpu_scale.sh N [spp_full] [spp_exact]
set -u
N="${1:-8}"; SPP="${2:-4096}"; SPPX="${3:-256}"
mkdir -p gpurun_out
run() { echo "== $*"; "$@"; echo "rc=$?"; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
run timeout 300 python bench.py --gpus 1 --steps 5 --warmup 3 --no-reference-gpu > gpurun_out/scale_bench_n1.json 2> gpurun_out/scale_bench_n1.err
run timeout 300 $TR bench.py --gpus $N --steps 5 --warmup 3 --no-reference-gpu > gpurun_out/scale_bench_n$N.json 2> gpurun_out/scale_bench_n$N.err
tail -1 gpurun_out/scale_bench_n1.json | cut -c1-400; tail -1 gpurun_out/scale_bench_n$N.json | cut -c1-400
run timeout 300 python -m cuda_raytracing_b200.render 3 s$((SPP / N)) b50 w3840 t2 --out=gpurun_out/c5.ppm 2> gpurun_out/scale_c5_fast_n1.json
run timeout 300 $TR -m cuda_raytracing_b200.render 3 s$SPP b50 w3840 t2 --out=gpurun_out/c5.ppm 2> gpurun_out/scale_c5_fast_samples_n$N.json
run timeout 300 python -m cuda_raytracing_b200.render 3 s$((SPPX / N)) b50 w3840 t0 --out=gpurun_out/c5x1.ppm 2> gpurun_out/scale_c5_exact_n1.json
run timeout 300 $TR -m cuda_raytracing_b200.render 3 s$((SPPX / N)) b50 w3840 t0 --shard=tiles --out=gpurun_out/c5xn.ppm 2> gpurun_out/scale_c5_exact_tiles_n$N.json
cmp gpurun_out/c5x1.ppm gpurun_out/c5xn.ppm && echo "exact mode, tiles x$N image == 1-GPU image (byte for byte)"
run timeout 300 $TR -m cuda_raytracing_b200.render 3 s$SPPX b50 w3840 t0 --shard=tiles --out=gpurun_out/c5xn.ppm 2> gpurun_out/scale_c5_exact_tiles_full_n$N.json
rm -f gpurun_out/*.ppm
for f in gpurun_out/scale_c5_*.json; do echo "$f: $(grep -h '^{' $f | tail -1 | cut -c1-420)"; done
