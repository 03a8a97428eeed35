#!/usr/bin/env bash
# Round 2, GPU call ZF (8 GPUs): bench at N = 2, 4 and 8 with the final kernels (N = 1: r2fin_bench_n1.json)
set -u
mkdir -p gpurun_out
for n in 2 4 8; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29520 + n)) bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/r2zf_bench_n$n.json 2> gpurun_out/r2zf_bench_n$n.err; echo "bench n$n rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/r2zf_bench_n$n.json')); print(d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['value'], d['per_rank'])"
done
