#!/usr/bin/env bash
# Round 2, GPU call ZD (2 GPUs): the rt_group tests and the CLI's gN path with the current build; bench at N = 2
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "group or two_gpus" > gpurun_out/r2zd_pytest_2gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2zd_pytest_2gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2zd_bench_n2.json 2> gpurun_out/r2zd_bench_n2.err; echo "bench n2 rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2zd_bench_n2.json')); print(d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['value'])"
