#!/usr/bin/env bash
# one short GPU visit: item-order experiment, then the parity suite and the bench with the automatic choice
set -u
mkdir -p gpurun_out
timeout 240 python tools/gpu_scatter.py > gpurun_out/scatter.jsonl 2> gpurun_out/scatter.err; echo "scatter rc=$?"; cat gpurun_out/scatter.jsonl
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/scatter_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/scatter_pytest.log
timeout 300 python bench.py --steps 5 --warmup 3 > gpurun_out/scatter_bench.json 2> gpurun_out/scatter_bench.err; echo "bench rc=$?"; cat gpurun_out/scatter_bench.json
timeout 300 python tools/gpu_configs.py C1 C3 C3b C4 > gpurun_out/scatter_configs.jsonl 2> gpurun_out/scatter_configs.err; echo "configs rc=$?"; cat gpurun_out/scatter_configs.jsonl
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/scatter_bench_ref.json 2> gpurun_out/scatter_bench_ref.err; echo "bench ref rc=$?"
