#!/usr/bin/env bash
# Round 2, GPU call ZE: the instrumented kernel on a scene of axis-aligned quads (it read reduced records with the general test), smoke with scene 1
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "instrumented or edge_cases" > gpurun_out/r2ze_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2ze_pytest.log
timeout 120 python __graft_entry__.py smoke > gpurun_out/r2ze_smoke.log 2>&1; echo "smoke rc=$?"; tail -5 gpurun_out/r2ze_smoke.log
timeout 300 python bench.py --steps 3 --warmup 3 --no-configs --no-reference-gpu > gpurun_out/r2ze_bench.json 2> gpurun_out/r2ze_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2ze_bench.json')); r=d['roofline']; print(d['ms_per_step'], r['frac'], r['flops_per_ray'], r['rays_per_path'], r['mrays_per_s'])"
