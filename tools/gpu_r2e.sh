#!/usr/bin/env bash
# Round 2, GPU call E (2 GPUs): whole parity suite incl. the rt_group tests, tier-2 matrix of scene 3 with instanced leaves,
# FAST timings, 2-GPU bench (torchrun) and the CLI on 2 GPUs.
set -u
mkdir -p gpurun_out
nvidia-smi -L
( time timeout 900 python -m pytest tests -m gpu -x -q -s --deselect "tests/test_gpu_parity.py::test_tier2_converged_images_vs_reference_gpu" ) > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2e_pytest.log
grep -h "fast-vs-oracle" gpurun_out/r2e_pytest.log | cut -c1-400
timeout 600 python tools/gpu_tier2_matrix.py 3 200 4096 2 > gpurun_out/r2e_tier2_matrix_scene3.jsonl 2>&1; cat gpurun_out/r2e_tier2_matrix_scene3.jsonl
timeout 300 python tools/gpu_fast_bias.py 3 200 1024 > gpurun_out/r2e_fast_bias_scene3.jsonl 2>&1; head -6 gpurun_out/r2e_fast_bias_scene3.jsonl
timeout 300 python tools/gpu_configs.py C5s C4 > gpurun_out/r2e_configs.jsonl 2>&1; cut -c1-400 gpurun_out/r2e_configs.jsonl
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 ) > gpurun_out/r2e_bench_n2.json 2> gpurun_out/r2e_bench_n2.err; echo "bench n2 rc=$?"; tail -3 gpurun_out/r2e_bench_n2.err; cut -c1-1500 gpurun_out/r2e_bench_n2.json
mkdir -p /tmp/cli && cd /tmp/cli
for extra in "" "g2" "g2 d1"; do $GRAFT_REPO_ROOT/bin/raytracing 1 s1024 b50 $extra q 2>&1 | tail -2 | cut -c1-400; done
for extra in "" "g2" "g2 d1"; do $GRAFT_REPO_ROOT/bin/raytracing 3 s64 b50 t2 $extra q 2>&1 | tail -1 | cut -c1-400; done
