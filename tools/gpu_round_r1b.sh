#!/usr/bin/env bash
# Record of the stackless reference-order walk (kernel mode 3) and the item-order rules: parity suite, per-config table
# against the reference binary, bench (both arms), one ncu capture of the Cornell box kernel.
set -u
mkdir -p gpurun_out
timeout 200 python -m pytest tests -m gpu -x -q > gpurun_out/r1b_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r1b_pytest.log
timeout 200 python tools/gpu_configs.py C1 C2 C3 C3b C4 C5s > gpurun_out/r1b_configs.jsonl 2> gpurun_out/r1b_configs.err; echo "configs rc=$?"
timeout 200 python bench.py --steps 5 --warmup 3 > gpurun_out/r1b_bench.json 2> gpurun_out/r1b_bench.err; echo "bench rc=$?"
timeout 100 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1b_bench_ref.json 2> gpurun_out/r1b_bench_ref.err; echo "bench ref rc=$?"
bash tools/gpu_prof.sh c4_ropes C4 64 > /dev/null 2>&1
timeout 60 python __graft_entry__.py smoke > gpurun_out/r1b_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r1b_smoke.log
cat gpurun_out/r1b_configs.jsonl
