#!/usr/bin/env bash
# Round record on one B200: parity tests, per-config table vs the reference binary, bench (both arms), and the ncu
# evidence for the bench command (launch list + one full capture of the render kernel).  Outputs under gpurun_out/final_*.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/final_pytest.log 2>&1; tail -2 gpurun_out/final_pytest.log
timeout 900 python tools/gpu_configs.py C1 C2 C3 C3b C4 C5s > gpurun_out/final_configs.jsonl 2> gpurun_out/final_configs.err
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/final_bench_ref.json 2> gpurun_out/final_bench_ref.err; echo "bench ref rc=$?"
timeout 600 python bench.py --steps 2 --warmup 3 --no-reference-gpu > gpurun_out/final_bench_short.json 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/final_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-reference-gpu > gpurun_out/final_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 3 -c 1 -o gpurun_out/final_render_kernel -f \
  python bench.py --steps 2 --warmup 3 --no-reference-gpu > gpurun_out/final_ncu_full.log 2>&1
ls -la gpurun_out/final_*
