#!/usr/bin/env bash
# Round 2, GPU call A: full parity suite (incl. the full-size BASELINE configs), bench (both arms), then experiments:
# how the Cornell box scales with spp (strong-scaling shares), the zero-numerator / op-prefetch variants of the linear
# program, the forward-throughput mode.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
( time timeout 1500 python -m pytest tests -m gpu -x -q -s ) > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2a_pytest.log
grep -h "^baseline\|^tier2" gpurun_out/r2a_pytest.log
( time timeout 600 python bench.py --steps 5 --warmup 3 ) > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"; tail -4 gpurun_out/r2a_bench.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2a_bench_ref.json 2> gpurun_out/r2a_bench_ref.err; echo "bench ref rc=$?"
K="python tools/gpu_knobs.py"
timeout 300 $K C4 C4h C4q C4s C4t C4s:RT_CHUNK=4 C4s:RT_CHUNK=16 C4s:RT_CHUNK=32 C4s:RT_CHUNK=128 C4:FLAGS=32 C2 C2:FLAGS=32 C1 C3 C3:FLAGS=32 > gpurun_out/r2a_knobs_default.jsonl 2>&1
for v in prefetch prefetch6 lin6 lin7; do
  RT_B200_LIB=build/variants/lib$v.so timeout 200 $K C4 C4s C1 > gpurun_out/r2a_knobs_$v.jsonl 2>&1
done
cat gpurun_out/r2a_knobs_*.jsonl | cut -c1-420
timeout 120 python __graft_entry__.py smoke > gpurun_out/r2a_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r2a_smoke.log
