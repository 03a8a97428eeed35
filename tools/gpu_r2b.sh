#!/usr/bin/env bash
# Round 2, GPU call B: parity suite after the guided chunk schedule + path-end noise resolution + group API, per-config
# timings against the per-bounce noise variant, the tier-2 PSNR matrix of scenes 3 and 1.
set -u
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q -s --deselect "tests/test_gpu_parity.py::test_tier2_converged_images_vs_reference_gpu" ) > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2b_pytest.log
grep -h "baseline C" gpurun_out/r2b_pytest.log | sed 's/^\.*//'
K="python tools/gpu_knobs.py"
timeout 300 $K C4 C4h C4q C4s C4t C2 C1 C3 C3b C5s C5m > gpurun_out/r2b_knobs_default.jsonl 2>&1
RT_B200_LIB=build/variants/libnoisenow.so timeout 200 $K C3 C3b C5s > gpurun_out/r2b_knobs_noisenow.jsonl 2>&1
cat gpurun_out/r2b_knobs_*.jsonl | cut -c1-330
timeout 600 python tools/gpu_tier2_matrix.py 3 200 4096 3 > gpurun_out/r2b_tier2_matrix_scene3.jsonl 2>&1; cat gpurun_out/r2b_tier2_matrix_scene3.jsonl
timeout 300 python tools/gpu_tier2_matrix.py 1 128 4096 3 > gpurun_out/r2b_tier2_matrix_scene1.jsonl 2>&1; cat gpurun_out/r2b_tier2_matrix_scene1.jsonl
( time timeout 600 python bench.py --steps 5 --warmup 3 ) > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; echo "bench rc=$?"; tail -4 gpurun_out/r2b_bench.err
