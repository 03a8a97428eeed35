#!/usr/bin/env bash
# Round 2, GPU call C: chunk rule (equal bulk + halving tail) against the alternatives, where the Perlin scenes spend their
# time (per-band times, ncu source-level captures), where RT_TRAVERSAL_FAST differs from the exact walk on scene 3.
set -u
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q -s --deselect "tests/test_gpu_parity.py::test_tier2_converged_images_vs_reference_gpu" ) > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2c_pytest.log
K="python tools/gpu_knobs.py"
timeout 400 $K C4 C4h C4q C4s C4t C2 C5s C5m C5m:RT_CHUNK=13 C5m:RT_CHUNK=8 C3b C3b:RT_CHUNK=92 C3b:RT_CHUNK=32 C3b:RT_CHUNK=16 C3b:RT_CHUNK=8 C3b:RT_SCATTER=1,RT_CHUNK=16 \
  C3b:RT_SCATTER=1,RT_CHUNK=64 C3 C3:RT_CHUNK=32 C3:RT_CHUNK=16 C3:RT_SCATTER=0,RT_CHUNK=16 C3:RT_BLOCKS_PER_SM=3 C3:RT_BLOCKS_PER_SM=2 C3b:RT_BLOCKS_PER_SM=4 > gpurun_out/r2c_knobs.jsonl 2>&1
cut -c1-300 gpurun_out/r2c_knobs.jsonl
timeout 120 python tools/gpu_band_times.py 17 64 16 > gpurun_out/r2c_bands_17.json 2>&1; cat gpurun_out/r2c_bands_17.json
timeout 120 python tools/gpu_band_times.py 19 64 16 > gpurun_out/r2c_bands_19.json 2>&1; cat gpurun_out/r2c_bands_19.json
timeout 600 python tools/gpu_fast_bias.py 3 200 2048 > gpurun_out/r2c_fast_bias_scene3.jsonl 2>&1; cat gpurun_out/r2c_fast_bias_scene3.jsonl
bash tools/gpu_prof.sh r2c_c3 C3 128 > /dev/null 2>&1; cat gpurun_out/plain_r2c_c3.log
bash tools/gpu_prof.sh r2c_c3b C3b 64 > /dev/null 2>&1; cat gpurun_out/plain_r2c_c3b.log
ls -la gpurun_out/*.ncu-rep
