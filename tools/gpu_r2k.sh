#!/usr/bin/env bash
# Round 2, GPU call K: four spheres in one basic block for the tiny-scene walk (scene 19) against the pair-by-pair build.
set -u
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q -s --deselect "tests/test_gpu_parity.py::test_tier2_converged_images_vs_reference_gpu" ) > gpurun_out/r2k_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2k_pytest.log
K="python tools/gpu_knobs.py"
timeout 300 $K C3b C3 C2 C4 C1 > gpurun_out/r2k_knobs.jsonl 2>&1
RT_B200_LIB=build/variants/libflat2.so timeout 300 $K C3b C3 >> gpurun_out/r2k_knobs.jsonl 2>&1
RT_B200_LIB=build/variants/libpairs.so timeout 300 $K C2 C2:RT_PAIR_LEAVES=0 C4 >> gpurun_out/r2k_knobs.jsonl 2>&1
RT_B200_LIB=build/variants/libpairs.so timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "baseline_configs or trace_probe or image_vs_oracle" > gpurun_out/r2k_pytest_pairs.log 2>&1; echo "pairs pytest rc=$?"; tail -2 gpurun_out/r2k_pytest_pairs.log
python - <<'PY'
import json
for ln in open('gpurun_out/r2k_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['lib'], d['knobs'], d['ms'], d['all_ms'], d['regs'], d['bps'], d['sha'])
PY
