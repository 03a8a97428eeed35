#!/usr/bin/env bash
# Round 2, GPU call D: asynchronous progress poll + sqrt-free rejection + first-chunk knob; FAST geometry against the oracle.
set -u
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q -s --deselect "tests/test_gpu_parity.py::test_tier2_converged_images_vs_reference_gpu" ) > gpurun_out/r2d_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2d_pytest.log
grep -h "fast-vs-oracle" gpurun_out/r2d_pytest.log | cut -c1-900
K="python tools/gpu_knobs.py"
timeout 400 $K C4 C4s C2 C1 C5s C5m C3 C3:RT_FIRST_CHUNK=8 C3b C3b:RT_FIRST_CHUNK=8 C3b:RT_FIRST_CHUNK=16 C3b:RT_FIRST_CHUNK=8,RT_CHUNK=64 C3b:RT_CHUNK=128 C3b:RT_CHUNK=256 > gpurun_out/r2d_knobs.jsonl 2>&1
python - <<'PY'
import json
for ln in open('gpurun_out/r2d_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['knobs'], d['ms'], d['all_ms'], d['regs'], d['bps'], d['identical_to_first'])
PY
