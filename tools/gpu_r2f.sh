#!/usr/bin/env bash
# Round 2, GPU call F: one vote per iteration at the loop top; Perlin scenes against chunk sizes and residency variants.
set -u
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q -s --deselect "tests/test_gpu_parity.py::test_tier2_converged_images_vs_reference_gpu" ) > gpurun_out/r2f_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2f_pytest.log
K="python tools/gpu_knobs.py"
timeout 400 $K C4 C4s C2 C1 C5s C3 C3:RT_CHUNK=256 C3:RT_CHUNK=512 C3:RT_CHUNK=64 C3b C3b:RT_CHUNK=184 C3b:RT_CHUNK=128 > gpurun_out/r2f_knobs.jsonl 2>&1
for v in occ5 occ4; do RT_B200_LIB=build/variants/lib$v.so timeout 200 $K C3 C3b C2 C1 >> gpurun_out/r2f_knobs.jsonl 2>&1; done
python - <<'PY'
import json
for ln in open('gpurun_out/r2f_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['lib'], d['knobs'], d['ms'], d['all_ms'], d['regs'], d['bps'], d['sha'])
PY
