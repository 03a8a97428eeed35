#!/usr/bin/env bash
# Round 2, GPU call I: program + records of the small scenes staged in shared memory (variant) against the default build.
set -u
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q -s --deselect "tests/test_gpu_parity.py::test_tier2_converged_images_vs_reference_gpu" ) > gpurun_out/r2i_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2i_pytest.log
K="python tools/gpu_knobs.py"
timeout 300 $K C4 C4s C1 C2 C3 > gpurun_out/r2i_knobs.jsonl 2>&1
RT_B200_LIB=build/variants/liblinsmem.so timeout 300 $K C4 C4s C1 >> gpurun_out/r2i_knobs.jsonl 2>&1
python - <<'PY'
import json
for ln in open('gpurun_out/r2i_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['lib'], d['knobs'], d['ms'], d['all_ms'], d['regs'], d['bps'], d['sha'])
PY
