#!/usr/bin/env bash
# Round 2, GPU call R: translate(rotate(x)) as one op each way in the linear program, against one op per wrapper side
set -u
mkdir -p gpurun_out
K="python tools/gpu_knobs.py"
timeout 600 $K C4:RT_FUSE_PROGRAM_WRAPPERS=0 C4 C4s:RT_FUSE_PROGRAM_WRAPPERS=0 C4s C1:RT_FUSE_PROGRAM_WRAPPERS=0 C1 > gpurun_out/r2r_knobs.jsonl 2>&1
python - <<'PY'
import json
for ln in open('gpurun_out/r2r_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['knobs'], d['ms'], d['all_ms'], d['identical_to_first'], d['sha'])
PY
( time timeout 1200 python -m pytest tests -m gpu -x -q --deselect "tests/test_gpu_parity.py::test_tier2_converged_images_vs_reference_gpu" ) > gpurun_out/r2r_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2r_pytest.log
