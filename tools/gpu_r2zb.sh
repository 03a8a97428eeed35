#!/usr/bin/env bash
# Round 2, GPU call ZB: traversal stack walked by address with a bottom sentinel (modes 0, 1, 2, 6); before: C2 94.3-95.7, C5s 105.9-107.2, C5m 389, C5w 1247 ms
set -u
mkdir -p gpurun_out
K="python tools/gpu_knobs.py"
timeout 600 $K C2 C5s C5m C5w C4 C1 > gpurun_out/r2zb_knobs.jsonl 2>&1
python - <<'PY'
import json
for ln in open('gpurun_out/r2zb_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['lib'], d['knobs'], d['ms'], d['all_ms'], d['sha'])
PY
( time timeout 1200 python -m pytest tests -m gpu -x -q --deselect "tests/test_gpu_parity.py::test_tier2_converged_images_vs_reference_gpu" ) > gpurun_out/r2zb_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2zb_pytest.log
RT_B200_LIB=build/variants/libstackcheck.so timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "image_vs_oracle or trace_probe or hybrid_walk or fast_traversal" > gpurun_out/r2zb_pytest_stackcheck.log 2>&1; echo "stackcheck pytest rc=$?"; tail -2 gpurun_out/r2zb_pytest_stackcheck.log
