#!/usr/bin/env bash
# Round 2, GPU call Q: the reduced inside test of axis-aligned quads (linear program of the Cornell box) against the general one
set -u
mkdir -p gpurun_out
K="python tools/gpu_knobs.py"
timeout 600 $K C4:RT_AXIS_QUADS=0 C4 C4s:RT_AXIS_QUADS=0 C4s C4t:RT_AXIS_QUADS=0 C4t > gpurun_out/r2q_knobs.jsonl 2>&1
python - <<'PY'
import json
for ln in open('gpurun_out/r2q_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['knobs'], d['ms'], d['all_ms'], d['identical_to_first'], d['sha'])
PY
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "baseline_configs or trace_probe or image_vs_oracle or all_scenes" > gpurun_out/r2q_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2q_pytest.log
