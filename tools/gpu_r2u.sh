#!/usr/bin/env bash
# Round 2, GPU call U (third run): linear walk addressed relative to the op at hand, end op instead of a bound; before: C4 279.0-279.6, C4s 38.06-38.25, C4t 20.63, C1 0.944-0.962 ms, general inside test 297.5
set -u
mkdir -p gpurun_out
K="python tools/gpu_knobs.py"
timeout 600 $K C4 C4s C4t C1 C4:RT_AXIS_QUADS=0 > gpurun_out/r2u_knobs.jsonl 2>&1
python - <<'PY'
import json
for ln in open('gpurun_out/r2u_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['lib'], d['knobs'], d['ms'], d['all_ms'], d['sha'])
PY
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "baseline_configs or trace_probe or image_vs_oracle or scheduling_knobs" > gpurun_out/r2u_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2u_pytest.log
