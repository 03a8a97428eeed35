#!/usr/bin/env bash
# Round 2, GPU call ZG: flag-bit dispatch of the two frequent ops in the linear walk, range-test node check in the own-BVH walk; before: C4 258.4-259.6, C4s 35.4, C1 0.94, C2 92.9-93.1, C3b 64.1 ms
set -u
mkdir -p gpurun_out
K="python tools/gpu_knobs.py"
timeout 600 $K C4 C4s C1 C2 C3b C5s:RT_TRAVERSAL=0 > gpurun_out/r2zg_knobs.jsonl 2>&1
python - <<'PY'
import json
for ln in open('gpurun_out/r2zg_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['lib'], d['knobs'], d['ms'], d['all_ms'], d['sha'])
PY
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "baseline_configs or trace_probe or image_vs_oracle or instrumented or fast_traversal" > gpurun_out/r2zg_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2zg_pytest.log
