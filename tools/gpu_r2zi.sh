#!/usr/bin/env bash
# Round 2, GPU call ZI: hybrid walk (mode 2): the reference's BVH node case in front of the switch (range test); before: C5s 102.4-102.8, C5m 376, C5w 1201 ms
set -u
mkdir -p gpurun_out
K="python tools/gpu_knobs.py"
timeout 600 $K C5s C5m C5w C1 C1 > gpurun_out/r2zi_knobs.jsonl 2>&1
python - <<'PY'
import json
for ln in open('gpurun_out/r2zi_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['lib'], d['knobs'], d['ms'], d['all_ms'], d['sha'])
PY
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "baseline_configs or hybrid_walk or image_vs_oracle or trace_probe" > gpurun_out/r2zi_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2zi_pytest.log
