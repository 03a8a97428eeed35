#!/usr/bin/env bash
# Round 2, GPU call ZJ: pooled claims (a warp takes 64 ... 256 adjacent items at once, its lanes draw from the pool as they finish) on C2;
# before: C2 91.4-91.8, C4 227.1-227.6, C1 0.92 ms
set -u
mkdir -p gpurun_out
K="python tools/gpu_knobs.py"
timeout 300 $K C2 C2:RT_POOL=64 C2:RT_POOL=96 C2:RT_POOL=128 C2:RT_POOL=256 C2:RT_POOL=64,RT_CHUNK=32 C4 C1 > gpurun_out/r2zj_knobs.jsonl 2>&1
python - <<'PY'
import json
for ln in open('gpurun_out/r2zj_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['knobs'], d['ms'], d['all_ms'], d['identical_to_first'])
PY
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "scheduling_knobs" > gpurun_out/r2zj_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2zj_pytest.log
