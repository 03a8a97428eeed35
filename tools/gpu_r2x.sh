#!/usr/bin/env bash
# Round 2, GPU call X: how far the halving tail of the chunk schedule goes down (8 samples today; 4, 2, 1)
set -u
mkdir -p gpurun_out
K="python tools/gpu_knobs.py"
: > gpurun_out/r2x_knobs.jsonl
for c in C4t C4s C4q C4 C2 C5s C3b; do
  timeout 600 $K $c $c:RT_TAIL_MIN=4 $c:RT_TAIL_MIN=2 $c:RT_TAIL_MIN=1 $c:RT_TAIL_MIN=16 >> gpurun_out/r2x_knobs.jsonl 2>&1
done
python - <<'PY'
import json
for ln in open('gpurun_out/r2x_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['knobs'], d['ms'], d['all_ms'], d['identical_to_first'])
PY
