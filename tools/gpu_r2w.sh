#!/usr/bin/env bash
# Round 2, GPU call W: 7 / 8 resident blocks for the own-BVH and tiny-scene kernels (C2, C3, C3b), with and without while-while
set -u
mkdir -p gpurun_out
K="python tools/gpu_knobs.py"
: > gpurun_out/r2w_knobs.jsonl
for v in default ww occ7 occ7ww occ8; do
  if [ $v = default ]; then unset RT_B200_LIB; else export RT_B200_LIB=build/variants/lib$v.so; fi
  timeout 600 $K C2 C3 C3b >> gpurun_out/r2w_knobs.jsonl 2>&1
done
python - <<'PY'
import json
for ln in open('gpurun_out/r2w_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['lib'], d['knobs'], d['ms'], d['all_ms'], d['regs'], d['bps'], d['sha'])
PY
