#!/usr/bin/env bash
# Round 2, GPU call ZA: C1 (1 ms, one chunk per pixel): does slicing pay now?  scattered order?  mode 3 instead of 4?
set -u
mkdir -p gpurun_out
K="python tools/gpu_knobs.py"
timeout 600 $K C1 C1:RT_CHUNK=8 C1:RT_CHUNK=4 C1:RT_CHUNK=12 C1:RT_SCATTER=1 C1:RT_SCATTER=1,RT_CHUNK=8 C1:RT_TILE_W=4 C1:RT_TILE_W=16 C1:RT_CLAIM_GROUP=32 C1:RT_CLAIM_GROUP=8 C1:RT_BLOCKS_PER_SM=4 C1:RT_BLOCKS_PER_SM=3 > gpurun_out/r2za_knobs.jsonl 2>&1
python - <<'PY'
import json
for ln in open('gpurun_out/r2za_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['knobs'], d['ms'], d['all_ms'], d['regs'], d['bps'], d['identical_to_first'])
PY
