#!/usr/bin/env bash
# Round 2, GPU call P: claim-group / tile-shape / chunk sweep on C2 with the current kernels (the ncu capture of call O shows
# 14.3 of 32 lanes holding a live pixel at a bounce on C2: lanes that finished their chunk wait for the warp's last one)
set -u
mkdir -p gpurun_out
K="python tools/gpu_knobs.py"
timeout 600 $K C2 C2:RT_CLAIM_GROUP=1 C2:RT_CLAIM_GROUP=4 C2:RT_CLAIM_GROUP=8 C2:RT_CLAIM_GROUP=16 C2:RT_CLAIM_GROUP=24 \
   C2:RT_CLAIM_GROUP=16,RT_TILE_W=4 C2:RT_CLAIM_GROUP=8,RT_TILE_W=4 C2:RT_TILE_W=4 C2:RT_TILE_W=16 \
   C2:RT_CHUNK=16 C2:RT_CHUNK=32 C2:RT_CHUNK=128 C2:RT_CLAIM_GROUP=8,RT_CHUNK=32 C2:RT_CLAIM_GROUP=16,RT_CHUNK=32 > gpurun_out/r2p_knobs.jsonl 2>&1
python - <<'PY'
import json
for ln in open('gpurun_out/r2p_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['knobs'], d['ms'], d['all_ms'], d['identical_to_first'])
PY
