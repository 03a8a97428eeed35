#!/usr/bin/env bash
# Round 2, GPU call V: own-BVH walk as while-while (leaf tests wait for the lanes that still descend) against the single loop; residency on C2
set -u
mkdir -p gpurun_out
K="python tools/gpu_knobs.py"
timeout 600 $K C2 C2:RT_BLOCKS_PER_SM=5 C2:RT_BLOCKS_PER_SM=4 C2:RT_CARVEOUT=10 C2:RT_CARVEOUT=30 > gpurun_out/r2v_knobs.jsonl 2>&1
RT_B200_LIB=build/variants/libww.so timeout 600 $K C2 C2:RT_CLAIM_GROUP=16 C2:RT_CLAIM_GROUP=1 >> gpurun_out/r2v_knobs.jsonl 2>&1
python - <<'PY'
import json
for ln in open('gpurun_out/r2v_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['lib'], d['knobs'], d['ms'], d['all_ms'], d['sha'])
PY
