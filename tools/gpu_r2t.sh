#!/usr/bin/env bash
# Round 2, GPU call T: the hybrid walk (kernel mode 2, scene 3) with one kind of item per step chosen by vote, against the switch
set -u
mkdir -p gpurun_out
K="python tools/gpu_knobs.py"
RT_B200_LIB=build/variants/libnovote.so timeout 600 $K C5s C5m C5w C4 C2 > gpurun_out/r2t_knobs.jsonl 2>&1
timeout 600 $K C5s C5m C5w C5s:RT_SCATTER=0 C4 C2 >> gpurun_out/r2t_knobs.jsonl 2>&1
python - <<'PY'
import json
for ln in open('gpurun_out/r2t_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['lib'], d['knobs'], d['ms'], d['all_ms'], d['sha'])
PY
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "baseline_configs or hybrid_walk or image_vs_oracle" > gpurun_out/r2t_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2t_pytest.log
