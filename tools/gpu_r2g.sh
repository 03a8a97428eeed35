#!/usr/bin/env bash
# Round 2, GPU call G: kernel mode 6 (instanced leaves out of the wrapper-free loop), Perlin chunk rules, a short first chunk
# for the strong-scaling shares of the Cornell box, C1 capture.
set -u
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q -s --deselect "tests/test_gpu_parity.py::test_tier2_converged_images_vs_reference_gpu" ) > gpurun_out/r2g_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2g_pytest.log
K="python tools/gpu_knobs.py"
timeout 400 $K C2 C4 C4:RT_FIRST_CHUNK=8 C4s C4s:RT_FIRST_CHUNK=8 C4s:RT_FIRST_CHUNK=16 C4q C4q:RT_FIRST_CHUNK=8 C1 C5s C3 C3b > gpurun_out/r2g_knobs.jsonl 2>&1
python - <<'PY'
import json
for ln in open('gpurun_out/r2g_knobs.jsonl'):
    try: d = json.loads(ln)
    except Exception: print(ln[:300]); continue
    print(d['config'], d['lib'], d['knobs'], d['ms'], d['all_ms'], d['regs'], d['bps'], d['sha'])
PY
timeout 200 python tools/gpu_configs.py C5s > gpurun_out/r2g_configs.jsonl 2>&1; cut -c1-330 gpurun_out/r2g_configs.jsonl
bash tools/gpu_prof.sh r2g_c1 C1 16 > /dev/null 2>&1; cat gpurun_out/plain_r2g_c1.log
