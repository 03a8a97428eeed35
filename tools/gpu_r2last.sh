#!/usr/bin/env bash
# Round 2, closing check: bench.py with no flags, as the driver runs it (roofline.traffic must come from the capture of the current sources)
set -u
mkdir -p gpurun_out
( time timeout 600 python bench.py ) > gpurun_out/r2last_bench_default.json 2> gpurun_out/r2last_bench_default.err; echo "bench rc=$?"; tail -3 gpurun_out/r2last_bench_default.err
python -c "
import json; d=json.load(open('gpurun_out/r2last_bench_default.json')); r=d['roofline']; print(d['ms_per_step'], d['value'], d['e2e']['value'], r['frac'], r['traffic'], r['ncu'].get('stale') if isinstance(r['ncu'], dict) else r['ncu'])"
