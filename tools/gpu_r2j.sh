#!/usr/bin/env bash
# Round 2, GPU call J (8 GPUs of one box): the strong-scaling curve of bench.py (C4 s1024) at 1/2/4/8 GPUs, C5 as BASELINE
# names it (4K, s4096, b50) sample-sharded over 8 GPUs through the CLI (rt_group + NCCL), the exact mode tile-sharded at 4K.
set -u
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 300 python bench.py --gpus 1 --steps 5 --warmup 3 --no-configs > gpurun_out/r2j_bench_n1.json 2> gpurun_out/r2j_bench_n1.err; echo "bench n1 rc=$?"
for n in 2 4 8; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/r2j_bench_n$n.json 2> gpurun_out/r2j_bench_n$n.err; echo "bench n$n rc=$?"
done
python - <<'PY'
import json
base = None
for n in (1, 2, 4, 8):
    try:
        d = json.loads(open('gpurun_out/r2j_bench_n%d.json' % n).read().strip().splitlines()[-1])
    except Exception as e:
        print(n, "no line", e); continue
    if n == 1: base = d['ms_per_step']
    pr = d['per_rank']
    print("N=%d  %.3f ms/step  %.4g paths/s  e2e %.4g  efficiency %.3f  render min/max %.2f/%.2f  reduce+resolve(rank0) %.3f" % (
        n, d['ms_per_step'], d['value'], d['e2e']['value'], base / (n * d['ms_per_step']) if base else 0, pr['render_ms_min'], pr['render_ms_max'], pr['reduce_resolve_ms_rank0']))
PY
mkdir -p /tmp/cli && cd /tmp/cli
R=$GRAFT_REPO_ROOT/bin/raytracing
run() { echo "== raytracing $*"; $R "$@" q 2> err.txt | tail -2 | head -1; tail -1 err.txt | cut -c1-420; }
run 1 s1024 b50 g8 d1
run 1 s1024 b50 g8 d0
run 3 s4096 b50 w3840 t2 g8 d1
run 3 s512 b50 w3840 t2
run 3 s1024 b50 w3840 t2
run 3 s64 b50 w3840 g8 d0
run 3 s64 b50 w3840
run 2 s256 b50 g8 d1
