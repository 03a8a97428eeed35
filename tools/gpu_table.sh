#!/usr/bin/env bash
# BASELINE metric "paths/s & Mrays/s per scene at 1/2/4/8 B200": every config at N GPUs (strong scaling: the config's
# total spp split over the ranks, sample sharding + NCCL reduce; C5 in the statistical mode at 4K).  usage: gpu_table.sh N
set -u
N="${1:-1}"
mkdir -p gpurun_out
OUT=gpurun_out/table_n$N.jsonl; : > $OUT
if [ "$N" = "1" ]; then RUN="python"; else RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"; fi
one() { name=$1; shift; timeout 400 $RUN -m cuda_raytracing_b200.render "$@" --out=gpurun_out/t.ppm 2> gpurun_out/t.err > /dev/null; grep -h '^{' gpurun_out/t.err | tail -1 | sed "s/^{/{\"config\": \"$name\", /" >> $OUT; }
one C1 0 s16 b8
one C2 2 s256 b50
one C3 17 s512 b50
one C4 1 s1024 b50
one C5 3 s4096 b50 w3840 t2
rm -f gpurun_out/t.ppm gpurun_out/t.err
cut -c1-260 $OUT
