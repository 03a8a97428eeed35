#!/usr/bin/env bash
# C4 (Cornell box, s1024 b50) sample-sharded at 1 and N GPUs.  usage: gpu_c4_scale.sh N
N="${1:-8}"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29547"
timeout 200 python -m cuda_raytracing_b200.render 1 s1024 b50 --out=gpurun_out/t.ppm 2> gpurun_out/c4_n1.json > /dev/null
timeout 200 $TR -m cuda_raytracing_b200.render 1 s1024 b50 --out=gpurun_out/t.ppm 2> gpurun_out/c4_n$N.json > /dev/null
timeout 200 python -m cuda_raytracing_b200.render 0 s16 b8 --out=gpurun_out/t.ppm 2> gpurun_out/c1_n1.json > /dev/null
rm -f gpurun_out/t.ppm
for f in gpurun_out/c4_n1.json gpurun_out/c4_n$N.json gpurun_out/c1_n1.json; do grep -h '^{' $f | tail -1 | cut -c1-300; done
