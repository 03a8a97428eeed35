#!/usr/bin/env bash
# exact mode (reference order + treelets) at 4K, tile-sharded over N GPUs vs 1 GPU.  usage: gpu_tiles.sh N [spp]
set -u
N="${1:-2}"; SPP="${2:-32}"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 300 python -m cuda_raytracing_b200.render 3 s$SPP b50 w3840 t0 --out=gpurun_out/c5x1.ppm 2> gpurun_out/tiles_n1.json > /dev/null
timeout 300 $TR -m cuda_raytracing_b200.render 3 s$SPP b50 w3840 t0 --shard=tiles --out=gpurun_out/c5xn.ppm 2> gpurun_out/tiles_n$N.json > /dev/null
cmp gpurun_out/c5x1.ppm gpurun_out/c5xn.ppm && echo "exact mode, tiles x$N image == 1-GPU image (byte for byte)"
rm -f gpurun_out/*.ppm
for f in gpurun_out/tiles_n1.json gpurun_out/tiles_n$N.json; do grep -h '^{' $f | tail -1 | cut -c1-330; done
