#!/bin/bash
# usage: tools/gpu_prof.sh NAME CONFIG SPP [TRAVERSAL]  -- plain run, then one ncu --set full capture of the render kernel
name=$1; cfg=$2; spp=$3; trav=${4:-0}
timeout 200 python tools/profile_target.py $cfg $spp 2 $trav > gpurun_out/plain_$name.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 1 -c 1 -o gpurun_out/prof_$name -f \
  python tools/profile_target.py $cfg $spp 2 $trav > gpurun_out/ncu_$name.log 2>&1
cat gpurun_out/plain_$name.log
