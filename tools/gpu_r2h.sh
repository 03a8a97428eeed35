#!/usr/bin/env bash
# Round 2, GPU call H: ncu evidence.  Launch list + one full capture of the bench command (C4), one full capture of the
# REFERENCE's own kernel (cuda_calc) on C2 and C4 at reduced spp, one of ours on the same reduced configs; residency variants on C1.
set -u
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-configs --no-reference-gpu"
timeout 300 $B > gpurun_out/r2h_bench_plain.json 2> gpurun_out/r2h_bench_plain.err; echo "plain bench rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2h_launches_bench.csv $B > gpurun_out/r2h_ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 4 -c 1 -o gpurun_out/r2h_bench_c4_render_kernel -f $B > gpurun_out/r2h_ncu_full.log 2>&1; echo "full capture rc=$?"
REF=oracle/_ref/raytracing_ref
mkdir -p /tmp/refrun && cd /tmp/refrun
timeout 600 ncu --set full --clock-control none -k regex:cuda_calc -c 1 -o $GRAFT_REPO_ROOT/gpurun_out/r2h_reference_cuda_calc_c2_s32 -f $GRAFT_REPO_ROOT/$REF 2 s32 b50 > $GRAFT_REPO_ROOT/gpurun_out/r2h_ncu_ref_c2.log 2>&1; echo "reference C2 capture rc=$?"
timeout 600 ncu --set full --clock-control none -k regex:cuda_calc -c 1 -o $GRAFT_REPO_ROOT/gpurun_out/r2h_reference_cuda_calc_c4_s64 -f $GRAFT_REPO_ROOT/$REF 1 s64 b50 > $GRAFT_REPO_ROOT/gpurun_out/r2h_ncu_ref_c4.log 2>&1; echo "reference C4 capture rc=$?"
cd $GRAFT_REPO_ROOT
timeout 400 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 1 -c 1 -o gpurun_out/r2h_ours_c2_s32 -f python tools/profile_target.py C2 32 2 > gpurun_out/r2h_ncu_ours_c2.log 2>&1; echo "ours C2 capture rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 1 -c 1 -o gpurun_out/r2h_ours_c4_s64 -f python tools/profile_target.py C4 64 2 > gpurun_out/r2h_ncu_ours_c4.log 2>&1; echo "ours C4 capture rc=$?"
K="python tools/gpu_knobs.py"
for v in occ6 occ7; do RT_B200_LIB=build/variants/lib$v.so timeout 200 $K C1 C5s > gpurun_out/r2h_knobs_$v.jsonl 2>&1; cut -c1-260 gpurun_out/r2h_knobs_$v.jsonl; done
ls -la gpurun_out/*.ncu-rep | tail -8
