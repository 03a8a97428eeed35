#!/usr/bin/env bash
# Round 2, final evidence with the frozen kernels: the whole GPU suite (tier 2 and the no-allocation entry point included), smoke,
# both bench arms as the driver runs them, the ncu launch list and one full capture of the bench command.
set -u
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q -s ) > gpurun_out/r2fin_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2fin_pytest.log
grep -h "baseline C\|^tier2\|coplanar\|fast-vs-oracle" gpurun_out/r2fin_pytest.log | sed 's/^\.*//' | cut -c1-330
timeout 120 python __graft_entry__.py smoke > gpurun_out/r2fin_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r2fin_smoke.log
timeout 400 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r2fin_bench_reference_arm.json 2> gpurun_out/r2fin_bench_reference_arm.err; echo "reference arm rc=$?"
( time timeout 900 python bench.py --steps 20 --warmup 5 ) > gpurun_out/r2fin_bench_n1.json 2> gpurun_out/r2fin_bench_n1.err; echo "bench rc=$?"; tail -4 gpurun_out/r2fin_bench_n1.err
B="python bench.py --steps 2 --warmup 3 --no-configs --no-reference-gpu"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2fin_launches_bench.csv $B > gpurun_out/r2fin_ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 4 -c 1 -o gpurun_out/r2fin_bench_c4_render_kernel -f $B > gpurun_out/r2fin_ncu_full.log 2>&1; echo "full capture rc=$?"
python -c "import bench; print('kernel_source_sha16', bench.kernel_source_hash())"
cp build/rt_kernels.o gpurun_out/r2fin_rt_kernels.o
