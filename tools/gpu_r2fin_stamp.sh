#!/usr/bin/env bash
# Round 2, last GPU call: the device sources' hash moved with a host-only addition (rt_scene_linear_image): one more full capture of the
# bench launch so that profiles/r2_ncu_traffic.json carries the current hash, the quick part of the GPU suite, smoke, one bench line
set -u
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q --deselect "tests/test_gpu_parity.py::test_tier2_converged_images_vs_reference_gpu" ) > gpurun_out/r2stamp_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2stamp_pytest.log
timeout 120 python __graft_entry__.py smoke > gpurun_out/r2stamp_smoke.log 2>&1; echo "smoke rc=$?"
B="python bench.py --steps 2 --warmup 3 --no-configs --no-reference-gpu"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 4 -c 1 -o gpurun_out/r2stamp_bench_c4_render_kernel -f $B > gpurun_out/r2stamp_ncu_full.log 2>&1; echo "full capture rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 --no-configs --no-reference-gpu > gpurun_out/r2stamp_bench.json 2> gpurun_out/r2stamp_bench.err; echo "bench rc=$?"
python -c "import bench; print('kernel_source_sha16', bench.kernel_source_hash())"
