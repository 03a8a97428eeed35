#!/bin/bash
# call N: the no-allocation entry point (rt_render_device_static): its test alone in a fresh process, then the whole GPU suite
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > gpurun_out/r2n_gpu.txt
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r2n_pytest_alone.log 2>&1
echo "alone exit $?" >> gpurun_out/r2n_pytest_alone.log
tail -5 gpurun_out/r2n_pytest_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2n_pytest_all.log 2>&1
echo "all exit $?" >> gpurun_out/r2n_pytest_all.log
tail -8 gpurun_out/r2n_pytest_all.log
