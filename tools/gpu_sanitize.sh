#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 150 compute-sanitizer --tool memcheck --print-limit 20 python tools/gpu_sanitize.py > gpurun_out/sanitize_memcheck.log 2>&1; echo "memcheck rc=$?"; tail -4 gpurun_out/sanitize_memcheck.log
timeout 150 compute-sanitizer --tool racecheck --print-limit 20 python tools/gpu_sanitize.py > gpurun_out/sanitize_racecheck.log 2>&1; echo "racecheck rc=$?"; tail -4 gpurun_out/sanitize_racecheck.log
