#!/usr/bin/env bash
# shared-memory carve-out sweep (env RT_CARVEOUT, percent; unset = what the resident blocks need)
for c in "" 12 25 50 100; do
  echo "== RT_CARVEOUT=${c:-auto}"
  if [ -n "$c" ]; then export RT_CARVEOUT=$c; else unset RT_CARVEOUT; fi
  timeout 300 python tools/gpu_sweep.py C2 C4:256 C3:128 C5s -- 0,0 2>&1 | cut -c1-60,100-175
done
