#!/usr/bin/env python3
"""Kernel time against samples per pixel on one scene: the fixed part of a render (ramp-up + tail) and the marginal cost.
usage: gpu_spp_sweep.py [scene=1] [bounces=50] [spp ...]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_raytracing_b200 as rt  # noqa: E402

scene = int(sys.argv[1]) if len(sys.argv) > 1 else 1
b = int(sys.argv[2]) if len(sys.argv) > 2 else 50
spps = [int(x) for x in sys.argv[3:]] or [1, 2, 4, 8, 16, 32, 64, 128, 256]
hs = rt.HostScene(scene)
ctx = rt.Context(0)
ctx.upload(hs, rt.RT_TRAVERSAL_AUTO)
for spp in spps:
    p = hs.params(0, spp, b)
    t, rays = [], 0
    for _ in range(4):
        _, _, st = ctx.render(p, want_float=False, want_rgb8=True)
        t.append(st.render_ms)
        rays = st.rays
    plan = hs.render_plan(0, spp, b)
    print(json.dumps(dict(scene=scene, spp=spp, b=b, ms=round(min(t), 3), all_ms=[round(x, 3) for x in t], rays=int(rays),
                          grays_s=round(rays / min(t) / 1e6, 3), chunks=int(plan.chunk_count), mode=int(plan.kernel_mode),
                          knobs={k: v for k, v in os.environ.items() if k.startswith("RT_") and k != "RT_B200_LIB"})), flush=True)
