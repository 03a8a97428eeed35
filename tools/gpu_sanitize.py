#!/usr/bin/env python3
"""Tiny renders of one scene per kernel mode (own BVH, stack walk, hybrid walk, linear program, Perlin scene with the
cooperative turbulence) for compute-sanitizer: `compute-sanitizer --tool memcheck|racecheck python tools/gpu_sanitize.py`."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_raytracing_b200 as rt  # noqa: E402

ctx = rt.Context(0)
for sid, trav, spp, b in ((2, rt.RT_TRAVERSAL_AUTO, 3, 12), (1, rt.RT_TRAVERSAL_REFERENCE, 3, 12), (3, rt.RT_TRAVERSAL_AUTO, 2, 12),
                          (1, rt.RT_TRAVERSAL_AUTO, 40, 12), (0, rt.RT_TRAVERSAL_AUTO, 3, 8), (17, rt.RT_TRAVERSAL_AUTO, 20, 12),
                          (19, rt.RT_TRAVERSAL_AUTO, 3, 12)):
    hs = rt.HostScene(sid)
    ctx.upload(hs, trav)
    p = hs.params(64, spp, b)
    img, rgb8, st = ctx.render(p)
    print("scene %d traversal %d: %dx%d s%d, %d rays, mode %d" % (sid, trav, p.img_width, p.img_height, spp, st.rays, hs.pack_info(trav).kernel_mode), flush=True)
    hs.close()
ctx.close()
