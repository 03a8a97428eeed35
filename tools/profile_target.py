#!/usr/bin/env python3
"""Short, deterministic target for ncu: a few launches of the render kernel on one config at reduced spp.
usage: profile_target.py [config=C2] [spp=32] [reps=2] [traversal=0]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_raytracing_b200 as rt  # noqa: E402

CFG = {"C1": (0, 8), "C2": (2, 50), "C3": (17, 50), "C3b": (19, 50), "C4": (1, 50), "C5": (3, 50)}
name = sys.argv[1] if len(sys.argv) > 1 else "C2"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 32
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
trav = int(sys.argv[4]) if len(sys.argv) > 4 else 0
scene, b = CFG[name]
ctx = rt.Context(0)
hs = rt.HostScene(scene)
ctx.upload(hs, trav)
p = hs.params(0, spp, b)
for _ in range(reps):
    _, _, st = ctx.render(p, want_float=False, want_rgb8=True)
    print("%s s%d: render %.3f ms, %.1f Mpaths/s, %.1f Mrays/s, regs %d, grid %d x %d" % (
        name, spp, st.render_ms, st.paths / st.render_ms / 1e3, st.rays / st.render_ms / 1e3, st.regs_per_thread, st.grid_blocks, st.block_threads))
