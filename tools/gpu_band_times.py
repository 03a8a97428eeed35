#!/usr/bin/env python3
"""Where is the critical path of a small image?  Renders SCENE at its native size in bands of BAND rows and prints the
time of each band alone (each band is far too small to fill the GPU, so its time is its slowest pixel's time).
usage: gpu_band_times.py SCENE SPP BAND [col0 col1]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_raytracing_b200 as rt  # noqa: E402

scene, spp, band = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
ctx = rt.Context(0)
hs = rt.HostScene(scene)
ctx.upload(hs, rt.RT_TRAVERSAL_AUTO)
p = hs.params(0, spp, 50)
out = []
for r0 in range(0, p.img_height, band):
    p.row_begin, p.row_end = r0, min(p.img_height, r0 + band)
    _, _, st = ctx.render(p, want_float=False, want_rgb8=True)
    _, _, st = ctx.render(p, want_float=False, want_rgb8=True)
    out.append((r0, round(st.render_ms, 3), round(st.rays / max(1, st.paths), 2)))
p.row_begin, p.row_end = 0, 0
_, _, st = ctx.render(p, want_float=False, want_rgb8=True)
print(json.dumps({"scene": scene, "spp": spp, "band": band, "full_ms": round(st.render_ms, 3), "bands(row, ms, rays/path)": out}))
