#!/usr/bin/env python3
"""Is a config bound by throughput or by its slowest pixels?  Same scene and spp at growing image widths: paths/s that
rises with the pixel count means the small image cannot fill the GPU (pixels are the unit of parallelism).
usage: gpu_width_scaling.py SCENE SPP W1 W2 ..."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_raytracing_b200 as rt  # noqa: E402

scene, spp = int(sys.argv[1]), int(sys.argv[2])
ctx = rt.Context(0)
hs = rt.HostScene(scene)
ctx.upload(hs, rt.RT_TRAVERSAL_AUTO)
for w in [int(x) for x in sys.argv[3:]]:
    p = hs.params(w, spp, 50)
    best = None
    for _ in range(2):
        _, _, st = ctx.render(p, want_float=False, want_rgb8=True)
        best = st if best is None or st.render_ms < best.render_ms else best
    print(json.dumps({"scene": scene, "spp": spp, "width": p.img_width, "height": p.img_height, "ms": round(best.render_ms, 3),
                      "mpaths_s": round(best.paths / best.render_ms / 1e3, 1), "mrays_s": round(best.rays / best.render_ms / 1e3, 1)}), flush=True)
