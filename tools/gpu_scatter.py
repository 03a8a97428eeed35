#!/usr/bin/env python3
"""Item order experiment on a B200: every BASELINE config under tile order (RT_SCATTER=0) and scattered order
(RT_SCATTER=1), plus chunk sizes around the automatic one for the Perlin scenes.  Kernel time (min of 3), and
whether the image is bit-identical to the first setting's (it must be: the order never changes a pixel's samples)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_raytracing_b200 as rt  # noqa: E402

CONFIGS = {"C1": (0, 16, 8), "C2": (2, 256, 50), "C3": (17, 512, 50), "C3b": (19, 512, 50), "C4": (1, 1024, 50), "C5s": (3, 16, 50),
           "S18": (18, 256, 50), "S16": (16, 64, 50)}
KNOBS = {
    "C3": [{"RT_SCATTER": "0"}, {"RT_SCATTER": "1"}, {"RT_SCATTER": "1", "RT_CHUNK": "8"}, {"RT_SCATTER": "1", "RT_CHUNK": "64"},
           {"RT_SCATTER": "1", "RT_CHUNK": "512"}, {}],
    "C3b": [{"RT_SCATTER": "0"}, {"RT_SCATTER": "1"}, {"RT_SCATTER": "1", "RT_CHUNK": "8"}, {"RT_SCATTER": "1", "RT_CHUNK": "64"}, {}],
    "C1": [{"RT_SCATTER": "0"}, {"RT_SCATTER": "1"}, {"RT_SCATTER": "1", "RT_CHUNK": "8"}, {"RT_SCATTER": "0", "RT_CHUNK": "8"}, {}],
}
DEFAULT = [{"RT_SCATTER": "0"}, {"RT_SCATTER": "1"}, {}]
NAMES = ("RT_SCATTER", "RT_CHUNK", "RT_CLAIM_GROUP", "RT_TILE_W")


def main():
    names = sys.argv[1:] or ["C3", "C3b", "C1", "C4", "C2", "C5s"]
    for name in names:
        scene, spp, b = CONFIGS[name]
        hs = rt.HostScene(scene)
        first = None
        for kn in KNOBS.get(name, DEFAULT):
            for k in NAMES:
                os.environ.pop(k, None)
            os.environ.update(kn)
            ctx = rt.Context(0)
            ctx.upload(hs, rt.RT_TRAVERSAL_AUTO)
            p = hs.params(0, spp, b)
            times = []
            for _ in range(3):
                img, _, st = ctx.render(p)
                times.append(st.render_ms)
            ctx.close()
            bits = img.view(np.uint32)
            if first is None:
                first = bits.copy()
            print(json.dumps(dict(config=name, scene=scene, spp=spp, b=b, w=p.img_width, h=p.img_height, knobs=kn,
                                  ms=round(min(times), 3), all_ms=[round(t, 3) for t in times],
                                  gpaths_s=round(st.paths / min(times) / 1e6, 3), identical_to_first=bool((bits == first).all()))), flush=True)
        hs.close()


if __name__ == "__main__":
    main()
