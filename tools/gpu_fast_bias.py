#!/usr/bin/env python3
"""Where does RT_TRAVERSAL_FAST differ systematically from the exact walk?  Renders SCENE at WIDTH twice per mode (disjoint
XORWOW blocks), averages, and prints the 8x8 blocks whose mean linear radiance differs most, in units of the block's
own Monte-Carlo noise (estimated from the two runs of each mode).  usage: gpu_fast_bias.py SCENE WIDTH SPP"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_raytracing_b200 as rt  # noqa: E402

scene, width, spp = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
ctx = rt.Context(0)
hs = rt.HostScene(scene)
B = 8


def blocks(a):
    h, w = (a.shape[0] // B) * B, (a.shape[1] // B) * B
    return a[:h, :w].astype(np.float64).reshape(h // B, B, w // B, B, 3).mean(axis=(1, 3, 4))


runs = {}
for trav, tag in ((rt.RT_TRAVERSAL_AUTO, "exact"), (rt.RT_TRAVERSAL_FAST, "fast")):
    ctx.upload(hs, trav)
    imgs = []
    for k in range(2):
        p = hs.params(width, spp, 50)
        p.subsequence_offset = (k + 1) * 13 * p.img_width * p.img_height
        img, _, st = ctx.render(p)
        imgs.append(np.minimum(img, 4.0))            # clip fireflies: the question is geometry, not caustic noise
    runs[tag] = [blocks(x) for x in imgs]
e = 0.5 * (runs["exact"][0] + runs["exact"][1]); f = 0.5 * (runs["fast"][0] + runs["fast"][1])
noise = np.sqrt(0.25 * ((runs["exact"][0] - runs["exact"][1]) ** 2 + (runs["fast"][0] - runs["fast"][1]) ** 2) + 1e-12)
z = (f - e) / noise
order = np.dstack(np.unravel_index(np.argsort(-np.abs(f - e), axis=None), e.shape))[0][:25]
print(json.dumps({"scene": scene, "width": width, "spp": spp, "rms_block_diff": float(np.sqrt(((f - e) ** 2).mean())),
                  "rms_block_noise": float(np.sqrt((noise ** 2).mean())), "blocks_over_5_sigma": int((np.abs(z) > 5).sum()), "blocks": int(z.size)}))
for (by, bx) in order:
    print(json.dumps({"block_row": int(by), "block_col": int(bx), "pixel": [int(bx * B), int(by * B)], "exact": round(float(e[by, bx]), 5),
                      "fast": round(float(f[by, bx]), 5), "sigma": round(float(z[by, bx]), 1)}))
