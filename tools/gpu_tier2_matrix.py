#!/usr/bin/env python3
"""How close can two 4096-spp estimates of a scene be, and is the flattened-instance walk (RT_TRAVERSAL_FAST) one of them?
Renders SCENE at WIDTH with the exact walk (RT_TRAVERSAL_AUTO: the reference's image bit for bit at offset 0) and with
RT_TRAVERSAL_FAST, each from several disjoint XORWOW subsequence blocks, and prints the pairwise PSNR of the 8-bit images,
the PSNR of their 8x8 block means (Monte-Carlo noise averages out, bias does not) and the mean linear radiance.
usage: gpu_tier2_matrix.py SCENE WIDTH [SPP=4096] [RUNS=3]"""
import itertools
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_raytracing_b200 as rt  # noqa: E402


def quant(img):
    g = np.where(img > 0, np.sqrt(np.maximum(img, 0)), 0).astype(np.float32)
    return (np.float32(256) * np.clip(g, 0, np.float32(0.999))).astype(np.uint8)


def psnr(a, b):
    mse = float(((a.astype(np.float64) - b.astype(np.float64)) ** 2).mean())
    return float("inf") if mse == 0 else 10 * np.log10(255.0 ** 2 / mse)


def blocks(a8):
    h, w = (a8.shape[0] // 8) * 8, (a8.shape[1] // 8) * 8
    return a8[:h, :w].astype(np.float64).reshape(h // 8, 8, w // 8, 8, 3).mean(axis=(1, 3))


def main():
    scene, width = int(sys.argv[1]), int(sys.argv[2])
    spp = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
    runs = int(sys.argv[4]) if len(sys.argv) > 4 else 3
    ctx = rt.Context(0)
    hs = rt.HostScene(scene)
    imgs = {}
    for trav, tag in ((rt.RT_TRAVERSAL_AUTO, "exact"), (rt.RT_TRAVERSAL_FAST, "fast")):
        ctx.upload(hs, trav)
        for k in range(runs):
            p = hs.params(width, spp, 50)
            p.subsequence_offset = k * 11 * p.img_width * p.img_height
            img, _, st = ctx.render(p)
            imgs["%s%d" % (tag, k)] = img
            print(json.dumps({"run": "%s%d" % (tag, k), "ms": round(st.render_ms, 1), "mean_radiance": float(img.mean()),
                              "rays_per_path": round(st.rays / st.paths, 4)}), flush=True)
    names = list(imgs)
    q = {n: quant(imgs[n]) for n in names}
    for a, b in itertools.combinations(names, 2):
        print(json.dumps({"pair": [a, b], "psnr8": round(psnr(q[a], q[b]), 2), "psnr8_block_means": round(psnr(blocks(q[a]), blocks(q[b])), 2)}), flush=True)


if __name__ == "__main__":
    main()
