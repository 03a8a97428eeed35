#!/usr/bin/env python3
"""Scheduling-knob experiments on a B200.  Each argument is CONFIG[:KNOB=VALUE,...]; the configs are BASELINE's
(C1..C5s) plus C5w (scene 3 at 3840x3840, s16) -- e.g. `gpu_knobs.py C3:RT_SCATTER=0 C3:RT_SCATTER=1,RT_CHUNK=512`.
Prints kernel time (min of 3) and whether the image is bit-identical to the first setting of the same config (it
must be: the knobs decide who renders a pixel's samples and when, never the samples)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_raytracing_b200 as rt  # noqa: E402

CONFIGS = {"C1": (0, 16, 8, 0), "C2": (2, 256, 50, 0), "C3": (17, 512, 50, 0), "C3b": (19, 512, 50, 0), "C4": (1, 1024, 50, 0),
           "C4h": (1, 512, 50, 0), "C4q": (1, 256, 50, 0), "C4s": (1, 128, 50, 0), "C4t": (1, 64, 50, 0),
           "C5s": (3, 16, 50, 0), "C5m": (3, 64, 50, 0), "C5w": (3, 16, 50, 3840)}
NAMES = ("RT_SCATTER", "RT_CHUNK", "RT_CLAIM_GROUP", "RT_TILE_W", "RT_FUSE_MIN", "RT_TREELET_MIN", "RT_CARVEOUT", "RT_LINEAR", "RT_BLOCKS_PER_SM", "RT_FLAT", "RT_AXIS_QUADS", "RT_FUSE_PROGRAM_WRAPPERS", "RT_TAIL_MIN")


def main():
    first = {}
    for arg in sys.argv[1:]:
        name, _, kn = arg.partition(":")
        knobs = dict(kv.split("=") for kv in kn.split(",") if kv)
        flags = int(knobs.pop("FLAGS", "0"))            # pseudo-knob: extra rt_render_params.flags (e.g. 32 = forward throughput)
        scene, spp, b, width = CONFIGS[name]
        for k in NAMES:
            os.environ.pop(k, None)
        os.environ.update(knobs)
        hs = rt.HostScene(scene)
        ctx = rt.Context(0)
        ctx.upload(hs, rt.RT_TRAVERSAL_AUTO)
        p = hs.params(width, spp, b)
        p.flags |= flags
        times = []
        for _ in range(3 if name != "C5w" else 2):
            img, _, st = ctx.render(p)
            times.append(st.render_ms)
        ctx.close()
        hs.close()
        bits = img.view(np.uint32)
        if name not in first:
            first[name] = bits.copy()
        import hashlib
        ref = first[name].view(np.float32)
        print(json.dumps(dict(config=name, scene=scene, spp=spp, b=b, w=p.img_width, h=p.img_height, knobs=knobs, flags=flags,
                              lib=os.path.basename(os.environ.get("RT_B200_LIB", "default")), sha=hashlib.sha256(bits.tobytes()).hexdigest()[:12],
                              max_rel_diff_to_first=float((np.abs(img - ref) / (np.abs(ref) + 1e-6)).max()),
                              ms=round(min(times), 3), all_ms=[round(t, 3) for t in times],
                              gpaths_s=round(st.paths / min(times) / 1e6, 3), regs=st.regs_per_thread, bps=st.blocks_per_sm,
                              identical_to_first=bool((bits == first[name]).all()))), flush=True)


if __name__ == "__main__":
    main()
