#!/usr/bin/env python3
"""Sweep the chunk size (env RT_CHUNK, 0 = automatic) on BASELINE configs; RT_B200_LIB selects a library variant
(tools/build_variants.sh); SWEEP_WIDTH / SWEEP_TRAVERSAL override the image width / traversal.  usage: gpu_sweep.py CONFIG[:spp] ... -- chunk[,claim_group[,tile_w]] ...   (0 = automatic)
Prints one JSON line per run; the image must not depend on the knobs."""
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_raytracing_b200 as rt  # noqa: E402

CONFIGS = {"C1": (0, 16, 8), "C2": (2, 256, 50), "C3": (17, 512, 50), "C3b": (19, 512, 50), "C4": (1, 1024, 50), "C5s": (3, 16, 50)}


def main():
    args = sys.argv[1:]
    cut = args.index("--")
    cfgs, knobs = args[:cut], args[cut + 1:]
    trav = int(os.environ.get("SWEEP_TRAVERSAL", "0"))
    for c in cfgs:
        name, _, spp_s = c.partition(":")
        scene, spp, b = CONFIGS[name]
        if spp_s:
            spp = int(spp_s)
        first = None
        for k in knobs:
            ch, grp, tw = (k.split(",") + ["0", "8"])[:3] if "," in k else (k, "0", "8")
            os.environ["RT_CHUNK"] = ch
            os.environ["RT_CLAIM_GROUP"] = grp
            os.environ["RT_TILE_W"] = tw
            ctx = rt.Context(0)
            hs = rt.HostScene(scene)
            ctx.upload(hs, trav)
            p = hs.params(int(os.environ.get("SWEEP_WIDTH", "0")), spp, b)
            ms = []
            for _ in range(3):
                img, _, st = ctx.render(p, want_rgb8=False)
                ms.append(st.render_ms)
            digest = hashlib.sha1(img.tobytes()).hexdigest()[:12]
            first = first or digest
            print(json.dumps(dict(config=name, spp=spp, chunk=int(ch), group=int(grp), tile_w=int(tw), lib=os.path.basename(os.environ.get("RT_B200_LIB", "default")), ms=round(min(ms), 3),
                                  mpaths_s=round(st.paths / min(ms) / 1e3, 1), mrays_s=round(st.rays / min(ms) / 1e3, 1),
                                  same_image=digest == first, regs=st.regs_per_thread, bps=st.blocks_per_sm)), flush=True)
            hs.close()
            ctx.close()


if __name__ == "__main__":
    main()
