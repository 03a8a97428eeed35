#!/usr/bin/env python3
"""BASELINE.json configs C1..C4 (and optionally others) at full size on a B200: the reference's own CUDA binary
vs our kernels -- kernel time, paths/s, Mrays/s, bit-identical pixels (f32), PPM byte equality."""
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_raytracing_b200 as rt  # noqa: E402

REF_EXT = os.path.join(ROOT, "oracle", "_ref", "raytracing_ref_ext")
CONFIGS = {  # name: (scene, spp, bounces, width override)
    "C1": (0, 16, 8, 0), "C2": (2, 256, 50, 0), "C3": (17, 512, 50, 0), "C3b": (19, 512, 50, 0), "C4": (1, 1024, 50, 0),
    "C5s": (3, 16, 50, 0),
}


def run_ref(scene, spp, b, width, height, reps=1):
    best = None
    img = None
    with tempfile.TemporaryDirectory() as d:
        dump = os.path.join(d, "img.f32")
        env = dict(os.environ, RT_DUMP_F32=dump, RT_NO_PPM="1")
        if width:
            env["RT_WIDTH"] = str(width)
        for _ in range(reps):
            out = subprocess.run([REF_EXT, str(scene), "s%d" % spp, "b%d" % b], cwd=d, env=env, capture_output=True,
                                 text=True, timeout=3000)
            for line in out.stdout.splitlines():
                if line.startswith("Rendering time:") and line.endswith("ms"):
                    ms = float(line.split()[2][:-2])
                    best = ms if best is None else min(best, ms)
        img = np.fromfile(dump, dtype=np.float32)
    return img, best


def main():
    names = sys.argv[1:] or ["C1", "C2", "C3", "C4"]
    ctx = rt.Context(0)
    for name in names:
        scene, spp, b, width = CONFIGS[name]
        hs = rt.HostScene(scene)
        for trav in (rt.RT_TRAVERSAL_AUTO, rt.RT_TRAVERSAL_FAST):
            ctx.upload(hs, trav)
            p = hs.params(width, spp, b)
            times = []
            for rep in range(3):
                img, rgb8, st = ctx.render(p)
                times.append(st.render_ms)
            ms = min(times)
            if trav == rt.RT_TRAVERSAL_AUTO:
                rimg, rms = run_ref(scene, spp, b, width, p.img_height, reps=2)
                rimg = rimg.reshape(img.shape)
            n = img.shape[0] * img.shape[1]
            same = int((img.view(np.uint32) == rimg.view(np.uint32)).all(axis=2).sum())
            d8 = np.abs(rgb8.astype(np.int32) - quant(rimg).astype(np.int32))
            mse = float(((rgb8.astype(np.float64) - quant(rimg).astype(np.float64)) ** 2).mean())
            rec = dict(config=name, scene=scene, spp=spp, b=b, w=p.img_width, h=p.img_height, traversal=trav,
                       used=st.traversal_used, ours_ms=round(ms, 3), ref_ms=rms, speedup=round(rms / ms, 2) if rms else None,
                       mpaths_s=round(st.paths / ms / 1e3, 1), mrays_s=round(st.rays / ms / 1e3, 1),
                       rays_per_path=round(st.rays / st.paths, 3), bit_identical=same, pixels=n,
                       frac_within_1_255=float((d8.max(axis=2) <= 1).mean()), psnr8=round(10 * np.log10(255.0 ** 2 / mse), 2) if mse > 0 else None,
                       rng_ms=round(st.rng_init_ms, 3), regs=st.regs_per_thread, grid=st.grid_blocks, bps=st.blocks_per_sm)
            print(json.dumps(rec), flush=True)
        hs.close()


def quant(img):
    g = np.where(img > 0, np.sqrt(np.maximum(img, 0)), 0).astype(np.float32)
    g = np.clip(g, 0, np.float32(0.999))
    return (np.float32(256) * g).astype(np.uint8)


if __name__ == "__main__":
    main()
