#!/usr/bin/env python3
"""First-light check on a B200: our kernels vs (a) the CPU oracle in the GPU FMA profile, (b) the reference's own
GPU binary (float dump from oracle/_ref/raytracing_ref_ext).  Prints per-scene counts of bit-identical pixels."""
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cuda_raytracing_b200 as rt  # noqa: E402
import oracleabi  # noqa: E402

REF_EXT = os.path.join(ROOT, "oracle", "_ref", "raytracing_ref_ext")


def ref_gpu_image(scene, width, spp, bounces, height):
    with tempfile.TemporaryDirectory() as d:
        dump = os.path.join(d, "img.f32")
        env = dict(os.environ, RT_WIDTH=str(width), RT_DUMP_F32=dump, RT_NO_PPM="1")
        t0 = time.time()
        out = subprocess.run([REF_EXT, str(scene), "s%d" % spp, "b%d" % bounces], cwd=d, env=env, capture_output=True,
                             text=True, timeout=600)
        wall = time.time() - t0
        ms = None
        for line in out.stdout.splitlines():
            if line.startswith("Rendering time:") and line.endswith("ms"):
                ms = float(line.split()[2][:-2])
        img = np.fromfile(dump, dtype=np.float32).reshape(height, width, 3)
        return img, ms, wall


def main():
    scenes = [int(a) for a in sys.argv[1:]] or list(range(20))
    ctx = rt.Context(0)
    results = []
    for sid in scenes:
        hs = rt.HostScene(sid)
        width = 200 if sid != 3 else 120
        spp, b = 4, 12
        for trav in (rt.RT_TRAVERSAL_AUTO, rt.RT_TRAVERSAL_REFERENCE, rt.RT_TRAVERSAL_FAST):
            ctx.upload(hs, trav)
            p = hs.params(width, spp, b)
            img, _, st = ctx.render(p)
            h, w, t = hs.arrays()
            osc = oracleabi.Scene(h, w, t, hs.world.root_wref)
            ocam = oracleabi.OrcCamera()
            for f in ("img_width", "img_height", "samples_per_pixel", "max_bounces", "defocus_angle"):
                setattr(ocam, f, getattr(p, f))
            for f in ("background", "pixel_delta_v", "pixel_delta_u", "camera_pos", "defocus_disk_u", "defocus_disk_v", "pixel00"):
                getattr(ocam, f)[:] = getattr(p, f)[:]
            ocam.pixel_samples_scale = np.float32(1.0) / np.float32(spp)
            if trav == rt.RT_TRAVERSAL_AUTO:
                oimg, orays = oracleabi.render(osc, ocam, gpu_profile=True)
                rimg, rms, rwall = ref_gpu_image(sid, width, spp, b, p.img_height) if os.path.exists(REF_EXT) else (None, None, None)
            n = img.shape[0] * img.shape[1]
            same_o = int((img.view(np.uint32) == oimg.view(np.uint32)).all(axis=2).sum())
            same_r = int((img.view(np.uint32) == rimg.view(np.uint32)).all(axis=2).sum()) if rimg is not None else -1
            oref = int((oimg.view(np.uint32) == rimg.view(np.uint32)).all(axis=2).sum()) if rimg is not None else -1
            rec = dict(scene=sid, traversal=trav, used=st.traversal_used, pixels=n, identical_vs_oracle_gpuprofile=same_o,
                       identical_vs_reference_gpu=same_r, oracle_vs_reference_gpu=oref, rays=st.rays, oracle_rays=orays,
                       render_ms=round(st.render_ms, 3), ref_ms=rms, maxdiff_ref=float(np.abs(img - rimg).max()) if rimg is not None else None)
            print(json.dumps(rec), flush=True)
            results.append(rec)
        hs.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
