"""Shared helpers for the parity tests (test infrastructure)."""
import ctypes as C
import hashlib
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def golden_json(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def golden_npz():
    return np.load(os.path.join(GOLDEN, "functions.npz"))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def mask_texture_pointers(tex):
    """Zero ImageTexture.pixels / NoiseTexture.noise (host addresses in the reference's arena)."""
    t = np.array(tex, dtype=np.uint8, copy=True)
    off = 0
    while off + 32 <= t.shape[0]:
        typ = int(t[off:off + 4].view(np.uint32)[0])
        if typ in (2, 3):
            t[off + 8:off + 16] = 0
        off += 32
        if typ == 3:
            off += 6144
        elif typ == 2:
            w, h = t[off - 8:off].view(np.int32)
            off += 12 * int(w) * int(h)
    return t


def oracle_scene(oracleabi, host_scene):
    """OrcScene over the arrays the PRODUCT's C scene builder made (they are byte-identical to the reference's)."""
    h, w, t = host_scene.arrays()
    return oracleabi.Scene(h, w, t, host_scene.world.root_wref)


def oracle_camera(oracleabi, params):
    cam = oracleabi.OrcCamera()
    for f in ("img_width", "img_height", "samples_per_pixel", "max_bounces", "defocus_angle"):
        setattr(cam, f, getattr(params, f))
    for f in ("background", "pixel_delta_v", "pixel_delta_u", "camera_pos", "defocus_disk_u", "defocus_disk_v", "pixel00"):
        getattr(cam, f)[:] = getattr(params, f)[:]
    cam.pixel_samples_scale = np.float32(1.0) / np.float32(max(params.samples_per_pixel, 1))
    return cam


def scene_rays(settings, rng, n):
    lf = np.array([settings.lookfrom.x, settings.lookfrom.y, settings.lookfrom.z], np.float32)
    la = np.array([settings.lookat.x, settings.lookat.y, settings.lookat.z], np.float32)
    span = max(1.0, float(np.linalg.norm(la - lf)))
    o = np.tile(lf, (n, 1)).astype(np.float32)
    d = ((la - lf)[None, :] + rng.normal(0, 0.6 * span, (n, 3))).astype(np.float32)
    d[:4] = np.array([[0, -1, 0], [1, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32)
    t = rng.random(n).astype(np.float32)
    return o, d, t


def psnr8(a, b):
    mse = float(((a.astype(np.float64) - b.astype(np.float64)) ** 2).mean())
    return float("inf") if mse == 0 else 10 * np.log10(255.0 ** 2 / mse)
