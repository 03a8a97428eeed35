#!/usr/bin/env python3
"""Generates the committed golden vectors from the REAL reference compiled for the host
(oracle/_ref/libref_host.so, built by oracle/build_ref.sh from /root/reference).  Run in the build container:

    python tests/golden/make_golden.py

Outputs (all small): xorwow.json, scenes.json, images.json, functions.npz.  The tests compare the plain-C oracle
(and the C scene builder) against these files, so the pin survives on machines without /root/reference.
"""
import ctypes as C
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import refabi  # noqa: E402

IMG_W, IMG_SPP, IMG_B = 64, 2, 8
FUNC_SCENES = [0, 1, 2, 3, 13, 15, 16, 17, 19]
N_RAYS = 256


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def mask_texture_pointers(tex):
    """Zero the host-pointer fields (ImageTexture.pixels / NoiseTexture.noise at +8) -- they hold process addresses."""
    t = tex.copy()
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


def rays_for(scene, rng, n):
    """Random rays from the scene's camera position towards jittered look directions (+ a few axis-aligned)."""
    lf = np.array(scene.lookfrom[:], np.float32)
    la = np.array(scene.lookat[:], np.float32)
    span = max(1.0, float(np.linalg.norm(la - lf)))
    o = np.tile(lf, (n, 1)).astype(np.float32)
    d = (la - lf)[None, :] + rng.normal(0, 0.6 * span, (n, 3)).astype(np.float32)
    d = d.astype(np.float32)
    d[:4] = np.array([[0, -1, 0], [1, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32)
    t = rng.random(n).astype(np.float32)
    return o, d, t


def main():
    L = refabi.lib()
    # ---- XORWOW known answers
    xs = []
    for seed, sub, off in [(1234, 0, 0), (1234, 1, 0), (1234, 359999, 0), (1234, 809999, 0), (1234, 14745599, 0),
                           (1234, 810000 * 7 + 3, 0), (99, 2 ** 33 + 5, 7), (2 ** 40 + 17, 123456, 2 ** 20 + 3), (0, 2 ** 32 - 1, 1)]:
        u = np.zeros(8, np.float32)
        st = (C.c_uint32 * 6)()
        L.ref_xorwow_uniforms(seed, sub, off, 8, refabi.fptr(u), st)
        xs.append(dict(seed=seed, subsequence=sub, offset=off, uniforms_hex=[int(x) for x in u.view(np.uint32)], state_after=[int(x) for x in st]))
    json.dump(xs, open(os.path.join(HERE, "xorwow.json"), "w"), indent=1)

    scenes, images, funcs = {}, {}, {}
    rng = np.random.default_rng(20261104)
    for sid in range(20):
        s = refabi.build_scene(sid)
        h, w, t = refabi.scene_arrays(s)
        scenes[str(sid)] = dict(
            hittable_count=int(s.hittable_count), world_count=int(s.world_count), root_wref=int(s.root_wref),
            texture_bytes=int(s.texture_bytes), texture_count=int(s.texture_count), max_bvh_depth=int(s.max_bvh_depth),
            hittables_sha256=sha(h), world_sha256=sha(w), textures_sha256=sha(mask_texture_pointers(t)),
            world_head=[int(x) for x in w[:8]], img_width=int(s.img_width), samples_per_pixel=int(s.samples_per_pixel),
            max_bounces=int(s.max_bounces), img_name=s.img_name.decode(),
            settings_hex=[int(x) for x in np.array([s.img_ratio, s.vfov, *s.lookfrom, *s.lookat, *s.view_up, s.defocus_angle, *s.background], np.float32).view(np.uint32)])
        cam = refabi.camera(s, IMG_W, IMG_SPP, IMG_B)
        img, rays = refabi.render_host(s, cam)
        images[str(sid)] = dict(width=cam.img_width, height=cam.img_height, spp=IMG_SPP, bounces=IMG_B, rays=int(rays),
                                image_sha256=sha(img), camera_hex=[int(x) for x in np.frombuffer(bytes(cam), np.uint32)])
        if sid in FUNC_SCENES:
            o, d, tm = rays_for(s, rng, N_RAYS)
            recs = np.zeros((N_RAYS, 24), np.uint32)   # hit, origin'(3), dir'(3), RefHit (16 words) , pad
            for i in range(N_RAYS):
                oo, dd = o[i].copy(), d[i].copy()
                rh = refabi.RefHit()
                hit = L.ref_hit(C.byref(s), refabi.fptr(oo), refabi.fptr(dd), float(tm[i]), 0.001, float("inf"), C.byref(rh))
                recs[i, 0] = hit
                recs[i, 1:4] = oo.view(np.uint32); recs[i, 4:7] = dd.view(np.uint32)
                recs[i, 7:23] = np.frombuffer(bytes(rh), np.uint32)
            funcs["s%d_rays_o" % sid] = o; funcs["s%d_rays_d" % sid] = d; funcs["s%d_rays_t" % sid] = tm
            funcs["s%d_hits" % sid] = recs
            # scatter: every hit record re-scattered with fixed random inputs
            uvec = rng.normal(size=(N_RAYS, 3)).astype(np.float32)
            uvec /= np.linalg.norm(uvec, axis=1, keepdims=True).astype(np.float32)
            rt_ = rng.random(N_RAYS).astype(np.float32)
            sc = np.zeros((N_RAYS, 10), np.uint32)     # scattered, dir(3), att(3), emit(3)
            for i in range(N_RAYS):
                if not recs[i, 0]:
                    continue
                rh = refabi.RefHit.from_buffer_copy(recs[i, 7:23].tobytes())
                dout = np.zeros(3, np.float32); att = np.zeros(3, np.float32); em = np.zeros(3, np.float32)
                din = recs[i, 4:7].view(np.float32).copy()
                ok = L.ref_scatter(C.byref(s), rh.mat, refabi.fptr(din), C.byref(rh), refabi.fptr(uvec[i].copy()), float(rt_[i]),
                                   refabi.fptr(dout), refabi.fptr(att))
                pt = np.array(rh.point[:], np.float32)
                L.ref_emit(C.byref(s), rh.mat, rh.u, rh.v, refabi.fptr(pt), refabi.fptr(em))
                sc[i, 0] = ok; sc[i, 1:4] = dout.view(np.uint32); sc[i, 4:7] = att.view(np.uint32); sc[i, 7:10] = em.view(np.uint32)
            funcs["s%d_scatter_u" % sid] = uvec; funcs["s%d_scatter_r" % sid] = rt_; funcs["s%d_scatter" % sid] = sc
        L.ref_scene_free(C.byref(s))

    # textures / perlin on scene 17 (noise at offset 0: header 32 B, table at 32) and scene 2 (checker at 96)
    s = refabi.build_scene(17)
    pts = (rng.normal(0, 6, (512, 3))).astype(np.float32)
    pts[0] = (1.5, 2.25, -3.75)
    nv = np.array([L.ref_perlin_noise(C.byref(s), 32, refabi.fptr(p.copy())) for p in pts], np.float32)
    tv = np.array([L.ref_perlin_turbulence(C.byref(s), 32, refabi.fptr(p.copy()), 7) for p in pts], np.float32)
    cv = np.zeros((512, 3), np.float32)
    for i, p in enumerate(pts):
        L.ref_texture_color(C.byref(s), 0, 0.3, 0.7, refabi.fptr(p.copy()), refabi.fptr(cv[i]))
    funcs["perlin_points"] = pts; funcs["perlin_noise"] = nv; funcs["perlin_turb"] = tv; funcs["noise_colour"] = cv
    L.ref_scene_free(C.byref(s))
    s = refabi.build_scene(2)
    cc = np.zeros((512, 3), np.float32)
    for i, p in enumerate(pts):
        L.ref_texture_color(C.byref(s), 96, 0.3, 0.7, refabi.fptr(p.copy()), refabi.fptr(cc[i]))
    funcs["checker_colour"] = cc
    L.ref_scene_free(C.byref(s))
    s = refabi.build_scene(16)
    ic = np.zeros((8, 3), np.float32)
    for i in range(8):
        L.ref_texture_color(C.byref(s), 0, float(i) / 7, 1 - float(i) / 7, refabi.fptr(pts[i].copy()), refabi.fptr(ic[i]))
    funcs["image_colour"] = ic
    L.ref_scene_free(C.byref(s))
    # aabb
    boxes = rng.normal(0, 3, (512, 3, 2)).astype(np.float32); boxes.sort(axis=2)
    ao = rng.normal(0, 4, (512, 3)).astype(np.float32); ad = rng.normal(0, 1, (512, 3)).astype(np.float32)
    ad[:16, 0] = 0.0; ad[16:32, 1] = 0.0
    ah = np.array([L.ref_aabb_hit(refabi.fptr(boxes[i].reshape(-1).copy()), refabi.fptr(ao[i].copy()), refabi.fptr(ad[i].copy()), 0.001, 1e30)
                   for i in range(512)], np.int32)
    funcs["aabb_boxes"] = boxes.reshape(512, 6); funcs["aabb_o"] = ao; funcs["aabb_d"] = ad; funcs["aabb_hit"] = ah

    json.dump(scenes, open(os.path.join(HERE, "scenes.json"), "w"), indent=1)
    json.dump(images, open(os.path.join(HERE, "images.json"), "w"), indent=1)
    np.savez_compressed(os.path.join(HERE, "functions.npz"), **funcs)
    print("golden written:", sorted(os.listdir(HERE)))


if __name__ == "__main__":
    main()
