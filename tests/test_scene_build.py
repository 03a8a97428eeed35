"""CPU: the host-side scene construction in C (rt_world.c / rt_scenes.c) against golden digests of the
reference's arrays, plus the known-answer values of SURVEY.md Appendix C.3."""
import ctypes as C

import numpy as np
import pytest

import common


@pytest.mark.parametrize("sid", list(range(20)))
def test_scene_digest(product_lib, sid):
    k = common.golden_json("scenes.json")[str(sid)]
    hs = product_lib.HostScene(sid)
    h, w, t = hs.arrays()
    assert (hs.world.hittable_count, hs.world.world_count, hs.world.root_wref, hs.world.texture_bytes) == \
        (k["hittable_count"], k["world_count"], k["root_wref"], k["texture_bytes"])
    assert common.sha(h) == k["hittables_sha256"]
    assert common.sha(w) == k["world_sha256"]
    assert common.sha(common.mask_texture_pointers(t)) == k["textures_sha256"]
    assert [int(x) for x in w[:8]] == k["world_head"]
    assert product_lib.lib().rt_world_stack_bound(C.byref(hs.world)) == k["max_bvh_depth"]
    s = hs.settings
    assert (s.img_width, s.samples_per_pixel, s.max_bounces, s.img_name.decode()) == \
        (k["img_width"], k["samples_per_pixel"], k["max_bounces"], k["img_name"])
    vals = np.array([s.img_ratio, s.vfov, s.lookfrom.x, s.lookfrom.y, s.lookfrom.z, s.lookat.x, s.lookat.y, s.lookat.z,
                     s.view_up.x, s.view_up.y, s.view_up.z, s.defocus_angle, s.background_color.x, s.background_color.y,
                     s.background_color.z], np.float32)
    assert [int(x) for x in vals.view(np.uint32)] == k["settings_hex"]
    hs.close()


def test_survey_known_answers(product_lib):
    hs = product_lib.HostScene(2)
    h, w, t = hs.arrays()
    assert list(w[:8]) == [0, 94, 95, 7, 28, 29, 73, 117] and hs.world.root_wref == 995
    f = h[1].view(np.float32)
    assert h[1].view(np.uint32)[0] == 0                                    # sphere
    assert np.allclose(f[2:5], [-10.6450558, 0.2, -10.2952108], rtol=0, atol=1e-6)
    assert np.allclose(f[12:16], [0.598775685, 0.95582366, 0.89921999, 0.167611375], rtol=0, atol=1e-7)   # metal albedo, fuzz
    f2 = h[2].view(np.float32)
    assert np.allclose(f2[5:8], [0, 0.317855865, 0], rtol=0, atol=1e-7)    # moving sphere "destination" is a direction
    hs.close()
    hs = product_lib.HostScene(17)
    _, _, t = hs.arrays()
    vec0 = t[32:44].view(np.float32)
    perm = t[32 + 3072:32 + 3072 + 16].view(np.int32)
    assert np.allclose(vec0, [0.622191846, -0.232123852, 0.74766022], rtol=0, atol=1e-7)
    assert list(perm) == [213, 228, 54, 47]
    hs.close()


def test_image_height_rule(product_lib):
    # H = (int)(W / ratio) in f32 (Q17): 800->450, 1200->675, 1080->607, 400->225
    for sid, wh in [(0, (800, 450)), (2, (1200, 675)), (19, (1080, 607)), (15, (400, 225)), (17, (400, 400)), (1, (600, 600))]:
        hs = product_lib.HostScene(sid, build_bvh=False)
        _, cam = hs.camera()
        assert (cam.img_width, cam.img_height) == wh
        hs.close()


def test_unknown_scene(product_lib):
    with pytest.raises(ValueError):
        product_lib.HostScene(20)
    with pytest.raises(ValueError):
        product_lib.HostScene(-1)


def test_image_texture_from_pixels_layout_and_fetch(product_lib, oracle):
    """rt_mat_lambert_image_pixels: the record is the reference's ImageTexture with the payload in the arena, padded by
    the row and texel its inclusive clamp can address (rt.cu:349-352); the oracle's fetch follows that rule exactly."""
    import ctypes as C
    rng = np.random.default_rng(5)
    tex = rng.random((4, 6, 3), dtype=np.float32)
    hs = product_lib.HostScene(None)
    mat = hs.mat_lambert_image_pixels(tex)
    hs.add_sphere((0, 0, -1), 0.5, mat)
    hs.add_sphere((0, -100.5, -1), 100.0, hs.mat_lambert_solid((0.5, 0.5, 0.5)))
    hs.build_bvh()
    h, r, t = hs.arrays()
    off = int(h[0, 48:56].copy().view(np.uint64)[0])
    rec = t[off:off + 32]
    assert int(rec[0:4].copy().view(np.uint32)[0]) == 2                               # RT_TEX_IMAGE
    assert int(rec[8:16].copy().view(np.uint64)[0]) != 0                              # "decoded": pixels != NULL
    px_off = int(rec[16:24].copy().view(np.uint64)[0])
    assert px_off == off + 32
    assert tuple(rec[24:32].copy().view(np.int32)) == (6, 4)
    payload = t[px_off:px_off + ((4 + 1) * 6 + 1) * 12].copy().view(np.float32).reshape(-1, 3)
    assert np.array_equal(payload[:24], tex.reshape(-1, 3)) and not payload[24:].any()
    sc = common.oracle_scene(oracle, hs)
    L = oracle.lib()
    for (u, v) in [(0.0, 1.0), (0.999, 0.001), (1.0, 0.0), (0.5, 0.5), (-3.0, 7.0), (1.0, 1.0), (0.17, 0.83)]:
        got = np.zeros(3, dtype=np.float32)
        L.orc_texture_color(C.byref(sc.c), off, u, v, oracle.fptr(np.zeros(3, dtype=np.float32)), oracle.fptr(got))
        uc, vc = min(max(u, 0.0), 1.0), 1.0 - min(max(v, 0.0), 1.0)
        x, y = min(max(int(np.float32(uc) * np.float32(6)), 0), 6), min(max(int(np.float32(vc) * np.float32(4)), 0), 4)
        assert np.array_equal(got, payload[y * 6 + x]), (u, v)
    # no pixels / empty image: the missing-file record (magenta), like rt_mat_lambert_image
    hs2 = product_lib.HostScene(None)
    m2 = hs2.mat_lambert_image_pixels(np.zeros((0, 0, 3), dtype=np.float32))
    hs2.add_sphere((0, 0, -1), 0.5, m2)
    hs2.build_bvh()
    h2, _, t2 = hs2.arrays()
    off2 = int(h2[0, 48:56].copy().view(np.uint64)[0])
    assert int(t2[off2 + 8:off2 + 16].copy().view(np.uint64)[0]) == 0
    hs.close(); hs2.close()
