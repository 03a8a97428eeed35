"""CPU: the plain-C oracle against the committed golden vectors (generated from the real reference's host build
by tests/golden/make_golden.py).  Bit-exact: profile 0 of the oracle IS the reference's host arithmetic."""
import ctypes as C

import numpy as np
import pytest

import common


def _golden_scene(oracle, product_lib, sid):
    hs = product_lib.HostScene(sid)
    return hs, common.oracle_scene(oracle, hs)


def test_xorwow_known_answers(oracle):
    L = oracle.lib()
    for k in common.golden_json("xorwow.json"):
        u = np.zeros(8, np.float32)
        st = (C.c_uint32 * 6)()
        L.orc_xorwow_uniforms(k["seed"], k["subsequence"], k["offset"], 8, oracle.fptr(u), st)
        assert [int(x) for x in u.view(np.uint32)] == k["uniforms_hex"], k
        assert [int(x) for x in st] == k["state_after"], k


def test_xorwow_survey_kat(oracle):
    # curand_init(1234, 0, 0) -> 0.145467639, 0.434898734, 0.870710075
    u = np.zeros(3, np.float32)
    oracle.lib().orc_xorwow_uniforms(1234, 0, 0, 3, oracle.fptr(u), None)
    assert np.allclose(u, [0.145467639, 0.434898734, 0.870710075], rtol=0, atol=1e-9)


@pytest.mark.parametrize("sid", [0, 1, 2, 3, 13, 15, 16, 17, 19])
def test_hit_and_scatter_vectors(oracle, product_lib, sid):
    g = common.golden_npz()
    hs, sc = _golden_scene(oracle, product_lib, sid)
    L = oracle.lib()
    o, d, t = g["s%d_rays_o" % sid], g["s%d_rays_d" % sid], g["s%d_rays_t" % sid]
    want = g["s%d_hits" % sid]
    uvec, rtest, wsc = g["s%d_scatter_u" % sid], g["s%d_scatter_r" % sid], g["s%d_scatter" % sid]
    n_hit = 0
    for i in range(o.shape[0]):
        oo, dd = o[i].copy(), d[i].copy()
        rec = oracle.OrcHit()
        hit = L.orc_hit(C.byref(sc.c), oracle.fptr(oo), oracle.fptr(dd), float(t[i]), 0.001, float("inf"), C.byref(rec))
        assert hit == want[i, 0]
        assert np.array_equal(oo.view(np.uint32), want[i, 1:4]) and np.array_equal(dd.view(np.uint32), want[i, 4:7])
        if not hit:
            continue
        n_hit += 1
        got = np.frombuffer(bytes(rec), np.uint32)
        ref = want[i, 7:23]
        # point, normal, t, front, material (6 words) are defined on a hit; u,v only matter for textures and are
        # compared too (the reference computes them for every accepted sphere candidate)
        assert np.array_equal(got, ref), (sid, i)
        dout = np.zeros(3, np.float32); att = np.zeros(3, np.float32); em = np.zeros(3, np.float32)
        ok = L.orc_scatter(C.byref(sc.c), rec.mat, oracle.fptr(dd), C.byref(rec), oracle.fptr(uvec[i].copy()), float(rtest[i]),
                           oracle.fptr(dout), oracle.fptr(att))
        pt = np.array(rec.point[:], np.float32)
        L.orc_emit(C.byref(sc.c), rec.mat, rec.u, rec.v, oracle.fptr(pt), oracle.fptr(em))
        assert ok == wsc[i, 0]
        assert np.array_equal(dout.view(np.uint32), wsc[i, 1:4]) and np.array_equal(att.view(np.uint32), wsc[i, 4:7])
        assert np.array_equal(em.view(np.uint32), wsc[i, 7:10])
    assert n_hit > 10
    hs.close()


def test_perlin_and_textures(oracle, product_lib):
    g = common.golden_npz()
    L = oracle.lib()
    pts = g["perlin_points"]
    hs, sc = _golden_scene(oracle, product_lib, 17)
    nv = np.array([L.orc_perlin_noise(C.byref(sc.c), 32, oracle.fptr(p.copy())) for p in pts], np.float32)
    tv = np.array([L.orc_perlin_turbulence(C.byref(sc.c), 32, oracle.fptr(p.copy()), 7) for p in pts], np.float32)
    cv = np.zeros((len(pts), 3), np.float32)
    for i, p in enumerate(pts):
        L.orc_texture_color(C.byref(sc.c), 0, 0.3, 0.7, oracle.fptr(p.copy()), oracle.fptr(cv[i]))
    assert np.array_equal(common.bits(nv), common.bits(g["perlin_noise"]))
    assert np.array_equal(common.bits(tv), common.bits(g["perlin_turb"]))
    assert np.array_equal(common.bits(cv), common.bits(g["noise_colour"]))
    # SURVEY.md Appendix C.3 KAT for scenes 17/19 at p = (1.5, 2.25, -3.75), scale 4
    assert abs(nv[0] - 0.291775733) < 1e-7 and abs(tv[0] - 0.383006454) < 1e-7 and abs(cv[0, 0] - 0.992418766) < 1e-7
    hs.close()
    hs, sc = _golden_scene(oracle, product_lib, 2)
    cc = np.zeros((len(pts), 3), np.float32)
    for i, p in enumerate(pts):
        L.orc_texture_color(C.byref(sc.c), 96, 0.3, 0.7, oracle.fptr(p.copy()), oracle.fptr(cc[i]))
    assert np.array_equal(common.bits(cc), common.bits(g["checker_colour"]))
    hs.close()
    hs, sc = _golden_scene(oracle, product_lib, 16)      # image texture whose file is missing: magenta (Q19)
    ic = np.zeros((8, 3), np.float32)
    for i in range(8):
        L.orc_texture_color(C.byref(sc.c), 0, float(i) / 7, 1 - float(i) / 7, oracle.fptr(pts[i].copy()), oracle.fptr(ic[i]))
    assert np.array_equal(ic, g["image_colour"]) and np.array_equal(ic[0], [1, 0, 1])
    hs.close()


def test_aabb_vectors(oracle):
    g = common.golden_npz()
    L = oracle.lib()
    got = np.array([L.orc_aabb_hit(oracle.fptr(g["aabb_boxes"][i].copy()), oracle.fptr(g["aabb_o"][i].copy()),
                                   oracle.fptr(g["aabb_d"][i].copy()), 0.001, 1e30) for i in range(512)], np.int32)
    assert np.array_equal(got, g["aabb_hit"])
    assert 0 < got.sum() < 512


@pytest.mark.parametrize("sid", list(range(20)))
def test_image_digests(oracle, product_lib, sid):
    """Whole-image render of every scene equals the reference host render bit for bit (sha256 of the f32 image)."""
    k = common.golden_json("images.json")[str(sid)]
    hs, sc = _golden_scene(oracle, product_lib, sid)
    cam = oracle.OrcCamera.from_buffer_copy(np.array(k["camera_hex"], np.uint32).tobytes())
    img, rays = oracle.render(sc, cam)
    assert rays == k["rays"]
    assert common.sha(img) == k["image_sha256"]
    hs.close()


def test_resolve_matches_reference_rule(oracle):
    x = np.array([[-1.0, 0.0, 1e-9], [0.25, 0.998, 0.999], [1.0, 7.0, np.nan]], np.float32).reshape(1, 3, 3)
    got = oracle.resolve_rgb8(x).reshape(-1)
    # sqrt gamma, clamp to [0, 0.999], (u8)(256*x)  (rt.cu:1391-1427)
    assert list(got) == [0, 0, 0, 128, 255, 255, 255, 255, 0]


def test_gpu_profile_differs_only_slightly(oracle, product_lib):
    """Profile 1 (the reference GPU kernel's FMA contraction) is a different rounding of the same algorithm."""
    hs, sc = _golden_scene(oracle, product_lib, 0)
    p = hs.params(64, 1, 1)
    cam = common.oracle_camera(oracle, p)
    a, ra = oracle.render(sc, cam, gpu_profile=False)
    b, rb = oracle.render(sc, cam, gpu_profile=True)
    assert ra == rb and np.abs(a - b).max() < 1e-5
    hs.close()
