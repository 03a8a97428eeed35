"""GPU (-m gpu): the CUDA path, called through the C ABI, against the CPU oracle on the same seeded inputs.

Bars (written here, per the north star):
  * integer work (XORWOW states): bit-exact.
  * traversal / scatter / textures / whole images against the oracle in its GPU FMA profile: bit-exact, except
    where CUDA's libm differs from glibc (sinf in the noise texture, powf in the dielectric): <= 2e-6 absolute.
  * against the reference's OWN GPU render (oracle/_ref/raytracing_ref_ext, travels as a binary): |delta| <= 1/255 on
    >= 99.9 % of pixels (tier 1) -- in practice the f32 images are bit-identical.
  * RT_TRAVERSAL_FAST on scenes with instance wrappers: statistical parity only (PSNR vs the exact mode).
"""
import ctypes as C
import os
import subprocess
import tempfile

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu

REF_EXT = os.path.join(common.ROOT, "oracle", "_ref", "raytracing_ref_ext")
WRAPPER_SCENES = {0, 1, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 14}
NOISE_SCENES = {3, 17, 19}


def test_xorwow_states_bit_exact(gpu_ctx, oracle):
    L = oracle.lib()
    for seed, first, n in [(1234, 0, 4099), (1234, 809990, 37), (77, 2 ** 33 + 11, 130), (1234, 14745600 - 5, 5)]:
        got = gpu_ctx.probe_xorwow(seed, first, n)
        for i in list(range(0, n, max(1, n // 97))) + [n - 1]:
            st = (C.c_uint32 * 6)()
            L.orc_xorwow_init(seed, first + i, 0, st)
            assert list(got[i]) == list(st), (seed, first, i)


@pytest.mark.parametrize("sid", list(range(20)))
def test_trace_probe_vs_oracle(gpu_ctx, oracle, product_lib, sid):
    hs = product_lib.HostScene(sid)
    sc = common.oracle_scene(oracle, hs)
    L = oracle.lib()
    L.orc_set_fma_profile(1)
    rng = np.random.default_rng(1000 + sid)
    o, d, t = common.scene_rays(hs.settings, rng, 3000)
    try:
        for trav in (product_lib.RT_TRAVERSAL_REFERENCE, product_lib.RT_TRAVERSAL_AUTO):
            gpu_ctx.upload(hs, trav)
            got = gpu_ctx.probe_trace(o, d, t)
            hits = 0
            for i in range(o.shape[0]):
                oo, dd = o[i].copy(), d[i].copy()
                rec = oracle.OrcHit()
                h = L.orc_hit(C.byref(sc.c), oracle.fptr(oo), oracle.fptr(dd), float(t[i]), 0.001, float("inf"), C.byref(rec))
                g = got[i]
                assert g.hit == h, (sid, trav, i)
                # the ray comes back edited by the wrappers exactly as in the reference (Q12)
                assert np.array_equal(common.bits(np.array(g.origin[:])), common.bits(oo))
                assert np.array_equal(common.bits(np.array(g.direction[:])), common.bits(dd))
                if h:
                    hits += 1
                    assert np.array_equal(common.bits(np.array(g.point[:])), common.bits(np.array(rec.point[:])))
                    assert np.array_equal(common.bits(np.array(g.normal[:])), common.bits(np.array(rec.normal[:])))
                    assert common.bits(np.float32(g.t)) == common.bits(np.float32(rec.t))
                    assert g.front == rec.front and g.material_type == rec.mat[0]
            assert hits > 100
    finally:
        L.orc_set_fma_profile(0)
        hs.close()


@pytest.mark.parametrize("sid", [0, 1, 2, 3, 12, 17])
def test_scatter_probe_vs_oracle(gpu_ctx, oracle, product_lib, sid):
    hs = product_lib.HostScene(sid)
    sc = common.oracle_scene(oracle, hs)
    L = oracle.lib()
    rng = np.random.default_rng(2000 + sid)
    o, d, t = common.scene_rays(hs.settings, rng, 2000)
    gpu_ctx.upload(hs, product_lib.RT_TRAVERSAL_REFERENCE)
    got = gpu_ctx.probe_trace(o, d, t)
    idx = [i for i in range(len(got)) if got[i].hit]
    L.orc_set_fma_profile(1)
    try:
        recs, mats = [], []
        for i in idx:
            oo, dd = o[i].copy(), d[i].copy()
            rec = oracle.OrcHit()
            L.orc_hit(C.byref(sc.c), oracle.fptr(oo), oracle.fptr(dd), float(t[i]), 0.001, float("inf"), C.byref(rec))
            recs.append(rec); mats.append(list(rec.mat))
        n = len(idx)
        U = rng.normal(size=(n, 3)).astype(np.float32)
        U /= np.linalg.norm(U, axis=1, keepdims=True).astype(np.float32)
        rt_ = rng.random(n).astype(np.float32)
        din = np.array([got[i].direction[:] for i in idx], np.float32)
        probe_recs = (product_lib.api.ProbeHit * n)(*[got[i] for i in idx])
        for k in range(n):      # texture u,v as the reference records them (the kernel computes them lazily)
            probe_recs[k].u = recs[k].u; probe_recs[k].v = recs[k].v
        do, ao, ok = gpu_ctx.probe_scatter(np.array(mats, np.uint32), din, probe_recs, U, rt_)
        worst = 0.0
        for k in range(n):
            edo = np.zeros(3, np.float32); eat = np.zeros(3, np.float32)
            eok = L.orc_scatter(C.byref(sc.c), recs[k].mat, oracle.fptr(din[k].copy()), C.byref(recs[k]), oracle.fptr(U[k].copy()),
                                float(rt_[k]), oracle.fptr(edo), oracle.fptr(eat))
            mtype = recs[k].mat[0]
            if mtype == 2:      # dielectric: powf (CUDA libm vs glibc) may flip reflect/refract right at the threshold
                if eok == ok[k] and np.abs(edo - do[k]).max() < 1e-5:
                    continue
                worst += 1
                continue
            assert eok == ok[k], (sid, k, mtype)
            assert np.array_equal(common.bits(do[k]), common.bits(edo)), (sid, k, mtype)
            assert np.abs(ao[k] - eat).max() <= 2e-6                      # noise texture: sinf
        assert worst <= max(1, n // 500)
    finally:
        L.orc_set_fma_profile(0)
        hs.close()


def test_texture_probe_vs_oracle(gpu_ctx, oracle, product_lib):
    L = oracle.lib()
    rng = np.random.default_rng(5)
    pts = rng.normal(0, 6, (4096, 3)).astype(np.float32)
    uv = rng.random((4096, 2)).astype(np.float32)
    L.orc_set_fma_profile(1)
    try:
        for sid, off, tol in [(2, 96, 0.0), (2, 0, 0.0), (16, 0, 0.0), (17, 0, 2e-6), (3, 128, 2e-6), (3, 96, 0.0)]:
            hs = product_lib.HostScene(sid)
            sc = common.oracle_scene(oracle, hs)
            gpu_ctx.upload(hs, product_lib.RT_TRAVERSAL_REFERENCE)
            got = gpu_ctx.probe_texture(off, uv, pts)
            want = np.zeros_like(got)
            for i in range(len(pts)):
                L.orc_texture_color(C.byref(sc.c), off, float(uv[i, 0]), float(uv[i, 1]), oracle.fptr(pts[i].copy()), oracle.fptr(want[i]))
            if tol == 0.0:
                assert np.array_equal(common.bits(got), common.bits(want)), (sid, off)
            else:
                assert np.abs(got - want).max() <= tol, (sid, off)
            hs.close()
    finally:
        L.orc_set_fma_profile(0)


def _require_reference_binary():
    """The reference's own GPU binary is the tier-1 pin (the oracle differs from it wherever CUDA's libm differs from
    glibc's): it travels to the box as a built file under oracle/_ref/.  Its absence FAILS the test -- a pin that
    quietly drops out would leave noise / dielectric scenes held to a 97 % bar only."""
    if not os.path.exists(REF_EXT):
        pytest.fail("oracle/_ref/raytracing_ref_ext is missing: run `bash oracle/build_ref.sh` where /root/reference exists "
                    "(the built binary travels with the snapshot)")


def _ref_gpu_image(scene, width, spp, bounces, height, want_ms=False):
    _require_reference_binary()
    with tempfile.TemporaryDirectory() as d:
        dump = os.path.join(d, "img.f32")
        env = dict(os.environ, RT_WIDTH=str(width), RT_DUMP_F32=dump, RT_NO_PPM="1")
        if not width:
            env.pop("RT_WIDTH")
        out = subprocess.run([REF_EXT, str(scene), "s%d" % spp, "b%d" % bounces], cwd=d, env=env, capture_output=True, text=True, timeout=900, check=True)
        img = np.fromfile(dump, dtype=np.float32).reshape(height, -1, 3)
        if want_ms:
            ms = [float(ln.split()[2][:-2]) for ln in out.stdout.splitlines() if ln.startswith("Rendering time:") and ln.endswith("ms")]
            return img, (ms[0] if ms else None)
        return img


@pytest.mark.parametrize("sid", list(range(20)))
def test_image_vs_oracle_and_reference_gpu(gpu_ctx, oracle, product_lib, sid):
    hs = product_lib.HostScene(sid)
    sc = common.oracle_scene(oracle, hs)
    width, spp, b = (160 if sid != 3 else 96), 4, 12
    p = hs.params(width, spp, b)
    cam = common.oracle_camera(oracle, p)
    want, want_rays = oracle.render(sc, cam, gpu_profile=True)
    n = p.img_width * p.img_height
    ref = _ref_gpu_image(sid, width, spp, b, p.img_height)       # mandatory: fails when the binary did not travel
    for trav in (product_lib.RT_TRAVERSAL_AUTO, product_lib.RT_TRAVERSAL_REFERENCE):
        gpu_ctx.upload(hs, trav)
        img, rgb8, st = gpu_ctx.render(p)
        if sid in NOISE_SCENES or sid in (8, 9, 10, 11, 12, 2):      # sinf / powf: libm may differ from glibc
            close = (np.abs(img - want).max(axis=2) <= 2e-6).mean()
            assert close >= 0.97, (sid, trav, close)
        else:
            assert st.rays == want_rays
            assert np.array_equal(common.bits(img), common.bits(want)), (sid, trav)
        assert np.array_equal(rgb8, oracle.resolve_rgb8(img))          # device resolve == the reference's gamma/quantise rule
        # tier 1 against the reference's own GPU render
        same = (common.bits(img) == common.bits(ref)).all(axis=2).mean()
        within = (np.abs(rgb8.astype(np.int32) - oracle.resolve_rgb8(ref).astype(np.int32)).max(axis=2) <= 1).mean()
        assert within >= 0.999, (sid, trav, within)
        assert same >= 0.999, (sid, trav, same)
    if sid in WRAPPER_SCENES:                                            # flattened instances: tier 2 only
        gpu_ctx.upload(hs, product_lib.RT_TRAVERSAL_AUTO)
        exact, e8, _ = gpu_ctx.render(hs.params(width, 64, b))
        gpu_ctx.upload(hs, product_lib.RT_TRAVERSAL_FAST)
        fast, f8, st = gpu_ctx.render(hs.params(width, 64, b))
        assert st.traversal_used == product_lib.RT_TRAVERSAL_FAST
        # two independent 64-spp estimates: compare 8x8 block means (Monte-Carlo noise averages out, bias would not)
        hb, wb = (e8.shape[0] // 8) * 8, (e8.shape[1] // 8) * 8
        eb = e8[:hb, :wb].reshape(hb // 8, 8, wb // 8, 8, 3).mean(axis=(1, 3))
        fb = f8[:hb, :wb].reshape(hb // 8, 8, wb // 8, 8, 3).mean(axis=(1, 3))
        assert common.psnr8(eb, fb) >= 32.0, (sid, common.psnr8(eb, fb))
        assert abs(float(exact.mean()) - float(fast.mean())) <= 0.01 * max(float(exact.mean()), 1e-3) + 1e-3
    hs.close()


BASELINE_CONFIGS = {   # BASELINE.json configs at their full sizes: name -> (scene, spp, bounces, width override, pixels)
    "C1": (0, 16, 8, 0, 800 * 450), "C2": (2, 256, 50, 0, 1200 * 675), "C3": (17, 512, 50, 0, 400 * 400),
    "C3b": (19, 512, 50, 0, 1080 * 607), "C4": (1, 1024, 50, 0, 600 * 600), "C5r": (3, 16, 50, 0, 800 * 800),
}


@pytest.mark.parametrize("name", list(BASELINE_CONFIGS))
def test_baseline_configs_vs_reference_gpu(gpu_ctx, oracle, product_lib, name):
    """Tier 1 at the FULL size of every BASELINE.json config (C5 at its built-in 800x800 and 1/256 of its samples: the
    reference needs 17.7 GB of colour stacks and hours at 4K s4096): the reference's own CUDA binary, run here on the
    same GPU with its seed 1234, against RT_TRAVERSAL_AUTO.  Bar (north_star): |delta| <= 1/255 on >= 99.9 % of the 8-bit
    pixels; in practice >= 99.9 % of the f32 pixels are bit-identical, which is asserted too."""
    sid, spp, b, width, pixels = BASELINE_CONFIGS[name]
    hs = product_lib.HostScene(sid)
    p = hs.params(width, spp, b)
    assert p.img_width * p.img_height == pixels
    ref, ref_ms = _ref_gpu_image(sid, width, spp, b, p.img_height, want_ms=True)
    assert ref.shape == (p.img_height, p.img_width, 3)
    gpu_ctx.upload(hs, product_lib.RT_TRAVERSAL_AUTO)
    img, rgb8, st = gpu_ctx.render(p)
    same = int((common.bits(img) == common.bits(ref)).all(axis=2).sum())
    d8 = np.abs(rgb8.astype(np.int32) - oracle.resolve_rgb8(ref).astype(np.int32)).max(axis=2)
    print("baseline %s: scene %d %dx%d s%d b%d: %d / %d f32 pixels bit-identical, %d 8-bit pixels differ by more than 1; ours %.2f ms "
          "(kernel mode %d), reference %.2f ms" % (name, sid, p.img_width, p.img_height, spp, b, same, pixels, int((d8 > 1).sum()),
                                                  st.render_ms, hs.pack_info().kernel_mode, ref_ms or -1))
    assert st.paths == pixels * spp
    assert same >= 0.999 * pixels, (name, same, pixels)
    assert (d8 <= 1).mean() >= 0.999, name
    assert np.array_equal(rgb8, oracle.resolve_rgb8(img))
    hs.close()


def test_full_size_properties_c2(gpu_ctx, oracle, product_lib):
    """BASELINE config C2 geometry at full resolution (1200x675, b50), reduced spp: size-independent properties."""
    hs = product_lib.HostScene(2)
    gpu_ctx.upload(hs, product_lib.RT_TRAVERSAL_AUTO)
    p = hs.params(0, 8, 50)
    a, a8, st = gpu_ctx.render(p)
    assert (p.img_width, p.img_height) == (1200, 675) and st.paths == 1200 * 675 * 8
    assert 2.6 < st.rays / st.paths < 3.2                               # 2.90 rays/path
    b, b8, _ = gpu_ctx.render(p)
    assert np.array_equal(common.bits(a), common.bits(b))                # idempotent / deterministic
    # tile sharding: rows rendered separately reproduce the whole image bit for bit (per-pixel streams)
    parts = np.zeros_like(a)
    for r0, r1 in [(0, 200), (200, 201), (201, 675)]:
        q = hs.params(0, 8, 50); q.row_begin, q.row_end = r0, r1
        part, _, _ = gpu_ctx.render(q)
        parts[r0:r1] = part[r0:r1]
        assert not part[:r0].any() and not part[r1:].any()
    assert np.array_equal(common.bits(parts), common.bits(a))
    # sum mode: the mean image is the sum image times 1/spp, rounded once
    q = hs.params(0, 8, 50, flags=product_lib.RT_RENDER_OUTPUT_SUM | product_lib.RT_RENDER_COUNT_RAYS)
    s, _, _ = gpu_ctx.render(q)
    assert np.array_equal(common.bits(s * np.float32(1.0 / 8)), common.bits(a))
    # sample sharding: disjoint subsequences give a statistically equal, different image
    q = hs.params(0, 8, 50, subsequence_offset=1200 * 675)
    c, c8, _ = gpu_ctx.render(q)
    assert not np.array_equal(a, c) and abs(float(a.mean()) - float(c.mean())) < 0.01
    # oracle spot check on a band of full-resolution rows
    cam = common.oracle_camera(oracle, p)
    want, _ = oracle.render(common.oracle_scene(oracle, hs), cam, row0=300, row1=304, gpu_profile=True)
    close = (np.abs(a[300:304] - want[300:304]).max(axis=2) <= 2e-6).mean()
    assert close >= 0.97
    hs.close()


def test_edge_cases(gpu_ctx, product_lib):
    hs = product_lib.HostScene(13)
    gpu_ctx.upload(hs)
    p = hs.params(33, 1, 1); p.max_bounces = 0  # b0: every path is black (Q14), but the camera draws still happen
    img, _, st = gpu_ctx.render(p)
    assert not img.any() and st.rays == 0
    p = hs.params(2, 3, 5)                      # 2x1 image (H = (int)(W / ratio), Q17)
    img, _, _ = gpu_ctx.render(p)
    assert img.shape[:2] == (1, 2) and np.isfinite(img).all()
    with pytest.raises(product_lib.RtError):    # W=1 gives H=0: rejected, not rendered
        gpu_ctx.render(hs.params(1, 3, 5))
    a, _, _ = gpu_ctx.render(hs.params(40, 2, 300))    # b > 255 wraps like the reference's u8 counter: b300 is b44
    w, _, _ = gpu_ctx.render(hs.params(40, 2, 44))
    assert np.array_equal(common.bits(a), common.bits(w))
    z, _, st = gpu_ctx.render(hs.params(40, 2, 256))   # ... and b256 is b0: black
    assert not z.any() and st.rays == 0
    q = hs.params(40, 2, 4, flags=product_lib.RT_RENDER_OUTPUT_SUM | product_lib.RT_RENDER_RESUME)
    with pytest.raises(product_lib.RtError):    # the host-buffer entry point cannot carry sums on: refused, not wrong
        gpu_ctx.render(q)
    with pytest.raises(product_lib.RtError):
        product_lib.Context(0).render(hs.params(8, 1, 1))    # no scene uploaded
    hs.close()


def test_cli_writes_reference_ppm(product_lib, oracle):
    exe = os.path.join(common.ROOT, "bin", "raytracing")
    with tempfile.TemporaryDirectory() as d:
        r = subprocess.run([exe, "13", "s2", "b4"], cwd=d, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr
        lines = r.stdout.splitlines()
        assert lines[0] == "cuda_calc launch error: no error" and lines[1] == "cuda_calc sync error: no error"
        assert lines[2].startswith("Rendering time: ") and lines[2].endswith("ms") and lines[3].endswith("s")
        data = open(os.path.join(d, "mblur.ppm"), "rb").read()
        assert data.startswith(b"P6\n800 450\n255\n") and len(data) == len(b"P6\n800 450\n255\n") + 800 * 450 * 3
        _require_reference_binary()
        if True:
            d2 = os.path.join(d, "ref"); os.mkdir(d2)
            subprocess.run([os.path.join(common.ROOT, "oracle", "_ref", "raytracing_ref"), "13", "s2", "b4"], cwd=d2, capture_output=True, timeout=600, check=True)
            assert open(os.path.join(d2, "mblur.ppm"), "rb").read() == data      # byte-identical file


TIER2_RECORD = os.path.join(common.ROOT, "gpurun_out", "tier2_psnr.jsonl")


@pytest.mark.gpu
@pytest.mark.parametrize("sid,width,abs_floor_db", [(0, 128, 40.0), (1, 128, None), (3, 200, None)])
def test_tier2_converged_images_vs_reference_gpu(gpu_ctx, oracle, product_lib, sid, width, abs_floor_db):
    """Tier 2 (north_star): at 4096 spp the statistical modes -- flattened instances (RT_TRAVERSAL_FAST) and sample
    sharding (2 virtual ranks x 2048 spp, disjoint XORWOW subsequences, summed) -- match the reference's own GPU render.
    PSNR on the 8-bit images.  What two 4096-spp estimates CAN agree to depends on the scene: the sky-lit scene 0 passes
    north_star's 40 dB outright; the Cornell box and the final scene (scene 3, the one C5 renders in statistical mode at 4K)
    are lit by one small quad on a black background and are noise-limited far below that at 4096 spp.  Their floor is
    therefore MEASURED in the same test: an independent run of the EXACT estimator (RT_TRAVERSAL_AUTO, the reference's
    arithmetic and visiting order, another XORWOW subsequence block) against the same reference image.  A statistical
    mode passes when it is within 1 dB of that -- i.e. it is as close to the reference as the reference's own algorithm
    is to itself with other random numbers.  The PSNRs are printed and written to gpurun_out/tier2_psnr.jsonl."""
    import json
    hs = product_lib.HostScene(sid)
    spp, b = 4096, 50
    p = hs.params(width, spp, b)
    n = p.img_width * p.img_height
    ref8 = oracle.resolve_rgb8(_ref_gpu_image(sid, width, spp, b, p.img_height))
    gpu_ctx.upload(hs, product_lib.RT_TRAVERSAL_AUTO)
    same8 = gpu_ctx.render(p)[1]                                         # tier 1 at 4096 spp, for the record
    _, indep8, _ = gpu_ctx.render(hs.params(width, spp, b, subsequence_offset=7 * n))
    db_floor = common.psnr8(indep8, ref8)
    gpu_ctx.upload(hs, product_lib.RT_TRAVERSAL_FAST)
    _, fast8, st = gpu_ctx.render(p)
    assert st.traversal_used == product_lib.RT_TRAVERSAL_FAST
    db_fast = common.psnr8(fast8, ref8)
    total = np.zeros((p.img_height, p.img_width, 3), np.float32)
    for g in range(2):                                                   # the reduce of bench.py / render.py, done on the host
        q = hs.params(width, spp // 2, b, subsequence_offset=g * n, flags=product_lib.RT_RENDER_OUTPUT_SUM)
        part, _, _ = gpu_ctx.render(q)
        total += part
    shard8 = oracle.resolve_rgb8(total * np.float32(1.0 / spp))
    db_shard = common.psnr8(shard8, ref8)
    rec = {"scene": sid, "width": p.img_width, "height": p.img_height, "spp": spp, "bounces": b,
           "psnr_exact_same_stream_db": common.psnr8(same8, ref8), "psnr_exact_independent_stream_db": db_floor,
           "psnr_fast_db": db_fast, "psnr_sample_sharded_db": db_shard}
    print("tier2 " + json.dumps(rec))
    try:
        os.makedirs(os.path.dirname(TIER2_RECORD), exist_ok=True)
        with open(TIER2_RECORD, "a") as f:
            f.write(json.dumps(rec) + "\n")
    except OSError:
        pass
    floor = db_floor - 1.0 if abs_floor_db is None else abs_floor_db
    assert db_fast >= floor and db_shard >= floor, rec
    hs.close()


@pytest.mark.gpu
@pytest.mark.parametrize("sid", [2, 17, 0, 1])
def test_image_does_not_depend_on_scheduling_knobs(product_lib, sid):
    """Chunk size, tile shape, claim grouping, item order (tile order / scattered) and which of the exact walks a scene
    gets only change WHO renders a pixel's samples, when and through which code, never the samples: the image is
    bit-identical whatever the knobs (contexts read them from the environment at creation)."""
    hs = product_lib.HostScene(sid)
    p = hs.params(200, 24, 12)
    images = []
    knobs = [{}, {"RT_CHUNK": "5"}, {"RT_CHUNK": "1000000"}, {"RT_CHUNK": "7", "RT_TILE_W": "32", "RT_CLAIM_GROUP": "1"},
             {"RT_CHUNK": "3", "RT_TILE_W": "4", "RT_CLAIM_GROUP": "32"}, {"RT_SCATTER": "0"}, {"RT_SCATTER": "1"},
             {"RT_SCATTER": "1", "RT_CHUNK": "4", "RT_TILE_W": "2", "RT_CLAIM_GROUP": "32"},
             {"RT_FLAT": "0"}, {"RT_LINEAR": "0"},                 # scene 17 through its BVH, scenes 0 / 1 through the stack walk
             {"RT_AXIS_QUADS": "0"}, {"RT_FUSE_PROGRAM_WRAPPERS": "0"},   # scene 1: general inside test; one op per wrapper side
             {"RT_AXIS_QUADS": "0", "RT_FUSE_PROGRAM_WRAPPERS": "0", "RT_CHUNK": "6"}]
    saved = {k: os.environ.get(k) for k in ("RT_CHUNK", "RT_TILE_W", "RT_CLAIM_GROUP", "RT_SCATTER", "RT_FLAT", "RT_LINEAR", "RT_AXIS_QUADS",
                                            "RT_FUSE_PROGRAM_WRAPPERS")}
    try:
        for kn in knobs:
            for k in saved:
                os.environ.pop(k, None)
            os.environ.update(kn)
            ctx = product_lib.Context(0)
            ctx.upload(hs, product_lib.RT_TRAVERSAL_AUTO)
            img, _, _ = ctx.render(p)
            images.append(img)
            ctx.close()
    finally:
        for k, v in saved.items():
            os.environ.pop(k, None)
            if v is not None:
                os.environ[k] = v
    for img in images[1:]:
        assert np.array_equal(common.bits(img), common.bits(images[0]))
    hs.close()


@pytest.mark.gpu
@pytest.mark.parametrize("sid", [1, 0, 2, 3, 17])
def test_instrumented_kernel_renders_the_same_image(gpu_ctx, product_lib, sid):
    """RT_RENDER_COUNT_OPS selects the instrumented build of the same walk (it feeds bench.py's roofline): same image, same
    ray count as the timed kernel -- on scene 1 that is the build with the reduced inside test of axis-aligned quads."""
    hs = product_lib.HostScene(sid)
    gpu_ctx.upload(hs, product_lib.RT_TRAVERSAL_AUTO)
    plain, _, st0 = gpu_ctx.render(hs.params(120, 12, 20, flags=product_lib.RT_RENDER_COUNT_RAYS))
    counted, _, st1 = gpu_ctx.render(hs.params(120, 12, 20, flags=product_lib.RT_RENDER_COUNT_OPS))
    assert np.array_equal(common.bits(counted), common.bits(plain))
    assert st0.rays == st1.rays > 0
    assert sum(st1.ops) > 0
    hs.close()


@pytest.mark.gpu
@pytest.mark.parametrize("sid", sorted(WRAPPER_SCENES))
def test_hybrid_walk_equals_reference_order(product_lib, sid):
    """RT_TRAVERSAL_AUTO on scenes with instance wrappers: treelets (own BVH inside wrapper-free subtrees) and fused
    instances, forced on even where the packer would not bother, and the linear program of the small scenes give the
    image of the literal reference-order walk bit for bit -- and probes return the same hits and the same edited rays."""
    hs = product_lib.HostScene(sid)
    p = hs.params(96, 6, 12)
    rng = np.random.default_rng(4242 + sid)
    o, d, t = common.scene_rays(hs.settings, rng, 4000)
    saved = {k: os.environ.get(k) for k in ("RT_FUSE_MIN", "RT_TREELET_MIN", "RT_LINEAR")}
    results = []
    try:
        # the literal stack walk; the hybrid walk forced two ways; the automatic choice (small scenes: the stackless linear
        # program, kernel mode 3); the automatic choice without the linear program
        for knobs, trav in (({}, product_lib.RT_TRAVERSAL_REFERENCE), ({"RT_FUSE_MIN": "1", "RT_TREELET_MIN": "2"}, product_lib.RT_TRAVERSAL_AUTO),
                            ({"RT_FUSE_MIN": "0", "RT_TREELET_MIN": "3"}, product_lib.RT_TRAVERSAL_AUTO),
                            ({}, product_lib.RT_TRAVERSAL_AUTO), ({"RT_LINEAR": "0"}, product_lib.RT_TRAVERSAL_AUTO)):
            for k in saved:
                os.environ.pop(k, None)
            os.environ.update(knobs)
            ctx = product_lib.Context(0)
            ctx.upload(hs, trav)
            img, _, st = ctx.render(p)
            assert st.traversal_used == product_lib.RT_TRAVERSAL_REFERENCE
            probes = ctx.probe_trace(o, d, t)
            results.append((img, probes))
            ctx.close()
    finally:
        for k, v in saved.items():
            os.environ.pop(k, None)
            if v is not None:
                os.environ[k] = v
    img0, pr0 = results[0]
    for img, pr in results[1:]:
        assert np.array_equal(common.bits(img), common.bits(img0)), sid
        for i in range(0, len(pr0), 7):
            a, b = pr0[i], pr[i]
            assert a.hit == b.hit, (sid, i)
            assert list(a.origin[:]) == list(b.origin[:]) and list(a.direction[:]) == list(b.direction[:]), (sid, i)
            if a.hit:
                assert a.t == b.t and list(a.point[:]) == list(b.point[:]) and list(a.normal[:]) == list(b.normal[:]), (sid, i)
    hs.close()


@pytest.mark.gpu
@pytest.mark.parametrize("sid,world,band", [(2, 3, 8), (1, 4, 5), (3, 2, 16)])
def test_interleaved_row_bands_assemble_the_full_image(gpu_ctx, product_lib, sid, world, band):
    """Tile sharding in one launch (rt_render_params.band_*): the `world` banded renders touch disjoint rows, leave
    the others alone, and together are the full render bit for bit (what the NCCL reduce over zero-initialised images
    assembles).  Height is not a multiple of the band on purpose."""
    hs = product_lib.HostScene(sid)
    gpu_ctx.upload(hs, product_lib.RT_TRAVERSAL_AUTO)
    p = hs.params(90, 5, 10)
    full, _, _ = gpu_ctx.render(p)
    total = np.zeros_like(full)
    paths = 0
    for rank in range(world):
        q = hs.params(90, 5, 10)
        q.band_rows, q.band_period, q.band_phase = band, world, rank
        img, _, st = gpu_ctx.render(q)
        rows = [r for r in range(full.shape[0]) if (r // band) % world == rank]
        others = [r for r in range(full.shape[0]) if (r // band) % world != rank]
        assert np.array_equal(common.bits(img[rows]), common.bits(full[rows])), (sid, rank)
        assert not img[others].any()
        assert st.paths == len(rows) * full.shape[1] * 5
        total += img
        paths += st.paths
    assert np.array_equal(common.bits(total), common.bits(full))
    assert paths == full.shape[0] * full.shape[1] * 5
    q = hs.params(90, 5, 10)
    q.band_rows, q.band_period, q.band_phase = band, world, world              # phase out of range
    with pytest.raises(product_lib.RtError):
        gpu_ctx.render(q)
    hs.close()


@pytest.mark.gpu
@pytest.mark.parametrize("sid,bands", [(2, False), (3, False), (1, True)])
def test_progressive_accumulation_equals_one_render(product_lib, sid, bands):
    """7 samples, then 9 more (RT_RENDER_RESUME), then 8 more after the state went through the host
    into a fresh context, is the 24-sample render bit for bit -- same XORWOW stream, same order of additions."""
    import torch
    hs = product_lib.HostScene(sid)
    SUM = product_lib.RT_RENDER_OUTPUT_SUM

    def params(spp, extra=0):
        q = hs.params(80, spp, 10, flags=SUM | extra)
        if bands:
            q.band_rows, q.band_period, q.band_phase = 6, 2, 1
        return q

    ctx = product_lib.Context(0)
    ctx.upload(hs, product_lib.RT_TRAVERSAL_AUTO)
    p24 = params(24)
    n = p24.img_width * p24.img_height
    whole = torch.zeros(n * 3, dtype=torch.float32, device="cuda:0")
    ctx.render_device(p24, whole.data_ptr())
    part = torch.zeros(n * 3, dtype=torch.float32, device="cuda:0")
    ctx.render_device(params(7), part.data_ptr())
    with pytest.raises(product_lib.RtError):                      # resuming needs the sums
        ctx.render_device(hs.params(80, 9, 10, flags=product_lib.RT_RENDER_RESUME), part.data_ptr())
    st = ctx.render_device(params(9, product_lib.RT_RENDER_RESUME), part.data_ptr())
    assert st.kernel_launches == 1                                # no XORWOW init, just the render
    state = ctx.save_state()
    assert state.shape == (st.paths // 9, 6)
    sums = part.cpu()
    ctx.close()

    ctx2 = product_lib.Context(0)
    ctx2.upload(hs, product_lib.RT_TRAVERSAL_AUTO)
    with pytest.raises(product_lib.RtError):                      # nothing to resume yet in this context
        ctx2.render_device(params(8, product_lib.RT_RENDER_RESUME), part.data_ptr())
    ctx2.load_state(params(8), state)
    again = sums.to("cuda:0")
    ctx2.render_device(params(8, product_lib.RT_RENDER_RESUME), again.data_ptr())
    assert np.array_equal(common.bits(again.cpu().numpy()), common.bits(whole.cpu().numpy()))
    with pytest.raises(product_lib.RtError):                      # a state of the wrong size is refused
        ctx2.load_state(params(8), state[:-1])
    ctx2.close()
    hs.close()


@pytest.mark.gpu
def test_image_texture_from_decoded_pixels(gpu_ctx, oracle, product_lib):
    """an image texture whose pixels the caller decoded (rt_mat_lambert_image_pixels) is fetched on
    the device with the reference's clamp semantics (inclusive upper clamp, Q19) -- probes against the oracle's
    restatement of image_texture_get_pixel (rt.cu:343-354), and a render against the oracle's render, bit for bit."""
    rng = np.random.default_rng(77)
    tex = rng.random((5, 7, 3), dtype=np.float32)
    hs = product_lib.HostScene(None)
    mat = hs.mat_lambert_image_pixels(tex)
    hs.add_sphere((0, 0, -1), 0.5, mat)
    hs.add_sphere((0, -100.5, -1), 100.0, hs.mat_lambert_solid((0.8, 0.8, 0.0)))
    hs.build_bvh()
    hs.settings.img_width = 64; hs.settings.samples_per_pixel = 4; hs.settings.max_bounces = 6
    hs.settings.lookfrom = product_lib.api.Vec3(0, 0, 1); hs.settings.lookat = product_lib.api.Vec3(0, 0, -1)
    h, r, t = hs.arrays()
    off = int(h[0, 48:56].copy().view(np.uint64)[0])                     # sphere 0: material at +40, its texture offset at +48
    sc = common.oracle_scene(oracle, hs)
    L = oracle.lib()
    L.orc_set_fma_profile(1)
    try:
        gpu_ctx.upload(hs, product_lib.RT_TRAVERSAL_AUTO)
        uv = rng.random((512, 2), dtype=np.float32)
        uv[:8] = [[0, 0], [1, 1], [1, 0], [0, 1], [0.999999, 0.5], [0.5, 1e-7], [1.5, -0.5], [-2, 3]]   # edges and beyond
        pts = np.zeros((512, 3), dtype=np.float32)
        got = gpu_ctx.probe_texture(off, uv, pts)
        want = np.zeros_like(got)
        for i in range(len(uv)):
            L.orc_texture_color(C.byref(sc.c), off, float(uv[i, 0]), float(uv[i, 1]), oracle.fptr(pts[i].copy()), oracle.fptr(want[i]))
        assert np.array_equal(common.bits(got), common.bits(want))
        assert got[8:].min() >= 0.0 and np.unique(got[8:], axis=0).shape[0] > 20      # really the image, not one colour
        p = hs.params(0, 0, 0)
        img, _, _ = gpu_ctx.render(p)
        cam = common.oracle_camera(oracle, p)
        ref, _ = oracle.render(sc, cam, gpu_profile=True)
        L.orc_set_fma_profile(1)
        # the sphere's (u, v) come from acosf / atan2f (CUDA libm vs glibc: a texel can flip at a boundary), nothing else differs
        same = (common.bits(img) == common.bits(ref)).all(axis=2).mean()
        assert same >= 0.99, same
    finally:
        L.orc_set_fma_profile(0)
    hs.close()


@pytest.mark.gpu
@pytest.mark.parametrize("sid", [2, 1, 17, 3])
def test_forward_throughput_mode_is_within_tolerance(gpu_ctx, oracle, product_lib, sid):
    """RT_RENDER_FORWARD_THROUGHPUT carries the product of a path's attenuations forwards in registers instead of unwinding
    the reference's stack backwards: the same paths (same rays, same RNG draws), the same factors, another association of
    the product.  Stated tolerance: relative 1e-5 per pixel on the f32 image (n * 2^-24 per sample with n <= 50 bounces),
    8-bit image within 1 everywhere -- far inside tier 1's 1/255, but not bit-identical, hence a flag and not the default."""
    hs = product_lib.HostScene(sid)
    gpu_ctx.upload(hs, product_lib.RT_TRAVERSAL_AUTO)
    p = hs.params(160 if sid != 3 else 96, 16, 50)
    exact, e8, st0 = gpu_ctx.render(p)
    q = hs.params(160 if sid != 3 else 96, 16, 50, flags=product_lib.RT_RENDER_COUNT_RAYS | product_lib.RT_RENDER_FORWARD_THROUGHPUT)
    fwd, f8, st1 = gpu_ctx.render(q)
    assert st0.rays == st1.rays                                        # the paths are the same paths
    assert np.all(np.abs(fwd - exact) <= 1e-5 * np.abs(exact) + 1e-9), float(np.abs(fwd - exact).max())
    assert np.abs(f8.astype(np.int32) - e8.astype(np.int32)).max() <= 1
    hs.close()


def _gpu_count():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.gpu
def test_group_of_one_device_is_the_single_gpu_path(product_lib, oracle):
    """rt_group with one device (no NCCL involved): rt_group_render_world == rt_render_world, bit for bit, for both
    shardings (with one device there is nothing to deal out)."""
    hs = product_lib.HostScene(1)
    s, _ = hs.camera(96, 8, 12)
    ctx = product_lib.Context(0)
    want, want8, st0 = ctx.render_world(hs, s)
    ctx.close()
    grp = product_lib.Group(1)
    for shard in (product_lib.RT_SHARD_TILES, product_lib.RT_SHARD_SAMPLES):
        got, got8, st = grp.render_world(hs, s, shard=shard)
        assert np.array_equal(common.bits(got), common.bits(want)) and np.array_equal(got8, want8)
        assert st.paths == st0.paths and st.rays == st0.rays
    grp.close()
    hs.close()


@pytest.mark.gpu
@pytest.mark.parametrize("sid", [1, 2, 3])
def test_group_of_two_devices(product_lib, oracle, sid):
    """Multi-GPU through the C ABI (rt_group: contexts + ncclCommInitAll + one ncclReduce onto device 0).  Tiles: the
    2-GPU image is byte-identical to the 1-GPU image.  Samples: it is the sum of the two shard renders (device 0 carries
    the reference's stream, device 1 subsequence pixel + n_pixels) scaled by 1/spp -- exactly what two virtual ranks on
    one GPU produce."""
    if _gpu_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    hs = product_lib.HostScene(sid)
    width, spp, b = 120, 10, 12
    s, cam = hs.camera(width, spp, b)
    ctx = product_lib.Context(0)
    one, one8, st1 = ctx.render_world(hs, s)
    grp = product_lib.Group(2)
    tiles, tiles8, st = grp.render_world(hs, s, shard=product_lib.RT_SHARD_TILES)
    assert np.array_equal(common.bits(tiles), common.bits(one)) and np.array_equal(tiles8, one8)
    assert st.paths == st1.paths and st.rays == st1.rays
    samp, samp8, st = grp.render_world(hs, s, shard=product_lib.RT_SHARD_SAMPLES)
    n = cam.img_width * cam.img_height
    ctx.upload(hs, product_lib.RT_TRAVERSAL_AUTO)
    total = np.zeros_like(one)
    rays = 0
    for g in range(2):
        q = hs.params(width, spp // 2, b, subsequence_offset=g * n, flags=product_lib.RT_RENDER_OUTPUT_SUM | product_lib.RT_RENDER_COUNT_RAYS)
        part, _, sp = ctx.render(q)
        total += part
        rays += sp.rays
    scale = np.float32(1.0) / np.float32(spp)
    assert np.array_equal(common.bits(samp), common.bits(total * scale))
    assert np.array_equal(samp8, oracle.resolve_rgb8(total * scale)) and st.rays == rays
    grp.close(); ctx.close(); hs.close()


@pytest.mark.gpu
def test_cli_two_gpus_writes_the_same_file(product_lib):
    if _gpu_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    exe = os.path.join(common.ROOT, "bin", "raytracing")
    with tempfile.TemporaryDirectory() as d:
        outs = []
        for extra in ([], ["g2"], ["g2", "d0"]):
            sub = os.path.join(d, "run%d" % len(outs)); os.mkdir(sub)
            r = subprocess.run([exe, "1", "s6", "b10", "w150"] + extra, cwd=sub, capture_output=True, text=True, timeout=600)
            assert r.returncode == 0, r.stderr
            assert r.stdout.splitlines()[0] == "cuda_calc launch error: no error"
            outs.append(open(os.path.join(sub, "cornell_box.ppm"), "rb").read())
        assert outs[0] == outs[1] == outs[2]
        sub = os.path.join(d, "samples"); os.mkdir(sub)
        r = subprocess.run([exe, "1", "s6", "b10", "w150", "g2", "d1"], cwd=sub, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0 and len(open(os.path.join(sub, "cornell_box.ppm"), "rb").read()) == len(outs[0])


@pytest.mark.gpu
@pytest.mark.parametrize("sid", [3, 1, 0, 12])
def test_fast_traversal_geometry_vs_oracle(gpu_ctx, oracle, product_lib, sid):
    """RT_TRAVERSAL_FAST flattens instances to world space: its hits are not bit-identical to the reference's (which edits the
    ray in place and undoes it inexactly), but they must be the SAME GEOMETRY -- same object hit or missed, t, point and
    normal equal to rounding -- for primary rays and for bounce rays that start on the instanced objects."""
    hs = product_lib.HostScene(sid)
    sc = common.oracle_scene(oracle, hs)
    L = oracle.lib()
    rng = np.random.default_rng(900 + sid)
    o, d, t = common.scene_rays(hs.settings, rng, 6000)
    gpu_ctx.upload(hs, product_lib.RT_TRAVERSAL_REFERENCE)
    first = gpu_ctx.probe_trace(o, d, t)
    # second generation: lambertian bounces off whatever the first generation hit
    idx = [i for i in range(len(first)) if first[i].hit][:4000]
    o2 = np.array([first[i].point[:] for i in idx], np.float32)
    u = rng.normal(size=(len(idx), 3)).astype(np.float32); u /= np.linalg.norm(u, axis=1, keepdims=True)
    d2 = (np.array([first[i].normal[:] for i in idx], np.float32) + u).astype(np.float32)
    t2 = rng.random(len(idx)).astype(np.float32)
    O, D, T = np.concatenate([o, o2]), np.concatenate([d, d2]), np.concatenate([t, t2])
    gpu_ctx.upload(hs, product_lib.RT_TRAVERSAL_FAST)
    got = gpu_ctx.probe_trace(O, D, T)
    bad = []
    for i in range(len(O)):
        oo, dd = O[i].copy(), D[i].copy()
        rec = oracle.OrcHit()
        h = L.orc_hit(C.byref(sc.c), oracle.fptr(oo), oracle.fptr(dd), float(T[i]), 0.001, float("inf"), C.byref(rec))
        g = got[i]
        if bool(g.hit) != bool(h):
            bad.append((i, "hit", g.hit, h)); continue
        if h:
            scale = max(1.0, abs(rec.t))
            if abs(g.t - rec.t) > 2e-3 * scale or np.abs(np.array(g.normal[:]) - np.array(rec.normal[:])).max() > 2e-2 \
                    or g.material_type != rec.mat[0]:
                bad.append((i, "rec", g.t, rec.t, list(g.normal[:]), list(rec.normal[:]), g.material_type, rec.mat[0]))
    print("fast-vs-oracle geometry, scene %d: %d rays, %d differ: %s" % (sid, len(O), len(bad), bad[:6]))
    assert len(bad) <= len(O) // 500, (sid, len(bad))           # grazing rays / ties may fall either way
    hs.close()


@pytest.mark.gpu
def test_coplanar_quads_tie_is_the_documented_difference(gpu_ctx, oracle, product_lib):
    """rt_api.h, RT_TRAVERSAL_AUTO: on a scene without wrappers the own BVH returns the reference's closest hit except for
    candidates at exactly the same t.  Two coplanar, overlapping quads of different materials (lambertian over metal):
    RT_TRAVERSAL_REFERENCE returns the oracle's hit bit for bit (the reference's inclusive planar test lets the quad it
    visits last win); RT_TRAVERSAL_AUTO returns the same t / point / normal and may name the other quad."""
    hs = product_lib.HostScene(None)
    lam = hs.mat_lambert_solid((0.2, 0.7, 0.2))
    met = product_lib.lib().rt_mat_metal(product_lib.api.Vec3(0.8, 0.8, 0.8), 0.0)
    hs.add_quad((-1, -1, -2), (2, 0, 0), (0, 2, 0), lam)
    hs.add_quad((-0.5, -0.5, -2), (2, 0, 0), (0, 2, 0), met)          # same plane z = -2, overlapping the first
    hs.add_sphere((0, -101.5, -2), 100.0, hs.mat_lambert_solid((0.5, 0.5, 0.5)))
    hs.build_bvh()
    sc = common.oracle_scene(oracle, hs)
    L = oracle.lib()
    L.orc_set_fma_profile(1)
    rng = np.random.default_rng(11)
    n = 3000
    o = np.zeros((n, 3), np.float32)
    d = np.stack([rng.uniform(-0.9, 1.4, n), rng.uniform(-0.9, 1.4, n), np.full(n, -2.0)], axis=1).astype(np.float32)
    t = np.zeros(n, np.float32)
    try:
        gpu_ctx.upload(hs, product_lib.RT_TRAVERSAL_REFERENCE)
        ref = gpu_ctx.probe_trace(o, d, t)
        gpu_ctx.upload(hs, product_lib.RT_TRAVERSAL_AUTO)
        own = gpu_ctx.probe_trace(o, d, t)
        both = other = 0
        for i in range(n):
            oo, dd = o[i].copy(), d[i].copy()
            rec = oracle.OrcHit()
            h = L.orc_hit(C.byref(sc.c), oracle.fptr(oo), oracle.fptr(dd), 0.0, 0.001, float("inf"), C.byref(rec))
            assert ref[i].hit == h and own[i].hit == h
            if not h:
                continue
            assert common.bits(np.float32(ref[i].t)) == common.bits(np.float32(rec.t)) and ref[i].material_type == rec.mat[0]
            assert common.bits(np.float32(own[i].t)) == common.bits(np.float32(rec.t))
            assert list(own[i].point[:]) == list(rec.point[:]) and list(own[i].normal[:]) == list(rec.normal[:])
            inside_both = -0.5 <= d[i, 0] <= 1.0 and -0.5 <= d[i, 1] <= 1.0
            if inside_both:
                both += 1
                other += own[i].material_type != rec.mat[0]
            else:
                assert own[i].material_type == rec.mat[0]
        print("coplanar quads: %d rays in the overlap, the own BVH names the other quad on %d" % (both, other))
        assert both > 300
    finally:
        L.orc_set_fma_profile(0)
    hs.close()
