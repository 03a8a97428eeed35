"""CPU, only where oracle/_ref/libref_host.so exists (built from /root/reference by oracle/build_ref.sh): the
oracle against the REAL reference host code, live, bit for bit.  This is the pin of record; the golden files
carry it to machines without the reference."""
import ctypes as C

import numpy as np
import pytest

import common
import refabi

pytestmark = pytest.mark.skipif(not refabi.available(), reason="oracle/_ref/libref_host.so not built")


def test_struct_sizes_match_survey():
    L = refabi.lib()
    sizes = [L.ref_sizeof(i) for i in range(15)]
    assert sizes == [120, 32, 6144, 24, 80, 112, 28, 40, 36, 48, 64, 28, 80, 48, 64]


def test_xorwow_random_streams(oracle):
    R, O = refabi.lib(), oracle.lib()
    rng = np.random.default_rng(7)
    for _ in range(24):
        seed = int(rng.integers(0, 2 ** 63)); sub = int(rng.integers(0, 2 ** 36)); off = int(rng.integers(0, 2 ** 20))
        a = np.zeros(6, np.float32); b = np.zeros(6, np.float32)
        sa = (C.c_uint32 * 6)(); sb = (C.c_uint32 * 6)()
        R.ref_xorwow_uniforms(seed, sub, off, 6, refabi.fptr(a), sa)
        O.orc_xorwow_uniforms(seed, sub, off, 6, oracle.fptr(b), sb)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)) and list(sa) == list(sb)


@pytest.mark.parametrize("sid", list(range(20)))
def test_scene_arrays_and_image(oracle, product_lib, sid):
    """Product C scene builder == reference scene builder (bytes), and oracle image == reference host image (bits)."""
    rs = refabi.build_scene(sid)
    rh, rw, rt_ = refabi.scene_arrays(rs)
    hs = product_lib.HostScene(sid)
    h, w, t = hs.arrays()
    assert np.array_equal(h, rh) and np.array_equal(w, rw)
    assert np.array_equal(common.mask_texture_pointers(t), common.mask_texture_pointers(rt_))
    assert hs.world.root_wref == rs.root_wref
    assert product_lib.lib().rt_world_stack_bound(C.byref(hs.world)) == rs.max_bvh_depth
    width = 96 if sid != 3 else 64
    cam = refabi.camera(rs, width, 2, 10)
    ref_img, ref_rays = refabi.render_host(rs, cam)
    sc = common.oracle_scene(oracle, hs)
    img, rays = oracle.render(sc, oracle.OrcCamera.from_buffer_copy(bytes(cam)))
    assert rays == ref_rays
    assert np.array_equal(common.bits(img), common.bits(ref_img))
    # product camera basis == render()'s (rt.cu:1657-1692)
    p = hs.params(width, 2, 10)
    ocam = common.oracle_camera(oracle, p)
    assert bytes(ocam) == bytes(cam)
    refabi.lib().ref_scene_free(C.byref(rs))
    hs.close()
