"""rt_render_device_static: the render through the library's static device scratch (no allocation inside the call).

Collected LAST on purpose: the entry point has no counterpart in the reference;
the test checks it against rt_render_device through the C ABI, bit for
bit (the tolerance of the other same-stream tests in test_gpu_parity.py), on two calls in a row that share the array.
"""
import numpy as np
import pytest

import common


def _render_both(rt, ctx, hs, params):
    """(image of rt_render_device, image of rt_render_device_static) as float32 arrays; the buffers are the caller's."""
    import torch
    n = params.img_width * params.img_height * 3
    want = torch.zeros(n, dtype=torch.float32, device="cuda:0")
    got = torch.full((n,), -7.0, dtype=torch.float32, device="cuda:0")      # the call writes every pixel it renders
    ctx.render_device(params, want.data_ptr(), stats=False)
    ctx.render_device_static(params, got.data_ptr())
    ctx.synchronize()
    return want.cpu().numpy(), got.cpu().numpy()


@pytest.mark.gpu
def test_static_scratch_render_matches_rt_render_device_twice(product_lib):
    rt = product_lib
    ctx = rt.Context(0)
    try:
        # first call: Cornell box (reference order as a linear program), several chunks per pixel -> progress + parked states
        hs1 = rt.HostScene(1)
        ctx.upload(hs1, rt.RT_TRAVERSAL_AUTO)
        p1 = hs1.params(96, 96, 12, flags=0)
        assert hs1.render_plan(96, 96, 12).chunk_count > 1
        want1, got1 = _render_both(rt, ctx, hs1, p1)
        assert np.isfinite(want1).all() and want1.max() > 0
        assert np.array_equal(common.bits(got1), common.bits(want1))

        # second call in a row on the same array, which still holds the first call's counters, progress and states:
        # another scene (own BVH walk, traversal stack), a larger image, another seed, row bands
        hs2 = rt.HostScene(2)
        ctx.upload(hs2, rt.RT_TRAVERSAL_FAST)
        p2 = hs2.params(160, 32, 20, seed=99, flags=0)
        want2, got2 = _render_both(rt, ctx, hs2, p2)
        assert np.array_equal(common.bits(got2), common.bits(want2))
        p2.band_rows, p2.band_period, p2.band_phase = 4, 3, 1
        want3, got3 = _render_both(rt, ctx, hs2, p2)
        rows = np.arange(p2.img_height)
        mine = ((rows // 4) % 3) == 1
        got3 = got3.reshape(p2.img_height, p2.img_width, 3)
        want3 = want3.reshape(p2.img_height, p2.img_width, 3)
        assert np.array_equal(common.bits(got3[mine]), common.bits(want3[mine]))
        assert (got3[~mine] == -7.0).all()                 # rows of the other bands are left untouched

        # and the first render once more: nothing of the calls in between leaks into it
        ctx.upload(hs1, rt.RT_TRAVERSAL_AUTO)
        want4, got4 = _render_both(rt, ctx, hs1, p1)
        assert np.array_equal(common.bits(got4), common.bits(want1))
        hs1.close(); hs2.close()
    finally:
        ctx.close()


@pytest.mark.gpu
def test_static_scratch_limits_are_refused_with_a_clear_error(product_lib):
    import torch
    rt = product_lib
    ctx = rt.Context(0)
    try:
        hs = rt.HostScene(0)
        ctx.upload(hs, rt.RT_TRAVERSAL_AUTO)
        # 800 x 450 = 360000 pixels > 262144
        p = hs.params(800, 1, 2, flags=0)
        assert p.img_width * p.img_height > rt.RT_STATIC_SCRATCH_MAX_PIXELS
        buf = torch.zeros(p.img_width * p.img_height * 3, dtype=torch.float32, device="cuda:0")
        with pytest.raises(rt.RtError, match="360000 pixels.*262144"):
            ctx.render_device_static(p, buf.data_ptr())
        # the same image by rows fits: two halves through the static scratch == the whole through rt_render_device
        whole = torch.zeros_like(buf)
        ctx.render_device(p, whole.data_ptr(), stats=False)
        half = p.img_height // 2
        for r0, r1 in ((0, half), (half, p.img_height)):
            p.row_begin, p.row_end = r0, r1
            ctx.render_device_static(p, buf.data_ptr())
        ctx.synchronize()
        assert np.array_equal(common.bits(buf.cpu().numpy()), common.bits(whole.cpu().numpy()))
        p.row_begin, p.row_end = 0, 0
        with pytest.raises(rt.RtError, match="RESUME"):
            q = hs.params(64, 2, 2, flags=rt.RT_RENDER_OUTPUT_SUM | rt.RT_RENDER_RESUME)
            ctx.render_device_static(q, buf.data_ptr())
        hs.close()
    finally:
        ctx.close()
