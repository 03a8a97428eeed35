"""The linear program of the reference-order walk (kernel mode 3, rt_pack.cpp::emit_program) against an independent
replay of the reference's stack machine (rt.cu:1079-1251 as trace_reference mirrors it) over the scene's own arrays:
whatever set of boxes a ray misses, the program must visit the same things in the same order.  Host only."""
import struct

import numpy as np
import pytest

SPHERE, BVH, TRIANGLE, QUAD, TRANSLATION, YROTATION, STRUCTURE = range(7)      # rt.cu:145-153


def parse(h, refs):
    """hittable array (n x 120 bytes) -> list of dicts (only what the walk needs)."""
    out, n_sph, n_pl = [], 0, 0
    for row in h:
        b = row.tobytes()
        t = struct.unpack_from("<I", b, 0)[0]
        o = dict(type=t)
        if t == SPHERE:
            o["prim"] = n_sph; n_sph += 1
        elif t in (TRIANGLE, QUAD):
            o["prim"] = n_pl; n_pl += 1
        elif t == BVH:
            left, right = struct.unpack_from("<HH", b, 8)
            o.update(left=int(refs[left]), right=int(refs[right]), box=b[12:36])
        elif t == TRANSLATION:
            o.update(child=struct.unpack_from("<H", b, 8)[0], rec=b[12:24])
        elif t == YROTATION:
            o.update(child=struct.unpack_from("<H", b, 8)[0], rec=b[12:20])
        elif t == STRUCTURE:
            o.update(first=struct.unpack_from("<H", b, 8)[0], count=struct.unpack_from("<Q", b, 48)[0])
        out.append(o)
    return out


def enters(box_bytes, seed):
    """the 'ray': a deterministic coin per (box bounds, seed)"""
    return (hash((box_bytes, seed)) & 3) != 0


def stack_machine(objs, root, seed):
    """visit order of the reference's walk: explicit stack, wrappers pushed again as their own way out"""
    events, stack, in_t, in_r = [], [root], False, False
    while stack:
        cur = stack.pop()
        o = objs[cur]
        t = o["type"]
        if t in (SPHERE, TRIANGLE, QUAD):
            events.append(("prim", t, o["prim"], in_t or in_r))
        elif t == BVH:
            e = enters(o["box"], seed)
            events.append(("box", o["box"], e))
            if e:
                stack.append(o["right"]); stack.append(o["left"])          # left is popped first
        elif t == TRANSLATION:
            if not in_t:
                events.append(("t_in", o["rec"])); stack.append(cur); stack.append(o["child"]); in_t = True
            else:
                events.append(("t_out", o["rec"])); in_t = False
        elif t == YROTATION:
            if not in_r:
                events.append(("r_in", o["rec"])); stack.append(cur); stack.append(o["child"]); in_r = True
            else:
                events.append(("r_out", o["rec"])); in_r = False
        elif t == STRUCTURE:
            for k in range(o["count"]):
                stack.append(o["first"] + k)                               # the last child is popped first
    return events


def run_program(ops, seed):
    events, pc = [], 0
    f = ops.view(np.float32)
    while pc < len(ops):
        code, arg = int(ops[pc, 7]), int(ops[pc, 6])
        kind = code & 7
        nxt = pc + 1
        if kind in (SPHERE, TRIANGLE, QUAD):
            events.append(("prim", kind, arg, bool(code & 0x100)))
        elif kind == BVH:
            box = f[pc, :6].tobytes()
            e = enters(box, seed)
            events.append(("box", box, e))
            if not e:
                assert pc < arg <= len(ops)                                 # a skip only goes forward
                nxt = arg
        elif kind == TRANSLATION:
            events.append(("t_in", f[pc, :3].tobytes()))
            if code & 0x200:                                                # translate(rotate(x)) as one op: {offset,0} {sin,cos,..}
                events.append(("r_in", f[pc, 4:6].tobytes()))
        elif kind == 6:
            events.append(("t_out", f[pc, :3].tobytes()))
        elif kind == YROTATION:
            events.append(("r_in", f[pc, :2].tobytes()))
        else:
            events.append(("r_out", f[pc, :2].tobytes()))
            if code & 0x200:                                                # ... and the way out: {sin,cos,offset.xy} {offset.z,..}
                events.append(("t_out", f[pc, 2:5].tobytes()))
        pc = nxt
    return events


@pytest.mark.parametrize("sid", list(range(20)))
def test_program_replays_the_stack_machine(product_lib, sid):
    hs = product_lib.HostScene(sid)
    info = hs.pack_info()
    ops = hs.program()
    assert len(ops) == info.program_ops
    if info.kernel_mode != 3:
        assert len(ops) == 0                                                # own BVH (no wrappers), or scene 3 (hybrid walk)
        hs.close()
        return
    h, refs, _ = hs.arrays()
    objs = parse(h, refs)
    root = int(refs[hs.world.root_wref])
    for seed in range(40):
        assert run_program(ops, seed) == stack_machine(objs, root, seed), (sid, seed)
    # the plane of a quad / triangle rides in its op: the first 16 bytes of its geometry are q..., so compare with normal, d
    planars = [b for b, o in zip(h, objs) if o["type"] in (TRIANGLE, QUAD)]
    for pc in range(len(ops)):
        if int(ops[pc, 7]) & 7 in (TRIANGLE, QUAD):
            g = planars[int(ops[pc, 6])].tobytes()[8:72]                    # rt_geometry: q, u, v, d, w, normal
            q_u_v_d_w_n = struct.unpack("<16f", g)
            normal, d = q_u_v_d_w_n[13:16], q_u_v_d_w_n[9]
            assert tuple(ops[pc, :4].view(np.float32)) == tuple(np.float32(x) for x in (*normal, d))
    hs.close()


def test_program_export_rejects_bad_input(product_lib):
    import ctypes as C
    n = C.c_uint32(0)
    assert product_lib.lib().rt_scene_program(None, None, 0, C.byref(n)) != 0


def test_wrapper_inside_a_wrapper_of_its_own_kind_keeps_the_stack_walk(product_lib):
    """The reference's in_translation / in_rotation are flags, not counters: a translation met inside a translation is
    taken for the way OUT.  Only the stack walk mirrors that, so such a scene gets no program (the shipped scenes never
    nest like this; the array is patched by hand: scene 0's translate(rotate(pyramid)) becomes translate(translate()))."""
    import ctypes as C
    hs = product_lib.HostScene(0)
    h, refs, arena = hs.arrays()
    types = h[:, :4].copy().view(np.uint32)[:, 0]
    rot = [i for i in range(len(h)) if types[i] == YROTATION]
    assert len(rot) == 1
    h[rot[0], :4] = np.frombuffer(struct.pack("<I", TRANSLATION), dtype=np.uint8)      # same layout: child u16 at +8, 12 bytes of record at +12
    refs = np.ascontiguousarray(refs)
    arena = np.ascontiguousarray(arena)
    from cuda_raytracing_b200 import api
    d = api.SceneDesc(h.ctypes.data, len(h), refs.ctypes.data, len(refs), hs.world.root_wref, arena.ctypes.data, len(arena))
    info = api.PackInfo()
    assert product_lib.lib().rt_scene_pack_info(C.byref(d), product_lib.RT_TRAVERSAL_AUTO, 8, 8, C.byref(info)) == 0
    assert (info.kernel_mode, info.program_ops, info.has_wrappers) == (0, 0, 1)
    n = C.c_uint32(7)
    assert product_lib.lib().rt_scene_program(C.byref(d), None, 0, C.byref(n)) == 0 and n.value == 0
    hs.close()


@pytest.mark.parametrize("sid", list(range(20)))
def test_image_of_the_program_is_the_program_in_relative_addresses(product_lib, sid):
    """What the kernels walk (rt_scene_linear_image, built at upload): the program's ops with their indices turned into byte
    offsets relative to the op, an end op, then the records the ops test -- so that the walk needs no base address.  Every
    offset must land where the index pointed; the two frequent kinds carry their flag bit."""
    hs = product_lib.HostScene(sid)
    ops = hs.program()
    for axis in (False, True):
        img = hs.linear_image(axis)
        if len(ops) == 0:
            assert len(img) == 0
            continue
        info = hs.pack_info()
        n = len(ops)
        assert len(img) == 2 * (n + 1) + 4 * info.planars + 2 * info.spheres
        planar0, sphere0 = 2 * (n + 1), 2 * (n + 1) + 4 * info.planars            # first 16-byte word of the record arrays
        iops = img[:2 * (n + 1)].reshape(n + 1, 8)
        for i in range(n):
            code, kind = int(iops[i, 7]), int(ops[i, 7]) & 7
            assert code & ~0x18 == int(ops[i, 7])                                  # same code, plus the flag bits
            assert bool(code & 0x8) == (kind in (TRIANGLE, QUAD)) and bool(code & 0x10) == (kind == BVH)
            assert np.array_equal(iops[i, :4], ops[i, :4])
            if kind == BVH:
                delta = int(iops[i, 6].astype(np.int32))
                assert delta % 32 == 0 and i + delta // 32 == int(ops[i, 6]) and delta > 0
                assert np.array_equal(iops[i, 4:6], ops[i, 4:6])
            elif kind in (TRIANGLE, QUAD, SPHERE):
                assert iops[i, 6] == ops[i, 6]                                      # the primitive index stays (it names the material)
                delta = int(iops[i, 4].astype(np.int32))
                assert delta % 16 == 0
                word = 2 * i + delta // 16
                if kind == SPHERE:
                    assert word == sphere0 + 2 * int(ops[i, 6])
                else:
                    assert word == planar0 + 4 * int(ops[i, 6])
                    assert np.array_equal(img[word], iops[i, :4])                  # {n, d}: the first word of the record rides in the op
            else:
                assert np.array_equal(iops[i, 4:7], ops[i, 4:7])                   # wrappers: untouched
        assert int(iops[n, 7]) == 0x400 and not iops[n, :7].any()                   # the end op
        if axis and info.program_axis_quads:                                        # reduced records: one non-zero in each S
            rec = img[planar0:planar0 + 4 * info.planars].view(np.float32).reshape(info.planars, 4, 4)
            assert (np.count_nonzero(rec[:, 2, :3], axis=1) == 1).all() and (np.count_nonzero(rec[:, 3, :3], axis=1) == 1).all()
            assert (rec[:, 2, 3] == 0).all() and (rec[:, 3, 3] == 0).all()
            general = hs.linear_image(False)[planar0:planar0 + 4 * info.planars].view(np.float32).reshape(info.planars, 4, 4)
            assert np.array_equal(rec[:, 0], general[:, 0]) and np.array_equal(rec[:, 1, :3], general[:, 1, :3])
        elif not axis:
            pass
    hs.close()


def _f32(x):
    return np.asarray(x, dtype=np.float32)


def _fma(a, b, c):
    """float32 fma through float64: the product of two float32 is exact in float64; adding an exact zero (all this test
    needs) is exact as well"""
    return _f32(_f32(a).astype(np.float64) * _f32(b).astype(np.float64) + _f32(c).astype(np.float64))


def _mul(a, b):
    return _f32(_f32(a).astype(np.float64) * _f32(b).astype(np.float64))


def _dot_yfirst(a, b):                                     # rt_math.cuh: fma(z,z', fma(x,x', y*y'))
    return _fma(a[..., 2], b[..., 2], _fma(a[..., 0], b[..., 0], _mul(a[..., 1], b[..., 1])))


def _cross(a, b):                                          # rt_math.cuh: fma(a, b, -(c*d)) per component
    return np.stack([_fma(a[..., 1], b[..., 2], -_mul(a[..., 2], b[..., 1])), _fma(a[..., 2], b[..., 0], -_mul(a[..., 0], b[..., 2])),
                     _fma(a[..., 0], b[..., 1], -_mul(a[..., 1], b[..., 0]))], axis=-1)


def test_reduced_inside_test_of_axis_aligned_quads_has_the_general_roundings(product_lib):
    """Cornell box: alpha / beta of the reduced records (w.k * dot(pl, S)) against the reference's formula
    (dot(w, cross(pl, v)), dot(w, cross(u, pl)), rt.cu:1069-1070) with the kernels' fma chains, bit for bit, on random
    points around every quad -- the terms the reduced form leaves out are products with an exact zero."""
    hs = product_lib.HostScene(1)
    info = hs.pack_info()
    assert info.program_axis_quads == 1
    n = info.program_ops
    first = 2 * (n + 1)
    gen = hs.linear_image(False)[first:first + 4 * info.planars].view(np.float32).reshape(info.planars, 4, 4)
    red = hs.linear_image(True)[first:first + 4 * info.planars].view(np.float32).reshape(info.planars, 4, 4)
    rng = np.random.default_rng(5)
    for k in range(info.planars):
        q, w = gen[k, 1, :3], np.array([gen[k, 1, 3], gen[k, 2, 3], gen[k, 3, 3]], dtype=np.float32)
        v, u = gen[k, 2, :3], gen[k, 3, :3]
        pts = _f32(q + rng.uniform(-1.2, 1.2, (4000, 1)) * u + rng.uniform(-1.2, 1.2, (4000, 1)) * v + rng.normal(0, 1e-3, (4000, 3)))
        pl = _f32(pts - q)
        alpha = _dot_yfirst(np.broadcast_to(w, pl.shape), _cross(pl, np.broadcast_to(v, pl.shape)))
        beta = _dot_yfirst(np.broadcast_to(w, pl.shape), _cross(np.broadcast_to(u, pl.shape), pl))
        wk = red[k, 1, 3]
        alpha_r = _mul(wk, _dot_yfirst(pl, np.broadcast_to(red[k, 2, :3], pl.shape)))
        beta_r = _mul(wk, _dot_yfirst(pl, np.broadcast_to(red[k, 3, :3], pl.shape)))
        # identical bits, up to the sign of an exact zero (which no comparison of the inside test sees)
        for a, b in ((alpha, alpha_r), (beta, beta_r)):
            same = (a.view(np.uint32) == b.view(np.uint32)) | ((a == 0) & (b == 0))
            assert same.all(), (k, a[~same][:3], b[~same][:3])
        assert ((alpha >= 0) & (alpha <= 1)).any() and ((alpha < 0) | (alpha > 1)).any()      # the sample straddles the quad
    hs.close()
