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
