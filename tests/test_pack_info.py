"""Host logic of the packer (no GPU): which walk a scene gets, what the treelet / fused-instance planting finds, and
that every built-in scene fits the device limits in every traversal mode (rt_scene_pack_info, host only)."""
import os

import pytest


@pytest.mark.parametrize("sid", list(range(20)))
def test_every_scene_packs_in_every_mode(product_lib, sid):
    hs = product_lib.HostScene(sid)
    for trav in (product_lib.RT_TRAVERSAL_AUTO, product_lib.RT_TRAVERSAL_REFERENCE, product_lib.RT_TRAVERSAL_FAST):
        i = hs.pack_info(trav)
        assert 1 <= i.stack_need <= i.stack_capacity
        assert i.spheres + i.planars >= 1
        tiny = not i.has_wrappers and i.spheres <= 4 and i.planars <= 4       # walked without its BVH (kernel mode 5)
        if trav == product_lib.RT_TRAVERSAL_FAST:           # one own BVH; instanced leaves where the scene has wrappers (kernel mode 6)
            assert (i.traversal_used, i.kernel_mode) == (product_lib.RT_TRAVERSAL_FAST, 5 if tiny else (6 if i.instanced_leaves else 1)) and i.own_nodes >= 1
            assert (i.instanced_leaves > 0) == bool(i.has_wrappers)
        if trav == product_lib.RT_TRAVERSAL_REFERENCE:          # the literal walk: no stand-ins
            assert (i.traversal_used, i.kernel_mode) == (product_lib.RT_TRAVERSAL_REFERENCE, 0)
            assert i.treelets == i.fused_instances == i.own_nodes == 0
            assert i.program_ops == 0
        if trav == product_lib.RT_TRAVERSAL_AUTO:               # exact results: own BVH only where order cannot matter
            assert i.traversal_used == (product_lib.RT_TRAVERSAL_REFERENCE if i.has_wrappers else product_lib.RT_TRAVERSAL_FAST)
            assert i.kernel_mode == ((5 if tiny else 1) if not i.has_wrappers else (2 if i.treelets + i.fused_instances else 3))
    hs.close()


def test_axis_aligned_quads_are_found_where_every_planar_of_the_program_is_one(product_lib):
    """Cornell box (scene 1): 6 walls + 2 instanced boxes of 6 quads, all axis-aligned in their own space -> reduced inside test;
    scene 0 has a pyramid of triangles in its program, the others spheres only or no program at all."""
    for sid in range(20):
        hs = product_lib.HostScene(sid)
        i = hs.pack_info()
        assert i.program_axis_quads == (1 if sid == 1 else 0), sid
        if sid == 1:
            assert (i.kernel_mode, i.planars, i.spheres) == (3, 18, 0)
            assert hs.pack_info(product_lib.RT_TRAVERSAL_REFERENCE).program_axis_quads == 0      # no program, no reduced form
        hs.close()


def test_final_scene_gets_treelets_and_fused_instances(product_lib):
    """Scene 3: 400 ground boxes (2400 quads) + light + 5 spheres at top level, 1000 spheres each in translate(rotate())."""
    hs = product_lib.HostScene(3)
    i = hs.pack_info()
    assert (i.traversal_used, i.kernel_mode) == (product_lib.RT_TRAVERSAL_REFERENCE, 2)
    assert (i.spheres, i.planars) == (1005, 2401)
    assert i.fused_instances == 1000
    assert i.treelets == 3 and i.treelet_prims == 2395          # the three wrapper-free subtrees with >= 8 primitives
    assert i.own_nodes == i.treelet_prims - i.treelets          # a binary tree over n leaves has n - 1 inner nodes
    off = hs.pack_info(treelet_min=0, fuse_min=0)
    assert off.kernel_mode == 0 and off.treelets == 0 and off.fused_instances == 0 and off.own_nodes == 0
    small = hs.pack_info(treelet_min=2)                         # smaller subtrees qualify too: more treelets, more primitives
    assert small.treelets > i.treelets and small.treelet_prims > i.treelet_prims
    fast = hs.pack_info(product_lib.RT_TRAVERSAL_FAST)          # one BVH over everything; the 1000 instanced spheres keep their own
    assert fast.spheres == 1005 and fast.instanced_leaves == 1000          # records and are tested in their own space
    assert fast.own_nodes == 2401 + 5 + 1000 - 1                # one leaf per primitive
    assert i.instanced_leaves == 0
    hs.close()


def test_small_instance_counts_keep_the_plain_walk(product_lib):
    """Cornell box: two instances, no wrapper-free subtree of 8 primitives -> plain reference-order kernel (measured
    faster); the knobs can force the hybrid walk (the GPU parity test does)."""
    hs = product_lib.HostScene(1)
    i = hs.pack_info()
    assert (i.kernel_mode, i.treelets, i.fused_instances) == (3, 0, 0)
    forced = hs.pack_info(treelet_min=2, fuse_min=1)
    assert forced.kernel_mode == 2 and forced.fused_instances == 2 and forced.treelets >= 1
    assert forced.stack_need <= forced.stack_capacity
    hs.close()


def test_linear_program_of_the_cornell_box(product_lib):
    """Scenes with wrappers that are too small for treelets get their reference-order walk laid out as a linear program
    (kernel mode 3).  Cornell box: 8 top-level objects under the 7 nodes of the reference's BVH; each of the two boxes is
    translate(rotate(structure of 6 quads)), a structure costs no op (its children are simply walked)."""
    hs = product_lib.HostScene(1)
    i = hs.pack_info()
    assert i.kernel_mode == 3
    assert i.program_boxes == 7 and i.program_wrappers == 2 * 2 and i.program_prims == 6 + 2 * 6
    # a wrapper is a way in and a way out; translate(rotate(x)) is ONE op each way (the two ways in follow each other at once)
    assert i.program_ops == i.program_boxes + i.program_prims + i.program_wrappers == 29
    saved = os.environ.get("RT_FUSE_PROGRAM_WRAPPERS")
    os.environ["RT_FUSE_PROGRAM_WRAPPERS"] = "0"                # one op per wrapper side, as before
    try:
        j = hs.pack_info()
    finally:
        os.environ.pop("RT_FUSE_PROGRAM_WRAPPERS")
        if saved is not None:
            os.environ["RT_FUSE_PROGRAM_WRAPPERS"] = saved
    assert j.program_ops == j.program_boxes + j.program_prims + 2 * j.program_wrappers == 33
    big = product_lib.HostScene(3).pack_info()                  # 7593 objects: far beyond the program limit, and it has treelets
    assert big.kernel_mode == 2 and big.program_ops == 0
    hs.close()


def test_tiny_scenes_skip_their_bvh(product_lib):
    """At most 4 spheres and 4 quads / triangles and no wrappers: every primitive is tested, two spheres at a time
    (kernel mode 5).  The Perlin scenes 17 and 19 are such scenes; scene 18 (5 quads) and scene 2 (485 spheres) are not."""
    modes = {}
    for sid in (13, 15, 16, 17, 19, 18, 2):
        hs = product_lib.HostScene(sid)
        modes[sid] = hs.pack_info().kernel_mode
        hs.close()
    assert modes == {13: 5, 15: 5, 16: 5, 17: 5, 19: 5, 18: 1, 2: 1}


def test_pack_info_rejects_bad_input(product_lib):
    hs = product_lib.HostScene(2)
    with pytest.raises(product_lib.RtError):
        hs.pack_info(traversal=7)
    hs.close()


def _desc_from_arrays(product_lib, h, r, t, root):
    import ctypes as C
    import numpy as np
    h = np.ascontiguousarray(h, dtype=np.uint8); r = np.ascontiguousarray(r, dtype=np.uint16); t = np.ascontiguousarray(t, dtype=np.uint8)
    d = product_lib.api.SceneDesc(h.ctypes.data, h.shape[0], r.ctypes.data, r.shape[0], root, t.ctypes.data, t.shape[0])
    return d, (h, r, t)


def _pack_info(product_lib, desc, traversal=0):
    import ctypes as C
    info = product_lib.api.PackInfo()
    rc = product_lib.lib().rt_scene_pack_info(C.byref(desc), traversal, 8, 8, C.byref(info))
    return rc, info


def test_orphan_hittable_keeps_the_bvh(product_lib):
    """The tiny-scene walk tests every primitive RECORD, so it is only chosen when every record is reachable from the
    world list: a caller-supplied array with an orphan hittable goes through the BVH (which never sees the orphan),
    like the reference."""
    import numpy as np
    hs = product_lib.HostScene(17)
    h, r, t = hs.arrays()
    desc, keep = _desc_from_arrays(product_lib, h, r, t, hs.world.root_wref)
    rc, info = _pack_info(product_lib, desc)
    assert rc == 0 and info.kernel_mode == 5 and info.spheres == 2
    orphan = np.concatenate([h, h[:1]])                     # a copy of sphere 0 that no world ref names
    desc, keep = _desc_from_arrays(product_lib, orphan, r, t, hs.world.root_wref)
    rc, info = _pack_info(product_lib, desc)
    assert rc == 0 and info.spheres == 3 and info.kernel_mode == 1
    hs.close()


def test_nesting_beyond_the_guard_is_refused(product_lib):
    """70 translations around one sphere: deeper than the packer's recursion guard.  Refused with RT_ERR_UNSUPPORTED in
    every mode instead of silently dropping the primitive."""
    import numpy as np
    hs = product_lib.HostScene(13)
    h, r, t = hs.arrays()
    sphere = h[0].copy()
    assert int(sphere[:4].view(np.uint32)[0]) == 0            # RT_OBJ_SPHERE
    depth = 70
    arr = np.zeros((depth + 2, 120), np.uint8)
    arr[0] = sphere
    for k in range(1, depth + 1):                             # translation k wraps hittable k - 1
        arr[k, :4] = np.array([4], np.uint32).view(np.uint8)
        arr[k, 8:10] = np.array([k - 1], np.uint16).view(np.uint8)
        arr[k, 12:24] = np.array([0.01, 0.0, 0.0], np.float32).view(np.uint8)
    arr[depth + 1, :4] = np.array([1], np.uint32).view(np.uint8)          # BVH node over world ref 0 (left == right)
    arr[depth + 1, 8:12] = np.array([0, 0], np.uint16).view(np.uint8)
    arr[depth + 1, 12:36] = np.array([-1e3, 1e3, -1e3, 1e3, -1e3, 1e3], np.float32).view(np.uint8)
    refs = np.array([depth, depth + 1], np.uint16)            # world: the outermost wrapper, then the node
    for trav in (0, 1, 2):
        desc, keep = _desc_from_arrays(product_lib, arr, refs, t, 1)
        rc, info = _pack_info(product_lib, desc, trav)
        assert rc == -4, (trav, rc)                           # RT_ERR_UNSUPPORTED
    shallow = arr.copy()
    shallow[depth + 1, 8:12] = np.array([0, 0], np.uint16).view(np.uint8)
    refs2 = np.array([10, depth + 1], np.uint16)              # ten wrappers: fine
    desc, keep = _desc_from_arrays(product_lib, shallow, refs2, t, 1)
    rc, info = _pack_info(product_lib, desc, 0)
    assert rc == 0 and info.has_wrappers == 1
    hs.close()
