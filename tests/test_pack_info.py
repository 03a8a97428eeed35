"""Host logic of the packer (no GPU): which walk a scene gets, what the treelet / fused-instance planting finds, and
that every built-in scene fits the device limits in every traversal mode (rt_scene_pack_info, host only)."""
import pytest


@pytest.mark.parametrize("sid", list(range(20)))
def test_every_scene_packs_in_every_mode(product_lib, sid):
    hs = product_lib.HostScene(sid)
    for trav in (product_lib.RT_TRAVERSAL_AUTO, product_lib.RT_TRAVERSAL_REFERENCE, product_lib.RT_TRAVERSAL_FAST):
        i = hs.pack_info(trav)
        assert 1 <= i.stack_need <= i.stack_capacity
        assert i.spheres + i.planars >= 1
        tiny = not i.has_wrappers and i.spheres <= 4 and i.planars <= 4       # walked without its BVH (kernel mode 5)
        if trav == product_lib.RT_TRAVERSAL_FAST:
            assert (i.traversal_used, i.kernel_mode) == (product_lib.RT_TRAVERSAL_FAST, 5 if tiny else 1) and i.own_nodes >= 1
        if trav == product_lib.RT_TRAVERSAL_REFERENCE:          # the literal walk: no stand-ins
            assert (i.traversal_used, i.kernel_mode) == (product_lib.RT_TRAVERSAL_REFERENCE, 0)
            assert i.treelets == i.fused_instances == i.own_nodes == 0
            assert i.program_ops == 0
        if trav == product_lib.RT_TRAVERSAL_AUTO:               # exact results: own BVH only where order cannot matter
            assert i.traversal_used == (product_lib.RT_TRAVERSAL_REFERENCE if i.has_wrappers else product_lib.RT_TRAVERSAL_FAST)
            assert i.kernel_mode == ((5 if tiny else 1) if not i.has_wrappers else (2 if i.treelets + i.fused_instances else 3))
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
    fast = hs.pack_info(product_lib.RT_TRAVERSAL_FAST)          # instances flattened to world space: 1000 more sphere records
    assert fast.spheres == 2005 and fast.own_nodes == 2401 + 5 + 1000 - 1   # one leaf per primitive in world space
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
    assert i.kernel_mode == 3 and i.program_ops == i.program_boxes + i.program_prims + 2 * i.program_wrappers
    assert i.program_boxes == 7 and i.program_wrappers == 2 * 2 and i.program_prims == 6 + 2 * 6
    assert i.program_ops == 33                                  # every wrapper is a way in and a way out
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
