"""Host logic of the render scheduler (rt_render_plan_query, no GPU): which kernel build a render gets, how a pixel's
samples are cut into chunks (big chunks while there is plenty of work, a short tail), claim grouping, item order."""
import pytest

CONFIGS = {"C1": (0, 16, 8, 0), "C2": (2, 256, 50, 0), "C3": (17, 512, 50, 0), "C3b": (19, 512, 50, 0), "C4": (1, 1024, 50, 0),
           "C4/8": (1, 128, 50, 0), "C5r": (3, 16, 50, 0), "C5": (3, 4096, 50, 3840)}


def _plan(product_lib, name, **kw):
    sid, spp, b, w = CONFIGS[name]
    hs = product_lib.HostScene(sid)
    pl = hs.render_plan(w, spp, b, **kw)
    hs.close()
    return pl, spp


@pytest.mark.parametrize("name", list(CONFIGS))
def test_chunk_table_is_a_partition_of_the_samples(product_lib, name):
    pl, spp = _plan(product_lib, name)
    assert pl.chunk_samples == 0 and 1 <= pl.chunk_count <= 32
    begin = list(pl.chunk_begin)
    assert begin[0] == 0 and begin[pl.chunk_count] == spp
    sizes = [begin[k + 1] - begin[k] for k in range(pl.chunk_count)]
    assert all(s > 0 for s in sizes)
    assert all(sizes[k] >= sizes[k + 1] for k in range(len(sizes) - 1))      # never growing: the tail is the shortest
    assert all(x == spp for x in begin[pl.chunk_count:])


def test_plans_of_the_baseline_configs(product_lib):
    got = {}
    for name in CONFIGS:
        pl, spp = _plan(product_lib, name)
        got[name] = (pl.kernel_mode, pl.resident_lanes, list(pl.chunk_begin)[:pl.chunk_count + 1], pl.claim_group, pl.scattered)
    assert got["C1"] == (4, 148 * 640, [0, 16], 1, 0)                        # short-lived pixels: 5 x 96 build, one chunk
    assert got["C2"] == (1, 148 * 768, [0, 57, 114, 171, 214, 235, 246, 256], 32, 0)
    assert got["C3"] == (5, 148 * 768, [0, 256, 512], 1, 1)                  # tiny image: 2 equal slices, scattered order
    assert got["C4"][:2] == (3, 148 * 1024) and got["C4"][2][:13] == [76 * k for k in range(13)]   # 12 equal chunks of 1/32 of a lane's share ...
    assert got["C4"][2][13:] == [968, 996, 1010, 1024]                       # ... then a halving tail
    assert got["C4/8"] == (3, 148 * 1024, [0, 32, 64, 96, 112, 120, 128], 1, 0)   # an 8-GPU share of C4: never below 32 samples in the bulk
    assert got["C5r"] == (2, 148 * 640, [0, 8, 16], 1, 1)                    # short hybrid render: two chunks of 8, scattered
    assert got["C5"][0] == 2 and got["C5"][4] == 0 and len(got["C5"][2]) <= 3


def test_plan_edge_cases(product_lib):
    hs = product_lib.HostScene(2)
    info = hs.pack_info()
    import ctypes as C
    L = product_lib.lib()
    pl = product_lib.api.RenderPlan()
    assert L.rt_render_plan_query(C.byref(info), 1200, 675, 0, 50, 148, C.byref(pl)) == 0
    assert (pl.chunk_samples, pl.chunk_count) == (1, 1)                      # spp 0: one (empty) chunk, the kernel writes zeros
    assert L.rt_render_plan_query(C.byref(info), 1200, 675, 1, 300, 148, C.byref(pl)) == 0
    assert list(pl.chunk_begin)[:2] == [0, 1] and pl.chunk_count == 1
    assert L.rt_render_plan_query(C.byref(info), 0, 675, 1, 50, 148, C.byref(pl)) != 0
    assert L.rt_render_plan_query(C.byref(info), 1200, 675, 1, -1, 148, C.byref(pl)) != 0
    for spp in (7, 8, 9, 17, 100, 1000, 65536):                               # any spp: a partition, at most 32 chunks
        assert L.rt_render_plan_query(C.byref(info), 64, 36, spp, 50, 148, C.byref(pl)) == 0
        b = list(pl.chunk_begin)
        assert b[0] == 0 and b[pl.chunk_count] == spp and pl.chunk_count <= 32
        assert all(b[k + 1] > b[k] for k in range(pl.chunk_count))
    hs.close()
