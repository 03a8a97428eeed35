"""CPU: host logic of the multi-GPU render front end (cuda_raytracing_b200/render.py): the reference's argv grammar
(src/main.cu:30-33, 118-129) plus extensions, and the per-rank work plan for both shardings."""
import pytest

from cuda_raytracing_b200 import render


def test_reference_grammar():
    o = render.parse_args(["2", "s256", "b50"])
    assert (o["scene"], o["spp"], o["bounces"], o["width"], o["traversal"]) == (2, 256, 50, 0, 0)
    o = render.parse_args(["1", "b8", "s16"])                     # either order (README.md:30-45)
    assert (o["scene"], o["spp"], o["bounces"]) == (1, 16, 8)
    assert render.parse_args(["abc"])["scene"] == 0               # atoi
    assert render.parse_args([])["scene"] == 0
    # only argv[2] and argv[3] are examined for b/s (main.cu:118-129); later tokens are extensions only
    o = render.parse_args(["3", "s4096", "b50", "w3840", "t2", "s1"])
    assert (o["spp"], o["bounces"], o["width"], o["traversal"]) == (4096, 50, 3840, 2)
    assert render.parse_args(["0", "x9", "s4"])["spp"] == 4       # unknown letters are ignored


def test_options():
    o = render.parse_args(["3", "--shard=tiles", "--out=/tmp/x.ppm", "--band=16"])
    assert (o["shard"], o["out"], o["band"]) == ("tiles", "/tmp/x.ppm", 16)
    with pytest.raises(SystemExit):
        render.parse_args(["3", "--shard=rows"])
    with pytest.raises(SystemExit):
        render.parse_args(["3", "--frobnicate"])


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_plans_cover_the_image_once(world):
    W, H, spp = 40, 30, 21
    o = render.parse_args(["0"])
    total = 0
    for r in range(world):
        calls, tot = render.plan(o, r, world, W, H, spp)
        assert tot == spp
        for (a, b, s, off) in calls:
            assert (a, b) == (0, H) and off == r * W * H
            total += s
    assert total == spp                                            # every sample rendered exactly once
    o = render.parse_args(["0", "--shard=tiles", "--band=4"])
    rows = []
    for r in range(world):
        calls, tot = render.plan(o, r, world, W, H, spp)
        for (a, b, s, off) in calls:
            assert s == spp and (off == 0 or world == 1)
            rows += list(range(a, b))
    assert sorted(rows) == list(range(H))                          # every row rendered exactly once, reference stream
