"""CPU: the N>1 plumbing on 2 gloo ranks -- sample shards rendered by the CPU oracle (the product cannot run
without a GPU), summed with torch.distributed.reduce exactly as bench.py does with NCCL, compared with a
single-process evaluation of the same shards."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import common


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, out_path):
    sys.path.insert(0, common.ROOT); sys.path.insert(0, os.path.join(common.ROOT, "tests"))
    import oracleabi
    import cuda_raytracing_b200 as rt
    from cuda_raytracing_b200 import sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    hs = rt.HostScene(13)
    p = hs.params(48, 3, 6)
    n = p.img_width * p.img_height
    off, spp, total = sharding.sample_shard(rank, world, n, per_rank_spp=3)
    cam = common.oracle_camera(oracleabi, p)
    cam.pixel_samples_scale = 1.0                       # SUM output, like RT_RENDER_OUTPUT_SUM
    img, _ = oracleabi.render(common.oracle_scene(oracleabi, hs), cam, subseq_offset=off, nthreads=1)
    t = torch.from_numpy(img.copy())
    dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)
    if rank == 0:
        np.save(out_path, (t.numpy() * np.float32(sharding.reduce_scale(total))))
    rows = sharding.tile_rows(rank, world, p.img_height, band=4)
    gathered = [None] * world
    dist.all_gather_object(gathered, rows)
    if rank == 0:
        covered = sorted(r for part in gathered for (a, b) in part for r in range(a, b))
        assert covered == list(range(p.img_height))
    dist.barrier()
    dist.destroy_process_group()


def test_sample_shard_reduce_two_ranks(tmp_path, oracle, product_lib):
    from cuda_raytracing_b200 import sharding
    out = str(tmp_path / "reduced.npy")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = np.load(out)
    hs = product_lib.HostScene(13)
    p = hs.params(48, 3, 6)
    n = p.img_width * p.img_height
    cam = common.oracle_camera(oracle, p); cam.pixel_samples_scale = 1.0
    sc = common.oracle_scene(oracle, hs)
    a, _ = oracle.render(sc, cam, subseq_offset=0, nthreads=1)
    b, _ = oracle.render(sc, cam, subseq_offset=n, nthreads=1)
    want = (a + b) * np.float32(sharding.reduce_scale(6))
    assert np.array_equal(common.bits(got), common.bits(want))
    # rank 0's shard is the reference stream itself
    cam1 = common.oracle_camera(oracle, p)
    ref, _ = oracle.render(sc, cam1, nthreads=1)
    assert np.array_equal(common.bits(a * np.float32(1.0 / 3)), common.bits(ref))
    hs.close()


def test_shard_arithmetic():
    from cuda_raytracing_b200 import sharding
    assert [sharding.sample_shard(r, 4, 100, total_spp=10)[1] for r in range(4)] == [3, 3, 2, 2]
    assert sharding.sample_shard(3, 8, 810000, per_rank_spp=256) == (3 * 810000, 256, 2048)
    rows = [sharding.tile_rows(r, 3, 50, band=8) for r in range(3)]
    flat = sorted(x for part in rows for (a, b) in part for x in range(a, b))
    assert flat == list(range(50))
    with pytest.raises(ValueError):
        sharding.sample_shard(4, 4, 10, total_spp=4)
