"""`python -m cuda_raytracing_b200.render NRSCENY [bN] [sN] [wN] [tN] [--shard samples|tiles]` -- the reference's CLI
(src/main.cu:29-131) on 1..8 B200s of one box: one process per GPU (torchrun), NCCL over NVLink for the one exchange.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        -m cuda_raytracing_b200.render 3 s4096 b50 w3840 t2

Sharding (cuda_raytracing_b200/sharding.py):
  samples  rank g renders its share of the samples of EVERY pixel (XORWOW subsequence = pixel + g * n_pixels, so
           rank 0 carries the reference's stream) into an f32 SUM image; one `reduce(SUM)` onto rank 0; rank 0
           scales by 1/spp, resolves (gamma + 8 bit) and writes the PPM.  Statistical (tier-2) parity.
  tiles    rank g renders interleaved bands of rows (~32 per rank, ONE launch: rt_render_params.band_*) with the
           reference's stream (subsequence = pixel); bands are disjoint, so the same `reduce(SUM)` over zero-initialised
           images assembles the picture exactly (x + 0 == x): bit-identical to the 1-GPU image.
Everything else (scene construction in C, packing, the render kernel) is the single-GPU path of the C ABI.
Prints the reference's four stdout lines (rt.cu:1759-1768) on rank 0 and one JSON line with the statistics on stderr.
"""
import json
import os
import sys
import time


def parse_args(argv):
    """The reference's grammar (argv[1] scene via atoi, 'bN'/'sN' in argv[2..3]) plus the extensions it ignores."""
    opts = {"scene": 0, "bounces": 0, "spp": 0, "width": 0, "traversal": 0, "shard": "samples", "out": None, "band": 0}
    pos = [a for a in argv if not a.startswith("--")]
    flags = [a for a in argv if a.startswith("--")]
    if pos:
        try:
            opts["scene"] = int(pos[0])
        except ValueError:
            opts["scene"] = 0                             # atoi("abc") == 0
    for i, a in enumerate(pos[1:], start=2):
        if not a:
            continue
        num = a[1:]
        val = int(num) if num.lstrip("-").isdigit() else 0
        if i <= 3 and a[0] == "b":
            opts["bounces"] = val
        elif i <= 3 and a[0] == "s":
            opts["spp"] = val
        elif a[0] == "w" and val > 0:
            opts["width"] = val
        elif a[0] == "t":
            opts["traversal"] = val
    it = iter(flags)
    for f in it:
        key, _, val = f[2:].partition("=")
        if key == "shard":
            if val not in ("samples", "tiles"):
                raise SystemExit("--shard=samples|tiles")
            opts["shard"] = val
        elif key == "out":
            opts["out"] = val
        elif key == "band":
            opts["band"] = max(1, int(val))
        else:
            raise SystemExit("unknown option %s" % f)
    return opts


def band_rows(opts, world, height):
    """Rows per interleaved band of the tile sharding: ~32 bands per rank balance sky against interior."""
    return opts["band"] or max(4, height // (world * 32))


def plan(opts, rank, world, width, height, spp):
    """What this rank renders: a list of (row_begin, row_end, spp, subsequence_offset) and the total spp.  For the tile
    sharding the list names the rank's bands; main() renders them in ONE launch (rt_render_params.band_*)."""
    from . import sharding
    if opts["shard"] == "tiles" and world > 1:
        return [(a, b, spp, 0) for (a, b) in sharding.tile_rows(rank, world, height, band_rows(opts, world, height))], spp
    off, mine, total = sharding.sample_shard(rank, world, width * height, total_spp=spp)
    return ([(0, height, mine, off)] if mine > 0 else []), total


def main(argv=None):
    opts = parse_args(sys.argv[1:] if argv is None else argv)
    import torch
    import torch.distributed as dist
    from . import api as rt
    from . import sharding

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("cuda_raytracing_b200.render: no CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        hs = rt.HostScene(opts["scene"])
    except ValueError:
        if rank == 0:
            print("Unknown scene")
        return 1
    t_wall = time.perf_counter()
    ctx = rt.Context(local)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    ctx.upload(hs, opts["traversal"])
    s, cam = hs.camera(opts["width"], opts["spp"], opts["bounces"])
    W, H, spp = cam.img_width, cam.img_height, s.samples_per_pixel
    n = W * H
    calls, total_spp = plan(opts, rank, world, W, H, spp)
    pixels = torch.zeros(n * 3, dtype=torch.float32, device=dev)
    rgb8 = torch.zeros(n * 3, dtype=torch.uint8, device=dev) if rank == 0 else None
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    # one-off costs that are not rendering: the CUDA module load behind the first launch, the device allocations for
    # this image size (a 1-sample, 1-bounce render of this rank's own rows into a scratch image) and the NCCL
    # communicator set-up (a 1-element all_reduce).  The timed render recomputes its XORWOW states (NO_RNG_CACHE).
    tiles = opts["shard"] == "tiles" and world > 1
    scratch = torch.zeros(n * 3, dtype=torch.float32, device=dev)
    if calls:
        wp = hs.params(opts["width"], 1, 1, subsequence_offset=calls[0][3], flags=0)
        if tiles:
            wp.band_rows, wp.band_period, wp.band_phase = band_rows(opts, world, H), world, rank
        ctx.render_device(wp, scratch.data_ptr(), stats=True)
    del scratch
    if world > 1:
        warm = torch.zeros(1, device=dev)
        dist.all_reduce(warm)
    torch.cuda.synchronize()
    paths = rays = 0
    e0.record(stream)
    st = None
    if tiles and calls:                                                  # all of this rank's bands in one launch
        p = hs.params(opts["width"], spp, opts["bounces"], flags=rt.RT_RENDER_COUNT_RAYS | rt.RT_RENDER_NO_RNG_CACHE)
        p.band_rows, p.band_period, p.band_phase = band_rows(opts, world, H), world, rank
        st = ctx.render_device(p, pixels.data_ptr(), stats=True)
        paths, rays = st.paths, st.rays
    elif not tiles:
        for (r0, r1, my_spp, off) in calls:
            p = hs.params(opts["width"], my_spp, opts["bounces"], subsequence_offset=off,
                          flags=rt.RT_RENDER_COUNT_RAYS | rt.RT_RENDER_OUTPUT_SUM | rt.RT_RENDER_NO_RNG_CACHE)
            p.row_begin, p.row_end = r0, r1
            st = ctx.render_device(p, pixels.data_ptr(), stats=True)    # stats: ray counter read-back
            paths += st.paths
            rays += st.rays
    e1.record(stream)
    if world > 1:
        dist.reduce(pixels, dst=0, op=dist.ReduceOp.SUM)                 # NCCL over NVLink: the one exchange
    if rank == 0:
        scale = 1.0 if tiles else sharding.reduce_scale(total_spp)
        ctx.resolve_device(pixels.data_ptr(), n, scale, rgb8.data_ptr())
    e2.record(stream)
    torch.cuda.synchronize()
    stats = torch.tensor([e0.elapsed_time(e1), e0.elapsed_time(e2), float(paths), float(rays)], dtype=torch.float64, device=dev)
    if world > 1:
        mx = stats.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
        render_ms, total_ms = float(mx[0]), float(mx[1])                 # time = max over ranks
    else:
        render_ms, total_ms = float(stats[0]), float(stats[1])
    all_paths, all_rays = float(stats[2]), float(stats[3])
    rc = 0
    if rank == 0:
        out = opts["out"] or s.img_name.decode()
        host = rgb8.cpu().numpy().reshape(H, W, 3)
        rc = rt.lib().rt_write_ppm(out.encode(), host.ctypes.data, W, H)
        print("cuda_calc launch error: no error")
        print("cuda_calc sync error: no error")
        print("Rendering time: %fms" % render_ms)
        print("Rendering time: %fs" % (render_ms / 1e3))
        sys.stderr.write(json.dumps({
            "scene": opts["scene"], "width": W, "height": H, "spp": total_spp, "bounces": s.max_bounces, "gpus": world,
            "shard": opts["shard"] if world > 1 else "none", "traversal": int(st.traversal_used) if calls else opts["traversal"],
            "render_ms": round(render_ms, 3), "render_reduce_resolve_ms": round(total_ms, 3),
            "wall_s": round(time.perf_counter() - t_wall, 3), "paths": all_paths, "rays": all_rays,
            "paths_per_s": all_paths / (total_ms / 1e3), "mrays_per_s": all_rays / (total_ms / 1e3) / 1e6, "output": out}) + "\n")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    hs.close()
    ctx.close()
    return 0 if rc == 0 else 2


if __name__ == "__main__":
    sys.exit(main())
