#!/usr/bin/env python3
"""bench.py -- camera paths/s of the path-tracing hot path on N B200s of one node (BASELINE.json metric).

Workload (config.workload): BASELINE.json configs[3] = `./raytracing 1 s1024 b50` -- the Cornell box with emissive quad
(cornell_box_scene, rt.cu:2113) at its built-in 600x600, 1024 samples/pixel, bounce depth 50: the largest configuration
that BASELINE.json names for ONE GPU and one of the two "high-spp scenes" north_star's scaling target is about.
A step = one full render of that image by the whole job.  N>1: STRONG scaling -- the 1024 samples of every pixel are
split by sample index (rank g renders 1024/N of them from XORWOW subsequence pixel + g*n_pixels; rank 0 carries the
reference's stream), the f32 sum images are summed onto rank 0 with one NCCL reduce over NVLink, rank 0 resolves
(1/spp scale + gamma + 8 bit).

  value    paths/s of the whole job with scene, BVH / program and RNG states resident in HBM (the region the reference
           times, rt.cu:1752-1767, plus our reduce + resolve), CUDA events on the launching stream, max over ranks.
  e2e      the same metric through the reference-facing C ABI with HOST buffers: rt_render_world() = scene upload (H2D),
           RNG init, render, resolve, image download (D2H) -- what render(World*, RenderSettings*) does.
  configs  (N=1) every BASELINE.json config at full size (C5 at its built-in 800x800 and 1/256 of its samples) measured in
           this process, the reference's own CUDA binary timed beside each on the same GPU.
  --impl reference   the reference's own __host__ __device__ code on this box's host cores (oracle/_ref/libref_host.so,
           else the C oracle port), all threads, bounded sample of the same workload.
"""
import argparse
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SCENE_ID, SPP, BOUNCES = 1, 1024, 50
WORKLOAD = "C4: ./raytracing 1 s1024 b50 (cornell_box_scene, 600x600, emissive quad, 2 rotated+translated boxes)"
METRIC, UNIT = "camera_paths_per_second", "paths/s"
FP32_PEAK_NOTE = "148 SMs x 128 FP32 lanes x 2 flop x sm_max_mhz (MEASURED_PEAKS.json has no FP32 figure)"
# FP32 operations per unit of work, from the reference's formulas (FMA = 2)
F_CONST = dict(ray=3, box=25, sphere=40, planar=58, translation=3, rotation=12, lambert=45, metal=80, glass=100,
               noise=1430, checker=10, arena=10)
# BASELINE.json configs at full size: name -> (scene, spp, bounces, label)
CONFIGS = {
    "C1": (0, 16, 8, "default scene 800x450 s16 b8"),
    "C2": (2, 256, 50, "many random spheres 1200x675 s256 b50"),
    "C3": (17, 512, 50, "Perlin spheres 400x400 s512 b50"),
    "C3b": (19, 512, 50, "Perlin spheres + emissive quad/sphere 1080x607 s512 b50"),
    "C4": (1, 1024, 50, "Cornell box 600x600 s1024 b50"),
    "C5r": (3, 16, 50, "final scene 800x800 s16 b50 (C5 at 1/23 of the pixels and 1/256 of the samples)"),
}
KERNEL_MODES = {0: "reference order, stack walk", 1: "own SAH BVH", 2: "reference order + treelets + fused instances",
                3: "reference order as a linear program (stackless)", 4: "linear program, short-lived pixels",
                5: "tiny scene: every primitive, no BVH", 6: "own SAH BVH with instanced leaves"}


def host_threads():
    """Threads this process may run on.  NOT omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


def kernel_source_hash():
    """sha256 over the device sources: ties a committed ncu capture to the kernels it was taken from."""
    h = hashlib.sha256()
    d = os.path.join(ROOT, "cuda_raytracing_b200", "csrc", "device")
    for name in sorted(os.listdir(d)):
        if name.endswith((".cu", ".cuh", ".h", ".cpp")):
            h.update(name.encode())
            with open(os.path.join(d, name), "rb") as f:
                h.update(f.read())
    return h.hexdigest()[:16]


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons while the timed region runs (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=6)
        sm = [float(s[0]) for s in self.samples if s and s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if len(s) > 1 and s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for s in self.samples for i in range(4) if len(s) > 2 + i and s[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------ CPU reference arm
def cpu_reference_render(spp, nthreads=0, scene_id=SCENE_ID, bounces=BOUNCES):
    """One bounded sample of the workload on the host cores.  Returns (paths, seconds, kind, cores)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refabi
    import cuda_raytracing_b200 as rt
    cores = nthreads or host_threads()
    if refabi.available():
        s = refabi.build_scene(scene_id)
        cam = refabi.camera(s, 0, spp, bounces)
        t0 = time.perf_counter()
        refabi.render_host(s, cam, nthreads=cores)
        dt = time.perf_counter() - t0
        paths = cam.img_width * cam.img_height * spp
        refabi.lib().ref_scene_free(C.byref(s))
        return paths, dt, "reference", cores
    import oracleabi
    hs = rt.HostScene(scene_id)
    h, w, t = hs.arrays()
    sc = oracleabi.Scene(h, w, t, hs.world.root_wref)
    p = hs.params(0, spp, bounces)
    cam = oracleabi.OrcCamera()
    for f in ("img_width", "img_height", "samples_per_pixel", "max_bounces", "defocus_angle"):
        setattr(cam, f, getattr(p, f))
    for f in ("background", "pixel_delta_v", "pixel_delta_u", "camera_pos", "defocus_disk_u", "defocus_disk_v", "pixel00"):
        getattr(cam, f)[:] = getattr(p, f)[:]
    cam.pixel_samples_scale = 1.0 / spp
    t0 = time.perf_counter()
    oracleabi.render(sc, cam, nthreads=cores)
    dt = time.perf_counter() - t0
    return p.img_width * p.img_height * spp, dt, "port", cores


CPU_SAMPLE_SPP = 16           # of the 1024: paths/s does not depend on spp (per-pixel XORWOW init amortised as in the full run)


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    spp = CPU_SAMPLE_SPP
    for _ in range(args.warmup):
        cpu_reference_render(1)
    paths = secs = 0.0
    kind, cores = "port", 1
    for _ in range(args.steps):
        p, dt, kind, cores = cpu_reference_render(spp)
        paths += p; secs += dt
    value = paths / secs
    sample = "%d steps x (600x600, b50, s%d) = %.2f Mpaths per step on %d host threads" % (args.steps, spp, 600 * 600 * spp / 1e6, cores)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "reference built-in scene (deterministic, glibc rand seed 1)",
            "config": {"workload": WORKLOAD, "sample": "s%d of s%d per step" % (spp, SPP)},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    emit(line)
    return 0


def reference_gpu_binary(scene_id, spp, bounces, reps=2):
    """The reference's own CUDA binary on this GPU (its printed kernel time), best of `reps`; None when it did not travel."""
    exe = os.path.join(ROOT, "oracle", "_ref", "raytracing_ref")
    if not os.path.exists(exe):
        return None
    best = None
    try:
        with tempfile.TemporaryDirectory() as d:
            for _ in range(reps):
                out = subprocess.run([exe, str(scene_id), "s%d" % spp, "b%d" % bounces], cwd=d, capture_output=True, text=True, timeout=600).stdout
                for ln in out.splitlines():
                    if ln.startswith("Rendering time:") and ln.endswith("ms"):
                        ms = float(ln.split()[2][:-2])
                        best = ms if best is None else min(best, ms)
    except Exception:  # noqa: BLE001
        return None
    return best


def roofline_of(rt, ctx, hs, spp_probe, bounces, kernel_ms, paths, fp32_peak):
    """Algorithmic FP32 work of one launch / its measured duration.  Per-ray operation counts come from the instrumented
    build of the same kernel (RT_RENDER_COUNT_OPS), run outside any timed region."""
    small = hs.params(0, spp_probe, bounces, flags=rt.RT_RENDER_COUNT_OPS)
    _, _, st = ctx.render(small, want_float=False, want_rgb8=False)
    ops = dict(zip(rt.OP_NAMES, [int(x) for x in st.ops]))
    per_ray = {k: v / max(st.rays, 1) for k, v in ops.items()}
    f_ray = F_CONST["ray"] + sum(F_CONST[k] * per_ray[k] for k in ("box", "sphere", "planar", "translation", "rotation", "lambert", "metal",
                                                                    "glass", "noise", "checker", "arena"))
    rays_per_path = st.rays / max(st.paths, 1)
    rays = rays_per_path * paths
    achieved = rays * f_ray / (kernel_ms / 1e3) / 1e12
    return {"achieved": achieved, "frac": achieved / fp32_peak, "flops_per_ray": f_ray, "rays_per_path": rays_per_path,
            "mrays_per_s": rays / (kernel_ms / 1e3) / 1e6, "per_ray_ops": per_ray}


def measure_configs(rt, ctx, fp32_peak, skip_reference_gpu):
    """Every BASELINE.json config, exact mode (RT_TRAVERSAL_AUTO), on this GPU; the reference's CUDA binary beside each."""
    out = {}
    for name, (sid, spp, b, label) in CONFIGS.items():
        hs = rt.HostScene(sid)
        ctx.upload(hs, rt.RT_TRAVERSAL_AUTO)
        p = hs.params(0, spp, b, flags=0)
        times = []
        st = None
        for _ in range(2 + 3):                                   # 2 warm-up renders, best of 3
            _, _, st = ctx.render(p, want_float=False, want_rgb8=False)
            times.append(st.render_ms)
        ms = min(times[2:])
        paths = p.img_width * p.img_height * spp
        roof = roofline_of(rt, ctx, hs, min(spp, 8), b, ms, paths, fp32_peak)
        mode = int(hs.pack_info().kernel_mode)
        ref_ms = None if skip_reference_gpu else reference_gpu_binary(sid, spp, b)
        out[name] = {"invocation": "./raytracing %d s%d b%d" % (sid, spp, b), "what": label, "width": p.img_width, "height": p.img_height,
                     "ours_ms": ms, "paths_per_s": paths / (ms / 1e3), "mrays_per_s": roof["mrays_per_s"], "rays_per_path": roof["rays_per_path"],
                     "kernel_mode": mode, "walk": KERNEL_MODES.get(mode), "regs": int(st.regs_per_thread), "blocks_per_sm": int(st.blocks_per_sm),
                     "reference_gpu_ms": ref_ms, "x_vs_reference_gpu": (ref_ms / ms) if ref_ms else None,
                     "flops_per_ray": roof["flops_per_ray"], "roofline_frac": roof["frac"]}
        hs.close()
    return out


# ------------------------------------------------------------------------------------------------ the GPU arm
def run_gpu_arm(args):
    import torch
    import torch.distributed as dist
    import cuda_raytracing_b200 as rt
    from cuda_raytracing_b200 import sharding

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)

    ctx = rt.Context(local)
    stream = torch.cuda.Stream(device=dev)                              # kernels, NCCL and events all on this stream
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    hs = rt.HostScene(SCENE_ID)
    ctx.upload(hs, rt.RT_TRAVERSAL_AUTO)
    base = hs.params(0, SPP, BOUNCES)
    W, H = base.img_width, base.img_height
    n = W * H
    offset, spp_local, spp_total = sharding.sample_shard(rank, world, n, total_spp=SPP)      # strong scaling: SPP is the job's
    p = hs.params(0, spp_local, BOUNCES, subsequence_offset=offset, flags=rt.RT_RENDER_OUTPUT_SUM)
    scale = sharding.reduce_scale(spp_total)

    pixels = torch.zeros(n * 3, dtype=torch.float32, device=dev)
    rgb8 = torch.zeros(n * 3, dtype=torch.uint8, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)       # > 126 MB L2

    def step():
        ctx.render_device(p, pixels.data_ptr(), stats=False)            # kernel 1: persistent path tracer
        if world > 1:
            dist.reduce(pixels, dst=0, op=dist.ReduceOp.SUM)            # NCCL over NVLink
        if rank == 0:
            ctx.resolve_device(pixels.data_ptr(), n, scale, rgb8.data_ptr())   # kernel 2: scale + gamma + quantise

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    warmup = max(args.warmup, 3)
    for _ in range(warmup):
        step()
    sync_all()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    total_ms = 0.0
    render_ms, tail_ms = [], []
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    for _ in range(args.steps):
        flush.fill_(1)                                                  # L2 flush between timed iterations
        sync_all()
        e0.record(stream)
        ctx.render_device(p, pixels.data_ptr(), stats=False)
        e1.record(stream)
        if world > 1:
            dist.reduce(pixels, dst=0, op=dist.ReduceOp.SUM)
        if rank == 0:
            ctx.resolve_device(pixels.data_ptr(), n, scale, rgb8.data_ptr())
        e2.record(stream)
        sync_all()
        total_ms += e0.elapsed_time(e2)
        render_ms.append(e0.elapsed_time(e1))
        tail_ms.append(e1.elapsed_time(e2))
    clocks = sampler.summary() if sampler else None
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    mine = torch.tensor([float(np.mean(render_ms)), float(np.mean(tail_ms))], dtype=torch.float64, device=dev)
    per_rank = [mine.clone() for _ in range(world)]
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_gather(per_rank, mine)
    total_ms = float(t.item())
    per_rank = [[float(x[0]), float(x[1])] for x in per_rank]
    paths_per_step = n * spp_total
    value = paths_per_step * args.steps / (total_ms / 1e3)

    # ---- end to end through the C ABI with host buffers (scene H2D + RNG init + render + resolve + image D2H) ----
    ctx.set_stream(None)
    s_e2e, _cam = hs.camera(0, spp_local, BOUNCES)
    e2e_ms = 0.0
    h2d = hs.world.hittable_count * 120 + hs.world.world_count * 2 + hs.world.texture_bytes
    d2h = n * 12 + n * 3
    for i in range(warmup + args.steps):
        sync_all()
        t0 = time.perf_counter()
        if world == 1:
            ctx.render_world(hs, s_e2e, rt.RT_TRAVERSAL_AUTO)
        else:                                                           # every rank: upload + render its shard; reduce; rank 0 downloads
            ctx.upload(hs, rt.RT_TRAVERSAL_AUTO)
            q = hs.params(0, spp_local, BOUNCES, subsequence_offset=offset, flags=rt.RT_RENDER_OUTPUT_SUM | rt.RT_RENDER_NO_RNG_CACHE)
            ctx.set_stream(stream.cuda_stream)
            ctx.render_device(q, pixels.data_ptr(), stats=False)
            dist.reduce(pixels, dst=0, op=dist.ReduceOp.SUM)
            if rank == 0:
                ctx.resolve_device(pixels.data_ptr(), n, scale, rgb8.data_ptr())
                pixels.cpu(); rgb8.cpu()
            ctx.set_stream(None)
        sync_all()
        if i >= warmup:
            e2e_ms += (time.perf_counter() - t0) * 1e3
    te = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = paths_per_step * args.steps / (float(te.item()) / 1e3)

    if rank != 0:
        if world > 1:
            dist.barrier(); dist.destroy_process_group()
        return 0

    # ---- roofline of the dominant kernel (render_kernel): algorithmic FP32 work / measured duration ----
    peaks, peak_src = measured_peaks()
    ctx.set_stream(None)
    fp32_peak = 148 * 128 * 2 * peaks.get("sm_max_mhz", 1965.0) * 1e6 / 1e12
    kernel_ms = float(np.mean(render_ms))
    roof = roofline_of(rt, ctx, hs, 8, BOUNCES, kernel_ms, n * spp_local, fp32_peak)
    info = hs.pack_info()
    st_probe = ctx.render_device(p, pixels.data_ptr(), stats=True)      # one more step, untimed: launch shape of the timed kernel
    kernel_name = "render_kernel<mode %d: %s; %d blocks/SM x 128 threads x %d registers>" % (
        info.kernel_mode, KERNEL_MODES.get(int(info.kernel_mode), "?"), st_probe.blocks_per_sm, st_probe.regs_per_thread)
    # the HBM side of the path: accumulate/resolve buffers (12 B written by the render, 12 B read + 3 B written by resolve)
    e3, e4 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx.set_stream(stream.cuda_stream)
    reps = 20
    flush.fill_(1); torch.cuda.synchronize()
    e3.record(stream)
    for _ in range(reps):
        ctx.resolve_device(pixels.data_ptr(), n, scale, rgb8.data_ptr())
    e4.record(stream); torch.cuda.synchronize()
    resolve_ms = e3.elapsed_time(e4) / reps
    hbm = {"kernel": "resolve_kernel", "bytes_per_launch": n * 15, "ms": resolve_ms, "achieved_gbs": n * 15 / (resolve_ms / 1e3) / 1e9,
           "peak_gbs": peaks.get("hbm_gbs"), "note": "4.3 MB working set is L2-resident when timed back to back; HBM is <0.1% of the step by design"}
    ctx.set_stream(None)

    # DRAM bytes of one launch etc. come from a COMMITTED ncu capture of this command (a run under ncu is never a bench
    # value); they are only reported while the device sources still hash to what the capture was taken from
    traffic, ncu_note = None, None
    src_hash = kernel_source_hash()
    try:
        with open(os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")) as f:
            tr = json.load(f)
        if WORKLOAD.startswith(tr.get("workload", "?")) and tr.get("kernel_source_sha16") == src_hash:
            traffic = int(tr["dram_bytes_read"]) + int(tr["dram_bytes_write"])
            ncu_note = {k: tr[k] for k in tr if k not in ("dram_bytes_read", "dram_bytes_write")}
            ncu_note["source_of_these_fields"] = "committed capture profiles/r2_ncu_traffic.json (not measured in this run)"
        else:
            ncu_note = {"stale": "profiles/r2_ncu_traffic.json was captured from other device sources (%s, now %s): not reported"
                        % (tr.get("kernel_source_sha16"), src_hash)}
    except (OSError, ValueError, KeyError):
        traffic, ncu_note = None, None

    cpu = None
    configs = None
    if world == 1:
        paths, dt, kind, cores = cpu_reference_render(CPU_SAMPLE_SPP)
        cpu = {"value": paths / dt, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": "600x600, b50, s%d (%.2f Mpaths) of the s1024 workload, OpenMP over rows, %.1f s wall" % (CPU_SAMPLE_SPP, paths / 1e6, dt)}
        if not args.no_configs:
            ctx.set_stream(None)
            configs = measure_configs(rt, ctx, fp32_peak, args.no_reference_gpu)
    renders = [r[0] for r in per_rank]
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warmup,
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "reference built-in scene (deterministic: glibc rand() seed 1 geometry, XORWOW seed 1234 per-pixel streams)",
        "config": {"workload": WORKLOAD, "width": W, "height": H, "total_spp": spp_total, "spp_per_gpu": [sharding.sample_shard(g, world, n, total_spp=SPP)[1] for g in range(world)],
                   "max_bounces": BOUNCES,
                   "parallelism": "sample-sharded x%d (strong: the job's 1024 spp split by sample index) + NCCL reduce" % world if world > 1 else "single GPU",
                   "l2": "flushed between timed iterations (256 MiB write)", "rng_stream": "reference (subsequence = pixel + rank*n)",
                   "traversal": "auto (exact): the reference's visiting order as a stackless linear program, bit-identical to the reference's GPU render at N=1"},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "what": "rt_render_world(): scene H2D + XORWOW init (recomputed every call) + render + resolve + f32/rgb8 image D2H"},
        "gpu_launches": (world + 1) * args.steps,
        "per_rank": {"render_ms": renders, "render_ms_min": min(renders), "render_ms_max": max(renders),
                     "reduce_resolve_ms_rank0": per_rank[0][1],
                     "note": "step = max over ranks of (render + wait for the slowest rank inside the reduce + reduce + resolve); "
                             "reduce_resolve_ms_rank0 includes rank 0's wait for the slowest render"},
        "roofline": {"bound": "fp32", "achieved": roof["achieved"], "peak": fp32_peak, "unit": "TFLOP/s", "frac": roof["frac"], "traffic": traffic,
                     "kernel": kernel_name, "kernel_ms": kernel_ms, "flops_per_ray": roof["flops_per_ray"], "rays_per_path": roof["rays_per_path"],
                     "mrays_per_s": roof["mrays_per_s"], "per_ray_ops": roof["per_ray_ops"], "peak_source": FP32_PEAK_NOTE + "; " + peak_src,
                     "note": "SURVEY.md 8(d): this path is bound by the FP32 pipe / issue rate, not HBM or tensor cores", "hbm": hbm,
                     "ncu": ncu_note, "kernel_source_sha16": src_hash},
        "cpu_baseline": cpu,
    }
    if configs is not None:
        line["configs"] = configs
        c4 = configs.get("C4") or {}
        line["reference_gpu"] = {"kernel_ms": c4.get("reference_gpu_ms"),
                                 "value": (n * SPP / (c4["reference_gpu_ms"] / 1e3)) if c4.get("reference_gpu_ms") else None, "unit": UNIT,
                                 "what": "oracle/_ref/raytracing_ref (unmodified reference, nvcc -arch sm_100 -O3), its own 'Rendering time', same GPU, same run"}
    emit(line)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()
    return 0


def claim_stdout():
    """stdout carries ONE JSON line.  Anything a library prints to fd 1 (NCCL's version line at NCCL_DEBUG=VERSION, for one)
    goes to stderr instead: fd 1 is pointed at fd 2 for the whole run and the line is written to the saved descriptor."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    return os.fdopen(real, "w")


OUT = None


def emit(line):
    (OUT or sys.stdout).write(json.dumps(line) + "\n")
    (OUT or sys.stdout).flush()


def main():
    global OUT
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-reference-gpu", action="store_true", help="skip timing the reference's own CUDA binary")
    ap.add_argument("--no-configs", action="store_true", help="skip the per-config block (N=1)")
    args = ap.parse_args()
    OUT = claim_stdout()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
