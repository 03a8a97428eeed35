#!/usr/bin/env python3
"""bench.py -- camera paths/s of the path-tracing hot path on N B200s of one node (BASELINE.json metric).

Workload (config.workload): BASELINE.json configs[1] = `./raytracing 2 s256 b50` -- the many-random-spheres scene
(bouncing_spheres_scene, rt.cu:1979) at its built-in 1200x675, 256 samples/pixel, bounce depth 50.
A step = one full render of that image per GPU.  N>1: sample sharding, weak scaling -- every rank renders its own
256 samples/pixel (XORWOW subsequence = pixel + rank*n_pixels; rank 0 is the reference stream), the f32 sum
buffers are summed onto rank 0 with one NCCL reduce over NVLink, rank 0 resolves (scale + gamma + 8 bit).

  value   paths/s with scene, BVH and RNG states resident in HBM (the region the reference times: rt.cu:1752-1767,
          plus our reduce + resolve), CUDA events on the launching stream, max over ranks.
  e2e     the same metric through the reference-facing C ABI with HOST buffers: rt_render_world() = scene upload
          (H2D), RNG init, render, resolve, image download (D2H) -- what render(World*, RenderSettings*) does.
  --impl reference   the reference's own __host__ __device__ code on this box's host cores (oracle/_ref/libref_host.so,
          else the C oracle port), all threads, bounded sample of the same workload.
"""
import argparse
import ctypes as C
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

SCENE_ID, SPP, BOUNCES = 2, 256, 50
WORKLOAD = "C2: ./raytracing 2 s256 b50 (bouncing_spheres_scene, 1200x675, 485 spheres)"
METRIC, UNIT = "camera_paths_per_second", "paths/s"
FP32_PEAK_NOTE = "148 SMs x 128 FP32 lanes x 2 flop x sm_max_mhz (MEASURED_PEAKS.json has no FP32 figure)"
# FP32 operations per unit of work, from the reference's formulas (FMA = 2)
F_CONST = dict(ray=3, box=25, sphere=40, planar=58, translation=3, rotation=12, lambert=45, metal=80, glass=100,
               noise=1430, checker=10, arena=10)


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


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
def cpu_reference_render(spp, nthreads=0):
    """One bounded sample of the workload on the host cores.  Returns (paths, seconds, kind, cores)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refabi
    import cuda_raytracing_b200 as rt
    if refabi.available():
        s = refabi.build_scene(SCENE_ID)
        cam = refabi.camera(s, 0, spp, BOUNCES)
        cores = nthreads or refabi.lib().ref_max_threads()
        t0 = time.perf_counter()
        refabi.render_host(s, cam, nthreads=cores)
        dt = time.perf_counter() - t0
        paths = cam.img_width * cam.img_height * spp
        refabi.lib().ref_scene_free(C.byref(s))
        return paths, dt, "reference", cores
    import oracleabi
    hs = rt.HostScene(SCENE_ID)
    h, w, t = hs.arrays()
    sc = oracleabi.Scene(h, w, t, hs.world.root_wref)
    p = hs.params(0, spp, BOUNCES)
    cam = oracleabi.OrcCamera()
    for f in ("img_width", "img_height", "samples_per_pixel", "max_bounces", "defocus_angle"):
        setattr(cam, f, getattr(p, f))
    for f in ("background", "pixel_delta_v", "pixel_delta_u", "camera_pos", "defocus_disk_u", "defocus_disk_v", "pixel00"):
        getattr(cam, f)[:] = getattr(p, f)[:]
    cam.pixel_samples_scale = 1.0 / spp
    cores = nthreads or oracleabi.lib().orc_max_threads()
    t0 = time.perf_counter()
    oracleabi.render(sc, cam, nthreads=cores)
    dt = time.perf_counter() - t0
    return p.img_width * p.img_height * spp, dt, "port", cores


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    spp = 16                      # a bounded sample per step; the per-pixel XORWOW init of the reference is amortised like in the full run
    for _ in range(args.warmup):
        cpu_reference_render(1)
    paths = secs = 0.0
    kind, cores = "port", 1
    for _ in range(args.steps):
        p, dt, kind, cores = cpu_reference_render(spp)
        paths += p; secs += dt
    value = paths / secs
    sample = "%d steps x (1200x675, b50, s%d) = %.2f Mpaths per step on %d host threads" % (args.steps, spp, 1200 * 675 * spp / 1e6, cores)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "reference built-in scene (deterministic, glibc rand seed 1)",
            "config": {"workload": WORKLOAD, "sample": "s%d of s%d per step" % (spp, SPP)},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def reference_gpu_binary():
    """The reference's own CUDA binary on this GPU (its printed kernel time), once; None when it did not travel."""
    exe = os.path.join(ROOT, "oracle", "_ref", "raytracing_ref")
    if not os.path.exists(exe):
        return None
    try:
        best = None
        with tempfile.TemporaryDirectory() as d:
            for _ in range(2):
                out = subprocess.run([exe, str(SCENE_ID), "s%d" % SPP, "b%d" % BOUNCES], cwd=d, capture_output=True, text=True, timeout=300).stdout
                for ln in out.splitlines():
                    if ln.startswith("Rendering time:") and ln.endswith("ms"):
                        ms = float(ln.split()[2][:-2])
                        best = ms if best is None else min(best, ms)
        if best is None:
            return None
        return {"kernel_ms": best, "value": 1200 * 675 * SPP / (best / 1e3), "unit": UNIT,
                "what": "oracle/_ref/raytracing_ref (unmodified reference, nvcc -arch sm_100 -O3), its own 'Rendering time'"}
    except Exception as e:  # noqa: BLE001
        return {"error": str(e)}


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
    offset, spp_local, spp_total = sharding.sample_shard(rank, world, n, per_rank_spp=SPP)
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

    for _ in range(max(args.warmup, 3)):
        step()
    sync_all()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    total_ms = 0.0
    render_ms = []
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
    clocks = sampler.summary() if sampler else None
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    paths_per_step = n * SPP * world
    value = paths_per_step * args.steps / (total_ms / 1e3)

    # ---- end to end through the C ABI with host buffers (scene H2D + RNG init + render + resolve + image D2H) ----
    ctx.set_stream(None)
    s_e2e, _cam = hs.camera(0, spp_local, BOUNCES)
    e2e_ms = 0.0
    h2d = hs.world.hittable_count * 120 + hs.world.world_count * 2 + hs.world.texture_bytes
    d2h = n * 12 + n * 3
    for i in range(args.warmup + args.steps):
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
        if i >= args.warmup:
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
    small = hs.params(0, 8, BOUNCES, flags=rt.RT_RENDER_COUNT_OPS)       # instrumented variant, outside the timed region
    _, _, st = ctx.render(small, want_float=False, want_rgb8=False)
    ops = dict(zip(rt.OP_NAMES, [int(x) for x in st.ops]))
    per_ray = {k: v / st.rays for k, v in ops.items()}
    f_ray = F_CONST["ray"] + sum(F_CONST[k] * per_ray[k] for k in ("box", "sphere", "planar", "translation", "rotation", "lambert", "metal",
                                                                    "glass", "noise", "checker", "arena"))
    rays_per_path = st.rays / st.paths
    kernel_ms = float(np.mean(render_ms))
    rays_per_launch = rays_per_path * n * SPP
    achieved = rays_per_launch * f_ray / (kernel_ms / 1e3) / 1e12
    fp32_peak = 148 * 128 * 2 * peaks.get("sm_max_mhz", 1965.0) * 1e6 / 1e12
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
           "peak_gbs": peaks.get("hbm_gbs"), "note": "12 MB working set is L2-resident when timed back to back; HBM is <0.1% of the step by design"}

    traffic, ncu_note = None, None                   # DRAM bytes of one launch etc. from the committed ncu capture of this command
    try:
        with open(os.path.join(ROOT, "profiles", "r1_ncu_traffic.json")) as f:
            tr = json.load(f)
        if WORKLOAD.startswith(tr.get("workload", "?")):
            traffic = int(tr["dram_bytes_read"]) + int(tr["dram_bytes_write"])
            ncu_note = {k: tr[k] for k in ("issue_slot_frac", "lanes_per_warp_instruction", "warp_instructions_per_ray", "l1_hit_pct",
                                            "l2_hit_pct", "fma_pipe_pct", "source") if k in tr}
    except (OSError, ValueError, KeyError):
        traffic, ncu_note = None, None
    cpu = None
    if world == 1:
        paths, dt, kind, cores = cpu_reference_render(16)
        cpu = {"value": paths / dt, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": "1200x675, b50, s16 (12.96 Mpaths) of the s256 workload, OpenMP over rows, %.1f s wall" % dt}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "reference built-in scene (deterministic: glibc rand() seed 1 geometry, XORWOW seed 1234 per-pixel streams)",
        "config": {"workload": WORKLOAD, "width": W, "height": H, "spp_per_gpu": SPP, "total_spp": spp_total, "max_bounces": BOUNCES,
                   "parallelism": "sample-sharded x%d + NCCL reduce" % world if world > 1 else "single GPU",
                   "l2": "flushed between timed iterations (256 MiB write)", "rng_stream": "reference (subsequence = pixel + rank*n)",
                   "traversal": "auto (exact): SAH BVH, results bit-identical to the reference's GPU render"},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "what": "rt_render_world(): scene H2D + RNG init + render + resolve + f32/rgb8 image D2H"},
        "gpu_launches": 2 * args.steps,
        "roofline": {"bound": "fp32", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved / fp32_peak, "traffic": traffic,
                     "kernel": "render_kernel<fast,64>", "kernel_ms": kernel_ms, "flops_per_ray": f_ray, "rays_per_path": rays_per_path,
                     "mrays_per_s": rays_per_launch / (kernel_ms / 1e3) / 1e6, "per_ray_ops": per_ray, "peak_source": FP32_PEAK_NOTE + "; " + peak_src,
                     "note": "SURVEY.md 8(d): this path is bound by the FP32 pipe / issue rate, not HBM or tensor cores", "hbm": hbm,
                     "ncu": ncu_note},
        "cpu_baseline": cpu,
    }
    if world == 1 and not args.no_reference_gpu:
        line["reference_gpu"] = reference_gpu_binary()
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-reference-gpu", action="store_true", help="skip timing the reference's own CUDA binary")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
