"""ctypes mirror of the C ABI (include/rt_api.h, include/rt_scene.h) of libraytracing_b200.so.

Host-side plumbing only: the path tracer itself is the CUDA library.  There is no CPU fallback --
loading fails loudly when the library is missing, and Context() fails when there is no GPU.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RT_B200_LIB") or os.path.join(_HERE, "libraytracing_b200.so")   # env: kernel-variant experiments

RT_TRAVERSAL_AUTO, RT_TRAVERSAL_REFERENCE, RT_TRAVERSAL_FAST = 0, 1, 2
RT_RENDER_OUTPUT_SUM, RT_RENDER_NO_RNG_CACHE, RT_RENDER_COUNT_RAYS, RT_RENDER_COUNT_OPS, RT_RENDER_RESUME = 1, 2, 4, 8, 16
RT_RENDER_FORWARD_THROUGHPUT = 32
RT_STATIC_SCRATCH_MAX_PIXELS = 262144
RT_SHARD_TILES, RT_SHARD_SAMPLES = 0, 1
OP_NAMES = ["box", "sphere", "planar", "translation", "rotation", "lambert", "metal", "glass", "light", "noise", "checker", "arena"]


class RtError(RuntimeError):
    pass


class Vec3(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float)]


class Material(C.Structure):  # rt_material (24 B): type, pad, 16-byte union
    _fields_ = [("type", C.c_uint32), ("pad_", C.c_uint32), ("u", C.c_uint8 * 16)]


class Xform(C.Structure):  # rt_xform
    _fields_ = [("rotate", C.c_int), ("angle_deg", C.c_float), ("translate", C.c_int), ("offset", Vec3)]


class World(C.Structure):  # rt_world
    _fields_ = [("hittables", C.c_void_p), ("hittable_count", C.c_uint64), ("hittable_cap", C.c_uint64),
                ("world_refs", C.POINTER(C.c_uint16)), ("world_count", C.c_uint64), ("world_cap", C.c_uint64),
                ("texture_arena", C.c_void_p), ("texture_bytes", C.c_uint64), ("texture_cap", C.c_uint64),
                ("texture_count", C.c_uint64), ("root_wref", C.c_uint32), ("bvh_built", C.c_int)]


class RenderSettings(C.Structure):  # rt_render_settings == the reference's RenderSettings (rt.cu:1509-1521)
    _fields_ = [("img_ratio", C.c_float), ("img_width", C.c_int32), ("samples_per_pixel", C.c_int32),
                ("max_bounces", C.c_int32), ("vfov", C.c_float), ("lookfrom", Vec3), ("lookat", Vec3),
                ("view_up", Vec3), ("defocus_angle", C.c_float), ("background_color", Vec3),
                ("img_name", C.c_char_p)]


class Camera(C.Structure):  # rt_camera
    _fields_ = [("img_width", C.c_int32), ("img_height", C.c_int32), ("pixel_delta_v", Vec3), ("pixel_delta_u", Vec3),
                ("camera_pos", Vec3), ("defocus_disk_u", Vec3), ("defocus_disk_v", Vec3), ("pixel00", Vec3),
                ("pixel_samples_scale", C.c_float)]


class SceneDesc(C.Structure):  # rt_scene_desc
    _fields_ = [("hittables", C.c_void_p), ("hittable_count", C.c_uint64), ("world_refs", C.c_void_p),
                ("world_count", C.c_uint64), ("root_wref", C.c_uint32), ("texture_arena", C.c_void_p),
                ("texture_bytes", C.c_uint64)]


class RenderParams(C.Structure):  # rt_render_params
    _fields_ = [("img_width", C.c_int32), ("img_height", C.c_int32), ("samples_per_pixel", C.c_int32),
                ("max_bounces", C.c_int32), ("defocus_angle", C.c_float), ("background", C.c_float * 3),
                ("pixel_delta_v", C.c_float * 3), ("pixel_delta_u", C.c_float * 3), ("camera_pos", C.c_float * 3),
                ("defocus_disk_u", C.c_float * 3), ("defocus_disk_v", C.c_float * 3), ("pixel00", C.c_float * 3),
                ("seed", C.c_uint64), ("subsequence_offset", C.c_uint64), ("row_begin", C.c_int32),
                ("row_end", C.c_int32), ("flags", C.c_uint32), ("band_rows", C.c_uint32), ("band_period", C.c_uint32),
                ("band_phase", C.c_uint32)]


class Stats(C.Structure):  # rt_stats
    _fields_ = [("rng_init_ms", C.c_double), ("render_ms", C.c_double), ("resolve_ms", C.c_double),
                ("h2d_ms", C.c_double), ("d2h_ms", C.c_double), ("total_ms", C.c_double),
                ("paths", C.c_uint64), ("rays", C.c_uint64), ("kernel_launches", C.c_uint32),
                ("traversal_used", C.c_uint32), ("grid_blocks", C.c_uint32), ("block_threads", C.c_uint32),
                ("regs_per_thread", C.c_uint32), ("blocks_per_sm", C.c_uint32), ("ops", C.c_uint64 * 12)]


class PackInfo(C.Structure):  # rt_pack_info
    _fields_ = [(n, C.c_uint32) for n in ("traversal_used", "kernel_mode", "spheres", "planars", "own_nodes", "treelets",
                                          "treelet_prims", "fused_instances", "stack_need", "stack_capacity", "perlin_tables",
                                          "has_wrappers", "program_ops", "program_boxes", "program_prims",
                                          "program_wrappers", "instanced_leaves", "program_axis_quads")]


class RenderPlan(C.Structure):  # rt_render_plan
    _fields_ = [("kernel_mode", C.c_uint32), ("resident_lanes", C.c_uint32), ("chunk_count", C.c_uint32),
                ("chunk_samples", C.c_int32), ("chunk_begin", C.c_int32 * 33), ("claim_group", C.c_uint32),
                ("scattered", C.c_uint32)]


class ProbeHit(C.Structure):  # rt_probe_hit
    _fields_ = [("origin", C.c_float * 3), ("direction", C.c_float * 3), ("point", C.c_float * 3),
                ("normal", C.c_float * 3), ("t", C.c_float), ("u", C.c_float), ("v", C.c_float),
                ("hit", C.c_int32), ("front", C.c_int32), ("material_type", C.c_int32)]


EXPORTS = [
    # rt_api.h
    "rt_context_create", "rt_context_destroy", "rt_last_error_string", "rt_context_set_stream", "rt_context_reset_stream", "rt_scene_upload",
    "rt_render_params_from_camera", "rt_render_device", "rt_render_device_static", "rt_resolve_device", "rt_render", "rt_synchronize",
    "rt_render_state_save", "rt_render_state_load",
    "rt_group_create", "rt_group_destroy", "rt_group_size", "rt_group_context", "rt_group_last_error", "rt_group_render_world",
    "rt_scene_pack_info", "rt_render_plan_query", "rt_scene_program", "rt_scene_linear_image", "rt_render_world", "rt_probe_trace", "rt_probe_scatter", "rt_probe_texture", "rt_probe_xorwow",
    # rt_scene.h
    "rt_world_init", "rt_world_free", "rt_mat_lambert_solid", "rt_mat_light_solid", "rt_mat_lambert_image",
    "rt_mat_lambert_image_pixels",
    "rt_mat_lambert_noise", "rt_mat_lambert_checker", "rt_mat_metal", "rt_mat_dielectric", "rt_add_sphere",
    "rt_add_moving_sphere", "rt_add_quad", "rt_add_triangle", "rt_add_box", "rt_add_pyramid", "rt_host_random",
    "rt_host_random_range", "rt_host_random_int", "rt_world_build_bvh", "rt_world_stack_bound",
    "rt_default_render_settings", "rt_scene_by_id", "rt_host_rand_reset", "rt_camera_setup", "rt_quantise_rgb8",
    "rt_write_ppm",
]

_lib = None


def lib():
    """Loads libraytracing_b200.so; raises if it has not been built (run `make` or __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RtError("%s is missing: build it with `make` (there is no CPU fallback)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.rt_last_error_string.restype = C.c_char_p
        L.rt_last_error_string.argtypes = [C.c_void_p]
        L.rt_context_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
        L.rt_context_destroy.argtypes = [C.c_void_p]
        L.rt_context_destroy.restype = None
        L.rt_context_set_stream.argtypes = [C.c_void_p, C.c_void_p]
        L.rt_context_reset_stream.argtypes = [C.c_void_p]
        L.rt_scene_upload.argtypes = [C.c_void_p, C.POINTER(SceneDesc), C.c_uint32]
        L.rt_render_params_from_camera.argtypes = [C.POINTER(RenderSettings), C.POINTER(Camera), C.POINTER(RenderParams)]
        L.rt_render_params_from_camera.restype = None
        L.rt_render_device.argtypes = [C.c_void_p, C.POINTER(RenderParams), C.c_void_p, C.POINTER(Stats)]
        L.rt_render_device_static.argtypes = [C.c_void_p, C.POINTER(RenderParams), C.c_void_p]
        L.rt_resolve_device.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_float, C.c_void_p]
        L.rt_render.argtypes = [C.c_void_p, C.POINTER(RenderParams), C.c_void_p, C.c_void_p, C.POINTER(Stats)]
        L.rt_synchronize.argtypes = [C.c_void_p]
        L.rt_scene_pack_info.argtypes = [C.POINTER(SceneDesc), C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(PackInfo)]
        L.rt_render_plan_query.argtypes = [C.POINTER(PackInfo), C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(RenderPlan)]
        L.rt_scene_program.argtypes = [C.POINTER(SceneDesc), C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32)]
        L.rt_scene_linear_image.argtypes = [C.POINTER(SceneDesc), C.c_uint32, C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32)]
        L.rt_render_state_save.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
        L.rt_render_state_load.argtypes = [C.c_void_p, C.POINTER(RenderParams), C.c_void_p, C.c_uint64]
        L.rt_render_world.argtypes = [C.c_void_p, C.POINTER(World), C.POINTER(RenderSettings), C.c_uint32, C.c_void_p,
                                      C.c_void_p, C.POINTER(Stats)]
        L.rt_group_create.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_void_p)]
        L.rt_group_destroy.argtypes = [C.c_void_p]
        L.rt_group_destroy.restype = None
        L.rt_group_size.argtypes = [C.c_void_p]
        L.rt_group_context.argtypes = [C.c_void_p, C.c_int]
        L.rt_group_context.restype = C.c_void_p
        L.rt_group_last_error.argtypes = [C.c_void_p]
        L.rt_group_last_error.restype = C.c_char_p
        L.rt_group_render_world.argtypes = [C.c_void_p, C.POINTER(World), C.POINTER(RenderSettings), C.c_uint32, C.c_uint32,
                                            C.c_void_p, C.c_void_p, C.POINTER(Stats)]
        L.rt_probe_trace.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float,
                                     C.c_void_p]
        L.rt_probe_scatter.argtypes = [C.c_void_p, C.c_uint64] + [C.c_void_p] * 8
        L.rt_probe_texture.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.rt_probe_xorwow.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p]
        L.rt_world_init.argtypes = [C.POINTER(World)]
        L.rt_world_free.argtypes = [C.POINTER(World)]
        L.rt_world_free.restype = None
        L.rt_world_build_bvh.argtypes = [C.POINTER(World)]
        L.rt_world_build_bvh.restype = C.c_uint32
        L.rt_mat_lambert_solid.argtypes = [C.POINTER(World), Vec3]
        L.rt_mat_lambert_solid.restype = Material
        L.rt_mat_lambert_image_pixels.argtypes = [C.POINTER(World), C.c_void_p, C.c_int32, C.c_int32]
        L.rt_mat_lambert_image_pixels.restype = Material
        L.rt_add_sphere.argtypes = [C.POINTER(World), Xform, Vec3, C.c_float, Material]
        L.rt_add_sphere.restype = C.c_uint32
        L.rt_add_quad.argtypes = [C.POINTER(World), Xform, Vec3, Vec3, Vec3, Material]
        L.rt_add_quad.restype = C.c_uint32
        L.rt_mat_metal.argtypes = [Vec3, C.c_float]
        L.rt_mat_metal.restype = Material
        L.rt_world_stack_bound.argtypes = [C.POINTER(World)]
        L.rt_world_stack_bound.restype = C.c_uint64
        L.rt_default_render_settings.argtypes = [C.POINTER(RenderSettings)]
        L.rt_default_render_settings.restype = None
        L.rt_scene_by_id.argtypes = [C.c_int, C.POINTER(World), C.POINTER(RenderSettings)]
        L.rt_host_rand_reset.restype = None
        L.rt_camera_setup.argtypes = [C.POINTER(RenderSettings), C.POINTER(Camera)]
        L.rt_camera_setup.restype = None
        L.rt_quantise_rgb8.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p]
        L.rt_quantise_rgb8.restype = None
        L.rt_write_ppm.argtypes = [C.c_char_p, C.c_void_p, C.c_uint32, C.c_uint32]
        _lib = L
    return _lib


def _check(rc, ctx=None):
    if rc != 0:
        msg = lib().rt_last_error_string(ctx)
        raise RtError("rt error %d: %s" % (rc, msg.decode() if msg else "?"))


class HostScene:
    """One of the 20 built-in scenes built by the C host layer (rt_scenes.c), BVH included."""

    def __init__(self, scene_id, build_bvh=True):
        L = lib()
        self.world = World()
        self.settings = RenderSettings()
        if L.rt_world_init(C.byref(self.world)) != 0:
            raise MemoryError("rt_world_init")
        L.rt_default_render_settings(C.byref(self.settings))
        L.rt_host_rand_reset()
        if scene_id is None:                              # an empty world for the add_* / mat_* helpers below
            self.scene_id = None
            return
        if L.rt_scene_by_id(scene_id, C.byref(self.world), C.byref(self.settings)) != 0:
            L.rt_world_free(C.byref(self.world))
            raise ValueError("Unknown scene %r" % (scene_id,))
        self.scene_id = scene_id
        if build_bvh:
            L.rt_world_build_bvh(C.byref(self.world))

    # a small part of the C scene-construction surface, for tests that need a world the 20 built-in scenes lack
    def mat_lambert_solid(self, rgb):
        return lib().rt_mat_lambert_solid(C.byref(self.world), Vec3(*[float(x) for x in rgb]))

    def mat_lambert_image_pixels(self, pixels):
        """pixels: float32 [height, width, 3], top row first (what stbi_loadf returns)."""
        px = np.ascontiguousarray(pixels, dtype=np.float32)
        return lib().rt_mat_lambert_image_pixels(C.byref(self.world), px.ctypes.data, px.shape[1], px.shape[0])

    def add_sphere(self, center, radius, material):
        return lib().rt_add_sphere(C.byref(self.world), Xform(), Vec3(*[float(x) for x in center]), float(radius), material)

    def add_quad(self, q, u, v, material):
        vec = lambda a: Vec3(*[float(x) for x in a])
        return lib().rt_add_quad(C.byref(self.world), Xform(), vec(q), vec(u), vec(v), material)

    def build_bvh(self):
        return lib().rt_world_build_bvh(C.byref(self.world))

    def arrays(self):
        w = self.world
        h = np.ctypeslib.as_array(C.cast(w.hittables, C.POINTER(C.c_uint8)), shape=(w.hittable_count, 120)).copy()
        r = np.ctypeslib.as_array(w.world_refs, shape=(w.world_count,)).copy()
        t = np.ctypeslib.as_array(C.cast(w.texture_arena, C.POINTER(C.c_uint8)), shape=(max(w.texture_bytes, 1),)).copy()
        return h, r, t[:w.texture_bytes]

    def desc(self):
        w = self.world
        return SceneDesc(w.hittables, w.hittable_count, C.cast(w.world_refs, C.c_void_p), w.world_count, w.root_wref,
                         w.texture_arena, w.texture_bytes)

    def camera(self, width=0, spp=0, bounces=0):
        """(settings copy, rt_camera) with optional overrides -- the CLI's wN / sN / bN."""
        s = RenderSettings.from_buffer_copy(bytes(self.settings))
        s.img_name = self.settings.img_name
        if width:
            s.img_width = width
        if spp:
            s.samples_per_pixel = spp
        if bounces:
            s.max_bounces = bounces
        cam = Camera()
        lib().rt_camera_setup(C.byref(s), C.byref(cam))
        return s, cam

    def params(self, width=0, spp=0, bounces=0, seed=1234, subsequence_offset=0, flags=RT_RENDER_COUNT_RAYS):
        s, cam = self.camera(width, spp, bounces)
        p = RenderParams()
        lib().rt_render_params_from_camera(C.byref(s), C.byref(cam), C.byref(p))
        p.seed = seed
        p.subsequence_offset = subsequence_offset
        p.flags = flags
        return p

    def pack_info(self, traversal=RT_TRAVERSAL_AUTO, treelet_min=8, fuse_min=8):
        """What rt_scene_upload would build (host only, no GPU needed)."""
        info = PackInfo()
        d = self.desc()
        rc = lib().rt_scene_pack_info(C.byref(d), traversal, treelet_min, fuse_min, C.byref(info))
        if rc != 0:
            raise RtError("rt_scene_pack_info: %d %s" % (rc, (lib().rt_last_error_string(None) or b"").decode()))
        return info

    def render_plan(self, width=0, spp=0, bounces=0, sm_count=148, traversal=RT_TRAVERSAL_AUTO, rows=0):
        """The scheduling decisions of a render of this scene (host only, no GPU needed)."""
        s, cam = self.camera(width, spp, bounces)
        info = self.pack_info(traversal)
        plan = RenderPlan()
        _check(lib().rt_render_plan_query(C.byref(info), cam.img_width, rows or cam.img_height, s.samples_per_pixel, s.max_bounces,
                                          sm_count, C.byref(plan)))
        return plan

    def program(self):
        """The linear program of the scene's reference-order walk (host only): uint32 view, shape (ops, 8); empty if none."""
        d = self.desc()
        n = C.c_uint32(0)
        _check(lib().rt_scene_program(C.byref(d), None, 0, C.byref(n)))
        ops = np.zeros((n.value, 8), dtype=np.uint32)
        if n.value:
            _check(lib().rt_scene_program(C.byref(d), ops.ctypes.data, n.value, C.byref(n)))
        return ops

    def linear_image(self, axis_quads=True):
        """The image of the linear program as the kernels walk it (host only): uint32 view, shape (16-byte words, 4)."""
        d = self.desc()
        n = C.c_uint32(0)
        _check(lib().rt_scene_linear_image(C.byref(d), 1 if axis_quads else 0, None, 0, C.byref(n)))
        words = np.zeros((n.value, 4), dtype=np.uint32)
        if n.value:
            _check(lib().rt_scene_linear_image(C.byref(d), 1 if axis_quads else 0, words.ctypes.data, n.value, C.byref(n)))
        return words

    def close(self):
        if self.world.hittables:
            lib().rt_world_free(C.byref(self.world))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Context:
    """rt_context: one GPU."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        _check(lib().rt_context_create(device, C.byref(self._h)))
        self.device = device

    def close(self):
        if self._h:
            lib().rt_context_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, stream_ptr):
        """Enqueue on this cudaStream_t (0 = the CUDA default stream); None = back to the context's own stream."""
        if stream_ptr is None:
            _check(lib().rt_context_reset_stream(self._h), self._h)
        else:
            _check(lib().rt_context_set_stream(self._h, C.c_void_p(stream_ptr)), self._h)

    def upload(self, scene, traversal=RT_TRAVERSAL_AUTO):
        d = scene.desc() if isinstance(scene, HostScene) else scene
        _check(lib().rt_scene_upload(self._h, C.byref(d), traversal), self._h)

    def render(self, params, want_float=True, want_rgb8=True):
        """Host-buffer render (H2D of nothing, D2H of the image): returns (float image | None, rgb8 | None, Stats)."""
        n = params.img_width * params.img_height
        f = np.empty((params.img_height, params.img_width, 3), dtype=np.float32) if want_float else None
        b = np.empty((params.img_height, params.img_width, 3), dtype=np.uint8) if want_rgb8 else None
        st = Stats()
        _check(lib().rt_render(self._h, C.byref(params), f.ctypes.data if want_float else None,
                               b.ctypes.data if want_rgb8 else None, C.byref(st)), self._h)
        assert n >= 0
        return f, b, st

    def render_device(self, params, device_ptr, stats=True):
        st = Stats() if stats else None
        _check(lib().rt_render_device(self._h, C.byref(params), C.c_void_p(device_ptr), C.byref(st) if stats else None),
               self._h)
        return st

    def render_device_static(self, params, device_ptr):
        """rt_render_device without any allocation: the kernel's scratch is the library's static device array
        (at most RT_STATIC_SCRATCH_MAX_PIXELS pixels per call; calls on one GPU must not overlap)."""
        _check(lib().rt_render_device_static(self._h, C.byref(params), C.c_void_p(device_ptr)), self._h)

    def resolve_device(self, pixels_ptr, n_pixels, scale, rgb8_ptr):
        _check(lib().rt_resolve_device(self._h, C.c_void_p(pixels_ptr), n_pixels, scale, C.c_void_p(rgb8_ptr)), self._h)

    def synchronize(self):
        _check(lib().rt_synchronize(self._h), self._h)

    def save_state(self):
        """XORWOW states of the last render (uint32[n, 6]) -- with the caller's sum image, all a resume needs."""
        need = C.c_uint64(0)
        _check(lib().rt_render_state_save(self._h, None, 0, C.byref(need)), self._h)
        out = np.empty((need.value // 24, 6), dtype=np.uint32)
        _check(lib().rt_render_state_save(self._h, out.ctypes.data, out.nbytes, C.byref(need)), self._h)
        return out

    def load_state(self, params, state):
        state = np.ascontiguousarray(state, dtype=np.uint32)
        _check(lib().rt_render_state_load(self._h, C.byref(params), state.ctypes.data, state.nbytes), self._h)

    def render_world(self, scene, settings=None, traversal=RT_TRAVERSAL_AUTO):
        """The reference's render(World*, RenderSettings*) in one call."""
        s = settings if settings is not None else scene.settings
        cam = Camera()
        lib().rt_camera_setup(C.byref(s), C.byref(cam))
        f = np.empty((cam.img_height, cam.img_width, 3), dtype=np.float32)
        b = np.empty((cam.img_height, cam.img_width, 3), dtype=np.uint8)
        st = Stats()
        _check(lib().rt_render_world(self._h, C.byref(scene.world), C.byref(s), traversal, f.ctypes.data, b.ctypes.data,
                                     C.byref(st)), self._h)
        return f, b, st

    # ---- probes ----
    def probe_trace(self, origins, directions, times=None, tmin=0.001, tmax=float("inf")):
        o = np.ascontiguousarray(origins, dtype=np.float32)
        d = np.ascontiguousarray(directions, dtype=np.float32)
        n = o.shape[0]
        t = np.ascontiguousarray(times, dtype=np.float32) if times is not None else None
        out = (ProbeHit * n)()
        _check(lib().rt_probe_trace(self._h, n, o.ctypes.data, d.ctypes.data, t.ctypes.data if t is not None else None,
                                    tmin, tmax, C.cast(out, C.c_void_p)), self._h)
        return out

    def probe_scatter(self, materials6, dir_in, recs, unit_vecs, rtests):
        m = np.ascontiguousarray(materials6, dtype=np.uint32)
        di = np.ascontiguousarray(dir_in, dtype=np.float32)
        uv = np.ascontiguousarray(unit_vecs, dtype=np.float32)
        rt = np.ascontiguousarray(rtests, dtype=np.float32)
        n = di.shape[0]
        do = np.empty((n, 3), np.float32)
        ao = np.empty((n, 3), np.float32)
        sc = np.empty(n, np.int32)
        _check(lib().rt_probe_scatter(self._h, n, m.ctypes.data, di.ctypes.data, C.cast(recs, C.c_void_p), uv.ctypes.data,
                                      rt.ctypes.data, do.ctypes.data, ao.ctypes.data, sc.ctypes.data), self._h)
        return do, ao, sc

    def probe_texture(self, offset, uv, points):
        u = np.ascontiguousarray(uv, dtype=np.float32)
        p = np.ascontiguousarray(points, dtype=np.float32)
        n = p.shape[0]
        out = np.empty((n, 3), np.float32)
        _check(lib().rt_probe_texture(self._h, n, offset, u.ctypes.data, p.ctypes.data, out.ctypes.data), self._h)
        return out

    def probe_xorwow(self, seed, first_subsequence, n):
        out = np.empty((n, 6), np.uint32)
        _check(lib().rt_probe_xorwow(self._h, seed, first_subsequence, n, out.ctypes.data), self._h)
        return out


class Group:
    """rt_group: several GPUs of one box behind one call (NCCL reduce onto device 0 inside the library)."""

    def __init__(self, devices):
        ids = list(range(devices)) if isinstance(devices, int) else list(devices)
        arr = (C.c_int * len(ids))(*ids)
        self._h = C.c_void_p()
        _check(lib().rt_group_create(arr, len(ids), C.byref(self._h)))
        self.devices = ids

    def close(self):
        if self._h:
            lib().rt_group_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def render_world(self, scene, settings=None, traversal=RT_TRAVERSAL_AUTO, shard=RT_SHARD_TILES):
        s = settings if settings is not None else scene.settings
        cam = Camera()
        lib().rt_camera_setup(C.byref(s), C.byref(cam))
        f = np.empty((cam.img_height, cam.img_width, 3), dtype=np.float32)
        b = np.empty((cam.img_height, cam.img_width, 3), dtype=np.uint8)
        st = Stats()
        rc = lib().rt_group_render_world(self._h, C.byref(scene.world), C.byref(s), traversal, shard, f.ctypes.data, b.ctypes.data, C.byref(st))
        if rc != 0:
            raise RtError("rt error %d: %s" % (rc, (lib().rt_group_last_error(self._h) or b"?").decode()))
        return f, b, st


def write_ppm(path, rgb8):
    a = np.ascontiguousarray(rgb8, dtype=np.uint8)
    _check(lib().rt_write_ppm(path.encode(), a.ctypes.data, a.shape[1], a.shape[0]))
