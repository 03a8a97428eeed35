"""ctypes bindings for oracle/_ref/libref_host.so (the real reference compiled for the host).

TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may use this.
"""
import ctypes as C
import os

import numpy as np

REF_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref")
REF_LIB = os.path.join(REF_DIR, "libref_host.so")


class RefScene(C.Structure):
    _fields_ = [
        ("hittables", C.c_void_p), ("hittable_count", C.c_uint64),
        ("world", C.POINTER(C.c_uint16)), ("world_count", C.c_uint64),
        ("root_wref", C.c_uint32),
        ("texture_arena", C.c_void_p), ("texture_bytes", C.c_uint64), ("texture_count", C.c_uint64),
        ("max_bvh_depth", C.c_uint64),
        ("img_ratio", C.c_float), ("img_width", C.c_int32), ("samples_per_pixel", C.c_int32),
        ("max_bounces", C.c_int32), ("vfov", C.c_float),
        ("lookfrom", C.c_float * 3), ("lookat", C.c_float * 3), ("view_up", C.c_float * 3),
        ("defocus_angle", C.c_float), ("background", C.c_float * 3),
        ("img_name", C.c_char * 64),
        ("priv_world", C.c_void_p),
    ]


class RefHit(C.Structure):
    _fields_ = [("point", C.c_float * 3), ("normal", C.c_float * 3), ("t", C.c_float), ("front", C.c_int32),
                ("mat", C.c_uint32 * 6), ("u", C.c_float), ("v", C.c_float)]


class RefCamera(C.Structure):
    _fields_ = [("img_width", C.c_int32), ("img_height", C.c_int32), ("samples_per_pixel", C.c_int32),
                ("max_bounces", C.c_int32), ("defocus_angle", C.c_float), ("background", C.c_float * 3),
                ("pixel_delta_v", C.c_float * 3), ("pixel_delta_u", C.c_float * 3), ("camera_pos", C.c_float * 3),
                ("defocus_disk_u", C.c_float * 3), ("defocus_disk_v", C.c_float * 3), ("pixel00", C.c_float * 3),
                ("pixel_samples_scale", C.c_float)]


def available():
    return os.path.exists(REF_LIB)


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(REF_LIB)
        L.ref_sizeof.restype = C.c_uint64
        L.ref_scene_build.argtypes = [C.c_int, C.POINTER(RefScene)]
        L.ref_scene_free.argtypes = [C.POINTER(RefScene)]
        L.ref_hit.argtypes = [C.POINTER(RefScene), C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_float,
                              C.c_float, C.c_float, C.POINTER(RefHit)]
        L.ref_aabb_hit.argtypes = [C.POINTER(C.c_float)] * 3 + [C.c_float, C.c_float]
        L.ref_texture_color.argtypes = [C.POINTER(RefScene), C.c_uint64, C.c_float, C.c_float,
                                        C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.ref_perlin_noise.argtypes = [C.POINTER(RefScene), C.c_uint64, C.POINTER(C.c_float)]
        L.ref_perlin_noise.restype = C.c_float
        L.ref_perlin_turbulence.argtypes = [C.POINTER(RefScene), C.c_uint64, C.POINTER(C.c_float), C.c_int]
        L.ref_perlin_turbulence.restype = C.c_float
        L.ref_scatter.argtypes = [C.POINTER(RefScene), C.POINTER(C.c_uint32), C.POINTER(C.c_float),
                                  C.POINTER(RefHit), C.POINTER(C.c_float), C.c_float,
                                  C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.ref_emit.argtypes = [C.POINTER(RefScene), C.POINTER(C.c_uint32), C.c_float, C.c_float,
                               C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.ref_xorwow_init.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint32)]
        L.ref_xorwow_uniforms.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, C.POINTER(C.c_float),
                                          C.POINTER(C.c_uint32)]
        L.ref_camera.argtypes = [C.POINTER(RefScene), C.c_int, C.c_int, C.c_int, C.POINTER(RefCamera)]
        L.ref_render_host.argtypes = [C.POINTER(RefScene), C.POINTER(RefCamera), C.c_uint64, C.c_uint64,
                                      C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_uint64), C.c_int]
        L.ref_write_ppm.argtypes = [C.POINTER(C.c_float), C.c_uint32, C.c_uint32, C.c_char_p]
        _lib = L
    return _lib


def build_scene(scene_id):
    s = RefScene()
    rc = lib().ref_scene_build(scene_id, C.byref(s))
    if rc != 0:
        raise ValueError("unknown scene %d" % scene_id)
    return s


def scene_arrays(s):
    """Copies of the three flat arrays as numpy (hittables as (count,120) uint8)."""
    h = np.ctypeslib.as_array(C.cast(s.hittables, C.POINTER(C.c_uint8)), shape=(s.hittable_count, 120)).copy()
    w = np.ctypeslib.as_array(s.world, shape=(s.world_count,)).copy()
    t = np.ctypeslib.as_array(C.cast(s.texture_arena, C.POINTER(C.c_uint8)), shape=(s.texture_bytes,)).copy()
    return h, w, t


def fptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def camera(s, width=0, spp=0, bounces=0):
    c = RefCamera()
    lib().ref_camera(C.byref(s), width, spp, bounces, C.byref(c))
    return c


def render_host(s, cam, seed=1234, subseq_offset=0, row0=0, row1=0, nthreads=0):
    out = np.zeros((cam.img_height, cam.img_width, 3), dtype=np.float32)
    rays = C.c_uint64(0)
    lib().ref_render_host(C.byref(s), C.byref(cam), seed, subseq_offset, row0, row1, fptr(out), C.byref(rays),
                          nthreads)
    return out, rays.value
