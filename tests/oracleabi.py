"""ctypes bindings for oracle/liboracle.so (plain-C CPU oracle).  TEST INFRASTRUCTURE."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_LIB = os.path.join(ROOT, "oracle", "liboracle.so")


class OrcScene(C.Structure):
    _fields_ = [("hittables", C.c_void_p), ("hittable_count", C.c_uint64),
                ("world", C.c_void_p), ("world_count", C.c_uint64), ("root_wref", C.c_uint32),
                ("texture_arena", C.c_void_p), ("texture_bytes", C.c_uint64)]


class OrcCamera(C.Structure):
    _fields_ = [("img_width", C.c_int32), ("img_height", C.c_int32), ("samples_per_pixel", C.c_int32),
                ("max_bounces", C.c_int32), ("defocus_angle", C.c_float), ("background", C.c_float * 3),
                ("pixel_delta_v", C.c_float * 3), ("pixel_delta_u", C.c_float * 3), ("camera_pos", C.c_float * 3),
                ("defocus_disk_u", C.c_float * 3), ("defocus_disk_v", C.c_float * 3), ("pixel00", C.c_float * 3),
                ("pixel_samples_scale", C.c_float)]


class OrcHit(C.Structure):
    _fields_ = [("point", C.c_float * 3), ("normal", C.c_float * 3), ("t", C.c_float), ("front", C.c_int32),
                ("mat", C.c_uint32 * 6), ("u", C.c_float), ("v", C.c_float)]


_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_LIB):
            build()
        L = C.CDLL(ORACLE_LIB)
        F = C.POINTER(C.c_float)
        L.orc_set_fma_profile.argtypes = [C.c_int]
        L.orc_xorwow_init.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint32)]
        L.orc_xorwow_uniforms.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, F, C.POINTER(C.c_uint32)]
        L.orc_aabb_hit.argtypes = [F, F, F, C.c_float, C.c_float]
        L.orc_hit.argtypes = [C.POINTER(OrcScene), F, F, C.c_float, C.c_float, C.c_float, C.POINTER(OrcHit)]
        L.orc_texture_color.argtypes = [C.POINTER(OrcScene), C.c_uint64, C.c_float, C.c_float, F, F]
        L.orc_perlin_noise.argtypes = [C.POINTER(OrcScene), C.c_uint64, F]
        L.orc_perlin_noise.restype = C.c_float
        L.orc_perlin_turbulence.argtypes = [C.POINTER(OrcScene), C.c_uint64, F, C.c_int]
        L.orc_perlin_turbulence.restype = C.c_float
        L.orc_scatter.argtypes = [C.POINTER(OrcScene), C.POINTER(C.c_uint32), F, C.POINTER(OrcHit), F, C.c_float, F, F]
        L.orc_emit.argtypes = [C.POINTER(OrcScene), C.POINTER(C.c_uint32), C.c_float, C.c_float, F, F]
        L.orc_render.argtypes = [C.POINTER(OrcScene), C.POINTER(OrcCamera), C.c_uint64, C.c_uint64, C.c_int, C.c_int,
                                 F, C.POINTER(C.c_uint64), C.c_int]
        L.orc_resolve_rgb8.argtypes = [F, C.c_uint64, C.POINTER(C.c_uint8)]
        _lib = L
    return _lib


class Scene:
    """Owns numpy copies of the three flat arrays and the OrcScene view over them."""

    def __init__(self, hittables, world, textures, root_wref):
        self.h = np.ascontiguousarray(hittables, dtype=np.uint8).reshape(-1, 120)
        self.w = np.ascontiguousarray(world, dtype=np.uint16)
        self.t = np.ascontiguousarray(textures, dtype=np.uint8)
        self.root = int(root_wref)
        self.c = OrcScene(self.h.ctypes.data, self.h.shape[0], self.w.ctypes.data, self.w.shape[0], self.root,
                          self.t.ctypes.data, self.t.shape[0])


def fptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def render(scene, cam, seed=1234, subseq_offset=0, row0=0, row1=0, nthreads=0, gpu_profile=False):
    L = lib()
    L.orc_set_fma_profile(1 if gpu_profile else 0)
    out = np.zeros((cam.img_height, cam.img_width, 3), dtype=np.float32)
    rays = C.c_uint64(0)
    L.orc_render(C.byref(scene.c), C.byref(cam), seed, subseq_offset, row0, row1, fptr(out), C.byref(rays), nthreads)
    L.orc_set_fma_profile(0)
    return out, rays.value


def resolve_rgb8(img):
    flat = np.ascontiguousarray(img, dtype=np.float32).reshape(-1)
    out = np.zeros(flat.shape[0], dtype=np.uint8)
    lib().orc_resolve_rgb8(fptr(flat), flat.shape[0] // 3, out.ctypes.data_as(C.POINTER(C.c_uint8)))
    return out.reshape(img.shape)
