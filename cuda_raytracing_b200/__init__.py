"""B200-native path tracer: Python plumbing over libraytracing_b200.so (C ABI in include/rt_api.h).

Only what the hot path needs lives here: csrc/ (CUDA kernels + C ABI + host scene construction in C) and
api.py (ctypes mirror).  The CUDA library is the product; importing this package never falls back to CPU code.
"""
from .api import (Context, Group, HostScene, RtError, lib, LIB_PATH, EXPORTS, write_ppm,  # noqa: F401
                  RT_TRAVERSAL_AUTO, RT_TRAVERSAL_REFERENCE, RT_TRAVERSAL_FAST,
                  RT_RENDER_OUTPUT_SUM, RT_RENDER_NO_RNG_CACHE, RT_RENDER_COUNT_RAYS, RT_RENDER_COUNT_OPS, RT_RENDER_RESUME,
                  RT_RENDER_FORWARD_THROUGHPUT, RT_STATIC_SCRATCH_MAX_PIXELS, RT_SHARD_TILES, RT_SHARD_SAMPLES, OP_NAMES)
