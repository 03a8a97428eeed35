/* rt_api.h -- the C ABI of libraytracing_b200.so: the drop-in boundary of the path-tracing hot path.
 *
 * The reference (luk-zm/cuda_raytracing) has no plugin/FFI layer; its de-facto boundary is
 *     void render(World *world, RenderSettings *settings)            src/raytracing.cu:1649-1773
 * which uploads three flat arrays, launches init_gpu_data + cuda_calc and downloads the image.
 * Each entry point below names the part of render() it replaces.  Plain pointers and sizes only:
 * no CUDA, C++ or torch types cross this boundary.  Every function returns 0 (RT_OK) or a negative
 * RT_ERR_* code; rt_last_error_string() gives the text.  One context = one GPU = one host thread at a
 * time (distinct contexts are independent), like the reference's single synchronous default stream.
 *
 * There is NO CPU fallback: without a CUDA device rt_context_create() fails with RT_ERR_CUDA.
 */
#ifndef RT_API_H
#define RT_API_H

#include <stdint.h>
#include "rt_scene.h"

#ifdef __cplusplus
extern "C" {
#endif

enum {
  RT_OK = 0,
  RT_ERR_INVALID = -1,      /* bad argument */
  RT_ERR_CUDA = -2,         /* CUDA runtime error (text in rt_last_error_string) */
  RT_ERR_NO_SCENE = -3,     /* rt_render* before rt_scene_upload */
  RT_ERR_UNSUPPORTED = -4,  /* scene exceeds a device limit (stack depth, u16 refs, ...) */
  RT_ERR_NOMEM = -5,
  RT_ERR_IO = -6
};

typedef struct rt_context rt_context;

/* replaces: implicit device 0 / default stream of the reference (rt.cu:1694-1748 never selects one) */
int  rt_context_create(int device_id, rt_context **out);
void rt_context_destroy(rt_context *ctx);                 /* frees every device allocation (the reference never does) */
const char *rt_last_error_string(const rt_context *ctx);  /* ctx may be NULL: last global error */
/* enqueue on the caller's CUDA stream (cudaStream_t passed as void*; NULL is the CUDA default stream, which is
 * what the reference uses); rt_context_reset_stream() goes back to the context's own non-blocking stream */
int  rt_context_set_stream(rt_context *ctx, void *cuda_stream);
int  rt_context_reset_stream(rt_context *ctx);

/* ---- scene upload: replaces the 4 cudaMalloc+cudaMemcpy blocks of render() (rt.cu:1699-1733) ----
 * Input is the reference's own scene format AFTER make_bvh(): Hittable[120 B], u16 world refs (prims
 * then BVH nodes), root world ref, texture byte arena.  The callee copies and repacks (SoA primitives,
 * 128-bit packed nodes); the caller keeps ownership of its arrays. */
typedef struct rt_scene_desc {
  const rt_hittable *hittables;  uint64_t hittable_count;
  const uint16_t *world_refs;    uint64_t world_count;
  uint32_t root_wref;
  const uint8_t *texture_arena;  uint64_t texture_bytes;
} rt_scene_desc;

enum {
  RT_TRAVERSAL_AUTO = 0,       /* EXACT results: own SAH BVH when the scene has no instance wrappers (closest hit is
                                  order-independent), the reference's BVH and visiting order otherwise.  Exact up to two
                                  things that depend on the reference's visiting order even without wrappers: (i) candidates
                                  at exactly the same t -- the reference's quad / triangle test is inclusive (tmin <= t <=
                                  tmax, rt.cu:1066: among equal-t planars the one it visits LAST wins), its sphere test strict
                                  (the first one wins); the own-BVH and tiny-scene walks resolve such ties in their own
                                  order, so coplanar overlapping quads of different materials can shade differently;
                                  (ii) the reference's own aabb_hit rounds 1/d per node and can cull a true hit that the
                                  conservative own walks keep (measured: 2 of 810 000 pixels on C2, 4 of 160 000 on C3
                                  through the BVH).  RT_TRAVERSAL_REFERENCE has neither. */
  RT_TRAVERSAL_REFERENCE = 1,  /* always walk the reference's own BVH in the reference's order */
  RT_TRAVERSAL_FAST = 2        /* one SAH BVH over world-space boxes of everything, instances included (an instanced
                                  primitive is tested in its own space with the reference's arithmetic): statistical
                                  (tier-2) parity only when the scene has wrappers -- the visiting order, and with it the
                                  rounding drift the reference's wrappers leave on the ray, is not the reference's --
                                  identical to AUTO otherwise */
};
int rt_scene_upload(rt_context *ctx, const rt_scene_desc *scene, uint32_t traversal);

/* ---- render: replaces init_gpu_data<<<>>> + cuda_calc<<<>>> (rt.cu:1739-1761) and the D2H (1770-1771) ----
 * The vectors are exactly the by-value arguments of cuda_calc (rt.cu:1544-1547); rt_camera_setup()
 * (rt_scene.h) derives them from an rt_render_settings the way render() does (rt.cu:1657-1692). */
enum {
  RT_RENDER_OUTPUT_SUM = 1u,      /* leave the per-pixel SUM over samples (for multi-GPU sample sharding + reduce) */
  RT_RENDER_NO_RNG_CACHE = 2u,    /* recompute the XORWOW states even if cached from the previous call */
  RT_RENDER_COUNT_RAYS = 4u,      /* fill rt_stats.rays (warp-aggregated counter; small cost) */
  RT_RENDER_COUNT_OPS = 8u,       /* instrumented kernel variant: fill rt_stats.ops (never use for timing) */
  RT_RENDER_FORWARD_THROUGHPUT = 32u, /* NOT bit-identical: carry the product of a path's attenuations forwards in three
                                     registers instead of unwinding the reference's attenuation stack backwards.  Emission
                                     only ever ends a path in this renderer, so the value is the same product associated the
                                     other way round (<= ~n * 2^-24 relative per sample); it is output-only arithmetic and
                                     cannot change a path.  Inside tier 1's 1/255 tolerance, outside bit-identity. */
  RT_RENDER_RESUME = 16u          /* progressive accumulation: carry on from the state the previous
                                     render of the same seed / rows / subsequence offset left behind -- the per-pixel XORWOW
                                     states where it stopped, and the per-pixel SUMS still in d_pixels_rgb (needs
                                     RT_RENDER_OUTPUT_SUM on both).  M samples, then N more == M + N samples in one call,
                                     bit for bit.  The state can go through the host: rt_render_state_save / _load. */
};
/* per-render operation totals of the instrumented variant; index names below */
enum { RT_OP_BOX = 0, RT_OP_SPHERE, RT_OP_PLANAR, RT_OP_TRANSLATION, RT_OP_ROTATION, RT_OP_LAMBERT, RT_OP_METAL, RT_OP_GLASS,
       RT_OP_LIGHT, RT_OP_NOISE, RT_OP_CHECKER, RT_OP_ARENA, RT_OP_COUNT };
typedef struct rt_render_params {
  int32_t img_width, img_height;
  int32_t samples_per_pixel, max_bounces;    /* max_bounces > 255 wraps modulo 256 like the reference's u8 counter (rt.cu:220) */
  float defocus_angle;
  float background[3];
  float pixel_delta_v[3], pixel_delta_u[3], camera_pos[3], defocus_disk_u[3], defocus_disk_v[3], pixel00[3];
  uint64_t seed;                  /* 1234 in the reference (rt.cu:1747) */
  uint64_t subsequence_offset;    /* XORWOW subsequence = pixel index + this; 0 = the reference's stream */
  int32_t row_begin, row_end;     /* render rows [row_begin,row_end) only (tile sharding); 0,0 = whole image */
  uint32_t flags;
  /* interleaved row bands (tile sharding in ONE launch): with band_rows > 0 and band_period > 1 the call renders the
   * rows r with (r / band_rows) % band_period == band_phase and leaves the others untouched (row_begin/row_end must
   * be 0,0 or the whole image).  Same stream, same pixels as rendering the bands one by one. */
  uint32_t band_rows, band_period, band_phase;
} rt_render_params;

typedef struct rt_stats {
  double rng_init_ms, render_ms, resolve_ms;   /* CUDA-event times on the context's stream */
  double h2d_ms, d2h_ms, total_ms;             /* host wall clock (rt_render only) */
  uint64_t paths, rays;
  uint32_t kernel_launches;
  uint32_t traversal_used;                     /* RT_TRAVERSAL_REFERENCE or RT_TRAVERSAL_FAST */
  uint32_t grid_blocks, block_threads, regs_per_thread, blocks_per_sm;
  uint64_t ops[12];                            /* RT_RENDER_COUNT_OPS only */
} rt_stats;

void rt_render_params_from_camera(const rt_render_settings *s, const rt_camera *c, rt_render_params *p);

/* device-resident: d_pixels_rgb is a DEVICE pointer to img_width*img_height*3 floats (caller-owned, e.g. a torch
 * tensor); rows outside [row_begin,row_end) are left untouched.  Asynchronous on the context's stream except
 * for stats (stats != NULL synchronises). */
int rt_render_device(rt_context *ctx, const rt_render_params *p, float *d_pixels_rgb, rt_stats *stats);
/* rt_render_device WITHOUT any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same kernels, same image bit for bit as rt_render_device; the render kernel's scratch (claim counters, per-pixel chunk
 * progress, per-pixel XORWOW states: 28 B per pixel) lives in one file-scope `static __device__` array of
 * 256 + 28 * RT_STATIC_SCRATCH_MAX_PIXELS bytes (7 MiB) that every GPU loading the library holds until exit.  The call makes
 * no cudaMalloc / cudaFree and no allocating library call; it only enqueues on the context's stream (the first call on a
 * device also resolves the array's device address with cudaGetSymbolAddress: make it outside a capture).
 * Limits: the rendered set (rows x img_width, or the rows of the bands) must not exceed RT_STATIC_SCRATCH_MAX_PIXELS
 * -> RT_ERR_UNSUPPORTED with the two numbers in the message (render by rows instead); no RT_RENDER_RESUME, no stats.
 * Every call on a device SHARES the array: it clears what must start at zero itself, but two calls whose kernels overlap on
 * one GPU (two contexts / streams on the same device) race -- serialise them.  The scene upload and rt_context_create
 * before it do allocate; d_pixels_rgb is the caller's buffer as in rt_render_device. */
#define RT_STATIC_SCRATCH_MAX_PIXELS 262144u               /* 512 x 512 */
int rt_render_device_static(rt_context *ctx, const rt_render_params *p, float *d_pixels_rgb);
/* gamma + clamp + 8-bit quantise on the device (rt.cu:1391-1433): rgb8 = (u8)(256*clamp(sqrt(scale*x),0,0.999)) */
int rt_resolve_device(rt_context *ctx, const float *d_pixels_rgb, uint64_t n_pixels, float scale, uint8_t *d_rgb8);
/* host buffers: renders, resolves and copies back (h_rgb8 and/or h_pixels_rgb may be NULL) */
int rt_render(rt_context *ctx, const rt_render_params *p, float *h_pixels_rgb, uint8_t *h_rgb8, rt_stats *stats);
int rt_synchronize(rt_context *ctx);
/* checkpoint / resume: the per-pixel XORWOW states of the last render (24 B per rendered pixel; the sums are the
 * caller's own pixel buffer).  _save: *bytes_out = size needed; copies when capacity suffices.  _load: makes `state`
 * (saved after a render with the same seed, rows, bands and subsequence offset as `p`) the one RT_RENDER_RESUME picks up. */
int rt_render_state_save(rt_context *ctx, void *h_state, uint64_t capacity, uint64_t *bytes_out);
int rt_render_state_load(rt_context *ctx, const rt_render_params *p, const void *h_state, uint64_t bytes);

/* ---- the whole of render(World*, RenderSettings*) in one call (BVH build if needed, camera, upload, render) ---- */
int rt_render_world(rt_context *ctx, rt_world *world, const rt_render_settings *settings, uint32_t traversal,
                    float *h_pixels_rgb, uint8_t *h_rgb8, rt_stats *stats);

/* ---- several GPUs of one box: one context per device inside, NCCL over NVLink for the one exchange ----
 * The reference renders on the default device only (rt.cu:1722-1723, 1754); this is the same call on N devices.  One host
 * thread drives all of them.  NCCL (libnccl.so.2) is bound at run time and only needed for n_devices > 1. */
typedef struct rt_group rt_group;
enum {
  RT_SHARD_TILES = 0,     /* device g renders interleaved bands of rows with the reference's XORWOW stream: the image is
                             byte-identical to the one-GPU image (tier 1 kept) */
  RT_SHARD_SAMPLES = 1    /* device g renders its share of the samples of every pixel (subsequence = pixel + g * n_pixels):
                             perfect load balance, statistical (tier-2) parity for n_devices > 1 */
};
int  rt_group_create(const int *device_ids, int n_devices, rt_group **out);   /* device_ids NULL: devices 0 .. n_devices-1 */
void rt_group_destroy(rt_group *g);
int  rt_group_size(const rt_group *g);
rt_context *rt_group_context(rt_group *g, int index);     /* the context of device `index` (owned by the group) */
const char *rt_group_last_error(const rt_group *g);
/* rt_render_world() on every device of the group: scene packed once and uploaded to each, shards rendered concurrently,
 * one ncclReduce(sum) onto device 0, resolve + download there.  stats: render_ms = slowest device (XORWOW init + render),
 * total_ms = device 0's clock from the first launch to the resolved image (render + reduce + resolve). */
int  rt_group_render_world(rt_group *g, rt_world *world, const rt_render_settings *settings, uint32_t traversal, uint32_t shard,
                           float *h_pixels_rgb, uint8_t *h_rgb8, rt_stats *stats);

/* ---- per-function probes used by the parity tests (same kernels' device functions, batch in / batch out) ---- */
typedef struct rt_probe_hit {          /* one per ray; mirrors the reference's HitRecord fields + the mutated ray */
  float origin[3], direction[3];       /* ray AFTER traversal (instance wrappers edit it in place, Q12) */
  float point[3], normal[3];
  float t, u, v;
  int32_t hit, front, material_type;
} rt_probe_hit;
int rt_probe_trace(rt_context *ctx, uint64_t n, const float *origins, const float *directions, const float *times,
                   float tmin, float tmax, rt_probe_hit *out);
int rt_probe_scatter(rt_context *ctx, uint64_t n, const uint32_t *material6, const float *dir_in, const rt_probe_hit *rec,
                     const float *unit_vec, const float *random_test, float *dir_out, float *atten_out, int32_t *scattered);
int rt_probe_texture(rt_context *ctx, uint64_t n, uint64_t texture_offset, const float *uv, const float *points, float *rgb_out);
int rt_probe_xorwow(rt_context *ctx, uint64_t seed, uint64_t first_subsequence, uint64_t n, uint32_t *states6_out);

/* what rt_scene_upload would build for this scene -- host only, no CUDA call, no context (works without a GPU):
 * the walk it would use and the sizes of the device structures.  treelet_min / fuse_min as in the RT_TREELET_MIN /
 * RT_FUSE_MIN knobs (8 / 8 by default). */
typedef struct rt_pack_info {
  uint32_t traversal_used;                 /* RT_TRAVERSAL_REFERENCE or RT_TRAVERSAL_FAST */
  uint32_t kernel_mode;                    /* 0 reference order, 1 own BVH, 2 reference order + treelets / fused instances,
                                              3 reference order as a stackless linear program, 5 tiny scene: every primitive,
                                              6 own BVH with instanced leaves (RT_TRAVERSAL_FAST on a scene with wrappers) */
  uint32_t spheres, planars;               /* primitive records (flattened instances included in FAST mode) */
  uint32_t own_nodes;                      /* 64-byte own-BVH nodes (whole scene, or all treelets) */
  uint32_t treelets, treelet_prims, fused_instances;
  uint32_t stack_need, stack_capacity;     /* traversal stack entries per thread */
  uint32_t perlin_tables, has_wrappers;
  uint32_t program_ops;                    /* ops of the linear program (0: none -- own BVH, too long, or odd wrapper nesting) */
  uint32_t program_boxes, program_prims, program_wrappers;   /* of which box tests, primitive tests, wrappers (ways in) */
  uint32_t instanced_leaves;               /* RT_TRAVERSAL_FAST: leaves of the own BVH that are tested in the primitive's own space
                                              (translation / y-rotation chains around one primitive): the reference's records and
                                              the reference's arithmetic, so the rounding behaviour of its images is kept */
  uint32_t program_axis_quads;             /* 1: every planar op of the linear program is a quad with axis-aligned normal, u and v
                                              (walls, boxes): the walk runs their inside test in reduced form -- the same
                                              roundings with the exact-zero terms left out (env RT_AXIS_QUADS=0 keeps the general form) */
} rt_pack_info;
int rt_scene_pack_info(const rt_scene_desc *scene, uint32_t traversal, uint32_t treelet_min, uint32_t fuse_min, rt_pack_info *out);

/* The scheduling decisions rt_render_device takes for a render of `rows` rows of an img_width-wide image of that scene on a
 * GPU with sm_count SMs (148 on a B200) -- host only, no CUDA call, no context (works without a GPU; environment knobs
 * such as RT_CHUNK are not applied).  A pixel's samples are produced in order, a chunk at a time, by whichever lane
 * claims the chunk: equal chunks for the bulk of the render, halving ones for its tail. */
typedef struct rt_render_plan {
  uint32_t kernel_mode;                    /* rt_pack_info.kernel_mode after the short-lived-pixel rule (3 -> 4) */
  uint32_t resident_lanes;                 /* threads the persistent grid keeps resident */
  uint32_t chunk_count;
  int32_t  chunk_samples;                  /* > 0: uniform chunks of that many samples; 0: the table below */
  int32_t  chunk_begin[33];                /* chunk k renders samples [chunk_begin[k], chunk_begin[k + 1]) */
  uint32_t claim_group;                    /* idle lanes of a warp claim work items together, this many at a time */
  uint32_t scattered;                      /* item order: 0 tile order, 1 golden-ratio permutation of the queue */
} rt_render_plan;
int rt_render_plan_query(const rt_pack_info *scene, int32_t img_width, int32_t rows, int32_t samples_per_pixel, int32_t max_bounces,
                         int32_t sm_count, rt_render_plan *out);

/* The linear program RT_TRAVERSAL_AUTO would upload for `scene` (host only, no GPU needed), for inspection and tests:
 * 8 floats per op -- {x0,x1,y0,y1, z0,z1, skip op, K_BVH} for a box (a missed box continues at `skip`, every other op at
 * the next one), {plane n.xyz,d | 0, 0,0, primitive index, kind | 0x100 inside a wrapper} for a primitive (kind: 0 sphere,
 * 2 triangle, 3 quad; the index counts spheres, or quads and triangles together, in the order of the hittable array),
 * {offset.xyz,0, 0,0,0, 4 way in | 6 way out} for a translation, {sin,cos,0,0, 0,0,0, 5 | 7} for a y-rotation; the last
 * two words of an op are raw 32-bit integers.  *n_ops_out = ops of the program (0: the scene has none); at most
 * capacity_ops ops are written to ops8_out (may be NULL with capacity 0). */
int rt_scene_program(const rt_scene_desc *scene, float *ops8_out, uint32_t capacity_ops, uint32_t *n_ops_out);
/* translate(rotate(x)) is one op each way (flag 0x200): way in {offset.xyz,0, sin,cos,0, 4 | 0x200}, way out
 * {sin,cos,offset.x,offset.y, offset.z,0,0, 7 | 0x200}.
 *
 * The IMAGE of that program as the kernels walk it (host only; what rt_scene_upload copies to the device and every block
 * to shared memory): the ops with their indices turned into byte offsets relative to the op -- a box {.., .., byte offset
 * to the op a missed box continues at, 1 | 0x10}, a primitive {.., byte offset to its record, 0, index, kind | 0x8 for a
 * quad / triangle} -- then the end op {0, 0,0,0, 0 | 0x400}, then 4 x 16 bytes per quad / triangle ({n,d} {q,w.x} {v,w.y}
 * {u,w.z}, or with axis_quads != 0 on a scene whose rt_pack_info says program_axis_quads the reduced {n,d} {q,w.k}
 * {S_alpha,0} {S_beta,0}) and 2 x 16 bytes per sphere.  *n_float4_out = size of the image in 16-byte words (0: no program). */
int rt_scene_linear_image(const rt_scene_desc *scene, uint32_t axis_quads, float *float4_out, uint32_t capacity_float4, uint32_t *n_float4_out);

#ifdef __cplusplus
}
#endif
#endif /* RT_API_H */
