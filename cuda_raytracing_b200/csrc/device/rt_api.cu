// rt_api.cu -- the C ABI (include/rt_api.h) over the kernels: context, scene upload, render, resolve, probes.
// No CPU fallback: every entry point needs a CUDA device and fails with RT_ERR_CUDA otherwise.
#include "rt_api.h"

#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>

#include <chrono>
#include <mutex>
#include <string>
#include <vector>

#include "rt_context.h"

namespace {

thread_local std::string g_last_error;     // per host thread: contexts used from different threads do not share it

using rtc::DeviceBuffer;

double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

template <class T>
cudaError_t to_device(DeviceBuffer &b, const T *src, size_t count, cudaStream_t s) { return b.upload(src, count * sizeof(T), s); }

}  // namespace

int rtc::fail(rt_context *ctx, int code, const std::string &msg) {
  g_last_error = msg;
  if (ctx) ctx->error = msg;
  return code;
}
using rtc::fail;
static int cuda_fail(rt_context *ctx, cudaError_t e, const char *what) {
  return fail(ctx, RT_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
#define RT_CUDA(ctx, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return cuda_fail(ctx, e_, #call); } while (0)

extern "C" {

const char *rt_last_error_string(const rt_context *ctx) { return ctx ? ctx->error.c_str() : g_last_error.c_str(); }

int rt_context_create(int device_id, rt_context **out) {
  if (!out) return fail(nullptr, RT_ERR_INVALID, "rt_context_create: out is NULL");
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    return fail(nullptr, RT_ERR_CUDA, std::string("rt_context_create: no CUDA device (") + cudaGetErrorString(e) + "); this library has no CPU fallback");
  if (device_id < 0 || device_id >= n) return fail(nullptr, RT_ERR_INVALID, "rt_context_create: device id out of range");
  RT_CUDA(nullptr, cudaSetDevice(device_id));
  rt_context *ctx = new rt_context();
  ctx->device = device_id;
  cudaDeviceProp prop;
  e = cudaGetDeviceProperties(&prop, device_id);
  if (e != cudaSuccess) { delete ctx; return cuda_fail(nullptr, e, "cudaGetDeviceProperties"); }
  ctx->sm_count = prop.multiProcessorCount;
  e = cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) { delete ctx; return cuda_fail(nullptr, e, "cudaStreamCreate"); }
  ctx->stream = ctx->own_stream;
  for (int i = 0; i < 4; ++i) cudaEventCreate(&ctx->ev[i]);
  if (const char *v = getenv("RT_CHUNK")) { const long x = strtol(v, nullptr, 10); if (x >= 0 && x <= (1 << 24)) ctx->chunk_samples = (int)x; }
  if (const char *v = getenv("RT_TILE_W")) { const long x = strtol(v, nullptr, 10); if (x >= 1 && x <= 32 && (x & (x - 1)) == 0) ctx->tile_w = (int)x; }
  if (const char *v = getenv("RT_FUSE_MIN")) { const long x = strtol(v, nullptr, 10); if (x >= 0 && x <= (1 << 30)) ctx->fuse_min = (uint32_t)x; }
  if (const char *v = getenv("RT_TREELET_MIN")) { const long x = strtol(v, nullptr, 10); if (x >= 0 && x <= (1 << 30)) ctx->treelet_min = (uint32_t)x; }
  if (const char *v = getenv("RT_LINEAR")) { const long x = strtol(v, nullptr, 10); if (x >= -1 && x <= 1) ctx->linear = (int)x; }
  if (const char *v = getenv("RT_FLAT")) { const long x = strtol(v, nullptr, 10); if (x >= -1 && x <= 1) ctx->flat = (int)x; }
  if (const char *v = getenv("RT_SCATTER")) { const long x = strtol(v, nullptr, 10); if (x >= -1 && x <= 1) ctx->scatter = (int)x; }
  if (const char *v = getenv("RT_AXIS_QUADS")) { const long x = strtol(v, nullptr, 10); if (x >= 0 && x <= 1) ctx->axis_quads = (int)x; }
  if (const char *v = getenv("RT_TAIL_MIN")) { const long x = strtol(v, nullptr, 10); if (x >= 1 && x <= 4096) ctx->tail_min = (int)x; }
  if (const char *v = getenv("RT_FIRST_CHUNK")) { const long x = strtol(v, nullptr, 10); if (x >= 0 && x <= (1 << 24)) ctx->first_chunk = (int)x; }
  if (const char *v = getenv("RT_CLAIM_GROUP")) { const long x = strtol(v, nullptr, 10); if (x >= 0 && x <= 32) ctx->claim_group = (int)x; }
  std::vector<uint32_t> mats;
  rtp::xorwow_sequence_matrices(mats);
  e = ctx->b_rng_mats.upload(mats.data(), mats.size() * sizeof(uint32_t), ctx->stream);
  if (e == cudaSuccess) e = ctx->b_counters.reserve(16 * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) { rt_context_destroy(ctx); return cuda_fail(nullptr, e, "rt_context_create upload"); }
  *out = ctx;
  return RT_OK;
}

void rt_context_destroy(rt_context *ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  DeviceBuffer *all[] = {&ctx->b_ref_nodes, &ctx->b_xforms, &ctx->b_instances, &ctx->b_structs, &ctx->b_struct_children, &ctx->b_fast_nodes,
                         &ctx->b_spheres, &ctx->b_planars, &ctx->b_materials, &ctx->b_checkers, &ctx->b_perlin_vec,
                         &ctx->b_perlin_perm, &ctx->b_arena, &ctx->b_program, &ctx->b_rng_mats, &ctx->b_rng_init, &ctx->b_rng_states,
                         &ctx->b_progress, &ctx->b_counters,
                         &ctx->b_pixels, &ctx->b_rgb8};
  for (DeviceBuffer *b : all) b->release();
  for (int i = 0; i < 4; ++i) if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
  if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
  delete ctx;
}

int rt_context_set_stream(rt_context *ctx, void *cuda_stream) {
  if (!ctx) return fail(nullptr, RT_ERR_INVALID, "rt_context_set_stream: ctx is NULL");
  ctx->stream = (cudaStream_t)cuda_stream;
  return RT_OK;
}

int rt_context_reset_stream(rt_context *ctx) {
  if (!ctx) return fail(nullptr, RT_ERR_INVALID, "rt_context_reset_stream: ctx is NULL");
  ctx->stream = ctx->own_stream;
  return RT_OK;
}

int rt_synchronize(rt_context *ctx) {
  if (!ctx) return fail(nullptr, RT_ERR_INVALID, "rt_synchronize: ctx is NULL");
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return RT_OK;
}

int rt_scene_upload(rt_context *ctx, const rt_scene_desc *desc, uint32_t traversal) {
  if (!ctx || !desc) return fail(ctx, RT_ERR_INVALID, "rt_scene_upload: NULL argument");
  rtp::PackedScene p;
  int rc = rtp::pack_scene(*desc, traversal, p, ctx->treelet_min, ctx->fuse_min, ctx->linear == 0 ? 0u : rtp::PROGRAM_MAX_DEFAULT);
  if (rc != RT_OK) { ctx->has_scene = false; return fail(ctx, rc, p.error); }
  return rtc::upload_packed(ctx, p);
}

}  // extern "C"

int rtc::upload_packed(rt_context *ctx, const rtp::PackedScene &packed) {
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  ctx->has_scene = false;
  ctx->live_valid = false;                                 // parked XORWOW states belong to the scene they were rendered with
  ctx->packed = packed;
  rtp::PackedScene &p = ctx->packed;
  if (ctx->flat == 0) p.flat = false;
  cudaStream_t s = ctx->stream;
#define UP(buf, vec) RT_CUDA(ctx, ctx->buf.upload((vec).data(), (vec).size() * sizeof((vec)[0]), s))
  UP(b_ref_nodes, p.ref_nodes); UP(b_xforms, p.xforms); UP(b_instances, p.instances); UP(b_structs, p.structures); UP(b_struct_children, p.struct_children);
  UP(b_fast_nodes, p.fast_nodes); UP(b_spheres, p.spheres); UP(b_planars, p.planars); UP(b_materials, p.materials);
  UP(b_checkers, p.checkers); UP(b_perlin_vec, p.perlin_vectors); UP(b_perlin_perm, p.perlin_perm); UP(b_arena, p.tex_arena);
  std::vector<float4> image;                               // the linear program as the kernels walk it (ops + records, one block)
  const bool axis = p.program_axis_quads && ctx->axis_quads != 0;
  rtp::build_linear_image(p, axis, image);
  UP(b_program, image);
#undef UP
  RT_CUDA(ctx, cudaStreamSynchronize(s));       // the host vectors may be reused after return
  rtd::Scene &sc = ctx->scene;
  sc.ref_nodes = (const float4 *)ctx->b_ref_nodes.ptr; sc.xforms = (const float4 *)ctx->b_xforms.ptr;
  sc.instances = (const float4 *)ctx->b_instances.ptr;
  sc.structures = (const uint2 *)ctx->b_structs.ptr; sc.struct_children = (const uint32_t *)ctx->b_struct_children.ptr;
  sc.program = (const float4 *)ctx->b_program.ptr; sc.program_ops = p.program.empty() ? 0u : (uint32_t)(p.program.size() / 2 + 1);
  sc.linear_float4s = (uint32_t)image.size(); sc.linear_axis = axis ? 1u : 0u;
  sc.ref_root = p.ref_root; sc.fast_nodes = (const float4 *)ctx->b_fast_nodes.ptr; sc.fast_root = p.fast_root;
  sc.spheres = (const float4 *)ctx->b_spheres.ptr; sc.planars = (const float4 *)ctx->b_planars.ptr;
  sc.sphere_count = p.sphere_count; sc.planar_count = p.planar_count; sc.flat_triangles = p.flat_triangles;
  sc.materials = (const uint4 *)ctx->b_materials.ptr; sc.checkers = (const float4 *)ctx->b_checkers.ptr;
  sc.perlin_vectors = (const float4 *)ctx->b_perlin_vec.ptr; sc.perlin_perm = (const uint8_t *)ctx->b_perlin_perm.ptr;
  sc.perlin_tables = p.perlin_tables; sc.tex_arena = (const uint8_t *)ctx->b_arena.ptr; sc.tex_bytes = p.tex_arena.size();
  {   // shared-memory traversal stack: what this scene's walk can need, in multiples of 4 entries
    const uint32_t need = p.traversal_used == RT_TRAVERSAL_FAST ? p.fast_stack_need : p.ref_stack_need;
    sc.stack_entries = std::min<uint32_t>((need + 3u) & ~3u, (uint32_t)rtd::STACK_CAPACITY);
    if (sc.stack_entries < 4) sc.stack_entries = 4;
  }
  ctx->has_scene = true;
  return RT_OK;
}

extern "C" {

// kernel variant: 1 = own BVH over the whole scene, 5 = tiny scene without wrappers (every primitive, no BVH),
// 3 = reference order as a linear program (small scenes with wrappers), 2 = reference order with treelets / fused
// instances, 0 = reference order through the stack
static int kernel_mode(const rtp::PackedScene &p) {
  if (p.traversal_used == RT_TRAVERSAL_FAST) return p.flat ? 5 : (p.instanced_leaves > 0 ? 6 : 1);
  if (p.treelets + p.fused_instances > 0) return 2;
  return p.program.empty() ? 0 : 3;
}

static int check_params(rt_context *ctx, const rt_render_params *p, int *row0, int *row1) {
  if (!ctx || !p) return fail(ctx, RT_ERR_INVALID, "rt_render: NULL argument");
  if (!ctx->has_scene) return fail(ctx, RT_ERR_NO_SCENE, "rt_render: no scene uploaded");
  if (p->img_width <= 0 || p->img_height <= 0) return fail(ctx, RT_ERR_INVALID, "rt_render: image size must be positive");
  if ((int64_t)p->img_width * p->img_height > 0x7fffffffLL / 3) return fail(ctx, RT_ERR_UNSUPPORTED, "rt_render: image too large");
  // max_bounces > 255 wraps modulo 256 exactly as in the reference, whose bounce-stack counter is a u8 (rt.cu:220, 1596:
  // `b300` renders 44 bounces); the kernels apply `& 255`
  if (p->max_bounces < 0) return fail(ctx, RT_ERR_INVALID, "rt_render: max_bounces must not be negative");
  int r0 = p->row_begin, r1 = p->row_end;
  if (r0 == 0 && r1 == 0) r1 = p->img_height;
  if (r0 < 0 || r1 > p->img_height || r0 >= r1) return fail(ctx, RT_ERR_INVALID, "rt_render: bad row range");
  *row0 = r0; *row1 = r1;
  return RT_OK;
}

// golden-ratio stride coprime to n (n >= 64): i -> i * mul mod n is a permutation with well-spread neighbours
static uint32_t scatter_multiplier(uint32_t n) {
  uint32_t m = (uint32_t)((double)n * 0.6180339887498949) | 1u;
  for (;; m += 2) {
    uint32_t x = m, y = n;
    while (y) { const uint32_t t = x % y; x = y; y = t; }
    if (x == 1) return m % n;
  }
}

// the rendered set of a call: a run of rows, or interleaved bands of rows (tile sharding in one launch)
struct RenderSet { bool banded; uint32_t rows; uint64_t first_pixel, count, first_sub, band_key; };
static int render_set(rt_context *ctx, const rt_render_params *p, int row0, int row1, RenderSet *out) {
  out->banded = p->band_rows > 0 && p->band_period > 1;
  out->rows = (uint32_t)(row1 - row0);
  if (out->banded) {
    if (p->band_phase >= p->band_period) return fail(ctx, RT_ERR_INVALID, "rt_render: band_phase must be below band_period");
    if (row0 != 0 || row1 != p->img_height) return fail(ctx, RT_ERR_INVALID, "rt_render: row bands need the whole row range");
    out->rows = 0;
    for (uint64_t b = p->band_phase; b * p->band_rows < (uint64_t)p->img_height; b += p->band_period) {
      const uint64_t lo = b * p->band_rows, hi = lo + p->band_rows;
      out->rows += (uint32_t)((hi < (uint64_t)p->img_height ? hi : (uint64_t)p->img_height) - lo);
    }
  }
  out->first_pixel = out->banded ? 0 : (uint64_t)row0 * p->img_width;
  out->count = (uint64_t)out->rows * p->img_width;
  out->first_sub = out->first_pixel + p->subsequence_offset;
  // the subsequence runs of a banded set start at lo * img_width, so the key of the cached / live states carries the image
  // size as well as the band parameters
  out->band_key = out->banded ? (((uint64_t)p->band_rows << 40) ^ ((uint64_t)p->band_period << 20) ^ p->band_phase ^ (1ull << 63)) +
                                    0x9e3779b97f4a7c15ull * (((uint64_t)(uint32_t)p->img_width << 32) | (uint32_t)p->img_height)
                              : 0;
  return RT_OK;
}

// ---- the scheduling decisions of a render: host logic only, shared by rt_render_device and rt_render_plan_query ----
struct PlanKnobs { int chunk_samples = 0, claim_group = 0, scatter = -1, tile_w = 8, first_chunk = 0, tail_min = 8; };
static void plan_render(int packed_mode, uint32_t traversal_used, uint32_t prims, int32_t img_width, uint32_t rows, int32_t samples_per_pixel,
                        int bounces, int sm_count, const PlanKnobs &kn, rt_render_plan *out) {
  memset(out, 0, sizeof(*out));
  int mode = packed_mode;
  const uint64_t count = (uint64_t)rows * (uint64_t)img_width;
  const int spp = samples_per_pixel > 0 ? samples_per_pixel : 1;
  const long long life = (long long)spp * (long long)(bounces > 0 ? bounces : 1);       // bounces of a pixel's whole life, at most
  // short-lived pixels (the regime that gets no time slices either) on the stackless walk: the 5 x 96-register build
  if (mode == 3 && bounces <= 64 && life <= 256) mode = 4;
  const double lanes = (double)sm_count * (mode == 3 ? 1024.0 : ((mode == 1 || mode == 5 || mode == 6) ? 768.0 : 640.0));   // resident threads: 8, 6 or 5 blocks of 128 per SM
  out->kernel_mode = (uint32_t)mode;
  out->resident_lanes = (uint32_t)lanes;
  // measured (profiles/r1_sweep_chunk_group.jsonl): big own-BVH scenes gain from claiming a whole warp of adjacent
  // pixels at once (coherent node visits), small scenes and the reference-order walk from claiming at once
  const bool big_fast = traversal_used == RT_TRAVERSAL_FAST && prims >= 64;
  out->claim_group = (uint32_t)(kn.claim_group > 0 ? kn.claim_group : (big_fast ? 32 : 1));
  // Item order (measured, profiles/r1_item_order_scatter.jsonl).  Tile order keeps neighbouring pixels in one warp and
  // neighbouring tiles on one SM (coherent walks, L1 hits): scattering costs C1 +40 %, C2 +70 %, scene 3 at 4K +34 %.
  // It pays where the render is short and bound by WHERE its slow pixels sit rather than by throughput:
  //  (a) a tiny image (every pixel resident at once, <= 2 per lane): the slow pixels otherwise fill whole warps whose
  //      every bounce then carries 32 Perlin evaluations (scene 17 at 400x400: 29.8 -> 27.2 ms with 4 time slices per
  //      pixel instead of 24 -- spreading the pixels already does what the many slices did);
  //  (b) the hybrid walk on a short render (scene 3 at 800x800: s16 136 -> 108 ms, s64 412 -> 390 ms): in tile order the
  //      queue ends with the image's heaviest rows.
  const uint32_t tw = (uint32_t)kn.tile_w, th = 32u / tw;
  const uint32_t slots = (((uint32_t)img_width + tw - 1) / tw) * ((rows + th - 1) / th) * 32u;
  const bool tiny_image = out->claim_group == 1 && (double)count <= 2.0 * lanes;
  const bool short_hybrid = mode == 2 && count <= (1u << 20) && (double)spp * (double)count <= 1024.0 * lanes;
  out->scattered = ((kn.scatter >= 0 ? kn.scatter == 1 : (tiny_image || short_hybrid)) && slots >= 64) ? 1u : 0u;

  // Chunks.  A pixel's samples are produced in order, a chunk at a time; between chunks its XORWOW state and running sum
  // are parked, so the render ends one CHUNK after the last claim, not one pixel.  Measured on the Cornell box
  // (profiles/r2a_c4_spp_chunk_prefetch_forward.jsonl, r2b_chunk_schedules.jsonl):
  //  * a hand-over costs a few iterations, so chunks below ~32 samples hurt (s128: 15 chunks of 9 samples 46.8 ms, 4 of 32
  //    44.0 ms) -- except where a pixel's samples are few and heavy (scene 3 at s16 needs its two chunks of 8);
  //  * chunks of more than ~1/16 of a resident lane's share of the render hurt too: a slow pixel's next chunk is claimed
  //    long before this one is done and the claimer waits (s1024: 14 chunks of 76 samples 318 ms, halving from 512 down
  //    340 ms; s128 as one chunk 52.4 ms);
  //  * so: equal chunks of 1/32 of a lane's share for the bulk, then halving ones down to 8 samples for the tail.
  if (samples_per_pixel <= 0) { out->chunk_samples = 1; out->chunk_count = 1; return; }   // spp <= 0: the kernel writes 0 * scale
  if (kn.chunk_samples > 0) {                              // RT_CHUNK: uniform chunks of that many samples
    const int c = kn.chunk_samples < spp ? kn.chunk_samples : spp;
    out->chunk_samples = c;
    out->chunk_count = (uint32_t)((spp + c - 1) / c);
    return;
  }
  const double per_lane = (double)spp * (double)count / lanes;
  // the tiny-scene walk (mode 5: the Perlin scenes) is bound by its slowest pixels' chains: fewer, longer chunks
  // (scene 19 s512: 92 samples 66.2 ms, 128 63.1, 184 63.7; scene 17 in scattered order: 4 slices 22.9 ms, 2 slices 22.3)
  int bulk = (int)(per_lane / (mode == 5 ? 24.0 : 32.0));
  // the 32-sample floor is the Cornell box's (cheap iterations, many of them); the hybrid walk over the big scene wants its
  // chunks short (scene 3 at s64: 13 samples 389 ms, 32 samples 409 ms), as before: never below 8
  const int floor_want = (mode == 3 || mode == 4) ? 32 : 8;
  const int floor_c = spp / 2 < 8 ? 8 : (spp / 2 < floor_want ? spp / 2 : floor_want);
  if (bulk < floor_c) bulk = floor_c;
  if (bulk < (spp + 23) / 24) bulk = (spp + 23) / 24;      // at most 24 equal chunks + the tail: the table holds 32
  // a pixel whose whole life is at most a few hundred bounces has no tail worth slicing: parking its state costs
  // more (C1, s16 b8: 1.31 -> 1.18 ms without slices; scene 3 at s16 b50 needs them: 143 vs 184 ms)
  if (life <= 256 || bulk >= spp) bulk = spp;
  int n = 0, done = 0;
  out->chunk_begin[0] = 0;
  if (bulk >= spp) {                                       // one chunk: nothing to slice
    out->chunk_begin[++n] = spp;
  } else if (out->scattered && tiny_image && kn.scatter < 0) {   // tiny image in scattered order: 2 equal slices (see (a) above)
    const int c = (spp + 1) / 2 > 8 ? (spp + 1) / 2 : 8;
    while (done < spp) { done = done + c < spp ? done + c : spp; out->chunk_begin[++n] = done; }
  } else {
    if (kn.first_chunk > 0 && kn.first_chunk < bulk) { done = kn.first_chunk; out->chunk_begin[++n] = done; }   // RT_FIRST_CHUNK (experiments)
    while (spp - done > 2 * bulk && n < rtd::MAX_CHUNK_TABLE - 6) { done += bulk; out->chunk_begin[++n] = done; }
    while (spp - done >= 2 * kn.tail_min && n < rtd::MAX_CHUNK_TABLE - 1) { done += (spp - done + 1) / 2; out->chunk_begin[++n] = done; }
    if (spp - done > 0) { done = spp; out->chunk_begin[++n] = done; }
  }
  out->chunk_count = (uint32_t)n;
  for (int k = n + 1; k <= rtd::MAX_CHUNK_TABLE; ++k) out->chunk_begin[k] = spp;
}

int rt_render_device(rt_context *ctx, const rt_render_params *p, float *d_pixels, rt_stats *stats) {
  int row0, row1;
  int rc = check_params(ctx, p, &row0, &row1);
  if (rc != RT_OK) return rc;
  if (!d_pixels) return fail(ctx, RT_ERR_INVALID, "rt_render_device: d_pixels_rgb is NULL");
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t s = ctx->stream;
  RenderSet set;
  rc = render_set(ctx, p, row0, row1, &set);
  if (rc != RT_OK) return rc;
  if (set.rows == 0) {                                     // more ranks than bands: nothing to do for this one
    if (stats) { memset(stats, 0, sizeof(*stats)); stats->traversal_used = ctx->packed.traversal_used; }
    return RT_OK;
  }
  const bool banded = set.banded;
  const uint32_t set_rows = set.rows;
  const uint64_t first_pixel = set.first_pixel, count = set.count, first_sub = set.first_sub, band_key = set.band_key;
  const bool resume = (p->flags & RT_RENDER_RESUME) != 0;
  if (resume) {
    if (!(p->flags & RT_RENDER_OUTPUT_SUM)) return fail(ctx, RT_ERR_INVALID, "rt_render: RT_RENDER_RESUME needs RT_RENDER_OUTPUT_SUM (the buffer must hold sums)");
    if (!(ctx->live_valid && ctx->live_seed == p->seed && ctx->live_first == first_sub && ctx->live_count == count && ctx->live_bands == band_key))
      return fail(ctx, RT_ERR_INVALID, "rt_render: RT_RENDER_RESUME without a previous render (or loaded state) of the same seed, rows and subsequence offset");
  }

  uint32_t launches = 0;
  if (stats) RT_CUDA(ctx, cudaEventRecord(ctx->ev[0], s));
  const bool cached = ctx->rng_valid && ctx->rng_seed == p->seed && ctx->rng_first == first_sub && ctx->rng_count == count &&
                      ctx->rng_bands == band_key && !(p->flags & RT_RENDER_NO_RNG_CACHE);
  if (!cached && !resume) {
    RT_CUDA(ctx, ctx->b_rng_init.reserve(count * 6 * sizeof(uint32_t)));
    if (!banded) {
      RT_CUDA(ctx, rtk::launch_rng_init(p->seed, first_sub, count, (const uint32_t *)ctx->b_rng_mats.ptr,
                                        (uint32_t *)ctx->b_rng_init.ptr, s));
      ++launches;
    } else {                                               // one run of consecutive subsequences per band
      uint64_t done = 0;
      for (uint64_t b = p->band_phase; b * p->band_rows < (uint64_t)p->img_height; b += p->band_period) {
        const uint64_t lo = b * p->band_rows, hi = lo + p->band_rows;
        const uint64_t n = ((hi < (uint64_t)p->img_height ? hi : (uint64_t)p->img_height) - lo) * p->img_width;
        RT_CUDA(ctx, rtk::launch_rng_init(p->seed, lo * p->img_width + p->subsequence_offset, n, (const uint32_t *)ctx->b_rng_mats.ptr,
                                          (uint32_t *)ctx->b_rng_init.ptr + 6 * done, s));
        done += n; ++launches;
      }
    }
    ctx->rng_seed = p->seed; ctx->rng_first = first_sub; ctx->rng_count = count; ctx->rng_bands = band_key; ctx->rng_valid = true;
  }
  // the render kernel advances the states chunk by chunk: it works on a copy, the initial states stay cached
  RT_CUDA(ctx, ctx->b_rng_states.reserve(count * 6 * sizeof(uint32_t)));
  RT_CUDA(ctx, ctx->b_progress.reserve(count * sizeof(uint32_t)));
  if (!resume) RT_CUDA(ctx, cudaMemcpyAsync(ctx->b_rng_states.ptr, ctx->b_rng_init.ptr, count * 6 * sizeof(uint32_t), cudaMemcpyDeviceToDevice, s));
  ctx->live_seed = p->seed; ctx->live_first = first_sub; ctx->live_count = count; ctx->live_bands = band_key; ctx->live_valid = true;
  RT_CUDA(ctx, cudaMemsetAsync(ctx->b_progress.ptr, 0, count * sizeof(uint32_t), s));
  RT_CUDA(ctx, cudaMemsetAsync(ctx->b_counters.ptr, 0, 16 * sizeof(unsigned long long), s));
  if (stats) RT_CUDA(ctx, cudaEventRecord(ctx->ev[1], s));

  rtd::RenderArgs a;
  memset(&a, 0, sizeof(a));
  a.scene = ctx->scene;
  a.pixels = d_pixels;
  a.rng_states = (uint32_t *)ctx->b_rng_states.ptr;
  a.progress = (uint32_t *)ctx->b_progress.ptr;
  const int bounces = p->max_bounces & 255;                // u8 counter of the reference (rt.cu:220)
  PlanKnobs knobs;
  knobs.chunk_samples = ctx->chunk_samples; knobs.claim_group = ctx->claim_group; knobs.scatter = ctx->scatter; knobs.tile_w = ctx->tile_w;
  knobs.first_chunk = ctx->first_chunk; knobs.tail_min = ctx->tail_min;
  rt_render_plan plan;
  plan_render(kernel_mode(ctx->packed), ctx->packed.traversal_used, ctx->packed.sphere_count + ctx->packed.planar_count, p->img_width, set_rows,
              p->samples_per_pixel, bounces, ctx->sm_count, knobs, &plan);
  const int mode = (int)plan.kernel_mode;
  a.chunk_samples = plan.chunk_samples;
  a.chunk_count = (int32_t)plan.chunk_count;
  for (int k = 0; k <= rtd::MAX_CHUNK_TABLE; ++k) a.chunk_begin[k] = plan.chunk_begin[k];
  a.claim_group = (int32_t)plan.claim_group;
  {
    const uint32_t tw = (uint32_t)ctx->tile_w, th = 32u / tw;
    a.tile_shift = 0; while ((1u << a.tile_shift) < tw) ++a.tile_shift;
    a.first_row = row0;
    if (banded) { a.band_rows = (int32_t)p->band_rows; a.band_period = (int32_t)p->band_period; a.band_phase = (int32_t)p->band_phase; }
    a.tiles_x = ((uint32_t)p->img_width + tw - 1) / tw;
    a.slots_per_chunk = a.tiles_x * ((set_rows + th - 1) / th) * 32u;
    a.item_count = (unsigned long long)a.slots_per_chunk * (unsigned long long)a.chunk_count;
    if (plan.scattered) a.scatter_mul = scatter_multiplier(a.slots_per_chunk);
  }
  a.counters = (unsigned long long *)ctx->b_counters.ptr;
  a.stack_entries = (int32_t)ctx->scene.stack_entries;
  a.img_width = p->img_width; a.img_height = p->img_height;
  a.samples_per_pixel = p->samples_per_pixel; a.max_bounces = bounces;
  a.first_pixel = (int32_t)first_pixel; a.pixel_count = (int32_t)count;
  a.defocus_angle = p->defocus_angle;
  memcpy(a.background, p->background, 12); memcpy(a.dv, p->pixel_delta_v, 12); memcpy(a.du, p->pixel_delta_u, 12);
  memcpy(a.cam, p->camera_pos, 12); memcpy(a.ddu, p->defocus_disk_u, 12); memcpy(a.ddv, p->defocus_disk_v, 12);
  memcpy(a.p00, p->pixel00, 12);
  a.sample_scale = (p->flags & RT_RENDER_OUTPUT_SUM) ? 1.0f : 1.0f / (float)p->samples_per_pixel;   // rt.cu:1657
  a.resume = resume ? 1u : 0u;
  a.count_rays = (p->flags & (RT_RENDER_COUNT_RAYS | RT_RENDER_COUNT_OPS)) ? 1u : 0u;
  a.count_ops = (p->flags & RT_RENDER_COUNT_OPS) ? 1u : 0u;
  a.forward = (p->flags & RT_RENDER_FORWARD_THROUGHPUT) ? 1u : 0u;
  if (mode >= 3 && mode <= 5) a.stack_entries = 0;         // neither the linear program nor the tiny-scene walk needs a traversal stack
  RT_CUDA(ctx, rtk::launch_render(a, mode, ctx->sm_count, s, &ctx->last_launch));
  ++launches;
  if (stats) {
    RT_CUDA(ctx, cudaEventRecord(ctx->ev[2], s));
    unsigned long long counters[16] = {0};
    RT_CUDA(ctx, cudaMemcpyAsync(counters, ctx->b_counters.ptr, sizeof(counters), cudaMemcpyDeviceToHost, s));
    RT_CUDA(ctx, cudaStreamSynchronize(s));
    float t_rng = 0, t_render = 0;
    cudaEventElapsedTime(&t_rng, ctx->ev[0], ctx->ev[1]);
    cudaEventElapsedTime(&t_render, ctx->ev[1], ctx->ev[2]);
    memset(stats, 0, sizeof(*stats));
    stats->rng_init_ms = t_rng; stats->render_ms = t_render;
    stats->paths = count * (uint64_t)(p->samples_per_pixel > 0 ? p->samples_per_pixel : 0);
    stats->rays = counters[1];
    for (int k = 0; k < 12; ++k) stats->ops[k] = counters[2 + k];
    stats->kernel_launches = launches;
    stats->traversal_used = ctx->packed.traversal_used;
    stats->grid_blocks = ctx->last_launch.grid; stats->block_threads = ctx->last_launch.block;
    stats->regs_per_thread = ctx->last_launch.regs; stats->blocks_per_sm = ctx->last_launch.blocks_per_sm;
  }
  return RT_OK;
}

int rt_resolve_device(rt_context *ctx, const float *d_pixels, uint64_t n_pixels, float scale, uint8_t *d_rgb8) {
  if (!ctx || !d_pixels || !d_rgb8) return fail(ctx, RT_ERR_INVALID, "rt_resolve_device: NULL argument");
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  RT_CUDA(ctx, rtk::launch_resolve(d_pixels, n_pixels, scale, d_rgb8, ctx->stream));
  return RT_OK;
}

int rt_render_state_save(rt_context *ctx, void *h_state, uint64_t capacity, uint64_t *bytes_out) {
  if (!ctx) return fail(nullptr, RT_ERR_INVALID, "rt_render_state_save: ctx is NULL");
  if (!ctx->live_valid) return fail(ctx, RT_ERR_INVALID, "rt_render_state_save: no render to save");
  const uint64_t bytes = ctx->live_count * 6 * sizeof(uint32_t);
  if (bytes_out) *bytes_out = bytes;
  if (!h_state || capacity < bytes) return h_state ? fail(ctx, RT_ERR_INVALID, "rt_render_state_save: buffer too small") : RT_OK;
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  RT_CUDA(ctx, cudaMemcpyAsync(h_state, ctx->b_rng_states.ptr, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return RT_OK;
}

int rt_render_state_load(rt_context *ctx, const rt_render_params *p, const void *h_state, uint64_t bytes) {
  int row0, row1;
  int rc = check_params(ctx, p, &row0, &row1);
  if (rc != RT_OK) return rc;
  if (!h_state) return fail(ctx, RT_ERR_INVALID, "rt_render_state_load: state is NULL");
  RenderSet set;
  rc = render_set(ctx, p, row0, row1, &set);
  if (rc != RT_OK) return rc;
  if (bytes != set.count * 6 * sizeof(uint32_t)) return fail(ctx, RT_ERR_INVALID, "rt_render_state_load: state size does not match the rendered set");
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  RT_CUDA(ctx, ctx->b_rng_states.reserve(bytes));
  RT_CUDA(ctx, cudaMemcpyAsync(ctx->b_rng_states.ptr, h_state, bytes, cudaMemcpyHostToDevice, ctx->stream));
  RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->live_seed = p->seed; ctx->live_first = set.first_sub; ctx->live_count = set.count; ctx->live_bands = set.band_key; ctx->live_valid = true;
  return RT_OK;
}

int rt_render(rt_context *ctx, const rt_render_params *p, float *h_pixels, uint8_t *h_rgb8, rt_stats *stats) {
  int row0, row1;
  int rc = check_params(ctx, p, &row0, &row1);
  if (rc != RT_OK) return rc;
  // the sums of a progressive render live in the CALLER's device buffer; this entry point renders into a scratch buffer
  // that it may clear, so it cannot carry them on
  if (p->flags & RT_RENDER_RESUME)
    return fail(ctx, RT_ERR_INVALID, "rt_render: RT_RENDER_RESUME needs rt_render_device (the running sums stay in the caller's device buffer)");
  const double t0 = now_ms();
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  const uint64_t n = (uint64_t)p->img_width * p->img_height;
  RT_CUDA(ctx, ctx->b_pixels.reserve(n * 3 * sizeof(float)));
  RT_CUDA(ctx, ctx->b_rgb8.reserve(n * 3 + 16));
  if (row1 - row0 != p->img_height || (p->band_rows > 0 && p->band_period > 1)) RT_CUDA(ctx, cudaMemsetAsync(ctx->b_pixels.ptr, 0, n * 3 * sizeof(float), ctx->stream));
  rt_stats local;
  rc = rt_render_device(ctx, p, (float *)ctx->b_pixels.ptr, &local);
  if (rc != RT_OK) return rc;
  cudaStream_t s = ctx->stream;
  float t_resolve = 0;
  if (h_rgb8) {
    RT_CUDA(ctx, cudaEventRecord(ctx->ev[2], s));
    RT_CUDA(ctx, rtk::launch_resolve((const float *)ctx->b_pixels.ptr, n, 1.0f, (uint8_t *)ctx->b_rgb8.ptr, s));
    RT_CUDA(ctx, cudaEventRecord(ctx->ev[3], s));
    local.kernel_launches++;
  }
  const double t1 = now_ms();
  if (h_rgb8) RT_CUDA(ctx, cudaMemcpyAsync(h_rgb8, ctx->b_rgb8.ptr, n * 3, cudaMemcpyDeviceToHost, s));
  if (h_pixels) RT_CUDA(ctx, cudaMemcpyAsync(h_pixels, ctx->b_pixels.ptr, n * 3 * sizeof(float), cudaMemcpyDeviceToHost, s));
  RT_CUDA(ctx, cudaStreamSynchronize(s));
  const double t2 = now_ms();
  if (h_rgb8) cudaEventElapsedTime(&t_resolve, ctx->ev[2], ctx->ev[3]);
  if (stats) {
    *stats = local;
    stats->resolve_ms = t_resolve; stats->d2h_ms = t2 - t1; stats->total_ms = t2 - t0;
  }
  return RT_OK;
}

// ---- probes ------------------------------------------------------------------------------------------

int rt_probe_trace(rt_context *ctx, uint64_t n, const float *origins, const float *directions, const float *times,
                   float tmin, float tmax, rt_probe_hit *out) {
  if (!ctx || !origins || !directions || !out) return fail(ctx, RT_ERR_INVALID, "rt_probe_trace: NULL argument");
  if (!ctx->has_scene) return fail(ctx, RT_ERR_NO_SCENE, "rt_probe_trace: no scene uploaded");
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t s = ctx->stream;
  DeviceBuffer o, d, t, r;
  RT_CUDA(ctx, to_device(o, origins, 3 * n, s)); RT_CUDA(ctx, to_device(d, directions, 3 * n, s));
  if (times) RT_CUDA(ctx, to_device(t, times, n, s));
  RT_CUDA(ctx, r.reserve(n * sizeof(rt_probe_hit)));
  const int mode = kernel_mode(ctx->packed) == 2 ? 0 : kernel_mode(ctx->packed);   // the stack walk covers treelets / instances
  cudaError_t e = rtk::launch_probe_trace(ctx->scene, mode, n, (const float *)o.ptr, (const float *)d.ptr,
                                          times ? (const float *)t.ptr : nullptr, tmin, tmax, (rt_probe_hit *)r.ptr, s);
  if (e == cudaSuccess) e = cudaMemcpyAsync(out, r.ptr, n * sizeof(rt_probe_hit), cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  o.release(); d.release(); t.release(); r.release();
  if (e != cudaSuccess) return cuda_fail(ctx, e, "rt_probe_trace");
  return RT_OK;
}

int rt_probe_scatter(rt_context *ctx, uint64_t n, const uint32_t *material6, const float *dir_in, const rt_probe_hit *rec,
                     const float *unit_vec, const float *random_test, float *dir_out, float *atten_out, int32_t *scattered) {
  if (!ctx || !material6 || !dir_in || !rec || !unit_vec || !random_test || !dir_out || !atten_out || !scattered)
    return fail(ctx, RT_ERR_INVALID, "rt_probe_scatter: NULL argument");
  if (!ctx->has_scene) return fail(ctx, RT_ERR_NO_SCENE, "rt_probe_scatter: no scene uploaded");
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t s = ctx->stream;
  // materials are packed exactly as rt_scene_upload packs them (same pre-resolution), against a scratch scene
  std::vector<uint4> mats(2 * n);
  rtp::PackedScene scratch;
  scratch.checkers = ctx->packed.checkers; scratch.checker_ids = ctx->packed.checker_ids;
  scratch.noise_ids = ctx->packed.noise_ids; scratch.perlin_tables = ctx->packed.perlin_tables;
  const uint8_t *arena = ctx->packed.tex_arena.data();
  for (uint64_t i = 0; i < n; ++i) {
    rt_material m; memcpy(&m, material6 + 6 * i, sizeof(m));
    rtp::pack_material(m, arena, ctx->packed.tex_arena.size(), scratch, &mats[2 * i]);
    const uint32_t kind = (mats[2 * i].x >> 8) & 0xffu;
    const bool known = (kind != rtd::TK_CHECKER_SOLID || mats[2 * i].z < ctx->packed.checkers.size() / 2) &&
                       (kind != rtd::TK_NOISE || mats[2 * i].z < ctx->packed.perlin_tables);
    if (!known) mats[2 * i].x = (mats[2 * i].x & ~0xff00u) | (rtd::TK_ARENA << 8);   // not in the uploaded tables: arena walk
  }
  DeviceBuffer bm, bd, br, bu, bt, od, oa, os;
  RT_CUDA(ctx, to_device(bm, mats.data(), 2 * n, s)); RT_CUDA(ctx, to_device(bd, dir_in, 3 * n, s));
  RT_CUDA(ctx, to_device(br, rec, n, s)); RT_CUDA(ctx, to_device(bu, unit_vec, 3 * n, s)); RT_CUDA(ctx, to_device(bt, random_test, n, s));
  RT_CUDA(ctx, od.reserve(12 * n)); RT_CUDA(ctx, oa.reserve(12 * n)); RT_CUDA(ctx, os.reserve(4 * n));
  cudaError_t e = rtk::launch_probe_scatter(ctx->scene, n, (const uint4 *)bm.ptr, (const float *)bd.ptr, (const rt_probe_hit *)br.ptr,
                                            (const float *)bu.ptr, (const float *)bt.ptr, (float *)od.ptr, (float *)oa.ptr,
                                            (int32_t *)os.ptr, s);
  if (e == cudaSuccess) e = cudaMemcpyAsync(dir_out, od.ptr, 12 * n, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaMemcpyAsync(atten_out, oa.ptr, 12 * n, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaMemcpyAsync(scattered, os.ptr, 4 * n, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  DeviceBuffer *all[] = {&bm, &bd, &br, &bu, &bt, &od, &oa, &os};
  for (DeviceBuffer *b : all) b->release();
  if (e != cudaSuccess) return cuda_fail(ctx, e, "rt_probe_scatter");
  return RT_OK;
}

int rt_probe_texture(rt_context *ctx, uint64_t n, uint64_t texture_offset, const float *uv, const float *points, float *rgb_out) {
  if (!ctx || !uv || !points || !rgb_out) return fail(ctx, RT_ERR_INVALID, "rt_probe_texture: NULL argument");
  if (!ctx->has_scene) return fail(ctx, RT_ERR_NO_SCENE, "rt_probe_texture: no scene uploaded");
  if (texture_offset + sizeof(rt_texture) > ctx->packed.tex_arena.size()) return fail(ctx, RT_ERR_INVALID, "rt_probe_texture: offset out of range");
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t s = ctx->stream;
  DeviceBuffer bu, bp, bo;
  RT_CUDA(ctx, to_device(bu, uv, 2 * n, s)); RT_CUDA(ctx, to_device(bp, points, 3 * n, s)); RT_CUDA(ctx, bo.reserve(12 * n));
  cudaError_t e = rtk::launch_probe_texture(ctx->scene, n, (uint32_t)texture_offset, (const float *)bu.ptr, (const float *)bp.ptr, (float *)bo.ptr, s);
  if (e == cudaSuccess) e = cudaMemcpyAsync(rgb_out, bo.ptr, 12 * n, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  bu.release(); bp.release(); bo.release();
  if (e != cudaSuccess) return cuda_fail(ctx, e, "rt_probe_texture");
  return RT_OK;
}

int rt_render_plan_query(const rt_pack_info *scene, int32_t img_width, int32_t rows, int32_t samples_per_pixel, int32_t max_bounces,
                         int32_t sm_count, rt_render_plan *out) {
  if (!scene || !out || img_width <= 0 || rows <= 0 || sm_count <= 0 || max_bounces < 0)
    return fail(nullptr, RT_ERR_INVALID, "rt_render_plan_query: bad argument");
  plan_render((int)scene->kernel_mode, scene->traversal_used, scene->spheres + scene->planars, img_width, (uint32_t)rows, samples_per_pixel,
              max_bounces & 255, sm_count, PlanKnobs(), out);
  return RT_OK;
}

int rt_scene_pack_info(const rt_scene_desc *scene, uint32_t traversal, uint32_t treelet_min, uint32_t fuse_min, rt_pack_info *out) {
  if (!scene || !out) return fail(nullptr, RT_ERR_INVALID, "rt_scene_pack_info: NULL argument");
  rtp::PackedScene p;
  const int rc = rtp::pack_scene(*scene, traversal, p, treelet_min, fuse_min);
  if (rc != RT_OK) return fail(nullptr, rc, p.error);
  memset(out, 0, sizeof(*out));
  out->traversal_used = p.traversal_used;
  out->kernel_mode = (uint32_t)kernel_mode(p);
  out->program_ops = (uint32_t)(p.program.size() / 2);
  out->program_boxes = p.program_boxes; out->program_prims = p.program_prims; out->program_wrappers = p.program_wrappers;
  out->spheres = p.sphere_count; out->planars = p.planar_count;
  out->own_nodes = (uint32_t)(p.fast_nodes.size() / 4);
  out->treelets = p.treelets; out->treelet_prims = p.treelet_prims; out->fused_instances = p.fused_instances;
  out->stack_need = p.traversal_used == RT_TRAVERSAL_FAST ? p.fast_stack_need : p.ref_stack_need;
  out->stack_capacity = (uint32_t)rtd::STACK_CAPACITY;
  out->perlin_tables = p.perlin_tables; out->has_wrappers = p.has_wrappers ? 1u : 0u;
  out->instanced_leaves = p.instanced_leaves;
  out->program_axis_quads = p.program_axis_quads ? 1u : 0u;
  return RT_OK;
}

int rt_scene_program(const rt_scene_desc *scene, float *ops8_out, uint32_t capacity_ops, uint32_t *n_ops_out) {
  if (!scene || !n_ops_out || (capacity_ops > 0 && !ops8_out)) return fail(nullptr, RT_ERR_INVALID, "rt_scene_program: NULL argument");
  rtp::PackedScene p;
  const int rc = rtp::pack_scene(*scene, RT_TRAVERSAL_AUTO, p);
  if (rc != RT_OK) return fail(nullptr, rc, p.error);
  const uint32_t n = kernel_mode(p) == 3 ? (uint32_t)(p.program.size() / 2) : 0u;
  *n_ops_out = n;
  if (n > 0 && capacity_ops > 0) memcpy(ops8_out, p.program.data(), sizeof(float) * 8 * (size_t)std::min(n, capacity_ops));
  return RT_OK;
}

int rt_scene_linear_image(const rt_scene_desc *scene, uint32_t axis_quads, float *float4_out, uint32_t capacity_float4, uint32_t *n_float4_out) {
  if (!scene || !n_float4_out || (capacity_float4 > 0 && !float4_out)) return fail(nullptr, RT_ERR_INVALID, "rt_scene_linear_image: NULL argument");
  rtp::PackedScene p;
  const int rc = rtp::pack_scene(*scene, RT_TRAVERSAL_AUTO, p);
  if (rc != RT_OK) return fail(nullptr, rc, p.error);
  std::vector<float4> image;
  if (kernel_mode(p) == 3) rtp::build_linear_image(p, axis_quads != 0, image);
  *n_float4_out = (uint32_t)image.size();
  if (!image.empty() && capacity_float4 > 0) memcpy(float4_out, image.data(), sizeof(float4) * (size_t)std::min<uint32_t>((uint32_t)image.size(), capacity_float4));
  return RT_OK;
}

int rt_probe_xorwow(rt_context *ctx, uint64_t seed, uint64_t first_subsequence, uint64_t n, uint32_t *states6_out) {
  if (!ctx || !states6_out) return fail(ctx, RT_ERR_INVALID, "rt_probe_xorwow: NULL argument");
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t s = ctx->stream;
  DeviceBuffer b;
  RT_CUDA(ctx, b.reserve(n * 24));
  cudaError_t e = rtk::launch_rng_init(seed, first_subsequence, n, (const uint32_t *)ctx->b_rng_mats.ptr, (uint32_t *)b.ptr, s);
  if (e == cudaSuccess) e = cudaMemcpyAsync(states6_out, b.ptr, n * 24, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  b.release();
  if (e != cudaSuccess) return cuda_fail(ctx, e, "rt_probe_xorwow");
  return RT_OK;
}

}  // extern "C"

// ---- rt_render_device_static: the render without any allocation ---------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops).  The render kernel's scratch -- the work-claim / statistics counters, the
// per-pixel chunk progress and the per-pixel XORWOW states it parks between chunks -- lives in ONE file-scope device array
// of fixed size instead of the context's grown buffers.  Every device that loads the library holds the array until exit
// (7 MiB); every call on a device shares it, so two calls that overlap on one GPU race.
namespace {
constexpr uint32_t kStaticMaxPixels = RT_STATIC_SCRATCH_MAX_PIXELS;
constexpr size_t kStaticCounterWords = 64;                                     // 16 u64 counters, padded to 256 B
constexpr size_t kStaticWords = kStaticCounterWords + 7 * (size_t)kStaticMaxPixels;   // + progress (1) + XORWOW state (6) per pixel
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mutex;
uint32_t *g_static_ptr[kStaticMaxDevices] = {};      // the array's DEVICE address on each device, resolved once per device
}  // namespace
static __device__ __align__(256) uint32_t g_static_scratch[kStaticWords];

// In host code `g_static_scratch` is a host-side stand-in for the symbol, not a device address: the address a kernel can use
// comes from cudaGetSymbolAddress, asked once per device with that device current.
static int static_scratch_address(rt_context *ctx, uint32_t **out) {
  if (ctx->device < 0 || ctx->device >= kStaticMaxDevices) return fail(ctx, RT_ERR_UNSUPPORTED, "rt_render_device_static: device index beyond 63");
  std::lock_guard<std::mutex> lock(g_static_mutex);
  if (!g_static_ptr[ctx->device]) {
    void *ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);        // the caller made ctx->device current
    if (e != cudaSuccess || !ptr) return cuda_fail(ctx, e, "rt_render_device_static: cudaGetSymbolAddress(static scratch)");
    g_static_ptr[ctx->device] = (uint32_t *)ptr;
  }
  *out = g_static_ptr[ctx->device];
  return RT_OK;
}

extern "C" int rt_render_device_static(rt_context *ctx, const rt_render_params *p, float *d_pixels) {
  int row0, row1;
  int rc = check_params(ctx, p, &row0, &row1);
  if (rc != RT_OK) return rc;
  if (!d_pixels) return fail(ctx, RT_ERR_INVALID, "rt_render_device_static: d_pixels_rgb is NULL");
  if (p->flags & RT_RENDER_RESUME)
    return fail(ctx, RT_ERR_INVALID, "rt_render_device_static: RT_RENDER_RESUME is not supported (the shared scratch does not keep a render's states)");
  RenderSet set;
  rc = render_set(ctx, p, row0, row1, &set);
  if (rc != RT_OK) return rc;
  if (set.rows == 0) return RT_OK;                         // more ranks than bands: nothing to do for this one
  if (set.count > kStaticMaxPixels)
    return fail(ctx, RT_ERR_UNSUPPORTED, "rt_render_device_static: the call renders " + std::to_string(set.count) + " pixels, the static scratch holds " +
                                             std::to_string(kStaticMaxPixels) + " (RT_STATIC_SCRATCH_MAX_PIXELS); use rt_render_device or render by rows");
  RT_CUDA(ctx, cudaSetDevice(ctx->device));
  uint32_t *scratch = nullptr;
  rc = static_scratch_address(ctx, &scratch);
  if (rc != RT_OK) return rc;
  unsigned long long *counters = (unsigned long long *)scratch;
  uint32_t *progress = scratch + kStaticCounterWords;
  uint32_t *states = progress + kStaticMaxPixels;
  cudaStream_t s = ctx->stream;
  const uint64_t count = set.count;

  // XORWOW states straight into the scratch (no cache of pristine states: that would be a second array)
  if (!set.banded) {
    RT_CUDA(ctx, rtk::launch_rng_init(p->seed, set.first_sub, count, (const uint32_t *)ctx->b_rng_mats.ptr, states, s));
  } else {
    uint64_t done = 0;
    for (uint64_t b = p->band_phase; b * p->band_rows < (uint64_t)p->img_height; b += p->band_period) {
      const uint64_t lo = b * p->band_rows, hi = lo + p->band_rows;
      const uint64_t n = ((hi < (uint64_t)p->img_height ? hi : (uint64_t)p->img_height) - lo) * p->img_width;
      RT_CUDA(ctx, rtk::launch_rng_init(p->seed, lo * p->img_width + p->subsequence_offset, n, (const uint32_t *)ctx->b_rng_mats.ptr, states + 6 * done, s));
      done += n;
    }
  }
  // the array keeps the last call's contents: the claim counter and the chunk progress must start at zero
  RT_CUDA(ctx, cudaMemsetAsync(progress, 0, count * sizeof(uint32_t), s));
  RT_CUDA(ctx, cudaMemsetAsync(counters, 0, 16 * sizeof(unsigned long long), s));

  rtd::RenderArgs a;
  memset(&a, 0, sizeof(a));
  a.scene = ctx->scene;
  a.pixels = d_pixels;
  a.rng_states = states;
  a.progress = progress;
  a.counters = counters;
  const int bounces = p->max_bounces & 255;
  PlanKnobs knobs;
  knobs.chunk_samples = ctx->chunk_samples; knobs.claim_group = ctx->claim_group; knobs.scatter = ctx->scatter; knobs.tile_w = ctx->tile_w;
  knobs.first_chunk = ctx->first_chunk; knobs.tail_min = ctx->tail_min;
  rt_render_plan plan;
  plan_render(kernel_mode(ctx->packed), ctx->packed.traversal_used, ctx->packed.sphere_count + ctx->packed.planar_count, p->img_width, set.rows,
              p->samples_per_pixel, bounces, ctx->sm_count, knobs, &plan);
  const int mode = (int)plan.kernel_mode;
  a.chunk_samples = plan.chunk_samples;
  a.chunk_count = (int32_t)plan.chunk_count;
  for (int k = 0; k <= rtd::MAX_CHUNK_TABLE; ++k) a.chunk_begin[k] = plan.chunk_begin[k];
  a.claim_group = (int32_t)plan.claim_group;
  const uint32_t tw = (uint32_t)ctx->tile_w, th = 32u / tw;
  while ((1u << a.tile_shift) < tw) ++a.tile_shift;
  a.first_row = row0;
  if (set.banded) { a.band_rows = (int32_t)p->band_rows; a.band_period = (int32_t)p->band_period; a.band_phase = (int32_t)p->band_phase; }
  a.tiles_x = ((uint32_t)p->img_width + tw - 1) / tw;
  a.slots_per_chunk = a.tiles_x * ((set.rows + th - 1) / th) * 32u;
  a.item_count = (unsigned long long)a.slots_per_chunk * (unsigned long long)a.chunk_count;
  if (plan.scattered) a.scatter_mul = scatter_multiplier(a.slots_per_chunk);
  a.stack_entries = (mode >= 3 && mode <= 5) ? 0 : (int32_t)ctx->scene.stack_entries;
  a.img_width = p->img_width; a.img_height = p->img_height;
  a.samples_per_pixel = p->samples_per_pixel; a.max_bounces = bounces;
  a.first_pixel = (int32_t)set.first_pixel; a.pixel_count = (int32_t)count;
  a.defocus_angle = p->defocus_angle;
  memcpy(a.background, p->background, 12); memcpy(a.dv, p->pixel_delta_v, 12); memcpy(a.du, p->pixel_delta_u, 12);
  memcpy(a.cam, p->camera_pos, 12); memcpy(a.ddu, p->defocus_disk_u, 12); memcpy(a.ddv, p->defocus_disk_v, 12);
  memcpy(a.p00, p->pixel00, 12);
  a.sample_scale = (p->flags & RT_RENDER_OUTPUT_SUM) ? 1.0f : 1.0f / (float)p->samples_per_pixel;
  a.forward = (p->flags & RT_RENDER_FORWARD_THROUGHPUT) ? 1u : 0u;
  rtk::LaunchInfo info;                                    // not kept: the context's record belongs to rt_render_device
  RT_CUDA(ctx, rtk::launch_render(a, mode, ctx->sm_count, s, &info));
  return RT_OK;
}
