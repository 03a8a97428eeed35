// rt_kernels.cu -- the sm_100a kernels of the path-tracing hot path and their launchers.
//
//   render_kernel<MODE>   persistent path tracer: replaces cuda_calc (rt.cu:1543-1589) and ray_color (1334-1371)
//   rng_init_kernel       XORWOW states for subsequence = pixel (+offset): replaces init_gpu_data (rt.cu:1591-1606)
//   resolve_kernel        1/spp scale + gamma + clamp + 8-bit quantise: replaces the host loop of pixels_to_ppm (1419-1433)
//   probe_* kernels       the same device functions on caller-supplied batches (parity tests through the C ABI)
//
// Design: one lane owns one pixel and generates its samples IN ORDER from the pixel's own
// XORWOW state (that is what makes the reference's RNG stream reproducible), but a lane never waits for the
// other 31 pixels of its warp: when its path ends it starts its next sample at once (path regeneration), and
// when its pixel is finished it claims the next pixel from a global queue with one warp-aggregated atomic
// (ballot + popc + shuffle).  The traversal stack lives in shared memory ([slot][thread], conflict-free), the
// attenuation stack in thread-local memory, the accumulator in registers (one 12-byte store per pixel instead
// of the reference's global read-modify-write per sample).
#include "rt_kernels.h"
#include <stdlib.h>

#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

#include "rt_device.h"
#include "rt_math.cuh"
#include "rt_shade.cuh"
#include "rt_trace.cuh"

namespace rtk {
using namespace rtm;
using rtd::RenderArgs;
using rtd::Scene;
using rts::PerlinSmem;
using rtt::Hit;
using rtt::Ray;

#ifndef RT_NOISE_AT_PATH_END
#define RT_NOISE_AT_PATH_END 1
#endif
constexpr int RT_BLOCK = 128;
// resident blocks per SM asked of the compiler: 6 x 128 threads x 80 registers for the own-BVH walk, 5 x 96 for the
// reference-order walk (measured both ways, profiles/r1_sweep_tile_occ.jsonl; tools/build_variants.sh overrides)
#ifdef RT_MIN_BLOCKS_PER_SM
constexpr int RT_MIN_BLOCKS_FAST = RT_MIN_BLOCKS_PER_SM, RT_MIN_BLOCKS_REF = RT_MIN_BLOCKS_PER_SM;
#else
constexpr int RT_MIN_BLOCKS_FAST = 6, RT_MIN_BLOCKS_REF = 5;
#endif
// the stackless walk (modes 3 and 4) needs no shared memory and gains from residency: Cornell box s1024 377 ms at 5 x 96
// registers, 345 at 6 x 80, 337 at 7 x 72, 327 at 8 x 64, 361 at 10 x 48 (profiles/r1b_linear_program_knobs.jsonl).  Short-
// lived pixels (scene 0, s16 b8) want the opposite -- 1.11 ms at 5 x 96, 1.18 at 8 x 64 resident -- and get mode 4.
// The stack walks lose at 7 or 8 blocks (C2 95 -> 103 ms, C3 27 -> 33, scene 3 unchanged).
#ifndef RT_MIN_BLOCKS_LINEAR
#define RT_MIN_BLOCKS_LINEAR 8
#endif
constexpr int min_blocks(int mode) { return (mode == 1 || mode == 5 || mode == 6) ? RT_MIN_BLOCKS_FAST : (mode == 3 ? RT_MIN_BLOCKS_LINEAR : RT_MIN_BLOCKS_REF); }

__device__ __forceinline__ void stage_perlin(const Scene &sc, PerlinSmem *dst) {
  for (int i = threadIdx.x; i < 256; i += blockDim.x) {
    dst->vec[i] = __ldg(sc.perlin_vectors + i);
    dst->perm[i] = __ldg(sc.perlin_perm + i);
  }
}

__device__ __forceinline__ void count_values(const rtt::NoCount &, unsigned long long *v) { for (int k = 0; k < 12; ++k) v[k] = 0; }
__device__ __forceinline__ void count_values(const rtt::OpCount &c, unsigned long long *v) {
  for (int k = 0; k < 12; ++k) v[k] = 0;
  v[0] = c.n_box; v[1] = c.n_sphere; v[2] = c.n_planar; v[3] = c.n_translation; v[4] = c.n_rotation;
}

__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t *p) { return *reinterpret_cast<const volatile uint32_t *>(p); }
__device__ __forceinline__ void st_volatile_u32(uint32_t *p, uint32_t v) { *reinterpret_cast<volatile uint32_t *>(p) = v; }

#ifndef RT_LINEAR_SMEM
#define RT_LINEAR_SMEM 1
#endif
// kernel modes 3 / 4 with RT_LINEAR_SMEM: the linear program, the planar and the sphere records of the (small) scene are
// staged in shared memory when they fit, and the walk reads them with generic loads (measured, r2i_linear_program_in_smem.jsonl:
// Cornell box s1024 309.6 -> 306.1 ms, s128 42.6 -> 41.6 ms, scene 0 1.100 -> 1.071 ms; RT_LINEAR_SMEM=0 builds the global-memory walk)
__host__ __device__ inline uint32_t linear_smem_bytes(const Scene &sc) {
  return (RT_LINEAR_SMEM && sc.program_ops > 0) ? 16u * sc.linear_float4s : 0u;
}

template <int MODE, bool AX, class C>
__device__ __forceinline__ bool trace(const Scene &sc, Ray &r, Hit &h, uint32_t *stk, int stride, C &cnt) {
  const float inf = __uint_as_float(0x7f800000u);
  if (MODE == 0) return rtt::trace_reference<false>(sc, r, 0.001f, inf, h, stk, stride, cnt);
  if (MODE == 2) return rtt::trace_reference<true>(sc, r, 0.001f, inf, h, stk, stride, cnt);   // reference order + treelets
  if (MODE == 5) return rtt::trace_flat(sc, r, 0.001f, inf, h, cnt);                            // tiny scene: every primitive
  if (MODE == 6) return rtt::trace_fast<true>(sc, r, 0.001f, inf, h, stk, stride, cnt);         // own BVH with instanced leaves
  if (MODE >= 3) return RT_LINEAR_SMEM ? rtt::trace_linear<true, AX>(sc, r, 0.001f, inf, h, cnt)    // reference order, stackless
                                        : rtt::trace_linear<false, AX>(sc, r, 0.001f, inf, h, cnt);
  return rtt::trace_fast<false>(sc, r, 0.001f, inf, h, stk, stride, cnt);
}

// rendered set -> image: the set is a run of rows, or interleaved bands of rows (tile sharding in one launch)
__device__ __forceinline__ int image_row(const RenderArgs &a, int y) {
  if (a.band_rows == 0) return a.first_row + y;
  const int k = y / a.band_rows;
  return (k * a.band_period + a.band_phase) * a.band_rows + (y - k * a.band_rows);
}
__device__ __forceinline__ size_t image_pixel(const RenderArgs &a, int pix) {
  if (a.band_rows == 0) return (size_t)a.first_pixel + (size_t)pix;
  const int y = pix / a.img_width;
  return (size_t)image_row(a, y) * (size_t)a.img_width + (size_t)(pix - y * a.img_width);
}

// AX (kernel modes 3 / 4 only): every planar op of the program is an axis-aligned quad; the program's image holds the reduced
// records and the walk runs the reduced inside test (planar_inside_axis in rt_trace.cuh) -- same roundings, fewer
// instructions, no triangle case
template <int MODE, int MAXB, bool COUNT, bool AX = false>
__global__ void __launch_bounds__(RT_BLOCK, min_blocks(MODE)) render_kernel(const RenderArgs a) {
  extern __shared__ __align__(16) uint32_t smem[];
  Scene sc = a.scene;
#if RT_LINEAR_SMEM
  // The image of the linear program (ops + the records they test, one block; the packer keeps a program only if it fits) is
  // copied to shared memory unconditionally: the walk's pointer is then a shared-memory pointer at compile time (LDS with
  // immediate offsets), and everything it reads is addressed relative to the op at hand -- no base address inside the walk.
  if (MODE == 3 || MODE == 4) {                                              // the staging area is the last part of smem
    float4 *dst = reinterpret_cast<float4 *>(smem + a.linear_smem_word);
    for (uint32_t i = threadIdx.x; i < sc.linear_float4s; i += blockDim.x) dst[i] = __ldg(a.scene.program + i);
    sc.program = dst;
    __syncthreads();
  }
#endif
  uint32_t *stk = smem + threadIdx.x;
  constexpr int stride = RT_BLOCK;                       // launch_render_t always launches RT_BLOCK threads
  PerlinSmem *perlin = nullptr;
  float *coop = nullptr;                                // per-warp scratch of the cooperative turbulence
  if (sc.perlin_tables > 0) {
    perlin = reinterpret_cast<PerlinSmem *>(smem + a.stack_entries * RT_BLOCK);
    coop = reinterpret_cast<float *>(perlin + 1) + (threadIdx.x >> 5) * rts::SCRATCH_WORDS_PER_WARP;
    stage_perlin(sc, perlin);
    __syncthreads();
  }
  const unsigned lane = threadIdx.x & 31u;
  const unsigned full = 0xffffffffu;

  const V3 du = v3(a.du[0], a.du[1], a.du[2]), dv = v3(a.dv[0], a.dv[1], a.dv[2]);
  const V3 cam = v3(a.cam[0], a.cam[1], a.cam[2]);
  const V3 bg = v3(a.background[0], a.background[1], a.background[2]);
  const int capacity = a.max_bounces & 255;            // RayColorStack.max_count is a u8 (rt.cu:220, 1596)

  // MAXB > 0: thread-local attenuation stack, unwound backwards like the reference's (result = att * result + emission, the
  // same roundings: bit-identical).  MAXB == 0 (RT_RENDER_FORWARD_THROUGHPUT): the product of the attenuations is carried
  // forwards in three registers instead.  Emission only ever ENDS a path here (lights do not scatter, rt.cu:452-490), so
  // the radiance is e * a_n * ... * a_1 either way; only the association of the product differs (~n * 2^-24 relative,
  // output-only arithmetic: it cannot change a path).
  constexpr bool FWD = MAXB == 0;
  float att_r[FWD ? 1 : MAXB], att_g[FWD ? 1 : MAXB], att_b[FWD ? 1 : MAXB];
  V3 thr = v3(1.0f, 1.0f, 1.0f);
  // noise-textured bounces park {point, scale} in their slot; evaluated when (and if) the path has something to unwind
  // (not in the stackless walk of the small scenes, modes 3 / 4: none of them has a noise texture and its 320 bytes per
  // thread would only widen a local-memory frame that already spills past L2 -- such a scene evaluates per bounce)
  constexpr bool DEFER = !FWD && (RT_NOISE_AT_PATH_END != 0) && MODE != 3 && MODE != 4;
  float att_s[DEFER ? MAXB : 1];
  uint8_t nslot[DEFER ? MAXB : 1];                       // the parked slots of the current path, in order
  int n_noise = 0;

  // Work item = (chunk of `chunk_samples` samples, pixel), handed out chunk-major: item = chunk * pixel_count + pixel.
  // A pixel's samples are produced strictly in order from its own XORWOW state (that is the reference's stream),
  // but not necessarily by one lane: between chunks the state and the running sum are parked in global memory,
  // so the tail of a render is one chunk long instead of one pixel long (parallelism = pixels).
  int pix = -1, chunk = 0;                               // pix: index inside the rendered range
  uint32_t seen = 0;                                     // progress[pix] as last polled (see below)
  bool exhausted = false, new_path = false, waiting = false;
  int samples_left = 0, budget = 0, depth = 0;
  Rng rng = {0, 0, 0, 0, 0, 0};
  Ray ray; ray.o = cam; ray.d = cam; ray.time = 0.0f;
  V3 base = cam, acc = v3(0.0f, 0.0f, 0.0f);
  unsigned long long rays = 0;
  typename std::conditional<COUNT, rtt::OpCount, rtt::NoCount>::type cnt;
  unsigned long long shade[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // COUNT only: lambert, metal, glass, light, noise, checker, arena, 0

  for (;;) {
    // ---- claim work items for idle lanes: one atomic per warp ----
    // Lanes claim together, `claim_group` at a time (fewer only when the rest of the warp has nothing to do): items
    // claimed together are adjacent pixels, whose rays visit the same nodes (coherence), at the price of a short wait.
    // One vote per iteration: which lanes hold a pixel.  `exhausted` is warp-uniform -- the queue is handed out in order,
    // so once one claim of this warp came back past its end every later one would.
    unsigned busy = __ballot_sync(full, pix >= 0);
    const unsigned m = ~busy;
    const bool want = pix < 0;
    if (m && !exhausted && (__popc(m) >= a.claim_group || busy == 0u)) {
      const int leader = __ffs(m) - 1;
      unsigned long long first = 0;
      if ((int)lane == leader) first = atomicAdd(a.counters, (unsigned long long)__popc(m));
      first = __shfl_sync(full, first, leader);
      if (want) {
        const unsigned long long item = first + __popc(m & ((1u << lane) - 1u));
        if (item < a.item_count) {
          // items walk the range in tiles of 32 pixels (tile_w x 32/tile_w): a warp that claims together gets a compact
          // block of the image, whose rays visit the same nodes.  Slots of a border tile that fall outside are no-ops.
          uint32_t slot;
          if (a.item_count <= 0xffffffffull) {                // the usual case: 32-bit divide
            chunk = (int)((uint32_t)item / a.slots_per_chunk);
            slot = (uint32_t)item - (uint32_t)chunk * a.slots_per_chunk;
          } else {
            chunk = (int)(item / (unsigned long long)a.slots_per_chunk);
            slot = (uint32_t)(item - (unsigned long long)chunk * (unsigned long long)a.slots_per_chunk);
          }
          // scattered order: a warp's pixels come from all over the image, so the pixels with the long paths (which
          // bound a small image: their samples are produced in sequence) do not sit in the same warps
          if (a.scatter_mul) slot = (uint32_t)(((unsigned long long)slot * a.scatter_mul) % a.slots_per_chunk);
          const uint32_t tile = slot >> 5, l = slot & 31u;
          const uint32_t ty = tile / a.tiles_x, tx = tile - ty * a.tiles_x;
          const uint32_t x = (tx << a.tile_shift) + (l & ((1u << a.tile_shift) - 1u)), y = (ty << (5u - a.tile_shift)) + (l >> a.tile_shift);
          if (x < (uint32_t)a.img_width && y * (uint32_t)a.img_width + x < (uint32_t)a.pixel_count) {
            pix = (int)(y * (uint32_t)a.img_width + x);
            waiting = true;
            seen = 0;
            const float fi = (float)image_row(a, (int)y), fj = (float)x;     // rt.cu:1561-1567
            base = v3(fadd(ffma(fj, du.x, fmul(fi, dv.x)), a.p00[0]), fadd(ffma(fj, du.y, fmul(fi, dv.y)), a.p00[1]),
                      fadd(ffma(fj, du.z, fmul(fi, dv.z)), a.p00[2]));
          }
        }
      }
      exhausted = first + (unsigned long long)__popc(m) > a.item_count;       // some claim of this round came back past the end
      busy = __ballot_sync(full, pix >= 0);
    }
    // chunk k of a pixel starts from the state chunk k-1 left behind (ordered by the fence + progress counter).  The poll is
    // asynchronous: the value tested is the one requested an iteration ago, the next request goes out now and nobody reads
    // it before the next iteration -- a warp with a waiting lane no longer pays an L2 round trip per iteration for as long
    // as the wait lasts (measured before: 8-10 % of all stall samples of the Perlin scenes on this one load).
    bool start = false;
    if (waiting) {
      start = chunk == 0 || seen >= (uint32_t)chunk;
      if (!start) seen = ld_volatile_u32(a.progress + pix);
    }
    if (start) {
      waiting = false;
      if (chunk != 0) __threadfence();
      const uint2 *st = reinterpret_cast<const uint2 *>(a.rng_states) + 3 * (size_t)pix;
      const uint2 s0 = __ldcg(st), s1 = __ldcg(st + 1), s2 = __ldcg(st + 2);
      rng.d = s0.x; rng.v0 = s0.y; rng.v1 = s1.x; rng.v2 = s1.y; rng.v3 = s2.x; rng.v4 = s2.y;
      float *out = a.pixels + 3 * image_pixel(a, pix);
      acc = (chunk || a.resume) ? v3(__ldcg(out), __ldcg(out + 1), __ldcg(out + 2)) : v3(0.0f, 0.0f, 0.0f);
      if (a.chunk_samples > 0) {
        const int done = chunk * a.chunk_samples;
        samples_left = a.samples_per_pixel - done < a.chunk_samples ? a.samples_per_pixel - done : a.chunk_samples;
      } else {
        samples_left = a.chunk_begin[chunk + 1] - a.chunk_begin[chunk];
      }
      new_path = true;
      if (samples_left <= 0) {                      // spp <= 0: the reference writes 0 * scale
        out[0] = 0.0f; out[1] = 0.0f; out[2] = 0.0f;
        pix = -1;
      }
    }
    if (busy == 0u) {                                     // nobody holds a pixel: done, or the slots claimed fell outside the image
      if (exhausted) break;
      continue;
    }
    const bool live = pix >= 0 && !waiting;               // lanes without work stay in step (the warp shares work below)

    // ---- (1) camera ray generation (rt.cu:1570-1578) ----
    if (live && new_path) {
      new_path = false;
      V3 o = cam;
      if (!(a.defocus_angle <= 0.0f)) {
        float px, py; rng_unit_disk(rng, px, py);
        o = v3(fadd(ffma(px, a.ddu[0], fmul(py, a.ddv[0])), cam.x), fadd(ffma(px, a.ddu[1], fmul(py, a.ddv[1])), cam.y),
               fadd(ffma(px, a.ddu[2], fmul(py, a.ddv[2])), cam.z));
      }
      const float xo = fadd(rng_uniform(rng), -0.5f);
      const float yo = fadd(rng_uniform(rng), -0.5f);
      V3 d = vsub(base, o);
      d = v3(ffma(xo, du.x, d.x), ffma(xo, du.y, d.y), ffma(xo, du.z, d.z));
      d = v3(ffma(yo, dv.x, d.x), ffma(yo, dv.y, d.y), ffma(yo, dv.z, d.z));
      ray.o = o; ray.d = d; ray.time = rng_uniform(rng);
      budget = capacity; depth = 0;
      if (FWD) thr = v3(1.0f, 1.0f, 1.0f);
    }

    // ---- (2) one bounce: closest hit, (3) emission + scatter ----
    V3 result = v3(0.0f, 0.0f, 0.0f);
    bool done = live && (budget <= 0);
    bool noise_pending = false;                           // this bounce's attenuation is a Perlin texture, evaluated below
    V3 noise_p = cam; float noise_scale = 0.0f;
    if (live && !done) {
      ++rays;
      Hit h; h.u = 0.0f; h.v = 0.0f; h.t = 0.0f; h.prim = 0; h.front = 0; h.p = cam; h.n = cam;
      if (trace<MODE, AX>(sc, ray, h, stk, stride, cnt)) {
        const uint4 m0 = __ldg(sc.materials + 2 * h.prim);
        if (COUNT) {
          const uint32_t ty = m0.x & 0xffu, tk = (m0.x >> 8) & 0xffu;
          ++shade[ty & 3];
          if (ty == rtd::M_LAMBERTIAN || ty == rtd::M_DIFFUSE_LIGHT) { if (tk == rtd::TK_NOISE) ++shade[4]; else if (tk == rtd::TK_CHECKER_SOLID) ++shade[5]; else if (tk == rtd::TK_ARENA) ++shade[6]; }
        }
        const float4 m1 = __ldg(reinterpret_cast<const float4 *>(sc.materials + 2 * h.prim + 1));
        const V3 U = rng_unit_vector(rng);               // rt.cu:500-501: drawn for EVERY hit, whatever the material
        const float rtest = rng_uniform(rng);
        V3 dout, att;
        if (rts::scatter(sc, m0, m1, ray.d, h, U, rtest, perlin, dout, att, &noise_pending)) {
          noise_p = h.p; noise_scale = m1.w;
          ray.o = h.p; ray.d = dout;
          if (FWD) { if (!noise_pending) thr = v3(fmul(thr.x, att.x), fmul(thr.y, att.y), fmul(thr.z, att.z)); }
          else if (depth != capacity) { att_r[depth] = att.x; att_g[depth] = att.y; att_b[depth] = att.z; ++depth; }
          --budget;
          done = (budget <= 0);                          // bounce budget spent: result stays (0,0,0) (Q14)
        } else {
          if ((m0.x & 0xffu) == rtd::M_DIFFUSE_LIGHT) result = rts::material_texture(sc, m0, m1, h, perlin);
          done = true;
        }
      } else {
        result = bg;
        done = true;
      }
    }

    // ---- (3b) Perlin textures: parked until the path ends (DEFER), or evaluated now by the whole warp ----
    if (DEFER) {
      if (noise_pending) {
        att_r[depth - 1] = noise_p.x; att_g[depth - 1] = noise_p.y; att_b[depth - 1] = noise_p.z; att_s[depth - 1] = noise_scale;
        nslot[n_noise++] = (uint8_t)(depth - 1);
      }
    } else if (perlin != nullptr) {                       // block-uniform
      const unsigned mn = __ballot_sync(full, noise_pending);
      if (mn) {
        const float turb = rts::coop_turbulence7(perlin, coop, mn, noise_pending, noise_p);
        if (noise_pending) {                              // the slot scatter pushed (depth - 1) gets the colour
          const V3 c = rts::noise_colour(turb, noise_scale, noise_p.z);
          if (FWD) thr = v3(fmul(thr.x, c.x), fmul(thr.y, c.y), fmul(thr.z, c.z));
          else { att_r[depth - 1] = c.x; att_g[depth - 1] = c.y; att_b[depth - 1] = c.z; }
        }
      }
    }

    // ---- (5) unwind, accumulate, resolve ----
    if (DEFER && perlin != nullptr) {                     // block-uniform
      const bool unwinds = done && !(result.x == 0.0f && result.y == 0.0f && result.z == 0.0f);
      if (__any_sync(full, unwinds && n_noise > 0))
        rts::coop_resolve_noise(perlin, coop, unwinds ? n_noise : 0, nslot, att_r, att_g, att_b, att_s);
      if (done) n_noise = 0;
    }
    if (done) {
      if (FWD) result = v3(fmul(result.x, thr.x), fmul(result.y, thr.y), fmul(result.z, thr.z));
      if (result.x == 0.0f && result.y == 0.0f && result.z == 0.0f) depth = 0;   // 0 stays +0 through every fma(0, att, 0)
      while (!FWD && depth > 0) {                                // result = att (.) result + emission, one fma per channel
        --depth;
        result = v3(ffma(result.x, att_r[depth], 0.0f), ffma(result.y, att_g[depth], 0.0f), ffma(result.z, att_b[depth], 0.0f));
      }
      acc = vadd(acc, result);
      new_path = true;
      if (--samples_left == 0) {                         // chunk done: final value, or running sum + state for the next owner
        float *out = a.pixels + 3 * image_pixel(a, pix);
        uint2 *st = reinterpret_cast<uint2 *>(a.rng_states + 6 * (size_t)pix);
        const bool last = (chunk + 1 == a.chunk_count);
        __stcg(out, last ? fmul(a.sample_scale, acc.x) : acc.x);
        __stcg(out + 1, last ? fmul(a.sample_scale, acc.y) : acc.y);
        __stcg(out + 2, last ? fmul(a.sample_scale, acc.z) : acc.z);
        __stcg(st, make_uint2(rng.d, rng.v0)); __stcg(st + 1, make_uint2(rng.v1, rng.v2)); __stcg(st + 2, make_uint2(rng.v3, rng.v4));
        if (!last) { __threadfence(); st_volatile_u32(a.progress + pix, (uint32_t)(chunk + 1)); }
        pix = -1;
      }
    }
  }

  if (a.count_rays) {                                    // warp-aggregated ray counter
    for (int o = 16; o > 0; o >>= 1) rays += __shfl_xor_sync(full, rays, o);
    if (lane == 0) atomicAdd(a.counters + 1, rays);
  }
  if (COUNT) {                                           // operation counters -> counters[2..13]
    unsigned long long v[12];
    count_values(cnt, v);
    for (int k = 0; k < 7; ++k) v[5 + k] = shade[k];
    for (int k = 0; k < 12; ++k) {
      unsigned long long x = v[k];
      for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(full, x, o);
      if (lane == 0) atomicAdd(a.counters + 2 + k, x);
    }
  }
}

// ======================================================================================================
// rng_init_kernel: XORWOW state after curand_init(seed, subsequence, 0) for a run of consecutive subsequences.
// The state words v[0..4] are a 160-bit vector over GF(2); skipping s*2^67 steps multiplies it by
// prod_k (M^(2^(67+k)))^(bit k of s).  One thread decomposes its first subsequence bit by bit, then walks
// RNG_CHUNK-1 neighbours with the k=0 matrix staged in shared memory.  (cuRAND's own curand_init does the
// same algebra with 2-bit windows and 218 KB of global tables per module.)
// ======================================================================================================
constexpr int RNG_CHUNK = 8;
constexpr int RNG_BLOCK = 128;

__device__ __forceinline__ void gf2_apply(const uint32_t *__restrict__ mat /* 160 x 5 */, uint32_t v[5]) {
  uint32_t r0 = 0, r1 = 0, r2 = 0, r3 = 0, r4 = 0;
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    uint32_t bits = v[i];
    while (bits) {
      const int j = __ffs(bits) - 1;
      bits &= bits - 1;
      const uint32_t *row = mat + 5 * (i * 32 + j);
      r0 ^= row[0]; r1 ^= row[1]; r2 ^= row[2]; r3 ^= row[3]; r4 ^= row[4];
    }
  }
  v[0] = r0; v[1] = r1; v[2] = r2; v[3] = r3; v[4] = r4;
}

__global__ void __launch_bounds__(RNG_BLOCK) rng_init_kernel(uint64_t seed, uint64_t first, uint64_t n,
                                                             const uint32_t *__restrict__ mats, uint32_t *__restrict__ out) {
  __shared__ uint32_t m0[160 * 5];
  for (int i = threadIdx.x; i < 160 * 5; i += blockDim.x) m0[i] = mats[i];
  __syncthreads();
  const uint64_t chunk = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t begin = chunk * RNG_CHUNK;
  if (begin >= n) return;
  // _curand_init_scratch (curand_kernel.h:772-790)
  const uint32_t s0 = ((uint32_t)seed) ^ 0xaad26b49u;
  const uint32_t s1 = (uint32_t)(seed >> 32) ^ 0xf7dcefddu;
  const uint32_t t0 = 1099087573u * s0;
  const uint32_t t1 = 2591861531u * s1;
  const uint32_t d = 6615241u + t1 + t0;
  uint32_t v[5] = {123456789u + t0, 362436069u ^ t0, 521288629u + t1, 88675123u ^ t1, 5783321u + t0};
  uint64_t sub = first + begin;
  for (int k = 0; sub; ++k, sub >>= 1)
    if (sub & 1ull) gf2_apply(mats + (size_t)k * 800, v);
  const uint64_t end = (begin + RNG_CHUNK < n) ? begin + RNG_CHUNK : n;
  for (uint64_t i = begin; i < end; ++i) {
    uint32_t *o = out + 6 * i;
    o[0] = d; o[1] = v[0]; o[2] = v[1]; o[3] = v[2]; o[4] = v[3]; o[5] = v[4];
    if (i + 1 < end) gf2_apply(m0, v);
  }
}

cudaError_t launch_rng_init(uint64_t seed, uint64_t first_subsequence, uint64_t n, const uint32_t *seq_matrices,
                            uint32_t *states_out, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  const uint64_t chunks = (n + RNG_CHUNK - 1) / RNG_CHUNK;
  const unsigned grid = (unsigned)((chunks + RNG_BLOCK - 1) / RNG_BLOCK);
  rng_init_kernel<<<grid, RNG_BLOCK, 0, stream>>>(seed, first_subsequence, n, seq_matrices, states_out);
  return cudaGetLastError();
}

// ======================================================================================================
// resolve_kernel: the accumulate -> resolve -> gamma tail (rt.cu:1586, 1391-1395, 1419-1427, utility.h:43-49).
// HBM-bound streaming kernel: 12 B read + 3 B written per pixel, 4 channel values per thread (128-bit loads).
// ======================================================================================================
__device__ __forceinline__ uint8_t resolve_channel(float x, float scale) {
  const float c = fmul(scale, x);
  float g = (c > 0.0f) ? fsqrt(c) : 0.0f;                // linear_to_gamma; NaN -> 0
  if (g < 0.0f) g = 0.0f;
  if (g > 0.999f) g = 0.999f;
  return (uint8_t)fmul(256.0f, g);
}
__global__ void __launch_bounds__(256) resolve_kernel(const float *__restrict__ px, uint64_t n_values, float scale,
                                                      uint8_t *__restrict__ out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t quads = n_values / 4;
  for (uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; q < quads; q += stride) {
    const float4 v = __ldg(reinterpret_cast<const float4 *>(px) + q);
    uchar4 o;
    o.x = resolve_channel(v.x, scale); o.y = resolve_channel(v.y, scale);
    o.z = resolve_channel(v.z, scale); o.w = resolve_channel(v.w, scale);
    reinterpret_cast<uchar4 *>(out)[q] = o;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n_values & 3)) {
    const uint64_t i = quads * 4 + threadIdx.x;
    out[i] = resolve_channel(px[i], scale);
  }
}
cudaError_t launch_resolve(const float *pixels, uint64_t n_pixels, float scale, uint8_t *rgb8, cudaStream_t stream) {
  const uint64_t n_values = 3 * n_pixels;
  if (n_values == 0) return cudaSuccess;
  uint64_t blocks = (n_values / 4 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks == 0) blocks = 1;
  resolve_kernel<<<(unsigned)blocks, 256, 0, stream>>>(pixels, n_values, scale, rgb8);
  return cudaGetLastError();
}

// ======================================================================================================
// probe kernels: the device functions above on caller-supplied batches, one thread per item.
// ======================================================================================================
}  // namespace rtk
#include "rt_api.h"
namespace rtk {

constexpr int PROBE_BLOCK = 128;

template <int MODE>
__global__ void __launch_bounds__(PROBE_BLOCK) probe_trace_kernel(const Scene sc, uint64_t n, const float *__restrict__ org,
                                                                  const float *__restrict__ dir, const float *__restrict__ times,
                                                                  float tmin, float tmax, rt_probe_hit *__restrict__ out) {
  extern __shared__ __align__(16) uint32_t smem[];
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Ray r;
  r.o = v3(org[3 * i], org[3 * i + 1], org[3 * i + 2]);
  r.d = v3(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]);
  r.time = times ? times[i] : 0.0f;
  Hit h; h.u = 0.0f; h.v = 0.0f; h.t = 0.0f; h.prim = 0; h.front = 0; h.p = v3(0, 0, 0); h.n = v3(0, 0, 0);
  bool hit;
  rtt::NoCount cnt;
  if (MODE == 0) hit = rtt::trace_reference<true>(sc, r, tmin, tmax, h, smem + threadIdx.x, blockDim.x, cnt);
  else if (MODE == 3) {
    if (sc.linear_axis) {                                  // the walk the render kernel runs on this scene (AX)
      hit = rtt::trace_linear<false, true>(sc, r, tmin, tmax, h, cnt);
    } else {
      hit = rtt::trace_linear(sc, r, tmin, tmax, h, cnt);
    }
  }
  else if (MODE == 5) hit = rtt::trace_flat(sc, r, tmin, tmax, h, cnt);
  else if (MODE == 6) hit = rtt::trace_fast<true>(sc, r, tmin, tmax, h, smem + threadIdx.x, blockDim.x, cnt);
  else hit = rtt::trace_fast<false>(sc, r, tmin, tmax, h, smem + threadIdx.x, blockDim.x, cnt);
  rt_probe_hit o;
  o.origin[0] = r.o.x; o.origin[1] = r.o.y; o.origin[2] = r.o.z;
  o.direction[0] = r.d.x; o.direction[1] = r.d.y; o.direction[2] = r.d.z;
  o.point[0] = h.p.x; o.point[1] = h.p.y; o.point[2] = h.p.z;
  o.normal[0] = h.n.x; o.normal[1] = h.n.y; o.normal[2] = h.n.z;
  o.t = h.t; o.u = h.u; o.v = h.v; o.hit = hit ? 1 : 0; o.front = h.front;
  o.material_type = hit ? (int)(__ldg(sc.materials + 2 * h.prim).x & 0xffu) : -1;
  out[i] = o;
}

__global__ void __launch_bounds__(PROBE_BLOCK) probe_scatter_kernel(const Scene sc, uint64_t n, const uint4 *__restrict__ mats,
                                                                    const float *__restrict__ din, const rt_probe_hit *__restrict__ rec,
                                                                    const float *__restrict__ uvec, const float *__restrict__ rtest,
                                                                    float *__restrict__ dout, float *__restrict__ aout,
                                                                    int32_t *__restrict__ scattered) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4 m0 = mats[2 * i];
  const uint4 m1u = mats[2 * i + 1];
  const float4 m1 = make_float4(__uint_as_float(m1u.x), __uint_as_float(m1u.y), __uint_as_float(m1u.z), __uint_as_float(m1u.w));
  Hit h;
  h.p = v3(rec[i].point[0], rec[i].point[1], rec[i].point[2]);
  h.n = v3(rec[i].normal[0], rec[i].normal[1], rec[i].normal[2]);
  h.t = rec[i].t; h.u = rec[i].u; h.v = rec[i].v; h.front = rec[i].front; h.prim = 0;
  V3 d = v3(0, 0, 0), a = v3(0, 0, 0);
  const bool ok = rts::scatter(sc, m0, m1, v3(din[3 * i], din[3 * i + 1], din[3 * i + 2]), h,
                               v3(uvec[3 * i], uvec[3 * i + 1], uvec[3 * i + 2]), rtest[i], nullptr, d, a);
  dout[3 * i] = d.x; dout[3 * i + 1] = d.y; dout[3 * i + 2] = d.z;
  aout[3 * i] = a.x; aout[3 * i + 1] = a.y; aout[3 * i + 2] = a.z;
  scattered[i] = ok ? 1 : 0;
}

__global__ void __launch_bounds__(PROBE_BLOCK) probe_texture_kernel(const Scene sc, uint64_t n, uint32_t off,
                                                                    const float *__restrict__ uv, const float *__restrict__ pts,
                                                                    float *__restrict__ rgb) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const V3 c = rts::arena_texture(sc, off, uv[2 * i], uv[2 * i + 1], v3(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]));
  rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z;
}

static inline unsigned probe_grid(uint64_t n) { return (unsigned)((n + PROBE_BLOCK - 1) / PROBE_BLOCK); }

cudaError_t launch_probe_trace(const rtd::Scene &scene, int mode, uint64_t n, const float *origins, const float *directions,
                               const float *times, float tmin, float tmax, rt_probe_hit *out, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  const size_t smem = sizeof(uint32_t) * rtd::STACK_CAPACITY * PROBE_BLOCK;
  if (mode == 0) probe_trace_kernel<0><<<probe_grid(n), PROBE_BLOCK, smem, stream>>>(scene, n, origins, directions, times, tmin, tmax, out);
  else if (mode == 3) probe_trace_kernel<3><<<probe_grid(n), PROBE_BLOCK, 0, stream>>>(scene, n, origins, directions, times, tmin, tmax, out);
  else if (mode == 5) probe_trace_kernel<5><<<probe_grid(n), PROBE_BLOCK, 0, stream>>>(scene, n, origins, directions, times, tmin, tmax, out);
  else if (mode == 6) probe_trace_kernel<6><<<probe_grid(n), PROBE_BLOCK, smem, stream>>>(scene, n, origins, directions, times, tmin, tmax, out);
  else probe_trace_kernel<1><<<probe_grid(n), PROBE_BLOCK, smem, stream>>>(scene, n, origins, directions, times, tmin, tmax, out);
  return cudaGetLastError();
}
cudaError_t launch_probe_scatter(const rtd::Scene &scene, uint64_t n, const uint4 *materials2, const float *dir_in,
                                 const rt_probe_hit *rec, const float *unit_vec, const float *random_test, float *dir_out,
                                 float *atten_out, int32_t *scattered, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  probe_scatter_kernel<<<probe_grid(n), PROBE_BLOCK, 0, stream>>>(scene, n, materials2, dir_in, rec, unit_vec, random_test,
                                                                   dir_out, atten_out, scattered);
  return cudaGetLastError();
}
cudaError_t launch_probe_texture(const rtd::Scene &scene, uint64_t n, uint32_t texture_offset, const float *uv,
                                 const float *points, float *rgb_out, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  probe_texture_kernel<<<probe_grid(n), PROBE_BLOCK, 0, stream>>>(scene, n, texture_offset, uv, points, rgb_out);
  return cudaGetLastError();
}

// ======================================================================================================
// launch_render: persistent grid = SMs x resident blocks per SM (occupancy API), so every SM is full and the
// work queue does the load balancing.
// ======================================================================================================
template <int MODE, int MAXB, bool COUNT, bool AX = false>
static cudaError_t launch_render_t(const RenderArgs &args, int sm_count, cudaStream_t stream, LaunchInfo *info) {
  size_t smem = sizeof(uint32_t) * (size_t)args.stack_entries * RT_BLOCK;
  if (args.scene.perlin_tables > 0) smem += sizeof(PerlinSmem) + sizeof(float) * rts::SCRATCH_WORDS_PER_WARP * (RT_BLOCK / 32);
  RenderArgs largs = args;
  if (MODE == 3 || MODE == 4) {
    smem = (smem + 15) & ~size_t(15);
    largs.linear_smem_word = (uint32_t)(smem / 4);
    smem += linear_smem_bytes(args.scene);
  }
  auto kernel = render_kernel<MODE, MAXB, COUNT, AX>;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  // shared memory carve-out: what the resident blocks need (the scene data and the attenuation stacks live in L1, so
  // everything else stays cache); env RT_CARVEOUT = percent overrides (tuning knob)
  int carve = 50;
  {
    constexpr int want_blocks = min_blocks(MODE);
    const size_t need = (size_t)want_blocks * (smem + 1024 + 1024);     // + static + the per-block reservation
    carve = (int)((need * 100 + 228 * 1024 - 1) / (228 * 1024)) + 3;
    if (carve > 100) carve = 100;
    const char *env = getenv("RT_CARVEOUT");
    if (env) { const long x = strtol(env, nullptr, 10); if (x >= 0 && x <= 100) carve = (int)x; }
  }
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
  if (e != cudaSuccess) return e;
  int per_sm = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, RT_BLOCK, smem);
  if (e != cudaSuccess) return e;
  if (const char *env = getenv("RT_BLOCKS_PER_SM")) {      // tuning knob: fewer resident blocks than fit
    const long x = strtol(env, nullptr, 10);
    if (x >= 1 && x < per_sm) per_sm = (int)x;
  }
  if (per_sm < 1) per_sm = 1;
  long long want = ((long long)args.pixel_count + RT_BLOCK - 1) / RT_BLOCK;
  long long grid = (long long)sm_count * per_sm;
  if (grid > want) grid = want;
  if (grid < 1) grid = 1;
  if (info) {
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, kernel);
    info->grid = (int)grid; info->block = RT_BLOCK; info->regs = fa.numRegs; info->blocks_per_sm = per_sm; info->smem = smem;
  }
  kernel<<<(unsigned)grid, RT_BLOCK, smem, stream>>>(largs);
  return cudaGetLastError();
}

cudaError_t launch_render(const rtd::RenderArgs &args, int mode, int sm_count, cudaStream_t stream, LaunchInfo *info) {
  const bool small = (args.max_bounces & 255) <= 64;
  if (args.count_ops) {                                  // instrumented build of the same kernel (not the timed one)
    if (mode == 0) return launch_render_t<0, 256, true>(args, sm_count, stream, info);
    if (mode == 2) return launch_render_t<2, 256, true>(args, sm_count, stream, info);
    if (mode == 5) return launch_render_t<5, 256, true>(args, sm_count, stream, info);
    if (mode == 6) return launch_render_t<6, 256, true>(args, sm_count, stream, info);
    if (mode >= 3) return args.scene.linear_axis ? launch_render_t<3, 256, true, true>(args, sm_count, stream, info)   // the image holds the reduced records
                                                 : launch_render_t<3, 256, true>(args, sm_count, stream, info);
    return launch_render_t<1, 256, true>(args, sm_count, stream, info);
  }
  if (args.scene.linear_axis && (mode == 3 || mode == 4)) {   // linear program of axis-aligned quads (see render_kernel, AX)
    if (args.forward) return mode == 4 ? launch_render_t<4, 0, false, true>(args, sm_count, stream, info) : launch_render_t<3, 0, false, true>(args, sm_count, stream, info);
    if (mode == 4 && small) return launch_render_t<4, 64, false, true>(args, sm_count, stream, info);
    return small ? launch_render_t<3, 64, false, true>(args, sm_count, stream, info) : launch_render_t<3, 256, false, true>(args, sm_count, stream, info);
  }
  if (args.forward) {                                    // RT_RENDER_FORWARD_THROUGHPUT: no attenuation stack (MAXB = 0)
    if (mode == 0) return launch_render_t<0, 0, false>(args, sm_count, stream, info);
    if (mode == 2) return launch_render_t<2, 0, false>(args, sm_count, stream, info);
    if (mode == 5) return launch_render_t<5, 0, false>(args, sm_count, stream, info);
    if (mode == 6) return launch_render_t<6, 0, false>(args, sm_count, stream, info);
    if (mode == 4) return launch_render_t<4, 0, false>(args, sm_count, stream, info);
    if (mode == 3) return launch_render_t<3, 0, false>(args, sm_count, stream, info);
    return launch_render_t<1, 0, false>(args, sm_count, stream, info);
  }
  if (mode == 0) return small ? launch_render_t<0, 64, false>(args, sm_count, stream, info) : launch_render_t<0, 256, false>(args, sm_count, stream, info);
  if (mode == 2) return small ? launch_render_t<2, 64, false>(args, sm_count, stream, info) : launch_render_t<2, 256, false>(args, sm_count, stream, info);
  if (mode == 5) return small ? launch_render_t<5, 64, false>(args, sm_count, stream, info) : launch_render_t<5, 256, false>(args, sm_count, stream, info);
  if (mode == 6) return small ? launch_render_t<6, 64, false>(args, sm_count, stream, info) : launch_render_t<6, 256, false>(args, sm_count, stream, info);
  if (mode == 4 && small) return launch_render_t<4, 64, false>(args, sm_count, stream, info);   // mode 3 at 5 x 96 registers
  if (mode >= 3) return small ? launch_render_t<3, 64, false>(args, sm_count, stream, info) : launch_render_t<3, 256, false>(args, sm_count, stream, info);
  return small ? launch_render_t<1, 64, false>(args, sm_count, stream, info) : launch_render_t<1, 256, false>(args, sm_count, stream, info);
}

}  // namespace rtk
