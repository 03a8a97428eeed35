// rt_shade.cuh -- texture evaluation (solid / checker / image / Perlin noise), emission and material scatter.
//
// Replaces texture_get_color, perlin_noise_val/perlin_turbulence, material_emit, material_scatter_common and
// reflect/refract/reflectance of the reference (rt.cu:343-398, 415-502; perlin_noise.cpp:1-46, 72-85).
#ifndef RT_SHADE_CUH
#define RT_SHADE_CUH

#include "rt_device.h"
#include "rt_math.cuh"
#include "rt_trace.cuh"

namespace rts {
using namespace rtm;
using rtd::Scene;

// Perlin tables of noise texture 0 staged in shared memory (north_star item 3); other tables stay in global.
struct PerlinSmem {
  float4 vec[256];
  uint8_t perm[256];
};

struct PerlinShared {
  const PerlinSmem *t;
  __device__ __forceinline__ V3 vec(int h) const { const float4 c = t->vec[h]; return v3(c.x, c.y, c.z); }
  __device__ __forceinline__ int perm(int i) const { return t->perm[i & 255]; }
};
struct PerlinGlobal {
  const float4 *v; const uint8_t *p;
  __device__ __forceinline__ V3 vec(int h) const { const float4 c = __ldg(v + h); return v3(c.x, c.y, c.z); }
  __device__ __forceinline__ int perm(int i) const { return __ldg(p + (i & 255)); }
};
struct PerlinArena {          // reference layout inside the texture arena: vec3[256] then i32 perm_x[256] at +3072
  const uint8_t *base;
  __device__ __forceinline__ V3 vec(int h) const {
    const float *f = reinterpret_cast<const float *>(base) + 3 * h;   // arena copy is 16-byte aligned, offsets are multiples of 4
    return v3(f[0], f[1], f[2]);
  }
  __device__ __forceinline__ int perm(int i) const { return reinterpret_cast<const int *>(base + 3072)[i & 255]; }
};

// perlin_noise_val + perlin_interp (perlin_noise.cpp:1-46).  perm_x hashes all three axes (lines 37-39).
template <class T>
__device__ __forceinline__ float perlin_noise(const T &tab, V3 p) {
  const float fx = floorf(p.x), fy = floorf(p.y), fz = floorf(p.z);
  const float u = fsub(p.x, fx), v = fsub(p.y, fy), w = fsub(p.z, fz);
  const int i = (int)fx, j = (int)fy, k = (int)fz;
  const float uu = fmul(fmul(u, u), fsub(3.0f, fadd(u, u)));
  const float vv = fmul(fmul(v, v), fsub(3.0f, fadd(v, v)));
  const float ww = fmul(fmul(w, w), fsub(3.0f, fadd(w, w)));
  const float wu[2] = {fsub(1.0f, uu), uu}, wv[2] = {fsub(1.0f, vv), vv}, wwt[2] = {fsub(1.0f, ww), ww};
  const float du[2] = {u, fadd(u, -1.0f)}, dv[2] = {v, fadd(v, -1.0f)}, dw[2] = {w, fadd(w, -1.0f)};
  const int pi_[2] = {tab.perm(i), tab.perm(i + 1)};
  const int pj_[2] = {tab.perm(j), tab.perm(j + 1)};
  const int pk_[2] = {tab.perm(k), tab.perm(k + 1)};
  float accum = 0.0f;
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int b = 0; b < 2; ++b)
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const V3 g = tab.vec(pi_[a] ^ pj_[b] ^ pk_[c]);
        const float d = ffma(dw[c], g.z, ffma(du[a], g.x, fmul(dv[b], g.y)));
        accum = ffma(fmul(fmul(wu[a], wv[b]), wwt[c]), d, accum);
      }
  return accum;
}
// perlin_turbulence (perlin_noise.cpp:72-85), depth fixed at 7 by the only call site (rt.cu:393)
template <class T>
__device__ __forceinline__ float perlin_turbulence7(const T &tab, V3 p) {
  float accum = 0.0f, weight = 1.0f;
#pragma unroll 1
  for (int o = 0; o < 7; ++o) {
    accum = ffma(weight, perlin_noise(tab, p), accum);
    weight = fmul(weight, 0.5f);
    p = v3(fadd(p.x, p.x), fadd(p.y, p.y), fadd(p.z, p.z));
  }
  return fabsf(accum);
}
// ---- warp-cooperative turbulence ----------------------------------------------------------------------------------
// The 7 octaves of perlin_turbulence are independent noise values at p, 2p, 4p, ... (exact doublings) that only
// meet in the final accumulate chain.  In a warp only the lanes that hit a noise-textured surface need one, so the
// (requests x 7) octave evaluations are dealt round-robin to ALL 32 lanes and each requester then runs the
// reference's accumulate chain over its 7 values: same bits, 32 lanes busy instead of the few that hit.
// scratch: 3 * 32 floats of points + 7 * 32 floats of octave values per warp.
constexpr int COOP_WORDS_PER_WARP = 3 * 32 + 7 * 32;
constexpr int SCRATCH_WORDS_PER_WARP = 4 * 32 + 7 * 32 + 32;      // the larger of this and coop_resolve_noise's (below)
__device__ __forceinline__ float coop_turbulence7(const PerlinSmem *table, float *scratch, unsigned requesters, bool mine, V3 p) {
  const unsigned lane = threadIdx.x & 31u;
  float *pts = scratch, *vals = scratch + 3 * 32;
  const int rank = __popc(requesters & ((1u << lane) - 1u));
  if (mine) { pts[3 * rank] = p.x; pts[3 * rank + 1] = p.y; pts[3 * rank + 2] = p.z; }
  __syncwarp();
  const int tasks = __popc(requesters) * 7;
  PerlinShared tab; tab.t = table;
  for (int t = (int)lane; t < tasks; t += 32) {
    const int req = t / 7, oct = t - req * 7;
    const float s = (float)(1 << oct);                                        // p * 2^oct == oct doublings, exactly
    vals[t] = perlin_noise(tab, v3(fmul(pts[3 * req], s), fmul(pts[3 * req + 1], s), fmul(pts[3 * req + 2], s)));
  }
  __syncwarp();
  float accum = 0.0f, weight = 1.0f;
  if (mine) {
#pragma unroll
    for (int o = 0; o < 7; ++o) { accum = ffma(weight, vals[7 * rank + o], accum); weight = fmul(weight, 0.5f); }
  }
  __syncwarp();                                                               // scratch is reused by the next call
  return fabsf(accum);
}

// ---- the noise textures of finished paths, all at once ----------------------------------------------------------------
// The attenuation of a bounce is not read before its path ends (the unwind), and not at all when the path ends with
// radiance 0 (bounce budget spent, or a black background): so a noise-textured bounce only parks its point in its
// attenuation slot, and the warp evaluates the parked points of the lanes whose path just ended with something to
// unwind -- up to 32 points per round, 7 octaves each dealt round-robin to all 32 lanes, then one lane per point folds the
// octaves in the reference's order and takes the sine.  Same perlin_noise / fold / noise_colour as the per-bounce
// evaluation: same bits.  What it buys: the slowest pixels of the Perlin scenes (between the spheres almost every path
// spends its 50 bounces and returns 0) never evaluate a texture at all, and a path that does is resolved with all lanes
// busy instead of 7 per request.
// scratch per warp: 32 x {x, y, z, scale} + 7 x 32 octave values + 32 colours

__device__ __forceinline__ V3 noise_colour(float turb, float scale, float z);
__device__ __forceinline__ void coop_resolve_noise(const PerlinSmem *table, float *scratch, int n_mine, const uint8_t *nslot,
                                                   float *att_r, float *att_g, float *att_b, const float *att_s) {
  const unsigned full = 0xffffffffu;
  const int lane = (int)(threadIdx.x & 31u);
  float *pts = scratch, *vals = scratch + 4 * 32, *cols = scratch + 4 * 32 + 7 * 32;
  PerlinShared tab; tab.t = table;
  int served = 0, remaining = n_mine;
  for (;;) {
    if (!__any_sync(full, remaining > 0)) break;
    const int want = remaining < 32 ? remaining : 32;
    int incl = want;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(full, incl, o); if (lane >= o) incl += v; }
    const int excl = incl - want;
    int total = __shfl_sync(full, incl, 31);
    if (total > 32) total = 32;
    int take = 32 - excl; take = take < 0 ? 0 : (take < want ? take : want);
    for (int k = 0; k < take; ++k) {
      const int j = nslot[served + k];
      float *q = pts + 4 * (excl + k);
      q[0] = att_r[j]; q[1] = att_g[j]; q[2] = att_b[j]; q[3] = att_s[j];
    }
    __syncwarp();
    const int tasks = total * 7;
    for (int t = lane; t < tasks; t += 32) {
      const int req = t / 7, oct = t - req * 7;
      const float s = (float)(1 << oct);                                      // p * 2^oct == oct doublings, exactly
      vals[t] = perlin_noise(tab, v3(fmul(pts[4 * req], s), fmul(pts[4 * req + 1], s), fmul(pts[4 * req + 2], s)));
    }
    __syncwarp();
    if (lane < total) {
      float accum = 0.0f, weight = 1.0f;
#pragma unroll
      for (int o = 0; o < 7; ++o) { accum = ffma(weight, vals[7 * lane + o], accum); weight = fmul(weight, 0.5f); }
      cols[lane] = noise_colour(fabsf(accum), pts[4 * lane + 3], pts[4 * lane + 2]).x;
    }
    __syncwarp();
    for (int k = 0; k < take; ++k) {
      const int j = nslot[served + k];
      const float c = cols[excl + k];
      att_r[j] = c; att_g[j] = c; att_b[j] = c;
    }
    served += take; remaining -= take;
    __syncwarp();                                                             // scratch is reused by the next round
  }
}

// noise texture colour (rt.cu:386-395): 0.5*(1 + sin(scale*z + 10*turb))
__device__ __forceinline__ V3 noise_colour(float turb, float scale, float z) {
  const float c = fmul(fadd(sinf(ffma(z, scale, fmul(turb, 10.0f))), 1.0f), 0.5f);
  return v3(c, c, c);
}
__device__ __forceinline__ bool checker_even(float inv_scale, V3 p) {       // rt.cu:365-371
  const int x = (int)floorf(fmul(p.x, inv_scale)), y = (int)floorf(fmul(p.y, inv_scale)), z = (int)floorf(fmul(p.z, inv_scale));
  return ((x + y + z) & 1) == 0;
}

// texture_get_color over the reference-format arena (generic fallback: nested checkers, image textures,
// extra noise tables).  The checker recursion of the reference becomes a loop.
__device__ __noinline__ V3 arena_texture(const Scene &sc, uint32_t off, float u, float v, V3 p) {
  for (int guard = 0; guard < 64; ++guard) {
    const uint8_t *t = sc.tex_arena + off;
    const uint32_t type = *reinterpret_cast<const uint32_t *>(t);
    if (type == 0) { const float *a = reinterpret_cast<const float *>(t + 8); return v3(a[0], a[1], a[2]); }
    if (type == 1) {
      const float inv = *reinterpret_cast<const float *>(t + 8);
      off = (uint32_t)(checker_even(inv, p) ? *reinterpret_cast<const uint64_t *>(t + 16) : *reinterpret_cast<const uint64_t *>(t + 24));
      continue;
    }
    if (type == 2) {                                                          // rt.cu:377-385, 343-354
      const int width = *reinterpret_cast<const int *>(t + 24), height = *reinterpret_cast<const int *>(t + 28);
      const float uc = u < 0.0f ? 0.0f : (u > 1.0f ? 1.0f : u);
      const float vc = fsub(1.0f, v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v));
      int x = (int)fmul(uc, (float)width), y = (int)fmul(vc, (float)height);
      if (*reinterpret_cast<const uint64_t *>(t + 8) == 0) return v3(1.0f, 0.0f, 1.0f);   // pixels == NULL
      x = x < 0 ? 0 : (x > width ? width : x);
      y = y < 0 ? 0 : (y > height ? height : y);
      const float *px = reinterpret_cast<const float *>(sc.tex_arena + *reinterpret_cast<const uint64_t *>(t + 16)) + 3 * (y * width + x);
      return v3(px[0], px[1], px[2]);
    }
    if (type == 3) {
      PerlinArena tab; tab.base = sc.tex_arena + *reinterpret_cast<const uint64_t *>(t + 16);
      return noise_colour(perlin_turbulence7(tab, p), *reinterpret_cast<const float *>(t + 24), p.z);
    }
    break;
  }
  return v3(0.0f, 0.0f, 0.0f);
}

// colour of the texture a lambertian / light material refers to
__device__ __forceinline__ V3 material_texture(const Scene &sc, const uint4 m0, const float4 m1, const rtt::Hit &h,
                                               const PerlinSmem *smem_perlin) {
  const uint32_t kind = (m0.x >> 8) & 0xffu;
  if (kind == rtd::TK_SOLID) return v3(m1.x, m1.y, m1.z);
  if (kind == rtd::TK_CHECKER_SOLID) {
    const float4 c0 = __ldg(sc.checkers + 2 * m0.z);
    if (checker_even(c0.x, h.p)) return v3(c0.y, c0.z, c0.w);
    const float4 c1 = __ldg(sc.checkers + 2 * m0.z + 1);
    return v3(c1.y, c1.z, c1.w);
  }
  if (kind == rtd::TK_NOISE) {
    float turb;
    if (m0.z == 0 && smem_perlin) { PerlinShared tab; tab.t = smem_perlin; turb = perlin_turbulence7(tab, h.p); }
    else { PerlinGlobal tab; tab.v = sc.perlin_vectors + 256 * m0.z; tab.p = sc.perlin_perm + 256 * m0.z; turb = perlin_turbulence7(tab, h.p); }
    return noise_colour(turb, m1.w, h.p.z);
  }
  return arena_texture(sc, m0.y, h.u, h.v, h.p);
}

__device__ __forceinline__ V3 unit_or_zero(V3 a, float len) {                 // vec3_to_unit_vec (vec3.h:105-114)
  if (len == 0.0f) return v3(0.0f, 0.0f, 0.0f);
  return v3(fdiv(a.x, len), fdiv(a.y, len), fdiv(a.z, len));
}

// material_scatter_common (rt.cu:452-490).  `U` and `rtest` are drawn by the caller for EVERY hit, whatever the
// material (rt.cu:500-501, Q5).  Returns false when the path ends here (absorbed metal bounce, or a light).
// `defer_noise` (optional): a lambertian whose texture is the shared-memory Perlin table does not evaluate it here but
// sets the flag; the render kernel then evaluates the pending textures of the whole warp cooperatively.
__device__ __forceinline__ bool scatter(const Scene &sc, const uint4 m0, const float4 m1, const V3 din, const rtt::Hit &h,
                                        const V3 U, const float rtest, const PerlinSmem *smem_perlin, V3 &dout, V3 &att,
                                        bool *defer_noise = nullptr) {
  const V3 n = h.n;
  const uint32_t type = m0.x & 0xffu;
  if (type == rtd::M_LAMBERTIAN) {
    V3 d = vadd(n, U);
    const float tiny = 1e-8f;                                                 // vec3_is_near_zero (vec3.h:177-181)
    if (fabsf(d.x) < tiny && fabsf(d.y) < tiny && fabsf(d.z) < tiny) d = n;
    dout = d;
    if (defer_noise && smem_perlin && ((m0.x >> 8) & 0xffu) == rtd::TK_NOISE && m0.z == 0) { *defer_noise = true; att = v3(0.0f, 0.0f, 0.0f); }
    else att = material_texture(sc, m0, m1, h, smem_perlin);
    return true;
  }
  if (type == rtd::M_METAL) {
    const float dn = dot_xfirst(n, din);                                      // reflect (rt.cu:424-429)
    const float two = fadd(dn, dn);
    const V3 r = v3(ffma(-n.x, two, din.x), ffma(-n.y, two, din.y), ffma(-n.z, two, din.z));
    const V3 ru = unit_or_zero(r, fsqrt(dot_yfirst(r, r)));
    dout = v3(ffma(m1.w, U.x, ru.x), ffma(m1.w, U.y, ru.y), ffma(m1.w, U.z, ru.z));
    att = v3(m1.x, m1.y, m1.z);
    return dot_yfirst(n, dout) > 0.0f;
  }
  if (type == rtd::M_DIELECTRIC) {
    att = v3(1.0f, 1.0f, 1.0f);
    const float ri = h.front ? frcp(m1.w) : m1.w;
    const V3 u = unit_or_zero(din, fsqrt(dot_xfirst(din, din)));
    const float nxux = fmul(n.x, u.x), nzuz = fmul(n.z, u.z);
    const float cos_theta = fminf(fsub(ffma(n.y, -u.y, -nxux), nzuz), 1.0f);  // fmin(dot(-u, n), 1)
    const float sin_theta = fsqrt(ffma(-cos_theta, cos_theta, 1.0f));
    bool reflect_it = fmul(ri, sin_theta) > 1.0f;
    if (!reflect_it) {                                                        // Schlick reflectance (rt.cu:431-437)
      const float r0 = fdiv(fsub(1.0f, ri), fadd(ri, 1.0f));
      const float r0sq = fmul(r0, r0);
      reflect_it = ffma(fsub(1.0f, r0sq), powf(fsub(1.0f, cos_theta), 5.0f), r0sq) > rtest;
    }
    if (reflect_it) {
      const float dn = fadd(nzuz, ffma(n.y, u.y, nxux));
      const float two = fadd(dn, dn);
      dout = v3(ffma(-n.x, two, u.x), ffma(-n.y, two, u.y), ffma(-n.z, two, u.z));
    } else {                                                                  // refract (rt.cu:415-422)
      const V3 perp = v3(fmul(ffma(n.x, cos_theta, u.x), ri), fmul(ffma(n.y, cos_theta, u.y), ri), fmul(ffma(n.z, cos_theta, u.z), ri));
      const float plen = fsqrt(dot_yfirst(perp, perp));
      const float k = fsqrt(fabsf(ffma(-plen, plen, 1.0f)));
      dout = v3(ffma(-n.x, k, perp.x), ffma(-n.y, k, perp.y), ffma(-n.z, k, perp.z));
    }
    return true;
  }
  return false;                                                               // DIFFUSE_LIGHT: no scatter
}

}  // namespace rts
#endif
