/* oracle/rt_oracle.c -- TEST INFRASTRUCTURE (CPU oracle), not product code.  See rt_oracle.h.
 *
 * Each function cites the reference lines it restates (rt.cu = /root/reference/src/raytracing.cu).
 * Build: gcc -O2 -ffp-contract=off (so that `a*b + c` is two roundings) -fopenmp.
 */
#include "rt_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------------------------------------
 * FMA profile.  FMA(a,b,c) is `a*b + c`: rounded twice in profile 0 (reference host semantics),
 * fused in profile 1 (where the reference's compiled sm_100 kernel has an FFMA).  Plain `*`, `+`, `-`
 * below are separately rounded in both profiles (FMUL/FADD in the reference's SASS).
 * ---------------------------------------------------------------------------------------------- */
static int g_gpu_profile = 0;
void orc_set_fma_profile(int gpu_profile) { g_gpu_profile = gpu_profile ? 1 : 0; }
int orc_get_fma_profile(void) { return g_gpu_profile; }
static inline float FMA(float a, float b, float c) { return g_gpu_profile ? fmaf(a, b, c) : (a * b + c); }

typedef struct { float x, y, z; } V3;
static inline V3 v3(float x, float y, float z) { V3 r = {x, y, z}; return r; }
static inline V3 ld3(const float *p) { return v3(p[0], p[1], p[2]); }
static inline void st3(float *p, V3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }
static inline V3 vadd(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }      /* vec3.h:22-29 */
static inline V3 vsub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }      /* vec3.h:31-38 */
/* vec3_dot_prod (vec3.h:88-91) is (x*x' + y*y') + z*z'.  The compiled reference fuses it as
 * fma(z,z', fma(x,x', y*y')) at most call sites ("y-first") and as fma(z,z', fma(y,y', x*x')) at a few
 * ("x-first"); in profile 0 both equal the plain left-to-right sum because f32 add commutes. */
static inline float dot_yfirst(V3 a, V3 b) { return FMA(a.z, b.z, FMA(a.x, b.x, a.y * b.y)); }
static inline float dot_xfirst(V3 a, V3 b) { return FMA(a.z, b.z, FMA(a.y, b.y, a.x * b.x)); }
/* vec3_cross_product (vec3.h:93-100): each component a*b - c*d compiles to fma(a, b, -(c*d)) */
static inline V3 cross(V3 a, V3 b) {
  return v3(FMA(a.y, b.z, -(a.z * b.y)), FMA(a.z, b.x, -(a.x * b.z)), FMA(a.x, b.y, -(a.y * b.x)));
}

/* ---- byte-offset accessors into the reference's records (layout: rt.cu:1-241) ---- */
#define HITTABLE_BYTES 120
static inline const uint8_t *hittable_at(const orc_scene *s, uint32_t idx) {
  return (const uint8_t *)s->hittables + (size_t)idx * HITTABLE_BYTES;
}
static inline uint32_t U32(const uint8_t *p, int off) { uint32_t v; memcpy(&v, p + off, 4); return v; }
static inline int32_t I32(const uint8_t *p, int off) { int32_t v; memcpy(&v, p + off, 4); return v; }
static inline uint16_t U16(const uint8_t *p, int off) { uint16_t v; memcpy(&v, p + off, 2); return v; }
static inline uint64_t U64(const uint8_t *p, int off) { uint64_t v; memcpy(&v, p + off, 8); return v; }
static inline float F32(const uint8_t *p, int off) { float v; memcpy(&v, p + off, 4); return v; }
static inline V3 VEC(const uint8_t *p, int off) { return v3(F32(p, off), F32(p, off + 4), F32(p, off + 8)); }

enum { OBJ_SPHERE = 0, OBJ_BVH = 1, OBJ_TRIANGLE = 2, OBJ_QUAD = 3, OBJ_TRANSLATION = 4, OBJ_YROT = 5, OBJ_STRUCT = 6 };
enum { MAT_LAMBERT = 0, MAT_METAL = 1, MAT_DIELECTRIC = 2, MAT_LIGHT = 3 };
enum { TEX_SOLID = 0, TEX_CHECKER = 1, TEX_IMAGE = 2, TEX_NOISE = 3 };

/* ================================================================================================
 * cuRAND XORWOW (third-party; CUDA Toolkit 12.9 curand_kernel.h -- not under /root/reference).
 * Published algorithm: Marsaglia's xorwow (five 32-bit xorshift words + a Weyl counter, step 362437)
 * [curand_kernel.h:863-874]; seeding scrambles (seed_lo, seed_hi) with two salts and two odd
 * multipliers [772-790]; subsequence s starts s*2^67 steps ahead, offset o another o steps, computed
 * with GF(2) matrix powers of the 160-bit linear state transition [646-700, 590-644].  Reference call
 * sites: curand_init(1234, pixel, 0) rt.cu:1604, curand_uniform rt.cu:1573-1578, 501, utility.h:36.
 * KATs: SURVEY.md Appendix C.3, tests/golden/xorwow.json (from the toolkit header compiled for the host).
 * ============================================================================================== */
#define XW 5                      /* state words */
#define XROWS (XW * 32)           /* 160 */
typedef struct { uint32_t row[XROWS][XW]; } xmat;   /* row r = image of unit vector e_r */

static inline uint32_t xorwow_step(uint32_t *st /* d, v0..v4 */) {
  uint32_t *v = st + 1;
  uint32_t t = v[0] ^ (v[0] >> 2);
  v[0] = v[1]; v[1] = v[2]; v[2] = v[3]; v[3] = v[4];
  v[4] = (v[4] ^ (v[4] << 4)) ^ (t ^ (t << 1));
  st[0] += 362437u;
  return v[4] + st[0];
}
static void xmat_apply(const xmat *m, uint32_t *v) {
  uint32_t r[XW] = {0};
  for (int i = 0; i < XW; ++i)
    for (int j = 0; j < 32; ++j)
      if (v[i] & (1u << j))
        for (int k = 0; k < XW; ++k) r[k] ^= m->row[i * 32 + j][k];
  memcpy(v, r, sizeof(r));
}
static void xmat_square(xmat *m) {
  xmat *t = (xmat *)malloc(sizeof(xmat));
  for (int r = 0; r < XROWS; ++r) { memcpy(t->row[r], m->row[r], sizeof(t->row[r])); xmat_apply(m, t->row[r]); }
  memcpy(m, t, sizeof(xmat));
  free(t);
}
static xmat *g_step_pow2;   /* [k] = M^(2^k),      k = 0..63  (offset skip)        */
static xmat *g_seq_pow2;    /* [k] = M^(2^(67+k)), k = 0..63  (subsequence skip)   */
__attribute__((constructor)) static void xorwow_tables_init(void) {
  g_step_pow2 = (xmat *)malloc(sizeof(xmat) * 64);
  g_seq_pow2 = (xmat *)malloc(sizeof(xmat) * 64);
  xmat m;
  for (int r = 0; r < XROWS; ++r) {               /* one-step matrix from unit states */
    uint32_t st[6] = {0};
    st[1 + r / 32] = 1u << (r & 31);
    xorwow_step(st);
    memcpy(m.row[r], st + 1, sizeof(m.row[r]));
  }
  for (int k = 0; k < 64; ++k) { g_step_pow2[k] = m; xmat_square(&m); }
  /* m is now M^(2^64) */
  xmat_square(&m); xmat_square(&m); xmat_square(&m);   /* 2^67 */
  for (int k = 0; k < 64; ++k) { g_seq_pow2[k] = m; xmat_square(&m); }
}
void orc_xorwow_init(uint64_t seed, uint64_t subsequence, uint64_t offset, uint32_t *st) {
  uint32_t s0 = ((uint32_t)seed) ^ 0xaad26b49u;
  uint32_t s1 = (uint32_t)(seed >> 32) ^ 0xf7dcefddu;
  uint32_t t0 = 1099087573u * s0;
  uint32_t t1 = 2591861531u * s1;
  st[0] = 6615241u + t1 + t0;
  st[1] = 123456789u + t0;
  st[2] = 362436069u ^ t0;
  st[3] = 521288629u + t1;
  st[4] = 88675123u ^ t1;
  st[5] = 5783321u + t0;
  for (int k = 0; k < 64; ++k) if (subsequence & (1ull << k)) xmat_apply(&g_seq_pow2[k], st + 1);
  for (int k = 0; k < 64; ++k) if (offset & (1ull << k)) xmat_apply(&g_step_pow2[k], st + 1);
  st[0] += 362437u * (uint32_t)offset;           /* the 2^67 spacing is a multiple of 2^32: no d update */
}
/* curand_uniform (curand_uniform.h:69-72): x * 2^-32 + 2^-33, in (0, 1]; the product is exact */
static inline float rng_uniform(uint32_t *st) {
  return FMA((float)xorwow_step(st), 2.3283064365386963e-10f, 1.1641532182693481e-10f);
}
void orc_xorwow_uniforms(uint64_t seed, uint64_t subsequence, uint64_t offset, int n, float *out, uint32_t *state6) {
  uint32_t st[6];
  orc_xorwow_init(seed, subsequence, offset, st);
  for (int i = 0; i < n; ++i) out[i] = rng_uniform(st);
  if (state6) memcpy(state6, st, sizeof(st));
}
/* cuda_random_f32_bound(rng, -1, 1) (utility.h:34-37): min + (max-min)*u */
static inline float rng_pm1(uint32_t *st) { return FMA(rng_uniform(st), 2.0f, -1.0f); }

/* cuda_vec3_random_unit_vector (vec3.h:152-161): draws x, y, z left to right (device order, Q5) */
static V3 rng_unit_vector(uint32_t *st) {
  for (;;) {
    float x = rng_pm1(st), y = rng_pm1(st), z = rng_pm1(st);
    float len = sqrtf(FMA(z, z, FMA(x, x, y * y)));
    float l2 = len * len;
    if (1e-50 < (double)l2 && l2 <= 1.0f) {
      float inv = 1.0f / len;
      return v3(x * inv, y * inv, z * inv);
    }
  }
}
/* cuda_random_in_unit_disk (rt.cu:1459-1469) */
static void rng_unit_disk(uint32_t *st, float *px, float *py) {
  for (;;) {
    float x = rng_pm1(st), y = rng_pm1(st);
    float len = sqrtf(FMA(x, x, y * y) + 0.0f);
    if (len * len < 1.0f) { *px = x; *py = y; return; }
  }
}

/* ================================================================================================
 * Textures and Perlin noise
 * ============================================================================================== */
/* perlin_noise_val + perlin_interp (perlin_noise.cpp:1-46).  NB: perm_x hashes all three axes (37-39). */
static float perlin_noise_at(const uint8_t *noise, V3 p) {
  float fx = floorf(p.x), fy = floorf(p.y), fz = floorf(p.z);
  float u = p.x - fx, v = p.y - fy, w = p.z - fz;
  int32_t i = (int32_t)fx, j = (int32_t)fy, k = (int32_t)fz;
  const int PERM_X = 3072;
  float uu = u * u * (3 - 2 * u), vv = v * v * (3 - 2 * v), ww = w * w * (3 - 2 * w);
  float accum = 0.0f;
  for (int di = 0; di < 2; ++di)
    for (int dj = 0; dj < 2; ++dj)
      for (int dk = 0; dk < 2; ++dk) {
        int32_t h = I32(noise, PERM_X + 4 * ((i + di) & 255)) ^ I32(noise, PERM_X + 4 * ((j + dj) & 255)) ^
                    I32(noise, PERM_X + 4 * ((k + dk) & 255));
        V3 c = VEC(noise, 12 * h);
        V3 wv = v3(u - (float)di, v - (float)dj, w - (float)dk);
        float wi = di * uu + (1 - di) * (1 - uu);
        float wj = dj * vv + (1 - dj) * (1 - vv);
        float wk = dk * ww + (1 - dk) * (1 - ww);
        accum = FMA(wi * wj * wk, dot_yfirst(c, wv), accum);
      }
  return accum;
}
static float perlin_turbulence_at(const uint8_t *noise, V3 p, int depth) {      /* perlin_noise.cpp:72-85 */
  float accum = 0.0f, weight = 1.0f;
  for (int i = 0; i < depth; ++i) {
    accum = FMA(weight, perlin_noise_at(noise, p), accum);
    weight *= 0.5f;
    p = v3(p.x * 2, p.y * 2, p.z * 2);
  }
  return fabsf(accum);
}
#define CLAMP_(val, lo, hi) ((val) < (lo) ? (lo) : ((val) > (hi) ? (hi) : (val)))   /* utility.h:8 */
static V3 texture_color(const orc_scene *s, uint64_t off, float u, float v, V3 p) {  /* rt.cu:356-398 */
  const uint8_t *t = s->texture_arena + off;
  switch (U32(t, 0)) {
    case TEX_SOLID: return VEC(t, 8);
    case TEX_CHECKER: {
      float inv = F32(t, 8);
      int32_t x = (int32_t)floorf(inv * p.x), y = (int32_t)floorf(inv * p.y), z = (int32_t)floorf(inv * p.z);
      int even = (x + y + z) % 2 == 0;            /* C '%': -1 % 2 == -1, i.e. "odd" */
      return texture_color(s, even ? U64(t, 16) : U64(t, 24), u, v, p);
    }
    case TEX_IMAGE: {                             /* rt.cu:377-385, 343-354 */
      int32_t width = I32(t, 24), height = I32(t, 28);
      u = CLAMP_(u, 0, 1);
      v = 1.0f - CLAMP_(v, 0, 1);
      int32_t x = (int32_t)(u * width), y = (int32_t)(v * height);
      if (U64(t, 8) == 0) return v3(1, 0, 1);     /* pixels == NULL: the missing-file magenta */
      x = CLAMP_(x, 0, width); y = CLAMP_(y, 0, height);
      return VEC(s->texture_arena + U64(t, 16), 12 * (y * width + x));
    }
    case TEX_NOISE: {                             /* rt.cu:386-395 */
      float turb = perlin_turbulence_at(s->texture_arena + U64(t, 16), p, 7);
      float c = (1 + sinf(FMA(p.z, F32(t, 24), 10 * turb))) * 0.5f;
      return v3(c, c, c);
    }
  }
  return v3(0, 0, 0);
}
void orc_texture_color(const orc_scene *s, uint64_t tex_off, float u, float v, const float *p, float *out3) {
  st3(out3, texture_color(s, tex_off, u, v, ld3(p)));
}
float orc_perlin_noise(const orc_scene *s, uint64_t noise_off, const float *p) {
  return perlin_noise_at(s->texture_arena + noise_off, ld3(p));
}
float orc_perlin_turbulence(const orc_scene *s, uint64_t noise_off, const float *p, int depth) {
  return perlin_turbulence_at(s->texture_arena + noise_off, ld3(p), depth);
}

/* ================================================================================================
 * Scene traversal
 * ============================================================================================== */
typedef struct { V3 o, d; float time; } Ray;
typedef struct { V3 point, normal; float t; int front; uint8_t mat[24]; float u, v; } Hit;

/* aabb_hit (rt.cu:989-1025).  MAX/MIN are the (a<b) selects of utility.h:6-7 with the running value as
 * first operand; 1/0 = +-inf and 0*inf = NaN flow through exactly as in the reference (Q9). */
static int aabb_hit(const float *b /* x0 x1 y0 y1 z0 z1 */, const Ray *r, float t0, float t1) {
  const float o[3] = {r->o.x, r->o.y, r->o.z}, d[3] = {r->d.x, r->d.y, r->d.z};
  for (int a = 0; a < 3; ++a) {
    float inv = 1.0f / d[a];
    float ta = (b[2 * a] - o[a]) * inv;
    float tb = (b[2 * a + 1] - o[a]) * inv;
    if (inv < 0.0f && ta > tb) { float tmp = ta; ta = tb; tb = tmp; }
    t0 = (t0 < ta) ? ta : t0;
    t1 = (t1 < tb) ? t1 : tb;
    if (t1 <= t0) return 0;
  }
  return 1;
}
int orc_aabb_hit(const float *box6, const float *origin, const float *direction, float t0, float t1) {
  Ray r; r.o = ld3(origin); r.d = ld3(direction); r.time = 0;
  return aabb_hit(box6, &r, t0, t1);
}

static inline void set_face_normal(Hit *h, float d_dot_n, V3 n) {               /* rt.cu:983-987 */
  h->front = d_dot_n < 0;
  h->normal = h->front ? n : v3(-n.x, -n.y, -n.z);
}
/* sphere_get_uv (rt.cu:583-590) */
static void sphere_uv(V3 p, float *u, float *v) {
  const float pi = 3.1415926535897932385f;
  float theta = acosf(-p.y);
  float phi = atan2f(-p.z, p.x) + pi;
  *u = phi / (2 * pi);
  *v = theta / pi;
}

/* hit() (rt.cu:1079-1251): explicit-stack closest hit over the tagged-union array.  The ray is edited in
 * place by the instance wrappers and NOT restored bit-exactly (Q12), so it is passed by pointer. */
static int scene_hit(const orc_scene *s, uint32_t root_href, float tmin, float tmax, Ray *ray, Hit *rec) {
  int result = 0, was_last_hit = 0, in_translation = 0, in_rotation = 0;
  uint16_t stack[1024]; int sp = 0;
  stack[sp++] = (uint16_t)root_href;
  while (sp > 0) {
    uint16_t cur = stack[--sp];
    const uint8_t *h = hittable_at(s, cur);
    switch (U32(h, 0)) {
      case OBJ_SPHERE: {                                                          /* rt.cu:1093-1123 */
        V3 c0 = VEC(h, 8), mv = VEC(h, 20); float radius = F32(h, 32);
        V3 c = v3(FMA(ray->time, mv.x, c0.x), FMA(ray->time, mv.y, c0.y), FMA(ray->time, mv.z, c0.z));
        V3 oc = vsub(c, ray->o);
        float a = dot_xfirst(ray->d, ray->d);
        float b = -2.0f * dot_yfirst(ray->d, oc);
        float cc = FMA(-radius, radius, dot_yfirst(oc, oc));
        float disc = FMA(b, b, (a * -4.0f) * cc);
        if (disc >= 0) {
          float sq = sqrtf(disc);
          float root = (-b - sq) / (a + a);
          if (tmin >= root || root >= tmax) root = (sq - b) / (a + a);
          if (tmin < root && root < tmax) {
            rec->t = root;
            rec->point = v3(FMA(ray->d.x, root, ray->o.x), FMA(ray->d.y, root, ray->o.y), FMA(ray->d.z, root, ray->o.z));
            memcpy(rec->mat, h + 40, 24);
            float inv_r = 1.0f / radius;
            V3 n = v3((rec->point.x - c.x) * inv_r, (rec->point.y - c.y) * inv_r, (rec->point.z - c.z) * inv_r);
            sphere_uv(n, &rec->u, &rec->v);
            set_face_normal(rec, dot_yfirst(ray->d, n), n);
            tmax = rec->t; result = 1;
            if (in_translation || in_rotation) was_last_hit = 1;
          }
        }
      } break;
      case OBJ_BVH: {                                                             /* rt.cu:1124-1130 */
        float box[6]; memcpy(box, h + 12, 24);
        if (aabb_hit(box, ray, tmin, tmax)) {
          stack[sp++] = s->world[U16(h, 10)];      /* right pushed first => left visited first (Q11) */
          stack[sp++] = s->world[U16(h, 8)];
        }
      } break;
      case OBJ_TRIANGLE: case OBJ_QUAD: {                                         /* rt.cu:1131-1175, 1060-1077 */
        V3 q = VEC(h, 8), eu = VEC(h, 20), ev = VEC(h, 32), w = VEC(h, 48), n = VEC(h, 60); float d = F32(h, 44);
        float alpha = INFINITY, beta = INFINITY, t = 0.0f; V3 p = v3(0, 0, 0);
        float denom = dot_yfirst(n, ray->d);
        if ((double)fabsf(denom) >= 1e-8) {
          t = (d - dot_yfirst(n, ray->o)) / denom;
          if (tmin <= t && t <= tmax) {
            p = v3(FMA(ray->d.x, t, ray->o.x), FMA(ray->d.y, t, ray->o.y), FMA(ray->d.z, t, ray->o.z));
            V3 pl = vsub(p, q);
            alpha = dot_yfirst(w, cross(pl, ev));
            beta = dot_yfirst(w, cross(eu, pl));
          }
        }
        int inside = (0.0f <= alpha && alpha <= 1.0f) && (0.0f <= beta && beta <= 1.0f);
        if (U32(h, 0) == OBJ_TRIANGLE) inside = inside && (alpha + beta <= 1.0f);
        if (inside) {
          rec->u = alpha; rec->v = beta; rec->t = t; rec->point = p;
          memcpy(rec->mat, h + 72, 24);
          set_face_normal(rec, denom, n);
          tmax = rec->t; result = 1;
          if (in_translation || in_rotation) was_last_hit = 1;
        }
      } break;
      case OBJ_TRANSLATION: {                                                     /* rt.cu:1176-1191 */
        V3 off = VEC(h, 12);
        if (in_translation == 0) {
          was_last_hit = 0;
          ray->o = vsub(ray->o, off);
          stack[sp++] = cur; stack[sp++] = U16(h, 8);
          in_translation++;
        } else {
          ray->o = vadd(ray->o, off);
          in_translation--;
          if (was_last_hit) rec->point = vadd(rec->point, off);
        }
      } break;
      case OBJ_YROT: {                                                            /* rt.cu:1192-1237 */
        float sn = F32(h, 12), cs = F32(h, 16);
        if (in_rotation == 0) {
          was_last_hit = 0;
          V3 o = ray->o, d = ray->d;
          ray->o = v3(FMA(o.x, cs, -(o.z * sn)), o.y, FMA(o.x, sn, o.z * cs));
          ray->d = v3(FMA(d.x, cs, -(d.z * sn)), d.y, FMA(d.z, cs, d.x * sn));
          stack[sp++] = cur; stack[sp++] = U16(h, 8);
          in_rotation++;
        } else {
          V3 o = ray->o, d = ray->d;
          ray->o = v3(FMA(o.x, cs, o.z * sn), o.y, FMA(o.z, cs, -(o.x * sn)));
          ray->d = v3(FMA(d.x, cs, d.z * sn), d.y, FMA(d.z, cs, -(d.x * sn)));
          in_rotation--;
          if (was_last_hit) {
            V3 p = rec->point, n = rec->normal;
            rec->point = v3(FMA(p.x, cs, p.z * sn), p.y, FMA(p.z, cs, -(p.x * sn)));
            rec->normal = v3(FMA(n.x, cs, n.z * sn), n.y, FMA(n.z, cs, -(n.x * sn)));
          }
        }
      } break;
      case OBJ_STRUCT: {                                                          /* rt.cu:1238-1247 */
        uint16_t first = U16(h, 8); uint64_t count = U64(h, 48);
        for (uint16_t i = 0; i < count; ++i) stack[sp++] = (uint16_t)(first + i);
      } break;
    }
  }
  return result;
}

static void hit_to_record(const Hit *h, orc_hit_record *r) {
  st3(r->point, h->point); st3(r->normal, h->normal); r->t = h->t; r->front = h->front;
  memcpy(r->mat, h->mat, 24); r->u = h->u; r->v = h->v;
}
int orc_hit(const orc_scene *s, float *origin, float *direction, float time, float tmin, float tmax, orc_hit_record *out) {
  Ray ray; ray.o = ld3(origin); ray.d = ld3(direction); ray.time = time;
  Hit h; memset(&h, 0, sizeof(h));
  int r = scene_hit(s, s->world[s->root_wref], tmin, tmax, &ray, &h);
  st3(origin, ray.o); st3(direction, ray.d);
  hit_to_record(&h, out);
  return r;
}

/* ================================================================================================
 * Materials
 * ============================================================================================== */
static V3 material_emit(const orc_scene *s, const uint8_t *mat, float u, float v, V3 p) {   /* rt.cu:439-450 */
  if (U32(mat, 0) == MAT_LIGHT) return texture_color(s, U64(mat, 8), u, v, p);
  return v3(0, 0, 0);
}
static inline V3 unit_or_zero(V3 a, float len) {                                             /* vec3.h:105-114 */
  if (len == 0.0f) return v3(0, 0, 0);
  return v3(a.x / len, a.y / len, a.z / len);
}
/* material_scatter_common (rt.cu:452-490) with reflect/refract/reflectance (rt.cu:415-437) inlined in the
 * shape the compiled reference has them. */
static int material_scatter(const orc_scene *s, const uint8_t *mat, V3 din, const Hit *rec, V3 U, float rtest,
                            V3 *dir_out, V3 *att_out) {
  V3 n = rec->normal;
  *dir_out = v3(0, 0, 0); *att_out = v3(0, 0, 0);
  switch (U32(mat, 0)) {
    case MAT_LAMBERT: {
      V3 d = vadd(n, U);
      const float tiny = 1e-8;                                                    /* vec3.h:177-181 */
      if (fabsf(d.x) < tiny && fabsf(d.y) < tiny && fabsf(d.z) < tiny) d = n;
      *dir_out = d;
      *att_out = texture_color(s, U64(mat, 8), rec->u, rec->v, rec->point);
      return 1;
    }
    case MAT_METAL: {
      float dn = dot_xfirst(n, din);                    /* reflect(): d - 2*dot(d,n)*n */
      float two = dn + dn;
      V3 r = v3(FMA(-n.x, two, din.x), FMA(-n.y, two, din.y), FMA(-n.z, two, din.z));
      float len = sqrtf(dot_yfirst(r, r));
      V3 ru = unit_or_zero(r, len);
      float fuzz = F32(mat, 20);
      V3 d = v3(FMA(fuzz, U.x, ru.x), FMA(fuzz, U.y, ru.y), FMA(fuzz, U.z, ru.z));
      *dir_out = d;
      *att_out = VEC(mat, 8);
      return dot_yfirst(n, d) > 0;
    }
    case MAT_DIELECTRIC: {
      *att_out = v3(1.0f, 1.0f, 1.0f);
      float ior = F32(mat, 8);
      float ri = rec->front ? (1.0f / ior) : ior;
      float len = sqrtf(dot_xfirst(din, din));
      V3 u = unit_or_zero(din, len);
      /* dot(-u, n): fma(ny, -uy, -(nx*ux)) - nz*uz in the compiled reference */
      float nxux = n.x * u.x, nzuz = n.z * u.z;
      float cos_theta = fminf(FMA(n.y, -u.y, -nxux) - nzuz, 1.0f);
      float sin_theta = sqrtf(FMA(-cos_theta, cos_theta, 1.0f));
      int cannot_refract = ri * sin_theta > 1.0f;
      int reflect_it = cannot_refract;
      if (!reflect_it) {                                 /* reflectance(): Schlick (rt.cu:431-437) */
        float r0 = (1.0f - ri) / (1.0f + ri);
        float r0sq = r0 * r0;
        float refl = FMA(1.0f - r0sq, powf(1.0f - cos_theta, 5.0f), r0sq);
        reflect_it = refl > rtest;
      }
      if (reflect_it) {
        float dn = FMA(n.y, u.y, nxux) + nzuz;
        float two = dn + dn;
        *dir_out = v3(FMA(-n.x, two, u.x), FMA(-n.y, two, u.y), FMA(-n.z, two, u.z));
      } else {                                           /* refract() (rt.cu:415-422) */
        V3 perp = v3(FMA(n.x, cos_theta, u.x) * ri, FMA(n.y, cos_theta, u.y) * ri, FMA(n.z, cos_theta, u.z) * ri);
        float plen = sqrtf(dot_yfirst(perp, perp));
        float k = sqrtf(fabsf(FMA(-plen, plen, 1.0f)));
        *dir_out = v3(FMA(-n.x, k, perp.x), FMA(-n.y, k, perp.y), FMA(-n.z, k, perp.z));
      }
      return 1;
    }
  }
  return 0;                                              /* DIFFUSE_LIGHT: no case => is_hit = 0 */
}
int orc_scatter(const orc_scene *s, const uint32_t *mat6, const float *dir_in, const orc_hit_record *rec,
                const float *unit_vec, float random_test, float *dir_out, float *atten_out) {
  Hit h; memset(&h, 0, sizeof(h));
  h.point = ld3(rec->point); h.normal = ld3(rec->normal); h.t = rec->t; h.front = rec->front; h.u = rec->u; h.v = rec->v;
  V3 d, a;
  int r = material_scatter(s, (const uint8_t *)mat6, ld3(dir_in), &h, ld3(unit_vec), random_test, &d, &a);
  st3(dir_out, d); st3(atten_out, a);
  return r;
}
void orc_emit(const orc_scene *s, const uint32_t *mat6, float u, float v, const float *p, float *out3) {
  st3(out3, material_emit(s, (const uint8_t *)mat6, u, v, ld3(p)));
}

/* ================================================================================================
 * Path integrator and the per-pixel sample loop
 * ============================================================================================== */
/* ray_color (rt.cu:1333-1371): bounce loop pushing (attenuation, emission), then the backward unwind
 * result = att (.) result + emission, which the compiled reference evaluates as one fma per channel. */
static V3 ray_color(const orc_scene *s, uint32_t *rng, Ray *ray, V3 background, int max_bounces, uint64_t *rays) {
  V3 att_stack[256], emi_stack[256];
  int depth = 0;
  const int capacity = max_bounces & 255;                /* RayColorStack.max_count is a u8 (rt.cu:220, 1596) */
  int budget = capacity;
  V3 result = v3(0, 0, 0);
  Hit rec; memset(&rec, 0, sizeof(rec));
  uint32_t root = s->world[s->root_wref];
  while (budget > 0) {
    ++*rays;
    if (scene_hit(s, root, 0.001f, INFINITY, ray, &rec)) {
      V3 emission = material_emit(s, rec.mat, rec.u, rec.v, rec.point);
      V3 U = rng_unit_vector(rng);                       /* rt.cu:500-501: drawn for EVERY material (Q5) */
      float rtest = rng_uniform(rng);
      V3 dir, att;
      if (material_scatter(s, rec.mat, ray->d, &rec, U, rtest, &dir, &att)) {
        ray->o = rec.point; ray->d = dir;
        if (depth != capacity) { att_stack[depth] = att; emi_stack[depth] = emission; depth++; }
        budget--;
      } else { result = emission; break; }
    } else { result = background; break; }
  }
  while (depth > 0) {
    --depth;
    V3 a = att_stack[depth], e = emi_stack[depth];
    result = v3(FMA(a.x, result.x, e.x), FMA(a.y, result.y, e.y), FMA(a.z, result.z, e.z));
  }
  return result;
}

/* cuda_calc (rt.cu:1543-1589) for rows [row0,row1) */
int orc_render(const orc_scene *s, const orc_camera *c, uint64_t seed, uint64_t subseq_offset,
               int row0, int row1, float *out, uint64_t *rays_out, int nthreads) {
  const int W = c->img_width, H = c->img_height;
  if (row0 < 0) row0 = 0;
  if (row1 > H || row1 <= 0) row1 = H;
  const V3 dv = ld3(c->pixel_delta_v), du = ld3(c->pixel_delta_u), cam = ld3(c->camera_pos);
  const V3 ddu = ld3(c->defocus_disk_u), ddv = ld3(c->defocus_disk_v), p00 = ld3(c->pixel00), bg = ld3(c->background);
  uint64_t total = 0;
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#else
  (void)nthreads;
#endif
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : total)
  for (int i = row0; i < row1; ++i) {
    uint64_t rays = 0;
    for (int j = 0; j < W; ++j) {
      int idx = i * W + j;
      uint32_t rng[6];
      orc_xorwow_init(seed, (uint64_t)idx + subseq_offset, 0, rng);
      float fi = (float)i, fj = (float)j;
      V3 base = v3(p00.x + FMA(du.x, fj, dv.x * fi), p00.y + FMA(du.y, fj, dv.y * fi), p00.z + FMA(du.z, fj, dv.z * fi));
      V3 acc = v3(0, 0, 0);
      for (int sn = 0; sn < c->samples_per_pixel; ++sn) {
        Ray ray;
        if (c->defocus_angle <= 0) ray.o = cam;
        else {
          float px, py; rng_unit_disk(rng, &px, &py);
          ray.o = v3(cam.x + FMA(ddu.x, px, ddv.x * py), cam.y + FMA(ddu.y, px, ddv.y * py), cam.z + FMA(ddu.z, px, ddv.z * py));
        }
        float xo = rng_uniform(rng) - 0.5f;
        float yo = rng_uniform(rng) - 0.5f;
        V3 d = vsub(base, ray.o);
        d = v3(FMA(du.x, xo, d.x), FMA(du.y, xo, d.y), FMA(du.z, xo, d.z));
        d = v3(FMA(dv.x, yo, d.x), FMA(dv.y, yo, d.y), FMA(dv.z, yo, d.z));
        ray.d = d;
        ray.time = rng_uniform(rng);
        V3 col = ray_color(s, rng, &ray, bg, c->max_bounces, &rays);
        acc = vadd(acc, col);
      }
      acc = v3(acc.x * c->pixel_samples_scale, acc.y * c->pixel_samples_scale, acc.z * c->pixel_samples_scale);
      st3(out + 3 * (size_t)idx, acc);
    }
    total += rays;
  }
  if (rays_out) *rays_out = total;
  return 0;
}

void orc_resolve_rgb8(const float *px, uint64_t n, uint8_t *rgb8) {              /* rt.cu:1391-1433, utility.h:43-49 */
  for (uint64_t i = 0; i < 3 * n; ++i) {
    float c = px[i];
    float g = (c > 0) ? sqrtf(c) : 0.0f;
    if (g < 0.0f) g = 0.0f;
    if (g > 0.999f) g = 0.999f;
    rgb8[i] = (uint8_t)(256 * g);
  }
}

int orc_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
