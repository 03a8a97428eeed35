// rt_trace.cuh -- ray/primitive intersection and the two closest-hit traversals.
//
// Replaces hit() / aabb_hit() / geom_calc_hit() of the reference (rt.cu:989-1025, 1060-1077, 1079-1251).
// Arithmetic that decides or produces a hit follows the reference kernel's roundings (rt_math.cuh).
#ifndef RT_TRACE_CUH
#define RT_TRACE_CUH

#include "rt_device.h"
#include "rt_math.cuh"

// a push onto the per-thread traversal stack; -DRT_STACK_CHECK builds trap when a walk needs more entries than the packer
// computed for the scene (tools/build_variants.sh stackcheck "-DRT_STACK_CHECK", then the GPU tests with RT_B200_LIB)
// The stack is walked by ADDRESS (`top`), not by index: a push is one store and one add, a pop one add and one load, and the
// bottom is a sentinel entry instead of a compare against the base -- with an index ptxas rebuilt the shared window's base
// from special registers at every push and pop (3 % of the warp-instructions of C2 and C5).
#ifdef RT_STACK_CHECK
#define RT_PUSH(v) do { if (top >= stk + (size_t)sc.stack_entries * stride) __trap(); *top = (v); top += stride; } while (0)
#else
#define RT_PUSH(v) (*top = (v), top += stride)
#endif
#define RT_POP() (top -= stride, *top)

namespace rtt {
using namespace rtm;
constexpr uint32_t STACK_BOTTOM = 0xffffffffu;             // no link looks like this (kind K_NONE, instance bit, index 2^28 - 1)
using rtd::Scene;

struct Ray { V3 o, d; float time; };

struct Hit {
  V3 p, n;          // point, normal (already flipped against the ray)
  float t, u, v;
  uint32_t prim;    // primitive id: sphere index, or sphere_count + planar index
  int front;
};

__device__ __forceinline__ float4 ldg4(const float4 *p) { return __ldg(p); }
// G = the record may have been staged in shared memory (kernel modes 3 / 4 with RT_LINEAR_SMEM): a generic load finds it
// wherever it is; everything else reads global memory through the read-only path
template <bool G> __device__ __forceinline__ float4 ldx(const float4 *p) { return G ? *p : __ldg(p); }

// Per-ray operation counters for the roofline's F_ray; NoCount compiles to nothing.
struct NoCount {
  __device__ __forceinline__ void box(int) {}
  __device__ __forceinline__ void sphere() {}
  __device__ __forceinline__ void planar() {}
  __device__ __forceinline__ void translation() {}
  __device__ __forceinline__ void rotation() {}
};
struct OpCount {
  unsigned long long n_box = 0, n_sphere = 0, n_planar = 0, n_translation = 0, n_rotation = 0;
  __device__ __forceinline__ void box(int k) { n_box += k; }
  __device__ __forceinline__ void sphere() { ++n_sphere; }
  __device__ __forceinline__ void planar() { ++n_planar; }
  __device__ __forceinline__ void translation() { ++n_translation; }
  __device__ __forceinline__ void rotation() { ++n_rotation; }
};

// ---- sphere (rt.cu:1093-1123): a = d.d is passed in (it only changes when a wrapper rotates the ray) ----
// The test (which root, if any, lies in (tmin,tmax)) and the hit record are separate so that the warp-scheduled
// kernel can carry only (t, prim) through a traversal and build the record once, for the closest hit.
// `rec` = the sphere's two words, wherever they are (Scene::spheres, or the image of a linear program)
template <bool G = false>
__device__ __forceinline__ bool sphere_root_rec(const float4 *rec, const Ray &r, float a, float tmin, float tmax, float &root_out) {
  const float4 s0 = ldx<G>(rec);
  const float4 s1 = ldx<G>(rec + 1);
  const V3 c = v3(ffma(r.time, s1.x, s0.x), ffma(r.time, s1.y, s0.y), ffma(r.time, s1.z, s0.z));   // movement_path_at
  const V3 oc = vsub(c, r.o);
  const float b = fmul(dot_yfirst(r.d, oc), -2.0f);
  const float cc = ffma(-s0.w, s0.w, dot_yfirst(oc, oc));
  const float disc = ffma(b, b, fmul(cc, fmul(a, -4.0f)));
  if (!(disc >= 0.0f)) return false;
  const float sq = fsqrt(disc);
  const float a2 = fadd(a, a);
  float root = fdiv(fsub(-b, sq), a2);
  if (tmin >= root || root >= tmax) root = fdiv(fsub(sq, b), a2);
  if (!(tmin < root && root < tmax)) return false;
  root_out = root;
  return true;
}
template <bool G = false>
__device__ __forceinline__ bool sphere_root(const Scene &sc, uint32_t idx, const Ray &r, float a, float tmin, float tmax,
                                            float &root_out) {
  return sphere_root_rec<G>(sc.spheres + 2 * idx, r, a, tmin, tmax, root_out);
}
template <bool G = false>
__device__ __forceinline__ void sphere_record_rec(const float4 *rec, uint32_t idx, const Ray &r, float root, Hit &h) {
  const float4 s0 = ldx<G>(rec);
  const float4 s1 = ldx<G>(rec + 1);
  const V3 c = v3(ffma(r.time, s1.x, s0.x), ffma(r.time, s1.y, s0.y), ffma(r.time, s1.z, s0.z));
  h.t = root;
  h.p = along(r.o, root, r.d);
  const float inv_r = frcp(s0.w);
  const V3 pc = vsub(h.p, c);
  const V3 n = v3(fmul(pc.x, inv_r), fmul(pc.y, inv_r), fmul(pc.z, inv_r));
  if (__float_as_uint(s1.w) & 1u) {                       // sphere_get_uv (rt.cu:583-590): only image textures read it
    const float pi = 3.1415926535897932385f;
    const float theta = acosf(-n.y);
    const float phi = fadd(atan2f(-n.z, n.x), pi);
    h.u = fdiv(phi, fmul(2.0f, pi));
    h.v = fdiv(theta, pi);
  }
  h.front = dot_yfirst(r.d, n) < 0.0f;                     // hit_record_set_face_normal (rt.cu:983-987)
  h.n = h.front ? n : vneg(n);
  h.prim = idx;
}
template <bool G = false>
__device__ __forceinline__ void sphere_record(const Scene &sc, uint32_t idx, const Ray &r, float root, Hit &h) {
  sphere_record_rec<G>(sc.spheres + 2 * idx, idx, r, root, h);
}
template <bool G = false>
__device__ __forceinline__ bool hit_sphere_rec(const float4 *rec, uint32_t idx, const Ray &r, float a, float tmin, float &tmax, Hit &h) {
  float root;
  if (!sphere_root_rec<G>(rec, r, a, tmin, tmax, root)) return false;
  sphere_record_rec<G>(rec, idx, r, root, h);
  tmax = root;
  return true;
}
template <bool G = false>
__device__ __forceinline__ bool hit_sphere(const Scene &sc, uint32_t idx, const Ray &r, float a, float tmin, float &tmax,
                                           Hit &h) {
  return hit_sphere_rec<G>(sc.spheres + 2 * idx, idx, r, a, tmin, tmax, h);
}

// ---- quad / triangle (rt.cu:1131-1175 with geom_calc_hit 1060-1077) ----
// EAGER requests the four 16-byte words of the record together, before the plane test decides whether the last three
// are needed: one exposed load latency per test instead of two, at the price of three loads that most tests do not
// use.  Measured: pays in the walk with treelets / fused instances (scene 3 +3 %), costs in the plain reference-order
// walk (Cornell box -5 %), so only the former asks for it.
__device__ __forceinline__ bool planar_inside_rec(const float4 p1, const float4 p2, const float4 p3, bool triangle, const Ray &r,
                                                  float t, V3 &p, float &alpha, float &beta) {
  p = along(r.o, t, r.d);
  const V3 pl = vsub(p, v3(p1.x, p1.y, p1.z));
  const V3 w = v3(p1.w, p2.w, p3.w);
  alpha = dot_yfirst(w, cross(pl, v3(p2.x, p2.y, p2.z)));
  beta = dot_yfirst(w, cross(v3(p3.x, p3.y, p3.z), pl));
  bool inside = (0.0f <= alpha && alpha <= 1.0f) && (0.0f <= beta && beta <= 1.0f);
  if (triangle) inside = inside && (fadd(alpha, beta) <= 1.0f);
  return inside;
}
// `rec` = the four words of the record, wherever they are (Scene::planars, or the image of a linear program)
template <bool G = false>
__device__ __forceinline__ bool planar_inside(const float4 *rec, bool triangle, const Ray &r, float t, V3 &p, float &alpha, float &beta) {
  const float4 p1 = ldx<G>(rec + 1);                       // q, w.x
  const float4 p2 = ldx<G>(rec + 2);                       // v, w.y
  const float4 p3 = ldx<G>(rec + 3);                       // u, w.z
  return planar_inside_rec(p1, p2, p3, triangle, r, t, p, alpha, beta);
}
// Axis-aligned quad (reduced record {n,d} {q, w.k} {S_alpha} {S_beta}; rt_pack.cpp axis_aligned_words has the
// derivation): with one non-zero component in each of w, u, v the reference's two cross products and dot products reduce,
// rounding for rounding, to alpha = fl(w.k * fl(pl.m * s)) -- fl(pl.m * s) evaluated as dot(pl, S), S zero except s at m,
// so that no component has to be picked at run time.  8 instead of 18 operations for the two coordinates.
template <bool G = false>
__device__ __forceinline__ bool planar_inside_axis(const float4 *rec, const Ray &r, float t, V3 &p, float &alpha, float &beta) {
  const float4 p1 = ldx<G>(rec + 1);                       // q, w.k
  const float4 p2 = ldx<G>(rec + 2);                       // S_alpha
  const float4 p3 = ldx<G>(rec + 3);                       // S_beta
  p = along(r.o, t, r.d);
  const V3 pl = vsub(p, v3(p1.x, p1.y, p1.z));
  alpha = fmul(p1.w, dot_yfirst(pl, v3(p2.x, p2.y, p2.z)));
  beta = fmul(p1.w, dot_yfirst(pl, v3(p3.x, p3.y, p3.z)));
  return (0.0f <= alpha && alpha <= 1.0f) && (0.0f <= beta && beta <= 1.0f);
}
// `p0` = first word of the record (normal, d): the linear program carries it inside the op, the other walks load it here
template <bool EAGER = false, bool G = false, bool AX = false>
__device__ __forceinline__ bool planar_t_p0(const float4 *rec, const float4 p0, bool triangle, const Ray &r, float tmin, float tmax,
                                            float &t_out) {
  float4 p1 = p0, p2 = p0, p3 = p0;
  if (EAGER) { p1 = ldg4(rec + 1); p2 = ldg4(rec + 2); p3 = ldg4(rec + 3); }
  const V3 n = v3(p0.x, p0.y, p0.z);
  const float denom = dot_yfirst(n, r.d);
  // reference: fabsf(denom) >= 1e-8 evaluated in double; (float)1e-8 < 1e-8, so in f32 this is a strict '>'
  if (!(fabsf(denom) > 1e-8f)) return false;
  // A ray that starts ON the plane (every bounce off a wall, about every other time) has a numerator of exactly 0.  0 / x is
  // a signed zero, but FCHK sends a zero numerator to the slow-path subroutine of div.rn.f32 and the whole warp waits for
  // it (measured: 25 % of the warp-level executions of this division on the Cornell box).  Those lanes divide 1 instead
  // and get their signed zero from the operands' signs: same bits, no call.
  const float num = fsub(p0.w, dot_yfirst(n, r.o));
  const bool zero_num = num == 0.0f;
  const float quot = fdiv(zero_num ? 1.0f : num, denom);
  const float t = zero_num ? __uint_as_float((__float_as_uint(num) ^ __float_as_uint(denom)) & 0x80000000u) : quot;
  if (!(tmin <= t && t <= tmax)) return false;
  V3 p; float alpha, beta;
  if (AX) { if (!planar_inside_axis<G>(rec, r, t, p, alpha, beta)) return false; }
  else if (EAGER ? !planar_inside_rec(p1, p2, p3, triangle, r, t, p, alpha, beta) : !planar_inside<G>(rec, triangle, r, t, p, alpha, beta)) return false;
  t_out = t;
  return true;
}
template <bool EAGER = false>
__device__ __forceinline__ bool planar_t(const Scene &sc, uint32_t idx, bool triangle, const Ray &r, float tmin, float tmax,
                                         float &t_out) {
  return planar_t_p0<EAGER>(sc.planars + 4 * idx, ldg4(sc.planars + 4 * idx), triangle, r, tmin, tmax, t_out);
}
// `prim` = the primitive id of the planar (sphere_count + its index)
template <bool G = false, bool AX = false>
__device__ __forceinline__ void planar_record_p0(const float4 *rec, const float4 p0, uint32_t prim, const Ray &r, float t, Hit &h) {
  const V3 n = v3(p0.x, p0.y, p0.z);
  const float denom = dot_yfirst(n, r.d);
  float alpha, beta;
  if (AX) planar_inside_axis<G>(rec, r, t, h.p, alpha, beta);
  else planar_inside<G>(rec, false, r, t, h.p, alpha, beta);
  h.u = alpha; h.v = beta; h.t = t;
  h.front = denom < 0.0f;
  h.n = h.front ? n : vneg(n);
  h.prim = prim;
}
__device__ __forceinline__ void planar_record(const Scene &sc, uint32_t idx, const Ray &r, float t, Hit &h) {
  planar_record_p0(sc.planars + 4 * idx, ldg4(sc.planars + 4 * idx), sc.sphere_count + idx, r, t, h);
}
template <bool EAGER = false>
__device__ __forceinline__ bool hit_planar(const Scene &sc, uint32_t idx, bool triangle, const Ray &r, float tmin, float &tmax,
                                           Hit &h) {
  float t;
  if (!planar_t<EAGER>(sc, idx, triangle, r, tmin, tmax, t)) return false;
  planar_record(sc, idx, r, t, h);
  tmax = t;
  return true;
}
template <bool EAGER = false, bool G = false, bool AX = false>
__device__ __forceinline__ bool hit_planar_p0(const float4 *rec, const float4 p0, uint32_t prim, bool triangle, const Ray &r, float tmin,
                                              float &tmax, Hit &h) {
  float t;
  if (!planar_t_p0<EAGER, G, AX>(rec, p0, triangle, r, tmin, tmax, t)) return false;
  planar_record_p0<G, AX>(rec, p0, prim, r, t, h);
  tmax = t;
  return true;
}

// ---- aabb_hit exactly as the reference (rt.cu:989-1025): (a<b) selects, per-axis early-out, NaNs flow through ----
__device__ __forceinline__ bool aabb_axis(float lo, float hi, float o, float inv, float &t0, float &t1) {
  float ta = fmul(fsub(lo, o), inv);
  float tb = fmul(fsub(hi, o), inv);
  if (inv < 0.0f && ta > tb) { float s = ta; ta = tb; tb = s; }
  t0 = (t0 < ta) ? ta : t0;
  t1 = (t1 < tb) ? t1 : tb;
  return !(t1 <= t0);
}

// ---- conservative slab test of the two child boxes of an own node (64 B: n0 = c0 x,y; n1 = c1 x,y; n2 = both z) ----
// Planes as one fma each: t = plane * inv - o * inv.  The rounding of o * inv moves a plane by up to |o * inv| * 2^-24
// in t, so the overlap test gets 8x that much slack (and a relative one for the fma itself): the cull stays
// conservative, which is all that exactness of the closest hit needs.  1/d is clamped to a finite value so that a
// ray parallel to a slab gets -huge/+huge plane distances instead of inf - inf; fminf/fmaxf drop NaNs (0 * inf).
__device__ __forceinline__ void slab_pair(const float4 n0, const float4 n1, const float4 n2, const V3 inv, const V3 noi,
                                          const float slack, const float tmin, const float tmax, float &e0, float &e1,
                                          bool &h0, bool &h1) {
  const float c0x0 = fmaf(n0.x, inv.x, noi.x), c0x1 = fmaf(n0.y, inv.x, noi.x);
  const float c0y0 = fmaf(n0.z, inv.y, noi.y), c0y1 = fmaf(n0.w, inv.y, noi.y);
  const float c0z0 = fmaf(n2.x, inv.z, noi.z), c0z1 = fmaf(n2.y, inv.z, noi.z);
  const float c1x0 = fmaf(n1.x, inv.x, noi.x), c1x1 = fmaf(n1.y, inv.x, noi.x);
  const float c1y0 = fmaf(n1.z, inv.y, noi.y), c1y1 = fmaf(n1.w, inv.y, noi.y);
  const float c1z0 = fmaf(n2.z, inv.z, noi.z), c1z1 = fmaf(n2.w, inv.z, noi.z);
  e0 = fmaxf(fmaxf(fminf(c0x0, c0x1), fminf(c0y0, c0y1)), fmaxf(fminf(c0z0, c0z1), tmin));
  float x0 = fminf(fminf(fmaxf(c0x0, c0x1), fmaxf(c0y0, c0y1)), fminf(fmaxf(c0z0, c0z1), tmax));
  e1 = fmaxf(fmaxf(fminf(c1x0, c1x1), fminf(c1y0, c1y1)), fmaxf(fminf(c1z0, c1z1), tmin));
  float x1 = fminf(fminf(fmaxf(c1x0, c1x1), fmaxf(c1y0, c1y1)), fminf(fmaxf(c1z0, c1z1), tmax));
  x0 = fmaf(x0, 1.000001f, slack); x1 = fmaf(x1, 1.000001f, slack);
  h0 = e0 <= x0; h1 = e1 <= x1;
}

// ================= reference-order traversal (instance wrappers, reference BVH, reference visiting order) ======
// Returns true when something was hit.  The ray is edited in place by the wrappers and, like the reference,
// comes back with the rounding drift of the un-rotate / un-translate (Q12).
//
// Treelets (K_NONE links with index > 0, planted by the packer): a subtree of the reference's BVH without wrappers
// below it cannot change the ray, and everything in it is visited before anything after it, so all that the rest
// of the walk sees of it is (tmax, hit record) when it is left.  That pair is the closest hit of the subtree's
// primitives within the tmax it was entered with -- independent of the order inside and of conservative culling --
// so the subtree is walked through an own SAH BVH, near child first, like trace_fast does for a whole scene.
//
// Fused instances (K_NONE links with INSTANCE_BIT): translate / rotate / test / un-rotate / un-translate of one wrapped
// primitive as one item -- the operations of the wrapper items, in their order, without five trips through the
// stack; around a structure the way in and the way out are one item each.
//
// TREELETS = false compiles the walk without the K_NONE case (scenes where the packer planted neither).
template <bool TREELETS, class C>
__device__ __forceinline__ bool trace_reference(const Scene &sc, Ray &r, float tmin, float tmax, Hit &h,
                                                uint32_t *stk, const int stride, C &cnt) {
  bool result = false, was_last_hit = false, in_translation = false, in_rotation = false;
  uint32_t *top = stk;
  RT_PUSH(STACK_BOTTOM);
  RT_PUSH(sc.ref_root);
  V3 inv = v3(frcp(r.d.x), frcp(r.d.y), frcp(r.d.z));
  float a = dot_xfirst(r.d, r.d);
  // slab constants of the own nodes (see slab_pair); recomputed after a wrapper moved the ray
  bool own_ok = false;
  V3 cinv = inv, noi = inv;
  float slack = 0.0f;
  // every item goes through the stack and is popped here: one place where the lanes of a warp meet again
  for (;;) {
    const uint32_t cur = RT_POP();
    if (cur == STACK_BOTTOM) break;
    const uint32_t idx = rtd::ref_index(cur);
    switch (rtd::ref_kind(cur)) {
      case rtd::K_SPHERE:
        cnt.sphere();
        if (hit_sphere(sc, idx, r, a, tmin, tmax, h)) { result = true; if (in_translation || in_rotation) was_last_hit = true; }
        break;
      case rtd::K_TRIANGLE:
      case rtd::K_QUAD:
        cnt.planar();
        if (hit_planar<TREELETS>(sc, idx, rtd::ref_kind(cur) == rtd::K_TRIANGLE, r, tmin, tmax, h)) {
          result = true; if (in_translation || in_rotation) was_last_hit = true;
        }
        break;
      case rtd::K_BVH: {
        const float4 n0 = ldg4(sc.ref_nodes + 2 * idx);
        const float4 n1 = ldg4(sc.ref_nodes + 2 * idx + 1);
        float t0 = tmin, t1 = tmax;
        cnt.box(1);
        // all three axes, then one branch (the reference leaves after the first axis that fails; evaluating the others
        // cannot change the verdict, and both node words are needed at once, so their loads go out together)
        const bool ax = aabb_axis(n0.x, n0.y, r.o.x, inv.x, t0, t1);
        const bool ay = aabb_axis(n0.z, n0.w, r.o.y, inv.y, t0, t1);
        const bool az = aabb_axis(n1.x, n1.y, r.o.z, inv.z, t0, t1);
        if (ax & ay & az) {
          RT_PUSH(__float_as_uint(n1.w));    // right first => left is visited first (Q11)
          RT_PUSH(__float_as_uint(n1.z));
        }
      } break;
      case rtd::K_TRANSLATION: {
        const float4 x = ldg4(sc.xforms + idx);
        const V3 off = v3(x.x, x.y, x.z);
        cnt.translation();
        own_ok = false;
        if (!in_translation) {
          was_last_hit = false;
          r.o = vsub(r.o, off);
          RT_PUSH(cur);                       // comes back as the undo
          RT_PUSH(__float_as_uint(x.w));
          in_translation = true;
        } else {
          r.o = vadd(r.o, off);
          in_translation = false;
          if (was_last_hit) h.p = vadd(h.p, off);
        }
      } break;
      case rtd::K_YROTATION: {
        const float4 x = ldg4(sc.xforms + idx);
        const float sn = x.x, cs = x.y;
        cnt.rotation();
        own_ok = false;
        if (!in_rotation) {                                 // rt.cu:1194-1209, roundings from the reference's SASS
          was_last_hit = false;
          const V3 o = r.o, d = r.d;
          r.o = v3(ffma(o.x, cs, -fmul(o.z, sn)), o.y, ffma(o.x, sn, fmul(o.z, cs)));
          r.d = v3(ffma(d.x, cs, -fmul(d.z, sn)), d.y, ffma(d.z, cs, fmul(d.x, sn)));
          RT_PUSH(cur);
          RT_PUSH(__float_as_uint(x.w));
          in_rotation = true;
        } else {                                            // rt.cu:1210-1236
          const V3 o = r.o, d = r.d;
          r.o = v3(ffma(o.x, cs, fmul(o.z, sn)), o.y, ffma(o.z, cs, -fmul(o.x, sn)));
          r.d = v3(ffma(d.x, cs, fmul(d.z, sn)), d.y, ffma(d.z, cs, -fmul(d.x, sn)));
          in_rotation = false;
          if (was_last_hit) {
            const V3 p = h.p, n = h.n;
            h.p = v3(ffma(p.x, cs, fmul(p.z, sn)), p.y, ffma(p.z, cs, -fmul(p.x, sn)));
            h.n = v3(ffma(n.x, cs, fmul(n.z, sn)), n.y, ffma(n.z, cs, -fmul(n.x, sn)));
          }
        }
        inv.x = frcp(r.d.x); inv.z = frcp(r.d.z);
        a = dot_xfirst(r.d, r.d);
      } break;
      case rtd::K_STRUCTURE: {                            // children in slots of 4 (packer pads): one 128-bit load per 4
        const uint2 s = __ldg(sc.structures + idx);
        const uint4 *kids = reinterpret_cast<const uint4 *>(sc.struct_children + s.x);
        for (uint32_t i = 0; i < s.y; i += 4) {
          const uint4 c = __ldg(kids + (i >> 2));
          RT_PUSH(c.x);
          if (i + 1 < s.y) RT_PUSH(c.y);
          if (i + 2 < s.y) RT_PUSH(c.z);
          if (i + 3 < s.y) RT_PUSH(c.w);
        }
      } break;
      default: {                                            // K_NONE: index 0 = nothing, else node (index - 1) of a treelet
        if (!TREELETS || idx == 0) break;
        if (idx & rtd::INSTANCE_BIT) {
          const float4 *rec = sc.instances + 2 * (size_t)(idx & ~rtd::INSTANCE_BIT);
          const float4 i0 = ldg4(rec), i1 = ldg4(rec + 1);
          const uint32_t fl = __float_as_uint(i1.z);
          const V3 off = v3(i0.x, i0.y, i0.z);
          const float sn = i1.x, cs = i1.y;
          bool hit_in;                                      // was_last_hit: cleared by the entries, set by a hit inside
          own_ok = false;
          if (!(in_translation || in_rotation)) {           // enter (instances are only planted outside any wrapper)
            if (fl & rtd::INST_TRANSLATE) { cnt.translation(); r.o = vsub(r.o, off); }
            if (fl & rtd::INST_ROTATE) {                    // rt.cu:1194-1209
              cnt.rotation();
              const V3 o = r.o, d = r.d;
              r.o = v3(ffma(o.x, cs, -fmul(o.z, sn)), o.y, ffma(o.x, sn, fmul(o.z, cs)));
              r.d = v3(ffma(d.x, cs, -fmul(d.z, sn)), d.y, ffma(d.z, cs, fmul(d.x, sn)));
              a = dot_xfirst(r.d, r.d);
            }
            const uint32_t child = __float_as_uint(i0.w);
            const uint32_t ck = rtd::ref_kind(child), ci = rtd::ref_index(child);
            if (ck == rtd::K_STRUCTURE) {
              // a structure's primitives go through the stack like any other (they share the primitive cases with the
              // rest of the warp); this link comes back after them as the way out
              const uint2 st = __ldg(sc.structures + ci);
              was_last_hit = false;
              in_translation = (fl & rtd::INST_TRANSLATE) != 0; in_rotation = (fl & rtd::INST_ROTATE) != 0;
              RT_PUSH(cur);
              const uint4 *kids = reinterpret_cast<const uint4 *>(sc.struct_children + st.x);
              for (uint32_t i = 0; i < st.y; i += 4) {
                const uint4 c = __ldg(kids + (i >> 2));
                RT_PUSH(c.x);
                if (i + 1 < st.y) RT_PUSH(c.y);
                if (i + 2 < st.y) RT_PUSH(c.z);
                if (i + 3 < st.y) RT_PUSH(c.w);
              }
              break;
            }
            if (ck == rtd::K_SPHERE) { cnt.sphere(); hit_in = hit_sphere(sc, ci, r, a, tmin, tmax, h); }
            else { cnt.planar(); hit_in = hit_planar<true>(sc, ci, ck == rtd::K_TRIANGLE, r, tmin, tmax, h); }
            if (hit_in) result = true;
          } else {                                          // the way out of a structure instance
            hit_in = was_last_hit;
            in_translation = false; in_rotation = false;
          }
          if (fl & rtd::INST_ROTATE) {                      // rt.cu:1210-1236
            cnt.rotation();
            const V3 o = r.o, d = r.d;
            r.o = v3(ffma(o.x, cs, fmul(o.z, sn)), o.y, ffma(o.z, cs, -fmul(o.x, sn)));
            r.d = v3(ffma(d.x, cs, fmul(d.z, sn)), d.y, ffma(d.z, cs, -fmul(d.x, sn)));
            if (hit_in) {
              const V3 p = h.p, n = h.n;
              h.p = v3(ffma(p.x, cs, fmul(p.z, sn)), p.y, ffma(p.z, cs, -fmul(p.x, sn)));
              h.n = v3(ffma(n.x, cs, fmul(n.z, sn)), n.y, ffma(n.z, cs, -fmul(n.x, sn)));
            }
            inv.x = frcp(r.d.x); inv.z = frcp(r.d.z);
            a = dot_xfirst(r.d, r.d);
          }
          if (fl & rtd::INST_TRANSLATE) {
            cnt.translation();
            r.o = vadd(r.o, off);
            if (hit_in) h.p = vadd(h.p, off);
          }
          was_last_hit = hit_in;
          break;
        }
        if (!own_ok) {
          const float big = 1.0e30f;
          cinv = v3(fminf(fmaxf(inv.x, -big), big), fminf(fmaxf(inv.y, -big), big), fminf(fmaxf(inv.z, -big), big));
          noi = v3(-(r.o.x * cinv.x), -(r.o.y * cinv.y), -(r.o.z * cinv.z));
          slack = fminf((fabsf(noi.x) + fabsf(noi.y) + fabsf(noi.z)) * 4.76837158e-7f, 3.0e38f);
          own_ok = true;
        }
        const float4 *nd = sc.fast_nodes + 4 * (size_t)(idx - 1u);
        const float4 n0 = ldg4(nd), n1 = ldg4(nd + 1), n2 = ldg4(nd + 2), n3 = ldg4(nd + 3);
        cnt.box(2);
        float e0, e1;
        bool h0, h1;
        slab_pair(n0, n1, n2, cinv, noi, slack, tmin, tmax, e0, e1, h0, h1);
        const uint32_t r0 = __float_as_uint(n3.x), r1 = __float_as_uint(n3.y);
        const bool swap = e1 < e0;                          // near child on top
        if (h0 && h1) {
          RT_PUSH(swap ? r0 : r1);
          RT_PUSH(swap ? r1 : r0);
        } else if (h0 || h1) {
          RT_PUSH(h0 ? r0 : r1);
        }
      } break;
    }
  }
  return result;
}

// ================= reference order as a linear program (small scenes with wrappers) ============================
// The reference's walk has a FIXED visiting order: a BVH node pushes right then left, a structure its children in order,
// a wrapper itself (the way out) and its child; nothing depends on the ray except WHETHER a node's box is entered.  The
// packer therefore lays the whole walk out as a program (rt_device.h, Scene::program): op i is followed by op i + 1,
// except that a box that is missed continues at its `skip` op (the first op after its subtree).  No stack: no pops, no
// pushes, a structure costs nothing (its children simply follow), a wrapper is two ops (way in, way out) whose record
// sits in the op, and so do a box's bounds and a quad's plane -- the first test of an op needs no second load.
// Per lane the operations and their order are those of trace_reference<false> (same helpers, same roundings, same
// drift of the ray through the wrappers); lanes meet by kind of op, as in the stack walk.
// Measured (profiles/r1_linear_program.md): Cornell box s1024 406 -> 345 ms, scene 0 1.17 -> 1.11 ms, bit-identical.
// A variant that walked the program in lockstep (every step the lowest op any lane has to do, redux.sync min; each op
// issued once per warp and ray) ran 14 % SLOWER than the stack walk: 7 % fewer warp-instructions, but every step a
// serial chain vote -> op load -> branch -> test, and the lanes' inside tests no longer overlap across quads.
constexpr uint32_t OP_T_EXIT = 6u, OP_R_EXIT = 7u, OP_IN_WRAPPER = 0x100u, OP_FUSED_WRAPPER = 0x200u, OP_END = 0x400u,
                   OP_IS_PLANAR = 0x8u, OP_IS_BOX = 0x10u;   // set in the image (build_linear_image) beside the kind
template <bool G = false, bool AX = false, class C>
__device__ __forceinline__ bool trace_linear(const Scene &sc, Ray &r, float tmin, float tmax, Hit &h, C &cnt) {
  bool result = false, was_last_hit = false;
  V3 inv = v3(frcp(r.d.x), frcp(r.d.y), frcp(r.d.z));
  float a = dot_xfirst(r.d, r.d);
  // The walk keeps the ADDRESS of its next op (in the kernels of modes 3 / 4 a 32-bit shared-memory address) and reaches
  // everything else -- a primitive's record, the op after a missed box -- by a byte offset from the op at hand that the op
  // itself carries (rt_device.h, Scene::program): two LDS.128 with immediate offsets and one add per step, no base address
  // and no bound to recompute (with both, ptxas rebuilt the shared window's address from special registers at every use:
  // 10.8 % of the warp-instructions of the Cornell box).  The last op of every program is the end op.
  const float4 *op = sc.program;
  for (;;) {
    const float4 *const self = op;
    const float4 w0 = ldx<G>(op), w1 = ldx<G>(op + 1);
    const uint32_t code = __float_as_uint(w1.w), arg = __float_as_uint(w1.z);
    op += 2;
    // the two kinds that make up nine ops in ten carry a flag bit of their own in the image (one LOP3 into a predicate and a
    // branch each; the switch over the kind was a chain of up to three compare-and-branch pairs in front of every op)
    if (code & OP_IS_PLANAR) {
      const float4 *rec = reinterpret_cast<const float4 *>(reinterpret_cast<const char *>(self) + __float_as_int(w1.x));
      cnt.planar();
      if (hit_planar_p0<false, G, AX>(rec, w0, sc.sphere_count + arg, !AX && (code & 7u) == rtd::K_TRIANGLE, r, tmin, tmax, h)) {
        result = true; if (code & OP_IN_WRAPPER) was_last_hit = true;
      }
      continue;
    }
    if (code & OP_IS_BOX) {
      float t0 = tmin, t1 = tmax;
      cnt.box(1);
      const bool ax = aabb_axis(w0.x, w0.y, r.o.x, inv.x, t0, t1);
      const bool ay = aabb_axis(w0.z, w0.w, r.o.y, inv.y, t0, t1);
      const bool az = aabb_axis(w1.x, w1.y, r.o.z, inv.z, t0, t1);
      if (!(ax & ay & az)) op = reinterpret_cast<const float4 *>(reinterpret_cast<const char *>(self) + (int)arg);
      continue;
    }
    switch (code & 7u) {
      case rtd::K_SPHERE: {
        if (code & OP_END) return result;
        const float4 *rec = reinterpret_cast<const float4 *>(reinterpret_cast<const char *>(self) + __float_as_int(w1.x));
        cnt.sphere();
        if (hit_sphere_rec<G>(rec, arg, r, a, tmin, tmax, h)) { result = true; if (code & OP_IN_WRAPPER) was_last_hit = true; }
      } break;
      case rtd::K_TRANSLATION:
        cnt.translation();
        was_last_hit = false;
        r.o = vsub(r.o, v3(w0.x, w0.y, w0.z));
        if (code & OP_FUSED_WRAPPER) {                        // translate(rotate(x)): the rotation's way in follows at once
          const float sn = w1.x, cs = w1.y;
          cnt.rotation();
          const V3 o = r.o, d = r.d;
          r.o = v3(ffma(o.x, cs, -fmul(o.z, sn)), o.y, ffma(o.x, sn, fmul(o.z, cs)));
          r.d = v3(ffma(d.x, cs, -fmul(d.z, sn)), d.y, ffma(d.z, cs, fmul(d.x, sn)));
          inv.x = frcp(r.d.x); inv.z = frcp(r.d.z);
          a = dot_xfirst(r.d, r.d);
        }
        break;
      case OP_T_EXIT: {
        const V3 off = v3(w0.x, w0.y, w0.z);
        cnt.translation();
        r.o = vadd(r.o, off);
        if (was_last_hit) h.p = vadd(h.p, off);
      } break;
      case rtd::K_YROTATION: {                              // rt.cu:1194-1209, roundings from the reference's SASS
        const float sn = w0.x, cs = w0.y;
        cnt.rotation();
        was_last_hit = false;
        const V3 o = r.o, d = r.d;
        r.o = v3(ffma(o.x, cs, -fmul(o.z, sn)), o.y, ffma(o.x, sn, fmul(o.z, cs)));
        r.d = v3(ffma(d.x, cs, -fmul(d.z, sn)), d.y, ffma(d.z, cs, fmul(d.x, sn)));
        inv.x = frcp(r.d.x); inv.z = frcp(r.d.z);
        a = dot_xfirst(r.d, r.d);
      } break;
      default: {                                            // OP_R_EXIT, rt.cu:1210-1236
        const float sn = w0.x, cs = w0.y;
        cnt.rotation();
        const V3 o = r.o, d = r.d;
        r.o = v3(ffma(o.x, cs, fmul(o.z, sn)), o.y, ffma(o.z, cs, -fmul(o.x, sn)));
        r.d = v3(ffma(d.x, cs, fmul(d.z, sn)), d.y, ffma(d.z, cs, -fmul(d.x, sn)));
        if (was_last_hit) {
          const V3 p = h.p, n = h.n;
          h.p = v3(ffma(p.x, cs, fmul(p.z, sn)), p.y, ffma(p.z, cs, -fmul(p.x, sn)));
          h.n = v3(ffma(n.x, cs, fmul(n.z, sn)), n.y, ffma(n.z, cs, -fmul(n.x, sn)));
        }
        inv.x = frcp(r.d.x); inv.z = frcp(r.d.z);
        a = dot_xfirst(r.d, r.d);
        if (code & OP_FUSED_WRAPPER) {                        // ... and the translation's way out follows the rotation's
          const V3 off = v3(w0.z, w0.w, w1.x);
          cnt.translation();
          r.o = vadd(r.o, off);
          if (was_last_hit) h.p = vadd(h.p, off);
        }
      } break;
    }
  }
  return result;
}

// ================= tiny scenes without wrappers: every primitive, no BVH, no stack ================================
// Scenes of at most 4 spheres and 4 quads / triangles (the Perlin scenes 17 and 19 among them) are bound by the latency
// of ONE lane: their slowest pixels are a serial chain of up to spp x 50 bounces (profiles/r1b_linear_program.md).  A BVH
// saves them no work; it only adds a node test and a stack trip in front of sphere tests that then run one after the
// other.  Here the spheres are tested two at a time in one basic block (independent sqrt / divisions interleave, see
// rt_math.cuh) against the tmax the walk started with; the closest hit is then picked in primitive order.  That is
// the result of testing them one after the other with a shrinking tmax: a sphere's candidate root does not depend on
// tmax except through the final `root < tmax`, because root2 >= root1 (same divisor > 0), so a root1 that a smaller
// tmax would reject is followed by a root2 that it rejects as well.
__device__ __forceinline__ void sphere_stage(const Scene &sc, uint32_t idx, const Ray &r, float a, float &b, float &disc) {
  const float4 s0 = ldg4(sc.spheres + 2 * idx);
  const float4 s1 = ldg4(sc.spheres + 2 * idx + 1);
  const V3 c = v3(ffma(r.time, s1.x, s0.x), ffma(r.time, s1.y, s0.y), ffma(r.time, s1.z, s0.z));   // movement_path_at
  const V3 oc = vsub(c, r.o);
  b = fmul(dot_yfirst(r.d, oc), -2.0f);
  const float cc = ffma(-s0.w, s0.w, dot_yfirst(oc, oc));
  disc = ffma(b, b, fmul(cc, fmul(a, -4.0f)));
}
// the two candidate roots of two spheres at once; returns false when a guard of the fast arithmetic fails (caller falls back)
__device__ __forceinline__ bool sphere_roots2_fast(float b0, float disc0, float b1, float disc1, float a2, float tmin, float tmax,
                                                   float &t0, bool &h0, float &t1, bool &h1) {
  const bool v0 = disc0 >= 0.0f, v1 = disc1 >= 0.0f;               // a negative discriminant is a miss: its lane computes on 1.0
  const float d0 = v0 ? disc0 : 1.0f, d1 = v1 ? disc1 : 1.0f;
  if (!(fsqrt_fast_ok(d0) && fsqrt_fast_ok(d1))) return false;
  const float q0 = fsqrt_fast(d0), q1 = fsqrt_fast(d1);
  const float n00 = fsub(-b0, q0), n01 = fsub(q0, b0), n10 = fsub(-b1, q1), n11 = fsub(q1, b1);
  if (!(fdiv_fast_ok(n00, a2) && fdiv_fast_ok(n01, a2) && fdiv_fast_ok(n10, a2) && fdiv_fast_ok(n11, a2))) return false;
  const float r00 = fdiv_fast(n00, a2), r01 = fdiv_fast(n01, a2), r10 = fdiv_fast(n10, a2), r11 = fdiv_fast(n11, a2);
  t0 = (tmin >= r00 || r00 >= tmax) ? r01 : r00;
  t1 = (tmin >= r10 || r10 >= tmax) ? r11 : r10;
  h0 = v0 && tmin < t0 && t0 < tmax;
  h1 = v1 && tmin < t1 && t1 < tmax;
  return true;
}
template <class C>
__device__ __forceinline__ bool trace_flat(const Scene &sc, const Ray &r, float tmin, float tmax, Hit &h, C &cnt) {
  const float a = dot_xfirst(r.d, r.d);
  const float a2 = fadd(a, a);
  const uint32_t ns = sc.sphere_count, np = sc.planar_count;
  uint32_t best = 0xffffffffu;
  float tbest = tmax;
  for (uint32_t i = 0; i < ns; i += 2) {
    const uint32_t j = (i + 1 < ns) ? i + 1 : i;                      // odd count: the last sphere stands in for its missing partner
    float b0, disc0, b1, disc1, t0 = 0.0f, t1 = 0.0f;
    bool h0 = false, h1 = false;
    cnt.sphere(); if (j != i) cnt.sphere();
    sphere_stage(sc, i, r, a, b0, disc0);
    sphere_stage(sc, j, r, a, b1, disc1);
    if (!sphere_roots2_fast(b0, disc0, b1, disc1, a2, tmin, tmax, t0, h0, t1, h1)) {   // rare: operands outside the guards
      h0 = sphere_root(sc, i, r, a, tmin, tmax, t0);
      h1 = sphere_root(sc, j, r, a, tmin, tmax, t1);
    }
    if (h0 && t0 < tbest) { tbest = t0; best = rtd::make_ref(rtd::K_SPHERE, i); }
    if (j != i && h1 && t1 < tbest) { tbest = t1; best = rtd::make_ref(rtd::K_SPHERE, j); }
  }
  for (uint32_t i = 0; i < np; ++i) {
    float tt;
    cnt.planar();
    if (planar_t(sc, i, (sc.flat_triangles >> i) & 1u, r, tmin, tbest, tt)) { tbest = tt; best = rtd::make_ref(rtd::K_QUAD, i); }
  }
  if (best == 0xffffffffu) return false;
  if (rtd::ref_kind(best) == rtd::K_SPHERE) sphere_record(sc, rtd::ref_index(best), r, tbest, h);
  else planar_record(sc, rtd::ref_index(best), r, tbest, h);
  return true;
}

// ---- instanced leaf of the own BVH (RT_TRAVERSAL_FAST on scenes with wrappers; rt_pack.cpp instanced_leaf) ----
// the ray goes into the primitive's space with the reference's roundings (rt.cu:1176-1209), the hit comes back with them
// (rt.cu:1210-1236, 1186-1188); `a` = d.d of the moved ray (a rotation changes its rounding)
__device__ __forceinline__ void instance_enter(const float4 i0, const float4 i1, Ray &q, float &a) {
  const uint32_t fl = __float_as_uint(i1.z);
  if (fl & rtd::INST_TRANSLATE) q.o = vsub(q.o, v3(i0.x, i0.y, i0.z));
  if (fl & rtd::INST_ROTATE) {
    const float sn = i1.x, cs = i1.y;
    const V3 o = q.o, d = q.d;
    q.o = v3(ffma(o.x, cs, -fmul(o.z, sn)), o.y, ffma(o.x, sn, fmul(o.z, cs)));
    q.d = v3(ffma(d.x, cs, -fmul(d.z, sn)), d.y, ffma(d.z, cs, fmul(d.x, sn)));
    a = dot_xfirst(q.d, q.d);
  }
}
__device__ __forceinline__ void instance_leave(const float4 i0, const float4 i1, Hit &h) {
  const uint32_t fl = __float_as_uint(i1.z);
  if (fl & rtd::INST_ROTATE) {
    const float sn = i1.x, cs = i1.y;
    const V3 p = h.p, n = h.n;
    h.p = v3(ffma(p.x, cs, fmul(p.z, sn)), p.y, ffma(p.z, cs, -fmul(p.x, sn)));
    h.n = v3(ffma(n.x, cs, fmul(n.z, sn)), n.y, ffma(n.z, cs, -fmul(n.x, sn)));
  }
  if (fl & rtd::INST_TRANSLATE) h.p = vadd(h.p, v3(i0.x, i0.y, i0.z));
}

// ================= own BVH (SAH, world-space primitives, near child first) =====================================
// Closest hit is independent of visiting order and of (conservative) box culling, so for scenes without
// instance wrappers this returns bit-identical hits to trace_reference at a fraction of the node visits.
// INSTANCED = the BVH has instanced leaves (RT_TRAVERSAL_FAST on a scene with wrappers; kernel mode 6).  A template
// parameter, not a run-time case: with the extra leaf kind in the loop the wrapper-free scenes ran 14 % slower (C2 95 -> 109 ms).
template <bool INSTANCED, class C>
__device__ __forceinline__ bool trace_fast(const Scene &sc, const Ray &r, float tmin, float tmax, Hit &h,
                                           uint32_t *stk, const int stride, C &cnt) {
  // only (tmax, primitive) are carried through the walk; the hit record is built once, for the closest hit, with
  // the same formulas hit_sphere / hit_planar use (identical bits)
  uint32_t best = 0xffffffffu;
  const float a = dot_xfirst(r.d, r.d);
  // slab constants: see slab_pair
  const float big = 1.0e30f;
  const V3 inv = v3(fminf(fmaxf(frcp(r.d.x), -big), big), fminf(fmaxf(frcp(r.d.y), -big), big), fminf(fmaxf(frcp(r.d.z), -big), big));
  const V3 noi = v3(-(r.o.x * inv.x), -(r.o.y * inv.y), -(r.o.z * inv.z));
  const float slack = fminf((fabsf(noi.x) + fabsf(noi.y) + fabsf(noi.z)) * 4.76837158e-7f, 3.0e38f);
  uint32_t *top = stk;
  RT_PUSH(STACK_BOTTOM);
  uint32_t cur = sc.fast_root;
  for (;;) {
    const uint32_t kind = rtd::ref_kind(cur);
    const uint32_t idx = rtd::ref_index(cur);
    // a range test on the link itself: the compiler lowers a chain of tests on `kind` to a search tree that reaches the
    // node case, nine visits in ten, through three compare-and-branch pairs
    if (cur - (rtd::K_BVH << rtd::KIND_SHIFT) < (1u << rtd::KIND_SHIFT)) {
      const float4 n0 = ldg4(sc.fast_nodes + 4 * idx);
      const float4 n1 = ldg4(sc.fast_nodes + 4 * idx + 1);
      const float4 n2 = ldg4(sc.fast_nodes + 4 * idx + 2);
      const float4 n3 = ldg4(sc.fast_nodes + 4 * idx + 3);
      cnt.box(2);
      float e0, e1;
      bool h0, h1;
      slab_pair(n0, n1, n2, inv, noi, slack, tmin, tmax, e0, e1, h0, h1);
      const uint32_t r0 = __float_as_uint(n3.x), r1 = __float_as_uint(n3.y);
      if (h0 && h1) {
        const bool swap = e1 < e0;
        cur = swap ? r1 : r0;
        RT_PUSH(swap ? r0 : r1);
        continue;
      } else if (h0 || h1) {
        cur = h0 ? r0 : r1;
        continue;
      }
    } else if (kind == rtd::K_SPHERE) {
      cnt.sphere();
      float root;
      if (sphere_root(sc, idx, r, a, tmin, tmax, root)) { tmax = root; best = cur; }
    } else if (kind == rtd::K_QUAD || kind == rtd::K_TRIANGLE) {
      cnt.planar();
      float tt;
      if (planar_t(sc, idx, kind == rtd::K_TRIANGLE, r, tmin, tmax, tt)) { tmax = tt; best = cur; }
    } else if (INSTANCED && kind == rtd::K_TRANSLATION) { // instanced leaf: the primitive's own record, tested in its own space
      const float4 i0 = ldg4(sc.instances + 2 * idx), i1 = ldg4(sc.instances + 2 * idx + 1);
      Ray q = r;
      float aq = a, tt;
      instance_enter(i0, i1, q, aq);
      const uint32_t child = __float_as_uint(i0.w), ck = rtd::ref_kind(child), ci = rtd::ref_index(child);
      bool hit_in;
      if (ck == rtd::K_SPHERE) { cnt.sphere(); hit_in = sphere_root(sc, ci, q, aq, tmin, tmax, tt); }
      else { cnt.planar(); hit_in = planar_t(sc, ci, ck == rtd::K_TRIANGLE, q, tmin, tmax, tt); }
      if (hit_in) { tmax = tt; best = cur; }
    }                                                     // K_NONE: the empty second child of a one-primitive scene
    cur = RT_POP();
    if (cur == STACK_BOTTOM) break;
  }
  if (best == 0xffffffffu) return false;
  if (rtd::ref_kind(best) == rtd::K_SPHERE) sphere_record(sc, rtd::ref_index(best), r, tmax, h);
  else if (INSTANCED && rtd::ref_kind(best) == rtd::K_TRANSLATION) {
    const uint32_t idx = rtd::ref_index(best);
    const float4 i0 = ldg4(sc.instances + 2 * idx), i1 = ldg4(sc.instances + 2 * idx + 1);
    Ray q = r;
    float aq = a;
    instance_enter(i0, i1, q, aq);
    const uint32_t child = __float_as_uint(i0.w);
    if (rtd::ref_kind(child) == rtd::K_SPHERE) sphere_record(sc, rtd::ref_index(child), q, tmax, h);
    else planar_record(sc, rtd::ref_index(child), q, tmax, h);
    instance_leave(i0, i1, h);
  } else planar_record(sc, rtd::ref_index(best), r, tmax, h);
  return true;
}

}  // namespace rtt
#endif
