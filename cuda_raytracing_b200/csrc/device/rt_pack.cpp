// rt_pack.cpp -- host-side packer (product code, plain C++): turns the reference's flat scene arrays
// (Hittable[120 B], u16 world refs, texture byte arena; rt.cu:1699-1733 uploads them verbatim) into the
// device layout of rt_device.h: SoA primitive records, tagged 32-bit links, pre-resolved materials, Perlin
// tables split out for shared-memory staging, and -- for RT_TRAVERSAL_FAST -- world-space flattened instances
// under a SAH BVH with 64-byte two-child nodes.
#include "rt_pack.h"

#include <math.h>
#include <cmath>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <map>

namespace rtp {
using namespace rtd;

static inline float4 f4(float x, float y, float z, float w) { float4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
static inline float as_float(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t as_uint(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

// ---------------------------------------------------------------------------------------- textures / materials
static const rt_texture *tex_at(const uint8_t *arena, uint64_t bytes, uint64_t off) {
  if (!arena || off + sizeof(rt_texture) > bytes) return nullptr;
  return reinterpret_cast<const rt_texture *>(arena + off);
}
static bool tex_has_image(const uint8_t *arena, uint64_t bytes, uint64_t off, int depth = 0) {
  const rt_texture *t = tex_at(arena, bytes, off);
  if (!t || depth > 32) return false;
  if (t->type == RT_TEX_IMAGE) return true;
  if (t->type == RT_TEX_CHECKER)
    return tex_has_image(arena, bytes, t->u.checker.even, depth + 1) || tex_has_image(arena, bytes, t->u.checker.odd, depth + 1);
  return false;
}


void pack_material(const rt_material &m, const uint8_t *arena, uint64_t bytes, PackedScene &out, uint4 rec[2]) {
  PackedScene &st = out;
  uint32_t type = m.type, kind = TK_SOLID, tex_off = 0, aux = 0, needs_uv = 0;
  float a = 0, b = 0, c = 0, d = 0;
  if (type == RT_MAT_METAL) { a = m.u.metal.albedo.x; b = m.u.metal.albedo.y; c = m.u.metal.albedo.z; d = m.u.metal.fuzz; }
  else if (type == RT_MAT_DIELECTRIC) { d = m.u.dielectric.refraction_index; }
  else {
    tex_off = (uint32_t)m.u.texture;
    const rt_texture *t = tex_at(arena, bytes, m.u.texture);
    kind = TK_ARENA;
    if (t && t->type == RT_TEX_SOLID) {
      kind = TK_SOLID; a = t->u.solid.albedo.x; b = t->u.solid.albedo.y; c = t->u.solid.albedo.z;
    } else if (t && t->type == RT_TEX_CHECKER) {
      const rt_texture *e = tex_at(arena, bytes, t->u.checker.even), *o = tex_at(arena, bytes, t->u.checker.odd);
      if (e && o && e->type == RT_TEX_SOLID && o->type == RT_TEX_SOLID) {
        kind = TK_CHECKER_SOLID;
        auto it = st.checker_ids.find(m.u.texture);
        if (it == st.checker_ids.end()) {
          aux = (uint32_t)(out.checkers.size() / 2);
          out.checkers.push_back(f4(t->u.checker.inverted_scale, e->u.solid.albedo.x, e->u.solid.albedo.y, e->u.solid.albedo.z));
          out.checkers.push_back(f4(0.0f, o->u.solid.albedo.x, o->u.solid.albedo.y, o->u.solid.albedo.z));
          st.checker_ids[m.u.texture] = aux;
        } else aux = it->second;
      }
    } else if (t && t->type == RT_TEX_NOISE && t->u.noise.noise_offset + sizeof(rt_perlin) <= bytes) {
      kind = TK_NOISE; d = t->u.noise.scale;
      auto it = st.noise_ids.find(t->u.noise.noise_offset);
      if (it == st.noise_ids.end()) {
        aux = out.perlin_tables++;
        const rt_perlin *pn = reinterpret_cast<const rt_perlin *>(arena + t->u.noise.noise_offset);
        for (int i = 0; i < RT_PERLIN_POINTS; ++i) {
          out.perlin_vectors.push_back(f4(pn->vectors[i].x, pn->vectors[i].y, pn->vectors[i].z, 0.0f));
          out.perlin_perm.push_back((uint8_t)(pn->perm_x[i] & 255));   // a permutation of 0..255
        }
        st.noise_ids[t->u.noise.noise_offset] = aux;
      } else aux = it->second;
    }
    if (kind == TK_ARENA) needs_uv = tex_has_image(arena, bytes, m.u.texture) ? 1u : 0u;
  }
  rec[0].x = type | (kind << 8) | (needs_uv << 16); rec[0].y = tex_off; rec[0].z = aux; rec[0].w = 0;
  rec[1].x = as_uint(a); rec[1].y = as_uint(b); rec[1].z = as_uint(c); rec[1].w = as_uint(d);
}

// ---------------------------------------------------------------------------------------- geometry helpers (f32)
struct V { float x, y, z; };
static inline V vv(const rt_vec3 &a) { return V{a.x, a.y, a.z}; }
struct Box { float lo[3], hi[3]; };
static inline Box box_of(const rt_aabb &b) { return Box{{b.x0, b.y0, b.z0}, {b.x1, b.y1, b.z1}}; }
static inline Box box_empty() { return Box{{INFINITY, INFINITY, INFINITY}, {-INFINITY, -INFINITY, -INFINITY}}; }
static inline void box_grow(Box &b, const Box &o) {
  for (int a = 0; a < 3; ++a) { b.lo[a] = std::min(b.lo[a], o.lo[a]); b.hi[a] = std::max(b.hi[a], o.hi[a]); }
}
static inline void box_point(Box &b, V p) {
  b.lo[0] = std::min(b.lo[0], p.x); b.hi[0] = std::max(b.hi[0], p.x);
  b.lo[1] = std::min(b.lo[1], p.y); b.hi[1] = std::max(b.hi[1], p.y);
  b.lo[2] = std::min(b.lo[2], p.z); b.hi[2] = std::max(b.hi[2], p.z);
}
static inline float box_area(const Box &b) {
  float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
  if (!(dx >= 0) || !(dy >= 0) || !(dz >= 0)) return 0.0f;
  return 2.0f * (dx * dy + dy * dz + dz * dx);
}
// object -> world transform of a (translation o y-rotation)* chain: world = R(p) + T,
// R: x = c*x' + s*z', z = -s*x' + c*z'   (the inverse of the ray rotation of rt.cu:1196-1206)
struct Xf {
  float c = 1.0f, s = 0.0f; V t{0, 0, 0}; bool identity = true;
  // the chain of wrappers as the reference would walk it, while it still has the shape [translation][y-rotation] (at most one
  // of each, the translation outermost): such a primitive stays in its OWN space and is tested there (instanced leaf)
  bool simple = true; uint32_t flags = 0; float off[3] = {0, 0, 0}, sn = 0.0f, cs = 1.0f;
};
static inline V xf_dir(const Xf &x, V p) { return V{x.c * p.x + x.s * p.z, p.y, -x.s * p.x + x.c * p.z}; }
static inline V xf_point(const Xf &x, V p) { V r = xf_dir(x, p); return V{r.x + x.t.x, r.y + x.t.y, r.z + x.t.z}; }

struct Leaf { uint32_t ref; Box box; float centroid[3]; };

// ---------------------------------------------------------------------------------------- SAH BVH (own traversal)
struct SahBuilder {
  std::vector<Leaf> &leaves;
  std::vector<float4> &nodes;
  uint32_t max_depth = 0;
  bool own_links;                                      // treelet: inner links are make_own_ref(), not K_BVH
  explicit SahBuilder(std::vector<Leaf> &l, std::vector<float4> &n, bool own = false) : leaves(l), nodes(n), own_links(own) {}

  // returns the tagged ref of the subtree over leaves[begin,end) and its box
  uint32_t build(size_t begin, size_t end, uint32_t depth, Box &box_out) {
    max_depth = std::max(max_depth, depth);
    if (end - begin == 1) { box_out = leaves[begin].box; return leaves[begin].ref; }
    size_t n = end - begin, best_split = begin + n / 2;
    int best_axis = -1;
    if (depth < 16) {                                  // full-sweep SAH; deeper levels fall back to a balanced median
      float best_cost = INFINITY;
      std::vector<float> right_area(n);
      for (int axis = 0; axis < 3; ++axis) {
        std::sort(leaves.begin() + begin, leaves.begin() + end,
                  [axis](const Leaf &a, const Leaf &b) { return a.centroid[axis] < b.centroid[axis]; });
        Box acc = box_empty();
        for (size_t i = n; i-- > 1;) { box_grow(acc, leaves[begin + i].box); right_area[i] = box_area(acc); }
        acc = box_empty();
        for (size_t i = 1; i < n; ++i) {
          box_grow(acc, leaves[begin + i - 1].box);
          float cost = box_area(acc) * (float)i + right_area[i] * (float)(n - i);
          if (cost < best_cost) { best_cost = cost; best_axis = axis; best_split = begin + i; }
        }
      }
    }
    if (best_axis < 0) {                               // median on the widest centroid axis
      Box cb = box_empty();
      for (size_t i = begin; i < end; ++i) box_point(cb, V{leaves[i].centroid[0], leaves[i].centroid[1], leaves[i].centroid[2]});
      best_axis = 0;
      for (int a = 1; a < 3; ++a) if (cb.hi[a] - cb.lo[a] > cb.hi[best_axis] - cb.lo[best_axis]) best_axis = a;
      best_split = begin + n / 2;
    }
    const int axis = best_axis;
    std::sort(leaves.begin() + begin, leaves.begin() + end,
              [axis](const Leaf &a, const Leaf &b) { return a.centroid[axis] < b.centroid[axis]; });
    const uint32_t me = (uint32_t)(nodes.size() / 4);
    nodes.resize(nodes.size() + 4);
    Box b0, b1;
    const uint32_t r0 = build(begin, best_split, depth + 1, b0);
    const uint32_t r1 = build(best_split, end, depth + 1, b1);
    nodes[4 * me + 0] = f4(b0.lo[0], b0.hi[0], b0.lo[1], b0.hi[1]);
    nodes[4 * me + 1] = f4(b1.lo[0], b1.hi[0], b1.lo[1], b1.hi[1]);
    nodes[4 * me + 2] = f4(b0.lo[2], b0.hi[2], b1.lo[2], b1.hi[2]);
    nodes[4 * me + 3] = f4(as_float(r0), as_float(r1), 0.0f, 0.0f);
    box_out = b0; box_grow(box_out, b1);
    return own_links ? make_own_ref(me) : make_ref(K_BVH, me);
  }
};

// ---------------------------------------------------------------------------------------- the packer proper
namespace {

struct Packer {
  const rt_scene_desc &d;
  PackedScene &out;
  std::vector<uint32_t> tag;                    // hittable index -> tagged ref (reference graph)
  std::vector<rt_material> sphere_mats, planar_mats;
  std::vector<Leaf> leaves;                     // FAST traversal leaves
  uint32_t treelet_min;                         // smallest wrapper-free subtree (primitives) worth an own BVH; 0 = never
  bool fuse_instances;
  uint32_t fuse_min;                            // fewer qualifying chains than this (and no treelet): keep the plain walk
  uint32_t program_max;                         // longest reference-order walk (ops) laid out as a linear program; 0 = never
  bool fuse_program_wrappers = true;            // env RT_FUSE_PROGRAM_WRAPPERS=0 keeps one op per wrapper side (bisecting)
  mutable bool too_deep = false;                // a recursion below hit its nesting guard: the scene is refused, never truncated
  Packer(const rt_scene_desc &desc, PackedScene &o, uint32_t tmin, uint32_t fmin, uint32_t pmax)
      : d(desc), out(o), treelet_min(tmin), fuse_instances(fmin > 0), fuse_min(fmin), program_max(pmax) {
    if (const char *v = getenv("RT_FUSE_PROGRAM_WRAPPERS")) fuse_program_wrappers = strtol(v, nullptr, 10) != 0;
  }

  const rt_hittable &H(uint32_t i) const { return d.hittables[i]; }

  uint32_t add_sphere(V c, V mv, float r, const rt_material &m) {
    uint32_t idx = (uint32_t)(out.spheres.size() / 2);
    uint32_t flags = 0;
    if ((m.type == RT_MAT_LAMBERTIAN || m.type == RT_MAT_DIFFUSE_LIGHT) &&
        tex_has_image(d.texture_arena, d.texture_bytes, m.u.texture)) flags |= 1u;
    out.spheres.push_back(f4(c.x, c.y, c.z, r));
    out.spheres.push_back(f4(mv.x, mv.y, mv.z, as_float(flags)));
    sphere_mats.push_back(m);
    return idx;
  }
  uint32_t add_planar(const rt_geometry &g, const rt_material &m) {
    uint32_t idx = (uint32_t)(out.planars.size() / 4);
    out.planars.push_back(f4(g.normal.x, g.normal.y, g.normal.z, g.d));
    out.planars.push_back(f4(g.q.x, g.q.y, g.q.z, g.w.x));
    out.planars.push_back(f4(g.v.x, g.v.y, g.v.z, g.w.y));
    out.planars.push_back(f4(g.u.x, g.u.y, g.u.z, g.w.z));
    planar_mats.push_back(m);
    return idx;
  }

  // exact stack need of the reference-order traversal (see trace_reference: every item is pushed, then popped and
  // worked on): `base` = entries below this subtree's root
  uint32_t stack_need(uint32_t href, uint32_t base, int guard) const {
    if (guard > 64) { too_deep = true; return base; }
    if (href >= d.hittable_count) return base;
    if (fused[href]) return base + fused[href] - 1;                        // nothing pushed, or the way out + a structure's children
    if (treelet_depth[href] > 0) return base + treelet_depth[href] + 1;    // one pending far child per level + the pair pushed last
    const rt_hittable &h = H(href);
    switch (h.type) {
      case RT_OBJ_BVH_NODE: {
        uint32_t l = d.world_refs[h.u.bvh.left], r = d.world_refs[h.u.bvh.right];
        return std::max(base + 2, std::max(stack_need(l, base + 1, guard + 1), stack_need(r, base, guard + 1)));
      }
      case RT_OBJ_TRANSLATION: return std::max(base + 2, stack_need(h.u.translation.child, base + 1, guard + 1));
      case RT_OBJ_Y_ROTATION: return std::max(base + 2, stack_need(h.u.y_rotation.child, base + 1, guard + 1));
      case RT_OBJ_STRUCTURE: {
        uint32_t need = base + (uint32_t)h.u.structure.count;
        for (uint32_t k = 0; k < h.u.structure.count; ++k)
          need = std::max(need, stack_need(h.u.structure.first + k, base + k, guard + 1));
        return need;
      }
      default: return base;
    }
  }

  // ---- treelets: wrapper-free subtrees of the reference BVH get an own SAH BVH (exactness argument in rt_trace.cuh) ----
  std::vector<int64_t> pure_prims;              // per hittable: primitives below it, or -1 when a wrapper is below it
  std::vector<uint32_t> treelet_depth;          // per hittable: > 0 on the BVH nodes that were replaced by a treelet
  int64_t purity(uint32_t href, int guard) {
    if (guard > 64 || href >= d.hittable_count) return -1;
    if (pure_prims[href] != -2) return pure_prims[href];
    const rt_hittable &h = H(href);
    int64_t n = -1;
    switch (h.type) {
      case RT_OBJ_SPHERE: case RT_OBJ_TRIANGLE: case RT_OBJ_QUAD: n = 1; break;
      case RT_OBJ_STRUCTURE:
        n = 0;
        if ((uint64_t)h.u.structure.first + h.u.structure.count > d.hittable_count) { n = -1; break; }
        for (uint32_t k = 0; k < h.u.structure.count && n >= 0; ++k) {
          const int64_t c = purity(h.u.structure.first + k, guard + 1);
          n = c < 0 ? -1 : n + c;
        }
        break;
      case RT_OBJ_BVH_NODE: {
        if (h.u.bvh.left >= d.world_count || h.u.bvh.right >= d.world_count) break;
        const int64_t a = purity(d.world_refs[h.u.bvh.left], guard + 1);
        const int64_t b = h.u.bvh.right == h.u.bvh.left ? 0 : purity(d.world_refs[h.u.bvh.right], guard + 1);
        n = (a < 0 || b < 0) ? -1 : a + b;
      } break;
      default: break;                           // wrappers (and anything unknown) end a pure subtree
    }
    pure_prims[href] = n;
    return n;
  }
  void gather(uint32_t href, std::vector<Leaf> &lv, int guard) {
    if (guard > 64) { too_deep = true; return; }
    const rt_hittable &h = H(href);
    switch (h.type) {
      case RT_OBJ_SPHERE: { Leaf lf; lf.ref = tag[href]; lf.box = box_of(h.u.sphere.bbox); center_leaf(lf); lv.push_back(lf); } break;
      case RT_OBJ_TRIANGLE: case RT_OBJ_QUAD: { Leaf lf; lf.ref = tag[href]; lf.box = box_of(h.u.planar.bbox); center_leaf(lf); lv.push_back(lf); } break;
      case RT_OBJ_STRUCTURE: for (uint32_t k = 0; k < h.u.structure.count; ++k) gather(h.u.structure.first + k, lv, guard + 1); break;
      case RT_OBJ_BVH_NODE:
        gather(d.world_refs[h.u.bvh.left], lv, guard + 1);
        if (h.u.bvh.right != h.u.bvh.left) gather(d.world_refs[h.u.bvh.right], lv, guard + 1);
        break;
      default: break;
    }
  }
  // top-down over the reference graph: the first BVH node on each path that is wrapper-free and big enough becomes a treelet
  void plant_treelets(uint32_t href, uint32_t min_prims, int guard) {
    if (guard > 64) { too_deep = true; return; }
    if (href >= d.hittable_count) return;
    const rt_hittable &h = H(href);
    if (h.type == RT_OBJ_BVH_NODE) {
      if (treelet_depth[href] > 0) return;
      const int64_t n = purity(href, 0);
      if (n >= (int64_t)min_prims && n >= 2) {
        std::vector<Leaf> lv;
        gather(href, lv, 0);
        if (lv.size() >= 2) {
          SahBuilder sb(lv, out.fast_nodes, true);
          Box rb;
          tag[href] = sb.build(0, lv.size(), 0, rb);
          treelet_depth[href] = sb.max_depth;
          ++out.treelets; out.treelet_prims += (uint32_t)lv.size();
          return;
        }
      }
      if (h.u.bvh.left >= d.world_count || h.u.bvh.right >= d.world_count) return;
      plant_treelets(d.world_refs[h.u.bvh.left], min_prims, guard + 1);
      if (h.u.bvh.right != h.u.bvh.left) plant_treelets(d.world_refs[h.u.bvh.right], min_prims, guard + 1);
    } else if (h.type == RT_OBJ_TRANSLATION) {
      plant_treelets(h.u.translation.child, min_prims, guard + 1);
    } else if (h.type == RT_OBJ_Y_ROTATION) {
      plant_treelets(h.u.y_rotation.child, min_prims, guard + 1);
    } else if (h.type == RT_OBJ_STRUCTURE && (uint64_t)h.u.structure.first + h.u.structure.count <= d.hittable_count) {
      for (uint32_t k = 0; k < h.u.structure.count; ++k) plant_treelets(h.u.structure.first + k, min_prims, guard + 1);
    }
  }
  // ---- fused instances: T(R(x)), T(x), R(x) around a primitive or a structure of primitives, reached from the BVH only ----
  std::vector<uint8_t> fused;
  bool is_prim(uint32_t href) const {
    return href < d.hittable_count && (H(href).type == RT_OBJ_SPHERE || H(href).type == RT_OBJ_TRIANGLE || H(href).type == RT_OBJ_QUAD);
  }
  bool fusable_child(uint32_t href) const {
    if (is_prim(href)) return true;
    if (href >= d.hittable_count || H(href).type != RT_OBJ_STRUCTURE) return false;
    const rt_structure &st = H(href).u.structure;
    if (st.count == 0 || st.count > 64 || (uint64_t)st.first + st.count > d.hittable_count) return false;
    for (uint32_t k = 0; k < st.count; ++k) if (!is_prim(st.first + k)) return false;
    return true;
  }
  uint32_t plant_instances(bool dry_run) {             // returns the number of chains that qualify
    uint32_t found = 0;
    const uint64_t n = d.hittable_count;
    std::vector<uint8_t> nested(n, 0);              // referenced from inside a wrapper or a structure
    for (uint64_t i = 0; i < n; ++i) {
      const rt_hittable &h = H((uint32_t)i);
      if (h.type == RT_OBJ_TRANSLATION && h.u.translation.child < n) nested[h.u.translation.child] = 1;
      else if (h.type == RT_OBJ_Y_ROTATION && h.u.y_rotation.child < n) nested[h.u.y_rotation.child] = 1;
      else if (h.type == RT_OBJ_STRUCTURE && (uint64_t)h.u.structure.first + h.u.structure.count <= n)
        for (uint32_t k = 0; k < h.u.structure.count; ++k) nested[h.u.structure.first + k] = 1;
    }
    for (uint64_t i = 0; i < n; ++i) {
      const rt_hittable &h = H((uint32_t)i);
      if (nested[i]) continue;
      uint32_t flags = 0, child = 0;
      float ox = 0, oy = 0, oz = 0, sn = 0, cs = 1;
      if (h.type == RT_OBJ_TRANSLATION && h.u.translation.child < n) {
        flags = INST_TRANSLATE; child = h.u.translation.child;
        ox = h.u.translation.offset.x; oy = h.u.translation.offset.y; oz = h.u.translation.offset.z;
        if (H(child).type == RT_OBJ_Y_ROTATION && H(child).u.y_rotation.child < n) {
          flags |= INST_ROTATE; sn = H(child).u.y_rotation.sin_theta; cs = H(child).u.y_rotation.cos_theta;
          child = H(child).u.y_rotation.child;
        }
      } else if (h.type == RT_OBJ_Y_ROTATION && h.u.y_rotation.child < n) {
        flags = INST_ROTATE; child = h.u.y_rotation.child; sn = h.u.y_rotation.sin_theta; cs = h.u.y_rotation.cos_theta;
      } else continue;
      if (!fusable_child(child)) continue;
      ++found;
      if (dry_run) continue;
      const uint32_t k = (uint32_t)(out.instances.size() / 2);
      out.instances.push_back(f4(ox, oy, oz, as_float(tag[child])));
      out.instances.push_back(f4(sn, cs, as_float(flags), 0.0f));
      tag[i] = make_instance_ref(k);
      fused[i] = (uint8_t)(1 + (is_prim(child) ? 0 : 1 + (uint32_t)H(child).u.structure.count));
      ++out.fused_instances;
    }
    return found;
  }
  // ---- axis-aligned quads of a linear program: the inside test in reduced form (planar_inside_axis in rt_trace.cuh) ----
  // alpha = dot(w, cross(p - q, v)), beta = dot(w, cross(u, p - q)) (rt.cu:1069-1070).  When w, u and v each have exactly one
  // non-zero component (w along axis k, u and v along the two others), every other term of the two cross products and of the
  // dot product is a product with an exact zero: the reference's fma chains then reduce, rounding for rounding, to
  //   alpha = fl(w.k * fl((p - q).m * s)),   with {m, s} = {a, v.b} or {b, -v.a}  ((k, a, b) cyclic),
  //   beta  = fl(w.k * fl((p - q).m' * s')), with {m', s'} = {b, u.a} or {a, -u.b}
  // (only the sign of an exact zero result can differ, which no comparison and no texture lookup sees).  The record keeps
  // the scale as a vector S with s at position m and zeros elsewhere, so the kernel needs no indexing: fl(pl.m * s) is
  // dot(pl, S) evaluated as the usual fma chain.  Words: {n, d} {q, w.k} {S_alpha, 0} {S_beta, 0}.
  static int single_axis(const rt_vec3 &a) {
    const float c[3] = {a.x, a.y, a.z};
    int k = -1;
    for (int i = 0; i < 3; ++i) {
      if (c[i] != c[i] || std::isinf(c[i])) return -1;
      if (c[i] != 0.0f) { if (k >= 0) return -1; k = i; }
    }
    return k;
  }
  void axis_aligned_words() {
    const size_t n_planar = out.planars.size() / 4;
    std::vector<uint8_t> seen(n_planar, 0);
    std::vector<float4> words(out.planars);
    size_t quads = 0;
    for (size_t op = 0; op + 1 < out.program.size(); op += 2) {
      uint32_t code, arg;
      memcpy(&code, &out.program[op + 1].w, 4); memcpy(&arg, &out.program[op + 1].z, 4);
      if ((code & 7u) == K_TRIANGLE) return;
      if ((code & 7u) != K_QUAD) continue;
      if (arg >= n_planar) return;
      ++quads;
      if (seen[arg]) continue;
      seen[arg] = 1;
      const float4 p0 = out.planars[4 * arg], p1 = out.planars[4 * arg + 1], p2 = out.planars[4 * arg + 2], p3 = out.planars[4 * arg + 3];
      const rt_vec3 nrm = {p0.x, p0.y, p0.z}, w = {p1.w, p2.w, p3.w}, v = {p2.x, p2.y, p2.z}, u = {p3.x, p3.y, p3.z};
      const int k = single_axis(w), kn = single_axis(nrm), jv = single_axis(v), iu = single_axis(u);
      if (k < 0 || kn != k || jv < 0 || iu < 0 || jv == k || iu == k || jv == iu) return;
      if (!std::isfinite(p1.x) || !std::isfinite(p1.y) || !std::isfinite(p1.z) || !std::isfinite(p0.w)) return;
      const int a = (k + 1) % 3, b = (k + 2) % 3;
      const float vc[3] = {v.x, v.y, v.z}, uc[3] = {u.x, u.y, u.z}, wc[3] = {w.x, w.y, w.z};
      float sa[3] = {0.0f, 0.0f, 0.0f}, sb[3] = {0.0f, 0.0f, 0.0f};
      // cross(pl, v).k = fma(pl.a, v.b, -(pl.b * v.a));  cross(u, pl).k = fma(u.a, pl.b, -(u.b * pl.a))
      if (jv == b) sa[a] = vc[b]; else sa[b] = -vc[a];
      if (iu == a) sb[b] = uc[a]; else sb[a] = -uc[b];
      words[4 * arg + 1] = f4(p1.x, p1.y, p1.z, wc[k]);
      words[4 * arg + 2] = f4(sa[0], sa[1], sa[2], 0.0f);
      words[4 * arg + 3] = f4(sb[0], sb[1], sb[2], 0.0f);
    }
    if (quads == 0) return;
    out.planars_axis.swap(words);
    out.program_axis_quads = true;
  }

  // ---- the reference-order walk as a linear program (trace_linear in rt_trace.cuh) ----
  // Static replay of the stack machine of trace_reference with every box entered: a BVH node pushes right then left (left
  // is walked first; a leaf node with left == right tests its object twice, as the reference does), a structure pushes
  // its children in order (the last is walked first), a wrapper is walked as way in / child / way out.  Returns false
  // when the walk does not fit `program_max` ops or a wrapper sits inside a wrapper of its own kind (the reference's
  // in_translation / in_rotation flags then mean something only the stack walk mirrors).
  void program_op(float4 a, float z0, float z1, uint32_t arg, uint32_t code) {
    out.program.push_back(a);
    out.program.push_back(f4(z0, z1, as_float(arg), as_float(code)));
  }
  bool emit_program(uint32_t href, bool in_t, bool in_r, int guard) {
    if (guard > 64 || href >= d.hittable_count || out.program.size() / 2 >= program_max) return false;
    const rt_hittable &h = H(href);
    const uint32_t inside = (in_t || in_r) ? 0x100u : 0u;
    switch (h.type) {
      case RT_OBJ_SPHERE:
        program_op(f4(0, 0, 0, 0), 0, 0, ref_index(tag[href]), ref_kind(tag[href]) | inside);
        ++out.program_prims;
        return true;
      case RT_OBJ_TRIANGLE: case RT_OBJ_QUAD:                  // the plane (first word of the record) rides in the op
        program_op(out.planars[4 * (size_t)ref_index(tag[href])], 0, 0, ref_index(tag[href]), ref_kind(tag[href]) | inside);
        ++out.program_prims;
        return true;
      case RT_OBJ_BVH_NODE: {
        if (h.u.bvh.left >= d.world_count || h.u.bvh.right >= d.world_count) return false;
        const size_t me = out.program.size() / 2;
        const rt_aabb &b = h.u.bvh.bbox;
        program_op(f4(b.x0, b.x1, b.y0, b.y1), b.z0, b.z1, 0, K_BVH);
        ++out.program_boxes;
        if (!emit_program(d.world_refs[h.u.bvh.left], in_t, in_r, guard + 1)) return false;
        if (!emit_program(d.world_refs[h.u.bvh.right], in_t, in_r, guard + 1)) return false;
        out.program[2 * me + 1].z = as_float((uint32_t)(out.program.size() / 2));
        return true;
      }
      case RT_OBJ_TRANSLATION: {
        if (in_t) return false;
        const rt_vec3 &o = h.u.translation.offset;
        // translate(rotate(x)) -- every instanced box of the reference's scenes -- as ONE op each way: the two ways in follow
        // each other without anything in between, and so do the two ways out, so the ops do the same arithmetic in the same
        // order and a lane spends two steps of the walk less per instance side (OP_FUSED_WRAPPER, rt_trace.cuh)
        if (fuse_program_wrappers && !in_r && h.u.translation.child < d.hittable_count && H(h.u.translation.child).type == RT_OBJ_Y_ROTATION) {
          const rt_hittable &rot = H(h.u.translation.child);
          const float sn = rot.u.y_rotation.sin_theta, cs = rot.u.y_rotation.cos_theta;
          program_op(f4(o.x, o.y, o.z, 0), sn, cs, 0, K_TRANSLATION | 0x200u);
          out.program_wrappers += 2;
          if (!emit_program(rot.u.y_rotation.child, true, true, guard + 2)) return false;
          program_op(f4(sn, cs, o.x, o.y), o.z, 0, 0, 7u | 0x200u);
          return out.program.size() / 2 <= program_max;
        }
        program_op(f4(o.x, o.y, o.z, 0), 0, 0, 0, K_TRANSLATION);
        ++out.program_wrappers;
        if (!emit_program(h.u.translation.child, true, in_r, guard + 1)) return false;
        program_op(f4(o.x, o.y, o.z, 0), 0, 0, 0, 6u);
        return out.program.size() / 2 <= program_max;
      }
      case RT_OBJ_Y_ROTATION: {
        if (in_r) return false;
        program_op(f4(h.u.y_rotation.sin_theta, h.u.y_rotation.cos_theta, 0, 0), 0, 0, 0, K_YROTATION);
        ++out.program_wrappers;
        if (!emit_program(h.u.y_rotation.child, in_t, true, guard + 1)) return false;
        program_op(f4(h.u.y_rotation.sin_theta, h.u.y_rotation.cos_theta, 0, 0), 0, 0, 0, 7u);
        return out.program.size() / 2 <= program_max;
      }
      case RT_OBJ_STRUCTURE:
        if ((uint64_t)h.u.structure.first + h.u.structure.count > d.hittable_count) return false;
        for (uint32_t k = h.u.structure.count; k-- > 0;)
          if (!emit_program(h.u.structure.first + k, in_t, in_r, guard + 1)) return false;
        return true;
      default: return false;
    }
  }
  static void center_leaf(Leaf &lf) { for (int a = 0; a < 3; ++a) lf.centroid[a] = 0.5f * (lf.box.lo[a] + lf.box.hi[a]); }

  // world-space copy of everything below `href` (FAST traversal)
  void flatten(uint32_t href, const Xf &xf, int guard) {
    if (guard > 64) { too_deep = true; return; }
    if (href >= d.hittable_count) return;
    const rt_hittable &h = H(href);
    switch (h.type) {
      case RT_OBJ_SPHERE: {
        Leaf lf;
        if (xf.identity) { lf.ref = tag[href]; lf.box = box_of(h.u.sphere.bbox); }
        else {
          V c = xf_point(xf, vv(h.u.sphere.center)), mv = xf_dir(xf, vv(h.u.sphere.direction));
          float r = h.u.sphere.radius, pad = 1e-4f * (1.0f + fabsf(r));
          lf.ref = xf.simple ? instanced_leaf(xf, tag[href]) : make_ref(K_SPHERE, add_sphere(c, mv, r, h.u.sphere.mat));
          lf.box = box_empty();
          for (int e = 0; e < 2; ++e) {
            V p{c.x + e * mv.x, c.y + e * mv.y, c.z + e * mv.z};
            box_point(lf.box, V{p.x - r - pad, p.y - r - pad, p.z - r - pad});
            box_point(lf.box, V{p.x + r + pad, p.y + r + pad, p.z + r + pad});
          }
        }
        push_leaf(lf);
      } break;
      case RT_OBJ_TRIANGLE: case RT_OBJ_QUAD: {
        Leaf lf;
        const bool tri = h.type == RT_OBJ_TRIANGLE;
        if (xf.identity) { lf.ref = tag[href]; lf.box = box_of(h.u.planar.bbox); }
        else {
          rt_geometry g;
          V q = xf_point(xf, vv(h.u.planar.geom.q)), u = xf_dir(xf, vv(h.u.planar.geom.u)), v = xf_dir(xf, vv(h.u.planar.geom.v));
          V n{u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x};       // make_quad (rt.cu:604-620)
          float nn = n.x * n.x + n.y * n.y + n.z * n.z, len = sqrtf(nn);
          g.q = rt_vec3{q.x, q.y, q.z}; g.u = rt_vec3{u.x, u.y, u.z}; g.v = rt_vec3{v.x, v.y, v.z};
          g.w = rt_vec3{n.x * (1.0f / nn), n.y * (1.0f / nn), n.z * (1.0f / nn)};
          g.normal = len == 0.0f ? rt_vec3{0, 0, 0} : rt_vec3{n.x / len, n.y / len, n.z / len};
          g.d = g.normal.x * q.x + g.normal.y * q.y + g.normal.z * q.z;
          lf.ref = xf.simple ? instanced_leaf(xf, tag[href]) : make_ref(tri ? K_TRIANGLE : K_QUAD, add_planar(g, h.u.planar.mat));
          lf.box = box_empty();
          box_point(lf.box, q); box_point(lf.box, V{q.x + u.x, q.y + u.y, q.z + u.z});
          box_point(lf.box, V{q.x + v.x, q.y + v.y, q.z + v.z});
          if (!tri) box_point(lf.box, V{q.x + u.x + v.x, q.y + u.y + v.y, q.z + u.z + v.z});
          for (int a = 0; a < 3; ++a) {
            float pad = 1e-3f + 1e-5f * std::max(fabsf(lf.box.lo[a]), fabsf(lf.box.hi[a]));
            lf.box.lo[a] -= pad; lf.box.hi[a] += pad;
          }
        }
        push_leaf(lf);
      } break;
      case RT_OBJ_TRANSLATION: {
        Xf x = xf; V o = xf_dir(xf, vv(h.u.translation.offset));
        x.t = V{xf.t.x + o.x, xf.t.y + o.y, xf.t.z + o.z}; x.identity = false;
        if (xf.simple && xf.flags == 0) {                  // the outermost wrapper: still [translation]
          x.flags = INST_TRANSLATE; x.off[0] = h.u.translation.offset.x; x.off[1] = h.u.translation.offset.y; x.off[2] = h.u.translation.offset.z;
        } else x.simple = false;
        flatten(h.u.translation.child, x, guard + 1);
      } break;
      case RT_OBJ_Y_ROTATION: {
        Xf x = xf; float c2 = h.u.y_rotation.cos_theta, s2 = h.u.y_rotation.sin_theta;
        x.c = xf.c * c2 - xf.s * s2; x.s = xf.s * c2 + xf.c * s2; x.identity = false;
        if (xf.simple && !(xf.flags & INST_ROTATE)) { x.flags |= INST_ROTATE; x.sn = s2; x.cs = c2; }   // [translation] y-rotation
        else x.simple = false;
        flatten(h.u.y_rotation.child, x, guard + 1);
      } break;
      case RT_OBJ_STRUCTURE:
        for (uint32_t k = 0; k < h.u.structure.count; ++k) flatten(h.u.structure.first + k, xf, guard + 1);
        break;
      default: break;
    }
  }
  // FAST traversal: a primitive under translation(y-rotation(.)), translation(.) or y-rotation(.) is NOT copied to world
  // space: the leaf names an instance record (same layout as the fused instances of the reference-order walk) and the walk
  // moves the ray into the primitive's space with the reference's own roundings, tests the reference's own record there and
  // moves the hit back.  The reference's images carry the rounding behaviour of that arithmetic (a 10-unit sphere 1000 units
  // from the camera is hit within ~0.02 units of its surface, inside or outside, and a bounce ray that starts inside hits the
  // sphere again from within): world-space copies had other roundings and a measurably different picture of the 1000-sphere
  // cluster of scene 3 (profiles/r2c_fast_bias_scene3.jsonl); this keeps the statistics.
  uint32_t instanced_leaf(const Xf &xf, uint32_t child_tag) {
    const uint32_t k = (uint32_t)(out.instances.size() / 2);
    out.instances.push_back(f4(xf.off[0], xf.off[1], xf.off[2], as_float(child_tag)));
    out.instances.push_back(f4(xf.sn, xf.cs, as_float(xf.flags), 0.0f));
    ++out.instanced_leaves;
    return make_ref(K_TRANSLATION, k);
  }
  void push_leaf(Leaf lf) {
    for (int a = 0; a < 3; ++a) lf.centroid[a] = 0.5f * (lf.box.lo[a] + lf.box.hi[a]);
    leaves.push_back(lf);
  }

  int run(uint32_t traversal) {
    const uint64_t n = d.hittable_count;
    if (!d.hittables || !d.world_refs || n == 0 || d.world_count == 0 || d.root_wref >= d.world_count) {
      out.error = "rt_scene_upload: empty or inconsistent scene description"; return RT_ERR_INVALID;
    }
    if (n > INDEX_MASK) { out.error = "rt_scene_upload: too many hittables"; return RT_ERR_UNSUPPORTED; }
    for (uint64_t i = 0; i < d.world_count; ++i)
      if (d.world_refs[i] >= n) { out.error = "rt_scene_upload: world ref out of range"; return RT_ERR_INVALID; }
    // 16-byte aligned copy of the arena (TK_ARENA fallback path)
    out.tex_arena.assign(d.texture_arena, d.texture_arena + d.texture_bytes);
    out.tex_arena.resize((out.tex_arena.size() + 31) & ~size_t(15), 0);

    // pass 1: one record per hittable, kinds in their own arrays
    tag.assign(n, make_ref(K_NONE, 0));
    uint32_t n_bvh = 0, n_xf = 0, n_st = 0;
    for (uint64_t i = 0; i < n; ++i) {
      const rt_hittable &h = H((uint32_t)i);
      switch (h.type) {
        case RT_OBJ_SPHERE: tag[i] = make_ref(K_SPHERE, add_sphere(vv(h.u.sphere.center), vv(h.u.sphere.direction), h.u.sphere.radius, h.u.sphere.mat)); break;
        case RT_OBJ_TRIANGLE: tag[i] = make_ref(K_TRIANGLE, add_planar(h.u.planar.geom, h.u.planar.mat)); break;
        case RT_OBJ_QUAD: tag[i] = make_ref(K_QUAD, add_planar(h.u.planar.geom, h.u.planar.mat)); break;
        case RT_OBJ_BVH_NODE: tag[i] = make_ref(K_BVH, n_bvh++); break;
        case RT_OBJ_TRANSLATION: tag[i] = make_ref(K_TRANSLATION, n_xf++); out.has_wrappers = true; break;
        case RT_OBJ_Y_ROTATION: tag[i] = make_ref(K_YROTATION, n_xf++); out.has_wrappers = true; break;
        case RT_OBJ_STRUCTURE: tag[i] = make_ref(K_STRUCTURE, n_st++); break;
        default: out.error = "rt_scene_upload: unknown hittable type"; return RT_ERR_INVALID;
      }
    }
    // wrapper-free subtrees of the reference BVH -> own SAH treelets (their roots' tags change before links are written)
    pure_prims.assign(n, -2); treelet_depth.assign(n, 0); fused.assign(n, 0);
    const bool want_fast = traversal == RT_TRAVERSAL_FAST || (traversal == RT_TRAVERSAL_AUTO && !out.has_wrappers);
    if (traversal == RT_TRAVERSAL_AUTO && out.has_wrappers && treelet_min > 0 && d.world_refs[d.root_wref] < n)
      plant_treelets(d.world_refs[d.root_wref], treelet_min, 0);
    // measured (profiles/r1_configs_treelets.jsonl): the walk with the extra cases only pays when it has work to do --
    // a Cornell box with two instances is faster through the plain reference-order kernel
    if (traversal == RT_TRAVERSAL_AUTO && out.has_wrappers && fuse_instances &&
        (out.treelets > 0 || plant_instances(true) >= fuse_min)) plant_instances(false);
    // pass 2: links
    out.ref_nodes.resize(2 * (size_t)n_bvh); out.xforms.resize(n_xf); out.structures.resize(n_st);
    for (uint64_t i = 0; i < n; ++i) {
      const rt_hittable &h = H((uint32_t)i);
      const uint32_t idx = ref_index(tag[i]);
      if (h.type == RT_OBJ_BVH_NODE) {
        if (treelet_depth[i] > 0) continue;                 // replaced: its links are never followed
        if (h.u.bvh.left >= d.world_count || h.u.bvh.right >= d.world_count) { out.error = "rt_scene_upload: BVH child out of range"; return RT_ERR_INVALID; }
        const rt_aabb &b = h.u.bvh.bbox;
        out.ref_nodes[2 * idx] = f4(b.x0, b.x1, b.y0, b.y1);
        out.ref_nodes[2 * idx + 1] = f4(b.z0, b.z1, as_float(tag[d.world_refs[h.u.bvh.left]]), as_float(tag[d.world_refs[h.u.bvh.right]]));
      } else if (fused[i]) {
        continue;                                          // a fused instance: its record is in out.instances
      } else if (h.type == RT_OBJ_TRANSLATION) {
        if (h.u.translation.child >= n) { out.error = "rt_scene_upload: wrapper child out of range"; return RT_ERR_INVALID; }
        out.xforms[idx] = f4(h.u.translation.offset.x, h.u.translation.offset.y, h.u.translation.offset.z, as_float(tag[h.u.translation.child]));
      } else if (h.type == RT_OBJ_Y_ROTATION) {
        if (h.u.y_rotation.child >= n) { out.error = "rt_scene_upload: wrapper child out of range"; return RT_ERR_INVALID; }
        out.xforms[idx] = f4(h.u.y_rotation.sin_theta, h.u.y_rotation.cos_theta, 0.0f, as_float(tag[h.u.y_rotation.child]));
      } else if (h.type == RT_OBJ_STRUCTURE) {
        if ((uint64_t)h.u.structure.first + h.u.structure.count > n) { out.error = "rt_scene_upload: structure out of range"; return RT_ERR_INVALID; }
        // each structure's children start on a 16-byte boundary and are padded to whole uint4s (the walk loads 4 at a time)
        uint2 s; s.x = (uint32_t)out.struct_children.size(); s.y = (uint32_t)h.u.structure.count;
        for (uint32_t k = 0; k < s.y; ++k) out.struct_children.push_back(tag[h.u.structure.first + k]);
        while (out.struct_children.size() & 3u) out.struct_children.push_back(make_ref(K_NONE, 0));
        out.structures[idx] = s;
      }
    }
    const uint32_t root_href = d.world_refs[d.root_wref];
    out.ref_root = tag[root_href];
    out.ref_stack_need = stack_need(root_href, 0, 0) + 2;        // + the root, + the bottom sentinel (rt_trace.cuh)
    // small scenes with wrappers (where neither treelets nor fused instances were worth planting): the same walk as a
    // linear program.  RT_TRAVERSAL_REFERENCE stays the literal stack walk (the mode to bisect against).
    if (traversal == RT_TRAVERSAL_AUTO && !want_fast && program_max > 0 && out.treelets + out.fused_instances == 0 &&
        !emit_program(root_href, false, false, 0)) {
      out.program.clear(); out.program_boxes = out.program_prims = out.program_wrappers = 0;
    }

    if (!out.program.empty() &&
        linear_stage_bytes((uint32_t)(out.program.size() / 2), (uint32_t)(out.planars.size() / 4), (uint32_t)(out.spheres.size() / 2)) > LINEAR_SMEM_MAX_BYTES) {
      out.program.clear(); out.program_boxes = out.program_prims = out.program_wrappers = 0;   // would not fit its staging area
    }
    if (!out.program.empty()) axis_aligned_words();

    // own BVH over the whole scene
    if (want_fast) {
      Xf id;
      for (uint64_t w = 0; w < d.world_count; ++w) {
        const uint32_t href = d.world_refs[w];
        if (H(href).type != RT_OBJ_BVH_NODE) flatten(href, id, 0);
      }
      if (too_deep) { out.error = "rt_scene_upload: objects nested deeper than 64 levels"; return RT_ERR_UNSUPPORTED; }
      if (leaves.empty()) { out.error = "rt_scene_upload: no primitives"; return RT_ERR_INVALID; }
      if (leaves.size() == 1) {                       // a root node with one real child and one empty box
        out.fast_nodes.resize(4);
        const Box &b = leaves[0].box;
        out.fast_nodes[0] = f4(b.lo[0], b.hi[0], b.lo[1], b.hi[1]);
        out.fast_nodes[1] = f4(INFINITY, -INFINITY, INFINITY, -INFINITY);
        out.fast_nodes[2] = f4(b.lo[2], b.hi[2], INFINITY, -INFINITY);
        out.fast_nodes[3] = f4(as_float(leaves[0].ref), as_float(make_ref(K_NONE, 0)), 0.0f, 0.0f);
        out.fast_root = make_ref(K_BVH, 0);
        out.fast_stack_need = 3;
      } else {
        SahBuilder sb(leaves, out.fast_nodes);
        Box rb;
        out.fast_root = sb.build(0, leaves.size(), 0, rb);
        out.fast_stack_need = sb.max_depth + 3;               // + the bottom sentinel
      }
      out.has_fast = true;
    }
    out.traversal_used = want_fast ? RT_TRAVERSAL_FAST : RT_TRAVERSAL_REFERENCE;
    if (too_deep) { out.error = "rt_scene_upload: objects nested deeper than 64 levels"; return RT_ERR_UNSUPPORTED; }
    // tiny scenes without wrappers whose every primitive record is reachable from the world list (so the primitive arrays
    // are exactly the scene; an orphan hittable would be tested here and by no other walk): no BVH at all
    if (want_fast && !out.has_wrappers && out.spheres.size() / 2 <= 4 && out.planars.size() / 4 <= 4 &&
        leaves.size() == out.spheres.size() / 2 + out.planars.size() / 4) {
      out.flat = true;
      uint32_t k = 0;
      for (uint64_t i = 0; i < n; ++i) {
        if (H((uint32_t)i).type == RT_OBJ_TRIANGLE) out.flat_triangles |= 1u << k;
        if (H((uint32_t)i).type == RT_OBJ_TRIANGLE || H((uint32_t)i).type == RT_OBJ_QUAD) ++k;
      }
    }
    const uint32_t need = want_fast ? out.fast_stack_need : out.ref_stack_need;
    if (need > (uint32_t)STACK_CAPACITY) {
      out.error = "rt_scene_upload: traversal stack need " + std::to_string(need) + " exceeds device capacity " + std::to_string(STACK_CAPACITY);
      return RT_ERR_UNSUPPORTED;
    }
    // materials: prim id = sphere index, or sphere_count + planar index
    out.sphere_count = (uint32_t)(out.spheres.size() / 2);
    out.planar_count = (uint32_t)(out.planars.size() / 4);
    out.materials.resize(2 * (size_t)(out.sphere_count + out.planar_count));
    for (uint32_t i = 0; i < out.sphere_count; ++i) pack_material(sphere_mats[i], d.texture_arena, d.texture_bytes, out, &out.materials[2 * i]);
    for (uint32_t i = 0; i < out.planar_count; ++i)
      pack_material(planar_mats[i], d.texture_arena, d.texture_bytes, out, &out.materials[2 * (size_t)(out.sphere_count + i)]);
    return RT_OK;
  }
};

}  // namespace

void build_linear_image(const PackedScene &p, bool axis, std::vector<float4> &image) {
  image.clear();
  if (p.program.empty()) return;
  const size_t n = p.program.size() / 2, n1 = n + 1;
  const std::vector<float4> &planars = (axis && p.program_axis_quads) ? p.planars_axis : p.planars;
  image.reserve(2 * n1 + planars.size() + p.spheres.size());
  image.assign(p.program.begin(), p.program.end());
  for (size_t i = 0; i < n; ++i) {
    float4 &w1 = image[2 * i + 1];
    uint32_t code, arg;
    memcpy(&code, &w1.w, 4); memcpy(&arg, &w1.z, 4);
    int32_t delta;
    switch (code & 7u) {
      case K_BVH:                                          // skip target: op index -> bytes from this op
        delta = ((int32_t)arg - (int32_t)i) * 32;
        memcpy(&w1.z, &delta, 4);
        code |= 0x10u;                                     // OP_IS_BOX (rt_trace.cuh): the two frequent kinds get a flag bit each
        memcpy(&w1.w, &code, 4);
        break;
      case K_TRIANGLE: case K_QUAD:
        delta = (int32_t)((n1 - i) * 32 + 64 * (size_t)arg);
        memcpy(&w1.x, &delta, 4);
        code |= 0x8u;                                      // OP_IS_PLANAR
        memcpy(&w1.w, &code, 4);
        break;
      case K_SPHERE:
        delta = (int32_t)((n1 - i) * 32 + 16 * planars.size() + 32 * (size_t)arg);
        memcpy(&w1.x, &delta, 4);
        break;
      default: break;
    }
  }
  const uint32_t end_code = K_SPHERE | 0x400u;
  float4 e0 = {0, 0, 0, 0}, e1 = {0, 0, 0, 0};
  memcpy(&e1.w, &end_code, 4);
  image.push_back(e0); image.push_back(e1);
  image.insert(image.end(), planars.begin(), planars.end());
  image.insert(image.end(), p.spheres.begin(), p.spheres.end());
}

int pack_scene(const rt_scene_desc &desc, uint32_t traversal, PackedScene &out, uint32_t treelet_min, uint32_t fuse_min,
               uint32_t program_max) {
  if (traversal > RT_TRAVERSAL_FAST) { out.error = "rt_scene_upload: unknown traversal mode"; return RT_ERR_INVALID; }
  Packer p(desc, out, treelet_min, fuse_min, program_max);
  return p.run(traversal);
}

// ---------------------------------------------------------------------------------------- XORWOW skip-ahead tables
// One XORWOW step is linear over GF(2) on the 160 state bits v[0..4] (the Weyl counter d is separate and a
// subsequence jump of 2^67 steps leaves it unchanged mod 2^32).  Row r of a matrix is the image of unit vector e_r.
void xorwow_sequence_matrices(std::vector<uint32_t> &out) {
  const int R = 160, W = 5;
  std::vector<uint32_t> m(R * W), t(R * W);
  for (int r = 0; r < R; ++r) {
    uint32_t v[5] = {0, 0, 0, 0, 0};
    v[r / 32] = 1u << (r & 31);
    uint32_t x = v[0] ^ (v[0] >> 2);
    v[0] = v[1]; v[1] = v[2]; v[2] = v[3]; v[3] = v[4];
    v[4] = (v[4] ^ (v[4] << 4)) ^ (x ^ (x << 1));
    for (int k = 0; k < W; ++k) m[r * W + k] = v[k];
  }
  auto square = [&]() {
    for (int r = 0; r < R; ++r) {
      uint32_t acc[5] = {0, 0, 0, 0, 0};
      for (int i = 0; i < W; ++i)
        for (int j = 0; j < 32; ++j)
          if (m[r * W + i] & (1u << j))
            for (int k = 0; k < W; ++k) acc[k] ^= m[(i * 32 + j) * W + k];
      for (int k = 0; k < W; ++k) t[r * W + k] = acc[k];
    }
    m.swap(t);
  };
  for (int i = 0; i < 67; ++i) square();
  out.resize((size_t)64 * R * W);
  for (int k = 0; k < 64; ++k) {
    memcpy(&out[(size_t)k * R * W], m.data(), sizeof(uint32_t) * R * W);
    square();
  }
}

}  // namespace rtp
