/* rt_world.c -- host-side scene construction in C (product code).
 *
 * Re-implements the BEHAVIOUR of the reference's L2 layer (luk-zm/cuda_raytracing
 * src/raytracing.cu:504-981: material/primitive constructors, AABBs, the merge sort and the
 * random-axis median-split BVH builder) so that the three flat arrays this file produces are
 * bit-identical to the reference's for the same glibc rand() stream.  Every f32 expression keeps
 * the reference's operation order (compiled with -ffp-contract=off; the reference's host pass is
 * g++ -O3 on x86-64 without FMA).  tests/test_scene_build.py diffs all 20 scenes against the real
 * reference host code (oracle/_ref/libref_host.so) and against committed digests.
 */
#include "rt_scene.h"

#include <assert.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

_Static_assert(sizeof(rt_vec3) == 12, "vec3");
_Static_assert(sizeof(rt_aabb) == 24, "aabb");
_Static_assert(sizeof(rt_texture) == 32, "Texture is 32 B in the reference");
_Static_assert(sizeof(rt_perlin) == 6144, "PerlinNoise is 6144 B in the reference");
_Static_assert(sizeof(rt_material) == 24, "Material is 24 B in the reference");
_Static_assert(sizeof(rt_sphere) == 80 && sizeof(rt_planar) == 112 && sizeof(rt_geometry) == 64, "prims");
_Static_assert(sizeof(rt_bvh_node) == 28 && sizeof(rt_translation) == 40, "bvh/translation");
_Static_assert(sizeof(rt_y_rotation) == 36 && sizeof(rt_structure) == 48, "yrot/structure");
_Static_assert(sizeof(rt_hittable) == 120, "Hittable is 120 B in the reference");
_Static_assert(offsetof(rt_hittable, u.sphere.mat) == 40 && offsetof(rt_hittable, u.sphere.bbox) == 64, "sphere");
_Static_assert(offsetof(rt_hittable, u.planar.geom.d) == 44 && offsetof(rt_hittable, u.planar.geom.normal) == 60, "geom");
_Static_assert(offsetof(rt_hittable, u.planar.mat) == 72 && offsetof(rt_hittable, u.planar.bbox) == 96, "planar");
_Static_assert(offsetof(rt_hittable, u.bvh.bbox) == 12 && offsetof(rt_hittable, u.translation.offset) == 12, "bvh");
_Static_assert(offsetof(rt_hittable, u.y_rotation.sin_theta) == 12 && offsetof(rt_hittable, u.structure.count) == 48, "yrot");
_Static_assert(offsetof(rt_texture, u.checker.even) == 16 && offsetof(rt_texture, u.noise.scale) == 24, "texture");
_Static_assert(offsetof(rt_material, u.metal.fuzz) == 20, "material");

/* ---------- f32 vector helpers, operation order as src/vec3.h ---------- */
static inline rt_vec3 v3(float x, float y, float z) { rt_vec3 r = {x, y, z}; return r; }
static inline rt_vec3 v3_add(rt_vec3 a, rt_vec3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline rt_vec3 v3_sub(rt_vec3 a, rt_vec3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline rt_vec3 v3_scale(float s, rt_vec3 a) { return v3(a.x * s, a.y * s, a.z * s); }
static inline rt_vec3 v3_neg(rt_vec3 a) { return v3(-a.x, -a.y, -a.z); }
static inline float v3_dot(rt_vec3 a, rt_vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline rt_vec3 v3_cross(rt_vec3 a, rt_vec3 b) {
  return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline float v3_len(rt_vec3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }
static inline rt_vec3 v3_unit(rt_vec3 a) {                                   /* vec3.h:105-114 */
  float len = v3_len(a);
  if (len == 0.0f) return v3(0.0f, 0.0f, 0.0f);
  return v3(a.x / len, a.y / len, a.z / len);
}
/* utility.h:6-7: MIN/MAX are (a<b) selects, which is what fixes their NaN/-0 behaviour */
#define RT_MIN(a, b) (((a) < (b)) ? (a) : (b))
#define RT_MAX(a, b) (((a) < (b)) ? (b) : (a))

/* ---------- host RNG: unseeded glibc rand() (utility.h:26-41) ---------- */
void rt_host_rand_reset(void) { srand(1); }
float rt_host_random(void) { return rand() / (RAND_MAX + 1.0f); }
float rt_host_random_range(float lo, float hi) { return lo + (hi - lo) * rt_host_random(); }
int32_t rt_host_random_int(int32_t lo, int32_t hi) { return (int)(rt_host_random_range((float)lo, (float)(hi + 1))); }

/* g++ evaluates the three arguments of the reference's Vec3(r(), r(), r()) right to left
 * (SURVEY.md Appendix A Q2): z takes the first draw. */
static rt_vec3 random_vec3(void) {
  rt_vec3 r; r.z = rt_host_random(); r.y = rt_host_random(); r.x = rt_host_random(); return r;
}
static rt_vec3 random_vec3_range(float lo, float hi) {
  rt_vec3 r; r.z = rt_host_random_range(lo, hi); r.y = rt_host_random_range(lo, hi); r.x = rt_host_random_range(lo, hi);
  return r;
}
rt_vec3 rt_host_random_vec3(void) { return random_vec3(); }
rt_vec3 rt_host_random_vec3_range(float lo, float hi) { return random_vec3_range(lo, hi); }

/* ---------- world container ---------- */
int rt_world_init(rt_world *w) {
  memset(w, 0, sizeof(*w));
  w->hittable_cap = RT_WORLD_CAPACITY;
  w->world_cap = RT_WORLD_CAPACITY;
  w->texture_cap = RT_TEXTURE_ARENA_BYTES;
  w->hittables = (rt_hittable *)calloc(w->hittable_cap, sizeof(rt_hittable));
  w->world_refs = (rt_href *)calloc(w->world_cap, sizeof(rt_href));
  w->texture_arena = (uint8_t *)calloc(w->texture_cap, 1);
  if (!w->hittables || !w->world_refs || !w->texture_arena) { rt_world_free(w); return -1; }
  return 0;
}

void rt_world_free(rt_world *w) {
  free(w->hittables); free(w->world_refs); free(w->texture_arena);
  memset(w, 0, sizeof(*w));
}

static void *arena_push(rt_world *w, uint64_t size) {                        /* memory.h:31-39: no alignment */
  if (w->texture_bytes + size > w->texture_cap) return NULL;
  void *p = w->texture_arena + w->texture_bytes;
  w->texture_bytes += size;
  return p;
}

static uint64_t texture_append(rt_world *w, const rt_texture *tex, const void *payload, uint64_t payload_bytes) {
  uint64_t offset = w->texture_bytes;
  rt_texture *slot = (rt_texture *)arena_push(w, sizeof(rt_texture));
  assert(slot);
  memcpy(slot, tex, sizeof(rt_texture));
  if (tex->type == RT_TEX_IMAGE) {
    slot->u.image.pixels_offset = w->texture_bytes;
  } else if (tex->type == RT_TEX_NOISE) {
    slot->u.noise.noise_offset = w->texture_bytes;
  }
  if (tex->type == RT_TEX_IMAGE || tex->type == RT_TEX_NOISE) {
    void *dst = arena_push(w, payload_bytes);
    assert(dst || payload_bytes == 0);
    if (payload_bytes) memcpy(dst, payload, payload_bytes);
  }
  w->texture_count++;
  return offset;
}

static uint64_t add_solid_texture(rt_world *w, rt_vec3 color) {
  rt_texture t; memset(&t, 0, sizeof(t));
  t.type = RT_TEX_SOLID; t.u.solid.albedo = color;
  return texture_append(w, &t, NULL, 0);
}

/* ---------- materials (rt.cu:1805-1854, 509-536) ---------- */
static rt_material material_zero(uint32_t type) { rt_material m; memset(&m, 0, sizeof(m)); m.type = type; return m; }

rt_material rt_mat_lambert_solid(rt_world *w, rt_vec3 color) {
  rt_material m = material_zero(RT_MAT_LAMBERTIAN);
  m.u.texture = add_solid_texture(w, color);
  return m;
}
rt_material rt_mat_light_solid(rt_world *w, rt_vec3 color) {
  rt_material m = material_zero(RT_MAT_DIFFUSE_LIGHT);
  m.u.texture = add_solid_texture(w, color);
  return m;
}
rt_material rt_mat_lambert_image(rt_world *w, const char *file_name) {
  /* The reference decodes with stb_image (third party, out of scope: SURVEY.md section 8 row f3) and its
   * repository ships no image, so stbi_loadf() returns NULL and width = height = 0 (rt.cu:332-341).
   * We produce that same record: pixels == NULL, no payload. */
  (void)file_name;
  rt_material m = material_zero(RT_MAT_LAMBERTIAN);
  rt_texture t; memset(&t, 0, sizeof(t));
  t.type = RT_TEX_IMAGE;
  m.u.texture = texture_append(w, &t, NULL, 0);
  return m;
}
rt_material rt_mat_lambert_image_pixels(rt_world *w, const float *rgb, int32_t width, int32_t height) {
  rt_material m = material_zero(RT_MAT_LAMBERTIAN);
  rt_texture t; memset(&t, 0, sizeof(t));
  t.type = RT_TEX_IMAGE;
  /* (height + 1) rows + 1 texel: everything the reference's inclusive clamp can address (rt.cu:349-352) */
  const uint64_t texels = (width > 0 && height > 0) ? ((uint64_t)height + 1) * (uint64_t)width + 1 : 0;
  const uint64_t bytes = texels * 3 * sizeof(float);
  if (!rgb || texels == 0 || w->texture_bytes + sizeof(rt_texture) + bytes > w->texture_cap) {
    m.u.texture = texture_append(w, &t, NULL, 0);
    return m;
  }
  float *padded = (float *)calloc(texels * 3, sizeof(float));
  memcpy(padded, rgb, (uint64_t)width * height * 3 * sizeof(float));
  t.u.image.width = width; t.u.image.height = height;
  t.u.image.pixels = (void *)(w->texture_arena + w->texture_bytes + sizeof(rt_texture));   /* non-NULL: "decoded" */
  m.u.texture = texture_append(w, &t, padded, bytes);
  free(padded);
  return m;
}
static void perlin_permutation(int32_t *arr, int32_t count) {               /* perlin_noise.cpp:48-59 */
  for (int32_t i = 0; i < count; ++i) arr[i] = i;
  for (int32_t i = count - 1; i > 0; --i) {
    int32_t target = rt_host_random_int(0, i);
    int32_t tmp = arr[i]; arr[i] = arr[target]; arr[target] = tmp;
  }
}
rt_material rt_mat_lambert_noise(rt_world *w, float scale) {
  rt_perlin *noise = (rt_perlin *)calloc(1, sizeof(rt_perlin));             /* perlin_noise.cpp:61-70 */
  for (int i = 0; i < RT_PERLIN_POINTS; ++i) noise->vectors[i] = v3_unit(random_vec3_range(-1.0f, 1.0f));
  perlin_permutation(noise->perm_x, RT_PERLIN_POINTS);
  perlin_permutation(noise->perm_y, RT_PERLIN_POINTS);
  perlin_permutation(noise->perm_z, RT_PERLIN_POINTS);
  rt_material m = material_zero(RT_MAT_LAMBERTIAN);
  rt_texture t; memset(&t, 0, sizeof(t));
  t.type = RT_TEX_NOISE; t.u.noise.scale = scale;
  m.u.texture = texture_append(w, &t, noise, sizeof(rt_perlin));
  free(noise);
  return m;
}
rt_material rt_mat_lambert_checker(rt_world *w, float scale, rt_vec3 even, rt_vec3 odd) {
  rt_material m = material_zero(RT_MAT_LAMBERTIAN);
  uint64_t e = add_solid_texture(w, even);
  uint64_t o = add_solid_texture(w, odd);
  rt_texture t; memset(&t, 0, sizeof(t));
  t.type = RT_TEX_CHECKER; t.u.checker.inverted_scale = 1.0f / scale; t.u.checker.even = e; t.u.checker.odd = o;
  m.u.texture = texture_append(w, &t, NULL, 0);
  return m;
}
rt_material rt_mat_metal(rt_vec3 albedo, float fuzz) {
  rt_material m = material_zero(RT_MAT_METAL); m.u.metal.albedo = albedo; m.u.metal.fuzz = fuzz; return m;
}
rt_material rt_mat_dielectric(float refraction_index) {
  rt_material m = material_zero(RT_MAT_DIELECTRIC); m.u.dielectric.refraction_index = refraction_index; return m;
}

/* ---------- AABBs (rt.cu:538-581) ---------- */
static void aabb_pad(rt_aabb *b) {
  const float delta = 0.0001f;
  const float padding = delta / 2;
  if ((b->x1 - b->x0) < delta) { b->x0 -= padding; b->x1 += padding; }
  if ((b->y1 - b->y0) < delta) { b->y0 -= padding; b->y1 += padding; }
  if ((b->z1 - b->z0) < delta) { b->z0 -= padding; b->z1 += padding; }
}
static rt_aabb aabb_from_points(rt_vec3 p1, rt_vec3 p2) {
  rt_aabb r;
  r.x0 = RT_MIN(p1.x, p2.x); r.x1 = RT_MAX(p1.x, p2.x);
  r.y0 = RT_MIN(p1.y, p2.y); r.y1 = RT_MAX(p1.y, p2.y);
  r.z0 = RT_MIN(p1.z, p2.z); r.z1 = RT_MAX(p1.z, p2.z);
  aabb_pad(&r);
  return r;
}
static rt_aabb aabb_union(const rt_aabb *a, const rt_aabb *b) {
  rt_aabb r;
  r.x0 = RT_MIN(a->x0, b->x0); r.x1 = RT_MAX(a->x1, b->x1);
  r.y0 = RT_MIN(a->y0, b->y0); r.y1 = RT_MAX(a->y1, b->y1);
  r.z0 = RT_MIN(a->z0, b->z0); r.z1 = RT_MAX(a->z1, b->z1);
  aabb_pad(&r);
  return r;
}
static const rt_aabb *hittable_bbox(const rt_hittable *h) {                  /* rt.cu:683-709 */
  switch (h->type) {
    case RT_OBJ_SPHERE: return &h->u.sphere.bbox;
    case RT_OBJ_BVH_NODE: return &h->u.bvh.bbox;
    case RT_OBJ_TRIANGLE: case RT_OBJ_QUAD: return &h->u.planar.bbox;
    case RT_OBJ_TRANSLATION: return &h->u.translation.bbox;
    case RT_OBJ_Y_ROTATION: return &h->u.y_rotation.bbox;
    case RT_OBJ_STRUCTURE: return &h->u.structure.bbox;
  }
  return NULL;
}
const rt_aabb *rt_hittable_bbox(const rt_hittable *h) { return hittable_bbox(h); }

/* ---------- primitive records (rt.cu:592-681) ---------- */
static rt_hittable hittable_zero(uint32_t type) { rt_hittable h; memset(&h, 0, sizeof(h)); h.type = type; return h; }

static rt_hittable make_sphere(rt_vec3 center, rt_vec3 direction, int moving, float radius, rt_material mat) {
  rt_hittable h = hittable_zero(RT_OBJ_SPHERE);
  rt_sphere *s = &h.u.sphere;
  s->center = center; s->direction = direction; s->radius = radius; s->mat = mat;
  rt_vec3 rv = v3(radius, radius, radius);
  if (!moving) {
    s->bbox = aabb_from_points(v3_sub(center, rv), v3_add(center, rv));
  } else {
    /* the "destination" argument of the reference is a DIRECTION: position(t) = center + t*direction */
    rt_vec3 p0 = v3_add(center, v3_scale(0.0f, direction));
    rt_vec3 p1 = v3_add(center, v3_scale(1.0f, direction));
    rt_aabb b0 = aabb_from_points(v3_sub(p0, rv), v3_add(p0, rv));
    rt_aabb b1 = aabb_from_points(v3_sub(p1, rv), v3_add(p1, rv));
    s->bbox = aabb_union(&b0, &b1);
  }
  return h;
}

static void planar_geometry(rt_geometry *g, rt_vec3 q, rt_vec3 u, rt_vec3 v) {
  g->q = q; g->u = u; g->v = v;
  rt_vec3 n = v3_cross(u, v);
  g->w = v3_scale(1.0f / v3_dot(n, n), n);
  g->normal = v3_unit(n);
  g->d = v3_dot(g->normal, q);
}
static rt_hittable make_quad(rt_vec3 q, rt_vec3 u, rt_vec3 v, rt_material mat) {
  rt_hittable h = hittable_zero(RT_OBJ_QUAD);
  planar_geometry(&h.u.planar.geom, q, u, v);
  h.u.planar.mat = mat;
  rt_aabb b1 = aabb_from_points(q, v3_add(q, v3_add(u, v)));
  rt_aabb b2 = aabb_from_points(v3_add(q, u), v3_add(q, v));
  h.u.planar.bbox = aabb_union(&b1, &b2);
  return h;
}
static rt_hittable make_triangle(rt_vec3 q, rt_vec3 u, rt_vec3 v, rt_material mat) {
  rt_hittable h = hittable_zero(RT_OBJ_TRIANGLE);
  planar_geometry(&h.u.planar.geom, q, u, v);
  h.u.planar.mat = mat;
  rt_vec3 p1 = v3_add(q, v), p2 = v3_add(q, u);
  rt_aabb b;
  b.x0 = RT_MIN(q.x, RT_MIN(p1.x, p2.x)); b.x1 = RT_MAX(q.x, RT_MAX(p1.x, p2.x));
  b.y0 = RT_MIN(q.y, RT_MIN(p1.y, p2.y)); b.y1 = RT_MAX(q.y, RT_MAX(p1.y, p2.y));
  b.z0 = RT_MIN(q.z, RT_MIN(p1.z, p2.z)); b.z1 = RT_MAX(q.z, RT_MAX(p1.z, p2.z));
  const float eps = 1e-3f;
  b.x0 -= eps; b.x1 += eps; b.y0 -= eps; b.y1 += eps; b.z0 -= eps; b.z1 += eps;
  h.u.planar.bbox = b;
  return h;
}

/* ---------- array appends ---------- */
static uint64_t hittables_append(rt_world *w, const rt_hittable *h) {        /* rt.cu:833-842 */
  uint64_t idx = w->hittable_count;
  assert(w->hittable_count < w->hittable_cap);
  if (w->hittable_count < w->hittable_cap) w->hittables[w->hittable_count++] = *h;
  return idx;
}
static uint64_t world_refs_append(rt_world *w, rt_href ref) {                /* rt.cu:256-261 */
  uint64_t idx = w->world_count;
  if (w->world_count < w->world_cap) w->world_refs[w->world_count++] = ref;
  return idx;
}

/* a group of `size` consecutive hittables (box = 6 quads, pyramid = quad + 4 triangles): rt.cu:908-927 */
static rt_hittable structure_begin(rt_world *w, uint64_t size) {
  rt_hittable h = hittable_zero(RT_OBJ_STRUCTURE);
  h.u.structure.first = (rt_href)w->hittable_count;
  h.u.structure.size = size;
  rt_hittable dummy = hittable_zero(0);
  for (uint64_t i = 0; i < size; ++i) hittables_append(w, &dummy);
  return h;
}
static void structure_push(rt_world *w, rt_hittable *st, const rt_hittable *child) {
  rt_structure *s = &st->u.structure;
  if (s->count <= s->size) {
    uint64_t idx = (uint64_t)s->first + s->count;
    if (idx < w->hittable_count) w->hittables[idx] = *child;
    /* note: the group box starts as all-zero and is only ever merged, so it always contains the origin */
    s->bbox = aabb_union(&s->bbox, hittable_bbox(child));
    s->count += 1;
  }
}

static rt_hittable make_box(rt_world *w, rt_vec3 a, rt_vec3 b, rt_material mat) {   /* rt.cu:929-952 */
  rt_hittable box = structure_begin(w, 6);
  rt_vec3 lo = v3(RT_MIN(a.x, b.x), RT_MIN(a.y, b.y), RT_MIN(a.z, b.z));
  rt_vec3 hi = v3(RT_MAX(a.x, b.x), RT_MAX(a.y, b.y), RT_MAX(a.z, b.z));
  rt_vec3 dx = v3(hi.x - lo.x, 0, 0), dy = v3(0, hi.y - lo.y, 0), dz = v3(0, 0, hi.z - lo.z);
  rt_hittable f;
  f = make_quad(v3(lo.x, lo.y, hi.z), dx, dy, mat);          structure_push(w, &box, &f);   /* front  */
  f = make_quad(v3(hi.x, lo.y, hi.z), v3_neg(dz), dy, mat);  structure_push(w, &box, &f);   /* right  */
  f = make_quad(v3(hi.x, lo.y, lo.z), v3_neg(dx), dy, mat);  structure_push(w, &box, &f);   /* back   */
  f = make_quad(v3(lo.x, lo.y, lo.z), dz, dy, mat);          structure_push(w, &box, &f);   /* left   */
  f = make_quad(v3(lo.x, hi.y, hi.z), dx, v3_neg(dz), mat);  structure_push(w, &box, &f);   /* top    */
  f = make_quad(v3(lo.x, lo.y, lo.z), dx, dz, mat);          structure_push(w, &box, &f);   /* bottom */
  return box;
}

static rt_hittable make_pyramid(rt_world *w, rt_vec3 a, rt_vec3 b, rt_vec3 tip, rt_material mat) {  /* rt.cu:954-980 */
  rt_hittable pyr = structure_begin(w, 5);
  rt_vec3 lo = v3(RT_MIN(a.x, b.x), RT_MIN(a.y, b.y), RT_MIN(a.z, b.z));
  rt_vec3 hi = v3(RT_MAX(a.x, b.x), RT_MAX(a.y, b.y), RT_MAX(a.z, b.z));
  rt_vec3 dx = v3(hi.x - lo.x, 0, 0), dz = v3(0, 0, hi.z - lo.z);
  rt_vec3 p1 = v3(lo.x, lo.y, lo.z), p2 = v3(hi.x, lo.y, lo.z), p3 = v3(hi.x, lo.y, hi.z), p4 = v3(lo.x, lo.y, hi.z);
  rt_hittable f;
  f = make_quad(p1, dx, dz, mat);                               structure_push(w, &pyr, &f);
  f = make_triangle(p1, dx, v3_sub(tip, p1), mat);              structure_push(w, &pyr, &f);
  f = make_triangle(p2, dz, v3_sub(tip, p2), mat);              structure_push(w, &pyr, &f);
  f = make_triangle(p3, v3_neg(dx), v3_sub(tip, p3), mat);      structure_push(w, &pyr, &f);
  f = make_triangle(p4, v3_neg(dz), v3_sub(tip, p4), mat);      structure_push(w, &pyr, &f);
  return pyr;
}

/* ---------- instance wrappers (rt.cu:787-831) ---------- */
static rt_hittable make_translation(const rt_world *w, rt_href child, rt_vec3 offset) {
  rt_hittable h = hittable_zero(RT_OBJ_TRANSLATION);
  h.u.translation.child = child; h.u.translation.offset = offset;
  rt_aabb b = *hittable_bbox(&w->hittables[child]);
  b.x0 += offset.x; b.x1 += offset.x; b.y0 += offset.y; b.y1 += offset.y; b.z0 += offset.z; b.z1 += offset.z;
  h.u.translation.bbox = b;
  return h;
}
static rt_hittable make_y_rotation(const rt_world *w, rt_href child, float angle_deg) {
  rt_hittable h = hittable_zero(RT_OBJ_Y_ROTATION);
  const float pi = 3.1415926535897932385f;
  float radians = angle_deg * (pi / 180);
  rt_y_rotation *r = &h.u.y_rotation;
  r->child = child; r->sin_theta = sinf(radians); r->cos_theta = cosf(radians);
  rt_aabb cb = *hittable_bbox(&w->hittables[child]);
  rt_vec3 lo = v3(INFINITY, INFINITY, INFINITY), hi = v3(-INFINITY, -INFINITY, -INFINITY);
  for (int i = 0; i < 2; ++i) for (int j = 0; j < 2; ++j) for (int k = 0; k < 2; ++k) {
    float x = i * cb.x1 + (1 - i) * cb.x0;
    float y = j * cb.y1 + (1 - j) * cb.y0;
    float z = k * cb.z1 + (1 - k) * cb.z0;
    float nx = r->cos_theta * x + r->sin_theta * z;
    float nz = -r->sin_theta * x + r->cos_theta * z;
    lo.x = RT_MIN(lo.x, nx); hi.x = RT_MAX(hi.x, nx);
    lo.y = RT_MIN(lo.y, y);  hi.y = RT_MAX(hi.y, y);
    lo.z = RT_MIN(lo.z, nz); hi.z = RT_MAX(hi.z, nz);
  }
  r->bbox = aabb_from_points(lo, hi);
  return h;
}

/* world_add / world_add_rotated / _translated / _rotated_translated (rt.cu:1775-1803) */
static uint32_t world_place(rt_world *w, const rt_hittable *obj, rt_xform xf) {
  rt_href ref = (rt_href)hittables_append(w, obj);
  if (xf.rotate) {
    rt_hittable rot = make_y_rotation(w, ref, xf.angle_deg);
    ref = (rt_href)hittables_append(w, &rot);
  }
  if (xf.translate) {
    rt_hittable tr = make_translation(w, ref, xf.offset);
    ref = (rt_href)hittables_append(w, &tr);
  }
  return (uint32_t)world_refs_append(w, ref);
}

uint32_t rt_add_sphere(rt_world *w, rt_xform xf, rt_vec3 center, float radius, rt_material mat) {
  rt_hittable h = make_sphere(center, v3(0.0f, 0.0f, 0.0f), 0, radius, mat);
  return world_place(w, &h, xf);
}
uint32_t rt_add_moving_sphere(rt_world *w, rt_xform xf, rt_vec3 center, rt_vec3 direction, float radius, rt_material mat) {
  rt_hittable h = make_sphere(center, direction, 1, radius, mat);
  return world_place(w, &h, xf);
}
uint32_t rt_add_quad(rt_world *w, rt_xform xf, rt_vec3 q, rt_vec3 u, rt_vec3 v, rt_material mat) {
  rt_hittable h = make_quad(q, u, v, mat);
  return world_place(w, &h, xf);
}
uint32_t rt_add_triangle(rt_world *w, rt_xform xf, rt_vec3 q, rt_vec3 u, rt_vec3 v, rt_material mat) {
  rt_hittable h = make_triangle(q, u, v, mat);
  return world_place(w, &h, xf);
}
uint32_t rt_add_box(rt_world *w, rt_xform xf, rt_vec3 a, rt_vec3 b, rt_material mat) {
  rt_hittable h = make_box(w, a, b, mat);
  return world_place(w, &h, xf);
}
uint32_t rt_add_pyramid(rt_world *w, rt_xform xf, rt_vec3 base_a, rt_vec3 base_b, rt_vec3 tip, rt_material mat) {
  rt_hittable h = make_pyramid(w, base_a, base_b, tip, mat);
  return world_place(w, &h, xf);
}

/* ---------- reference-compatible BVH (rt.cu:711-785, 850-906) ---------- */
static float bbox_min_on_axis(const rt_world *w, rt_href href, int axis) {
  const rt_aabb *b = hittable_bbox(&w->hittables[href]);
  return axis == 0 ? b->x0 : (axis == 1 ? b->y0 : b->z0);
}

/* Top-down merge sort over refs[lo..hi] (inclusive), split at (lo+hi)/2; on a tie the element of the
 * RIGHT run is emitted first (the reference takes left only on strict '<'), so the sort is not stable
 * and the resulting order is part of the scene's identity. */
static void sort_refs(const rt_world *w, rt_href *refs, rt_href *tmp, uint64_t lo, uint64_t hi, int axis) {
  if (hi <= lo) return;
  uint64_t mid = (lo + hi) / 2;
  sort_refs(w, refs, tmp, lo, mid, axis);
  sort_refs(w, refs, tmp, mid + 1, hi, axis);
  uint64_t a = lo, b = mid + 1, o = lo;
  while (a <= mid && b <= hi) {
    if (bbox_min_on_axis(w, refs[a], axis) < bbox_min_on_axis(w, refs[b], axis)) tmp[o++] = refs[a++];
    else tmp[o++] = refs[b++];
  }
  while (a <= mid) tmp[o++] = refs[a++];
  while (b <= hi) tmp[o++] = refs[b++];
  memcpy(refs + lo, tmp + lo, (hi - lo + 1) * sizeof(rt_href));
}

static uint32_t build_bvh_range(rt_world *w, rt_href *tmp, uint64_t start, uint64_t end) {
  rt_hittable node = hittable_zero(RT_OBJ_BVH_NODE);
  uint64_t span = end - start;
  if (span == 1) {
    node.u.bvh.left = node.u.bvh.right = (rt_href)start;
  } else if (span == 2) {
    node.u.bvh.left = (rt_href)start;
    node.u.bvh.right = (rt_href)(start + 1);
  } else {
    int axis = rt_host_random_int(0, 2);            /* drawn before recursing (pre-order) */
    sort_refs(w, w->world_refs, tmp, start, end - 1, axis);
    uint64_t mid = start + span / 2;
    node.u.bvh.left = (rt_href)build_bvh_range(w, tmp, start, mid);
    node.u.bvh.right = (rt_href)build_bvh_range(w, tmp, mid, end);
  }
  const rt_aabb *lb = hittable_bbox(&w->hittables[w->world_refs[node.u.bvh.left]]);
  const rt_aabb *rb = hittable_bbox(&w->hittables[w->world_refs[node.u.bvh.right]]);
  node.u.bvh.bbox = aabb_union(lb, rb);
  return (uint32_t)world_refs_append(w, (rt_href)hittables_append(w, &node));
}

uint32_t rt_world_build_bvh(rt_world *w) {
  rt_href *tmp = (rt_href *)malloc(sizeof(rt_href) * w->world_cap);
  w->root_wref = build_bvh_range(w, tmp, 0, w->world_count);
  w->bvh_built = 1;
  free(tmp);
  return w->root_wref;
}

static int32_t hittable_depth(const rt_world *w, rt_href href) {             /* rt.cu:1614-1637 */
  const rt_hittable *h = &w->hittables[href];
  switch (h->type) {
    case RT_OBJ_SPHERE: case RT_OBJ_TRIANGLE: case RT_OBJ_QUAD: return 1;
    case RT_OBJ_TRANSLATION: return 2 + hittable_depth(w, h->u.translation.child);
    case RT_OBJ_Y_ROTATION: return 2 + hittable_depth(w, h->u.y_rotation.child);
    case RT_OBJ_STRUCTURE: return (int32_t)h->u.structure.size;
    default: return 0;
  }
}
uint64_t rt_world_stack_bound(const rt_world *w) {                           /* rt.cu:1639-1647, 1735-1737 */
  int32_t deepest = 0;
  for (uint64_t i = 0; i < w->world_count; ++i) {
    int32_t d = hittable_depth(w, w->world_refs[i]);
    if (d > deepest) deepest = d;
  }
  uint64_t bound = (uint64_t)(logf((float)w->world_count) / logf(2.0f) + 1);
  return bound + (uint64_t)deepest;
}
