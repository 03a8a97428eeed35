// oracle/ref_host_driver.cu -- TEST INFRASTRUCTURE, not product code.
//
// Thin driver that compiles the UNMODIFIED reference sources (luk-zm/cuda_raytracing,
// mounted read-only at /root/reference) where they lie and exposes, as a C ABI:
//   * the reference's own host scene construction + BVH build (src/raytracing.cu:1775-2320,
//     850-906), so our C scene builder can be diffed field by field against it;
//   * the reference's own __host__ __device__ hot-path functions (hit, aabb_hit,
//     material_scatter_common, texture_get_color, perlin_*), so the plain-C oracle
//     (oracle/rt_oracle.c) and the CUDA kernels can be differential-tested against them;
//   * a host restatement of the ~110 __device__-only glue lines (ray_color rt.cu:1334-1371,
//     the cuda_calc sample loop rt.cu:1560-1587, the XORWOW-driven samplers vec3.h:140-161,
//     rt.cu:1459-1476, 497-502) over cuRAND's XORWOW compiled for the host, with the device's
//     left-to-right draw order.  OpenMP over image rows.
//
// Built ONLY by oracle/build_ref.sh into oracle/_ref/ (git-ignored).  No reference source is
// copied into this repository: the files are #included from REF_SRC at build time.
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <stdint.h>
#include <stdlib.h>
#include <time.h>
#include <assert.h>
#define QUALIFIERS static __forceinline__ __host__ __device__
#include <curand_kernel.h>
#define STB_IMAGE_IMPLEMENTATION
#include "stb_image.h"

#define REF_STR2(x) #x
#define REF_STR(x) REF_STR2(x)
#define REF_FILE(name) REF_STR(REF_SRC/name)

#include REF_FILE(utility.h)
#include REF_FILE(vec3.h)
#include REF_FILE(perlin_noise.h)
#include REF_FILE(memory.h)
#include REF_FILE(platform.h)
#include REF_FILE(perlin_noise.cpp)
#include REF_FILE(platform.cpp)
#include REF_FILE(raytracing.cu)

#ifdef _OPENMP
#include <omp.h>
#endif

#define REF_WORLD_SIZE 8000   // src/main.cu:24

typedef void (*ref_scene_fn)(World *, RenderSettings *);
static ref_scene_fn g_scene_table[20] = {   // src/main.cu:52-112
  fv1, cornell_box_scene, bouncing_spheres_scene, final_scene, fv2, fv3, fv4, fv5,
  refract1, refract2, refract3, refract4, refract5, motion_blur_scene, light_test_scene,
  checkered_spheres_scene, earth_scene, perlin_spheres_scene, quads_scene, simple_light_scene };

extern "C" {

typedef struct {
  void *hittables;            // Hittable[hittable_count], 120 B each
  uint64_t hittable_count;
  uint16_t *world;            // HittableRef[world_count]
  uint64_t world_count;
  uint32_t root_wref;
  uint8_t *texture_arena;
  uint64_t texture_bytes;
  uint64_t texture_count;
  uint64_t max_bvh_depth;     // rt.cu:1735-1737
  // RenderSettings, flattened (rt.cu:1509-1521)
  float img_ratio; int32_t img_width; int32_t samples_per_pixel; int32_t max_bounces; float vfov;
  float lookfrom[3], lookat[3], view_up[3]; float defocus_angle; float background[3];
  char img_name[64];
  // private
  void *priv_world;
} RefScene;

uint64_t ref_sizeof(int what) {
  switch (what) {
    case 0: return sizeof(Hittable); case 1: return sizeof(Texture); case 2: return sizeof(PerlinNoise);
    case 3: return sizeof(Material); case 4: return sizeof(Sphere); case 5: return sizeof(Quad);
    case 6: return sizeof(BvhNode); case 7: return sizeof(Translation); case 8: return sizeof(YRotation);
    case 9: return sizeof(HittableStructure); case 10: return sizeof(HitRecord); case 11: return sizeof(Ray);
    case 12: return sizeof(RenderSettings); case 13: return sizeof(curandState); case 14: return sizeof(Geometry);
  }
  return 0;
}

// Scene function + make_bvh, exactly as main() then render() do (main.cu:35-117, rt.cu:1653).
// glibc rand() is re-seeded to its default state (seed 1) first, which is the state a fresh
// `./raytracing` process starts in (utility.h:26-28 never calls srand).
int ref_scene_build(int scene_id, RefScene *out) {
  if (scene_id < 0 || scene_id >= 20) return -1;
  memset(out, 0, sizeof(*out));
  srand(1);
  World *world = (World *)calloc(1, sizeof(World));
  MemArena texture_arena = {0};
  mem_arena_alloc(&texture_arena, MB(10));
  memset(texture_arena.data, 0, MB(10));
  world->textures = make_texture_list(&texture_arena);
  MemArena ref_arena = {0};
  mem_arena_alloc(&ref_arena, sizeof(HittableRef) * REF_WORLD_SIZE);
  world->bvh_include = make_hittable_ref_list(&ref_arena, REF_WORLD_SIZE);
  world->all_hittables.objects = (Hittable *)calloc(REF_WORLD_SIZE, sizeof(Hittable));
  world->all_hittables.size = REF_WORLD_SIZE;
  RenderSettings settings = g_default_render_settings;
  g_scene_table[scene_id](world, &settings);
  WorldRef root = make_bvh(&world->all_hittables, &world->bvh_include);
  u64 max_bvh_depth = logf(world->bvh_include.count) / logf(2.0f) + 1;
  max_bvh_depth += world_max_depth(world);

  out->hittables = world->all_hittables.objects;
  out->hittable_count = world->all_hittables.count;
  out->world = world->bvh_include.data;
  out->world_count = world->bvh_include.count;
  out->root_wref = root;
  out->texture_arena = world->textures.arena.data;
  out->texture_bytes = world->textures.arena.current_pos;
  out->texture_count = world->textures.count;
  out->max_bvh_depth = max_bvh_depth;
  out->img_ratio = settings.img_ratio; out->img_width = settings.img_width;
  out->samples_per_pixel = settings.samples_per_pixel; out->max_bounces = settings.max_bounces;
  out->vfov = settings.vfov;
  memcpy(out->lookfrom, &settings.lookfrom, 12); memcpy(out->lookat, &settings.lookat, 12);
  memcpy(out->view_up, &settings.view_up, 12);
  out->defocus_angle = settings.defocus_angle; memcpy(out->background, &settings.background_color, 12);
  strncpy(out->img_name, settings.img_name, sizeof(out->img_name) - 1);
  out->priv_world = world;
  return 0;
}

void ref_scene_free(RefScene *s) {
  if (!s || !s->priv_world) return;
  World *world = (World *)s->priv_world;
  free(world->all_hittables.objects);
  free(world->bvh_include.data);
  free(world->textures.arena.data);
  free(world);
  memset(s, 0, sizeof(*s));
}

static void ref_make_scene_data(const RefScene *s, SceneData *sd) {
  memset(sd, 0, sizeof(*sd));
  sd->hittables.objects = (Hittable *)s->hittables;
  sd->hittables.count = s->hittable_count;
  sd->hittables.size = REF_WORLD_SIZE;
  sd->textures.arena.data = s->texture_arena;
  sd->textures.arena.current_pos = s->texture_bytes;
  sd->textures.count = s->texture_count;
  sd->world.data = s->world;
  sd->world.count = s->world_count;
  sd->root_bvh_node = (WorldRef)s->root_wref;
}

// ---- per-function entry points (the reference's own __host__ __device__ code) ----

// rec_out: point(3) normal(3) t front mat_type mat_words(5) u v  = 16 x 4 B, see tests/refabi.py
typedef struct {
  float point[3]; float normal[3]; float t; int32_t front; uint32_t mat[6]; float u, v;
} RefHit;

// Ray is passed in/out: hit() mutates it through the instance wrappers (rt.cu:1176-1237).
int ref_hit(const RefScene *s, float *origin, float *direction, float time, float tmin, float tmax,
            RefHit *rec_out) {
  SceneData sd; ref_make_scene_data(s, &sd);
  HittableRef stack_data[256];
  PerPixelGpuData gd; memset(&gd, 0, sizeof(gd));
  gd.hit_stack.data = stack_data; gd.hit_stack.max_count = 256; gd.hit_stack.count = 0;
  Ray ray; memcpy(&ray.origin, origin, 12); memcpy(&ray.direction, direction, 12);
  ray.intersection_time = time;
  HitRecord rec; memset(&rec, 0, sizeof(rec));
  b32 r = hit(&gd, &sd, wref_to_href(&sd.world, sd.root_bvh_node), tmin, tmax, &ray, &rec);
  memcpy(origin, &ray.origin, 12); memcpy(direction, &ray.direction, 12);
  memcpy(rec_out->point, &rec.point_hit, 12); memcpy(rec_out->normal, &rec.normal, 12);
  rec_out->t = rec.t; rec_out->front = rec.is_front_face;
  memcpy(rec_out->mat, &rec.mat, 24); rec_out->u = rec.u; rec_out->v = rec.v;
  return r;
}

int ref_aabb_hit(const float *box6, const float *origin, const float *direction, float t0, float t1) {
  Aabb b; memcpy(&b, box6, 24);
  Ray ray; memcpy(&ray.origin, origin, 12); memcpy(&ray.direction, direction, 12); ray.intersection_time = 0;
  return aabb_hit(&b, &ray, t0, t1);
}

void ref_texture_color(const RefScene *s, uint64_t tex_off, float u, float v, const float *p, float *out3) {
  SceneData sd; ref_make_scene_data(s, &sd);
  TextureRef tr; tr.idx = tex_off;
  vec3 pt; memcpy(&pt, p, 12);
  vec3 c = texture_get_color(&sd.textures, tr, u, v, pt);
  memcpy(out3, &c, 12);
}

float ref_perlin_noise(const RefScene *s, uint64_t noise_off, const float *p) {
  vec3 pt; memcpy(&pt, p, 12);
  return perlin_noise_val((PerlinNoise *)(s->texture_arena + noise_off), pt);
}

float ref_perlin_turbulence(const RefScene *s, uint64_t noise_off, const float *p, int depth) {
  vec3 pt; memcpy(&pt, p, 12);
  return perlin_turbulence((PerlinNoise *)(s->texture_arena + noise_off), pt, depth);
}

// material_scatter_common (rt.cu:452-490) with caller-supplied random inputs.
int ref_scatter(const RefScene *s, const uint32_t *mat6, const float *dir_in, const RefHit *rec_in,
                const float *unit_vec, float random_test, float *dir_out, float *atten_out) {
  SceneData sd; ref_make_scene_data(s, &sd);
  Material mat; memcpy(&mat, mat6, 24);
  HitRecord rec; memset(&rec, 0, sizeof(rec));
  memcpy(&rec.point_hit, rec_in->point, 12); memcpy(&rec.normal, rec_in->normal, 12);
  rec.t = rec_in->t; rec.is_front_face = rec_in->front; rec.mat = mat; rec.u = rec_in->u; rec.v = rec_in->v;
  vec3 d; memcpy(&d, dir_in, 12); vec3 uv; memcpy(&uv, unit_vec, 12);
  ScatterRes r = material_scatter_common(&sd.textures, &mat, d, &rec, uv, random_test);
  memcpy(dir_out, &r.direction, 12); memcpy(atten_out, &r.attenuation, 12);
  return r.is_hit;
}

void ref_emit(const RefScene *s, const uint32_t *mat6, float u, float v, const float *p, float *out3) {
  SceneData sd; ref_make_scene_data(s, &sd);
  Material mat; memcpy(&mat, mat6, 24);
  vec3 pt; memcpy(&pt, p, 12);
  vec3 c = material_emit(&sd.textures, &mat, u, v, pt);
  memcpy(out3, &c, 12);
}

// ---- cuRAND XORWOW on the host (toolkit header, QUALIFIERS override above) ----
// state6 = d, v[0..4]
void ref_xorwow_init(uint64_t seed, uint64_t subsequence, uint64_t offset, uint32_t *state6) {
  curandState st; curand_init(seed, subsequence, offset, &st);
  state6[0] = st.d; for (int i = 0; i < 5; ++i) state6[1 + i] = st.v[i];
}
void ref_xorwow_uniforms(uint64_t seed, uint64_t subsequence, uint64_t offset, int n, float *out, uint32_t *state6) {
  curandState st; curand_init(seed, subsequence, offset, &st);
  for (int i = 0; i < n; ++i) out[i] = curand_uniform(&st);
  if (state6) { state6[0] = st.d; for (int i = 0; i < 5; ++i) state6[1 + i] = st.v[i]; }
}

// ---- host restatement of the __device__-only glue, over the reference's HD functions ----

static inline f32 h_bound(curandState *rng, f32 mn, f32 mx) {          // utility.h:34-37
  return mn + (mx - mn) * curand_uniform(rng);
}
static inline vec3 h_unit_vector(curandState *rng) {                    // vec3.h:152-161, draws x,y,z L->R
  while (1) {
    f32 x = h_bound(rng, -1.0f, 1.0f);
    f32 y = h_bound(rng, -1.0f, 1.0f);
    f32 z = h_bound(rng, -1.0f, 1.0f);
    vec3 r = Vec3(x, y, z);
    f32 len = vec3_length(r);
    f32 l2 = len * len;
    if (1e-50 < l2 && l2 <= 1.0f) return vec3_scale(1.0f / len, r);
  }
}
static inline vec3 h_unit_disk(curandState *rng) {                      // rt.cu:1459-1469
  while (1) {
    f32 x = h_bound(rng, -1.0f, 1.0f);
    f32 y = h_bound(rng, -1.0f, 1.0f);
    vec3 r = Vec3(x, y, 0.0f);
    f32 len = vec3_length(r);
    if (len * len < 1.0f) return r;
  }
}

typedef struct { uint64_t rays; } RefCounters;

static vec3 h_ray_color(SceneData *sd, PerPixelGpuData *gd, Ray *ray, vec3 background, uint64_t *rays) {  // rt.cu:1334-1371
  i32 max_bounces = gd->rc_stack.max_count;
  vec3 result = {0};
  HitRecord record = {0};
  Ray *bounce = ray;
  while (max_bounces > 0) {
    ++*rays;
    if (hit(gd, sd, wref_to_href(&sd->world, sd->root_bvh_node), 0.001f, inf_f32(), bounce, &record)) {
      vec3 emission = material_emit(&sd->textures, &record.mat, record.u, record.v, record.point_hit);
      vec3 uvec = h_unit_vector(&gd->rng);                 // rt.cu:500-501: unit vector first,
      f32 rtest = curand_uniform(&gd->rng);                // then one uniform, for EVERY material
      ScatterRes sc = material_scatter_common(&sd->textures, &record.mat, bounce->direction, &record, uvec, rtest);
      if (sc.is_hit) {
        bounce->origin = record.point_hit;
        bounce->direction = sc.direction;
        ray_color_stack_push(&gd->rc_stack, sc.attenuation, emission);
        max_bounces--;
      } else { result = emission; break; }
    } else { result = background; break; }
  }
  while (!ray_color_stack_is_empty(&gd->rc_stack)) {
    vec3 att = {0}, em = {0};
    ray_color_stack_pop(&gd->rc_stack, &att, &em);
    result = vec3_add(vec3_comp_scale(att, result), em);
  }
  return result;
}

typedef struct {
  int32_t img_width, img_height, samples_per_pixel, max_bounces;
  float defocus_angle; float background[3];
  float pixel_delta_v[3], pixel_delta_u[3], camera_pos[3], defocus_disk_u[3], defocus_disk_v[3], pixel00[3];
  float pixel_samples_scale;
} RefCamera;

// Camera basis exactly as render() computes it (rt.cu:1657-1692).
void ref_camera(const RefScene *s, int width_override, int spp, int bounces, RefCamera *c) {
  RenderSettings st; memset(&st, 0, sizeof(st));
  st.img_ratio = s->img_ratio; st.img_width = width_override > 0 ? width_override : s->img_width;
  st.samples_per_pixel = spp > 0 ? spp : s->samples_per_pixel;
  st.max_bounces = bounces > 0 ? bounces : s->max_bounces; st.vfov = s->vfov;
  memcpy(&st.lookfrom, s->lookfrom, 12); memcpy(&st.lookat, s->lookat, 12); memcpy(&st.view_up, s->view_up, 12);
  st.defocus_angle = s->defocus_angle; memcpy(&st.background_color, s->background, 12);
  RenderSettings *settings = &st;

  f32 pixel_samples_scale = 1.0f / (f32)settings->samples_per_pixel;
  i32 img_width = settings->img_width;
  i32 img_height = (int)(img_width / settings->img_ratio);
  f32 focus_dist = 10.0f;
  f32 theta = settings->vfov * (pi / 180.0f);
  f32 h = tanf(theta / 2.0f);
  f32 viewport_height = 2.0f * h * focus_dist;
  f32 viewport_ratio = (f32)img_width / (f32)img_height;
  f32 viewport_width = viewport_height * viewport_ratio;
  vec3 camera_pos = settings->lookfrom;
  vec3 cam_w = vec3_to_unit_vec(vec3_sub(settings->lookfrom, settings->lookat));
  vec3 cam_u = vec3_to_unit_vec(vec3_cross_product(settings->view_up, cam_w));
  vec3 cam_v = vec3_cross_product(cam_w, cam_u);
  vec3 viewport_v = vec3_scale(viewport_height, vec3_sign_flip(cam_v));
  vec3 viewport_u = vec3_scale(viewport_width, cam_u);
  vec3 pixel_delta_v = vec3_scale(1.0f / (f32)img_height, viewport_v);
  vec3 pixel_delta_u = vec3_scale(1.0f / (f32)img_width, viewport_u);
  vec3 corner = camera_pos;
  vec3_inplace_sub(&corner, vec3_scale(focus_dist, cam_w));
  vec3_inplace_sub(&corner, vec3_scale(0.5f, viewport_u));
  vec3_inplace_sub(&corner, vec3_scale(0.5f, viewport_v));
  f32 defocus_radius = focus_dist * tan((settings->defocus_angle / 2.0f) * pi / 180.0f);
  vec3 ddu = vec3_scale(defocus_radius, cam_u);
  vec3 ddv = vec3_scale(defocus_radius, cam_v);
  vec3 p00 = corner;
  vec3_inplace_add(&p00, vec3_scale(0.5f, pixel_delta_v));
  vec3_inplace_add(&p00, vec3_scale(0.5f, pixel_delta_u));

  c->img_width = img_width; c->img_height = img_height;
  c->samples_per_pixel = settings->samples_per_pixel; c->max_bounces = settings->max_bounces;
  c->defocus_angle = settings->defocus_angle; memcpy(c->background, &settings->background_color, 12);
  memcpy(c->pixel_delta_v, &pixel_delta_v, 12); memcpy(c->pixel_delta_u, &pixel_delta_u, 12);
  memcpy(c->camera_pos, &camera_pos, 12); memcpy(c->defocus_disk_u, &ddu, 12); memcpy(c->defocus_disk_v, &ddv, 12);
  memcpy(c->pixel00, &p00, 12); c->pixel_samples_scale = pixel_samples_scale;
}

// The cuda_calc body (rt.cu:1560-1587) for every pixel, on the host.  Rows [row0,row1).
// out: img_width*img_height*3 floats (only the requested rows are written).
int ref_render_host(const RefScene *s, const RefCamera *c, uint64_t seed, uint64_t subseq_offset,
                    int row0, int row1, float *out, uint64_t *rays_out, int nthreads) {
  SceneData sd; ref_make_scene_data(s, &sd);
  const i32 W = c->img_width, H = c->img_height;
  if (row0 < 0) row0 = 0; if (row1 > H || row1 <= 0) row1 = H;
  vec3 dv, du, cam, ddu, ddv, p00, bg;
  memcpy(&dv, c->pixel_delta_v, 12); memcpy(&du, c->pixel_delta_u, 12); memcpy(&cam, c->camera_pos, 12);
  memcpy(&ddu, c->defocus_disk_u, 12); memcpy(&ddv, c->defocus_disk_v, 12); memcpy(&p00, c->pixel00, 12);
  memcpy(&bg, c->background, 12);
  uint64_t total_rays = 0;
  const i32 spp = c->samples_per_pixel;
  const i32 maxb = c->max_bounces;
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel reduction(+:total_rays)
  {
    HittableRef *hstack = (HittableRef *)malloc(sizeof(HittableRef) * 512);
    RayColorStackData *cstack = (RayColorStackData *)malloc(sizeof(RayColorStackData) * 256);
    PerPixelGpuData gd; memset(&gd, 0, sizeof(gd));
    gd.hit_stack.data = hstack; gd.hit_stack.max_count = 512; gd.hit_stack.count = 0;
    gd.rc_stack.data = cstack; gd.rc_stack.max_count = (u8)maxb; gd.rc_stack.count = 0;
    uint64_t rays = 0;
#pragma omp for schedule(dynamic, 1)
    for (i32 i = row0; i < row1; ++i) {
      for (i32 j = 0; j < W; ++j) {
        i32 idx = i * W + j;
        curand_init(seed, (uint64_t)idx + subseq_offset, 0, &gd.rng);
        vec3 base = vec3_add(p00, vec3_add(vec3_scale((f32)i, dv), vec3_scale((f32)j, du)));
        vec3 acc = {0};
        for (i32 sidx = 0; sidx < spp; ++sidx) {
          Ray ray = {0};
          if (c->defocus_angle <= 0) ray.origin = cam;
          else {
            vec3 p = h_unit_disk(&gd.rng);
            ray.origin = vec3_add(cam, vec3_add(vec3_scale(p.x, ddu), vec3_scale(p.y, ddv)));
          }
          f32 xo = curand_uniform(&gd.rng) - 0.5f;
          f32 yo = curand_uniform(&gd.rng) - 0.5f;
          ray.direction = vec3_sub(base, ray.origin);
          vec3_inplace_add(&ray.direction, vec3_scale(xo, du));
          vec3_inplace_add(&ray.direction, vec3_scale(yo, dv));
          ray.intersection_time = curand_uniform(&gd.rng);
          vec3_inplace_add(&acc, h_ray_color(&sd, &gd, &ray, bg, &rays));
        }
        vec3_inplace_scale(c->pixel_samples_scale, &acc);
        memcpy(out + 3 * (size_t)idx, &acc, 12);
      }
    }
    total_rays += rays;
    free(hstack); free(cstack);
  }
  if (rays_out) *rays_out = total_rays;
  return 0;
}

// pixels_to_ppm (rt.cu:1398-1436) on a caller-supplied float image.
void ref_write_ppm(const float *pixels, uint32_t w, uint32_t h, const char *name) {
  pixels_to_ppm((vec3 *)pixels, w, h, name);
}

int ref_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

}  // extern "C"
