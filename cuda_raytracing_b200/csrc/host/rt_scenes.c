/* rt_scenes.c -- the 20 built-in scenes, selected by the same ids as the reference CLI
 * (luk-zm/cuda_raytracing src/main.cu:52-112; scene bodies src/raytracing.cu:1979-2320).
 *
 * The scene CONTENT (positions, colours, camera, file names) is data that must equal the
 * reference's; how it is expressed here is ours: one table of entries, small composable builders,
 * and explicit sequencing of every host-RNG draw (the reference relies on g++ evaluating call
 * arguments right to left).
 */
#include "rt_scene.h"

#include <math.h>
#include <string.h>

rt_vec3 rt_host_random_vec3(void);
rt_vec3 rt_host_random_vec3_range(float lo, float hi);

static inline rt_vec3 P(double x, double y, double z) { rt_vec3 r = {(float)x, (float)y, (float)z}; return r; }
static const rt_vec3 ORIGIN = {0.0f, 0.0f, 0.0f};

static rt_xform at(rt_vec3 offset) { rt_xform x = {0, 0.0f, 1, offset}; return x; }
static rt_xform turned_at(float deg, rt_vec3 offset) { rt_xform x = {1, deg, 1, offset}; return x; }

void rt_default_render_settings(rt_render_settings *s) {                    /* rt.cu:1527-1529 */
  s->img_ratio = 1; s->img_width = 800; s->samples_per_pixel = 100; s->max_bounces = 50; s->vfov = 90;
  s->lookfrom = P(0, 0, 0); s->lookat = P(0, 0, -1); s->view_up = P(0, 1, 0);
  s->defocus_angle = 0.0f; s->background_color = P(0.7f, 0.8f, 1.0f); s->img_name = "image.ppm";
}

/* ---- the "feature view" stage shared by ids 0, 4-12, 14: floor, red box, blue pyramid ---- */
static void stage_floor_box_pyramid(rt_world *w) {
  rt_add_quad(w, RT_XFORM_NONE, P(-5, 0, 1), P(10, 0, 0), P(0, 0, -5), rt_mat_lambert_solid(w, P(0.73, 0.73, 0.73)));
  rt_add_box(w, at(P(-1.5, 0, -1)), ORIGIN, P(0.7, 0.7, 0.7), rt_mat_lambert_solid(w, P(0.5, 0, 0)));
  rt_add_pyramid(w, turned_at(30, P(1, 0, -1)), ORIGIN, P(0.7, 0, 0.7), P(0.35, 1.5, 0.35),
                 rt_mat_lambert_solid(w, P(0, 0, 0.5)));
}
static void stage_camera(rt_render_settings *s, rt_vec3 from, const char *name) {
  s->img_width = 800; s->img_ratio = (float)(16.0 / 9.0); s->samples_per_pixel = 100; s->max_bounces = 40;
  s->lookfrom = from; s->lookat = P(0, 0.5, -1); s->img_name = name;
}

static void scene_fv1(rt_world *w, rt_render_settings *s) {                 /* id 0, rt.cu:2206 */
  rt_add_sphere(w, RT_XFORM_NONE, P(0, 0.5, -1), 0.5f, rt_mat_lambert_solid(w, P(0, 0.5, 0)));
  stage_floor_box_pyramid(w);
  stage_camera(s, P(0, 0.9, 0.75), "fv1.ppm");
}
static void scene_metal(rt_world *w, rt_render_settings *s, float fuzz, const char *name) {   /* ids 4-7, rt.cu:2222 */
  rt_add_sphere(w, RT_XFORM_NONE, P(0, 0.5, -1), 0.5f, rt_mat_metal(P(0.8, 0.8, 0.8), fuzz));
  stage_floor_box_pyramid(w);
  stage_camera(s, P(0, 0.9, 0), name);
}
static void scene_refract(rt_world *w, rt_render_settings *s, float ior, const char *name) {  /* ids 8-12, rt.cu:2254 */
  rt_add_sphere(w, RT_XFORM_NONE, P(-0.4, 0.5, -2), 0.5f, rt_mat_lambert_solid(w, P(0.5, 0.5, 0)));
  rt_add_sphere(w, RT_XFORM_NONE, P(0, 0.5, -1), 0.5f, rt_mat_dielectric(ior));
  stage_floor_box_pyramid(w);
  stage_camera(s, P(0, 0.9, 0), name);
}
static void scene_motion_blur(rt_world *w, rt_render_settings *s) {         /* id 13, rt.cu:2291 */
  rt_add_moving_sphere(w, RT_XFORM_NONE, P(-1, 0.5, -1), P(0, 1, 0), 0.5f, rt_mat_lambert_solid(w, P(0.5, 0.5, 0)));
  rt_add_moving_sphere(w, RT_XFORM_NONE, P(1, 0.5, -1), P(0, 0.5, 0), 0.5f, rt_mat_lambert_solid(w, P(0.5, 0.5, 0.5)));
  rt_add_quad(w, RT_XFORM_NONE, P(-5, 0, 1), P(10, 0, 0), P(0, 0, -5), rt_mat_lambert_solid(w, P(0.73, 0.73, 0.73)));
  stage_camera(s, P(0, 0.9, 0.5), "mblur.ppm");
}
static void scene_light_test(rt_world *w, rt_render_settings *s) {          /* id 14, rt.cu:2305 */
  rt_add_sphere(w, RT_XFORM_NONE, P(0, 0.5, -1), 0.5f, rt_mat_light_solid(w, P(15, 15, 15)));
  stage_floor_box_pyramid(w);
  stage_camera(s, P(0, 0.9, 0.75), "lighttest.ppm");
  s->background_color = P(0, 0, 0);
}

static void scene_cornell(rt_world *w, rt_render_settings *s) {             /* id 1, rt.cu:2113 */
  rt_material red = rt_mat_lambert_solid(w, P(0.65, 0.05, 0.05));
  rt_material white = rt_mat_lambert_solid(w, P(0.73, 0.73, 0.73));
  rt_material green = rt_mat_lambert_solid(w, P(0.12, 0.45, 0.15));
  rt_material light = rt_mat_light_solid(w, P(15, 15, 15));
  rt_add_quad(w, RT_XFORM_NONE, P(555, 0, 0), P(0, 555, 0), P(0, 0, 555), green);
  rt_add_quad(w, RT_XFORM_NONE, P(0, 0, 0), P(0, 555, 0), P(0, 0, 555), red);
  rt_add_quad(w, RT_XFORM_NONE, P(343, 554, 332), P(-130, 0, 0), P(0, 0, -105), light);
  rt_add_quad(w, RT_XFORM_NONE, P(0, 0, 0), P(555, 0, 0), P(0, 0, 555), white);
  rt_add_quad(w, RT_XFORM_NONE, P(555, 555, 555), P(-555, 0, 0), P(0, 0, -555), white);
  rt_add_quad(w, RT_XFORM_NONE, P(0, 0, 555), P(555, 0, 0), P(0, 555, 0), white);
  rt_add_box(w, turned_at(15, P(265, 0, 295)), ORIGIN, P(165, 330, 165), white);
  rt_add_box(w, turned_at(-18, P(130, 0, 100)), ORIGIN, P(165, 165, 165), white);
  s->img_ratio = 1.0f; s->img_width = 600; s->samples_per_pixel = 200; s->max_bounces = 40; s->vfov = 40;
  s->lookfrom = P(278, 278, -800); s->lookat = P(278, 278, 0); s->view_up = P(0, 1, 0);
  s->defocus_angle = 0; s->background_color = P(0, 0, 0); s->img_name = "cornell_box.ppm";
}

static void scene_bouncing_spheres(rt_world *w, rt_render_settings *s) {    /* id 2, rt.cu:1979 */
  (void)rt_mat_lambert_solid(w, P(0.5, 0.5, 0.5));      /* the reference creates (and never uses) this texture */
  rt_add_sphere(w, RT_XFORM_NONE, P(0, -1000, 0), 1000,
                rt_mat_lambert_checker(w, (float)0.32, P(0.2, 0.3, 0.1), P(0.9, 0.9, 0.9)));
  for (int a = -11; a < 11; ++a) {
    for (int b = -11; b < 11; ++b) {
      float choice = rt_host_random();
      rt_vec3 c;                                        /* brace initialiser: x drawn before z */
      c.x = a + 0.9f * rt_host_random(); c.y = 0.2f; c.z = b + 0.9f * rt_host_random();
      rt_vec3 d = {c.x - 4.0f, c.y - 0.2f, c.z - 0.0f};
      if (sqrtf(d.x * d.x + d.y * d.y + d.z * d.z) > 0.9f) {
        if (choice < 0.8f) {
          rt_vec3 second = rt_host_random_vec3();       /* comp_scale(random(), random()): right operand first */
          rt_vec3 first = rt_host_random_vec3();
          rt_vec3 albedo = {first.x * second.x, first.y * second.y, first.z * second.z};
          rt_material m = rt_mat_lambert_solid(w, albedo);
          rt_vec3 drift = {0.0f, rt_host_random_range(0.0f, 0.5f), 0.0f};
          rt_add_moving_sphere(w, RT_XFORM_NONE, c, drift, (float)0.2, m);
        } else {
          rt_material m;
          if (choice < 0.95f) {
            rt_vec3 albedo = rt_host_random_vec3_range(0.5f, 1.0f);
            float fuzz = rt_host_random_range(0.0f, 0.5f);
            m = rt_mat_metal(albedo, fuzz);
          } else {
            m = rt_mat_dielectric(1.5f);
          }
          rt_add_sphere(w, RT_XFORM_NONE, c, (float)0.2, m);
        }
      }
    }
  }
  rt_add_sphere(w, RT_XFORM_NONE, P(0, 1, 0), 1, rt_mat_dielectric(1.5f));
  rt_add_sphere(w, RT_XFORM_NONE, P(-4, 1, 0), 1, rt_mat_lambert_solid(w, P(0.4, 0.2, 0.1)));
  rt_add_sphere(w, RT_XFORM_NONE, P(4, 1, 0), 1, rt_mat_metal(P(0.7f, 0.6f, 0.5f), 0.0f));
  /* rt.cu:1523-1525: this scene replaces the whole settings struct */
  s->img_ratio = 16.0f / 9.0f; s->img_width = 1200; s->samples_per_pixel = 100; s->max_bounces = 50; s->vfov = 20.0f;
  s->lookfrom = P(13, 2, 3); s->lookat = P(0, 0, 0); s->view_up = P(0, 1, 0);
  s->defocus_angle = 0.6f; s->background_color = P(0.7f, 0.8f, 1.0f); s->img_name = "image.ppm";
}

static void scene_final(rt_world *w, rt_render_settings *s) {               /* id 3, rt.cu:2151 */
  rt_material ground = rt_mat_lambert_solid(w, P(0.48f, 0.83f, 0.53f));
  const int per_side = 20;
  for (int i = 0; i < per_side; ++i) {
    for (int j = 0; j < per_side; ++j) {
      float width = 100;
      float x0 = -1000.0f + i * width, z0 = -1000.0f + j * width, y0 = 0;
      float x1 = x0 + width;
      float y1 = rt_host_random_range(1, 101);
      float z1 = z0 + width;
      rt_add_box(w, RT_XFORM_NONE, P(x0, y0, z0), P(x1, y1, z1), ground);
    }
  }
  rt_add_quad(w, RT_XFORM_NONE, P(123, 554, 147), P(300, 0, 0), P(0, 0, 265), rt_mat_light_solid(w, P(7, 7, 7)));
  rt_add_moving_sphere(w, RT_XFORM_NONE, P(400, 400, 200), P(30, 0, 0), 50, rt_mat_lambert_solid(w, P(0.7f, 0.3f, 0.1f)));
  rt_add_sphere(w, RT_XFORM_NONE, P(260, 150, 45), 50, rt_mat_dielectric(1.5f));
  rt_add_sphere(w, RT_XFORM_NONE, P(0, 150, 145), 50, rt_mat_metal(P(0.8f, 0.8f, 0.9f), 1.0f));
  rt_add_sphere(w, RT_XFORM_NONE, P(400, 200, 400), 100, rt_mat_lambert_image(w, "earthmap.jpg"));
  rt_add_sphere(w, RT_XFORM_NONE, P(220, 280, 300), 80, rt_mat_lambert_noise(w, 0.2f));
  rt_material white = rt_mat_lambert_solid(w, P(0.73f, 0.73f, 0.73f));
  for (int i = 0; i < 1000; ++i) {
    rt_vec3 c = rt_host_random_vec3_range(0, 165);
    rt_add_sphere(w, turned_at(15, P(-100, 270, 395)), c, 10, white);
  }
  s->img_ratio = 1.0f; s->img_width = 800; s->samples_per_pixel = 10000; s->max_bounces = 40; s->vfov = 40.0f;
  s->lookfrom = P(478, 278, -600); s->lookat = P(278, 278, 0); s->view_up = P(0.0f, 0.1f, 0.0f);
  s->defocus_angle = 0.0f; s->background_color = ORIGIN; s->img_name = "feature_test.ppm";
}

static void scene_checkered_spheres(rt_world *w, rt_render_settings *s) {   /* id 15, rt.cu:2022 */
  rt_material checker = rt_mat_lambert_checker(w, (float)0.32, P(0.2, 0.3, 0.1), P(0.9, 0.9, 0.9));
  rt_add_sphere(w, RT_XFORM_NONE, P(0, -10, 0), 10, checker);
  rt_add_sphere(w, RT_XFORM_NONE, P(0, 10, 0), 10, checker);
  s->img_ratio = 16.0f / 9.0f; s->img_width = 400; s->samples_per_pixel = 100; s->max_bounces = 50; s->vfov = 20.0f;
  s->lookfrom = P(13, 2, 3); s->lookat = ORIGIN; s->view_up = P(0.0f, 0.1f, 0.0f);
  s->defocus_angle = 0.0f; s->background_color = P(0.7f, 0.8f, 1.0f); s->img_name = "checkered_spheres.ppm";
}
static void scene_earth(rt_world *w, rt_render_settings *s) {               /* id 16, rt.cu:2042 */
  rt_add_sphere(w, RT_XFORM_NONE, ORIGIN, 2, rt_mat_lambert_image(w, "earthmap.jpg"));
  s->img_width = 400; s->samples_per_pixel = 100; s->lookfrom = P(0, 0, 12); s->lookat = ORIGIN;
  s->defocus_angle = 0; s->img_name = "earth.ppm";
}
static void scene_perlin_spheres(rt_world *w, rt_render_settings *s) {      /* id 17, rt.cu:2054 */
  rt_material noise = rt_mat_lambert_noise(w, 4);
  rt_add_sphere(w, RT_XFORM_NONE, P(0, -1000, 0), 1000, noise);
  rt_add_sphere(w, RT_XFORM_NONE, P(0, 2, 0), 2, noise);
  s->img_width = 400; s->samples_per_pixel = 100; s->lookfrom = P(13, 2, 3); s->lookat = ORIGIN;
  s->defocus_angle = 0; s->img_name = "perlin_spheres.ppm";
}
static void scene_quads(rt_world *w, rt_render_settings *s) {               /* id 18, rt.cu:2067 */
  rt_material left = rt_mat_lambert_solid(w, P(1.0f, 0.2f, 0.2f));
  rt_material back = rt_mat_lambert_solid(w, P(0.2f, 1.0f, 0.2f));
  rt_material right = rt_mat_lambert_solid(w, P(0.2f, 0.2f, 1.0f));
  rt_material upper = rt_mat_lambert_solid(w, P(1.0f, 0.5f, 0.0f));
  rt_material lower = rt_mat_lambert_solid(w, P(0.2f, 0.8f, 0.8f));
  rt_add_quad(w, RT_XFORM_NONE, P(-3, -2, 5), P(0, 0, -4), P(0, 4, 0), left);
  rt_add_quad(w, RT_XFORM_NONE, P(-2, -2, 0), P(4, 0, 0), P(0, 4, 0), back);
  rt_add_quad(w, RT_XFORM_NONE, P(3, -2, 1), P(0, 0, 4), P(0, 4, 0), right);
  rt_add_quad(w, RT_XFORM_NONE, P(-2, 3, 1), P(4, 0, 0), P(0, 0, 4), upper);
  rt_add_quad(w, RT_XFORM_NONE, P(-2, -3, 5), P(4, 0, 0), P(0, 0, -4), lower);
  s->img_ratio = 1.0f; s->img_width = 400; s->samples_per_pixel = 100; s->max_bounces = 50; s->vfov = 80;
  s->lookfrom = P(0, 0, 9); s->lookat = ORIGIN; s->view_up = P(0, 1, 0);
  s->defocus_angle = 0; s->background_color = P(0.7f, 0.8f, 1.0f); s->img_name = "quads.ppm";
}
static void scene_simple_light(rt_world *w, rt_render_settings *s) {        /* id 19, rt.cu:2093 */
  rt_material noise = rt_mat_lambert_noise(w, 4);
  rt_add_sphere(w, RT_XFORM_NONE, P(0, -1000, 0), 1000, noise);
  rt_add_sphere(w, RT_XFORM_NONE, P(0, 2, 0), 2, noise);
  rt_material light = rt_mat_light_solid(w, P(4, 4, 4));
  rt_add_quad(w, RT_XFORM_NONE, P(3, 1, -2), P(2, 0, 0), P(0, 2, 0), light);
  rt_add_sphere(w, RT_XFORM_NONE, P(0, 7, 0), 2, light);
  s->img_ratio = 16.0f / 9.0f; s->img_width = 1080; s->samples_per_pixel = 1000; s->max_bounces = 50; s->vfov = 20;
  s->lookfrom = P(26, 3, 6); s->lookat = P(0, 2, 0); s->view_up = P(0, 1, 0); s->img_name = "simple_light.ppm";
  /* background stays the inherited sky colour */
}

int rt_scene_by_id(int id, rt_world *w, rt_render_settings *s) {
  static const float fuzz_by_id[4] = {0.0f, (float)0.2, (float)0.5, 1.0f};
  static const char *fv_names[4] = {"fv2.ppm", "fv3.ppm", "fv4.ppm", "fv5.ppm"};
  static const float ior_by_id[5] = {1.0f, (float)1.05, (float)1.333, (float)1.5, (float)2.417};
  static const char *refract_names[5] = {"refract1.ppm", "refract2.ppm", "refract3.ppm", "refract4.ppm", "refract5.ppm"};
  switch (id) {
    case 0: scene_fv1(w, s); return 0;
    case 1: scene_cornell(w, s); return 0;
    case 2: scene_bouncing_spheres(w, s); return 0;
    case 3: scene_final(w, s); return 0;
    case 4: case 5: case 6: case 7: scene_metal(w, s, fuzz_by_id[id - 4], fv_names[id - 4]); return 0;
    case 8: case 9: case 10: case 11: case 12: scene_refract(w, s, ior_by_id[id - 8], refract_names[id - 8]); return 0;
    case 13: scene_motion_blur(w, s); return 0;
    case 14: scene_light_test(w, s); return 0;
    case 15: scene_checkered_spheres(w, s); return 0;
    case 16: scene_earth(w, s); return 0;
    case 17: scene_perlin_spheres(w, s); return 0;
    case 18: scene_quads(w, s); return 0;
    case 19: scene_simple_light(w, s); return 0;
    default: return -1;
  }
}

/* ---- camera basis, operation for operation as render() (rt.cu:1657-1692) ---- */
static inline rt_vec3 c_sub(rt_vec3 a, rt_vec3 b) { rt_vec3 r = {a.x - b.x, a.y - b.y, a.z - b.z}; return r; }
static inline rt_vec3 c_add(rt_vec3 a, rt_vec3 b) { rt_vec3 r = {a.x + b.x, a.y + b.y, a.z + b.z}; return r; }
static inline rt_vec3 c_scale(float k, rt_vec3 a) { rt_vec3 r = {a.x * k, a.y * k, a.z * k}; return r; }
static inline rt_vec3 c_cross(rt_vec3 a, rt_vec3 b) {
  rt_vec3 r = {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; return r;
}
static inline rt_vec3 c_unit(rt_vec3 a) {
  float len = sqrtf(a.x * a.x + a.y * a.y + a.z * a.z);
  if (len == 0.0f) { rt_vec3 z = {0.0f, 0.0f, 0.0f}; return z; }
  rt_vec3 r = {a.x / len, a.y / len, a.z / len}; return r;
}

void rt_camera_setup(const rt_render_settings *s, rt_camera *c) {
  const float pi = 3.1415926535897932385f;
  c->pixel_samples_scale = 1.0f / (float)s->samples_per_pixel;
  int32_t width = s->img_width;
  int32_t height = (int)(width / s->img_ratio);
  const float focus_dist = 10.0f;                       /* hard-coded in the reference (rt.cu:1663) */
  float theta = s->vfov * (pi / 180.0f);
  float h = tanf(theta / 2.0f);
  float viewport_height = 2.0f * h * focus_dist;
  float viewport_ratio = (float)width / (float)height;
  float viewport_width = viewport_height * viewport_ratio;
  rt_vec3 cam_w = c_unit(c_sub(s->lookfrom, s->lookat));
  rt_vec3 cam_u = c_unit(c_cross(s->view_up, cam_w));
  rt_vec3 cam_v = c_cross(cam_w, cam_u);
  rt_vec3 neg_v = {-cam_v.x, -cam_v.y, -cam_v.z};
  rt_vec3 viewport_v = c_scale(viewport_height, neg_v);
  rt_vec3 viewport_u = c_scale(viewport_width, cam_u);
  c->pixel_delta_v = c_scale(1.0f / (float)height, viewport_v);
  c->pixel_delta_u = c_scale(1.0f / (float)width, viewport_u);
  rt_vec3 corner = s->lookfrom;
  corner = c_sub(corner, c_scale(focus_dist, cam_w));
  corner = c_sub(corner, c_scale(0.5f, viewport_u));
  corner = c_sub(corner, c_scale(0.5f, viewport_v));
  float defocus_radius = focus_dist * tanf((s->defocus_angle / 2.0f) * pi / 180.0f);
  c->defocus_disk_u = c_scale(defocus_radius, cam_u);
  c->defocus_disk_v = c_scale(defocus_radius, cam_v);
  rt_vec3 p00 = corner;
  p00 = c_add(p00, c_scale(0.5f, c->pixel_delta_v));
  p00 = c_add(p00, c_scale(0.5f, c->pixel_delta_u));
  c->pixel00 = p00;
  c->camera_pos = s->lookfrom;
  c->img_width = width; c->img_height = height;
}

/* ---- resolve on the host (rt.cu:1391-1436): used by the CLI only when no device buffer exists ---- */
void rt_quantise_rgb8(const float *px, uint64_t n, uint8_t *rgb8) {
  for (uint64_t i = 0; i < 3 * n; ++i) {
    float c = px[i];
    float g = (c > 0) ? sqrtf(c) : 0.0f;
    if (g < 0.0f) g = 0.0f;
    if (g > 0.999f) g = 0.999f;
    rgb8[i] = (uint8_t)(256 * g);
  }
}
