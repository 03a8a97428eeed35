/* rt_scene.h -- host-side scene description, in C, for the B200 path tracer.
 *
 * These POD records are the INPUT FORMAT of the hot path: they have exactly the in-memory layout
 * of the reference's scene arrays (luk-zm/cuda_raytracing, src/raytracing.cu:1-241; sizes probed in
 * SURVEY.md Appendix C.3), so that the three flat arrays a reference `World` holds after
 * make_bvh() -- Hittable[count] (120 B each), HittableRef[count] (u16) and the texture byte arena --
 * can be handed to rt_scene_upload() (rt_api.h) unchanged, and so that the scenes this library
 * builds can be compared byte-for-byte with the reference's.
 *
 * Everything here is plain C99; no CUDA types.  The scene-construction API below replaces the
 * reference's L2/L4 host layers (rt.cu:504-981 constructors/AABB/BVH build, rt.cu:1775-2320 scenes).
 */
#ifndef RT_SCENE_H
#define RT_SCENE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct { float x, y, z; } rt_vec3;                               /* vec3.h:4-8, 12 B */
typedef struct { float x0, x1, y0, y1, z0, z1; } rt_aabb;                /* rt.cu:61-65, 24 B */
typedef uint16_t rt_href;                                                /* rt.cu:1 HittableRef */

/* ---- textures: 32-byte headers in a byte arena; references are BYTE OFFSETS (rt.cu:263-291) ---- */
enum { RT_TEX_SOLID = 0, RT_TEX_CHECKER = 1, RT_TEX_IMAGE = 2, RT_TEX_NOISE = 3 };   /* rt.cu:39-44 */

typedef struct rt_texture {                                              /* rt.cu:46-54, 32 B */
  uint32_t type;
  uint32_t pad_;
  union {
    struct { rt_vec3 albedo; } solid;
    struct { float inverted_scale; uint32_t pad_; uint64_t even; uint64_t odd; } checker;
    struct { void *pixels; uint64_t pixels_offset; int32_t width; int32_t height; } image;
    struct { void *noise; uint64_t noise_offset; float scale; uint32_t pad_; } noise;
  } u;
} rt_texture;

#define RT_PERLIN_POINTS 256
typedef struct rt_perlin {                                               /* perlin_noise.h:4-10, 6144 B */
  rt_vec3 vectors[RT_PERLIN_POINTS];
  int32_t perm_x[RT_PERLIN_POINTS];
  int32_t perm_y[RT_PERLIN_POINTS];
  int32_t perm_z[RT_PERLIN_POINTS];
} rt_perlin;

/* ---- materials (rt.cu:78-99), 24 B ---- */
enum { RT_MAT_LAMBERTIAN = 0, RT_MAT_METAL = 1, RT_MAT_DIELECTRIC = 2, RT_MAT_DIFFUSE_LIGHT = 3 };

typedef struct rt_material {
  uint32_t type;
  uint32_t pad_;
  union {
    uint64_t texture;                                   /* lambertian / diffuse light: arena offset */
    struct { rt_vec3 albedo; float fuzz; } metal;
    struct { float refraction_index; } dielectric;
  } u;
} rt_material;

/* ---- hittables (rt.cu:107-193): one 120-byte tagged union ---- */
enum {                                                                   /* rt.cu:145-153 */
  RT_OBJ_SPHERE = 0, RT_OBJ_BVH_NODE = 1, RT_OBJ_TRIANGLE = 2, RT_OBJ_QUAD = 3,
  RT_OBJ_TRANSLATION = 4, RT_OBJ_Y_ROTATION = 5, RT_OBJ_STRUCTURE = 6
};

typedef struct { rt_vec3 center; rt_vec3 direction; float radius; uint32_t pad_; rt_material mat; rt_aabb bbox; } rt_sphere;   /* 80 B */
typedef struct { rt_vec3 q, u, v; float d; rt_vec3 w; rt_vec3 normal; } rt_geometry;                                            /* 64 B */
typedef struct { rt_geometry geom; rt_material mat; rt_aabb bbox; } rt_planar;                                                  /* quad or triangle, 112 B */
typedef struct { rt_href left, right; rt_aabb bbox; } rt_bvh_node;                                                              /* 28 B; children are WORLD refs */
typedef struct { rt_href child; uint16_t pad_; rt_vec3 offset; rt_aabb bbox; } rt_translation;                                  /* 40 B; child is a hittable index */
typedef struct { rt_href child; uint16_t pad_; float sin_theta, cos_theta; rt_aabb bbox; } rt_y_rotation;                       /* 36 B */
typedef struct { rt_href first; uint16_t pad_; rt_aabb bbox; uint32_t pad2_; uint64_t size; uint64_t count; } rt_structure;     /* 48 B */

typedef struct rt_hittable {
  uint32_t type;
  uint32_t pad_;
  union {
    rt_sphere sphere;
    rt_bvh_node bvh;
    rt_planar planar;
    rt_translation translation;
    rt_y_rotation y_rotation;
    rt_structure structure;
  } u;
} rt_hittable;

/* ---- render settings: the reference's RenderSettings (rt.cu:1509-1521), same field meaning ---- */
typedef struct rt_render_settings {
  float img_ratio;
  int32_t img_width;
  int32_t samples_per_pixel;
  int32_t max_bounces;
  float vfov;
  rt_vec3 lookfrom, lookat, view_up;
  float defocus_angle;
  rt_vec3 background_color;
  const char *img_name;
} rt_render_settings;

/* ---- world: the reference's `World` (rt.cu:1608-1612) as three growable flat arrays ---- */
#define RT_WORLD_CAPACITY 8000              /* main.cu:24 WORLD_SIZE */
#define RT_TEXTURE_ARENA_BYTES (10u << 20)  /* main.cu:36 */

typedef struct rt_world {
  rt_hittable *hittables;  uint64_t hittable_count, hittable_cap;
  rt_href *world_refs;     uint64_t world_count, world_cap;      /* "bvh_include": objects the BVH is built over, then + nodes */
  uint8_t *texture_arena;  uint64_t texture_bytes, texture_cap, texture_count;
  uint32_t root_wref;      /* valid after rt_world_build_bvh */
  int bvh_built;
} rt_world;

int  rt_world_init(rt_world *w);
void rt_world_free(rt_world *w);

/* instance transform applied when adding an object (rt.cu:1775-1803) */
typedef struct rt_xform { int rotate; float angle_deg; int translate; rt_vec3 offset; } rt_xform;
static const rt_xform RT_XFORM_NONE = {0, 0.0f, 0, {0.0f, 0.0f, 0.0f}};

/* materials (rt.cu:1805-1854) */
rt_material rt_mat_lambert_solid(rt_world *w, rt_vec3 color);
rt_material rt_mat_light_solid(rt_world *w, rt_vec3 color);
rt_material rt_mat_lambert_image(rt_world *w, const char *file_name);   /* file never decoded here: NULL pixels => magenta (rt.cu:346-347) */
/* image texture from DECODED pixels (what stbi_loadf() would have returned: width*height RGB floats, row-major, top row
 * first): the decoder stays with the caller (stb_image is third party), the fetch with the reference's clamp semantics
 * -- x in [0,width], y in [0,height] inclusive, rt.cu:343-354 -- is done on the device.  The reference reads one row
 * and one texel past its allocation at u = 1 or v = 0; here those texels exist and are 0.  NULL / empty / too big for
 * the arena: returns the missing-file record (magenta). */
rt_material rt_mat_lambert_image_pixels(rt_world *w, const float *rgb, int32_t width, int32_t height);
rt_material rt_mat_lambert_noise(rt_world *w, float scale);             /* consumes rand() like make_perlin_noise */
rt_material rt_mat_lambert_checker(rt_world *w, float scale, rt_vec3 even, rt_vec3 odd);
rt_material rt_mat_metal(rt_vec3 albedo, float fuzz);
rt_material rt_mat_dielectric(float refraction_index);

/* primitives (rt.cu:1856-1977); return the world ref of the added object */
uint32_t rt_add_sphere(rt_world *w, rt_xform xf, rt_vec3 center, float radius, rt_material mat);
uint32_t rt_add_moving_sphere(rt_world *w, rt_xform xf, rt_vec3 center, rt_vec3 direction, float radius, rt_material mat);
uint32_t rt_add_quad(rt_world *w, rt_xform xf, rt_vec3 q, rt_vec3 u, rt_vec3 v, rt_material mat);
uint32_t rt_add_triangle(rt_world *w, rt_xform xf, rt_vec3 q, rt_vec3 u, rt_vec3 v, rt_material mat);
uint32_t rt_add_box(rt_world *w, rt_xform xf, rt_vec3 a, rt_vec3 b, rt_material mat);
uint32_t rt_add_pyramid(rt_world *w, rt_xform xf, rt_vec3 base_a, rt_vec3 base_b, rt_vec3 tip, rt_material mat);

/* host RNG used by scene construction and the BVH builder: glibc rand(), utility.h:26-41 */
float   rt_host_random(void);
float   rt_host_random_range(float lo, float hi);
int32_t rt_host_random_int(int32_t lo, int32_t hi);

/* reference-compatible BVH build (rt.cu:850-906): random split axis, merge sort, nodes appended to
 * both arrays; returns the root world ref and stores it in w->root_wref */
uint32_t rt_world_build_bvh(rt_world *w);
/* traversal-stack bound of the reference (rt.cu:1614-1647, 1735-1737) */
uint64_t rt_world_stack_bound(const rt_world *w);

/* the 20 built-in scenes (main.cu:52-112); returns 0, or -1 for an unknown id.  Re-seeds nothing:
 * call from a fresh rand() state (or rt_host_rand_reset()) to get the reference's scenes. */
void rt_default_render_settings(rt_render_settings *s);   /* rt.cu:1527-1529 */
int  rt_scene_by_id(int scene_id, rt_world *w, rt_render_settings *s);
void rt_host_rand_reset(void);                            /* srand(1): the state a fresh process has */

/* camera basis as render() derives it (rt.cu:1657-1692), pure f32 */
typedef struct rt_camera {
  int32_t img_width, img_height;
  rt_vec3 pixel_delta_v, pixel_delta_u, camera_pos, defocus_disk_u, defocus_disk_v, pixel00;
  float pixel_samples_scale;
} rt_camera;
void rt_camera_setup(const rt_render_settings *s, rt_camera *c);

/* gamma + quantise + P6 writer (rt.cu:1391-1436); rgb8 may be NULL (derived from pixels) */
void rt_quantise_rgb8(const float *pixels_rgb, uint64_t n_pixels, uint8_t *rgb8);
int  rt_write_ppm(const char *path, const uint8_t *rgb8, uint32_t width, uint32_t height);

#ifdef __cplusplus
}
#endif
#endif /* RT_SCENE_H */
