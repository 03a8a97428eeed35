/* oracle/rt_oracle.h -- TEST INFRASTRUCTURE (CPU oracle), not product code.
 *
 * Plain-C restatement of the per-pixel path-tracing hot path of luk-zm/cuda_raytracing
 * (src/raytracing.cu, src/perlin_noise.cpp, src/vec3.h, src/utility.h) plus cuRAND's XORWOW
 * generator (CUDA Toolkit 12.9 curand_kernel.h -- third party, absent from /root/reference;
 * restated from its published algorithm, see rt_oracle.c).  It consumes the reference's own
 * in-memory scene format (Hittable[120 B], u16 world refs, texture byte arena).
 *
 * PINNING: tests/test_oracle_vs_reference.py checks this oracle BIT FOR BIT against the real
 * reference compiled for the host (oracle/_ref/libref_host.so, built by oracle/build_ref.sh from
 * /root/reference) -- every exported function on randomised inputs, whole images of all 20 scenes --
 * and tests/golden/ holds vectors generated from that reference build (tests/golden/make_golden.py)
 * so the pin also holds where /root/reference is absent (the GPU box).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (libraytracing_b200.so) never does.
 *
 * The function signatures deliberately equal those of oracle/ref_host_driver.cu (ref_* -> orc_*), so
 * a test can swap one for the other.
 */
#ifndef RT_ORACLE_H
#define RT_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  const void *hittables;      /* reference `Hittable[hittable_count]`, 120 B each (rt.cu:182-193) */
  uint64_t hittable_count;
  const uint16_t *world;      /* HittableRef world[world_count] (rt.cu:1, 230-235) */
  uint64_t world_count;
  uint32_t root_wref;         /* SceneData.root_bvh_node */
  const uint8_t *texture_arena;
  uint64_t texture_bytes;
} orc_scene;

typedef struct {              /* same layout as RefCamera in ref_host_driver.cu */
  int32_t img_width, img_height, samples_per_pixel, max_bounces;
  float defocus_angle; float background[3];
  float pixel_delta_v[3], pixel_delta_u[3], camera_pos[3], defocus_disk_u[3], defocus_disk_v[3], pixel00[3];
  float pixel_samples_scale;
} orc_camera;

typedef struct {              /* same layout as RefHit */
  float point[3]; float normal[3]; float t; int32_t front; uint32_t mat[6]; float u, v;
} orc_hit_record;

/* 0 (default): every f32 operation rounded separately -- the semantics of the reference's host build
 *    (g++ -O3, x86-64, no FMA); this is the profile pinned bit-for-bit against libref_host.so.
 * 1: multiply-adds fused exactly where nvcc 12.9 / ptxas fused them in the reference's sm_100 cuda_calc
 *    (read from its SASS with tools/sass_expr.py; DESIGN.md "FMA profile") -- the arithmetic of the
 *    reference's GPU binary, up to CUDA-libm vs glibc differences in powf/sinf. */
void orc_set_fma_profile(int gpu_profile);
int  orc_get_fma_profile(void);

/* cuRAND XORWOW: state6 = {d, v[0..4]} */
void orc_xorwow_init(uint64_t seed, uint64_t subsequence, uint64_t offset, uint32_t *state6);
void orc_xorwow_uniforms(uint64_t seed, uint64_t subsequence, uint64_t offset, int n, float *out, uint32_t *state6);

int   orc_aabb_hit(const float *box6, const float *origin, const float *direction, float t0, float t1);
int   orc_hit(const orc_scene *s, float *origin, float *direction, float time, float tmin, float tmax, orc_hit_record *rec);
void  orc_texture_color(const orc_scene *s, uint64_t tex_off, float u, float v, const float *p, float *out3);
float orc_perlin_noise(const orc_scene *s, uint64_t noise_off, const float *p);
float orc_perlin_turbulence(const orc_scene *s, uint64_t noise_off, const float *p, int depth);
int   orc_scatter(const orc_scene *s, const uint32_t *mat6, const float *dir_in, const orc_hit_record *rec,
                  const float *unit_vec, float random_test, float *dir_out, float *atten_out);
void  orc_emit(const orc_scene *s, const uint32_t *mat6, float u, float v, const float *p, float *out3);

/* whole-image render of rows [row0,row1) (row1<=0: all), OpenMP over rows; returns 0 */
int orc_render(const orc_scene *s, const orc_camera *c, uint64_t seed, uint64_t subseq_offset,
               int row0, int row1, float *out_rgb, uint64_t *rays_out, int nthreads);
/* gamma + clamp + 8-bit quantise (rt.cu:1391-1433) */
void orc_resolve_rgb8(const float *pixels_rgb, uint64_t n_pixels, uint8_t *rgb8);
int  orc_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
