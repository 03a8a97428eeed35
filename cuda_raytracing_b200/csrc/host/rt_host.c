/* rt_host.c -- the host half of the drop-in boundary, in C: what the reference's render() does around its two
 * kernel launches (rt.cu:1649-1773), expressed over the C ABI of rt_api.h.  No CUDA calls here. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rt_api.h"

void rt_render_params_from_camera(const rt_render_settings *s, const rt_camera *c, rt_render_params *p) {
  memset(p, 0, sizeof(*p));
  p->img_width = c->img_width; p->img_height = c->img_height;
  p->samples_per_pixel = s->samples_per_pixel; p->max_bounces = s->max_bounces;
  p->defocus_angle = s->defocus_angle;
  memcpy(p->background, &s->background_color, 12);
  memcpy(p->pixel_delta_v, &c->pixel_delta_v, 12); memcpy(p->pixel_delta_u, &c->pixel_delta_u, 12);
  memcpy(p->camera_pos, &c->camera_pos, 12);
  memcpy(p->defocus_disk_u, &c->defocus_disk_u, 12); memcpy(p->defocus_disk_v, &c->defocus_disk_v, 12);
  memcpy(p->pixel00, &c->pixel00, 12);
  p->seed = 1234;                      /* rt.cu:1747 */
  p->subsequence_offset = 0;
}

/* render(World*, RenderSettings*) (rt.cu:1649-1773) minus the PPM write: BVH build (once), camera basis,
 * scene upload, RNG init + render + resolve, download. */
int rt_render_world(rt_context *ctx, rt_world *world, const rt_render_settings *settings, uint32_t traversal,
                    float *h_pixels_rgb, uint8_t *h_rgb8, rt_stats *stats) {
  if (!ctx || !world || !settings) return RT_ERR_INVALID;
  if (!world->bvh_built) rt_world_build_bvh(world);            /* rt.cu:1653 */
  rt_camera cam;
  rt_camera_setup(settings, &cam);
  rt_scene_desc desc;
  desc.hittables = world->hittables; desc.hittable_count = world->hittable_count;
  desc.world_refs = world->world_refs; desc.world_count = world->world_count;
  desc.root_wref = world->root_wref;
  desc.texture_arena = world->texture_arena; desc.texture_bytes = world->texture_bytes;
  int rc = rt_scene_upload(ctx, &desc, traversal);
  if (rc != RT_OK) return rc;
  rt_render_params p;
  rt_render_params_from_camera(settings, &cam, &p);
  /* like render(), which runs init_gpu_data on every call (rt.cu:1739-1748): the XORWOW states are recomputed, not reused */
  p.flags = RT_RENDER_COUNT_RAYS | RT_RENDER_NO_RNG_CACHE;
  return rt_render(ctx, &p, h_pixels_rgb, h_rgb8, stats);
}

/* P6 writer with the reference's header and byte order (rt.cu:1398-1436), one buffered write */
int rt_write_ppm(const char *path, const uint8_t *rgb8, uint32_t width, uint32_t height) {
  FILE *f = fopen(path, "wb");
  if (!f) { fprintf(stderr, "Couldn't open the file for writing\n"); return RT_ERR_IO; }
  fprintf(f, "P6\n%u %u\n255\n", width, height);
  size_t n = (size_t)width * height * 3;
  size_t w = fwrite(rgb8, 1, n, f);
  fclose(f);
  return w == n ? RT_OK : RT_ERR_IO;
}
