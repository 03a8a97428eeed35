/* main.c -- `raytracing NRSCENY [bN] [sN]`: the reference's CLI surface (src/main.cu:29-131, README.md:30-45).
 *
 * Same grammar: argv[1] = scene id via atoi (default 0, non-numeric => 0), argv[2] and argv[3] may each be
 * 'b<N>' (bounce depth) or 's<N>' (samples per pixel), applied after the scene function; unknown ids print
 * "Unknown scene" and exit 1.  Same stdout lines in the same order (rt.cu:1759-1768), same output file names.
 * Extensions the reference ignores (it never looks past argv[3] nor at other letters, main.cu:118-129): 'w<N>' image width,
 * 't<N>' traversal (0 auto-exact, 1 reference order, 2 fast), 'g<N>' render on GPUs 0..N-1 of this box (rt_group: one
 * NCCL reduce onto GPU 0), 'd<N>' how the render is dealt to them (0 tiles: interleaved row bands, the image stays
 * byte-identical to one GPU's; 1 samples: every GPU renders spp/N samples of every pixel, best balance, statistically
 * equal image), 'q' JSON stats line on stderr.
 */
#include <stdio.h>
#include <unistd.h>
#include <stdlib.h>
#include <string.h>

#include "rt_api.h"

int main(int argc, char *argv[]) {
  int scene_select = 0;
  if (argc > 1) scene_select = atoi(argv[1]);

  rt_world world;
  if (rt_world_init(&world) != 0) { fprintf(stderr, "out of memory\n"); return 1; }
  rt_render_settings settings;
  rt_default_render_settings(&settings);
  if (rt_scene_by_id(scene_select, &world, &settings) != 0) {
    printf("Unknown scene\n");
    return 1;
  }
  uint32_t traversal = RT_TRAVERSAL_AUTO, shard = RT_SHARD_TILES;
  int json = 0, gpus = 1;
  for (int i = 2; i < argc; ++i) {
    const char *a = argv[i];
    if (i <= 3 && a[0] == 'b') settings.max_bounces = atoi(a + 1);
    else if (i <= 3 && a[0] == 's') settings.samples_per_pixel = atoi(a + 1);
    else if (a[0] == 'w') { int w = atoi(a + 1); if (w > 0) settings.img_width = w; }
    else if (a[0] == 't') traversal = (uint32_t)atoi(a + 1);
    else if (a[0] == 'g') { int g = atoi(a + 1); if (g > 0) gpus = g; }
    else if (a[0] == 'd') shard = atoi(a + 1) == 1 ? RT_SHARD_SAMPLES : RT_SHARD_TILES;
    else if (a[0] == 'q') json = 1;
  }

  rt_context *ctx = NULL;
  rt_group *group = NULL;
  /* stdout is the reference's four lines (scripts scrape them).  NCCL prints its version line to stdout at NCCL_DEBUG=VERSION
   * whatever NCCL_DEBUG_FILE says, so while the group sets up its communicators fd 1 points at stderr. */
  int rc;
  if (gpus > 1) {
    fflush(stdout);
    const int saved = dup(1);
    if (saved >= 0) dup2(2, 1);
    rc = rt_group_create(NULL, gpus, &group);
    fflush(stdout);
    if (saved >= 0) { dup2(saved, 1); close(saved); }
  } else {
    rc = rt_context_create(0, &ctx);
  }
  if (rc != RT_OK) { fprintf(stderr, "raytracing: %s\n", rt_last_error_string(NULL)); return 2; }
  rt_camera cam;
  rt_camera_setup(&settings, &cam);
  size_t n = (size_t)cam.img_width * cam.img_height;
  uint8_t *rgb8 = (uint8_t *)malloc(n * 3);
  rt_stats st;
  memset(&st, 0, sizeof(st));
  if (group) rc = rt_group_render_world(group, &world, &settings, traversal, shard, NULL, rgb8, &st);
  else rc = rt_render_world(ctx, &world, &settings, traversal, NULL, rgb8, &st);
  const char *err = rc == RT_OK ? "no error" : (group ? rt_group_last_error(group) : rt_last_error_string(ctx));
  /* the reference prints these four lines whatever happened (rt.cu:1758-1768) */
  printf("cuda_calc launch error: %s\n", err);
  printf("cuda_calc sync error: %s\n", err);
  printf("Rendering time: %fms\n", st.render_ms);
  printf("Rendering time: %fs\n", st.render_ms / 1e3);
  if (rc != RT_OK) { rt_context_destroy(ctx); rt_group_destroy(group); return 2; }
  rc = rt_write_ppm(settings.img_name, rgb8, (uint32_t)cam.img_width, (uint32_t)cam.img_height);
  if (json) {
    fprintf(stderr,
            "{\"scene\": %d, \"gpus\": %d, \"shard\": \"%s\", \"width\": %d, \"height\": %d, \"spp\": %d, \"bounces\": %d, \"render_ms\": %.3f, "
            "\"rng_init_ms\": %.3f, \"resolve_ms\": %.3f, \"total_ms\": %.3f, \"paths\": %llu, \"rays\": %llu, "
            "\"paths_per_s\": %.1f, \"traversal\": %u, \"grid\": %u, \"block\": %u, \"regs\": %u}\n",
            scene_select, gpus, gpus > 1 ? (shard == RT_SHARD_SAMPLES ? "samples" : "tiles") : "none", cam.img_width, cam.img_height,
            settings.samples_per_pixel, settings.max_bounces, st.render_ms,
            st.rng_init_ms, st.resolve_ms, st.total_ms, (unsigned long long)st.paths, (unsigned long long)st.rays,
            st.render_ms > 0 ? st.paths / (st.render_ms / 1e3) : 0.0, st.traversal_used, st.grid_blocks, st.block_threads,
            st.regs_per_thread);
  }
  free(rgb8);
  rt_context_destroy(ctx);
  rt_group_destroy(group);
  rt_world_free(&world);
  return rc == RT_OK ? 0 : 2;
}
