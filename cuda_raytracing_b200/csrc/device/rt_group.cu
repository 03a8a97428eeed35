// rt_group.cu -- several GPUs of one box behind the C ABI (include/rt_api.h, rt_group_*): one context per device, driven by
// one host thread (every call below only enqueues), the shards summed onto device 0 with ONE ncclReduce over NVLink /
// NVSwitch, resolve and download on device 0.  The reference has no counterpart: it renders on the default device only
// (rt.cu:1722-1723, 1754).  SURVEY.md 8(b) / 8(e).
//
//   RT_SHARD_TILES    device g renders interleaved bands of rows with the REFERENCE stream (subsequence = pixel); the bands
//                     are disjoint, so the reduce over zero-initialised images assembles the picture exactly (x + 0 == x):
//                     byte-identical to the one-GPU image.
//   RT_SHARD_SAMPLES  device g renders its share of the samples of EVERY pixel (subsequence = pixel + g * n_pixels; device 0
//                     carries the reference's stream) into an f32 SUM image; the reduce adds them, the resolve scales by
//                     1 / spp.  Perfect load balance; statistical (tier-2) parity for more than one device.
//
// NCCL is bound at run time (dlopen of libnccl.so.2: the system library, or the one a host process such as PyTorch already
// loaded), so the library itself has no link-time dependency on it and a single-GPU group needs no NCCL at all.
#include <dlfcn.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "rt_context.h"

namespace {

// the slice of NCCL's C API used here (nccl.h; stable since NCCL 2.x)
typedef struct ncclComm *ncclComm_t;
typedef int ncclResult_t;
constexpr int kNcclFloat32 = 7, kNcclSum = 0;
struct NcclApi {
  void *handle = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*Reduce)(const void *, void *, size_t, int, int, int, ncclComm_t, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  bool load(std::string &err) {
    if (handle) return true;
    // NCCL prints its version / debug lines to stdout; the CLI's stdout is the reference's four lines (scripts scrape them):
    // send NCCL's output to stderr unless the user chose a file
    setenv("NCCL_DEBUG_FILE", "/dev/stderr", 0);
    handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!handle) handle = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!handle) { err = std::string("NCCL not found (dlopen libnccl.so.2): ") + dlerror(); return false; }
    CommInitAll = (decltype(CommInitAll))dlsym(handle, "ncclCommInitAll");
    CommDestroy = (decltype(CommDestroy))dlsym(handle, "ncclCommDestroy");
    GroupStart = (decltype(GroupStart))dlsym(handle, "ncclGroupStart");
    GroupEnd = (decltype(GroupEnd))dlsym(handle, "ncclGroupEnd");
    Reduce = (decltype(Reduce))dlsym(handle, "ncclReduce");
    GetErrorString = (decltype(GetErrorString))dlsym(handle, "ncclGetErrorString");
    if (!CommInitAll || !CommDestroy || !GroupStart || !GroupEnd || !Reduce || !GetErrorString) { err = "libnccl.so.2 lacks an expected symbol"; return false; }
    return true;
  }
};
NcclApi g_nccl;

}  // namespace

struct rt_group {
  std::vector<rt_context *> ctx;
  std::vector<int> devices;
  std::vector<ncclComm_t> comms;
  std::vector<rtc::DeviceBuffer> pixels;          // one f32 image per device
  std::vector<cudaEvent_t> ev0, ev1;              // around each device's render
  cudaEvent_t ev_all0 = nullptr, ev_all1 = nullptr;
  rtc::DeviceBuffer rgb8;                         // device 0
  std::string error;
};

static int group_fail(rt_group *g, int code, const std::string &msg) {
  if (g) g->error = msg;
  return rtc::fail(nullptr, code, msg);
}
#define RTG_CUDA(g, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return group_fail(g, RT_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)
#define RTG_NCCL(g, call) do { ncclResult_t r_ = (call); if (r_ != 0) return group_fail(g, RT_ERR_CUDA, std::string(#call) + ": " + g_nccl.GetErrorString(r_)); } while (0)

extern "C" {

const char *rt_group_last_error(const rt_group *g) { return g ? g->error.c_str() : rt_last_error_string(nullptr); }
int rt_group_size(const rt_group *g) { return g ? (int)g->ctx.size() : 0; }
rt_context *rt_group_context(rt_group *g, int index) { return (g && index >= 0 && index < (int)g->ctx.size()) ? g->ctx[index] : nullptr; }

void rt_group_destroy(rt_group *g) {
  if (!g) return;
  for (size_t i = 0; i < g->ctx.size(); ++i) {
    cudaSetDevice(g->devices[i]);
    if (g->ctx[i]) cudaStreamSynchronize(g->ctx[i]->stream);
    if (i < g->comms.size() && g->comms[i]) g_nccl.CommDestroy(g->comms[i]);
    if (i < g->pixels.size()) g->pixels[i].release();
    if (i < g->ev0.size() && g->ev0[i]) cudaEventDestroy(g->ev0[i]);
    if (i < g->ev1.size() && g->ev1[i]) cudaEventDestroy(g->ev1[i]);
    if (i == 0) {
      g->rgb8.release();
      if (g->ev_all0) cudaEventDestroy(g->ev_all0);
      if (g->ev_all1) cudaEventDestroy(g->ev_all1);
    }
    rt_context_destroy(g->ctx[i]);
  }
  delete g;
}

int rt_group_create(const int *device_ids, int n_devices, rt_group **out) {
  if (!out) return group_fail(nullptr, RT_ERR_INVALID, "rt_group_create: out is NULL");
  *out = nullptr;
  if (n_devices < 1 || n_devices > 64) return group_fail(nullptr, RT_ERR_INVALID, "rt_group_create: n_devices must be 1..64");
  rt_group *g = new rt_group();
  for (int i = 0; i < n_devices; ++i) {
    const int dev = device_ids ? device_ids[i] : i;
    for (int k = 0; k < i; ++k)
      if (g->devices[k] == dev) { rt_group_destroy(g); return group_fail(nullptr, RT_ERR_INVALID, "rt_group_create: a device is listed twice"); }
    rt_context *c = nullptr;
    const int rc = rt_context_create(dev, &c);
    if (rc != RT_OK) { rt_group_destroy(g); return rc; }
    g->ctx.push_back(c); g->devices.push_back(dev);
  }
  g->pixels.resize(n_devices); g->ev0.assign(n_devices, nullptr); g->ev1.assign(n_devices, nullptr);
  for (int i = 0; i < n_devices; ++i) {
    cudaSetDevice(g->devices[i]);
    if (cudaEventCreate(&g->ev0[i]) != cudaSuccess || cudaEventCreate(&g->ev1[i]) != cudaSuccess) {
      rt_group_destroy(g); return group_fail(nullptr, RT_ERR_CUDA, "rt_group_create: cudaEventCreate failed");
    }
  }
  cudaSetDevice(g->devices[0]);
  cudaEventCreate(&g->ev_all0); cudaEventCreate(&g->ev_all1);
  if (n_devices > 1) {
    std::string err;
    if (!g_nccl.load(err)) { rt_group_destroy(g); return group_fail(nullptr, RT_ERR_UNSUPPORTED, "rt_group_create: " + err); }
    g->comms.assign(n_devices, nullptr);
    const ncclResult_t r = g_nccl.CommInitAll(g->comms.data(), n_devices, g->devices.data());      // single process, all devices
    if (r != 0) {
      const std::string msg = std::string("rt_group_create: ncclCommInitAll: ") + g_nccl.GetErrorString(r);
      g->comms.clear(); rt_group_destroy(g);
      return group_fail(nullptr, RT_ERR_CUDA, msg);
    }
  }
  if (n_devices > 1) {
    // NCCL sets up its channels on the first collective (~0.2 s on 8 GPUs): pay that here, with the rest of the set-up,
    // not inside the first render
    bool ok = true;
    for (int i = 0; i < n_devices && ok; ++i) { cudaSetDevice(g->devices[i]); ok = g->pixels[i].reserve(256) == cudaSuccess; }
    if (ok) {
      ok = g_nccl.GroupStart() == 0;
      for (int i = 0; i < n_devices && ok; ++i)
        ok = g_nccl.Reduce(g->pixels[i].ptr, g->pixels[0].ptr, 16, kNcclFloat32, kNcclSum, 0, g->comms[i], g->ctx[i]->stream) == 0;
      ok = (g_nccl.GroupEnd() == 0) && ok;
    }
    for (int i = 0; i < n_devices; ++i) { cudaSetDevice(g->devices[i]); if (cudaStreamSynchronize(g->ctx[i]->stream) != cudaSuccess) ok = false; }
    if (!ok) { rt_group_destroy(g); return group_fail(nullptr, RT_ERR_CUDA, "rt_group_create: NCCL warm-up reduce failed"); }
  }
  *out = g;
  return RT_OK;
}

// interleaved bands of the tile sharding: ~32 bands per device balance sky against interior
static uint32_t group_band_rows(int n, int height) { const int b = height / (n * 32); return (uint32_t)(b < 4 ? 4 : b); }

int rt_group_render_world(rt_group *g, rt_world *world, const rt_render_settings *settings, uint32_t traversal, uint32_t shard,
                          float *h_pixels_rgb, uint8_t *h_rgb8, rt_stats *stats) {
  if (!g || !world || !settings) return group_fail(g, RT_ERR_INVALID, "rt_group_render_world: NULL argument");
  if (shard != RT_SHARD_TILES && shard != RT_SHARD_SAMPLES) return group_fail(g, RT_ERR_INVALID, "rt_group_render_world: unknown sharding");
  const int n = (int)g->ctx.size();
  if (!world->bvh_built) rt_world_build_bvh(world);            /* rt.cu:1653 */
  rt_camera cam;
  rt_camera_setup(settings, &cam);
  if (cam.img_width <= 0 || cam.img_height <= 0) return group_fail(g, RT_ERR_INVALID, "rt_group_render_world: image size must be positive");
  rt_scene_desc desc;
  desc.hittables = world->hittables; desc.hittable_count = world->hittable_count;
  desc.world_refs = world->world_refs; desc.world_count = world->world_count;
  desc.root_wref = world->root_wref;
  desc.texture_arena = world->texture_arena; desc.texture_bytes = world->texture_bytes;
  // pack once (host), upload to every device (<= 1 MB each)
  rtp::PackedScene packed;
  {
    rt_context *c0 = g->ctx[0];
    const int rc = rtp::pack_scene(desc, traversal, packed, c0->treelet_min, c0->fuse_min, c0->linear == 0 ? 0u : rtp::PROGRAM_MAX_DEFAULT);
    if (rc != RT_OK) return group_fail(g, rc, packed.error);
  }
  for (int i = 0; i < n; ++i) {
    const int rc = rtc::upload_packed(g->ctx[i], packed);
    if (rc != RT_OK) return group_fail(g, rc, g->ctx[i]->error);
  }
  const uint64_t n_pixels = (uint64_t)cam.img_width * cam.img_height;
  const int spp = settings->samples_per_pixel;
  rt_render_params base;
  rt_render_params_from_camera(settings, &cam, &base);
  base.flags = RT_RENDER_COUNT_RAYS | RT_RENDER_NO_RNG_CACHE;   // like render(): the XORWOW states are computed on every call
  const bool tiles = shard == RT_SHARD_TILES || n == 1;

  RTG_CUDA(g, cudaSetDevice(g->devices[0]));
  RTG_CUDA(g, cudaEventRecord(g->ev_all0, g->ctx[0]->stream));
  std::vector<int> my_spp(n, 0);
  for (int i = 0; i < n; ++i) {                                 // enqueue only: the devices render concurrently
    rt_context *c = g->ctx[i];
    RTG_CUDA(g, cudaSetDevice(g->devices[i]));
    RTG_CUDA(g, g->pixels[i].reserve(n_pixels * 3 * sizeof(float)));
    rt_render_params p = base;
    if (n > 1 && tiles) {
      p.band_rows = group_band_rows(n, cam.img_height); p.band_period = (uint32_t)n; p.band_phase = (uint32_t)i;
      my_spp[i] = spp;
    } else if (n > 1) {                                          // samples: the first spp % n devices take one extra
      my_spp[i] = spp / n + (i < spp % n ? 1 : 0);
      p.samples_per_pixel = my_spp[i];
      p.subsequence_offset = (uint64_t)i * n_pixels;
      p.flags |= RT_RENDER_OUTPUT_SUM;
    } else {
      my_spp[i] = spp;
    }
    if (n > 1) RTG_CUDA(g, cudaMemsetAsync(g->pixels[i].ptr, 0, n_pixels * 3 * sizeof(float), c->stream));
    RTG_CUDA(g, cudaEventRecord(g->ev0[i], c->stream));
    if (!(n > 1 && !tiles && my_spp[i] <= 0)) {                  // more devices than samples: this one contributes zeros
      const int rc = rt_render_device(c, &p, (float *)g->pixels[i].ptr, nullptr);
      if (rc != RT_OK) return group_fail(g, rc, c->error);
    }
    RTG_CUDA(g, cudaEventRecord(g->ev1[i], c->stream));
  }
  if (n > 1) {                                                   // the one exchange: sum onto device 0
    RTG_NCCL(g, g_nccl.GroupStart());
    for (int i = 0; i < n; ++i)
      RTG_NCCL(g, g_nccl.Reduce(g->pixels[i].ptr, g->pixels[0].ptr, (size_t)n_pixels * 3, kNcclFloat32, kNcclSum, 0, g->comms[i], g->ctx[i]->stream));
    RTG_NCCL(g, g_nccl.GroupEnd());
  }
  rt_context *c0 = g->ctx[0];
  RTG_CUDA(g, cudaSetDevice(g->devices[0]));
  const float scale = (n > 1 && !tiles) ? 1.0f / (float)spp : 1.0f;     // f32 like the reference's pixel_samples_scale (rt.cu:1657)
  if (h_rgb8) {
    RTG_CUDA(g, g->rgb8.reserve(n_pixels * 3 + 16));
    RTG_CUDA(g, rtk::launch_resolve((const float *)g->pixels[0].ptr, n_pixels, scale, (uint8_t *)g->rgb8.ptr, c0->stream));
  }
  RTG_CUDA(g, cudaEventRecord(g->ev_all1, c0->stream));
  if (h_rgb8) RTG_CUDA(g, cudaMemcpyAsync(h_rgb8, g->rgb8.ptr, n_pixels * 3, cudaMemcpyDeviceToHost, c0->stream));
  if (h_pixels_rgb) RTG_CUDA(g, cudaMemcpyAsync(h_pixels_rgb, g->pixels[0].ptr, n_pixels * 3 * sizeof(float), cudaMemcpyDeviceToHost, c0->stream));
  for (int i = n - 1; i >= 0; --i) {
    RTG_CUDA(g, cudaSetDevice(g->devices[i]));
    RTG_CUDA(g, cudaStreamSynchronize(g->ctx[i]->stream));
  }
  if (h_pixels_rgb && scale != 1.0f)                              // the float image a caller gets is the mean, as from one GPU
    for (uint64_t k = 0; k < n_pixels * 3; ++k) h_pixels_rgb[k] = scale * h_pixels_rgb[k];
  if (stats) {
    memset(stats, 0, sizeof(*stats));
    float worst = 0.0f, all = 0.0f;
    for (int i = 0; i < n; ++i) {
      RTG_CUDA(g, cudaSetDevice(g->devices[i]));
      float ms = 0.0f;
      cudaEventElapsedTime(&ms, g->ev0[i], g->ev1[i]);          // XORWOW init + render of device i
      if (ms > worst) worst = ms;
      unsigned long long counters[2] = {0, 0};
      RTG_CUDA(g, cudaMemcpy(counters, g->ctx[i]->b_counters.ptr, sizeof(counters), cudaMemcpyDeviceToHost));
      if (my_spp[i] > 0) stats->rays += counters[1];
    }
    RTG_CUDA(g, cudaSetDevice(g->devices[0]));
    cudaEventElapsedTime(&all, g->ev_all0, g->ev_all1);
    stats->render_ms = worst;                                    // slowest device
    stats->total_ms = all;                                       // device 0's clock: render + reduce + resolve
    stats->paths = n_pixels * (uint64_t)(spp > 0 ? spp : 0);
    stats->kernel_launches = (uint32_t)(2 * n + (h_rgb8 ? 1 : 0));
    stats->traversal_used = g->ctx[0]->packed.traversal_used;
    stats->grid_blocks = g->ctx[0]->last_launch.grid; stats->block_threads = g->ctx[0]->last_launch.block;
    stats->regs_per_thread = g->ctx[0]->last_launch.regs; stats->blocks_per_sm = g->ctx[0]->last_launch.blocks_per_sm;
  }
  return RT_OK;
}

}  // extern "C"
