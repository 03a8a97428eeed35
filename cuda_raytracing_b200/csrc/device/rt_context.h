// rt_context.h -- internal (C++): the context behind the C ABI, shared by rt_api.cu (one GPU) and rt_group.cu (several).
#ifndef RT_CONTEXT_H
#define RT_CONTEXT_H

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "rt_api.h"
#include "rt_device.h"
#include "rt_kernels.h"
#include "rt_pack.h"

namespace rtc {

struct DeviceBuffer {
  void *ptr = nullptr; size_t bytes = 0;
  cudaError_t upload(const void *src, size_t n, cudaStream_t s) {
    release();
    bytes = n;
    if (n == 0) { return cudaMalloc(&ptr, 16); }                 // keep pointers valid for empty tables
    cudaError_t e = cudaMalloc(&ptr, n);
    if (e != cudaSuccess) return e;
    return cudaMemcpyAsync(ptr, src, n, cudaMemcpyHostToDevice, s);
  }
  cudaError_t reserve(size_t n) {
    if (n <= bytes && ptr) return cudaSuccess;
    release();
    bytes = n;
    return cudaMalloc(&ptr, n ? n : 16);
  }
  void release() { if (ptr) cudaFree(ptr); ptr = nullptr; bytes = 0; }
};

// sets the per-thread error text (and the context's, when there is one); returns `code`
int fail(rt_context *ctx, int code, const std::string &msg);
// uploads an already packed scene (rt_scene_upload = pack_scene + this); the group packs once for all its devices
int upload_packed(rt_context *ctx, const rtp::PackedScene &packed);

}  // namespace rtc

struct rt_context {
  int device = 0, sm_count = 0;
  cudaStream_t own_stream = nullptr, stream = nullptr;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  std::string error;
  // scene
  bool has_scene = false;
  rtd::Scene scene{};
  rtp::PackedScene packed;
  rtc::DeviceBuffer b_ref_nodes, b_xforms, b_instances, b_structs, b_struct_children, b_fast_nodes, b_spheres, b_planars, b_materials,
      b_checkers, b_perlin_vec, b_perlin_perm, b_arena, b_program;
  // rng
  rtc::DeviceBuffer b_rng_mats, b_rng_init, b_rng_states, b_progress, b_counters;   // b_rng_init: pristine states (cache)
  int chunk_samples = 0;                                   // 0 = automatic; env RT_CHUNK overrides (tuning knob)
  int claim_group = 0;                                     // 0 = automatic; env RT_CLAIM_GROUP (tuning knob)
  int axis_quads = 1;                                      // env RT_AXIS_QUADS=0: keep the general inside test for axis-aligned quads (bisecting)
  int tail_min = 8;                                        // env RT_TAIL_MIN: the halving tail of the chunk schedule stops at this many samples
  int first_chunk = 0;                                     // env RT_FIRST_CHUNK: a short first chunk in front of the schedule (experiments)
  uint32_t fuse_min = 8;                                   // env RT_FUSE_MIN: wrapper chains around one object become one item when there are this many (0 = never)
  uint32_t treelet_min = 8;                                // env RT_TREELET_MIN: smallest wrapper-free subtree given an own BVH (0 = none)
  int tile_w = 8;                                          // work items walk 8x4-pixel tiles; env RT_TILE_W (1..32, power of 2)
  int linear = -1;                                         // small scenes with wrappers walk their linear program (kernel mode 3): -1 / 1 yes, 0 never; env RT_LINEAR
  int flat = -1;                                           // tiny scenes without wrappers skip their BVH (kernel mode 5): -1 / 1 yes, 0 never; env RT_FLAT
  int scatter = -1;                                        // item order: -1 automatic, 0 tile order, 1 scattered; env RT_SCATTER
  uint64_t rng_seed = 0, rng_first = 0, rng_count = 0, rng_bands = 0; bool rng_valid = false;   // what b_rng_init holds
  uint64_t live_seed = 0, live_first = 0, live_count = 0, live_bands = 0; bool live_valid = false;  // what b_rng_states carries on from
  // scratch for rt_render (host-buffer path)
  rtc::DeviceBuffer b_pixels, b_rgb8;
  rtk::LaunchInfo last_launch{};
};

#endif
