// rt_kernels.h -- internal (C++) launch interface between rt_api.cu and rt_kernels.cu.
#ifndef RT_KERNELS_H
#define RT_KERNELS_H

#include <cuda_runtime.h>
#include <stdint.h>

#include "rt_device.h"

struct rt_probe_hit;   // rt_api.h

namespace rtk {

struct LaunchInfo { int grid, block, regs, blocks_per_sm; size_t smem; };

// mode: 0 = reference-order traversal, 1 = own BVH, 2 = reference order with own-BVH treelets, 3 = reference order as a
// linear program, no stack (args.scene.program), 4 = the same for short-lived pixels (fewer resident blocks, more registers),
// 5 = tiny scene without wrappers: every primitive tested, no BVH, no stack.  Enqueues on `stream`; *info (optional) describes the launch.
cudaError_t launch_render(const rtd::RenderArgs &args, int mode, int sm_count, cudaStream_t stream, LaunchInfo *info);

// XORWOW states {d,v0..v4} for subsequences first..first+n-1 of `seed`.
// seq_matrices: 64 matrices M^(2^(67+k)), each 160 rows x 5 words (device).
cudaError_t launch_rng_init(uint64_t seed, uint64_t first_subsequence, uint64_t n, const uint32_t *seq_matrices,
                            uint32_t *states_out, cudaStream_t stream);

// rgb8 = (u8)(256 * clamp(sqrt(scale * x), 0, 0.999)) for 3*n_pixels floats
cudaError_t launch_resolve(const float *pixels, uint64_t n_pixels, float scale, uint8_t *rgb8, cudaStream_t stream);

cudaError_t launch_probe_trace(const rtd::Scene &scene, int mode, uint64_t n, const float *origins, const float *directions,
                               const float *times, float tmin, float tmax, rt_probe_hit *out, cudaStream_t stream);
// materials2: 2 x uint4 packed material records per item (rt_device.h), built by the packer
cudaError_t launch_probe_scatter(const rtd::Scene &scene, uint64_t n, const uint4 *materials2, const float *dir_in,
                                 const rt_probe_hit *rec, const float *unit_vec, const float *random_test, float *dir_out,
                                 float *atten_out, int32_t *scattered, cudaStream_t stream);
cudaError_t launch_probe_texture(const rtd::Scene &scene, uint64_t n, uint32_t texture_offset, const float *uv,
                                 const float *points, float *rgb_out, cudaStream_t stream);

}  // namespace rtk
#endif
