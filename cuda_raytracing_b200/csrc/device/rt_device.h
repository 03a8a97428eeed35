// rt_device.h -- device-side scene layout shared by the host packer (rt_pack.cpp) and the kernels.
//
// The reference keeps one 120-byte AoS tagged union per object and reaches it through two dependent
// indirections (u16 world ref -> u16 hittable ref -> struct; rt.cu:1127-1128, 1091).  Here every object
// kind has its own array of 16-byte-aligned records read with 128-bit ld.global.nc, and every link is a
// 32-bit TAGGED reference (kind in the top 3 bits, index below) resolved at pack time, so a stack pop
// yields the kind without touching memory.
#ifndef RT_DEVICE_H
#define RT_DEVICE_H

#include <stdint.h>

#if defined(__CUDACC__)
#include <vector_types.h>
#else
struct alignas(16) float4 { float x, y, z, w; };
struct alignas(16) uint4 { unsigned x, y, z, w; };
struct alignas(8) uint2 { unsigned x, y; };
#endif

namespace rtd {

// tagged reference ------------------------------------------------------------------------------------
enum Kind : uint32_t {
  K_SPHERE = 0, K_BVH = 1, K_TRIANGLE = 2, K_QUAD = 3, K_TRANSLATION = 4, K_YROTATION = 5, K_STRUCTURE = 6,
  K_NONE = 7
};
constexpr uint32_t KIND_SHIFT = 29;
constexpr uint32_t INDEX_MASK = (1u << KIND_SHIFT) - 1u;
#if defined(__CUDACC__)
#define RTD_HD __host__ __device__ __forceinline__
#else
#define RTD_HD inline
#endif
RTD_HD uint32_t make_ref(uint32_t kind, uint32_t index) { return (kind << KIND_SHIFT) | index; }
RTD_HD uint32_t ref_kind(uint32_t r) { return r >> KIND_SHIFT; }
RTD_HD uint32_t ref_index(uint32_t r) { return r & INDEX_MASK; }
// Treelets: inside the reference-order graph a link of kind K_NONE with index i > 0 points at node i - 1 of
// `fast_nodes` -- an own SAH BVH that stands in for a wrapper-free subtree of the reference's BVH (index 0 stays
// "nothing": the empty second child of a one-primitive scene).
RTD_HD uint32_t make_own_ref(uint32_t node) { return (K_NONE << KIND_SHIFT) | (node + 1u); }
// Fused instances: a K_NONE link with INSTANCE_BIT set points at record (index & ~INSTANCE_BIT) of `instances` -- a
// translation and/or y-rotation around one primitive or one structure of primitives, done as ONE item of the walk
// (enter, test, leave: the same operations in the same order as the wrapper items it replaces).
constexpr uint32_t INSTANCE_BIT = 1u << 28;
RTD_HD uint32_t make_instance_ref(uint32_t i) { return (K_NONE << KIND_SHIFT) | INSTANCE_BIT | i; }
constexpr uint32_t INST_TRANSLATE = 1u, INST_ROTATE = 2u;

// material record: 2 x 16 B per primitive (prim id = sphere index, or sphere_count + planar index) -------
enum MatType : uint32_t { M_LAMBERTIAN = 0, M_METAL = 1, M_DIELECTRIC = 2, M_DIFFUSE_LIGHT = 3 };   // rt.cu:79-84
enum TexKind : uint32_t {
  TK_SOLID = 0,          // colour inline in m1.xyz (pre-resolved at pack time)
  TK_CHECKER_SOLID = 1,  // checker of two solid colours: record `aux` in the checker table
  TK_NOISE = 2,          // Perlin table `aux`, scale in m1.w
  TK_ARENA = 3           // anything else: walk the reference-format texture arena from `tex_offset`
};
// m0 = { type | tex_kind << 8 | needs_uv << 16, tex_offset, aux, 0 }
// m1 = { albedo/colour xyz, fuzz | refraction index | noise scale }

struct Scene {
  // reference-order graph (RT_TRAVERSAL_REFERENCE)
  const float4 *ref_nodes;      // 2 per BVH node: {x0,x1,y0,y1} {z0,z1, left ref, right ref} (refs as raw bits)
  const float4 *xforms;         // 1 per wrapper: translation {ox,oy,oz, child ref}; y-rotation {sin,cos,0, child ref}
  const uint2 *structures;      // {first child slot, count}
  const uint32_t *struct_children;
  const float4 *instances;      // 2 per fused instance: {offset.xyz, child ref} {sin, cos, INST_* flags, 0}
  uint32_t ref_root;            // tagged ref of the reference's root BVH node
  // the same walk as a linear program (small scenes; trace_linear).  `program` points at the program's IMAGE (built at upload,
  // rt_pack.cpp build_linear_image): the ops, 2 x float4 each, op i followed by op i + 1, then the planar and the sphere records
  // the ops test -- one block that the kernels of modes 3 / 4 copy to shared memory, addressed RELATIVE to the op at hand
  // (byte offsets inside the ops), so that the walk needs no base address:
  //   box          {x0,x1,y0,y1} {z0,z1, byte offset to the op a missed box continues at, K_BVH}
  //   primitive    {n.xyz, d} of a quad / triangle (its record's first word), 0 for a sphere
  //                              {byte offset to its record, 0, sphere / planar index, K_SPHERE|K_TRIANGLE|K_QUAD (| OP_IN_WRAPPER)}
  //   translation  {offset,0}    {0,0,0, K_TRANSLATION (way in) | OP_T_EXIT (way out)}
  //   y-rotation   {sin,cos,0,0} {0,0,0, K_YROTATION (way in) | OP_R_EXIT (way out)}
  //   translate(rotate(x)) as one op each way (OP_FUSED_WRAPPER = 0x200):
  //                {offset,0}    {sin,cos,0, K_TRANSLATION | 0x200}         way in: translate, then rotate
  //                {sin,cos,offset.xy} {offset.z,0,0, OP_R_EXIT | 0x200}    way out: un-rotate, then un-translate
  //   end          {0} {0,0,0, K_SPHERE | OP_END (0x400)}                   the last op: the walk returns
  //   planar records: {n.xyz,d} {q.xyz,w.x} {v.xyz,w.y} {u.xyz,w.z}, or with linear_axis {n.xyz,d} {q.xyz,w.k} {S_alpha,0} {S_beta,0}
  const float4 *program;
  uint32_t program_ops;         // ops of the image, the end op included; 0: the scene has no program (too big, or wrappers nested in a way only the stack walk mirrors)
  uint32_t linear_float4s;      // size of the image in float4 (ops + records)
  uint32_t linear_axis;         // 1: every planar op is an axis-aligned quad and the image holds the reduced records (rt_pack.cpp axis_aligned_words)
  uint32_t stack_entries;       // traversal stack entries per thread this scene's walk can need (<= STACK_CAPACITY)
  // own BVH over world-space primitives (RT_TRAVERSAL_FAST): 4 x float4 per node
  //   n0 = {c0.x0, c0.x1, c0.y0, c0.y1}  n1 = {c1.x0, c1.x1, c1.y0, c1.y1}
  //   n2 = {c0.z0, c0.z1, c1.z0, c1.z1}  n3 = {child0 ref, child1 ref, 0, 0}
  //   child refs are K_BVH links in the whole-scene BVH, make_own_ref() links in a treelet of the reference graph
  const float4 *fast_nodes;
  uint32_t fast_root;
  // primitives
  const float4 *spheres;        // 2 per sphere: {cx,cy,cz,r} {mx,my,mz, flags (bit0: needs uv)}
  const float4 *planars;        // 4 per quad/triangle: {n.xyz,d} {q.xyz,w.x} {v.xyz,w.y} {u.xyz,w.z}
  uint32_t sphere_count, planar_count;
  uint32_t flat_triangles;      // trace_flat (tiny scenes, kernel mode 5): bit i set = planar i is a triangle
  // second primitive set for flattened instances (FAST mode on scenes with wrappers) lives in the same arrays
  const uint4 *materials;       // 2 x 16 B per primitive id
  const float4 *checkers;       // 2 per record: {inv_scale, even.rgb} {0, odd.rgb}
  const float4 *perlin_vectors; // 256 per noise table (xyz, 0)
  const uint8_t *perlin_perm;   // 256 per noise table (perm_x; the reference hashes all axes with it)
  uint32_t perlin_tables;
  const uint8_t *tex_arena;     // reference-format arena (TK_ARENA fallback), 16-byte aligned copy
  uint64_t tex_bytes;
};

// The linear program and the primitive records it reads are staged in shared memory by the kernels of modes 3 / 4; the
// packer only keeps a program whose staging area fits (256 ops of 32 B + their primitives: 24 KB at most)
constexpr uint32_t LINEAR_SMEM_MAX_BYTES = 24576;
RTD_HD uint32_t linear_stage_bytes(uint32_t program_ops, uint32_t planar_count, uint32_t sphere_count) {
  return 32u * (program_ops + 1u) + 64u * planar_count + 32u * sphere_count;      // + the end op
}

constexpr int MAX_CHUNK_TABLE = 32;

struct RenderArgs {
  Scene scene;
  float *pixels;                // img_width*img_height*3 f32 (device)
  uint32_t *rng_states;         // 6 words per pixel of the rendered row range: {d, v0..v4}; advanced chunk by chunk
  uint32_t *progress;           // per pixel of the range: chunks completed (zeroed before the launch)
  unsigned long long *counters; // [0] next work item, [1] rays traced, [2..13] operation counters (count_ops)
  unsigned long long item_count;// slots_per_chunk * chunk_count work items; item = chunk * slots_per_chunk + slot
  uint32_t slots_per_chunk;     // 32 * number of 32-pixel tiles covering the rendered rows (>= pixel_count)
  uint32_t tiles_x, tile_shift; // tiles per row of tiles; log2 of the tile width (tile = 2^s x 2^(5-s) pixels)
  uint32_t scatter_mul;         // 0: slots in tile order; else slot -> slot * scatter_mul mod slots_per_chunk (coprime: a
                                // permutation), so items claimed together are pixels far apart (scenes bound by slow pixels)
  int32_t first_row;            // first_pixel / img_width
  int32_t band_rows, band_period, band_phase;   // band_rows > 0: row y of the rendered set is image row
                                // ((y / band_rows) * band_period + band_phase) * band_rows + y % band_rows (interleaved bands)
  int32_t chunk_samples, chunk_count;   // a pixel's samples are produced in order, a chunk at a time: chunk_samples each
                                // (uniform), or -- chunk_samples == 0 -- samples [chunk_begin[k], chunk_begin[k + 1]) for chunk k
  int32_t chunk_begin[MAX_CHUNK_TABLE + 1];   // equal chunks for the bulk of the render, halving ones for its tail
  int32_t claim_group;          // idle lanes of a warp claim work items together, this many at a time
  int32_t stack_entries;        // traversal stack entries per thread that this scene needs (<= STACK_CAPACITY): sizes the shared memory
  int32_t img_width, img_height, samples_per_pixel, max_bounces;
  int32_t first_pixel, pixel_count;   // rendered range (row_begin*W .. row_end*W)
  float defocus_angle;
  float background[3];
  float dv[3], du[3], cam[3], ddu[3], ddv[3], p00[3];
  float sample_scale;           // 1/spp, or 1 with RT_RENDER_OUTPUT_SUM
  uint32_t resume;               // chunk 0 starts from the sums already in `pixels` (progressive accumulation)
  uint32_t count_rays;
  uint32_t count_ops;            // run the instrumented kernel variant (per-ray operation counts for F_ray)
  uint32_t forward;              // RT_RENDER_FORWARD_THROUGHPUT: attenuation product carried forwards in registers
  uint32_t linear_smem_word;     // kernel modes 3 / 4: word offset of the staged program / records inside dynamic shared memory
};

constexpr int STACK_CAPACITY = 32;     // traversal stack entries per thread (shared memory, [slot][thread])
constexpr int MAX_BOUNCE_STACK = 256;  // RayColorStack.max_count is a u8 in the reference (rt.cu:220)

}  // namespace rtd
#endif
