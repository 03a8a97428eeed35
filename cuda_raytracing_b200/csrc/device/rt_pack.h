// rt_pack.h -- host-side packer: reference-format scene arrays -> device layout (rt_device.h).
#ifndef RT_PACK_H
#define RT_PACK_H

#include <stdint.h>
#include <map>
#include <string>
#include <vector>

#include "rt_api.h"
#include "rt_device.h"

namespace rtp {

struct PackedScene {
  std::vector<float4> ref_nodes, xforms, instances, fast_nodes, spheres, planars, checkers, perlin_vectors;
  std::vector<float4> program;           // reference-order walk as a linear program (2 per op), empty when not available
  uint32_t program_boxes = 0, program_prims = 0, program_wrappers = 0;   // op counts by kind (wrappers: ways in)
  // every planar op of the program is a QUAD whose normal / u / v / w each have ONE non-zero component (walls, boxes):
  // planars_axis holds its inside test in reduced form, 4 words per planar like `planars` (see rt_pack.cpp axis_aligned_words)
  std::vector<float4> planars_axis;
  bool program_axis_quads = false;
  std::vector<uint2> structures;
  std::vector<uint32_t> struct_children;
  std::vector<uint4> materials;          // 2 per primitive id
  std::vector<uint8_t> perlin_perm, tex_arena;
  uint32_t ref_root = 0, fast_root = 0;
  uint32_t sphere_count = 0, planar_count = 0, perlin_tables = 0;
  bool has_wrappers = false;             // any translation / y-rotation in the scene
  bool has_fast = false;                 // fast_nodes valid
  bool flat = false;                     // own-BVH scene of at most 4 spheres and 4 planars: walked without the BVH (trace_flat)
  uint32_t flat_triangles = 0;           // bit i: planar i is a triangle
  uint32_t ref_stack_need = 0, fast_stack_need = 0;
  uint32_t treelets = 0, treelet_prims = 0;   // own-BVH stand-ins inside the reference graph
  uint32_t fused_instances = 0;               // wrapper chains around one primitive / structure done as one item
  uint32_t instanced_leaves = 0;              // FAST traversal: leaves of the own BVH tested in the primitive's own space
  uint32_t traversal_used = 0;           // RT_TRAVERSAL_REFERENCE or RT_TRAVERSAL_FAST
  std::map<uint64_t, uint32_t> checker_ids, noise_ids;   // arena offset -> table index (dedup)
  std::string error;
};

// returns RT_OK or an RT_ERR_* code (text in out.error)
// treelet_min: RT_TRAVERSAL_AUTO on a scene with wrappers replaces wrapper-free subtrees of the reference BVH that hold
// at least this many primitives by an own SAH BVH (0 = never)
// fuse_min: the same mode does translation / y-rotation chains around one object as one item of the walk when the scene
// has a treelet or at least this many such chains (0 = never)
// program_max: a scene with wrappers whose whole reference-order walk fits this many ops also gets the linear program
// (rt_device.h Scene::program; 0 = never)
constexpr uint32_t PROGRAM_MAX_DEFAULT = 256;
int pack_scene(const rt_scene_desc &desc, uint32_t traversal, PackedScene &out, uint32_t treelet_min = 8,
               uint32_t fuse_min = 8, uint32_t program_max = PROGRAM_MAX_DEFAULT);

// the image of the linear program the kernels walk (rt_device.h, Scene::program): p.program with its indices turned into
// byte offsets relative to the op, the end op, then the planar records (the reduced ones with `axis`) and the sphere records
void build_linear_image(const PackedScene &p, bool axis, std::vector<float4> &image);

// one reference Material (24 B) -> the two 16-byte words of a device material record, pre-resolving its
// texture against `arena`; may append to out.checkers / perlin tables.
void pack_material(const rt_material &m, const uint8_t *arena, uint64_t arena_bytes, PackedScene &out, uint4 rec[2]);

// GF(2) skip-ahead matrices M^(2^(67+k)), k = 0..63, 160 rows x 5 words each (XORWOW subsequence spacing)
void xorwow_sequence_matrices(std::vector<uint32_t> &out);

}  // namespace rtp
#endif
