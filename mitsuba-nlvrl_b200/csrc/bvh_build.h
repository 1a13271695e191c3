// bvh_build.h -- host BVH builder interface (internal).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <functional>
#include <vector>

#include "vrlb_device.cuh"

namespace vrlb {
// verts: n*9 floats (p0 p1 p2 per triangle), flags: n surface words (may be null).
// Produces the flattened binary BVH (root = node 0) and the triangles in leaf order.
void build_bvh(const float *verts, const uint32_t *flags, uint32_t n, std::vector<DNode> &nodes, std::vector<DTri> &tris);
// The same tree layout built on the device (bvh_lbvh.cu, LBVH): d_verts / d_flags are device arrays in ORIGINAL triangle
// order, lo / hi the geometry's bounding box, pad the host build's box margin; writes 2n - 1 nodes and n triangles.
// *max_depth = depth of the tree (root = 0).
cudaError_t build_bvh_device(const float *d_verts, const uint32_t *d_flags, uint32_t n, float pad, const float lo[3], const float hi[3],
                             DNode *d_nodes, DTri *d_tris, int *max_depth, uint64_t *launches, cudaStream_t stream);
int last_bvh_depth();      // depth (root = 0) of the tree the last build_bvh call produced; bounded by kMaxBvhDepth (bvh_build.cpp)
// Re-packs the binary BVH for the shadow traversal: one DWNode per inner node holding both children's boxes.
// Empty when the root is a leaf.
void build_wide_nodes(const std::vector<DNode> &nodes, std::vector<DWNode> &wnodes);
// Shadow (any-hit) structure: main BVH over ordinary geometry + per-shell BVHs (see bvh_build.cpp).
struct ShadowAccel {
    std::vector<DTri>   tris;      // own triangle order (prim = original index)
    std::vector<DWNode> wnodes;
    std::vector<DShell> shells;
    std::vector<DSolid> solids;    // largest non-shell components, each with its own BVH + proxies (occluder_proxies.cpp)
    uint32_t main_root = 0x80000000u;
    float    main_lo[3] = { 0, 0, 0 }, main_hi[3] = { 0, 0, 0 };
    bool     has_main = false;
    uint32_t n_main_tris = 0;
    int      max_depth = 0;        // deepest of the BVHs in here
};
// proxies of one component (triangle indices `sub` into verts); fills everything but `root`
void build_solid_proxies(const float *verts, const std::vector<uint32_t> &sub, float margin, DSolid &out);
// `custom`, when given, builds the binary tree of one component (host arrays in, nodes / triangles in leaf order and the
// tree's depth out; false = could not, use the host build): how the device LBVH (bvh_lbvh.cu) is plugged in.  A tree
// deeper than the shadow traversals' stacks allow (kMaxBvhDepth, bvh_build.cpp) is replaced by the host build.
using SubsetBuilder = std::function<bool(const float *verts, const uint32_t *flags, uint32_t n, std::vector<DNode> &nodes, std::vector<DTri> &tris, int &depth)>;
void build_shadow_accel(const float *verts, const uint32_t *flags, uint32_t n, ShadowAccel &out, const SubsetBuilder *custom = nullptr, uint32_t *n_custom = nullptr);
}  // namespace vrlb
