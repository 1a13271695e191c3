// bvh_build.h -- host BVH builder interface (internal).
#pragma once
#include <stdint.h>
#include <vector>

#include "vrlb_device.cuh"

namespace vrlb {
// verts: n*9 floats (p0 p1 p2 per triangle), flags: n surface words (may be null).
// Produces the flattened binary BVH (root = node 0) and the triangles in leaf order.
void build_bvh(const float *verts, const uint32_t *flags, uint32_t n, std::vector<DNode> &nodes, std::vector<DTri> &tris);
}  // namespace vrlb
