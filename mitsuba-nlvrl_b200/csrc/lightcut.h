// lightcut.h -- host-visible interface of the VRL light cut (lightcut.cu), internal to libvrl_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "vrlb_device.cuh"

namespace vrlb {

struct DLcNode {             // 96 B: six 16-byte rows.  A cluster of the light tree (VRLLightCut::Node, vrl_lightcut.h:35-63)
    float ox, oy, oz, length;            // the representative's geometry ...
    float dx, dy, dz; uint32_t medium;
    float fr, fg, fb; uint32_t rep_leaf; // ... carrying the cluster's flux; the VRL it comes from (PCG32 stream address)
    float lox, loy, loz, max_len;        // aabb of every VRL of the cluster; maxVRLLenght
    float hix, hiy, hiz; uint32_t n_leaves;
    int32_t child0, child1, same;        // -1 / -1 / -1 for a leaf; `same` = the child sharing the representative (isSame)
    uint32_t first_leaf;                 // the cluster's VRLs = leaf_order[first_leaf, first_leaf + n_leaves)
};

struct LightcutTree;
// Builds the tree over n VRL records resident on the device; *launches is incremented by the kernels launched.
cudaError_t lightcut_build(const DVrl *d_vrls, uint64_t n, uint64_t first_index, uint64_t base_seed, uint64_t seed_offset,
                           cudaStream_t stream, LightcutTree **out, uint64_t *launches);
void lightcut_free(LightcutTree *t);
const DLcNode *lightcut_nodes(const LightcutTree *t);
const uint32_t *lightcut_leaf_order(const LightcutTree *t);
const ulonglong2 *lightcut_jump(const LightcutTree *t);
const float4 *lightcut_leaf_geom(const LightcutTree *t);
uint32_t lightcut_root(const LightcutTree *t);
uint64_t lightcut_n_vrls(const LightcutTree *t);
void lightcut_build_stats(const LightcutTree *t, uint64_t *n_nodes, uint64_t *searches, float *ms);
uint64_t lightcut_dump(const LightcutTree *t, float *f13, int32_t *i6, uint64_t cap);

struct LightcutParams {
    DScene              scene;
    const vrlb_segment *segs; uint64_t n_segs;
    const DLcNode      *nodes; const uint32_t *leaf_order; const ulonglong2 *jump;
    const float4       *leaf_geom;       // (o, length), (d, -) of the VRLs in leaf order
    const DVrl         *vrls;            // the set itself: the exact sum of a query whose heap overflows
    uint64_t            n_vrls, first_index, total;
    uint32_t            root;
    float               scale, err;      // VRLMap::m_scale; m_errorRatio
    int32_t             better_dist;     // m_thresholdBetterDist
    uint32_t            q;               // samples_per_query
    unsigned long long *heap_key; float4 *slot_est; uint32_t cap;     // per-thread refinement heaps, interleaved across the grid's threads
    float              *film_rgb, *seg_out, *seg_stats;               // seg_stats: n_segs * 4 = evaluations, refinements, largest heap, overflowed
    unsigned long long *stats;           // 8 words, zeroed before the launch: queries, fresh evaluations, refinements, overflowed queries, largest heap, next query
};
// blocks of the query kernel such that its heaps fit `budget_bytes`
uint32_t lightcut_grid(uint64_t n_segs, uint32_t cap, int n_sms, size_t budget_bytes);
cudaError_t launch_lightcut_gather(const LightcutParams &p, uint32_t grid, bool uniform, bool vis, cudaStream_t stream);

}  // namespace vrlb
