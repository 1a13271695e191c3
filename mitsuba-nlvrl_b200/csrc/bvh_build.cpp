// bvh_build.cpp -- host-side construction of the device BVH (binned SAH, binary, flattened).
//
// Stands in for ShapeKDTree construction (include/mitsuba/render/kdtree.h, SAH kd-tree) of the
// reference.  Only the SET of triangle hits matters for parity: hits are decided by the exact
// Moeller-Trumbore test on the device (vrlb_device.cuh), node boxes are padded so the hierarchy can
// only admit more triangles to that test, never fewer.  A GPU LBVH/wide-BVH builder is the next step
// (DESIGN.md); scenes on this path have <= a few thousand primitives, so the host build is microseconds.
#include "bvh_build.h"

#include <algorithm>
#include <cmath>
#include <limits>

namespace vrlb {

namespace {
constexpr float INF = std::numeric_limits<float>::infinity();
struct Box {
    float lo[3] = { INF, INF, INF }, hi[3] = { -INF, -INF, -INF };
    void grow(const float *p) { for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); } }
    void grow(const Box &b) { for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); } }
    float area() const {
        float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        return (dx < 0 || dy < 0 || dz < 0) ? 0.f : 2.f * (dx * dy + dy * dz + dz * dx);
    }
};
}  // namespace

void build_bvh(const float *verts, const uint32_t *flags, uint32_t n, std::vector<DNode> &nodes, std::vector<DTri> &tris) {
    nodes.clear(); tris.clear();
    if (n == 0) return;
    std::vector<Box> tb(n);
    std::vector<float> cen((size_t) n * 3);
    float ext = 0.f;
    for (uint32_t i = 0; i < n; ++i) {
        const float *p = verts + (size_t) i * 9;
        tb[i].grow(p); tb[i].grow(p + 3); tb[i].grow(p + 6);
        for (int k = 0; k < 3; ++k) {
            cen[(size_t) i * 3 + k] = (p[k] + p[3 + k] + p[6 + k]) / 3.0f;
            ext = std::max(ext, std::max(std::fabs(tb[i].lo[k]), std::fabs(tb[i].hi[k])));
        }
    }
    const float pad = 1e-4f * ext + 1e-6f;
    std::vector<uint32_t> order(n);
    for (uint32_t i = 0; i < n; ++i) order[i] = i;

    struct Item { uint32_t begin, end; int node; };
    std::vector<Item> stack;
    nodes.push_back(DNode());
    stack.push_back({ 0, n, 0 });
    constexpr int NB = 16;
    while (!stack.empty()) {
        Item it = stack.back(); stack.pop_back();
        Box box, cbox;
        for (uint32_t i = it.begin; i < it.end; ++i) { box.grow(tb[order[i]]); cbox.grow(&cen[(size_t) order[i] * 3]); }
        DNode nd;
        for (int k = 0; k < 3; ++k) { nd.lo[k] = box.lo[k] - pad; nd.hi[k] = box.hi[k] + pad; }
        uint32_t count = it.end - it.begin;
        bool leaf = count <= 4;
        uint32_t mid = 0;
        if (!leaf) {
            int axis = -1, best_bin = -1; float best = INF;
            for (int k = 0; k < 3; ++k) {
                float e = cbox.hi[k] - cbox.lo[k];
                if (!(e > 0.f)) continue;
                Box bb[NB]; int cnt[NB] = { 0 };
                for (uint32_t i = it.begin; i < it.end; ++i) {
                    int b = std::min(NB - 1, (int) ((cen[(size_t) order[i] * 3 + k] - cbox.lo[k]) / e * NB));
                    cnt[b]++; bb[b].grow(tb[order[i]]);
                }
                float la[NB], ra[NB]; int lc[NB], rc[NB];
                Box acc; int c = 0;
                for (int b = 0; b < NB; ++b) { acc.grow(bb[b]); c += cnt[b]; la[b] = acc.area(); lc[b] = c; }
                acc = Box(); c = 0;
                for (int b = NB - 1; b >= 0; --b) { acc.grow(bb[b]); c += cnt[b]; ra[b] = acc.area(); rc[b] = c; }
                for (int b = 0; b + 1 < NB; ++b) {
                    if (lc[b] == 0 || rc[b + 1] == 0) continue;
                    float cost = la[b] * lc[b] + ra[b + 1] * rc[b + 1];
                    if (cost < best) { best = cost; axis = k; best_bin = b; }
                }
            }
            if (axis >= 0) {
                float e = cbox.hi[axis] - cbox.lo[axis];
                auto mid_it = std::stable_partition(order.begin() + it.begin, order.begin() + it.end, [&](uint32_t idx) {
                    return std::min(NB - 1, (int) ((cen[(size_t) idx * 3 + axis] - cbox.lo[axis]) / e * NB)) <= best_bin;
                });
                mid = (uint32_t) (mid_it - order.begin());
            } else {
                mid = (it.begin + it.end) / 2;
            }
            if (mid == it.begin || mid == it.end) mid = (it.begin + it.end) / 2;
            if (count > 16 || axis >= 0) {
                nd.a = (int32_t) nodes.size(); nodes.push_back(DNode());
                nd.b = (int32_t) nodes.size(); nodes.push_back(DNode());
                stack.push_back({ it.begin, mid, nd.a });
                stack.push_back({ mid, it.end, nd.b });
            } else {
                leaf = true;
            }
        }
        if (leaf) { nd.a = (int32_t) it.begin; nd.b = -(int32_t) count; }
        nodes[(size_t) it.node] = nd;
    }
    tris.resize(n);
    for (uint32_t i = 0; i < n; ++i) {
        uint32_t prim = order[i];
        const float *p = verts + (size_t) prim * 9;
        DTri t;
        for (int k = 0; k < 3; ++k) { t.p0[k] = p[k]; t.e1[k] = p[3 + k] - p[k]; t.e2[k] = p[6 + k] - p[k]; }
        t.prim = prim; t.flags = flags ? flags[prim] : 0u; t.pad = 0;
        tris[i] = t;
    }
}

}  // namespace vrlb
