// bvh_build.cpp -- host-side construction of the device BVH (binned SAH, binary, flattened).
//
// Stands in for ShapeKDTree construction (include/mitsuba/render/kdtree.h, SAH kd-tree) of the
// reference.  Only the SET of triangle hits matters for parity: hits are decided by the exact
// Moeller-Trumbore test on the device (vrlb_device.cuh), node boxes are padded so the hierarchy can
// only admit more triangles to that test, never fewer.  A GPU LBVH/wide-BVH builder is the next step
// (DESIGN.md); scenes on this path have <= a few thousand primitives, so the host build is microseconds.
#include "bvh_build.h"

#include <algorithm>
#include <cstring>
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

// Depth bound.  The device traversals keep fixed per-lane stacks (vrlb_device.cuh: closest_hit 48 entries, two pushes per
// level; traverse_any_hit 32, one push per level; gather_fast.cu trace_queue 48 incl. up to 14 pre-pushed single-leaf
// roots), and a binned-SAH build has no depth bound of its own (a geometric progression of triangle sizes peels one
// triangle per level).  So the build switches to balanced median splits as soon as the SAH's depth plus the levels a
// balanced tree still needs for the node's triangles would reach kMaxBvhDepth; 2^19 - 1 triangles need 18 levels.
constexpr int kMaxBvhDepth = 28;
static_assert(kMaxBvhDepth + 1 <= 48 && kMaxBvhDepth + 1 <= 32 && kMaxBvhDepth + VRLB_MAX_SHELLS + VRLB_MAX_SOLIDS <= 48,
              "traversal stacks of vrlb_device.cuh / gather_fast.cu");
static int g_last_depth = 0;
int last_bvh_depth() { return g_last_depth; }

void build_bvh(const float *verts, const uint32_t *flags, uint32_t n, std::vector<DNode> &nodes, std::vector<DTri> &tris) {
    nodes.clear(); tris.clear();
    g_last_depth = 0;
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

    struct Item { uint32_t begin, end; int node; int depth; };
    std::vector<Item> stack;
    nodes.push_back(DNode());
    stack.push_back({ 0, n, 0, 0 });
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
        g_last_depth = std::max(g_last_depth, it.depth);
        int balanced_levels = 0;                                  // levels a balanced tree needs below this node
        while ((4u << balanced_levels) < count) ++balanced_levels;
        const bool force_median = it.depth + balanced_levels + 1 >= kMaxBvhDepth;
        if (!leaf && force_median) {
            // object median along the widest centroid axis: the subtree below is balanced, so the bound holds
            int axis = 0;
            for (int k = 1; k < 3; ++k) if (cbox.hi[k] - cbox.lo[k] > cbox.hi[axis] - cbox.lo[axis]) axis = k;
            mid = (it.begin + it.end) / 2;
            std::nth_element(order.begin() + it.begin, order.begin() + mid, order.begin() + it.end,
                             [&](uint32_t a, uint32_t b) { return cen[(size_t) a * 3 + axis] < cen[(size_t) b * 3 + axis]; });
            nd.a = (int32_t) nodes.size(); nodes.push_back(DNode());
            nd.b = (int32_t) nodes.size(); nodes.push_back(DNode());
            stack.push_back({ it.begin, mid, nd.a, it.depth + 1 });
            stack.push_back({ mid, it.end, nd.b, it.depth + 1 });
        } else if (!leaf) {
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
                stack.push_back({ it.begin, mid, nd.a, it.depth + 1 });
                stack.push_back({ mid, it.end, nd.b, it.depth + 1 });
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

void build_wide_nodes(const std::vector<DNode> &nodes, std::vector<DWNode> &wnodes) {
    wnodes.clear();
    if (nodes.empty() || nodes[0].b < 0) return;
    std::vector<int32_t> windex(nodes.size(), -1);
    int32_t n_inner = 0;
    for (size_t i = 0; i < nodes.size(); ++i) if (nodes[i].b >= 0) windex[i] = n_inner++;   // root (node 0) -> wide node 0
    wnodes.resize((size_t) n_inner);
    auto ref = [&](int32_t child) -> uint32_t {
        const DNode &c = nodes[(size_t) child];
        if (c.b >= 0) return (uint32_t) windex[(size_t) child];
        return 0x80000000u | ((uint32_t) c.a << 12) | ((uint32_t) (-c.b) << 7);      // leaf: first tri, count (<= 16)
    };
    for (size_t i = 0; i < nodes.size(); ++i) {
        if (nodes[i].b < 0) continue;
        DWNode w{};
        const DNode &l = nodes[(size_t) nodes[i].a], &r = nodes[(size_t) nodes[i].b];
        for (int k = 0; k < 3; ++k) { w.lo0[k] = l.lo[k]; w.hi0[k] = l.hi[k]; w.lo1[k] = r.lo[k]; w.hi1[k] = r.hi[k]; }
        w.c0 = ref(nodes[i].a); w.c1 = ref(nodes[i].b);
        wnodes[(size_t) windex[i]] = w;
    }
}

// ---- shadow structure: main BVH + "shell" components -------------------------------------------------------
// A connected component whose every triangle lies in a face plane of the component's own AABB (a fog container
// box, the walls of a room) cannot be hit by a segment whose two end points are strictly inside that AABB, so
// such components get their own small BVH which the shadow query enters only when an end point is within
// `margin` of (or outside) the box.  Everything else goes into the main BVH, whose SAH is then no longer
// dominated by scene-sized triangles.  Exactness: skipped triangles are at least `margin` away from the segment.
void build_shadow_accel(const float *verts, const uint32_t *flags, uint32_t n, ShadowAccel &out, const SubsetBuilder *custom, uint32_t *n_custom) {
    out = ShadowAccel();
    if (n_custom) *n_custom = 0;
    if (n == 0) return;
    // union-find over triangles sharing a vertex position (exact fp32 equality)
    std::vector<uint32_t> parent(n);
    for (uint32_t i = 0; i < n; ++i) parent[i] = i;
    auto find = [&](uint32_t x) { while (parent[x] != x) { parent[x] = parent[parent[x]]; x = parent[x]; } return x; };
    struct VKey { uint32_t x, y, z; uint32_t tri; };
    std::vector<VKey> keys; keys.reserve((size_t) n * 3);
    auto bits = [](float f) { if (f == 0.f) f = 0.f; uint32_t u; std::memcpy(&u, &f, 4); return u; };
    for (uint32_t i = 0; i < n; ++i)
        for (int v = 0; v < 3; ++v) {
            const float *p = verts + (size_t) i * 9 + v * 3;
            keys.push_back({ bits(p[0]), bits(p[1]), bits(p[2]), i });
        }
    std::sort(keys.begin(), keys.end(), [](const VKey &a, const VKey &b) {
        if (a.x != b.x) return a.x < b.x; if (a.y != b.y) return a.y < b.y; if (a.z != b.z) return a.z < b.z; return a.tri < b.tri; });
    for (size_t i = 1; i < keys.size(); ++i)
        if (keys[i].x == keys[i - 1].x && keys[i].y == keys[i - 1].y && keys[i].z == keys[i - 1].z) {
            uint32_t a = find(keys[i].tri), b = find(keys[i - 1].tri);
            if (a != b) parent[std::max(a, b)] = std::min(a, b);
        }
    // component boxes
    std::vector<Box> cbox(n);
    for (uint32_t i = 0; i < n; ++i) { uint32_t r = find(i); const float *p = verts + (size_t) i * 9; cbox[r].grow(p); cbox[r].grow(p + 3); cbox[r].grow(p + 6); }
    std::vector<char> is_shell(n, 0);
    for (uint32_t i = 0; i < n; ++i) if (find(i) == i) is_shell[i] = 1;
    for (uint32_t i = 0; i < n; ++i) {
        uint32_t r = find(i); const Box &b = cbox[r]; const float *p = verts + (size_t) i * 9;
        bool on_face = false;
        for (int k = 0; k < 3 && !on_face; ++k)
            on_face = (p[k] == b.lo[k] && p[3 + k] == b.lo[k] && p[6 + k] == b.lo[k]) || (p[k] == b.hi[k] && p[3 + k] == b.hi[k] && p[6 + k] == b.hi[k]);
        if (!on_face) is_shell[r] = 0;
    }
    std::vector<uint32_t> shell_roots;
    for (uint32_t i = 0; i < n; ++i) if (find(i) == i && is_shell[i] && shell_roots.size() < VRLB_MAX_SHELLS) shell_roots.push_back(i);
    // "solids": the largest remaining components (by box area, at least 12 triangles) get their own BVH + proxies
    std::vector<uint32_t> solid_roots;
    {
        std::vector<uint32_t> ccount(n, 0);
        for (uint32_t i = 0; i < n; ++i) ccount[find(i)]++;
        std::vector<std::pair<float, uint32_t>> cand;
        for (uint32_t i = 0; i < n; ++i) {
            if (find(i) != i || is_shell[i] || ccount[i] < 12) continue;
            const Box &b = cbox[i]; float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
            cand.push_back({ -(dx * dy + dy * dz + dz * dx), i });
        }
        std::sort(cand.begin(), cand.end());
        for (size_t i = 0; i < cand.size() && i < VRLB_MAX_SOLIDS; ++i) solid_roots.push_back(cand[i].second);
    }
    const size_t n_sh = shell_roots.size();
    // subset index: 0 = main, 1 .. n_sh = shells, n_sh + 1 .. = solids
    auto subset_of = [&](uint32_t tri) -> size_t {
        uint32_t r = find(tri);
        for (size_t s = 0; s < shell_roots.size(); ++s) if (shell_roots[s] == r) return 1 + s;
        for (size_t s = 0; s < solid_roots.size(); ++s) if (solid_roots[s] == r) return 1 + n_sh + s;
        return 0;
    };
    // gather subsets
    std::vector<std::vector<uint32_t>> subsets(1 + shell_roots.size() + solid_roots.size());
    for (uint32_t i = 0; i < n; ++i) subsets[subset_of(i)].push_back(i);
    float ext = 0.f;
    for (uint32_t i = 0; i < n * 9; ++i) ext = std::max(ext, std::fabs(verts[i]));
    // Safety margin of the shell boxes and the solids' proxies.  It has to cover the fp32 error of the device's tests and of
    // the fast path's end points (a few 1e-6 * ext) and the graze margins of intersect_tri_graze (2e-4 + 1e-5 * length);
    // every eval with an end point closer than this to a shell goes to the BVH, and importance sampling puts ~1 % of the
    // end points within 1e-3 * ext of a wall (VRLs and camera parts END on walls), so it is not chosen any larger.
    const float margin = 1e-4f * ext + 3e-4f;
    for (size_t s = 0; s < subsets.size(); ++s) {
        const std::vector<uint32_t> &sub = subsets[s];
        uint32_t root_ref = 0x80000000u;          // empty leaf
        if (!sub.empty()) {
            std::vector<float> v((size_t) sub.size() * 9); std::vector<uint32_t> f(sub.size());
            for (size_t i = 0; i < sub.size(); ++i) { std::memcpy(&v[i * 9], verts + (size_t) sub[i] * 9, 36); f[i] = flags ? flags[sub[i]] : 0u; }
            std::vector<DNode> nodes; std::vector<DTri> tris; std::vector<DWNode> w;
            int depth = 0;
            if (custom && (*custom)(v.data(), f.data(), (uint32_t) sub.size(), nodes, tris, depth) && depth + 2 <= kMaxBvhDepth) {
                if (n_custom) ++*n_custom;
            } else {
                build_bvh(v.data(), f.data(), (uint32_t) sub.size(), nodes, tris);
                depth = last_bvh_depth();
            }
            out.max_depth = std::max(out.max_depth, depth);
            build_wide_nodes(nodes, w);
            const uint32_t tri0 = (uint32_t) out.tris.size(), node0 = (uint32_t) out.wnodes.size();
            for (DTri &t : tris) { t.prim = sub[t.prim]; out.tris.push_back(t); }
            auto fix = [&](uint32_t ref) -> uint32_t {
                if (ref & 0x80000000u) return 0x80000000u | ((((ref & 0x7fffffffu) >> 12) + tri0) << 12) | (ref & 0xf80u);
                return ref + node0;
            };
            for (DWNode &wn : w) { wn.c0 = fix(wn.c0); wn.c1 = fix(wn.c1); out.wnodes.push_back(wn); }
            root_ref = w.empty() ? (0x80000000u | (tri0 << 12) | ((uint32_t) tris.size() << 7)) : node0;
        }
        if (s == 0) {
            out.main_root = root_ref; out.has_main = !sub.empty(); out.n_main_tris = (uint32_t) sub.size();
            Box mb; for (uint32_t t : sub) { const float *p = verts + (size_t) t * 9; mb.grow(p); mb.grow(p + 3); mb.grow(p + 6); }
            if (!sub.empty()) for (int k = 0; k < 3; ++k) { out.main_lo[k] = mb.lo[k] - margin; out.main_hi[k] = mb.hi[k] + margin; }
        }
        else if (s <= n_sh) {
            DShell sh{}; const Box &b = cbox[shell_roots[s - 1]];
            for (int k = 0; k < 3; ++k) { sh.lo[k] = b.lo[k] + margin; sh.hi[k] = b.hi[k] - margin; }
            sh.root = root_ref;
            out.shells.push_back(sh);
        } else {
            DSolid so; build_solid_proxies(verts, sub, margin, so);
            so.root = root_ref;
            out.solids.push_back(so);
        }
    }
}

}  // namespace vrlb
