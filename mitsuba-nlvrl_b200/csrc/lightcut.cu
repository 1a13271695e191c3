// lightcut.cu -- the VRL light cut.
//
// Replaces VRLLightCut (src/integrators/vrl/vrl_lightcut.h): the light tree over the VRL set (constructor :196-212 ->
// buildLightTree :515-627, Node::dist :101-109, the merge :111-151), the per-query cut refinement (query :243-393), its
// cluster bound (getClusterUpperBound :429-512 with BoundingBox::getMinDistanceSqr, include/mitsuba/core/bbox.h:335-816,
// and sqrDistanceVRL :772-838), dispatched from VRLMap::query's ELightCutAcceleration branch (vrl_map.h:198-201).
//
// Build.  The reference's clustering is greedy and sequential: every live cluster keeps one queue entry (itself, its
// cheapest partner at the time the entry was made); the cheapest entry is popped and merges if both ends are still live,
// otherwise the live end searches again.  The order of the merges IS the tree, so the sequence is kept: the queue and the
// merge arithmetic (a few flops per merge) run on the host exactly as written there, and the O(N) partner search -- the
// O(N^2) part, 2-3 N searches of N candidates -- runs on the device: one launch per search over the 48-byte cluster
// records resident in L2, min-reduced on (metric bits << 32 | slot) so that the lowest slot wins a tie as in the
// reference's scan, the result handed back through mapped pinned memory (the host spins on a sequence word: no stream
// synchronisation in the loop).  The first N searches are one all-pairs launch.
//
// Query.  One thread per camera segment: the refinement is a priority queue walk whose every step depends on the running
// estimate, so it is kept sequential per query and parallel across queries (32 neighbouring segments per warp walk
// similar cuts).  Each step pops the cluster with the largest bound, evaluates the child that does not share the parent's
// representative (getContrib in the reference's operation order, gather_exact.cuh), scales the parent's estimate for the
// one that does, bounds both.  The heap is 4-ary: its first three levels sit in shared memory, the rest in the thread's own
// run of global memory where four siblings are one 32-byte sector; keys are (bound bits << 32 | slot), payloads (estimate,
// node) sit in slots recycled through a free list.
// PCG32 addressing: the evaluation of a cluster whose representative is VRL j, in query sample s, reads the two uniforms
// at ((s * total + first_index + j) * 2) -- every VRL represents exactly one chain of clusters and only the top of a chain
// is evaluated, so no position is read twice; the jump (multiplier, increment) per VRL is tabulated at build time.
// A query whose heap outgrows its capacity finishes with the exact sum over the VRL set (counted in the stats).
#include <cfloat>
#include <chrono>
#include <cstring>
#include <queue>
#include <vector>

#include "gather_exact.cuh"
#include "lightcut.h"

namespace vrlb {

// ---------------------------------------------------------------------------------------------------------------------
//  Build: partner search
// ---------------------------------------------------------------------------------------------------------------------
struct DLcSlot {             // 48 B: what Node::dist reads of a live cluster
    float lox, loy, loz, fr;
    float hix, hiy, hiz, fg;
    float fb; uint32_t alive; uint32_t pad0, pad1;
};
constexpr unsigned long long LC_KEY_INIT = ((unsigned long long) 0x7f7fffffu) << 32;   // (FLT_MAX, slot 0): "nothing cheaper found"

// Node::dist, vrl_lightcut.h:101-109: squared diagonal of the joint box times the largest channel of the summed flux
__device__ __forceinline__ float lc_metric(const float4 &a0, const float4 &a1, float afb, const float4 &b0, const float4 &b1, float bfb) {
    float ex = fmaxf(a1.x, b1.x) - fminf(a0.x, b0.x);
    float ey = fmaxf(a1.y, b1.y) - fminf(a0.y, b0.y);
    float ez = fmaxf(a1.z, b1.z) - fminf(a0.z, b0.z);
    float total = fmaxf(fmaxf(a0.w + b0.w, a1.w + b1.w), afb + bfb);
    return (ex * ex + ey * ey + ez * ez) * total;
}
__device__ __forceinline__ unsigned long long lc_key(float d, uint32_t j) { return ((unsigned long long) __float_as_uint(d) << 32) | j; }
__device__ __forceinline__ unsigned long long umin64(unsigned long long a, unsigned long long b) { return a < b ? a : b; }

// the first N searches (vrl_lightcut.h:566-578): every cluster against every other
__global__ void __launch_bounds__(256) lc_init_pairs_kernel(const DLcSlot *slots, uint32_t n, unsigned long long *best) {
    __shared__ float4 t0[256], t1[256]; __shared__ float t2[256];
    const uint32_t i = blockIdx.x * 256 + threadIdx.x;
    float4 a0 = make_float4(0, 0, 0, 0), a1 = a0; float afb = 0.f;
    if (i < n) { const float4 *s = reinterpret_cast<const float4 *>(slots + i); a0 = s[0]; a1 = s[1]; afb = s[2].x; }
    unsigned long long key = LC_KEY_INIT;
    for (uint32_t base = 0; base < n; base += 256) {
        const uint32_t j = base + threadIdx.x;
        if (j < n) { const float4 *s = reinterpret_cast<const float4 *>(slots + j); t0[threadIdx.x] = s[0]; t1[threadIdx.x] = s[1]; t2[threadIdx.x] = s[2].x; }
        __syncthreads();
        const uint32_t cnt = min(256u, n - base);
        for (uint32_t k = 0; k < cnt; ++k) {
            if (base + k == i) continue;
            key = umin64(key, lc_key(lc_metric(a0, a1, afb, t0[k], t1[k], t2[k]), base + k));
        }
        __syncthreads();
    }
    if (i < n) best[i] = key;
}

// one search (min_pair, vrl_lightcut.h:538-554) for the cluster `q` sitting in slot qi; the same launch stores q into its
// slot and retires slot `kill` (both are skipped by the search itself).  The last block to finish publishes the result.
__global__ void __launch_bounds__(256) lc_scan_kernel(DLcSlot *slots, uint32_t n, uint32_t qi, DLcSlot q, uint32_t write_q, uint32_t kill,
                                                      unsigned long long *dev_best, uint32_t *dev_ticket,
                                                      volatile unsigned long long *host_out, unsigned long long seq) {
    __shared__ unsigned long long wmin[8];
    const float4 a0 = make_float4(q.lox, q.loy, q.loz, q.fr), a1 = make_float4(q.hix, q.hiy, q.hiz, q.fg);
    unsigned long long key = LC_KEY_INIT;
    for (uint32_t j = blockIdx.x * 256 + threadIdx.x; j < n; j += gridDim.x * 256) {
        if (j == qi || j == kill) continue;
        const float4 *s = reinterpret_cast<const float4 *>(slots + j);
        float4 b2 = s[2];
        if (__float_as_uint(b2.y) == 0u) continue;
        key = umin64(key, lc_key(lc_metric(a0, a1, q.fb, s[0], s[1], b2.x), j));
    }
    for (int o = 16; o > 0; o >>= 1) key = umin64(key, __shfl_xor_sync(0xffffffffu, key, o));
    if ((threadIdx.x & 31) == 0) wmin[threadIdx.x >> 5] = key;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) key = umin64(key, wmin[w]);
        atomicMin(dev_best, key);
        if (blockIdx.x == 0) {
            if (write_q) slots[qi] = q;
            if (kill != 0xffffffffu) slots[kill].alive = 0u;
        }
        __threadfence();
        const uint32_t ticket = atomicAdd(dev_ticket, 1u);
        if (ticket == gridDim.x - 1) {
            const unsigned long long v = atomicExch(dev_best, LC_KEY_INIT);
            *dev_ticket = 0u;
            host_out[0] = v;
            __threadfence_system();
            host_out[1] = seq;
        }
    }
}

// PCG32 jump of 2 * (first_index + j) draws as (multiplier, increment): state' = m * state + c (enoki random.h:265-286)
__global__ void lc_jump_kernel(ulonglong2 *jump, uint32_t n, uint64_t first_index) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    uint64_t delta = (first_index + j) * 2u, cur_mult = PCG32_MULT, cur_plus = PCG32_INC, acc_mult = 1, acc_plus = 0;
    while (delta != 0) {
        if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
        cur_plus = (cur_mult + 1) * cur_plus;
        cur_mult *= cur_mult;
        delta >>= 1;
    }
    jump[j] = make_ulonglong2(acc_mult, acc_plus);
}

// ---------------------------------------------------------------------------------------------------------------------
//  Build: the host side of buildLightTree
// ---------------------------------------------------------------------------------------------------------------------
namespace {

struct HNode {
    float lo[3], hi[3], flux[3], add_flux[3];
    float o[3], d[3], len; uint32_t medium;
    int32_t rep_leaf, child[2], same, parent;
    float max_len; uint32_t n_leaves, first_leaf;
};
inline float hmax3(const float *f) { return std::max(std::max(f[0], f[1]), f[2]); }

// Node(const VRL &), vrl_lightcut.h:56-63
HNode leaf_node(const DVrl &v, int32_t index) {
    HNode n{};
    const float o[3] = { v.ox, v.oy, v.oz }, d[3] = { v.dx, v.dy, v.dz };
    for (int k = 0; k < 3; ++k) {
        float end = o[k] + d[k] * v.length;
        n.lo[k] = std::min(o[k], end); n.hi[k] = std::max(o[k], end);
        n.o[k] = o[k]; n.d[k] = d[k];
    }
    n.flux[0] = v.fr; n.flux[1] = v.fg; n.flux[2] = v.fb;
    n.len = v.length; n.medium = v.medium; n.rep_leaf = index; n.child[0] = n.child[1] = -1; n.same = -1; n.parent = -1;
    n.max_len = v.length; n.n_leaves = 1;
    return n;
}
// Node(Node *n1, Node *n2, Float v, bool), vrl_lightcut.h:111-151
HNode merge_nodes(const HNode &n1, int32_t i1, const HNode &n2, int32_t i2, float v) {
    HNode m{};
    for (int k = 0; k < 3; ++k) { m.lo[k] = std::min(n1.lo[k], n2.lo[k]); m.hi[k] = std::max(n1.hi[k], n2.hi[k]); }
    m.max_len = std::max(n1.max_len, n2.max_len);
    m.child[0] = i1; m.child[1] = i2; m.parent = -1;
    m.n_leaves = n1.n_leaves + n2.n_leaves;
    const float leftWM = hmax3(n1.flux) * n1.len, rightWM = hmax3(n2.flux) * n2.len;
    float ratio = leftWM / (leftWM + rightWM);
    const int index = v < ratio ? 0 : 1;
    const HNode &keep = index == 0 ? n1 : n2, &other = index == 0 ? n2 : n1;
    for (int k = 0; k < 3; ++k) { m.o[k] = keep.o[k]; m.d[k] = keep.d[k]; }
    m.len = keep.len; m.medium = keep.medium; m.rep_leaf = keep.rep_leaf;
    ratio = other.len / keep.len;
    for (int k = 0; k < 3; ++k) { m.add_flux[k] = other.flux[k] * ratio; m.flux[k] = keep.flux[k] + m.add_flux[k]; }
    m.same = index;
    return m;
}
inline DLcSlot slot_of(const HNode &n) {
    DLcSlot s{};
    s.lox = n.lo[0]; s.loy = n.lo[1]; s.loz = n.lo[2]; s.hix = n.hi[0]; s.hiy = n.hi[1]; s.hiz = n.hi[2];
    s.fr = n.flux[0]; s.fg = n.flux[1]; s.fb = n.flux[2]; s.alive = 1u;
    return s;
}
struct HostPcg {             // the sampler handed to VRLMap::build (IndependentSampler: seed(base + offset, default stream))
    Pcg32 r;
    float next_1d() { uint32_t b = (r.next_u32() >> 9) | 0x3f800000u; float f; std::memcpy(&f, &b, 4); return f - 1.0f; }
};

}  // namespace

struct LightcutTree {
    std::vector<HNode> nodes;        // leaves 0..n-1 in VRL order, then the merges in creation order; root = last
    int32_t root = -1;
    DLcNode *d_nodes = nullptr; uint32_t *d_leaf_order = nullptr; ulonglong2 *d_jump = nullptr; float4 *d_leaf_geom = nullptr;
    uint64_t n_vrls = 0, searches = 0, merges = 0;
    float build_ms = 0.f;
};

void lightcut_free(LightcutTree *t) {
    if (!t) return;
    if (t->d_nodes) cudaFree(t->d_nodes);
    if (t->d_leaf_order) cudaFree(t->d_leaf_order);
    if (t->d_jump) cudaFree(t->d_jump);
    if (t->d_leaf_geom) cudaFree(t->d_leaf_geom);
    delete t;
}

#define LC_CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { err = e_; goto done; } } while (0)

cudaError_t lightcut_build(const DVrl *d_vrls, uint64_t n64, uint64_t first_index, uint64_t base_seed, uint64_t seed_offset,
                           cudaStream_t stream, LightcutTree **out, uint64_t *launches) {
    *out = nullptr;
    if (n64 == 0 || n64 > 0x3fffffffull) return cudaErrorInvalidValue;
    const uint32_t n = (uint32_t) n64;
    const auto t_start = std::chrono::steady_clock::now();
    cudaError_t err = cudaSuccess;
    LightcutTree *tree = new LightcutTree();
    std::vector<DVrl> vrls(n);
    std::vector<DLcSlot> slots(n);
    std::vector<unsigned long long> first(n);
    std::vector<int32_t> slot_node(n);
    DLcSlot *d_slots = nullptr; unsigned long long *d_best = nullptr, *d_one = nullptr; uint32_t *d_ticket = nullptr;
    unsigned long long *h_out = nullptr, *h_out_dev = nullptr;
    uint64_t n_launch = 0;
    {
        LC_CU(cudaMemcpyAsync(vrls.data(), d_vrls, (size_t) n * sizeof(DVrl), cudaMemcpyDeviceToHost, stream));
        LC_CU(cudaStreamSynchronize(stream));
        tree->n_vrls = n;
        tree->nodes.reserve(2 * (size_t) n);
        for (uint32_t j = 0; j < n; ++j) { tree->nodes.push_back(leaf_node(vrls[j], (int32_t) j)); slots[j] = slot_of(tree->nodes[j]); slot_node[j] = (int32_t) j; }
        LC_CU(cudaMalloc((void **) &d_slots, (size_t) n * sizeof(DLcSlot)));
        LC_CU(cudaMalloc((void **) &d_best, (size_t) n * sizeof(unsigned long long)));
        LC_CU(cudaMalloc((void **) &d_one, sizeof(unsigned long long)));
        LC_CU(cudaMalloc((void **) &d_ticket, sizeof(uint32_t)));
        LC_CU(cudaHostAlloc((void **) &h_out, 2 * sizeof(unsigned long long), cudaHostAllocMapped));
        LC_CU(cudaHostGetDevicePointer((void **) &h_out_dev, h_out, 0));
        h_out[0] = 0; h_out[1] = 0;
        LC_CU(cudaMemcpyAsync(d_slots, slots.data(), (size_t) n * sizeof(DLcSlot), cudaMemcpyHostToDevice, stream));
        {
            unsigned long long init = LC_KEY_INIT;
            LC_CU(cudaMemcpyAsync(d_one, &init, sizeof init, cudaMemcpyHostToDevice, stream));
            LC_CU(cudaMemsetAsync(d_ticket, 0, sizeof(uint32_t), stream));
        }
        lc_init_pairs_kernel<<<(n + 255) / 256, 256, 0, stream>>>(d_slots, n, d_best);
        ++n_launch;
        LC_CU(cudaGetLastError());
        LC_CU(cudaMemcpyAsync(first.data(), d_best, (size_t) n * sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream));
        LC_CU(cudaStreamSynchronize(stream));
    }
    {
        // buildLightTree, vrl_lightcut.h:515-627
        struct QE { uint32_t id1; int32_t n1; uint32_t id2; int32_t n2; float distance; bool operator>(const QE &o) const { return distance > o.distance; } };
        std::priority_queue<QE, std::vector<QE>, std::greater<QE>> pairs;
        auto unpack = [](unsigned long long key, uint32_t &o, float &d) { o = (uint32_t) (key & 0xffffffffu); uint32_t b = (uint32_t) (key >> 32); std::memcpy(&d, &b, 4); };
        for (uint32_t i = 0; i < n; ++i) {
            uint32_t o; float d; unpack(first[i], o, d);
            pairs.push(QE{ i, slot_node[i], o, n > 1 || o != i ? slot_node[o] : -1, d });
        }
        tree->searches = n;
        std::vector<char> has_parent(2 * (size_t) n, 0);
        const uint32_t grid = std::min<uint32_t>((n + 255) / 256, 296u);
        unsigned long long seq = 0;
        int dev = 0; cudaGetDevice(&dev);
        // search for the cluster `node` living in slot `slot`; optionally store it / retire a slot in the same launch
        auto search = [&](int32_t node, uint32_t slot, bool write, uint32_t kill, uint32_t &o, float &d) -> cudaError_t {
            ++seq;
            lc_scan_kernel<<<grid, 256, 0, stream>>>(d_slots, n, slot, slot_of(tree->nodes[(size_t) node]), write ? 1u : 0u, kill, d_one, d_ticket, h_out_dev, seq);
            ++n_launch; ++tree->searches;
            cudaError_t e = cudaGetLastError();
            if (e != cudaSuccess) return e;
            volatile unsigned long long *flag = h_out + 1;
            uint64_t spins = 0;
            while (*flag != seq) {
                if ((++spins & 0xfffffu) == 0) {                      // every ~1M polls: has the stream died?
                    e = cudaStreamQuery(stream);
                    if (e != cudaSuccess && e != cudaErrorNotReady) return e;
                    if (e == cudaSuccess && *flag != seq) return cudaErrorUnknown;
                }
            }
            unpack(*(volatile unsigned long long *) h_out, o, d);
            return cudaSuccess;
        };
        HostPcg sampler; sampler.r.seed(base_seed + seed_offset);
        uint32_t live = n;
        while (live != 1) {
            QE p = pairs.top(); pairs.pop();
            const bool free1 = !has_parent[(size_t) p.n1], free2 = p.n2 >= 0 && !has_parent[(size_t) p.n2];
            if (free1 && free2) {
                const float v = sampler.next_1d();
                const int32_t id = (int32_t) tree->nodes.size();
                tree->nodes.push_back(merge_nodes(tree->nodes[(size_t) p.n1], p.n1, tree->nodes[(size_t) p.n2], p.n2, v));
                tree->nodes[(size_t) p.n1].parent = id; tree->nodes[(size_t) p.n2].parent = id;
                has_parent[(size_t) p.n1] = has_parent[(size_t) p.n2] = 1;
                slot_node[p.id1] = id; slot_node[p.id2] = -1;
                ++tree->merges; --live;
                uint32_t o; float d;
                LC_CU(search(id, p.id1, true, p.id2, o, d));
                pairs.push(QE{ p.id1, id, o, slot_node[o], d });
            } else {
                if (free1) { uint32_t o; float d; LC_CU(search(p.n1, p.id1, false, 0xffffffffu, o, d)); pairs.push(QE{ p.id1, p.n1, o, slot_node[o], d }); }
                if (free2) { uint32_t o; float d; LC_CU(search(p.n2, p.id2, false, 0xffffffffu, o, d)); pairs.push(QE{ p.id2, p.n2, o, slot_node[o], d }); }
            }
        }
        for (uint32_t i = 0; i < n; ++i) if (slot_node[i] >= 0) tree->root = slot_node[i];
    }
    {
        // flatten: leaves in depth-first order (Node::nodes of a cluster = a contiguous range), device records
        std::vector<uint32_t> leaf_order; leaf_order.reserve(n);
        std::vector<int32_t> stack{ tree->root };
        while (!stack.empty()) {
            const int32_t id = stack.back(); stack.pop_back();
            HNode &nd = tree->nodes[(size_t) id];
            nd.first_leaf = (uint32_t) leaf_order.size();      // valid for a leaf; inner nodes: fixed below
            if (nd.child[0] < 0) leaf_order.push_back((uint32_t) id);
            else { stack.push_back(nd.child[1]); stack.push_back(nd.child[0]); }
        }
        // an inner node's range starts where its leftmost leaf sits: creation order is bottom-up, children precede parents
        for (size_t id = n; id < tree->nodes.size(); ++id) tree->nodes[id].first_leaf = tree->nodes[(size_t) tree->nodes[id].child[0]].first_leaf;
        std::vector<DLcNode> dn(tree->nodes.size());
        for (size_t id = 0; id < dn.size(); ++id) {
            const HNode &h = tree->nodes[id]; DLcNode &q = dn[id];
            q.ox = h.o[0]; q.oy = h.o[1]; q.oz = h.o[2]; q.length = h.len;
            q.dx = h.d[0]; q.dy = h.d[1]; q.dz = h.d[2]; q.medium = h.medium;
            q.fr = h.flux[0]; q.fg = h.flux[1]; q.fb = h.flux[2]; q.rep_leaf = (uint32_t) h.rep_leaf;
            q.lox = h.lo[0]; q.loy = h.lo[1]; q.loz = h.lo[2]; q.max_len = h.max_len;
            q.hix = h.hi[0]; q.hiy = h.hi[1]; q.hiz = h.hi[2]; q.n_leaves = h.n_leaves;
            q.child0 = h.child[0]; q.child1 = h.child[1]; q.same = h.same; q.first_leaf = h.first_leaf;
        }
        LC_CU(cudaMalloc((void **) &tree->d_nodes, dn.size() * sizeof(DLcNode)));
        LC_CU(cudaMalloc((void **) &tree->d_leaf_order, (size_t) n * sizeof(uint32_t)));
        LC_CU(cudaMalloc((void **) &tree->d_jump, (size_t) n * sizeof(ulonglong2)));
        LC_CU(cudaMalloc((void **) &tree->d_leaf_geom, (size_t) n * 2 * sizeof(float4)));
        {   // the VRLs' geometry in leaf order: the exact-distance bound of a small cluster reads a contiguous run
            std::vector<float4> geom((size_t) n * 2);
            for (uint32_t k = 0; k < n; ++k) {
                const HNode &h = tree->nodes[(size_t) leaf_order[k]];
                geom[2 * (size_t) k] = make_float4(h.o[0], h.o[1], h.o[2], h.len);
                geom[2 * (size_t) k + 1] = make_float4(h.d[0], h.d[1], h.d[2], 0.f);
            }
            LC_CU(cudaMemcpyAsync(tree->d_leaf_geom, geom.data(), geom.size() * sizeof(float4), cudaMemcpyHostToDevice, stream));
            LC_CU(cudaStreamSynchronize(stream));
        }
        LC_CU(cudaMemcpyAsync(tree->d_nodes, dn.data(), dn.size() * sizeof(DLcNode), cudaMemcpyHostToDevice, stream));
        LC_CU(cudaMemcpyAsync(tree->d_leaf_order, leaf_order.data(), (size_t) n * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
        lc_jump_kernel<<<(n + 255) / 256, 256, 0, stream>>>(tree->d_jump, n, first_index);
        ++n_launch;
        LC_CU(cudaGetLastError());
        LC_CU(cudaStreamSynchronize(stream));
    }
done:
    if (d_slots) cudaFree(d_slots);
    if (d_best) cudaFree(d_best);
    if (d_one) cudaFree(d_one);
    if (d_ticket) cudaFree(d_ticket);
    if (h_out) cudaFreeHost(h_out);
    if (launches) *launches += n_launch;
    if (err != cudaSuccess) { lightcut_free(tree); return err; }
    tree->build_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t_start).count();
    *out = tree;
    return cudaSuccess;
}

const DLcNode *lightcut_nodes(const LightcutTree *t) { return t->d_nodes; }
const uint32_t *lightcut_leaf_order(const LightcutTree *t) { return t->d_leaf_order; }
const ulonglong2 *lightcut_jump(const LightcutTree *t) { return t->d_jump; }
const float4 *lightcut_leaf_geom(const LightcutTree *t) { return t->d_leaf_geom; }
uint32_t lightcut_root(const LightcutTree *t) { return (uint32_t) t->root; }
uint64_t lightcut_n_vrls(const LightcutTree *t) { return t->n_vrls; }
void lightcut_build_stats(const LightcutTree *t, uint64_t *n_nodes, uint64_t *searches, float *ms) {
    *n_nodes = t->nodes.size(); *searches = t->searches; *ms = t->build_ms;
}
// preorder dump in the oracle's format (tests): f13 = flux, add_flux, lo, hi, max_len; i6 = leaf, rep VRL, same, n_leaves, id, parent
uint64_t lightcut_dump(const LightcutTree *t, float *f13, int32_t *i6, uint64_t cap) {
    std::vector<int32_t> stack{ t->root };
    uint64_t k = 0;
    while (!stack.empty()) {
        const int32_t id = stack.back(); stack.pop_back();
        const HNode &n = t->nodes[(size_t) id];
        if (k < cap) {
            float *f = f13 + 13 * k; int32_t *i = i6 + 6 * k;
            for (int c = 0; c < 3; ++c) { f[c] = n.flux[c]; f[3 + c] = n.add_flux[c]; f[6 + c] = n.lo[c]; f[9 + c] = n.hi[c]; }
            f[12] = n.max_len;
            i[0] = n.child[0] < 0 ? 1 : 0; i[1] = n.rep_leaf; i[2] = n.same; i[3] = (int32_t) n.n_leaves; i[4] = id; i[5] = n.parent;
        }
        ++k;
        if (n.child[0] >= 0) { stack.push_back(n.child[1]); stack.push_back(n.child[0]); }
    }
    return k;
}

// ---------------------------------------------------------------------------------------------------------------------
//  Query: the cluster bound
// ---------------------------------------------------------------------------------------------------------------------
// BoundingBox::getMinDistanceSqr(ray), bbox.h:335-816: the published line / box distance query (reflect the line into the
// positive octant, find the face / edge / corner it passes closest to), with the fork's quirks kept because the bound is
// computed from them: the slab test in front is for the whole LINE (a part whose supporting line meets the box is at
// distance 0), and the line query is handed the box's FULL extents where it expects half extents.
struct LineBox { float sqr, t; };

template <int I0, int I1, int I2>
__device__ inline void lb_face(const float *p, const float *d, const float *pme, const float *e, LineBox &r) {
    float ppe[3] = { 0.f, 0.f, 0.f };
    ppe[I1] = p[I1] + e[I1];
    ppe[I2] = p[I2] + e[I2];
    auto close = [&](float len_sqr, float x1, float x2) {
        float delta = d[I0] * pme[I0] + d[I1] * x1 + d[I2] * x2;
        float param = -delta / len_sqr;
        r.sqr += pme[I0] * pme[I0] + x1 * x1 + x2 * x2 + delta * param;
        r.t = param;
    };
    // the edge of the face running along axis ia (ib = the face's other axis)
    auto edge = [&](bool along_i1, bool only_if_ahead) -> bool {
        const float dia = along_i1 ? d[I1] : d[I2], dib = along_i1 ? d[I2] : d[I1];
        const float ppa = along_i1 ? ppe[I1] : ppe[I2], ppb = along_i1 ? ppe[I2] : ppe[I1];
        const float ea = along_i1 ? e[I1] : e[I2], pma = along_i1 ? pme[I1] : pme[I2];
        float len_sqr = d[I0] * d[I0] + dib * dib;
        float tmp = len_sqr * ppa - dia * (d[I0] * pme[I0] + dib * ppb);
        if (only_if_ahead && !(tmp >= 0.f)) return false;
        float xa;
        if (tmp <= 2.f * len_sqr * ea) { float t = tmp / len_sqr; len_sqr += dia * dia; xa = ppa - t; }
        else { len_sqr += dia * dia; xa = pma; }
        if (along_i1) close(len_sqr, xa, ppe[I2]); else close(len_sqr, ppe[I1], xa);
        return true;
    };
    const bool in1 = d[I0] * ppe[I1] >= d[I1] * pme[I0];
    if (in1) {
        if (d[I0] * ppe[I2] >= d[I2] * pme[I0]) { float inv = 1.f / d[I0]; r.t = -pme[I0] * inv; return; }
        edge(true, false);
        return;
    }
    if (d[I0] * ppe[I2] >= d[I2] * pme[I0]) { edge(false, false); return; }
    if (edge(true, true)) return;
    if (edge(false, true)) return;
    float len_sqr = d[I0] * d[I0] + d[I1] * d[I1];
    len_sqr += d[I2] * d[I2];
    close(len_sqr, ppe[I1], ppe[I2]);
}
__device__ inline void lb_clamp_axis(float p, float e, LineBox &r) {
    if (p < -e) { float delta = p + e; r.sqr += delta * delta; }
    else if (p > e) { float delta = p - e; r.sqr += delta * delta; }
}
template <int I0, int I1, int I2>
__device__ inline void lb_one_zero(const float *p, const float *d, const float *e, LineBox &r) {
    float pme0 = p[I0] - e[I0], pme1 = p[I1] - e[I1];
    float prod0 = d[I1] * pme0, prod1 = d[I0] * pme1;
    if (prod0 >= prod1) {
        float ppe1 = p[I1] + e[I1];
        float delta = prod0 - d[I0] * ppe1;
        if (delta >= 0.f) {
            float inv_l = 1.f / (d[I0] * d[I0] + d[I1] * d[I1]);
            r.sqr += delta * delta * inv_l;
            r.t = -(d[I0] * pme0 + d[I1] * ppe1) * inv_l;
        } else { float inv = 1.f / d[I0]; r.t = -pme0 * inv; }
    } else {
        float ppe0 = p[I0] + e[I0];
        float delta = prod1 - d[I1] * ppe0;
        if (delta >= 0.f) {
            float inv_l = 1.f / (d[I0] * d[I0] + d[I1] * d[I1]);
            r.sqr += delta * delta * inv_l;
            r.t = -(d[I0] * ppe0 + d[I1] * pme1) * inv_l;
        } else { float inv = 1.f / d[I1]; r.t = -pme1 * inv; }
    }
    lb_clamp_axis(p[I2], e[I2], r);
}
template <int I0, int I1, int I2>
__device__ inline void lb_two_zeros(const float *p, const float *d, const float *e, LineBox &r) {
    r.t = (e[I0] - p[I0]) / d[I0];
    lb_clamp_axis(p[I1], e[I1], r);
    lb_clamp_axis(p[I2], e[I2], r);
}
__device__ inline float box_point_sqr(V3 lo, V3 hi, V3 p) {                   // bbox.h:221-224
    V3 a((p.x < lo.x) ? lo.x - p.x : 0.f, (p.y < lo.y) ? lo.y - p.y : 0.f, (p.z < lo.z) ? lo.z - p.z : 0.f);
    V3 b((p.x > hi.x) ? p.x - hi.x : 0.f, (p.y > hi.y) ? p.y - hi.y : 0.f, (p.z > hi.z) ? p.z - hi.z : 0.f);
    return squared_norm(a + b);
}
__device__ __noinline__ float box_ray_sqr(V3 lo, V3 hi, V3 ro, V3 rd, float rmaxt) {
    {   // the slab test on the line, bbox.h:302-325 (:806)
        float rx = 1.0f / rd.x, ry = 1.0f / rd.y, rz = 1.0f / rd.z;
        float ax = (lo.x - ro.x) * rx, bx = (hi.x - ro.x) * rx;
        float ay = (lo.y - ro.y) * ry, by = (hi.y - ro.y) * ry;
        float az = (lo.z - ro.z) * rz, bz = (hi.z - ro.z) * rz;
        // std::min / std::max semantics of the oracle (NaN from 0 * inf keeps the first operand)
        auto mn = [](float a, float b) { return b < a ? b : a; };
        auto mx = [](float a, float b) { return a < b ? b : a; };
        float mint = mx(mx(mn(ax, bx), mn(ay, by)), mn(az, bz));
        float maxt = mn(mn(mx(ax, bx), mx(ay, by)), mx(az, bz));
        if (maxt >= mint) return 0.0f;
    }
    V3 c = (hi + lo) * 0.5f, ext = hi - lo;
    V3 pt = ro - c;
    float p[3] = { pt.x, pt.y, pt.z }, d[3] = { rd.x, rd.y, rd.z };
    const float e[3] = { ext.x, ext.y, ext.z };
    int positive = 0;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        if (d[i] < 0.f) { p[i] = -p[i]; d[i] = -d[i]; }
        if (d[i] > 0.f) positive |= 1 << i;
    }
    LineBox r; r.sqr = 0.f; r.t = 0.f;
    if (positive == 7) {
        const float pme[3] = { p[0] - e[0], p[1] - e[1], p[2] - e[2] };
        if (d[1] * pme[0] >= d[0] * pme[1]) {
            if (d[2] * pme[0] >= d[0] * pme[2]) lb_face<0, 1, 2>(p, d, pme, e, r);
            else lb_face<2, 0, 1>(p, d, pme, e, r);
        } else {
            if (d[2] * pme[1] >= d[1] * pme[2]) lb_face<1, 2, 0>(p, d, pme, e, r);
            else lb_face<2, 0, 1>(p, d, pme, e, r);
        }
    }
    else if (positive == 3) lb_one_zero<0, 1, 2>(p, d, e, r);
    else if (positive == 5) lb_one_zero<0, 2, 1>(p, d, e, r);
    else if (positive == 6) lb_one_zero<1, 2, 0>(p, d, e, r);
    else if (positive == 1) lb_two_zeros<0, 1, 2>(p, d, e, r);
    else if (positive == 2) lb_two_zeros<1, 0, 2>(p, d, e, r);
    else if (positive == 4) lb_two_zeros<2, 0, 1>(p, d, e, r);
    else { lb_clamp_axis(p[0], e[0], r); lb_clamp_axis(p[1], e[1], r); lb_clamp_axis(p[2], e[2], r); }
    if (r.t < 0.f) return box_point_sqr(lo, hi, ro);
    if (r.t > rmaxt) return box_point_sqr(lo, hi, ro + rd * rmaxt);
    return r.sqr;
}

// VRLLightCut::sqrDistanceVRL, vrl_lightcut.h:772-838 (the clamped closest points of two segments; thresholds compared
// in double as the reference's `< 0.0001` literals are)
__device__ __noinline__ float seg_seg_sqr(V3 ro, V3 rd, float rmaxt, V3 vo, V3 vd, float vlen) {
    V3 u = vd * vlen, v = rd * rmaxt, w = vo - ro;
    float a = dot(u, u), b = dot(u, v), c = dot(v, v), d = dot(u, w), e = dot(v, w);
    float D = a * c - b * b;
    float sN, sD = D, tN, tD = D;
    if ((double) D < 0.0001) { sN = 0.f; sD = 1.f; tN = e; tD = c; }
    else {
        sN = b * e - c * d;
        tN = a * e - b * d;
        if (sN < 0.f) { sN = 0.f; tN = e; tD = c; }
        else if (sN > sD) { sN = sD; tN = e + b; tD = c; }
    }
    if (tN < 0.f) {
        tN = 0.f;
        if (-d < 0.f) sN = 0.f;
        else if (-d > a) sN = sD;
        else { sN = -d; sD = a; }
    } else if (tN > tD) {
        tN = tD;
        if ((-d + b) < 0.f) sN = 0.f;
        else if ((-d + b) > a) sN = sD;
        else { sN = -d + b; sD = a; }
    }
    float sc = ((double) fabsf(sN) < 0.0001) ? 0.f : sN / sD;
    float tc = ((double) fabsf(tN) < 0.0001) ? 0.f : tN / tD;
    V3 dP = w + (u * sc) - (v * tc);
    return squared_norm(dP);
}

// getClusterUpperBound, vrl_lightcut.h:429-512
__device__ __noinline__ float lc_bound(const LightcutParams &p, const DLcNode *nd, const SegRegs &ray) {
    const float4 *row = reinterpret_cast<const float4 *>(nd);
    const float4 r3 = __ldg(row + 3), r4 = __ldg(row + 4);
    const uint32_t n_leaves = __float_as_uint(r4.w);
    float best;
    if ((int) n_leaves > p.better_dist) {
        best = box_ray_sqr(V3(r3.x, r3.y, r3.z), V3(r4.x, r4.y, r4.z), ray.o, ray.d, ray.maxt);
    } else {
        const uint32_t first = __ldg(&nd->first_leaf);
        best = FLT_MAX;
        const float4 *lf = p.leaf_geom + 2 * (size_t) first;
        for (uint32_t k0 = 0; k0 < n_leaves; k0 += 4) {                          // four records in flight per round
            float4 a[4], b[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) if (k0 + k < n_leaves) { a[k] = __ldg(lf + 2 * (k0 + k)); b[k] = __ldg(lf + 2 * (k0 + k) + 1); }
#pragma unroll
            for (int k = 0; k < 4; ++k) if (k0 + k < n_leaves) best = fminf(best, seg_seg_sqr(ray.o, ray.d, ray.maxt, V3(a[k].x, a[k].y, a[k].z), V3(b[k].x, b[k].y, b[k].z), a[k].w));
        }
    }
    const float min_length = safe_sqrt(best);
    if ((double) min_length < 0.0001) return FLT_MAX;
    const float4 r1 = __ldg(row + 1), r2 = __ldg(row + 2);
    const DMedium &m = p.scene.media[__float_as_uint(r1.w)];
    const float tr = homo_transmittance(m.max_density, min_length);
    const double k = (m.phase == 0 ? 1.0 : 2.0) / (16 * 3.14159265358979323846 * 3.14159265358979323846);
    float mat[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) { float s = 1.f * m.sigma_s[c]; s = s * m.sigma_s[c]; mat[c] = (float) ((double) s * k); }
    const float invVRLPDF = (float) (2.0 * atan((double) ray.maxt * 0.5 / (double) min_length) / (double) min_length);
    const float invPdf = r3.w * invVRLPDF;                                   // invRayPDF = maxVRLLenght
    const float tx = r2.x * invPdf * mat[0] * tr, ty = r2.y * invPdf * mat[1] * tr, tz = r2.z * invPdf * mat[2] * tr;
    return fmaxf(fmaxf(tx, ty), tz);
}

// ---------------------------------------------------------------------------------------------------------------------
//  Query: VRLLightCut::query, vrl_lightcut.h:243-393 (stochastic = false), wrapped as VRLMap::query does (vrl_map.h:143, 198-205)
// ---------------------------------------------------------------------------------------------------------------------
constexpr int LC_THREADS = 128;
constexpr int LC_BLOCKS_PER_SM = 5;      // 640 resident queries per SM (<= 102 registers)
constexpr uint32_t LC_SMEM_TOP = 21;     // levels 0-2 of the 4-ary heap live in shared memory (21 KB per block)

// getContrib of a cluster's representative carrying the cluster's flux, on the uniforms addressed by the representative's VRL
template <bool UNIFORM, bool VIS>
__device__ __noinline__ void lc_eval(const LightcutParams &p, const DLcNode *nd, const SegRegs &ray, uint64_t base_state, float *est) {
    const float4 *row = reinterpret_cast<const float4 *>(nd);
    const float4 a = __ldg(row), b = __ldg(row + 1), c = __ldg(row + 2);
    const ulonglong2 j = __ldg(p.jump + __float_as_uint(c.w));
    Pcg32 r; r.state = j.x * base_state + j.y;
    const float u1 = r.next_f32(), u2 = r.next_f32();
    const DMedium &m = p.scene.media[__float_as_uint(b.w)];
    const float f = eval_contrib<UNIFORM, VIS, false, true>(p.scene, ray, V3(a.x, a.y, a.z), V3(b.x, b.y, b.z), a.w,
                                                      m.max_density, m.phase, m.g, u1, u2, &m);
    est[0] = c.x * f * m.sigma_s[0] * m.sigma_s[0];
    est[1] = c.y * f * m.sigma_s[1] * m.sigma_s[1];
    est[2] = c.z * f * m.sigma_s[2] * m.sigma_s[2];
}
// the exact sum over the set, as the no-acceleration branch (vrl_map.h:151-169) with its stream addressing
template <bool UNIFORM, bool VIS>
__device__ __noinline__ void lc_exact_sum(const LightcutParams &p, const SegRegs &ray, uint64_t base_state, float *Li) {
    Pcg32 r; r.state = base_state;
    r.advance(p.first_index * 2u);
    Li[0] = Li[1] = Li[2] = 0.f;
    for (uint64_t j = 0; j < p.n_vrls; ++j) {
        const float4 *vp = reinterpret_cast<const float4 *>(p.vrls + j);
        const float4 a = __ldg(vp), b = __ldg(vp + 1), c = __ldg(vp + 2);
        const DMedium &m = p.scene.media[__float_as_uint(b.w)];
        const float u1 = r.next_f32(), u2 = r.next_f32();
        const float f = eval_contrib<UNIFORM, VIS, false, true>(p.scene, ray, V3(a.x, a.y, a.z), V3(b.x, b.y, b.z), a.w,
                                                          m.max_density, m.phase, m.g, u1, u2, &m);
        const float cr = c.x * f * m.sigma_s[0] * m.sigma_s[0], cg = c.y * f * m.sigma_s[1] * m.sigma_s[1], cb = c.z * f * m.sigma_s[2] * m.sigma_s[2];
        if (isfinite(cr) && isfinite(cg) && isfinite(cb)) { Li[0] += cr; Li[1] += cg; Li[2] += cb; }
    }
}

// Per-thread refinement heap: a 4-ary max-heap of keys (bound bits << 32 | slot).  Entry i's children are 4i+1 .. 4i+4;
// entries 0-20 (three levels) sit in shared memory, the rest in the thread's own contiguous run of global memory where every
// group of four siblings is one aligned 32-byte sector (entry i >= 21 at word i - 21: 4i + 1 - 21 is a multiple of 4).
struct LcHeap {
    unsigned long long *g;       // thread's global run
    unsigned long long *s;       // thread's column of the block's shared array (stride LC_THREADS)
    uint32_t n;
    __device__ __forceinline__ unsigned long long load(uint32_t i) const { return i < LC_SMEM_TOP ? s[i * LC_THREADS] : g[i - LC_SMEM_TOP]; }
    __device__ __forceinline__ void store(uint32_t i, unsigned long long v) { if (i < LC_SMEM_TOP) s[i * LC_THREADS] = v; else g[i - LC_SMEM_TOP] = v; }
    __device__ inline void push(unsigned long long key) {
        uint32_t i = n++;
        while (i > 0) {
            const uint32_t parent = (i - 1) >> 2;
            const unsigned long long pk = load(parent);
            if (pk >= key) break;
            store(i, pk);
            i = parent;
        }
        store(i, key);
    }
    __device__ inline unsigned long long pop() {
        const unsigned long long top = load(0);
        const unsigned long long last = load(--n);
        uint32_t i = 0;
        while (true) {
            const uint32_t c0 = 4 * i + 1;
            if (c0 >= n) break;
            unsigned long long k0, k1, k2, k3;
            if (c0 < LC_SMEM_TOP) { k0 = s[c0 * LC_THREADS]; k1 = s[(c0 + 1) * LC_THREADS]; k2 = s[(c0 + 2) * LC_THREADS]; k3 = s[(c0 + 3) * LC_THREADS]; }
            else {
                const ulonglong2 *q = reinterpret_cast<const ulonglong2 *>(g + (c0 - LC_SMEM_TOP));
                const ulonglong2 lo = q[0], hi = q[1];
                k0 = lo.x; k1 = lo.y; k2 = hi.x; k3 = hi.y;
            }
            // entries at or beyond n are stale: masked to 0, which no live key is below
            if (c0 + 1 >= n) k1 = 0;
            if (c0 + 2 >= n) k2 = 0;
            if (c0 + 3 >= n) k3 = 0;
            unsigned long long mk = k0; uint32_t mi = c0;
            if (k1 > mk) { mk = k1; mi = c0 + 1; }
            if (k2 > mk) { mk = k2; mi = c0 + 2; }
            if (k3 > mk) { mk = k3; mi = c0 + 3; }
            if (mk <= last) break;
            store(i, mk);
            i = mi;
        }
        if (n) store(i, last);
        return top;
    }
};

template <bool UNIFORM, bool VIS>
__global__ void __launch_bounds__(LC_THREADS, LC_BLOCKS_PER_SM)
lightcut_gather_kernel(const __grid_constant__ LightcutParams p) {
    __shared__ unsigned long long top_keys[LC_SMEM_TOP * LC_THREADS];
    const uint32_t T = gridDim.x * LC_THREADS, t = blockIdx.x * LC_THREADS + threadIdx.x;
    LcHeap heap; heap.g = p.heap_key + (size_t) t * p.cap; heap.s = top_keys + threadIdx.x; heap.n = 0;
    float4 *sl = p.slot_est + (size_t) t * p.cap;    // payload slots: estimate + node (while live) / next free slot
    unsigned long long n_evals = 0, n_pops = 0, n_over = 0, n_queries = 0; uint32_t max_heap = 0;

    for (uint64_t seg_i = t; seg_i < p.n_segs; seg_i += T) {
        SegRegs ray; uint64_t rng0; uint32_t pixel; float w[3];
        {
            const float4 *sp = reinterpret_cast<const float4 *>(p.segs + seg_i);
            float4 r0 = __ldg(sp), r1 = __ldg(sp + 1), r2 = __ldg(sp + 2);
            ulonglong2 r3 = __ldg(reinterpret_cast<const ulonglong2 *>(sp + 3));
            ray.o = V3(r0.x, r0.y, r0.z); ray.maxt = r0.w;
            ray.d = V3(r1.x, r1.y, r1.z); pixel = __float_as_uint(r1.w);
            w[0] = r2.x; w[1] = r2.y; w[2] = r2.z;
            rng0 = r3.x;
        }
        float tot[3] = { 0.f, 0.f, 0.f };
        uint32_t q_evals = 0, q_pops = 0, q_heap = 0; bool q_over = false;
        for (uint32_t s = 0; s < p.q; ++s) {
            Pcg32 base; base.state = rng0;
            if (s) base.advance((uint64_t) s * p.total * 2u);
            heap.n = 0;
            uint32_t bump = 0; int32_t free_head = -1;
            // payload slot + key; false when the thread's run is full
            auto push = [&](float bound, uint32_t node, const float *est) -> bool {
                int32_t slot;
                if (free_head >= 0) { slot = free_head; free_head = __float_as_int(sl[slot].w); }
                else if (bump < p.cap) slot = (int32_t) bump++;
                else return false;
                sl[slot] = make_float4(est[0], est[1], est[2], __uint_as_float(node));
                heap.push(((unsigned long long) __float_as_uint(bound) << 32) | (uint32_t) slot);
                return true;
            };
            float Li[3];
            bool overflow = false;
            {
                const DLcNode *root = p.nodes + p.root;
                const float b = lc_bound(p, root, ray);
                float e[3]; lc_eval<UNIFORM, VIS>(p, root, ray, base.state, e);
                push(b, p.root, e);
                Li[0] = e[0]; Li[1] = e[1]; Li[2] = e[2];
            }
            while (heap.n > 0) {
                const unsigned long long top = heap.pop();
                const float bound = __uint_as_float((uint32_t) (top >> 32));
                const int32_t slot = (int32_t) (top & 0xffffffffu);
                const float4 pay = sl[slot];
                sl[slot].w = __int_as_float(free_head); free_head = slot;
                const float thr = fmaxf(fmaxf(Li[0] * p.err, Li[1] * p.err), Li[2] * p.err);
                if (bound < thr) break;
                ++q_pops;
                Li[0] = Li[0] - pay.x; Li[1] = Li[1] - pay.y; Li[2] = Li[2] - pay.z;
                const DLcNode *cl = p.nodes + __float_as_uint(pay.w);
                const float4 c2 = __ldg(reinterpret_cast<const float4 *>(cl) + 2);
                const int4 topo = __ldg(reinterpret_cast<const int4 *>(cl) + 5);       // child0, child1, same, first_leaf
                if (topo.x < 0) continue;                                             // a one-VRL tree: the root has no children (:309-310)
                const int same = topo.z;
                const DLcNode *ch_same = p.nodes + (same == 0 ? topo.x : topo.y), *ch_fresh = p.nodes + (same == 0 ? topo.y : topo.x);
                float es[3], ef[3];
                {   // the child that keeps the representative: the parent's estimate scaled by the flux ratio (:320-339)
                    const float4 f = __ldg(reinterpret_cast<const float4 *>(ch_same) + 2);
                    float ex = f.x, ey = f.y, ez = f.z;
                    if (c2.x > 0.0f) ex /= c2.x;
                    if (c2.y > 0.0f) ey /= c2.y;
                    if (c2.z > 0.0f) ez /= c2.z;
                    ex *= pay.x; ey *= pay.y; ez *= pay.z;
                    if (isnan(ex) || isnan(ey) || isnan(ez)) { ex = ey = ez = 0.f; }
                    es[0] = ex; es[1] = ey; es[2] = ez;
                }
                lc_eval<UNIFORM, VIS>(p, ch_fresh, ray, base.state, ef);
                ++q_evals;
                // children in the reference's order (0 then 1): Li += estimate; an inner child enters the heap with its bound
#pragma unroll 1
                for (int c = 0; c < 2; ++c) {
                    const bool is_same = c == same;
                    const float *e = is_same ? es : ef;
                    const DLcNode *ch = is_same ? ch_same : ch_fresh;
                    Li[0] = Li[0] + e[0]; Li[1] = Li[1] + e[1]; Li[2] = Li[2] + e[2];
                    if (__ldg(&ch->child0) >= 0) {
                        if (!push(lc_bound(p, ch, ray), (uint32_t) (c == 0 ? topo.x : topo.y), e)) { overflow = true; break; }
                    }
                }
                if (overflow) break;
                q_heap = max(q_heap, heap.n);
            }
            if (overflow) {
                q_over = true;
                lc_exact_sum<UNIFORM, VIS>(p, ray, base.state, Li);
                q_evals += (uint32_t) p.n_vrls;
            }
            tot[0] = tot[0] + fmaxf(Li[0], 0.f) * p.scale;
            tot[1] = tot[1] + fmaxf(Li[1], 0.f) * p.scale;
            tot[2] = tot[2] + fmaxf(Li[2], 0.f) * p.scale;
        }
        if (p.q > 1) { tot[0] = tot[0] / (float) p.q; tot[1] = tot[1] / (float) p.q; tot[2] = tot[2] / (float) p.q; }
        if (p.seg_out) { p.seg_out[seg_i * 3 + 0] = tot[0]; p.seg_out[seg_i * 3 + 1] = tot[1]; p.seg_out[seg_i * 3 + 2] = tot[2]; }
        if (p.seg_stats) { p.seg_stats[seg_i * 4 + 0] = (float) q_evals; p.seg_stats[seg_i * 4 + 1] = (float) q_pops; p.seg_stats[seg_i * 4 + 2] = (float) q_heap; p.seg_stats[seg_i * 4 + 3] = q_over ? 1.f : 0.f; }
        if (p.film_rgb) {
            atomicAdd(p.film_rgb + (size_t) pixel * 3 + 0, tot[0] * w[0]);
            atomicAdd(p.film_rgb + (size_t) pixel * 3 + 1, tot[1] * w[1]);
            atomicAdd(p.film_rgb + (size_t) pixel * 3 + 2, tot[2] * w[2]);
        }
        n_evals += q_evals; n_pops += q_pops; n_over += q_over ? 1 : 0; ++n_queries; max_heap = max(max_heap, q_heap);
    }
    // stats: queries, fresh evaluations, refinements, overflowed queries, largest heap
    for (int o = 16; o > 0; o >>= 1) {
        n_evals += __shfl_xor_sync(0xffffffffu, n_evals, o); n_pops += __shfl_xor_sync(0xffffffffu, n_pops, o);
        n_over += __shfl_xor_sync(0xffffffffu, n_over, o); n_queries += __shfl_xor_sync(0xffffffffu, n_queries, o);
        max_heap = max(max_heap, __shfl_xor_sync(0xffffffffu, max_heap, o));
    }
    if ((threadIdx.x & 31) == 0 && p.stats) {
        atomicAdd(p.stats + 0, n_queries); atomicAdd(p.stats + 1, n_evals); atomicAdd(p.stats + 2, n_pops); atomicAdd(p.stats + 3, n_over);
        atomicMax(p.stats + 4, (unsigned long long) max_heap);
    }
}

uint32_t lightcut_grid(uint64_t n_segs, uint32_t cap, int n_sms, size_t budget_bytes) {
    uint32_t per_sm = LC_BLOCKS_PER_SM;
    while (per_sm > 1 && (size_t) n_sms * per_sm * LC_THREADS * cap * 24ull > budget_bytes) --per_sm;
    uint64_t want = (n_segs + LC_THREADS - 1) / LC_THREADS;
    return (uint32_t) std::max<uint64_t>(1, std::min<uint64_t>(want, (uint64_t) n_sms * per_sm));
}

cudaError_t launch_lightcut_gather(const LightcutParams &p, uint32_t grid, bool uniform, bool vis, cudaStream_t stream) {
    if (p.n_segs == 0 || p.n_vrls == 0) return cudaSuccess;
    if (uniform) { if (vis) lightcut_gather_kernel<true, true><<<grid, LC_THREADS, 0, stream>>>(p); else lightcut_gather_kernel<true, false><<<grid, LC_THREADS, 0, stream>>>(p); }
    else { if (vis) lightcut_gather_kernel<false, true><<<grid, LC_THREADS, 0, stream>>>(p); else lightcut_gather_kernel<false, false><<<grid, LC_THREADS, 0, stream>>>(p); }
    return cudaGetLastError();
}

}  // namespace vrlb
