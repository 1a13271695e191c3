// bvh_lbvh.cu -- the closest-hit BVH built on the device (first slice).
//
// Stands where the reference builds its ShapeKDTree on the host (include/mitsuba/render/kdtree.h:2069-2172 is what
// Scene::ray_intersect walks).  Linear BVH after Karras 2012 ("Maximizing parallelism in the construction of BVHs,
// octrees and k-d trees"): 30-bit Morton codes of the triangle centroids, made distinct by the triangle index in the low
// word; one radix sort; every inner node finds its key range and split with count-leading-zeros searches, all nodes in
// parallel; boxes are fitted bottom-up, the second thread to reach a node fits it.  Output is the layout closest_hit
// already walks (vrlb_device.cuh): DNode[2n - 1] -- inner node i at index i (root 0), the leaf of the k-th sorted
// triangle at n - 1 + k with one triangle -- and DTri[n] in sorted order, built with the host build's arithmetic
// (e1 = p1 - p0, e2 = p2 - p0 rounded once; boxes padded by the same margin).  The closest hit does not depend on the
// tree (ties go to the lowest original triangle index), so camera segments and light paths are identical to the host
// build's: that is the test (tests/test_bvh_gpu.py).  Opt-in (vrlb_set_bvh_builder): the binned-SAH host build stays the
// default -- its trees are shallower (depth bound 28) and cheaper to walk, and the shadow BVHs with their proxies are
// still built there.  A tree deeper than the traversal stack allows is refused and the caller falls back.
#include <cub/cub.cuh>

#include "vrlb_device.cuh"
#include "bvh_build.h"

namespace vrlb {

namespace {

__device__ __forceinline__ uint32_t expand10(uint32_t v) {       // 10 bits -> every third bit
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

__global__ void lbvh_keys_kernel(const float *verts, uint32_t n, float3 lo, float3 inv_ext, unsigned long long *keys) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *p = verts + (size_t) i * 9;
    float c[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) c[k] = (p[k] + p[3 + k] + p[6 + k]) / 3.0f;
    const float x = fminf(fmaxf((c[0] - lo.x) * inv_ext.x, 0.f), 1.f), y = fminf(fmaxf((c[1] - lo.y) * inv_ext.y, 0.f), 1.f),
                z = fminf(fmaxf((c[2] - lo.z) * inv_ext.z, 0.f), 1.f);
    const uint32_t xi = min((uint32_t) (x * 1024.f), 1023u), yi = min((uint32_t) (y * 1024.f), 1023u), zi = min((uint32_t) (z * 1024.f), 1023u);
    const uint32_t morton = (expand10(xi) << 2) | (expand10(yi) << 1) | expand10(zi);
    keys[i] = ((unsigned long long) morton << 32) | i;
}

// sorted triangle k: its DTri record and its leaf node
__global__ void lbvh_leaves_kernel(const float *verts, const uint32_t *flags, const unsigned long long *keys, uint32_t n, float pad,
                                   DTri *tris, DNode *nodes) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint32_t prim = (uint32_t) (keys[k] & 0xffffffffu);
    const float *p = verts + (size_t) prim * 9;
    DTri t; DNode nd;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        t.p0[a] = p[a]; t.e1[a] = __fsub_rn(p[3 + a], p[a]); t.e2[a] = __fsub_rn(p[6 + a], p[a]);
        nd.lo[a] = fminf(p[a], fminf(p[3 + a], p[6 + a])) - pad;
        nd.hi[a] = fmaxf(p[a], fmaxf(p[3 + a], p[6 + a])) + pad;
    }
    t.prim = prim; t.flags = flags ? flags[prim] : 0u; t.pad = 0u;
    tris[k] = t;
    nd.a = (int32_t) k; nd.b = -1;
    nodes[(size_t) (n - 1) + k] = nd;
}

__device__ __forceinline__ int lbvh_delta(const unsigned long long *keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    return __clzll((long long) (keys[i] ^ keys[j]));              // keys are distinct: never 64
}

// Karras 2012, section 4: range and split of inner node i
__global__ void lbvh_hierarchy_kernel(const unsigned long long *keys, int n, DNode *nodes, int32_t *parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (lbvh_delta(keys, n, i, i + 1) - lbvh_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int delta_min = lbvh_delta(keys, n, i, i - d);
    int l_max = 2;
    while (lbvh_delta(keys, n, i, i + l_max * d) > delta_min) l_max *= 2;
    int l = 0;
    for (int t = l_max / 2; t >= 1; t /= 2)
        if (lbvh_delta(keys, n, i, i + (l + t) * d) > delta_min) l += t;
    const int j = i + l * d;
    const int delta_node = lbvh_delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) / 2; ; t = (t + 1) / 2) {
        if (lbvh_delta(keys, n, i, i + (s + t) * d) > delta_node) s += t;
        if (t == 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int left = (min(i, j) == gamma) ? (n - 1) + gamma : gamma;
    const int right = (max(i, j) == gamma + 1) ? (n - 1) + gamma + 1 : gamma + 1;
    nodes[i].a = left; nodes[i].b = right;
    parent[left] = i; parent[right] = i;
    if (i == 0) parent[0] = -1;
}

// bottom-up fit: the second arrival at a node owns it
__global__ void lbvh_fit_kernel(DNode *nodes, const int32_t *parent, uint32_t *arrived, int32_t *depth, int n, int32_t *max_depth) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    int cur = parent[(n - 1) + k];
    while (cur >= 0) {
        if (atomicAdd(&arrived[cur], 1u) == 0u) return;          // the sibling subtree is not done: its thread will carry on
        __threadfence();
        const int a = nodes[cur].a, b = nodes[cur].b;
        const volatile DNode *ca = nodes + a, *cb = nodes + b;
#pragma unroll
        for (int x = 0; x < 3; ++x) { nodes[cur].lo[x] = fminf(ca->lo[x], cb->lo[x]); nodes[cur].hi[x] = fmaxf(ca->hi[x], cb->hi[x]); }
        const int da = a >= n - 1 ? 0 : ((const volatile int32_t *) depth)[a], db = b >= n - 1 ? 0 : ((const volatile int32_t *) depth)[b];
        depth[cur] = 1 + max(da, db);
        if (cur == 0) *max_depth = depth[0];
        __threadfence();
        cur = parent[cur];
    }
}

}  // namespace

cudaError_t build_bvh_device(const float *d_verts, const uint32_t *d_flags, uint32_t n, float pad, const float lo[3], const float hi[3],
                             DNode *d_nodes, DTri *d_tris, int *max_depth, uint64_t *launches, cudaStream_t stream) {
    *max_depth = 0;
    if (n == 0) return cudaSuccess;
    cudaError_t err = cudaSuccess;
    unsigned long long *keys = nullptr, *keys_sorted = nullptr; void *tmp = nullptr; size_t tmp_bytes = 0;
    int32_t *parent = nullptr, *depth = nullptr, *d_max = nullptr; uint32_t *arrived = nullptr;
    const uint32_t grid = (n + 255) / 256;
#define LB_CU(call) do { err = (call); if (err != cudaSuccess) goto done; } while (0)
    LB_CU(cudaMalloc((void **) &keys, (size_t) n * 8)); LB_CU(cudaMalloc((void **) &keys_sorted, (size_t) n * 8));
    LB_CU(cudaMalloc((void **) &parent, (size_t) (2 * n) * 4)); LB_CU(cudaMalloc((void **) &depth, (size_t) n * 4));
    LB_CU(cudaMalloc((void **) &arrived, (size_t) n * 4)); LB_CU(cudaMalloc((void **) &d_max, 4));
    LB_CU(cudaMemsetAsync(arrived, 0, (size_t) n * 4, stream)); LB_CU(cudaMemsetAsync(depth, 0, (size_t) n * 4, stream));
    LB_CU(cudaMemsetAsync(d_max, 0, 4, stream)); LB_CU(cudaMemsetAsync(parent, 0xff, (size_t) (2 * n) * 4, stream));
    {
        float3 l = make_float3(lo[0], lo[1], lo[2]);
        float3 ie = make_float3(hi[0] > lo[0] ? 1.f / (hi[0] - lo[0]) : 0.f, hi[1] > lo[1] ? 1.f / (hi[1] - lo[1]) : 0.f, hi[2] > lo[2] ? 1.f / (hi[2] - lo[2]) : 0.f);
        lbvh_keys_kernel<<<grid, 256, 0, stream>>>(d_verts, n, l, ie, keys);
        LB_CU(cudaGetLastError());
    }
    LB_CU(cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, keys, keys_sorted, (int) n, 0, 64, stream));
    LB_CU(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 1));
    LB_CU(cub::DeviceRadixSort::SortKeys(tmp, tmp_bytes, keys, keys_sorted, (int) n, 0, 64, stream));
    lbvh_leaves_kernel<<<grid, 256, 0, stream>>>(d_verts, d_flags, keys_sorted, n, pad, d_tris, d_nodes);
    LB_CU(cudaGetLastError());
    if (n > 1) {
        lbvh_hierarchy_kernel<<<grid, 256, 0, stream>>>(keys_sorted, (int) n, d_nodes, parent);
        LB_CU(cudaGetLastError());
        lbvh_fit_kernel<<<grid, 256, 0, stream>>>(d_nodes, parent, arrived, depth, (int) n, d_max);
        LB_CU(cudaGetLastError());
        int32_t h = 0;
        LB_CU(cudaMemcpyAsync(&h, d_max, 4, cudaMemcpyDeviceToHost, stream));
        LB_CU(cudaStreamSynchronize(stream));
        *max_depth = h;
        if (launches) *launches += 5;
    } else {
        LB_CU(cudaStreamSynchronize(stream));
        if (launches) *launches += 3;
    }
done:
#undef LB_CU
    if (keys) cudaFree(keys); if (keys_sorted) cudaFree(keys_sorted); if (tmp) cudaFree(tmp);
    if (parent) cudaFree(parent); if (depth) cudaFree(depth); if (arrived) cudaFree(arrived); if (d_max) cudaFree(d_max);
    return err;
}

}  // namespace vrlb
