// surface_photons.cu -- what the camera path collects at its first smooth surface (vrl.cpp:671-680):
//     radiance += throughput * (causticMap.estimateCausticRadiance(si, r_c, k_c) + globalMap.estimateRadiance(si, r_g, k_g))
// PhotonMap::estimateRadiance / estimateCausticRadiance (photonmap.h:131-211) take the k nearest photons inside the lookup
// radius from the kd-tree (PointKDTree::nnSearch, kdtree.h:319-400: once more than k points lie inside, the search radius
// shrinks to the k-th nearest distance and that shrunk radius is what the density estimate divides by), weight each with
// bsdf->eval(si, wo = to_local(-photon.direction)) -- SmoothDiffuse::eval: reflectance / pi * cos_theta_o where
// cos_theta_i and cos_theta_o are positive -- and scale by m_scale / (pi r^2); the caustic estimate adds the cone filter
// max(0, 1 - d / r) and a factor 3.
// Here the photons are binned into a uniform grid on the host (cell 2.5 R, the volume map's layout) and one thread serves
// one query in two sweeps over the <= 8 cells its sphere meets: sweep 1 keeps the k smallest squared distances in a
// max-heap (global-memory scratch, one stripe per query) and yields r^2; sweep 2 sums the photons inside r^2 in cell
// order.  Sums are deterministic; the reference's order is the heap's, so parity is to rounding (tests: 2e-5).
#include "vrlb_device.cuh"
#include "vrlb_kernels.h"

namespace vrlb {

namespace {

struct CellRange { int lo[3], hi[3]; bool any; };

__device__ inline CellRange cells_of(const PhotonGridDev &g, const float p[3], float radius) {
    CellRange c; c.any = true;
    for (int k = 0; k < 3; ++k) {
        c.lo[k] = max((int) floorf((p[k] - radius - g.grid_lo[k]) * g.inv_cell), 0);
        c.hi[k] = min((int) floorf((p[k] + radius - g.grid_lo[k]) * g.inv_cell), g.grid_res[k] - 1);
        c.any = c.any && c.lo[k] <= c.hi[k];
    }
    return c;
}

template <class F>
__device__ inline void for_photons(const PhotonGridDev &g, const CellRange &c, F &&f) {
    if (!c.any) return;
    for (int cz = c.lo[2]; cz <= c.hi[2]; ++cz) for (int cy = c.lo[1]; cy <= c.hi[1]; ++cy) {
        // the cells of one x-run are consecutive: one range of photons
        const uint32_t row = ((uint32_t) cz * (uint32_t) g.grid_res[1] + (uint32_t) cy) * (uint32_t) g.grid_res[0];
        const uint32_t b = __ldg(g.cell_start + row + (uint32_t) c.lo[0]), e = __ldg(g.cell_start + row + (uint32_t) c.hi[0] + 1u);
        for (uint32_t q = b; q < e; ++q) f(q);
    }
}

__device__ inline float dist2(const DPhoton *ph, uint32_t q, const float p[3]) {
    const float4 a = __ldg(reinterpret_cast<const float4 *>(ph + q));
    const float dx = xsub(a.x, p[0]), dy = xsub(a.y, p[1]), dz = xsub(a.z, p[2]);      // node.getPosition() - p
    return xadd(xadd(xmul(dx, dx), xmul(dy, dy)), xmul(dz, dz));
}

// max-heap of the k smallest values seen, in heap[0 .. n)
__device__ inline void sift_down(float *heap, int n, int i) {
    const float v = heap[i];
    for (;;) {
        int c = 2 * i + 1;
        if (c >= n) break;
        if (c + 1 < n && heap[c + 1] > heap[c]) ++c;
        if (!(heap[c] > v)) break;
        heap[i] = heap[c]; i = c;
    }
    heap[i] = v;
}

// r^2 of the estimate and how the membership test reads: d2 < R^2 while the map holds at most k photons in range,
// d2 <= (k-th smallest d2) once it holds more
__device__ inline float knn_radius2(const PhotonGridDev &g, const CellRange &c, const float p[3], float *heap, bool &shrunk) {
    const float R2 = xmul(g.lookup_radius, g.lookup_radius);
    const int k = g.lookup_size;
    int n = 0; bool is_heap = false; float r2 = R2;
    for_photons(g, c, [&](uint32_t q) {
        const float d2 = dist2(g.photons, q, p);
        if (!(d2 < r2)) return;
        if (n < k) { heap[n++] = d2; return; }
        if (!is_heap) { for (int i = n / 2 - 1; i >= 0; --i) sift_down(heap, n, i); is_heap = true; }
        // push d2, pop the largest: the largest of (heap, d2) leaves (kdtree.h:379-392)
        if (d2 < heap[0]) { heap[0] = d2; sift_down(heap, n, 0); }
        r2 = heap[0];
    });
    shrunk = is_heap;
    return r2;
}

__device__ inline void estimate(const PhotonGridDev &g, const vrlb_surface_query &sq, bool caustic, float *heap, float out[3]) {
    out[0] = out[1] = out[2] = 0.f;
    if (g.n == 0) return;
    const float p[3] = { sq.p[0], sq.p[1], sq.p[2] };
    V3 n(sq.n[0], sq.n[1], sq.n[2]);
    float cos_i = xdot(V3(sq.wi[0], sq.wi[1], sq.wi[2]), n);
    if (cos_i < 0.f && (sq.flags & VRLB_SURF_TWOSIDED)) { n = -n; cos_i = -cos_i; }      // twosided.cpp flips the frame
    if (!(cos_i > 0.f)) return;                                                          // SmoothDiffuse::eval: active needs cos_theta_i > 0
    const CellRange c = cells_of(g, p, g.lookup_radius);
    bool shrunk = false;
    const float r2 = knn_radius2(g, c, p, heap, shrunk);
    const float inv_r2 = __fdiv_rn(1.0f, r2);
    float acc[3] = { 0.f, 0.f, 0.f };
    for_photons(g, c, [&](uint32_t q) {
        const float d2 = dist2(g.photons, q, p);
        if (shrunk ? !(d2 <= r2) : !(d2 < r2)) return;
        const float4 b = __ldg(reinterpret_cast<const float4 *>(g.photons + q) + 1), w = __ldg(reinterpret_cast<const float4 *>(g.photons + q) + 2);
        const float cos_o = -xdot(V3(b.x, b.y, b.z), n);                                 // wo = to_local(-photon.direction)
        if (!(cos_o > 0.f)) return;
        float f = xmul(0.31830988618379067154f, cos_o);                                  // reflectance * INV_PI * cos_theta_o
        if (caustic) f = xmul(f, fmaxf(0.f, xsub(1.0f, __fsqrt_rn(xmul(d2, inv_r2)))));  // wp, photonmap.h:204
        acc[0] = xadd(acc[0], xmul(w.x, xmul(sq.reflectance[0], f)));
        acc[1] = xadd(acc[1], xmul(w.y, xmul(sq.reflectance[1], f)));
        acc[2] = xadd(acc[2], xmul(w.z, xmul(sq.reflectance[2], f)));
    });
    const float s = xmul(xmul(xmul(caustic ? 3.0f : 1.0f, g.scale), 0.31830988618379067154f), inv_r2);   // (3) m_scale * INV_PI * invSquaredRadius
    for (int ch = 0; ch < 3; ++ch) out[ch] = xmul(acc[ch], s);
}

__global__ void __launch_bounds__(128) surface_gather_kernel(SurfaceGatherParams p) {
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n_queries) return;
    const vrlb_surface_query sq = p.queries[i];
    float *heap = p.heap + i * (uint64_t) p.heap_stride;
    float c[3], g[3];
    estimate(p.caustic, sq, true, heap, c);
    estimate(p.global, sq, false, heap, g);
    const float out[3] = { xadd(c[0], g[0]), xadd(c[1], g[1]), xadd(c[2], g[2]) };
    if (p.query_out) { p.query_out[i * 3 + 0] = out[0]; p.query_out[i * 3 + 1] = out[1]; p.query_out[i * 3 + 2] = out[2]; }
    if (p.film_rgb && (out[0] != 0.f || out[1] != 0.f || out[2] != 0.f)) {
        atomicAdd(p.film_rgb + (size_t) sq.pixel * 3 + 0, xmul(out[0], sq.weight[0]));
        atomicAdd(p.film_rgb + (size_t) sq.pixel * 3 + 1, xmul(out[1], sq.weight[1]));
        atomicAdd(p.film_rgb + (size_t) sq.pixel * 3 + 2, xmul(out[2], sq.weight[2]));
    }
}

}  // namespace

cudaError_t launch_surface_gather(const SurfaceGatherParams &p, cudaStream_t stream) {
    if (p.n_queries == 0) return cudaSuccess;
    surface_gather_kernel<<<(uint32_t) ((p.n_queries + 127) / 128), 128, 0, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace vrlb
