// photons.cu -- the direct in-scattering term of VRLIntegrator::sample (vrl.cpp:523-611): along every in-medium camera
// segment, gather points every 2 r (the first at r, r = m_volumeLookupRadius * lerp(0.75, 1.25, u)) collect the volume
// photon map with PhotonMap::estimateRadianceVolume (photonmap.h:216-271, kernel 3/pi (1 - d^2/r^2)^2 / r^2), weighted by
// the grey transmittance accumulated from the segment origin (Medium::homoEvalTransmittance, medium.h:103-107) and the
// path throughput; the sum is scaled by the map's 1 / directVolumePathCount.
// The reference finds photons with a kd-tree range query; here the photons are binned into a uniform grid of cell size
// 2 r_max on the host when the set becomes resident (sorted by cell, order inside a cell = set order: deterministic
// sums) and a gather point visits the <= 8 cells its sphere overlaps.  One thread per camera segment.
// The jitter draw u is the draw just before the segment's query position in the pixel's PCG32 stream (vrl.cpp:527 then
// :619); it is recovered by stepping the LCG back once, so the 64-byte segment record does not grow.
#include "vrlb_device.cuh"
#include "vrlb_kernels.h"

namespace vrlb {

constexpr uint64_t PCG32_MULT_INV = 0xc097ef87329e28a5ULL;      // PCG32_MULT * PCG32_MULT_INV = 1 (mod 2^64)

__device__ __forceinline__ float phase_iso_or_hg(const DMedium &m, V3 wi, V3 wo) {
    if (m.phase != VRLB_PHASE_HG) return kInvFourPi;                      // isotropic.cpp:142-146
    float g = m.g;
    float cos_theta = xdot(wo, wi);                                       // hg.cpp:77-81: eval_hg(dot(wo, mi.wi))
    float temp = xadd(xadd(1.0f, xmul(g, g)), xmul(xmul(2.0f, g), cos_theta));
    return __fdiv_rn(xmul(kInvFourPi, xsub(1.0f, xmul(g, g))), xmul(temp, __fsqrt_rn(temp)));
}

__global__ void photon_gather_kernel(PhotonGatherParams p) {
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n_segs) return;
    const float4 *sp = reinterpret_cast<const float4 *>(p.segs + i);
    const float4 r0 = __ldg(sp), r1 = __ldg(sp + 1), r2 = __ldg(sp + 2);
    const ulonglong2 r3 = __ldg(reinterpret_cast<const ulonglong2 *>(sp + 3));
    const V3 o(r0.x, r0.y, r0.z), d(r1.x, r1.y, r1.z);
    const float maxt = r0.w;
    const uint32_t pixel = __float_as_uint(r1.w);
    const int medium = __float_as_int(r2.w);
    const uint32_t part = (uint32_t) (r3.y >> 32), sample = (uint32_t) r3.y;
    float out[3] = { 0.f, 0.f, 0.f };
    const DMedium &m = p.scene.media[medium];
    // A curved camera path (graded-IOR medium) is n_parts consecutive records (segments.cu; n_parts in the high half of
    // `part`, or of `sample` under use_nl_atomic_query): the thread of its first record walks all of them in order, with
    // the step phase carried from part to part (leftOver, vrl.cpp:553-577), the transmittance taken over the straight
    // distance between consecutive gather points and the phase function against the first part's direction
    // (mediumRay.d, vrl.cpp:530, 604-606).  A straight segment is the one-part case of the same walk.
    if ((part & 0xffffu) == 0 && p.n_photons > 0) {
        uint32_t n_parts = max(max(part >> 16, sample >> 16), 1u);
        if (i + n_parts > p.n_segs) n_parts = (uint32_t) (p.n_segs - i);
        Pcg32 rng; rng.state = (r3.x - PCG32_INC) * PCG32_MULT_INV;      // one draw back: the radius jitter, vrl.cpp:527
        const float u = rng.next_f32();
        const float radius = xmul(p.lookup_radius, xadd(xmul(1.25f, u), xsub(0.75f, xmul(0.75f, u))));   // enoki::lerp(a, b, t) = b t + (a - a t)
        const float r2q = xmul(radius, radius);
        float step = radius, left = 0.f;
        V3 last = o;
        float tr[3] = { r2.x, r2.y, r2.z };                               // throughput at the path's start
        float acc[3] = { 0.f, 0.f, 0.f };
        for (uint32_t k = 0; k < n_parts; ++k) {
            V3 po = o, pd = d; float pmaxt = maxt;
            if (k > 0) {
                const float4 *sk = reinterpret_cast<const float4 *>(p.segs + i + k);
                const float4 k0 = __ldg(sk), k1 = __ldg(sk + 1);
                po = V3(k0.x, k0.y, k0.z); pd = V3(k1.x, k1.y, k1.z); pmaxt = k0.w;
            }
            float t = 0.f;
            for (int guard = 0; guard < 10000000 && t <= xsub(pmaxt, xsub(step, left)); ++guard) {
                t = xadd(t, xsub(step, left));
                left = 0.f;
                const V3 gp(xadd(po.x, xmul(pd.x, t)), xadd(po.y, xmul(pd.y, t)), xadd(po.z, xmul(pd.z, t)));     // gatherRay(gather_t), unfused like the oracle
                step = xmul(radius, 2.0f);
                const V3 dl(xsub(gp.x, last.x), xsub(gp.y, last.y), xsub(gp.z, last.z));
                const float len = __fsqrt_rn(xdot(dl, dl));
                const float T = m.max_density != 0.f ? expf(xmul(m.max_density, -len)) : 1.0f;   // homoEvalTransmittance
                tr[0] = xmul(tr[0], T); tr[1] = xmul(tr[1], T); tr[2] = xmul(tr[2], T);
                // estimateRadianceVolume(gatherPoint, mediumRay.d, medium, sampler, radius, M)
                float est[3] = { 0.f, 0.f, 0.f };
                int lo[3], hi[3];
                const float g3[3] = { gp.x, gp.y, gp.z };
                bool any = true;
                for (int c = 0; c < 3; ++c) {
                    lo[c] = (int) floorf((g3[c] - radius - p.grid_lo[c]) * p.inv_cell);
                    hi[c] = (int) floorf((g3[c] + radius - p.grid_lo[c]) * p.inv_cell);
                    lo[c] = max(lo[c], 0); hi[c] = min(hi[c], p.grid_res[c] - 1);
                    any = any && lo[c] <= hi[c];
                }
                if (any)
                    for (int cz = lo[2]; cz <= hi[2]; ++cz) for (int cy = lo[1]; cy <= hi[1]; ++cy) for (int cx = lo[0]; cx <= hi[0]; ++cx) {
                        const uint32_t c = ((uint32_t) cz * (uint32_t) p.grid_res[1] + (uint32_t) cy) * (uint32_t) p.grid_res[0] + (uint32_t) cx;
                        for (uint32_t q = __ldg(p.cell_start + c); q < __ldg(p.cell_start + c + 1); ++q) {
                            const float4 *pp = reinterpret_cast<const float4 *>(p.photons + q);
                            const float4 a = __ldg(pp), b = __ldg(pp + 1), w = __ldg(pp + 2);
                            const V3 dp(xsub(gp.x, a.x), xsub(gp.y, a.y), xsub(gp.z, a.z));
                            const float l2 = xdot(dp, dp);
                            if (xsub(r2q, l2) < 0.f) continue;                                     // photonmap.h:226-229
                            const float pf = phase_iso_or_hg(m, V3(-b.x, -b.y, -b.z), d);       // mi1.wi = -photon.direction, wo = ray.d
                            const float q1 = xsub(1.0f, __fdiv_rn(l2, r2q));
                            // power * photonPF * sqr(1 - l2 / r2) / r2 * INV_PI * 3, left to right (photonmap.h:245)
                            const float wgt = xmul(q1, q1);
                            const float pw[3] = { w.x, w.y, w.z };
                            for (int ch = 0; ch < 3; ++ch)
                                est[ch] = xadd(est[ch], xmul(xmul(__fdiv_rn(xmul(xmul(pw[ch], pf), wgt), r2q), 0.31830988618379067154f), 3.0f));
                        }
                    }
                for (int ch = 0; ch < 3; ++ch) acc[ch] = xadd(acc[ch], xmul(tr[ch], est[ch]));
                last = gp;
            }
            if (k + 1 < n_parts) left = xadd(left, xsub(pmaxt, t));           // leftOver += nli.t - gather_t, vrl.cpp:570
        }
        for (int ch = 0; ch < 3; ++ch) out[ch] = xmul(acc[ch], p.scale);   // directIllum *= getScaleFactor()
    }
    if (p.seg_out) { p.seg_out[i * 3 + 0] = out[0]; p.seg_out[i * 3 + 1] = out[1]; p.seg_out[i * 3 + 2] = out[2]; }
    if (p.film_rgb && (out[0] != 0.f || out[1] != 0.f || out[2] != 0.f)) {
        atomicAdd(p.film_rgb + (size_t) pixel * 3 + 0, out[0]);           // the throughput is already inside `tr`
        atomicAdd(p.film_rgb + (size_t) pixel * 3 + 1, out[1]);
        atomicAdd(p.film_rgb + (size_t) pixel * 3 + 2, out[2]);
    }
}

cudaError_t launch_photon_gather(const PhotonGatherParams &p, cudaStream_t stream) {
    if (p.n_segs == 0) return cudaSuccess;
    photon_gather_kernel<<<(uint32_t) ((p.n_segs + 127) / 128), 128, 0, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace vrlb
