// vrl_post.cu -- the two VRL-set transforms VRLIntegrator::preprocess applies after light tracing and before the gather
// (vrl.cpp:408-415):
//   * long beams   VRLMap::toLong -> VRL::convertToLong      (vrl_map.h:56-62, vrl_struct.h:146-154)
//   * dicing       VRLMap::dicingVRL -> VRL::dice            (vrl_map.h:104-128, vrl_struct.h:88-128)
// One thread per VRL.  Dicing is two passes (count, then write at host-prefixed offsets) so that the diced set keeps the
// reference's order: VRL-major, pieces from the origin outwards.
#include "vrlb_device.cuh"
#include "vrlb_kernels.h"

namespace vrlb {

// VRL::convertToLong: Ray3f r(origin, direction, 0) keeps the default [RayEpsilon, inf) interval (ray.h:34-35);
// length = si.t of the closest surface of ANY kind (the medium's own null-BSDF boundary included).
__global__ void vrls_to_long_kernel(DScene s, DVrl *v, uint64_t n, unsigned long long *missed) {
    uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    DVrl r = v[j];
    Hit h = closest_hit(s, V3(r.ox, r.oy, r.oz), V3(r.dx, r.dy, r.dz), kRayEpsilon, VRLB_INF);
    if (h.tri < 0) { atomicAdd(missed, 1ull); return; }      // reference: Log(Error, "Impossible to convert to long vrl...")
    v[j].length = h.t;
}

// Number of pieces of VRL::dice: `current_distance` is a double that starts at 0 and grows by the float avgLength
// while (current_distance <= length); sums of one fp32 value are exact in fp64, so current_distance = k * avgLength.
__device__ __forceinline__ uint32_t dice_count(float length, float dice_len) {
    const double L = (double) dice_len, len = (double) length;
    if (!(L > 0.0) || !(len >= 0.0)) return 0u;
    double kf = floor(len / L);
    if (kf > 4.0e6) kf = 4.0e6;                                // a VRL a million times the mean length is a broken set
    uint32_t k = (uint32_t) kf;
    while (__dmul_rn((double) (k + 1), L) <= len) ++k;        // division rounding: settle by the reference's own comparison
    while (k > 0 && __dmul_rn((double) k, L) > len) --k;
    return k + 1;                                              // pieces 0 .. k
}

__global__ void vrls_dice_kernel(const DVrl *in, uint64_t n, const DMedium *media, float dice_len, uint32_t *counts,
                                 const uint64_t *offsets, DVrl *out) {
    uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const DVrl v = in[j];
    const uint32_t cnt = dice_count(v.length, dice_len);
    if (!out) { counts[j] = cnt; return; }
    const DMedium m = media[v.medium];
    DVrl *dst = out + offsets[j];
    const double L = (double) dice_len;
    for (uint32_t k = 0; k < cnt; ++k) {
        const double cur = __dmul_rn((double) k, L);
        const float next = (float) fmin(__dsub_rn((double) v.length, cur), L);   // min(length - current_distance, avgLength)
        // vrl_trans (vrl_struct.h:97-107): 1 at distance 0, else Medium::evalTransmittance over [0, dist] -- homogeneous
        // branch exp(max_density * (mint - maxt)), medium.h:112-117
        float tr = 1.0f;
        if (cur != 0.0) {
            const float neg = __fsub_rn(0.0f, (float) cur);
            tr = m.max_density != 0.f ? expf(__fmul_rn(m.max_density, neg)) : 1.0f;
        }
        DVrl o = v;
        o.fr = __fmul_rn(v.fr, tr); o.fg = __fmul_rn(v.fg, tr); o.fb = __fmul_rn(v.fb, tr);
        // origin + direction * current_distance: float vector times a double is evaluated in fp64 and rounded once
        o.ox = (float) __dadd_rn((double) v.ox, __dmul_rn((double) v.dx, cur));
        o.oy = (float) __dadd_rn((double) v.oy, __dmul_rn((double) v.dy, cur));
        o.oz = (float) __dadd_rn((double) v.oz, __dmul_rn((double) v.dz, cur));
        o.length = next;
        dst[k] = o;                                            // medium, scatter depth (pad) copied; vrl_struct.h:118-121
    }
}

cudaError_t launch_vrls_to_long(const DScene &s, DVrl *v, uint64_t n, unsigned long long *missed, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    vrls_to_long_kernel<<<(uint32_t) ((n + 127) / 128), 128, 0, stream>>>(s, v, n, missed);
    return cudaGetLastError();
}
cudaError_t launch_vrls_dice(const DVrl *in, uint64_t n, const DMedium *media, float dice_len, uint32_t *counts,
                             const uint64_t *offsets, DVrl *out, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    vrls_dice_kernel<<<(uint32_t) ((n + 127) / 128), 128, 0, stream>>>(in, n, media, dice_len, counts, offsets, out);
    return cudaGetLastError();
}

}  // namespace vrlb
