// gather.cu -- the camera-segment x VRL gather (north-star subsystem 2).
//
// Replaces VRLMap::query (src/integrators/vrl/vrl_map.h:133-211, no-acceleration branch) and
// VRL::getContrib / samplingVRL / findClosetPoint / evalTransmittance
// (src/integrators/vrl/vrl_struct.h:163-228, 239-360, 509-572, 584-662).
//
// Mapping: one thread per camera segment, 32 neighbouring (Morton-ordered) segments per warp, so a
// warp evaluates one VRL against 32 coherent camera segments per step; the VRL record is a
// shared-memory broadcast.  The VRL set (48 B records) is streamed through shared memory in tiles of
// TILE records by the TMA engine (cp.async.bulk + mbarrier complete_tx), double buffered.  The path is
// FP32/SFU-issue bound: ~0 HBM bytes per eval, all operands in registers/shared memory.
//
// PCG32 addressing: eval (sample s, global VRL j) of a segment uses the draws at offset
// ((s*total + j) * dpe) from the segment's rng_state (dpe = 2, or 3 with distance roulette); a thread
// jumps to its shard's first draw with PCG32::advance and then steps the LCG.
#include "vrlb_device.cuh"
#include "vrlb_kernels.h"

namespace vrlb {

constexpr int GATHER_THREADS = 256;
constexpr int TILE           = 256;   // VRL records per shared-memory tile (12 KB)

// ---- mbarrier / TMA bulk-copy PTX (sm_90+; SASS: UBLKCP + SYNCS) -----------------------------------
__device__ inline uint32_t smem_u32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ inline void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ inline void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ inline void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ inline void tma_bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ---- VRL::findClosetPoint, vrl_struct.h:163-228 -----------------------------------------------------
__device__ inline void find_closest_point(V3 vo, V3 vd, float vlen, V3 ro, V3 rd, float rmaxt, float &tCam, float &tVRL) {
    V3 u = vd * vlen;
    V3 v = rd * rmaxt;
    V3 w = vo - ro;
    float a = dot(u, u), b = dot(u, v), c = dot(v, v), d = dot(u, w), e = dot(v, w);
    float D = a * c - b * b;
    float sN, sD = D, tN, tD = D;
    if (D < 0.0001f) {
        sN = 0.0f; sD = 1.0f; tN = e; tD = c;
    } else {
        sN = (b * e - c * d);
        tN = (a * e - b * d);
        if (sN < 0.0f) { sN = 0.0f; tN = e; tD = c; }
        else if (sN > sD) { sN = sD; tN = e + b; tD = c; }
    }
    if (tN < 0.0f) {
        tN = 0.0f;
        if (-d < 0.0f) sN = 0.0f;
        else if (-d > a) sN = sD;
        else { sN = -d; sD = a; }
    } else if (tN > tD) {
        tN = tD;
        if ((-d + b) < 0.0f) sN = 0;
        else if ((-d + b) > a) sN = sD;
        else { sN = (-d + b); sD = a; }
    }
    float sc = (fabsf(sN) < 0.0001f ? 0.0f : sN / sD);
    float tc = (fabsf(tN) < 0.0001f ? 0.0f : tN / tD);
    tCam = tc * rmaxt;
    tVRL = sc * vlen;
}

struct SegRegs { V3 o, d; float maxt; };

// ---- VRL::getContrib (vrl_struct.h:584-662) with samplingVRL (239-360) inlined.  Returns the scalar
//      factor that multiplies flux * sigma_s(cam) * sigma_s(vrl); 0 when the reference returns 0. ----
template <bool UNIFORM, bool VIS>
__device__ inline float eval_contrib(const DScene &scene, const SegRegs &ray, V3 vo, V3 vd, float vlen,
                                     float max_density, int phase, float g, float u1, float u2) {
    V3 pCam, pVRL; float tCam, tVRL, invPDF;
    if (UNIFORM) {
        // VRL::sampleMC, vrl_struct.h:239-243 (single-part NLRay::at, :44-56)
        tCam = u1 * ray.maxt;
        tVRL = u2 * vlen;
        pCam = ray.o + ray.d * tCam;
        pVRL = vo + vd * tVRL;
        invPDF = vlen * ray.maxt;
    } else {
        float cpCam, cpVRL;
        find_closest_point(vo, vd, vlen, ray.o, ray.d, ray.maxt, cpCam, cpVRL);
        V3 vh = vo + vd * cpVRL;
        V3 uh = ray.o + ray.d * cpCam;
        float h = squared_norm(uh - vh);
        if (h == 0.0f) return 0.f;                               // vrl_struct.h:264
        h = safe_sqrt(h);
        float v0Hat = -cpVRL;
        float v1Hat = vlen + v0Hat;
        float sinTheta = squared_norm(cross(vd, ray.d));
        if (sinTheta == 0) return 0.f;                           // vrl_struct.h:275
        sinTheta = safe_sqrt(sinTheta);
        // inverseCDF_A, vrl_struct.h:293-310 (Novak et al. 2012, eq. 10-13)
        float z0 = (v0Hat / h) * sinTheta, z1 = (v1Hat / h) * sinTheta;
        float a0 = logf(z0 + sqrtf(z0 * z0 + 1));
        float a1 = logf(z1 + sqrtf(z1 * z1 + 1));
        float vHat = (h * sinhf(lerpf(a0, a1, u1))) / sinTheta;
        float invPDF_A = (a1 - a0) * sqrtf(h * h + vHat * vHat * sinTheta * sinTheta);
        invPDF_A /= sinTheta;
        tVRL = (vHat + cpVRL);
        pVRL = vo + vd * tVRL;
        // equi-angular sampling of the camera part, vrl_struct.h:318-336 (Kulla & Fajardo 2012)
        float uHat  = dot(ray.d, (pVRL - ray.o));
        float u0Hat = -uHat;
        float u1Hat = ray.maxt + u0Hat;
        float hPoint = norm(ray.o + ray.d * uHat - pVRL);
        float thetaA = atanf(u0Hat / hPoint);
        float thetaB = atanf(u1Hat / hPoint);
        float uS     = hPoint * tanf(lerpf(thetaA, thetaB, u2));
        float invPDF_B = (thetaB - thetaA) * (hPoint * hPoint + uS * uS) / hPoint;
        tCam = (uS - u0Hat);
        pCam = ray.o + ray.d * tCam;
        tCam = fminf(fmaxf(tCam, 0.0f), ray.maxt);               // vrl_struct.h:340-341
        tVRL = fminf(fmaxf(tVRL, 0.0f), vlen);
        invPDF = invPDF_A * invPDF_B;
    }
    V3 dir = pVRL - pCam;
    float lengthPtoP = norm(dir);
    if (lengthPtoP == 0) return 0.f;                             // vrl_struct.h:589
    if (isnan(invPDF)) return 0.f;                               // vrl_struct.h:594
    dir = dir / lengthPtoP;
    float rayPF = phase_eval(phase, g, -ray.d, dir);             // wi = -part.d (quirk 1)
    if (rayPF == 0) return 0.f;
    float vrlPF = phase_eval(phase, g, -vd, -dir);
    if (vrlPF == 0) return 0.f;
    // VRL::evalTransmittance pCam -> pVRL, vrl_struct.h:509-572
    float vrlToRayTrans;
    {
        V3 dd = pVRL - pCam;
        float remaining = norm(dd);
        dd = normalize(dd);
        if (VIS && remaining > 0 && any_hit(scene, pCam, dd, 0.f, remaining)) return 0.f;
        vrlToRayTrans = remaining > 0 ? homo_transmittance(max_density, remaining) : 1.0f;
    }
    if (vrlToRayTrans == 0.0f) return 0.f;
    float rayTrans = homo_transmittance(max_density, tCam);
    if (rayTrans == 0.0f) return 0.f;
    float vrlTrans = homo_transmittance(max_density, tVRL);
    if (vrlTrans == 0.0f) return 0.f;
    float fallOff = 1.0f / (lengthPtoP * lengthPtoP);
    return fallOff * vrlPF * rayPF * vrlTrans * rayTrans * vrlToRayTrans * invPDF;
}

// compensated (Kahan) fp32 accumulation: the per-pixel sum over up to 1e6 VRLs stays at fp64-sum
// quality without fp64 issue slots ("FP32 accumulation order")
struct Kahan3 {
    float s[3], c[3];
    __device__ inline void init() { s[0] = s[1] = s[2] = 0.f; c[0] = c[1] = c[2] = 0.f; }
    __device__ inline void add(int k, float x) {
        float y = x - c[k];
        float t = s[k] + y;
        c[k]    = (t - s[k]) - y;
        s[k]    = t;
    }
};

template <bool UNIFORM, bool RR, bool VIS>
__global__ void __launch_bounds__(GATHER_THREADS)
gather_kernel(GatherParams p) {
    __shared__ alignas(128) DVrl tile[2][TILE];
    __shared__ alignas(8) uint64_t full_bar[2];

    const uint32_t tid = threadIdx.x;
    const uint64_t seg_i = (uint64_t) blockIdx.x * GATHER_THREADS + tid;
    const bool live = seg_i < p.n_segs;

    SegRegs ray; uint64_t rng0 = 0; uint32_t pixel = 0; float w[3] = { 0.f, 0.f, 0.f };
    if (live) {
        const float4 *sp = reinterpret_cast<const float4 *>(p.segs + seg_i);
        float4 r0 = __ldg(sp), r1 = __ldg(sp + 1), r2 = __ldg(sp + 2);
        ulonglong2 r3 = __ldg(reinterpret_cast<const ulonglong2 *>(sp + 3));
        ray.o = V3(r0.x, r0.y, r0.z); ray.maxt = r0.w;
        ray.d = V3(r1.x, r1.y, r1.z); pixel = __float_as_uint(r1.w);
        w[0] = r2.x; w[1] = r2.y; w[2] = r2.z;
        rng0 = r3.x;
    } else {
        ray.o = V3(0.f); ray.d = V3(0.f, 0.f, 1.f); ray.maxt = 1.f;
    }

    const uint32_t n_tiles = (uint32_t) ((p.n_vrls + TILE - 1) / TILE);
    const uint32_t total_tiles = n_tiles * p.q;     // the tile sequence repeats once per query sample
    constexpr uint32_t DPE = RR ? 3u : 2u;

    if (tid == 0) {
        mbar_init(&full_bar[0], 1);
        mbar_init(&full_bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto issue = [&](uint32_t it) {
        uint32_t t = it % n_tiles, buf = it & 1u;
        uint64_t first = (uint64_t) t * TILE;
        uint32_t cnt = (uint32_t) min((uint64_t) TILE, p.n_vrls - first);
        uint32_t bytes = cnt * (uint32_t) sizeof(DVrl);
        mbar_expect_tx(&full_bar[buf], bytes);
        tma_bulk_g2s(&tile[buf][0], p.vrls + first, bytes, &full_bar[buf]);
    };
    if (tid == 0) {
        if (total_tiles > 0) issue(0);
        if (total_tiles > 1) issue(1);
    }

    const DMedium *media = p.scene.media;
    Kahan3 acc; acc.init();
    Pcg32 rng; rng.state = rng0;

    for (uint32_t it = 0; it < total_tiles; ++it) {
        const uint32_t t = it % n_tiles, s = it / n_tiles, buf = it & 1u;
        if (t == 0 && live) {
            rng.state = rng0;
            rng.advance(((uint64_t) s * p.total + p.first_index) * DPE);
        }
        mbar_wait(&full_bar[buf], (it >> 1) & 1u);
        const uint32_t cnt = (uint32_t) min((uint64_t) TILE, p.n_vrls - (uint64_t) t * TILE);
        if (live) {
            for (uint32_t j = 0; j < cnt; ++j) {
                const float4 *vp = reinterpret_cast<const float4 *>(&tile[buf][j]);
                float4 a = vp[0], b = vp[1], c = vp[2];
                V3 vo(a.x, a.y, a.z), vd(b.x, b.y, b.z);
                float vlen = a.w;
                const DMedium &m = media[__float_as_uint(b.w)];
                float rrWeight = 1.f; bool survive = true;
                if (RR) {
                    // EDistanceRoulette, vrl_map.h:171-190
                    float cc, cv;
                    find_closest_point(vo, vd, vlen, ray.o, ray.d, ray.maxt, cc, cv);
                    float d2 = squared_norm((ray.o + ray.d * cc) - (vo + vd * cv));
                    float min_distance = sqrtf(d2);
                    rrWeight = fminf(1 / (min_distance * p.scale_rr), 1.0f);
                    survive  = rrWeight > rng.next_f32();
                }
                float u1 = rng.next_f32(), u2 = rng.next_f32();
                if (!survive) continue;
                float f = eval_contrib<UNIFORM, VIS>(p.scene, ray, vo, vd, vlen, m.max_density, m.phase, m.g, u1, u2);
                if (RR) f = f / rrWeight;
                // flux * f * sigma_s(cam) * sigma_s(vrl) * m_scale (vrl_struct.h:652-659, vrl_map.h:160)
                float cr = c.x * f * m.sigma_s[0] * m.sigma_s[0] * p.scale;
                float cg = c.y * f * m.sigma_s[1] * m.sigma_s[1] * p.scale;
                float cb = c.z * f * m.sigma_s[2] * m.sigma_s[2] * p.scale;
                // NaN / Inf rejection, vrl_map.h:162
                if (isfinite(cr) && isfinite(cg) && isfinite(cb)) { acc.add(0, cr); acc.add(1, cg); acc.add(2, cb); }
            }
        }
        __syncthreads();                       // everyone is done with tile[buf]
        if (tid == 0 && it + 2 < total_tiles) issue(it + 2);
    }

    if (live) {
        float inv_q = p.q > 1 ? 1.0f / (float) p.q : 1.0f;
        float lr = acc.s[0], lg = acc.s[1], lb = acc.s[2];
        if (p.q > 1) { lr = lr / (float) p.q; lg = lg / (float) p.q; lb = lb / (float) p.q; }
        (void) inv_q;
        if (p.seg_out) { p.seg_out[seg_i * 3 + 0] = lr; p.seg_out[seg_i * 3 + 1] = lg; p.seg_out[seg_i * 3 + 2] = lb; }
        if (p.film_rgb) {
            // indirectIllum += color * vrlThroughput (vrl.cpp:617, 626); film pixel owned via fp32 RED
            atomicAdd(p.film_rgb + (size_t) pixel * 3 + 0, lr * w[0]);
            atomicAdd(p.film_rgb + (size_t) pixel * 3 + 1, lg * w[1]);
            atomicAdd(p.film_rgb + (size_t) pixel * 3 + 2, lb * w[2]);
        }
    }
}

template <bool U, bool R, bool V>
static cudaError_t launch_t(const GatherParams &p, cudaStream_t stream) {
    uint32_t grid = (uint32_t) ((p.n_segs + GATHER_THREADS - 1) / GATHER_THREADS);
    if (grid == 0 || p.n_vrls == 0) return cudaSuccess;
    gather_kernel<U, R, V><<<grid, GATHER_THREADS, 0, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_gather(const GatherParams &p, bool uniform, bool rr, bool vis, cudaStream_t stream) {
    int key = (uniform ? 4 : 0) | (rr ? 2 : 0) | (vis ? 1 : 0);
    switch (key) {
        case 0: return launch_t<false, false, false>(p, stream);
        case 1: return launch_t<false, false, true>(p, stream);
        case 2: return launch_t<false, true, false>(p, stream);
        case 3: return launch_t<false, true, true>(p, stream);
        case 4: return launch_t<true, false, false>(p, stream);
        case 5: return launch_t<true, false, true>(p, stream);
        case 6: return launch_t<true, true, false>(p, stream);
        default: return launch_t<true, true, true>(p, stream);
    }
}

}  // namespace vrlb
