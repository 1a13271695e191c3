// gather_fast.cu -- the camera-segment x VRL gather, production arithmetic (vrlb_options::math_mode = 1).
//
// Same algorithm, stream addressing and tile pipeline as gather.cu (which restates the reference's IEEE
// operation order for bit-level comparison with the oracle); here every well-conditioned step uses the
// issue-cheapest sm_100a form: FMA contraction, MUFU.RCP / RSQ / LG2 / EX2 instead of IEEE division, sqrt,
// logf, sinhf, expf, and two short polynomials for atan / tan (tools/fit_polys.py, <= 1.5e-7 relative).
// What stays exactly rounded, because the reference's own formula is ill-conditioned there and a
// different rounding would produce a different (equally noisy) value:
//   * asinh(z) = log(z + sqrt(z*z + 1)) for z < 0 (vrl_struct.h:297-301): z*z, +1, sqrt and the sum are
//     single IEEE operations, only the final log is MUFU.LG2;
//   * the Moeller-Trumbore shadow test (vrlb_device.cuh::intersect_tri).
// Parity: tests/test_gather_gpu.py checks this kernel against the oracle at the north-star tolerance
// (per-pixel relative error <= 1e-4); per-eval relative error is ~1e-6.
//
// Per-eval cost drops from ~710 to ~260 warp instructions (profiles/), which is what moves the FP32-issue-bound
// kernel; the VRL records are pre-multiplied (flux * sigma_s^2 * m_scale, extinction in the pad lane) by
// prep_fast_vrls so the inner loop reads no medium table when every medium is isotropic.
#include "gather_exact.cuh"

namespace vrlb {

__device__ __forceinline__ float f_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float f_sqrt(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float f_rsqrt(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float f_lg2(float x) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float f_ex2(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

constexpr float kLn2 = 0.693147180559945309f, kLog2e = 1.44269504088896341f, kHalfPi = 1.57079632679489662f;

// atan(x): |x| <= 1 -> x P(x^2); else +-pi/2 - atan(1/x)
__device__ __forceinline__ float f_atan(float x) {
    float ax = fabsf(x);
    bool big = ax > 1.0f;
    float t = big ? f_rcp(ax) : ax;
    float s = t * t;
    float p = 0.002849772f;
    p = fmaf(p, s, -0.01606816f); p = fmaf(p, s, 0.042690758f); p = fmaf(p, s, -0.07504231f);
    p = fmaf(p, s, 0.10640904f);  p = fmaf(p, s, -0.14203636f); p = fmaf(p, s, 0.19992618f);
    p = fmaf(p, s, -0.33333072f);
    float r = fmaf(p * s, t, t);
    r = big ? kHalfPi - r : r;
    return copysignf(r, x);
}
// tan(x) for |x| < pi/2: |x| <= pi/4 -> x Q(x^2); else 1 / tan(pi/2 - |x|)
__device__ __forceinline__ float f_tan(float x) {
    float ax = fabsf(x);
    bool big = ax > 0.78539816339744831f;
    float t = big ? kHalfPi - ax : ax;
    float s = t * t;
    float q = 0.009499237f;
    q = fmaf(q, s, 0.0028941673f); q = fmaf(q, s, 0.02460186f); q = fmaf(q, s, 0.05334885f);
    q = fmaf(q, s, 0.13339895f);   q = fmaf(q, s, 0.33333075f);
    float r = fmaf(q * s, t, t);
    r = big ? f_rcp(r) : r;
    return copysignf(r, x);
}

// This translation unit is compiled with -fmad=false (the exact slow path lives here too), so the fast
// path spells its FMAs out.
__device__ __forceinline__ float fdot(V3 a, V3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
__device__ __forceinline__ V3 fmad3(V3 o, V3 d, float t) { return V3(fmaf(d.x, t, o.x), fmaf(d.y, t, o.y), fmaf(d.z, t, o.z)); }
__device__ __forceinline__ V3 fcross(V3 a, V3 b) {
    return V3(fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x)));
}

struct FastSeg {
    V3 o, d;          // part origin / direction
    float maxt;
    V3 v;             // d * maxt
    float c;          // dot(v, v)
    __device__ __forceinline__ float dotv(V3 x) const { return fmaf(v.x, x.x, fmaf(v.y, x.y, v.z * x.z)); }
    __device__ __forceinline__ float cc() const { return c; }
};
// The same without the hoisted v and c: four registers fewer across the VRL loop for three more multiplies per eval
// (the visibility kernel sits at its register cap; d is unit length, so dot(v, v) = maxt^2 to fp32 rounding).
struct FastSegC {
    V3 o, d;
    float maxt;
    __device__ __forceinline__ float dotv(V3 x) const { return maxt * fmaf(d.x, x.x, fmaf(d.y, x.y, d.z * x.z)); }
    __device__ __forceinline__ float cc() const { return maxt * maxt; }
};

// VRL::findClosetPoint, vrl_struct.h:163-228 (v, c hoisted per camera segment)
template <class Seg>
__device__ __forceinline__ void closest_fast(V3 vo, V3 vd, float vlen, const Seg &r, float &tCam, float &tVRL) {
    V3 u = vd * vlen;
    V3 w = vo - r.o;
    float a = fdot(u, u), b = r.dotv(u), c = r.cc(), d = fdot(u, w), e = r.dotv(w);
    float D = fmaf(a, c, -(b * b));
    float sN, sD = D, tN, tD = D;
    if (D < 0.0001f) {
        sN = 0.0f; sD = 1.0f; tN = e; tD = c;
    } else {
        sN = fmaf(b, e, -(c * d));
        tN = fmaf(a, e, -(b * d));
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
    float sc = (fabsf(sN) < 0.0001f ? 0.0f : sN * f_rcp(sD));
    float tc = (fabsf(tN) < 0.0001f ? 0.0f : tN * f_rcp(tD));
    tCam = tc * r.maxt;
    tVRL = sc * vlen;
}

__device__ __forceinline__ float hg_fast(float g, float cos_theta) {
    float temp = fmaf(2.0f * g, cos_theta, fmaf(g, g, 1.0f));
    float rs = f_rsqrt(temp);
    return kInvFourPi * (1.0f - g * g) * (rs * rs * rs);
}

// Conditioning guard.  The estimator is ill-conditioned when the VRL passes close to the camera part: the
// reference's asinh(z) = log(z + sqrt(z*z + 1)) cancels for z << 0 (relative noise 2 z^2 eps, an exact 0 -> NaN ->
// rejected eval beyond |z| ~ 4096), and the equi-angular pdf ~ 1/hPoint amplifies any position error by
// 1/h.  Those evals carry the largest weights of their pixel, so they are evaluated by the exactly-rounded path
// (gather_exact.cuh), which reproduces the reference's rounding; everything else takes the fast path, whose
// per-eval relative error (~1e-6) is then not amplified.  Thresholds come from GatherParams.
struct Guard { float h2min, zmax; };

// VRL::getContrib (vrl_struct.h:584-662) with samplingVRL (239-360) inlined; returns the scalar that multiplies
// flux * sigma_s(cam) * sigma_s(vrl) * m_scale BEFORE the shadow test (the segment to trace is returned in so / sd /
// sdist when VIS), 0 where the reference returns 0.  `slow` is set (and nothing
// else computed) when the guard wants the exact path.
// What the visibility kernel needs of an eval's connection segment pCam -> pVRL: its start, the unnormalised direction,
// |dir|^2 and 1/|dir|, and the two sampled parameters BEFORE the clamp of vrl_struct.h:340-341 (pCam / pVRL are computed
// from the unclamped values there, so these are the parameters of the segment's end points on the camera part / the VRL).
struct Conn { V3 pCam, dir; float d2, rd, tCam, tVRL; };

template <bool UNIFORM, bool VIS, bool ISO, bool GUARD, class Seg>
__device__ __forceinline__ float contrib_fast(const DScene &scene, const Seg &ray, V3 vo, V3 vd, float vlen,
                                              float max_density, float g, float u1, float u2, Guard guard, bool &slow,
                                              Conn &cn) {
    V3 pCam, pVRL; float tCam, tVRL, invPDF;
    slow = false;
    if (UNIFORM) {
        tCam = u1 * ray.maxt;                                    // VRL::sampleMC, vrl_struct.h:239-243
        tVRL = u2 * vlen;
        pCam = fmad3(ray.o, ray.d, tCam);
        pVRL = fmad3(vo, vd, tVRL);
        invPDF = vlen * ray.maxt;
        if (VIS) { cn.tCam = tCam; cn.tVRL = tVRL; }
    } else {
        float cpCam, cpVRL;
        closest_fast(vo, vd, vlen, ray, cpCam, cpVRL);
        V3 vh = fmad3(vo, vd, cpVRL);
        V3 uh = fmad3(ray.o, ray.d, cpCam);
        V3 dh = uh - vh;
        float h2 = fdot(dh, dh);
        V3 cr = fcross(vd, ray.d);
        float st2 = fdot(cr, cr);
        float rh = f_rsqrt(h2);
        float rst = f_rsqrt(st2);
        float k  = st2 * rst * rh;                               // sin(theta) / h
        float z0 = -cpVRL * k, z1 = (vlen - cpVRL) * k;
        if (GUARD && !(h2 >= guard.h2min && z0 >= -guard.zmax && st2 > 1e-6f)) { slow = true; return 0.f; }
        if (h2 == 0.0f) return 0.f;                              // vrl_struct.h:264
        if (st2 == 0.f) return 0.f;                              // vrl_struct.h:275
        float h  = h2 * rh;
        // inverseCDF_A, vrl_struct.h:293-310; asinh(z) = log(z + sqrt(z^2 + 1)), cancellation-free for z < 0
        float r0 = f_sqrt(fmaf(z0, z0, 1.0f)), r1 = f_sqrt(fmaf(z1, z1, 1.0f));
        float e0 = z0 < 0.f ? f_rcp(r0 - z0) : z0 + r0;
        float e1 = z1 < 0.f ? f_rcp(r1 - z1) : z1 + r1;
        float l0 = f_lg2(e0), l1 = f_lg2(e1);                    // a = l * ln2
        float E  = f_ex2(fmaf(l1 - l0, u1, l0));                 // exp(lerp(a0, a1, u1))
        float sh = 0.5f * (E - f_rcp(E));                        // sinh
        float vHat = h * sh * rst;
        float invPDF_A = (l1 - l0) * kLn2 * f_sqrt(fmaf(vHat * vHat, st2, h2)) * rst;
        tVRL = vHat + cpVRL;
        pVRL = fmad3(vo, vd, tVRL);
        // equi-angular sampling of the camera part, vrl_struct.h:318-336
        float uHat  = fdot(ray.d, pVRL - ray.o);
        float u0Hat = -uHat;
        float u1Hat = ray.maxt + u0Hat;
        V3 perp = fmad3(ray.o, ray.d, uHat) - pVRL;
        float hp2 = fdot(perp, perp);
        float rhp = f_rsqrt(hp2);
        float hPoint = hp2 * rhp;
        float thetaA = f_atan(u0Hat * rhp);
        float thetaB = f_atan(u1Hat * rhp);
        float uS     = hPoint * f_tan(fmaf(thetaB - thetaA, u2, thetaA));
        float invPDF_B = (thetaB - thetaA) * fmaf(uS, uS, hp2) * rhp;
        tCam = uS - u0Hat;
        pCam = fmad3(ray.o, ray.d, tCam);
        if (VIS) { cn.tCam = tCam; cn.tVRL = tVRL; }
        tCam = fminf(fmaxf(tCam, 0.0f), ray.maxt);               // vrl_struct.h:340-341
        tVRL = fminf(fmaxf(tVRL, 0.0f), vlen);
        invPDF = invPDF_A * invPDF_B;
    }
    V3 dir = pVRL - pCam;
    float d2 = fdot(dir, dir);
    if (d2 == 0.f) return 0.f;                                   // vrl_struct.h:589
    if (isnan(invPDF)) return 0.f;                               // vrl_struct.h:594
    float rd = f_rsqrt(d2);
    float dist = d2 * rd;
    float pf = kInvFourPi * kInvFourPi;
    if (!ISO) {
        V3 dn = dir * rd;
        float rayPF = hg_fast(g, -fdot(dn, ray.d));              // wi = -part.d (quirk 1), wo = dir
        float vrlPF = hg_fast(g, fdot(dn, vd));                  // wi = -vrl.dir, wo = -dir
        pf = rayPF * vrlPF;
        if (pf == 0.f) return 0.f;
    }
    if (VIS) { cn.pCam = pCam; cn.dir = dir; cn.d2 = d2; cn.rd = rd; }   // shadow segment, classified by the caller (vrl_struct.h:509-572)
    // three homogeneous transmittances, medium.h:112-117: exp(-sigma (dist + tCam + tVRL)) in one MUFU.EX2
    float tr = f_ex2(-(max_density * kLog2e) * (dist + tCam + tVRL));
    if (tr == 0.f) return 0.f;
    return (rd * rd) * pf * tr * invPDF;
}

// ---- kernel: one thread per camera segment, VRL tiles through shared memory by TMA (as gather.cu) ----
template <bool UNIFORM, bool RR, bool VIS, bool ISO, bool GUARD>
__global__ void __launch_bounds__(GATHER_THREADS)
gather_fast_kernel(GatherParams p) {
    __shared__ alignas(128) DVrl tile[2][TILE];
    __shared__ alignas(8) uint64_t full_bar[2];

    const uint32_t tid = threadIdx.x;
    const uint64_t seg_i = (uint64_t) blockIdx.x * GATHER_THREADS + tid;
    const bool live = seg_i < p.n_segs;

    FastSeg ray; uint64_t rng0 = 0; uint32_t pixel = 0; float w[3] = { 0.f, 0.f, 0.f };
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
    ray.v = ray.d * ray.maxt;
    ray.c = fdot(ray.v, ray.v);

    const uint32_t n_tiles = (uint32_t) ((p.n_vrls + TILE - 1) / TILE);
    const uint32_t total_tiles = n_tiles * p.q;
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
        tma_bulk_g2s(&tile[buf][0], p.vrls_fast + first, bytes, &full_bar[buf]);
    };
    if (tid == 0) {
        if (total_tiles > 0) issue(0);
        if (total_tiles > 1) issue(1);
    }

    const DMedium *media = p.scene.media;
    const Guard guard = { p.guard_h2, p.guard_z };
    float sum[3] = { 0.f, 0.f, 0.f }, comp[3] = { 0.f, 0.f, 0.f };     // Kahan across tiles
    Pcg32 rng; rng.state = rng0;

    for (uint32_t it = 0; it < total_tiles; ++it) {
        const uint32_t t = it % n_tiles, s = it / n_tiles, buf = it & 1u;
        if (t == 0) {
            rng.state = rng0;
            rng.advance(((uint64_t) s * p.total + p.first_index) * DPE);
        }
        mbar_wait(&full_bar[buf], (it >> 1) & 1u);
        const uint32_t cnt = (uint32_t) min((uint64_t) TILE, p.n_vrls - (uint64_t) t * TILE);
        float pr = 0.f, pg = 0.f, pb = 0.f;                               // plain fp32 inside a tile
        if (live) {
#pragma unroll 2
            for (uint32_t j = 0; j < cnt; ++j) {
                const float4 *vp = reinterpret_cast<const float4 *>(&tile[buf][j]);
                float4 a = vp[0], b = vp[1], c = vp[2];
                V3 vo(a.x, a.y, a.z), vd(b.x, b.y, b.z);
                float vlen = a.w;
                float rrWeight = 1.f; bool survive = true;
                if (RR) {
                    float cc, cv;                                         // EDistanceRoulette, vrl_map.h:171-190
                    closest_fast(vo, vd, vlen, ray, cc, cv);
                    float d2 = squared_norm((ray.o + ray.d * cc) - (vo + vd * cv));
                    rrWeight = fminf(f_rcp(f_sqrt(d2) * p.scale_rr), 1.0f);
                    survive  = rrWeight > rng.next_f32();
                }
                float u1 = rng.next_f32(), u2 = rng.next_f32();
                if (!survive) continue;
                float g = 0.f;
                if (!ISO) g = media[__float_as_uint(b.w)].phase == VRLB_PHASE_HG ? media[__float_as_uint(b.w)].g : 0.f;
                bool slow;
                Conn cn;
                float f = contrib_fast<UNIFORM, false, ISO, GUARD>(p.scene, ray, vo, vd, vlen, c.w, g, u1, u2, guard, slow, cn);
                if (GUARD && slow) {                                      // ill-conditioned eval: exactly-rounded path
                    const DMedium &m = media[__float_as_uint(b.w)];
                    SegRegs er; er.o = ray.o; er.d = ray.d; er.maxt = ray.maxt;
                    f = eval_contrib<UNIFORM, false>(p.scene, er, vo, vd, vlen, m.max_density, m.phase, m.g, u1, u2);
                }
                if (RR) f = f * f_rcp(rrWeight);
                float cr = c.x * f, cg = c.y * f, cb = c.z * f;           // flux pre-multiplied by sigma_s^2 * m_scale
                if (fabsf(cr) + fabsf(cg) + fabsf(cb) < VRLB_INF) { pr += cr; pg += cg; pb += cb; }   // vrl_map.h:162
            }
        }
        float part[3] = { pr, pg, pb };
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            float y = xsub(part[k], comp[k]);
            float tt = xadd(sum[k], y);
            comp[k] = xsub(xsub(tt, sum[k]), y);
            sum[k] = tt;
        }
        __syncthreads();
        if (tid == 0 && it + 2 < total_tiles) issue(it + 2);
    }

    if (live) {
        float inv_q = 1.0f / (float) p.q;
        float lr = sum[0] * inv_q, lg = sum[1] * inv_q, lb = sum[2] * inv_q;
        if (p.seg_out) { p.seg_out[seg_i * 3 + 0] = lr; p.seg_out[seg_i * 3 + 1] = lg; p.seg_out[seg_i * 3 + 2] = lb; }
        if (p.film_rgb) {
            const float4 *sp = reinterpret_cast<const float4 *>(p.segs + seg_i);
            const uint32_t pixel = __float_as_uint(__ldg(sp + 1).w);
            const float4 w = __ldg(sp + 2);
            atomicAdd(p.film_rgb + (size_t) pixel * 3 + 0, lr * w.x);
            atomicAdd(p.film_rgb + (size_t) pixel * 3 + 1, lg * w.y);
            atomicAdd(p.film_rgb + (size_t) pixel * 3 + 2, lb * w.z);
        }
    }
}

// ---- shadow phase: convergent classification + a persistent, compacted trace buffer ------------------------
// The shadow query is the divergent part of an eval: most connection segments miss every occluder's bounding volume,
// some need the finer proxies, ~1 % need the BVH (2..80 node visits).  Done inline with per-lane branches, every lane
// waits for the slowest lane of its warp at every VRL.  Here the VRL loop keeps all 32 lanes on ONE instruction
// stream: the levels of a solid's proxies are gated by warp votes and evaluated predicated (classify_segment), the
// shells cost four compares against precomputed parameter intervals, and a segment that is still undecided is
// appended -- ballot-compacted -- to the warp's TRACE BUFFER together with its contribution.  The buffer persists
// across VRLs and tiles and is traced through the BVHs only when >= 32 rays are waiting (drain_trace, out of line):
// persistent lanes, a lane that finishes a ray pulls the next one; node steps and triangle tests are separated, leaf
// hits go to a leaf queue and are tested 32 at a time.  Contributions are added by their owner lane in buffer order,
// so the sum of a lane is deterministic.  B200 has no RT cores: plain SIMT traversal with warp-level ray compaction.
constexpr int VTILE = 32, NBUF = 2;      // VRL staging of the visibility kernel: per WARP, NBUF tiles of VTILE 64-byte records
constexpr int DRAIN_WINDOW = 16;         // tiles between forced drains of the trace buffer (512 VRLs)
constexpr int TCAP  = 64;                // trace buffer: rays waiting for the BVH (traced when >= 32 are pending)
constexpr int QF    = 11;                // floats per ray: o(3) d(3) dist alive r g b
constexpr float Q_REDO = -2.f;           // dist marker: hit decision inside the graze margins, re-evaluate exactly

struct WarpBuf {
    float    v[QF][TCAP];                // trace buffer, SoA: consecutive rays -> consecutive banks
    uint32_t jg[TCAP];                   // per ray: index of the VRL in this rank's shard (exact re-evaluation)
    uint16_t roots[TCAP];                // per ray: which roots to visit
    uint16_t own[TCAP];                  // per ray: owner lane | query sample index << 5
    uint32_t lq[128];                    // pending leaf entries: (first triangle << 12) | (count << 7) | ray
};

struct RayState {
    V3 o, d, inv; float dist;
    uint32_t cur;        // current wide node
    int sp;              // stack pointer
    uint32_t roots;      // bit 0: main BVH pending, bit 1 + i: shell i pending, bit 9 + i: solid i pending
};

__device__ __forceinline__ bool shell_inside(const DShell &sh, V3 o, V3 e) {
    return o.x > sh.lo[0] && o.x < sh.hi[0] && o.y > sh.lo[1] && o.y < sh.hi[1] && o.z > sh.lo[2] && o.z < sh.hi[2] &&
           e.x > sh.lo[0] && e.x < sh.hi[0] && e.y > sh.lo[1] && e.y < sh.hi[1] && e.z > sh.lo[2] && e.z < sh.hi[2];
}

// Classification of one connection segment o + t dir, t in [0, 1] (dir NOT normalised, inv_d2 = 1 / |dir|^2), with all
// lanes of the warp on one instruction stream.  Returns which roots the segment has to visit -- bit 0 = main BVH (its
// root box is met), bit 1 + i = shell i (an end point is not strictly inside it), bit 9 + i = solid i (undecided by its
// proxies); `occluded` = a closed solid's inner proxy already decides the query.  Per solid (occluder_proxies.cpp):
// bounding sphere missed -> cannot hit it; then, only if some lane of the warp got that far (warp vote), the finer
// proxies the solid carries: inscribed sphere met by a segment that enters from outside -> occluded; outer oriented box
// missed -> cannot hit; inscribed box met -> occluded.  `cs` is the scene as the kernel PARAMETER (constant bank).
__device__ __forceinline__ uint32_t classify_segment(const DScene &cs, const DSolid *solids, const float4 *bsph, uint32_t sflags,
                                                     V3 o, V3 dir, float inv_d2, bool notsafe, bool active, bool &occluded) {
    uint32_t m = 0;
    bool occ = false;
    const V3 e = o + dir;
    const int n_solids = cs.n_solids;
#pragma unroll 1
    for (int i = 0; i < n_solids; ++i) {
        const float4 bs = bsph[i];
        const V3 oc(bs.x - o.x, bs.y - o.y, bs.z - o.z);
        const float t = fminf(fmaxf(fdot(oc, dir) * inv_d2, 0.f), 1.f);   // closest approach of the segment to the centre
        const V3 q = fmad3(oc, dir, -t);
        const bool touch = active && fdot(q, q) <= bs.w;
        if (!__any_sync(0xffffffffu, touch)) continue;
        const uint32_t fl = (sflags >> (3 * i)) & 7u;
        const DSolid &S = solids[i];
        // inner proxies count only for a segment that enters from outside: an end point outside the bounding sphere
        const V3 ec(bs.x - e.x, bs.y - e.y, bs.z - e.z);
        const bool enters = fdot(oc, oc) > bs.w || fdot(ec, ec) > bs.w;
        bool und = touch, hit = false;
        if (fl & VRLB_SOLID_INNER_SPHERE) {
            const V3 oi(S.ci[0] - o.x, S.ci[1] - o.y, S.ci[2] - o.z);
            const float ti = fminf(fmaxf(fdot(oi, dir) * inv_d2, 0.f), 1.f);
            const V3 qi = fmad3(oi, dir, -ti);
            if (und && enters && fdot(qi, qi) <= S.r_in2) { hit = true; und = false; }
        }
        if ((fl & (VRLB_SOLID_OUTER_OBB | VRLB_SOLID_INNER_OBB)) && __any_sync(0xffffffffu, und)) {
            const V3 r(o.x - S.oc[0], o.y - S.oc[1], o.z - S.oc[2]);
            const V3 ol(fdot(V3(S.ax[0], S.ax[1], S.ax[2]), r), fdot(V3(S.ax[3], S.ax[4], S.ax[5]), r), fdot(V3(S.ax[6], S.ax[7], S.ax[8]), r));
            const V3 dl(fdot(V3(S.ax[0], S.ax[1], S.ax[2]), dir), fdot(V3(S.ax[3], S.ax[4], S.ax[5]), dir), fdot(V3(S.ax[6], S.ax[7], S.ax[8]), dir));
            const V3 il(f_rcp(dl.x), f_rcp(dl.y), f_rcp(dl.z));
            bool enters_b = enters;
            if (fl & VRLB_SOLID_OUTER_OBB) {
                if (!slab2(-S.h_out[0], -S.h_out[1], -S.h_out[2], S.h_out[0], S.h_out[1], S.h_out[2], il, ol, 0.f, 1.f)) und = false;
                // an end point outside the outer box is outside the solid as well (importance sampling favours short
                // connections: both ends inside the bounding sphere of a box-like solid is common)
                const V3 el = ol + dl;
                enters_b = enters_b || fabsf(ol.x) > S.h_out[0] || fabsf(ol.y) > S.h_out[1] || fabsf(ol.z) > S.h_out[2] ||
                           fabsf(el.x) > S.h_out[0] || fabsf(el.y) > S.h_out[1] || fabsf(el.z) > S.h_out[2];
            }
            if (fl & VRLB_SOLID_INNER_OBB) {
                const float ax = (-S.h_in[0] - ol.x) * il.x, bx = (S.h_in[0] - ol.x) * il.x;
                const float ay = (-S.h_in[1] - ol.y) * il.y, by = (S.h_in[1] - ol.y) * il.y;
                const float az = (-S.h_in[2] - ol.z) * il.z, bz = (S.h_in[2] - ol.z) * il.z;
                const float t0 = fmaxf(fmaxf(0.f, fminf(ax, bx)), fmaxf(fminf(ay, by), fminf(az, bz)));
                const float t1 = fminf(fminf(1.f, fmaxf(ax, bx)), fminf(fmaxf(ay, by), fmaxf(az, bz)));
                if (und && enters_b && t0 <= t1) { hit = true; und = false; }   // strict: the box is already pulled in by the margin
            }
        }
        occ = occ || hit;
        if (und) m |= 512u << i;
    }
    if (cs.main_root != 0x80000000u) {
        const V3 inv(f_rcp(dir.x), f_rcp(dir.y), f_rcp(dir.z));
        if (active && slab2(cs.main_lo[0], cs.main_lo[1], cs.main_lo[2], cs.main_hi[0], cs.main_hi[1], cs.main_hi[2], inv, o, 0.f, 1.f)) m |= 1u;
    }
    // shells: only a segment with an end point outside the safe box has to look at the individual shells
    if (__any_sync(0xffffffffu, notsafe && active)) {
        if (notsafe && active)
            for (int i = 0; i < cs.n_shells; ++i) if (!shell_inside(cs.shells[i], o, e)) m |= 2u << i;
    }
    occluded = occ;
    return occ ? 0u : m;
}
__device__ __forceinline__ bool next_root(const DScene &s, RayState &r, uint32_t &root) {
    if (r.roots == 0) return false;
    int i = __ffs((int) r.roots) - 1;
    r.roots &= r.roots - 1;
    root = i == 0 ? s.main_root : (i <= VRLB_MAX_SHELLS ? s.shells[i - 1].root : s.solids[i - 1 - VRLB_MAX_SHELLS].root);
    return true;
}

// Tracing.  `cnt` rays in the buffer; on return q.v[7][ray] = 0 for occluded segments and q.v[6][ray] = Q_REDO for
// segments whose hit decision fell inside the graze margins.  Two kinds of work, each done by all 32 lanes at
// once: node steps (one DWNode visit per active lane; lanes that finish a ray pull the next one) and
// leaf entries (ray, leaf) that the node steps append to q.lq -- as soon as 32 are pending every lane tests one
// leaf's triangles.  Triangle tests at 1-3 active lanes were what held the inline traversal at 10 of 32 lanes.
__device__ __forceinline__ void process_leaves(const DScene &s, WarpBuf &q, int first, int n, unsigned lane) {
    if ((int) lane < n) {
        uint32_t ent = q.lq[first + lane];
        int slot = (int) (ent & 127u);
        uint32_t cntl = (ent >> 7) & 31u, tri0 = (ent & 0x7fffffffu) >> 12;
        V3 o(q.v[0][slot], q.v[1][slot], q.v[2][slot]), d(q.v[3][slot], q.v[4][slot], q.v[5][slot]);
        float dist = q.v[6][slot];
        if (dist > 0.f) {                                     // not already marked for exact re-evaluation
            int res = 0;
            for (uint32_t i = tri0; i < tri0 + cntl; ++i) {
                int h = intersect_tri_graze(s.stris[i], o, d, 0.f, dist);
                res |= h;
                if (h == 1) break;
            }
            if (res & 1) q.v[7][slot] = 0.f;                  // certain hit: contribution is 0 (benign same-value races)
            else if (res & 2) q.v[6][slot] = Q_REDO;
        }
    }
}

// Traversal stack: build_shadow_accel bounds the depth of every BVH it emits (bvh_build.cpp, kMaxShadowDepth) so
// that depth + the single-leaf roots pushed up front stays below this.
constexpr int TRACE_STACK = 48;
template <bool WATCHDOG>
__device__ __forceinline__ void trace_queue(const DScene &s, WarpBuf &q, int cnt, uint32_t *wd) {
    const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    int head = 0, slot = 0, lqn = 0;
    bool active = false;
    RayState r;
    uint32_t stack[TRACE_STACK];
    r.sp = 0; r.cur = 0; r.roots = 0; r.dist = 0.f;
    for (uint32_t steps = 0; ; ++steps) {
        if (WATCHDOG && steps > (1u << 22)) {                      // a batch takes a few hundred steps
            if (wd && atomicCAS(wd, 0u, 2u) == 0u) { wd[1] = blockIdx.x; wd[2] = threadIdx.x; wd[3] = (uint32_t) cnt; wd[4] = (uint32_t) head; wd[5] = r.cur; wd[6] = (uint32_t) r.sp; __threadfence_system(); }
            __trap();
        }
        // (1) leaf work: a full warp's worth pending
        while (lqn >= 32) {
            lqn -= 32;
            __syncwarp();
            process_leaves(s, q, lqn, 32, lane);
            __syncwarp();
        }
        // (2) retire rays found occluded, refill idle lanes
        if (active && q.v[7][slot] == 0.f) active = false;
        unsigned idle = __ballot_sync(0xffffffffu, !active);
        uint32_t leaf0 = 0, leaf1 = 0;                        // leaf entries produced by this step
        if (idle) {
            int avail = cnt - head;
            if (avail > 0) {
                int rank = __popc(idle & lt);
                if (!active && rank < avail) {
                    slot = head + rank;
                    r.o = V3(q.v[0][slot], q.v[1][slot], q.v[2][slot]);
                    r.d = V3(q.v[3][slot], q.v[4][slot], q.v[5][slot]);
                    r.dist = q.v[6][slot];
                    r.inv = V3(f_rcp(r.d.x), f_rcp(r.d.y), f_rcp(r.d.z));      // conservative slabs: approximate reciprocal is enough
                    r.sp = 0; r.roots = q.roots[slot];
                    uint32_t root;
                    active = false;
                    while (next_root(s, r, root)) {
                        if (root & 0x80000000u) { stack[r.sp++] = root; continue; }      // single-leaf component
                        r.cur = root; active = true; break;
                    }
                    if (!active && r.sp) { r.cur = stack[--r.sp]; active = true; }
                }
                head += min(__popc(idle), avail);
            } else if (idle == 0xffffffffu) break;
        }
        // (3) one node step per active lane
        if (active) {
            if (r.cur & 0x80000000u) {                        // a leaf reference popped from the stack (root leaves)
                leaf0 = r.cur;
                bool more = true;
                if (r.sp) r.cur = stack[--r.sp];
                else {
                    uint32_t root; more = false;
                    while (next_root(s, r, root)) {
                        if (root & 0x80000000u) { stack[r.sp++] = root; continue; }
                        r.cur = root; more = true; break;
                    }
                    if (!more && r.sp) { r.cur = stack[--r.sp]; more = true; }
                }
                active = more;
            } else {
                const float4 *np = reinterpret_cast<const float4 *>(s.wnodes + r.cur);
                float4 a = __ldg(np), b = __ldg(np + 1), c = __ldg(np + 2), e = __ldg(np + 3);
                bool h0 = slab2(a.x, a.y, a.z, a.w, b.x, b.y, r.inv, r.o, 0.f, r.dist);
                bool h1 = slab2(b.z, b.w, c.x, c.y, c.z, c.w, r.inv, r.o, 0.f, r.dist);
                uint32_t c0 = __float_as_uint(e.x), c1 = __float_as_uint(e.y);
                if (h0 && (c0 & 0x80000000u)) { leaf0 = c0; h0 = false; }
                if (h1 && (c1 & 0x80000000u)) { leaf1 = c1; h1 = false; }
                bool more = true;
                if (h0) { if (h1) stack[r.sp++] = c1; r.cur = c0; }
                else if (h1) r.cur = c1;
                else if (r.sp) r.cur = stack[--r.sp];
                else {
                    uint32_t root; more = false;
                    while (next_root(s, r, root)) {
                        if (root & 0x80000000u) { stack[r.sp++] = root; continue; }
                        r.cur = root; more = true; break;
                    }
                    if (!more && r.sp) { r.cur = stack[--r.sp]; more = true; }
                }
                active = more;
            }
        }
        // append leaf entries: leaf reference | ray
        {
            unsigned m0 = __ballot_sync(0xffffffffu, leaf0 != 0);
            if (leaf0) q.lq[lqn + __popc(m0 & lt)] = leaf0 | (uint32_t) slot;
            lqn += __popc(m0);
            unsigned m1 = __ballot_sync(0xffffffffu, leaf1 != 0);
            if (leaf1) q.lq[lqn + __popc(m1 & lt)] = leaf1 | (uint32_t) slot;
            lqn += __popc(m1);
        }
    }
    __syncwarp();
    if (lqn > 0) process_leaves(s, q, 0, lqn, lane);
}

// the exactly-rounded eval, out of line: it runs for < 1 % of the evals (instruction-cache footprint).
// exact_eval_seg: without the shadow query; seg[0..6] = pCam, direction, length of the connection
template <bool UNIFORM>
__device__ __noinline__ float exact_eval_seg(const DScene &scene, const SegRegs &ray, V3 vo, V3 vd, float vlen,
                                             float max_density, int phase, float g, float u1, float u2, float *seg) {
    seg[6] = 0.f;
    return eval_contrib<UNIFORM, false>(scene, ray, vo, vd, vlen, max_density, phase, g, u1, u2, nullptr, 0, nullptr, seg);
}
template <bool UNIFORM>
__device__ __noinline__ float exact_eval_vis(const DScene &scene, const SegRegs &ray, V3 vo, V3 vd, float vlen,
                                             float max_density, int phase, float g, float u1, float u2) {
    return eval_contrib<UNIFORM, true>(scene, ray, vo, vd, vlen, max_density, phase, g, u1, u2);
}

// Watchdog for the ring (debug builds, -DVRLB_WATCHDOG): a wait that outlasts ~2 s (a tile takes ~0.1 ms) writes the
// CTA's ring state to pinned host memory and traps, so that a lost hand-over surfaces as an error with evidence
// instead of a silent hang.  Off in production: the clock reads were 3 % of the issued instructions.
#ifdef VRLB_WATCHDOG
constexpr bool kWatchdog = true;
#else
constexpr bool kWatchdog = false;
#endif
__device__ __noinline__ void ring_watchdog_trip(uint32_t *wd, uint32_t code, uint32_t it, uint32_t buf,
                                                const uint32_t *warp_it, const uint64_t *bars) {
    if (wd && atomicCAS(wd, 0u, code) == 0u) {
        wd[1] = blockIdx.x; wd[2] = threadIdx.x; wd[3] = it; wd[4] = buf;
        for (int b = 0; b < 8; ++b) { wd[5 + 2 * b] = (uint32_t) bars[b]; wd[6 + 2 * b] = (uint32_t) (bars[b] >> 32); }   // full[0..3], empty[0..3]
        for (int w = 0; w < 8; ++w) wd[21 + w] = warp_it[w];
        __threadfence_system();
    }
    __trap();
}
__device__ __forceinline__ void mbar_wait_wd(uint64_t *bar, uint32_t parity, uint32_t *wd, uint32_t code, uint32_t it, uint32_t buf,
                                             const uint32_t *warp_it, const uint64_t *bars) {
    uint32_t ok = 0;
    long long t0 = 0;
    for (uint32_t spins = 0; ; ++spins) {
        // try_wait with a suspend-time hint: a waiting warp sleeps in the barrier unit instead of burning issue slots
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\nselp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity), "r"(20000u) : "memory");
        if (ok) return;
        // a warp that ran ahead of the producer waits here for a tile (~0.1 ms of work): sleep instead of spinning
        __nanosleep(spins < 2 ? 2000u : 10000u);               // a tile is ~150 us of work for a warp
        if (kWatchdog) {
            if (spins == 4) t0 = clock64();
            else if (spins > 4 && (spins & 15u) == 0 && clock64() - t0 > 4000000000LL) ring_watchdog_trip(wd, code, it, buf, warp_it, bars);
        }
    }
}

struct Acc3 { float r, g, b; };

// What a drain needs beyond the warp's buffer, in one struct so the out-of-line call passes one pointer.
struct DrainCtx {
    const DScene  *S;
    const DMedium *media;
    const DVrlV   *vrls;         // this rank's shard (global memory): exact re-evaluation of traced rays
    uint64_t first_index, total; // stream addressing (GatherParams)
    float    scale_rr;
    uint32_t *wd;
};

// Trace the n rays waiting in the buffer and let every lane add what survives of its own (buffer order).  Out of
// line: it runs once per ~100 VRLs and keeps the traversal's registers out of the shading loop.
template <bool UNIFORM, bool RR>
__device__ __noinline__ Acc3 drain_trace(const DrainCtx &fc, WarpBuf &q, int n, const SegRegs &er, uint64_t rng0) {
    const unsigned lane = threadIdx.x & 31u;
    constexpr uint32_t DPE = RR ? 3u : 2u;
    Acc3 acc = { 0.f, 0.f, 0.f };
    __syncwarp();
    trace_queue<kWatchdog>(*fc.S, q, n, fc.wd);
    __syncwarp();
    // lane-parallel: occluded rays contribute nothing; graze decisions are re-evaluated exactly by their owner below
    bool any_redo = false;
    for (int i = (int) lane; i < n; i += 32) {
        const bool redo = q.v[6][i] == Q_REDO;
        if (redo || q.v[7][i] == 0.f) { q.v[8][i] = 0.f; q.v[9][i] = 0.f; q.v[10][i] = 0.f; }
        any_redo = any_redo || redo;
    }
    any_redo = __any_sync(0xffffffffu, any_redo);
    __syncwarp();
    for (int i = 0; i < n; ++i) {                  // uniform loop, broadcast reads: the owner adds
        const bool mine = (q.own[i] & 31u) == lane;
        const float cr = q.v[8][i], cg = q.v[9][i], cb = q.v[10][i];
        if (mine && fabsf(cr) + fabsf(cg) + fabsf(cb) < VRLB_INF) { acc.r += cr; acc.g += cg; acc.b += cb; }   // vrl_map.h:162
    }
    if (any_redo) {
        for (int i = 0; i < n; ++i) {
            const uint32_t own = q.own[i];
            if ((own & 31u) != lane || q.v[6][i] != Q_REDO) continue;
            // graze margin: exactly-rounded eval + exact trace
            const uint32_t jg = q.jg[i], s = own >> 5;
            const float4 *vp = reinterpret_cast<const float4 *>(fc.vrls + jg);
            float4 a = __ldg(vp), b = __ldg(vp + 1), c = __ldg(vp + 2);
            const DMedium &m = fc.media[__float_as_uint(b.w)];
            Pcg32 rr; rr.state = rng0;
            rr.advance(((uint64_t) s * fc.total + fc.first_index + jg) * DPE + (RR ? 1u : 0u));
            float ru1 = rr.next_f32(), ru2 = rr.next_f32();
            float fe = exact_eval_vis<UNIFORM>(*fc.S, er, V3(a.x, a.y, a.z), V3(b.x, b.y, b.z), a.w, m.max_density, m.phase, m.g, ru1, ru2);
            if (RR) {
                FastSegC ray; ray.o = er.o; ray.d = er.d; ray.maxt = er.maxt;
                float cc, cv;
                closest_fast(V3(a.x, a.y, a.z), V3(b.x, b.y, b.z), a.w, ray, cc, cv);
                V3 dd = fmad3(ray.o, ray.d, cc) - fmad3(V3(a.x, a.y, a.z), V3(b.x, b.y, b.z), cv);
                fe = fe * f_rcp(fminf(f_rcp(f_sqrt(fdot(dd, dd)) * fc.scale_rr), 1.0f));
            }
            const float cr = c.x * fe, cg = c.y * fe, cb = c.z * fe;
            if (fabsf(cr) + fabsf(cg) + fabsf(cb) < VRLB_INF) { acc.r += cr; acc.g += cg; acc.b += cb; }
        }
    }
    __syncwarp();
    return acc;
}

// Parameter interval (lo, hi) of the points o + t d that lie strictly inside the box `safe` (lo[3], hi[3]).
__device__ __forceinline__ void safe_interval(V3 o, V3 d, const float *safe, float &lo, float &hi) {
    lo = -VRLB_INF; hi = VRLB_INF;
    const float oo[3] = { o.x, o.y, o.z }, dd[3] = { d.x, d.y, d.z };
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        if (fabsf(dd[k]) > 1e-12f) {
            const float inv = 1.0f / dd[k];
            const float t0 = (safe[k] - oo[k]) * inv, t1 = (safe[3 + k] - oo[k]) * inv;
            lo = fmaxf(lo, fminf(t0, t1)); hi = fminf(hi, fmaxf(t0, t1));
        } else if (!(oo[k] > safe[k] && oo[k] < safe[3 + k])) { lo = VRLB_INF; hi = -VRLB_INF; }
    }
}

// ---- kernel with visibility ----
#ifndef VRLB_VIS_THREADS
#define VRLB_VIS_THREADS 256
#endif
#ifndef VRLB_VIS_MINBLOCKS
#define VRLB_VIS_MINBLOCKS 3
#endif
constexpr int VIS_THREADS = VRLB_VIS_THREADS, VIS_MINBLOCKS = VRLB_VIS_MINBLOCKS;
template <bool UNIFORM, bool RR, bool ISO, bool GUARD>
__global__ void __launch_bounds__(VIS_THREADS, VIS_MINBLOCKS)
gather_fast_vis_kernel(GatherParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    // VRL tiles are staged PER WARP (TMA bulk copies into the warp's own double buffer, completion on the warp's own
    // mbarriers): warps of a CTA do different amounts of shadow work, and a CTA-wide ring made the fast ones wait for the
    // slow ones (issue slots 66 % busy on a frame without solids, profiles/r02d); the tiles come from L2 either way.
    DVrlV (*tile)[VTILE] = reinterpret_cast<DVrlV (*)[VTILE]>(smem_raw) + (threadIdx.x >> 5) * NBUF;  // [warp][NBUF][VTILE]
    WarpBuf *bufs = reinterpret_cast<WarpBuf *>(smem_raw + (VIS_THREADS / 32) * NBUF * VTILE * sizeof(DVrlV));           // [warps]
    __shared__ alignas(8) uint64_t ring_bar[(VIS_THREADS / 32) * NBUF];  // per warp: "full" (TMA complete_tx) of its two buffers
    uint64_t *full_bar = ring_bar + (threadIdx.x >> 5) * NBUF;
    __shared__ uint32_t warp_it[VIS_THREADS / 32];   // watchdog evidence: the tile each warp is on
    __shared__ DSolid sol[VRLB_MAX_SOLIDS];             // the solids' proxies
    __shared__ float4 bsph[VRLB_MAX_SOLIDS];            // their bounding spheres (centre, r^2): the first test of every eval, one LDS.128
    __shared__ uint32_t s_sflags;                       // proxy flags, 3 bits per solid
    __shared__ DScene s_scene;                          // the out-of-line helpers take the scene by reference: shared, not a local copy
    __shared__ DrainCtx s_dc;
    if (threadIdx.x == 0) {
        s_scene = p.scene;
        s_dc.S = &s_scene; s_dc.media = p.scene.media; s_dc.vrls = p.vrls_vis;
        s_dc.first_index = p.first_index; s_dc.total = p.total; s_dc.scale_rr = p.scale_rr; s_dc.wd = p.wd;
    }
    __syncthreads();
    const DScene &S = s_scene;

    const uint32_t tid = threadIdx.x, lane = tid & 31u, lt = (1u << lane) - 1u;
    WarpBuf &q = bufs[tid >> 5];
    const uint64_t seg_i = (uint64_t) blockIdx.x * VIS_THREADS + tid;
    const bool live = seg_i < p.n_segs;

    // Only what every eval needs stays in registers (origin, direction, maxt, the safe interval, the stream); pixel,
    // weight and the stream's start are re-read from the segment record where they are used.
    FastSegC ray;
    if (live) {
        const float4 *sp = reinterpret_cast<const float4 *>(p.segs + seg_i);
        float4 r0 = __ldg(sp), r1 = __ldg(sp + 1);
        ray.o = V3(r0.x, r0.y, r0.z); ray.maxt = r0.w;
        ray.d = V3(r1.x, r1.y, r1.z);
    } else {
        ray.o = V3(0.f); ray.d = V3(0.f, 0.f, 1.f); ray.maxt = 1.f;
    }
    SegRegs er; er.o = ray.o; er.d = ray.d; er.maxt = ray.maxt;
    float t_lo, t_hi;                                   // camera-part parameters strictly inside the shells' safe box
    safe_interval(ray.o, ray.d, p.safe, t_lo, t_hi);

    const uint32_t n_tiles = (uint32_t) ((p.n_vrls + VTILE - 1) / VTILE);
    const uint32_t total_tiles = n_tiles * p.q;
    constexpr uint32_t DPE = RR ? 3u : 2u;

    for (uint32_t i = tid; i < (uint32_t) S.n_solids * (sizeof(DSolid) / 4); i += VIS_THREADS)
        reinterpret_cast<uint32_t *>(sol)[i] = reinterpret_cast<const uint32_t *>(S.solids)[i];
    if (tid < (uint32_t) S.n_solids) bsph[tid] = make_float4(S.solids[tid].c[0], S.solids[tid].c[1], S.solids[tid].c[2], S.solids[tid].r_out2);
    if (tid == 32) {
        uint32_t f = 0;
        for (int i = 0; i < S.n_solids; ++i) f |= (S.solids[i].flags & 7u) << (3 * i);
        s_sflags = f;
    }
    if (lane == 0) {
        for (int b = 0; b < NBUF; ++b) mbar_init(&full_bar[b], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto issue = [&](uint32_t it) {
        uint32_t t = it % n_tiles, buf = it % NBUF;
        uint64_t first = (uint64_t) t * VTILE;
        uint32_t cnt = (uint32_t) min((uint64_t) VTILE, p.n_vrls - first);
        uint32_t bytes = cnt * (uint32_t) sizeof(DVrlV);
        mbar_expect_tx(&full_bar[buf], bytes);
        tma_bulk_g2s(&tile[buf][0], p.vrls_vis + first, bytes, &full_bar[buf]);
    };
    if (lane == 0 && total_tiles > 0) issue(0);

    const DMedium *media = S.media;
    const Guard guard = { p.guard_h2, p.guard_z };
    const uint32_t sflags = s_sflags;
    float sum[3] = { 0.f, 0.f, 0.f }, comp[3] = { 0.f, 0.f, 0.f };     // Kahan across tiles
    Pcg32 rng; rng.state = 0;
    int tb_n = 0;                                          // rays in the trace buffer (warp-uniform)
    auto kahan_add = [&](float r, float g, float b) {
        const float part[3] = { r, g, b };
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            float y = xsub(part[k], comp[k]);
            float tt = xadd(sum[k], y);
            comp[k] = xsub(xsub(tt, sum[k]), y);
            sum[k] = tt;
        }
    };
    auto rng_start = [&]() { return live ? __ldg(reinterpret_cast<const ulonglong2 *>(p.segs + seg_i) + 3).x : 0ull; };

    for (uint32_t it = 0; it < total_tiles; ++it) {
        const uint32_t t = it % n_tiles, s = it / n_tiles, buf = it % NBUF;
        if (t == 0) {
            rng.state = rng_start();
            rng.advance(((uint64_t) s * p.total + p.first_index) * DPE);
        }
        if (kWatchdog && lane == 0) warp_it[tid >> 5] = it;
        // the next tile goes into the other buffer, which the warp finished reading at the end of the previous iteration
        __syncwarp();
        if (lane == 0 && it + 1 < total_tiles) issue(it + 1);
        mbar_wait_wd(&full_bar[buf], (it / NBUF) & 1u, p.wd, 1u, it, buf, warp_it, ring_bar);
        const uint32_t cnt = (uint32_t) min((uint64_t) VTILE, p.n_vrls - (uint64_t) t * VTILE);
        float pr = 0.f, pg = 0.f, pb = 0.f;
        // The (out-of-line) drain is called from the OUTER of two loops: the inner loop shades VRLs until 32 rays wait in the
        // trace buffer and contains no call except the rare exactly-rounded eval, so its state stays in registers.
        for (uint32_t j = 0; ; ) {
          for (; j < cnt; ++j) {
            // shade VRL j for the warp's 32 camera segments (converged); only undecided shadow segments are parked
            const float4 *vp = reinterpret_cast<const float4 *>(&tile[buf][j]);
            float4 a = vp[0], b = vp[1];
            V3 vo(a.x, a.y, a.z), vd(b.x, b.y, b.z);
            float vlen = a.w;
            const float max_density = reinterpret_cast<const float *>(vp)[11];   // the flux is read after the eval (registers)
            float rrWeight = 1.f; bool survive = true;           // lanes past the last segment shade a dummy segment, never written
            if (RR) {
                float cc, cv;                                     // EDistanceRoulette, vrl_map.h:171-190
                closest_fast(vo, vd, vlen, ray, cc, cv);
                V3 dd = fmad3(ray.o, ray.d, cc) - fmad3(vo, vd, cv);
                rrWeight = fminf(f_rcp(f_sqrt(fdot(dd, dd)) * p.scale_rr), 1.0f);
                survive  = rrWeight > rng.next_f32();
            }
            float u1 = rng.next_f32(), u2 = rng.next_f32();
            float f = 0.f; bool notsafe = false; Conn cn;
            cn.pCam = V3(0.f); cn.dir = V3(0.f, 0.f, 1.f); cn.d2 = 1.f; cn.rd = 1.f; cn.tCam = 0.f; cn.tVRL = 0.f;
            if (survive) {
                float g = 0.f;
                if (!ISO) g = media[__float_as_uint(b.w)].phase == VRLB_PHASE_HG ? media[__float_as_uint(b.w)].g : 0.f;
                bool slow;
                f = contrib_fast<UNIFORM, true, ISO, GUARD>(S, ray, vo, vd, vlen, max_density, g, u1, u2, guard, slow, cn);
                if (GUARD && slow) {                              // ill-conditioned eval: exactly-rounded arithmetic; its
                    const DMedium &m = media[__float_as_uint(b.w)];       // connection segment is classified like any other
                    float seg[7];
                    f = exact_eval_seg<UNIFORM>(S, er, vo, vd, vlen, m.max_density, m.phase, m.g, u1, u2, seg);
                    const float dist = seg[6];
                    if (!(dist > 0.f)) f = 0.f;
                    cn.pCam = V3(seg[0], seg[1], seg[2]); cn.dir = V3(seg[3], seg[4], seg[5]) * dist;
                    cn.d2 = dist * dist; cn.rd = f_rcp(dist);
                    notsafe = true;                               // no parameters: the shells look at the end points
                } else {
                    // shells: both end points strictly inside the safe box <=> both parameters inside their intervals
                    const float2 vi = *reinterpret_cast<const float2 *>(vp + 3);
                    notsafe = !(cn.tCam > t_lo && cn.tCam < t_hi && cn.tVRL > vi.x && cn.tVRL < vi.y);
                }
                if (RR) f = f * f_rcp(rrWeight);
            }
            bool occluded;
            const uint32_t roots = classify_segment(p.scene, sol, bsph, sflags, cn.pCam, cn.dir, cn.rd * cn.rd, notsafe, f != 0.f, occluded);
            if (occluded) f = 0.f;                                // an inner proxy of a closed solid decides it
            float4 c;                                             // flux pre-multiplied by sigma_s^2 * m_scale; read here, not
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"    // at the top: four registers fewer across the eval
                         : "=f"(c.x), "=f"(c.y), "=f"(c.z), "=f"(c.w) : "r"(smem_u32(vp + 2)));
            const float cr = c.x * f, cg = c.y * f, cb = c.z * f;
            const bool undec = roots != 0;
            const unsigned um = __ballot_sync(0xffffffffu, undec);
            if (undec) {                                          // park the segment: compacted, (VRL, lane) order
                const int r = tb_n + __popc(um & lt);
                const float dist = cn.d2 * cn.rd;
                q.v[0][r] = cn.pCam.x; q.v[1][r] = cn.pCam.y; q.v[2][r] = cn.pCam.z;
                q.v[3][r] = cn.dir.x * cn.rd; q.v[4][r] = cn.dir.y * cn.rd; q.v[5][r] = cn.dir.z * cn.rd;
                q.v[6][r] = dist; q.v[7][r] = 1.f;
                q.v[8][r] = cr; q.v[9][r] = cg; q.v[10][r] = cb;
                q.roots[r] = (uint16_t) roots; q.own[r] = (uint16_t) (lane | (s << 5)); q.jg[r] = t * VTILE + j;
            } else if (fabsf(cr) + fabsf(cg) + fabsf(cb) < VRLB_INF) { pr += cr; pg += cg; pb += cb; }   // vrl_map.h:162
            tb_n += __popc(um);
            if (tb_n >= 32) { ++j; break; }
          }
          // ... and at the end of every DRAIN_WINDOW-th tile (GLOBAL tile index) whatever is waiting: VRL shards cut at
          // a window boundary then form the same trace batches as the whole set and reproduce its sums bit for bit
          const bool window_end = j >= cnt && (((p.first_index / VTILE + t + 1) % DRAIN_WINDOW) == 0 || t + 1 == n_tiles);
          if (tb_n >= 32 || (window_end && tb_n > 0)) {
              // the tile's partial sums go into the Kahan accumulators first: nothing of the inner loop's state is live
              // across the call except what the accumulators hold
              kahan_add(pr, pg, pb);
              pr = pg = pb = 0.f;
              Acc3 r = drain_trace<UNIFORM, RR>(s_dc, q, tb_n, er, rng_start());
              kahan_add(r.r, r.g, r.b);
              tb_n = 0;
          }
          if (j >= cnt) break;
        }
        kahan_add(pr, pg, pb);
        __syncwarp();
    }
    // rays still waiting in the trace buffer
    if (tb_n > 0) {
        Acc3 r = drain_trace<UNIFORM, RR>(s_dc, q, tb_n, er, rng_start());
        kahan_add(r.r, r.g, r.b);
    }

    if (live) {
        float inv_q = 1.0f / (float) p.q;
        float lr = sum[0] * inv_q, lg = sum[1] * inv_q, lb = sum[2] * inv_q;
        if (p.seg_out) { p.seg_out[seg_i * 3 + 0] = lr; p.seg_out[seg_i * 3 + 1] = lg; p.seg_out[seg_i * 3 + 2] = lb; }
        if (p.film_rgb) {
            const float4 *sp = reinterpret_cast<const float4 *>(p.segs + seg_i);
            const uint32_t pixel = __float_as_uint(__ldg(sp + 1).w);
            const float4 w = __ldg(sp + 2);
            atomicAdd(p.film_rgb + (size_t) pixel * 3 + 0, lr * w.x);
            atomicAdd(p.film_rgb + (size_t) pixel * 3 + 1, lg * w.y);
            atomicAdd(p.film_rgb + (size_t) pixel * 3 + 2, lb * w.z);
        }
    }
}

// ---- VRL records for the fast kernel: flux * sigma_s(medium)^2 * m_scale, extinction in the pad lane ----
__global__ void prep_fast_vrls_kernel(const DVrl *in, uint64_t n, const DMedium *media, float scale, DVrl *out) {
    uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    DVrl v = in[j];
    const DMedium m = media[v.medium];
    v.fr = v.fr * m.sigma_s[0] * m.sigma_s[0] * scale;      // vrl_struct.h:652-659, vrl_map.h:160
    v.fg = v.fg * m.sigma_s[1] * m.sigma_s[1] * scale;
    v.fb = v.fb * m.sigma_s[2] * m.sigma_s[2] * scale;
    v.pad = m.max_density;
    out[j] = v;
}
cudaError_t launch_prep_fast_vrls(const DVrl *in, uint64_t n, const DMedium *media, float scale, DVrl *out, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    prep_fast_vrls_kernel<<<(uint32_t) ((n + 255) / 256), 256, 0, stream>>>(in, n, media, scale, out);
    return cudaGetLastError();
}

// ---- 64-byte records of the visibility kernel: the same three rows + the VRL's parameter interval inside the shells'
//      safe box (gather_fast_vis_kernel tests "end point strictly inside every shell" with two compares) ----
struct SafeBox { float v[6]; };
__global__ void prep_vis_vrls_kernel(const DVrl *in, uint64_t n, const DMedium *media, float scale, SafeBox safe, DVrlV *out) {
    uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const DVrl v = in[j];
    const DMedium m = media[v.medium];
    DVrlV o;
    o.ox = v.ox; o.oy = v.oy; o.oz = v.oz; o.length = v.length;
    o.dx = v.dx; o.dy = v.dy; o.dz = v.dz; o.medium = v.medium;
    o.fr = v.fr * m.sigma_s[0] * m.sigma_s[0] * scale;      // vrl_struct.h:652-659, vrl_map.h:160
    o.fg = v.fg * m.sigma_s[1] * m.sigma_s[1] * scale;
    o.fb = v.fb * m.sigma_s[2] * m.sigma_s[2] * scale;
    o.ext = m.max_density;
    safe_interval(V3(v.ox, v.oy, v.oz), V3(v.dx, v.dy, v.dz), safe.v, o.v_lo, o.v_hi);
    o.pad0 = o.pad1 = 0u;
    out[j] = o;
}
void shell_safe_box(const DScene &scene, float safe[6]) {
    for (int k = 0; k < 3; ++k) { safe[k] = -INFINITY; safe[3 + k] = INFINITY; }
    for (int i = 0; i < scene.n_shells; ++i)
        for (int k = 0; k < 3; ++k) { safe[k] = fmaxf(safe[k], scene.shells[i].lo[k]); safe[3 + k] = fminf(safe[3 + k], scene.shells[i].hi[k]); }
}
cudaError_t launch_prep_vis_vrls(const DVrl *in, uint64_t n, const DScene &scene, float scale, DVrlV *out, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    SafeBox sb; shell_safe_box(scene, sb.v);
    prep_vis_vrls_kernel<<<(uint32_t) ((n + 255) / 256), 256, 0, stream>>>(in, n, scene.media, scale, sb, out);
    return cudaGetLastError();
}

template <bool U, bool R, bool I, bool P>
static cudaError_t launch_f(const GatherParams &p, cudaStream_t stream) {
    uint32_t grid = (uint32_t) ((p.n_segs + GATHER_THREADS - 1) / GATHER_THREADS);
    if (grid == 0 || p.n_vrls == 0) return cudaSuccess;
    GatherParams pp = p;
    pp.guard_h2 = p.guard_h * p.guard_h;
    gather_fast_kernel<U, R, false, I, P><<<grid, GATHER_THREADS, 0, stream>>>(pp);
    return cudaGetLastError();
}
template <bool U, bool R, bool I, bool P>
static cudaError_t launch_fv(const GatherParams &p, cudaStream_t stream) {
    uint32_t grid = (uint32_t) ((p.n_segs + VIS_THREADS - 1) / VIS_THREADS);
    if (grid == 0 || p.n_vrls == 0) return cudaSuccess;
    constexpr size_t smem = (VIS_THREADS / 32) * (NBUF * VTILE * sizeof(DVrlV) + sizeof(WarpBuf));
    GatherParams pp = p;                   // intersection of the shell boxes: a segment with both ends strictly inside it meets no shell
    pp.guard_h2 = p.guard_h * p.guard_h;
    shell_safe_box(p.scene, pp.safe);
    // per device, so set before every launch (it is cheap): a process may hold contexts on several GPUs
    cudaError_t e = cudaFuncSetAttribute(gather_fast_vis_kernel<U, R, I, P>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    if (e != cudaSuccess) return e;
    gather_fast_vis_kernel<U, R, I, P><<<grid, VIS_THREADS, smem, stream>>>(pp);
    return cudaGetLastError();
}
template <bool U, bool R, bool V>
static cudaError_t launch_f2(const GatherParams &p, bool iso, bool guard, cudaStream_t stream) {
    if (V) {
        if (iso) return guard ? launch_fv<U, R, true, true>(p, stream) : launch_fv<U, R, true, false>(p, stream);
        return guard ? launch_fv<U, R, false, true>(p, stream) : launch_fv<U, R, false, false>(p, stream);
    }
    if (iso) return guard ? launch_f<U, R, true, true>(p, stream) : launch_f<U, R, true, false>(p, stream);
    return guard ? launch_f<U, R, false, true>(p, stream) : launch_f<U, R, false, false>(p, stream);
}

cudaError_t launch_gather_fast(const GatherParams &p, bool uniform, bool rr, bool vis, bool iso, bool guard, cudaStream_t stream) {
    int key = (uniform ? 4 : 0) | (rr ? 2 : 0) | (vis ? 1 : 0);
    switch (key) {
        case 0: return launch_f2<false, false, false>(p, iso, guard, stream);
        case 1: return launch_f2<false, false, true>(p, iso, guard, stream);
        case 2: return launch_f2<false, true, false>(p, iso, guard, stream);
        case 3: return launch_f2<false, true, true>(p, iso, guard, stream);
        case 4: return launch_f2<true, false, false>(p, iso, guard, stream);
        case 5: return launch_f2<true, false, true>(p, iso, guard, stream);
        case 6: return launch_f2<true, true, false>(p, iso, guard, stream);
        default: return launch_f2<true, true, true>(p, iso, guard, stream);
    }
}

}  // namespace vrlb
