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
#include "gather_exact.cuh"

namespace vrlb {

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

template <bool UNIFORM, bool RR, bool VIS, bool GRID>
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
    // use_nl_atomic_query (vrl.cpp:613-617): this part's place in its NLRay and the first part (distance roulette)
    float nla_off = 0.f, nla_total = -1.f; SegRegs ray0 = ray;
    if (UNIFORM && p.nla && live) {
        const vrlb_segment &me = p.segs[seg_i];
        const uint32_t n_all = me.sample >> 16, part = me.part;
        if (n_all > 0) {
            const vrlb_segment *g0 = p.segs + (seg_i - part);
            float tot = 0.f;
            for (uint32_t k = 0; k < n_all; ++k) { if (k == part) nla_off = tot; tot = tot + __ldg(&g0[k].maxt); }   // NLRay::push_back order
            nla_total = tot;
            ray0.o = V3(g0->o[0], g0->o[1], g0->o[2]); ray0.d = V3(g0->d[0], g0->d[1], g0->d[2]); ray0.maxt = g0->maxt;
        }
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
                const uint64_t eval_state = rng.state;      // keys the grid-medium sub-stream
                if (RR) {
                    // EDistanceRoulette, vrl_map.h:171-190
                    float cc, cv;
                    find_closest_point(vo, vd, vlen, ray0.o, ray0.d, ray0.maxt, cc, cv);
                    float d2 = squared_norm((ray0.o + ray0.d * cc) - (vo + vd * cv));
                    float min_distance = sqrtf(d2);
                    rrWeight = fminf(1 / (min_distance * p.scale_rr), 1.0f);
                    survive  = rrWeight > rng.next_f32();
                }
                float u1 = rng.next_f32(), u2 = rng.next_f32();
                if (!survive) continue;
                float st2 = 0.f;
                float f = eval_contrib<UNIFORM, VIS, GRID, true>(p.scene, ray, vo, vd, vlen, m.max_density, m.phase, m.g, u1, u2,
                                                           &m, eval_state, &st2, nullptr, nla_off, nla_total);
                if (RR) f = f / rrWeight;
                // flux * f * sigma_s(cam) * sigma_s(vrl) * m_scale (vrl_struct.h:652-659, vrl_map.h:160)
                float cr, cg, cb;
                if (GRID && m.kind == VRLB_MEDIUM_GRID) {
                    cr = c.x * f * (m.albedo[0] * m.albedo[0]) * st2 * p.scale;
                    cg = c.y * f * (m.albedo[1] * m.albedo[1]) * st2 * p.scale;
                    cb = c.z * f * (m.albedo[2] * m.albedo[2]) * st2 * p.scale;
                } else {
                    cr = c.x * f * m.sigma_s[0] * m.sigma_s[0] * p.scale;
                    cg = c.y * f * m.sigma_s[1] * m.sigma_s[1] * p.scale;
                    cb = c.z * f * m.sigma_s[2] * m.sigma_s[2] * p.scale;
                }
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

template <bool U, bool R, bool V, bool G>
static cudaError_t launch_t(const GatherParams &p, cudaStream_t stream) {
    uint32_t grid = (uint32_t) ((p.n_segs + GATHER_THREADS - 1) / GATHER_THREADS);
    if (grid == 0 || p.n_vrls == 0) return cudaSuccess;
    gather_kernel<U, R, V, G><<<grid, GATHER_THREADS, 0, stream>>>(p);
    return cudaGetLastError();
}
template <bool G>
static cudaError_t launch_g(const GatherParams &p, bool uniform, bool rr, bool vis, cudaStream_t stream) {
    int key = (uniform ? 4 : 0) | (rr ? 2 : 0) | (vis ? 1 : 0);
    switch (key) {
        case 0: return launch_t<false, false, false, G>(p, stream);
        case 1: return launch_t<false, false, true, G>(p, stream);
        case 2: return launch_t<false, true, false, G>(p, stream);
        case 3: return launch_t<false, true, true, G>(p, stream);
        case 4: return launch_t<true, false, false, G>(p, stream);
        case 5: return launch_t<true, false, true, G>(p, stream);
        case 6: return launch_t<true, true, false, G>(p, stream);
        default: return launch_t<true, true, true, G>(p, stream);
    }
}

cudaError_t launch_gather(const GatherParams &p, bool uniform, bool rr, bool vis, bool grid, cudaStream_t stream) {
    return grid ? launch_g<true>(p, uniform, rr, vis, stream) : launch_g<false>(p, uniform, rr, vis, stream);
}

}  // namespace vrlb
