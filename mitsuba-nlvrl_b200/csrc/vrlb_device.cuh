// vrlb_device.cuh -- device-side building blocks shared by the kernels of libvrl_b200.so.
//
// Arithmetic contract ("exact" mode): this translation unit is compiled with
//   -fmad=false -prec-div=true -prec-sqrt=true -ftz=false
// so every + - * / sqrtf below is one IEEE binary32 operation in the written order, the same
// arithmetic the reference's scalar_rgb variant performs statement by statement
// (src/integrators/vrl/vrl_struct.h).  Transcendentals (logf, sinhf, atanf, tanf, expf) are the CUDA
// libm versions (<= 4 ulp); the reference reaches the host libm through Enoki's scalar fallbacks.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace vrlb {

// ------------------------------------------------------------------------------------------------
struct V3 {
    float x, y, z;
    __host__ __device__ V3() {}
    __host__ __device__ V3(float a) : x(a), y(a), z(a) {}
    __host__ __device__ V3(float a, float b, float c) : x(a), y(b), z(c) {}
};
__host__ __device__ inline V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
__host__ __device__ inline V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
__host__ __device__ inline V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
__host__ __device__ inline V3 operator*(V3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
__host__ __device__ inline V3 operator*(V3 a, V3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
__host__ __device__ inline V3 operator/(V3 a, float s) { return V3(a.x / s, a.y / s, a.z / s); }
__host__ __device__ inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__host__ __device__ inline V3 cross(V3 a, V3 b) {
    return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__host__ __device__ inline float squared_norm(V3 a) { return dot(a, a); }
__device__ inline float norm(V3 a) { return sqrtf(dot(a, a)); }
__device__ inline V3 normalize(V3 a) { return a / norm(a); }
__device__ inline float safe_sqrt(float x) { return sqrtf(fmaxf(x, 0.0f)); }
// enoki::lerp = fmadd(b, t, fnmadd(a, t, a)), ext/enoki/include/enoki/array_math.h:1352 (uncontracted)
__device__ inline float lerpf(float a, float b, float t) { return b * t + (a - a * t); }
__device__ inline float hmax_abs(V3 p) { return fmaxf(fabsf(p.x), fmaxf(fabsf(p.y), fabsf(p.z))); }

constexpr float kEpsilon    = 5.9604644775390625e-08f;  // math::Epsilon<float>, include/mitsuba/core/math.h:36
constexpr float kRayEpsilon = kEpsilon * 1500.0f;       // math.h:37
constexpr float kInvFourPi  = 0.07957747154594766788f;
#define VRLB_INF __int_as_float(0x7f800000)

// ------------------------------------------------------------------------------------------------
// PCG32, ext/enoki/include/enoki/random.h:62-119, 265-286.  All pixel streams share the default
// stream constant (sampler.cpp:93), so `inc` is a compile-time constant.
constexpr uint64_t PCG32_MULT = 0x5851f42d4c957f2dULL;
constexpr uint64_t PCG32_INC  = (0xda3e39cb94b95bdbULL << 1) | 1ULL;

struct Pcg32 {
    uint64_t state;
    __host__ __device__ inline uint32_t next_u32() {
        uint64_t old = state;
        state        = old * PCG32_MULT + PCG32_INC;
        uint32_t xs  = (uint32_t) (((old >> 18) ^ old) >> 27);
        uint32_t rot = (uint32_t) (old >> 59);
        return (xs >> rot) | (xs << ((32 - rot) & 31));
    }
    __device__ inline float next_f32() { return __uint_as_float((next_u32() >> 9) | 0x3f800000u) - 1.0f; }
    __host__ __device__ inline void seed(uint64_t initstate) {   // seed(initstate, PCG32_DEFAULT_STREAM)
        state = 0; next_u32(); state += initstate; next_u32();
    }
    __host__ __device__ inline void advance(uint64_t delta) {
        uint64_t cur_mult = PCG32_MULT, cur_plus = PCG32_INC, acc_mult = 1, acc_plus = 0;
        while (delta != 0) {
            if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
            cur_plus = (cur_mult + 1) * cur_plus;
            cur_mult *= cur_mult;
            delta >>= 1;
        }
        state = acc_mult * state + acc_plus;
    }
};

// ------------------------------------------------------------------------------------------------
// Device-resident scene records
struct DTri {            // 48 B, BVH order.  e1 = p1 - p0, e2 = p2 - p0 rounded once on the host (same fp32 op)
    float p0[3]; uint32_t prim;     // original triangle index (tie-break, flags lookup)
    float e1[3]; uint32_t flags;
    float e2[3]; uint32_t pad;
};
struct DNode {           // 32 B.  inner: a = left child, b = right child; leaf: a = first tri, b = -(count)
    float lo[3]; int32_t a;
    float hi[3]; int32_t b;
};
struct DMedium {         // what the gather needs of a medium (homogeneous / nonlinear coefficients)
    float max_density;   // scale * mean(sigma_t): grey extinction (homogeneous.cpp:26, constant3d.cpp:37)
    float sigma_s[3];    // scale * sigma_t * albedo (homogeneous.cpp:42-50)
    float g;
    int32_t phase;
    int32_t kind;
    int32_t pad;
};
struct DVrl {            // 48 B record: three 16-byte rows, staged tile-wise into shared memory by TMA
    float ox, oy, oz, length;
    float dx, dy, dz; uint32_t medium;
    float fr, fg, fb, pad;
};

struct DScene {
    const DTri  *tris;
    const float *verts;      // n_tris * 9, ORIGINAL order and values (si.p re-interpolation, mesh.cpp:484)
    const DNode *nodes;
    int32_t      n_tris;
    int32_t      n_nodes;
    const DMedium *media;
    int32_t      n_media;
    const float *emitters;   // n_emitters * 4
    int32_t      n_emitters;
};

// Mesh::ray_intersect_triangle, include/mitsuba/render/mesh.h:195-225
__device__ inline bool intersect_tri(const DTri &tr, V3 o, V3 d, float mint, float maxt, float &t, float &u, float &v) {
    V3 p0(tr.p0[0], tr.p0[1], tr.p0[2]), e1(tr.e1[0], tr.e1[1], tr.e1[2]), e2(tr.e2[0], tr.e2[1], tr.e2[2]);
    V3 pvec       = cross(d, e2);
    float inv_det = 1.0f / dot(e1, pvec);
    V3 tvec       = o - p0;
    u             = dot(tvec, pvec) * inv_det;
    if (!(u >= 0.f && u <= 1.f)) return false;
    V3 qvec = cross(tvec, e1);
    v       = dot(d, qvec) * inv_det;
    if (!(v >= 0.f && u + v <= 1.f)) return false;
    t = dot(e2, qvec) * inv_det;
    return (t >= mint && t <= maxt);
}

// conservative slab test (boxes are padded on the host); NaNs from 0*inf are dropped by fminf/fmaxf
__device__ inline bool slab(const float4 lo, const float4 hi, V3 o, V3 inv, float mint, float maxt) {
    float ax = (lo.x - o.x) * inv.x, bx = (hi.x - o.x) * inv.x;
    float ay = (lo.y - o.y) * inv.y, by = (hi.y - o.y) * inv.y;
    float az = (lo.z - o.z) * inv.z, bz = (hi.z - o.z) * inv.z;
    float t0 = fmaxf(fmaxf(mint, fminf(ax, bx)), fmaxf(fminf(ay, by), fminf(az, bz)));
    float t1 = fminf(fminf(maxt, fmaxf(ax, bx)), fminf(fmaxf(ay, by), fmaxf(az, bz)));
    return t0 <= t1 * 1.00001f + 1e-6f;
}

struct Hit { float t, u, v; int tri; };   // tri = index into DScene::tris (BVH order), -1 = miss

// Scene::ray_intersect (src/librender/scene.cpp:114-131): closest hit, ties -> lowest original index
__device__ inline Hit closest_hit(const DScene &s, V3 o, V3 d, float mint, float maxt) {
    Hit best; best.t = VRLB_INF; best.u = best.v = 0.f; best.tri = -1;
    if (s.n_nodes == 0) return best;
    V3 inv(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    int stack[48]; int sp = 0; stack[sp++] = 0;
    uint32_t best_prim = 0xffffffffu;
    while (sp) {
        int ni = stack[--sp];
        const float4 *np = reinterpret_cast<const float4 *>(s.nodes + ni);
        float4 lo = __ldg(np), hi = __ldg(np + 1);
        float far = best.tri >= 0 ? best.t : maxt;
        if (!slab(lo, hi, o, inv, mint, far)) continue;
        int a = __float_as_int(lo.w), b = __float_as_int(hi.w);
        if (b < 0) {
            for (int i = a; i < a - b; ++i) {
                const DTri &tr = s.tris[i];
                float t, u, v;
                if (intersect_tri(tr, o, d, mint, maxt, t, u, v)) {
                    if (t < best.t || (t == best.t && tr.prim < best_prim)) { best.t = t; best.u = u; best.v = v; best.tri = i; best_prim = tr.prim; }
                }
            }
        } else {
            stack[sp++] = a; stack[sp++] = b;
        }
    }
    return best;
}

// Shadow query of VRL::evalTransmittance (vrl_struct.h:509-572).  Null::eval() returns 0
// (src/bsdfs/null.cpp:59-62), so crossing ANY surface -- occluder or index-matched boundary -- zeroes
// the transmittance: the reference's closest-hit loop reduces to "is there any hit in [0, maxt]".
__device__ inline bool any_hit(const DScene &s, V3 o, V3 d, float mint, float maxt) {
    if (s.n_nodes == 0) return false;
    V3 inv(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    int stack[48]; int sp = 0; stack[sp++] = 0;
    while (sp) {
        int ni = stack[--sp];
        const float4 *np = reinterpret_cast<const float4 *>(s.nodes + ni);
        float4 lo = __ldg(np), hi = __ldg(np + 1);
        if (!slab(lo, hi, o, inv, mint, maxt)) continue;
        int a = __float_as_int(lo.w), b = __float_as_int(hi.w);
        if (b < 0) {
            for (int i = a; i < a - b; ++i) {
                float t, u, v;
                if (intersect_tri(s.tris[i], o, d, mint, maxt, t, u, v)) return true;
            }
        } else {
            stack[sp++] = a; stack[sp++] = b;
        }
    }
    return false;
}

// Medium::evalTransmittance homogeneous branch (include/mitsuba/render/medium.h:112-117)
__device__ inline float homo_transmittance(float max_density, float maxt) {
    float val = expf(max_density * (0.f - maxt));
    return max_density != 0.f ? val : 1.0f;
}

// IsotropicPhaseFunction::eval (src/phase/isotropic.cpp:40-44) / HGPhaseFunction::eval (src/phase/hg.cpp:51-54,77-81)
__device__ inline float phase_eval(int phase, float g, V3 wi, V3 wo) {
    if (phase == 0) return kInvFourPi;
    float cos_theta = dot(wo, wi);
    float temp      = 1.0f + g * g + 2.0f * g * cos_theta;
    return kInvFourPi * (1 - g * g) / (temp * sqrtf(temp));
}

}  // namespace vrlb
