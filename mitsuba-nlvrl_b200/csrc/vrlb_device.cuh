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

#include "../../include/vrl_b200.h"

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
constexpr uint64_t PCG32_DEFAULT_STREAM = 0xda3e39cb94b95bdbULL;
constexpr uint64_t PCG32_INC  = (PCG32_DEFAULT_STREAM << 1) | 1ULL;

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

// PCG32 on an arbitrary stream (the grid-medium transmittance sub-stream, SURVEY 8d "addressed" mode)
struct Pcg32s {
    uint64_t state, inc;
    __device__ inline uint32_t next_u32() {
        uint64_t old = state;
        state        = old * PCG32_MULT + inc;
        uint32_t xs  = (uint32_t) (((old >> 18) ^ old) >> 27);
        uint32_t rot = (uint32_t) (old >> 59);
        return (xs >> rot) | (xs << ((32 - rot) & 31));
    }
    __device__ inline float next_f32() { return __uint_as_float((next_u32() >> 9) | 0x3f800000u) - 1.0f; }
    __device__ inline void seed(uint64_t initstate, uint64_t initseq) {      // enoki/random.h:79-86
        state = 0; inc = (initseq << 1u) | 1u; next_u32(); state += initstate; next_u32();
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
struct DWNode {          // 64 B: the boxes of BOTH children of an inner node, for the shadow (any-hit) traversal.
    float lo0[3], hi0[3], lo1[3], hi1[3];
    uint32_t c0, c1;     // child reference: inner node index, or leaf = 0x80000000 | first_tri << 12 | count << 7 (low 7 bits free)
    uint32_t pad[2];
};
#define VRLB_MAX_SHELLS 8
struct DShell {          // a component lying in the faces of its own AABB (fog container, room walls)
    float lo[3], hi[3];  // AABB shrunk by the safety margin
    uint32_t root;       // child reference as in DWNode
    uint32_t pad;
};
// A connected component with its own BVH and conservative proxies (occluder_proxies.cpp): outer = bounding sphere +
// oriented box (a segment missing one misses the mesh); inner = inscribed sphere + oriented box of a CLOSED component
// (a segment with an end point outside the AABB that meets one crosses the surface).
#define VRLB_MAX_SOLIDS 6
#define VRLB_SOLID_OUTER_OBB    1u
#define VRLB_SOLID_INNER_SPHERE 2u
#define VRLB_SOLID_INNER_OBB    4u
struct DSolid {
    float lo[3], hi[3];      // AABB grown by the safety margin
    uint32_t root;           // child reference as in DWNode
    uint32_t flags;
    float c[3], r_out2;      // bounding sphere: centre, (radius + margin)^2
    float ci[3], r_in2;      // inscribed sphere: centre, (radius - margin)^2
    float ax[9];             // oriented-box frame, rows = unit axes
    float oc[3];             // oriented-box centre
    float h_out[3];          // outer half extents (+ margin)
    float h_in[3];           // inscribed half extents (- margin)
};
struct DMedium {         // what the gather needs of a medium (homogeneous / nonlinear coefficients)
    float max_density;   // scale * mean(sigma_t): grey extinction (homogeneous.cpp:26, constant3d.cpp:37)
    float sigma_s[3];    // scale * sigma_t * albedo (homogeneous.cpp:42-50)
    float g;
    int32_t phase;
    int32_t kind;
    int32_t from_bottom; // nonlinear.cpp:149-150: 1 = height measured from the box floor; 0 = n(y) = lerp(bottom, top, max(0, y / (height / 2))), world y
    // graded refractive index (VRLB_MEDIUM_NONLINEAR, nonlinear.cpp:129-153 method 0): n(y) = lerp(bottom, top, (y - y0) / height)
    float nl_lo[3], nl_hi[3];
    float ior_bottom, ior_top;
    float nl_step;       // eikonal step length = height / res_y (one IOR layer of the reference's cell grid)
    float albedo[3];
    float scale;         // medium `scale`
    // grid medium (VRLB_MEDIUM_GRID, heterogeneous.cpp + grid3d.cpp): scalar sigma_t grid, x fastest; max_density = scale * grid max
    const float *grid;   // device copy, one allocation per scene
    int32_t g_res[3];
    float   g_lo[3], g_hi[3];
    float   pad2[2];
};
struct DVrl {            // 48 B record: three 16-byte rows, staged tile-wise into shared memory by TMA
    float ox, oy, oz, length;
    float dx, dy, dz; uint32_t medium;
    float fr, fg, fb, pad;
};
struct DVrlV {           // 64 B record of the visibility kernel: DVrl rows (flux pre-multiplied, extinction) + the parameter
    float ox, oy, oz, length;    // interval (v_lo, v_hi) of the VRL's points strictly inside the shells' safe box
    float dx, dy, dz; uint32_t medium;
    float fr, fg, fb, ext;
    float v_lo, v_hi; uint32_t pad0, pad1;
};

struct DScene {
    const DTri  *tris;
    const float *verts;      // n_tris * 9, ORIGINAL order and values (si.p re-interpolation, mesh.cpp:484)
    const DNode *nodes;
    // shadow structure (bvh_build.cpp::build_shadow_accel)
    const DTri   *stris;
    const DWNode *wnodes;
    uint32_t     main_root;  // child reference; 0x80000000 = no main geometry
    float        main_lo[3], main_hi[3];   // root box of the main BVH
    int32_t      n_shells;
    DShell       shells[VRLB_MAX_SHELLS];
    int32_t      n_solids;
    const DSolid *solids;    // device array (dynamic indexing of a by-value parameter array would go through local memory)
    int32_t      n_tris;
    int32_t      n_nodes;
    const DMedium *media;
    int32_t      n_media;
    const float *emitters;   // n_emitters * 4
    int32_t      n_emitters;
};

// IOR profile of the `nonlinear` medium (nonlinear.cpp:129-153, ior_method 0): the index n at p and its vertical gradient
// (0 outside the medium's box).  from_bottom = false (mirage.xml): the lower half of the world (y <= 0) has the bottom index.
__device__ inline void nl_profile(const DMedium &m, V3 p, float &n, float &g) {
    const float h = m.nl_hi[1] - m.nl_lo[1];
    const bool in = p.x >= m.nl_lo[0] && p.x <= m.nl_hi[0] && p.y >= m.nl_lo[1] && p.y <= m.nl_hi[1] && p.z >= m.nl_lo[2] && p.z <= m.nl_hi[2];
    if (m.from_bottom) {
        const float t = fminf(fmaxf((p.y - m.nl_lo[1]) / h, 0.f), 1.f);
        n = m.ior_bottom + (m.ior_top - m.ior_bottom) * t;
        g = in ? (m.ior_top - m.ior_bottom) / h : 0.f;
    } else {
        const float hh = h * 0.5f;
        const float t = fmaxf(0.f, p.y / hh);
        n = m.ior_bottom + (m.ior_top - m.ior_bottom) * t;
        g = (in && p.y > 0.f) ? (m.ior_top - m.ior_bottom) / hh : 0.f;
    }
}
// one eikonal step of length ds: v = n dir, v.y += ds * dn/dy, dir = v / |v| (bends `dir` towards increasing n)
__device__ inline V3 nl_bend(const DMedium &m, V3 p, V3 dir, float ds) {
    float n, g;
    nl_profile(m, p, n, g);
    V3 v = dir * n;
    v.y = v.y + ds * g;
    return normalize(v);
}

// ---- delta BSDFs of the specular chains (vrl.cpp:318-358 light side, 682-716 camera side) ----
// The translation units are built with -fmad=false -prec-div=true -prec-sqrt=true, so the plain expressions below are
// one IEEE operation per written operation, like the oracle's.
__device__ inline void coordinate_system(V3 n, V3 &s, V3 &t) {            // core/vector.h (Duff et al.)
    float sign = copysignf(1.f, n.z), a = -1.f / (sign + n.z), b = n.x * n.y * a;
    s = V3(1.f + sign * n.x * n.x * a, sign * b, -sign * n.x);
    t = V3(b, sign + n.y * n.y * a, -n.y);
}
struct DeltaSample { V3 wo; V3 weight; bool valid, transmitted; };
// bsdf->sample(ctx, si, sample1, sample2) of SmoothConductor / SmoothDielectric with the default BSDFContext
// (TransportMode::Radiance on both the camera and the light side, as vrl.cpp constructs it).  `rd` = ray direction,
// `n` = geometric normal = shading frame, prm = the triangle's three tri_reflectance floats.
__device__ inline DeltaSample sample_delta(uint32_t kind, float p0, float p1, float p2, bool twosided, V3 n, V3 rd, float sample1) {
    DeltaSample r; r.valid = false; r.transmitted = false; r.weight = V3(0.f); r.wo = rd;
    V3 s, t; coordinate_system(n, s, t);
    const V3 w = -rd;
    const V3 wi(dot(w, s), dot(w, t), dot(w, n));                          // si.wi = to_local(-ray.d)
    V3 wo;
    if (kind == VRLB_SURF_BSDF_MIRROR) {                                   // conductor.cpp: active &= cos_theta_i > 0; wo = reflect(wi)
        if (!(wi.z > 0.f) && !(twosided && wi.z < 0.f)) return r;
        wo = V3(-wi.x, -wi.y, wi.z);
        r.weight = V3(p0, p1, p2);
    } else {                                                               // dielectric.cpp + fresnel(), core/fresnel.h
        const float eta = p0, cos_i = wi.z;
        const bool outside = cos_i >= 0.f;
        const float rcp_eta = 1.f / eta, eta_it = outside ? eta : rcp_eta, eta_ti = outside ? rcp_eta : eta;
        const float cos_t_sqr = 1.f - (1.f - cos_i * cos_i) * (eta_ti * eta_ti);
        const float cos_i_abs = fabsf(cos_i), cos_t_abs = sqrtf(fmaxf(cos_t_sqr, 0.f));
        const bool special = eta == 1.f || cos_i_abs == 0.f;
        const float a_s = (cos_i_abs - eta_it * cos_t_abs) / (cos_i_abs + eta_it * cos_t_abs);
        const float a_p = (cos_t_abs - eta_it * cos_i_abs) / (cos_t_abs + eta_it * cos_i_abs);
        float r_i = 0.5f * (a_s * a_s + a_p * a_p);
        if (special) r_i = eta == 1.f ? 0.f : 1.f;
        const float cos_t = cos_i >= 0.f ? -cos_t_abs : cos_t_abs;         // mulsign_neg(cos_theta_t_abs, cos_theta_i)
        if (sample1 <= r_i) { wo = V3(-wi.x, -wi.y, wi.z); r.weight = V3(p1); }
        else {
            wo = V3(-eta_ti * wi.x, -eta_ti * wi.y, cos_t);                // refract(wi, cos_theta_t, eta_ti)
            r.weight = V3(p2 * (eta_ti * eta_ti));                         // radiance scaling, dielectric.cpp (mode == Radiance)
            r.transmitted = true;
        }
    }
    r.wo = s * wo.x + t * wo.y + n * wo.z;                                 // si.to_world(bs.wo)
    r.valid = true;
    return r;
}

// Exactly-rounded fp32 building blocks: one IEEE operation each, immune to FMA contraction whatever the
// translation unit's -fmad setting.  Hit / no-hit decisions (and the exact-mode gather) are built from these so
// they agree with the oracle's -ffp-contract=off arithmetic.
__device__ __forceinline__ float xmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float xadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float xsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float xdot(V3 a, V3 b) { return xadd(xadd(xmul(a.x, b.x), xmul(a.y, b.y)), xmul(a.z, b.z)); }
__device__ __forceinline__ V3 xcross(V3 a, V3 b) {
    return V3(xsub(xmul(a.y, b.z), xmul(a.z, b.y)), xsub(xmul(a.z, b.x), xmul(a.x, b.z)), xsub(xmul(a.x, b.y), xmul(a.y, b.x)));
}
__device__ __forceinline__ V3 xsub3(V3 a, V3 b) { return V3(xsub(a.x, b.x), xsub(a.y, b.y), xsub(a.z, b.z)); }

// Mesh::ray_intersect_triangle, include/mitsuba/render/mesh.h:195-225
__device__ inline bool intersect_tri(const DTri &tr, V3 o, V3 d, float mint, float maxt, float &t, float &u, float &v) {
    V3 p0(tr.p0[0], tr.p0[1], tr.p0[2]), e1(tr.e1[0], tr.e1[1], tr.e1[2]), e2(tr.e2[0], tr.e2[1], tr.e2[2]);
    V3 pvec       = xcross(d, e2);
    float inv_det = __fdiv_rn(1.0f, xdot(e1, pvec));
    V3 tvec       = xsub3(o, p0);
    u             = xmul(xdot(tvec, pvec), inv_det);
    if (!(u >= 0.f && u <= 1.f)) return false;
    V3 qvec = xcross(tvec, e1);
    v       = xmul(xdot(d, qvec), inv_det);
    if (!(v >= 0.f && xadd(u, v) <= 1.f)) return false;
    t = xmul(xdot(e2, qvec), inv_det);
    return (t >= mint && t <= maxt);
}

// Shadow-test variant for rays whose end points carry fast-arithmetic error (~1e-5 absolute): 0 = certain miss,
// 1 = certain hit, 2 = within the margins of an edge / end point: the caller re-evaluates that eval with the
// exactly-rounded path, whose ray matches the oracle's.
__device__ inline int intersect_tri_graze(const DTri &tr, V3 o, V3 d, float mint, float maxt) {
    const float eu = 1e-4f, et = 2e-4f + 1e-5f * maxt;
    V3 p0(tr.p0[0], tr.p0[1], tr.p0[2]), e1(tr.e1[0], tr.e1[1], tr.e1[2]), e2(tr.e2[0], tr.e2[1], tr.e2[2]);
    V3 pvec       = xcross(d, e2);
    float inv_det = __fdiv_rn(1.0f, xdot(e1, pvec));
    V3 tvec       = xsub3(o, p0);
    float u       = xmul(xdot(tvec, pvec), inv_det);
    if (!(u >= -eu && u <= 1.f + eu)) return 0;
    V3 qvec = xcross(tvec, e1);
    float v = xmul(xdot(d, qvec), inv_det);
    if (!(v >= -eu && xadd(u, v) <= 1.f + eu)) return 0;
    float t = xmul(xdot(e2, qvec), inv_det);
    if (!(t >= mint - et && t <= maxt + et)) return 0;
    bool sure = u >= eu && v >= eu && xadd(u, v) <= 1.f - eu && t >= mint + et && t <= maxt - et;
    return sure ? 1 : 2;
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
// Traversal over DWNode: one 64 B fetch gives both children's boxes; the slab tests are conservative (padded
// boxes, FMA form), the triangle test is the exactly-rounded one, so hit / no-hit equals the oracle's.
__device__ __forceinline__ bool slab2(float lox, float loy, float loz, float hix, float hiy, float hiz,
                                      V3 inv, V3 o, float mint, float maxt) {
    // (plane - o) * inv, not fma(plane, inv, -o * inv): the latter loses |o * inv| * eps absolutely when a
    // direction component is tiny, enough to skip a box that is hit
    float ax = (lox - o.x) * inv.x, bx = (hix - o.x) * inv.x;
    float ay = (loy - o.y) * inv.y, by = (hiy - o.y) * inv.y;
    float az = (loz - o.z) * inv.z, bz = (hiz - o.z) * inv.z;
    float t0 = fmaxf(fmaxf(mint, fminf(ax, bx)), fmaxf(fminf(ay, by), fminf(az, bz)));
    float t1 = fminf(fminf(maxt, fmaxf(ax, bx)), fminf(fmaxf(ay, by), fmaxf(az, bz)));
    return t0 <= fmaf(t1, 1.00001f, 1e-5f);
}
__device__ __forceinline__ bool leaf_any_hit(const DScene &s, uint32_t ref, V3 o, V3 d, float mint, float maxt) {
    uint32_t first = (ref & 0x7fffffffu) >> 12, cnt = (ref >> 7) & 31u;
    for (uint32_t i = first; i < first + cnt; ++i) {
        float t, u, v;
        if (intersect_tri(s.stris[i], o, d, mint, maxt, t, u, v)) return true;
    }
    return false;
}
__device__ inline bool traverse_any_hit(const DScene &s, uint32_t root, V3 o, V3 d, V3 inv, float mint, float maxt) {
    if (root & 0x80000000u) return leaf_any_hit(s, root, o, d, mint, maxt);
    // 0 * inf = NaN when a direction component is 0 and the origin lies on a slab plane: fminf/fmaxf drop NaNs,
    // which only widens the interval (conservative)
    uint32_t stack[32]; int sp = 0;
    uint32_t cur = root;
    while (true) {
        const float4 *np = reinterpret_cast<const float4 *>(s.wnodes + cur);
        float4 a = __ldg(np), b = __ldg(np + 1), c = __ldg(np + 2), e = __ldg(np + 3);
        bool h0 = slab2(a.x, a.y, a.z, a.w, b.x, b.y, inv, o, mint, maxt);
        bool h1 = slab2(b.z, b.w, c.x, c.y, c.z, c.w, inv, o, mint, maxt);
        uint32_t c0 = __float_as_uint(e.x), c1 = __float_as_uint(e.y);
        if (h0 && (c0 & 0x80000000u)) { if (leaf_any_hit(s, c0, o, d, mint, maxt)) return true; h0 = false; }
        if (h1 && (c1 & 0x80000000u)) { if (leaf_any_hit(s, c1, o, d, mint, maxt)) return true; h1 = false; }
        if (h0) { if (h1) stack[sp++] = c1; cur = c0; }
        else if (h1) cur = c1;
        else { if (sp == 0) return false; cur = stack[--sp]; }
    }
}
__device__ inline bool any_hit(const DScene &s, V3 o, V3 d, float mint, float maxt) {
    if (s.n_tris == 0) return false;
    V3 inv(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    if (s.main_root != 0x80000000u && traverse_any_hit(s, s.main_root, o, d, inv, mint, maxt)) return true;
    V3 e = o + d * maxt;
    for (int i = 0; i < s.n_shells; ++i) {
        const DShell &sh = s.shells[i];
        bool inside = o.x > sh.lo[0] && o.x < sh.hi[0] && o.y > sh.lo[1] && o.y < sh.hi[1] && o.z > sh.lo[2] && o.z < sh.hi[2] &&
                      e.x > sh.lo[0] && e.x < sh.hi[0] && e.y > sh.lo[1] && e.y < sh.hi[1] && e.z > sh.lo[2] && e.z < sh.hi[2];
        if (!inside && traverse_any_hit(s, sh.root, o, d, inv, mint, maxt)) return true;
    }
    // solids: exact traversal only (the proxies are the fast path's business; this is the parity anchor)
    for (int i = 0; i < s.n_solids; ++i) {
        const DSolid &so = s.solids[i];
        if (slab(make_float4(so.lo[0], so.lo[1], so.lo[2], 0.f), make_float4(so.hi[0], so.hi[1], so.hi[2], 0.f), o, inv, mint, maxt) && traverse_any_hit(s, so.root, o, d, inv, mint, maxt)) return true;
    }
    return false;
}

// any_hit with the solids' conservative proxies in front of their BVHs (occluder_proxies.cpp): a segment that misses a
// solid's bounding sphere or outer oriented box cannot hit it; a segment that enters from outside and meets the inscribed
// sphere or box of a CLOSED solid is occluded.  The proxies carry the build's safety margin, so the answer is any_hit's;
// only segments between the inner and the outer proxies walk the triangles.  For callers whose lanes trace unrelated
// segments (the light cut's per-query refinement), where a BVH walk runs one or two lanes wide.
__device__ inline bool any_hit_proxied(const DScene &s, V3 o, V3 d, float mint, float maxt) {
    if (s.n_tris == 0) return false;
    V3 inv(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    if (s.main_root != 0x80000000u && traverse_any_hit(s, s.main_root, o, d, inv, mint, maxt)) return true;
    const V3 a = o + d * mint, e = o + d * maxt;
    for (int i = 0; i < s.n_shells; ++i) {
        const DShell &sh = s.shells[i];
        bool inside = o.x > sh.lo[0] && o.x < sh.hi[0] && o.y > sh.lo[1] && o.y < sh.hi[1] && o.z > sh.lo[2] && o.z < sh.hi[2] &&
                      e.x > sh.lo[0] && e.x < sh.hi[0] && e.y > sh.lo[1] && e.y < sh.hi[1] && e.z > sh.lo[2] && e.z < sh.hi[2];
        if (!inside && traverse_any_hit(s, sh.root, o, d, inv, mint, maxt)) return true;
    }
    const V3 dir = e - a;                                      // the segment a + t dir, t in [0, 1]
    const float inv_d2 = 1.0f / dot(dir, dir);
    for (int i = 0; i < s.n_solids; ++i) {
        const DSolid &so = s.solids[i];
        if (!slab(make_float4(so.lo[0], so.lo[1], so.lo[2], 0.f), make_float4(so.hi[0], so.hi[1], so.hi[2], 0.f), o, inv, mint, maxt)) continue;
        const V3 oc(so.c[0] - a.x, so.c[1] - a.y, so.c[2] - a.z);
        {
            const float t = fminf(fmaxf(dot(oc, dir) * inv_d2, 0.f), 1.f);
            const V3 q = oc - dir * t;
            if (dot(q, q) > so.r_out2) continue;                                    // misses the bounding sphere
        }
        const V3 ec(so.c[0] - e.x, so.c[1] - e.y, so.c[2] - e.z);
        bool enters = dot(oc, oc) > so.r_out2 || dot(ec, ec) > so.r_out2;          // an end point outside the solid
        if ((so.flags & VRLB_SOLID_INNER_SPHERE) && enters) {
            const V3 oi(so.ci[0] - a.x, so.ci[1] - a.y, so.ci[2] - a.z);
            const float t = fminf(fmaxf(dot(oi, dir) * inv_d2, 0.f), 1.f);
            const V3 q = oi - dir * t;
            if (dot(q, q) <= so.r_in2) return true;
        }
        if (so.flags & (VRLB_SOLID_OUTER_OBB | VRLB_SOLID_INNER_OBB)) {
            const V3 r(a.x - so.oc[0], a.y - so.oc[1], a.z - so.oc[2]);
            const V3 ol(dot(V3(so.ax[0], so.ax[1], so.ax[2]), r), dot(V3(so.ax[3], so.ax[4], so.ax[5]), r), dot(V3(so.ax[6], so.ax[7], so.ax[8]), r));
            const V3 dl(dot(V3(so.ax[0], so.ax[1], so.ax[2]), dir), dot(V3(so.ax[3], so.ax[4], so.ax[5]), dir), dot(V3(so.ax[6], so.ax[7], so.ax[8]), dir));
            const V3 il(1.0f / dl.x, 1.0f / dl.y, 1.0f / dl.z);
            if (so.flags & VRLB_SOLID_OUTER_OBB) {
                if (!slab2(-so.h_out[0], -so.h_out[1], -so.h_out[2], so.h_out[0], so.h_out[1], so.h_out[2], il, ol, 0.f, 1.f)) continue;
                const V3 el = ol + dl;
                enters = enters || fabsf(ol.x) > so.h_out[0] || fabsf(ol.y) > so.h_out[1] || fabsf(ol.z) > so.h_out[2] ||
                         fabsf(el.x) > so.h_out[0] || fabsf(el.y) > so.h_out[1] || fabsf(el.z) > so.h_out[2];
            }
            if ((so.flags & VRLB_SOLID_INNER_OBB) && enters) {
                const float ax = (-so.h_in[0] - ol.x) * il.x, bx = (so.h_in[0] - ol.x) * il.x;
                const float ay = (-so.h_in[1] - ol.y) * il.y, by = (so.h_in[1] - ol.y) * il.y;
                const float az = (-so.h_in[2] - ol.z) * il.z, bz = (so.h_in[2] - ol.z) * il.z;
                const float t0 = fmaxf(fmaxf(0.f, fminf(ax, bx)), fmaxf(fminf(ay, by), fminf(az, bz)));
                const float t1 = fminf(fminf(1.f, fmaxf(ax, bx)), fminf(fmaxf(ay, by), fmaxf(az, bz)));
                if (t0 <= t1) return true;
            }
        }
        if (traverse_any_hit(s, so.root, o, d, inv, mint, maxt)) return true;
    }
    return false;
}

__device__ inline V3 tri_normal(const DTri &t) {
    return normalize(cross(V3(t.e1[0], t.e1[1], t.e1[2]), V3(t.e2[0], t.e2[1], t.e2[2])));
}
__device__ inline int target_medium(const DTri &t, V3 d) {      // interaction.h:186-200
    return dot(d, tri_normal(t)) > 0.f ? VRLB_SURF_GET_EXTERIOR(t.flags) : VRLB_SURF_GET_INTERIOR(t.flags);
}

// Medium::evalTransmittance homogeneous branch (include/mitsuba/render/medium.h:112-117)
__device__ inline float homo_transmittance(float max_density, float maxt) {
    float val = expf(max_density * (0.f - maxt));
    return max_density != 0.f ? val : 1.0f;
}

// ---- grid ("heterogeneous") medium ------------------------------------------------------------------
// GridVolume::eval, 1 channel, trilinear, clamp wrap (src/textures/grid3d.cpp:219-340); no domain test, as there
__device__ inline float grid_eval(const DMedium &m, V3 p) {
    float pw[3] = { p.x, p.y, p.z }, w1[3], w0[3];
    int i0[3], i1[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float l  = __fmul_rn(__fsub_rn(pw[k], m.g_lo[k]), __fdiv_rn(1.0f, __fsub_rn(m.g_hi[k], m.g_lo[k])));
        float q  = fmaf(l, (float) m.g_res[k], -.5f);
        float fl = floorf(q);
        int   i  = (int) fl;
        w1[k] = __fsub_rn(q, fl); w0[k] = __fsub_rn(1.f, w1[k]);
        i0[k] = min(max(i, 0), m.g_res[k] - 1);
        i1[k] = min(max(i + 1, 0), m.g_res[k] - 1);
    }
    const float *g = m.grid;
    const size_t nx = (size_t) m.g_res[0], ny = (size_t) m.g_res[1];
    const size_t r00 = ((size_t) i0[2] * ny + i0[1]) * nx, r10 = ((size_t) i0[2] * ny + i1[1]) * nx,
                 r01 = ((size_t) i1[2] * ny + i0[1]) * nx, r11 = ((size_t) i1[2] * ny + i1[1]) * nx;
    float d000 = __ldg(g + r00 + i0[0]), d100 = __ldg(g + r00 + i1[0]), d010 = __ldg(g + r10 + i0[0]), d110 = __ldg(g + r10 + i1[0]);
    float d001 = __ldg(g + r01 + i0[0]), d101 = __ldg(g + r01 + i1[0]), d011 = __ldg(g + r11 + i0[0]), d111 = __ldg(g + r11 + i1[0]);
    float v00 = fmaf(w0[0], d000, __fmul_rn(w1[0], d100)), v01 = fmaf(w0[0], d001, __fmul_rn(w1[0], d101)),
          v10 = fmaf(w0[0], d010, __fmul_rn(w1[0], d110)), v11 = fmaf(w0[0], d011, __fmul_rn(w1[0], d111));
    float v0 = fmaf(w0[1], v00, __fmul_rn(w1[1], v10)), v1 = fmaf(w0[1], v01, __fmul_rn(w1[1], v11));
    return fmaf(w0[2], v0, __fmul_rn(w1[2], v1));
}

// BoundingBox::ray_intersect, include/mitsuba/core/bbox.h:302-325
__device__ inline bool bbox_ray_intersect(const float lo[3], const float hi[3], V3 o, V3 d, float &mint, float &maxt) {
    float oo[3] = { o.x, o.y, o.z }, dd[3] = { d.x, d.y, d.z };
    bool active = true;
    float t1p[3], t2p[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        active = active && ((dd[k] != 0.f) || ((oo[k] > lo[k]) || (oo[k] < hi[k])));
        float rcp = __fdiv_rn(1.0f, dd[k]);
        float t1 = __fmul_rn(__fsub_rn(lo[k], oo[k]), rcp), t2 = __fmul_rn(__fsub_rn(hi[k], oo[k]), rcp);
        t1p[k] = fminf(t1, t2); t2p[k] = fmaxf(t1, t2);
    }
    mint = fmaxf(t1p[0], fmaxf(t1p[1], t1p[2]));
    maxt = fminf(t2p[0], fminf(t2p[1], t2p[2]));
    return active && maxt >= mint;
}

// Medium::evalTransmittance, non-homogeneous branch (include/mitsuba/render/medium.h:119-154) with
// Medium::sample_interaction (src/librender/medium.cpp:36-75) and eval_tr_and_pdf (:77-89) inlined: the product of
// majorant transmittances over the accepted free-flight steps (the grid values do not enter).
// Draws: 1 channel + one per step, from the eval's sub-stream.
__device__ inline float grid_transmittance(const DMedium &m, V3 o_, V3 d, float maxt_, Pcg32s &sub) {
    const float kInfF = __int_as_float(0x7f800000);
    float t = 0.f, maxt = maxt_;
    const float majorant = m.max_density;
    (void) sub.next_u32();                             // channel draw; the majorant is grey
    float si_t = maxt_;
    float optical = 0.f;                               // sum of the accepted steps' majorant depths: ONE exp at the end (each
                                                       // factor exp(-tt * majorant) of the reference rounds separately; the
                                                       // decisions below -- what the estimator's value hangs on -- are untouched)
    V3 ro = o_; float rmint = 0.f, rmaxt = maxt_;
    // BoundingBox::ray_intersect per step (bbox.h:302-325): the direction does not change, its reciprocals are hoisted
    const float dd[3] = { d.x, d.y, d.z };
    const float rcp[3] = { __fdiv_rn(1.0f, d.x), __fdiv_rn(1.0f, d.y), __fdiv_rn(1.0f, d.z) };
    // A segment whose end points both lie inside the grid's box by a margin (1e-4 of the box + 1e-5; positions along it
    // carry ~1e-6 of rounding) keeps every step's origin inside and every step's exit distance beyond the remaining
    // length: the box test returns active with bmin <= 0 <= mint and bmax >= rmaxt, i.e. it changes nothing and is skipped.
    bool well_inside;
    {
        const V3 e = o_ + d * maxt_;
        const float mg[3] = { 1e-4f * (m.g_hi[0] - m.g_lo[0]) + 1e-5f, 1e-4f * (m.g_hi[1] - m.g_lo[1]) + 1e-5f, 1e-4f * (m.g_hi[2] - m.g_lo[2]) + 1e-5f };
        well_inside = fminf(o_.x, e.x) > m.g_lo[0] + mg[0] && fmaxf(o_.x, e.x) < m.g_hi[0] - mg[0] &&
                      fminf(o_.y, e.y) > m.g_lo[1] + mg[1] && fmaxf(o_.y, e.y) < m.g_hi[1] - mg[1] &&
                      fminf(o_.z, e.z) > m.g_lo[2] + mg[2] && fmaxf(o_.z, e.z) < m.g_hi[2] - mg[2];
    }
    while (t < maxt) {
        float sample = sub.next_f32();
        float bmin = -1.f, bmax = kInfF;
        bool active = true;
        if (!well_inside) {
            const float oo[3] = { ro.x, ro.y, ro.z };
            float t1p[3], t2p[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                active = active && ((dd[k] != 0.f) || ((oo[k] > m.g_lo[k]) || (oo[k] < m.g_hi[k])));
                float t1 = __fmul_rn(__fsub_rn(m.g_lo[k], oo[k]), rcp[k]), t2 = __fmul_rn(__fsub_rn(m.g_hi[k], oo[k]), rcp[k]);
                t1p[k] = fminf(t1, t2); t2p[k] = fmaxf(t1, t2);
            }
            bmin = fmaxf(t1p[0], fmaxf(t1p[1], t1p[2]));
            bmax = fminf(t2p[0], fminf(t2p[1], t2p[2]));
            active = active && bmax >= bmin;
            active = active && (isfinite(bmin) || isfinite(bmax));
            if (!active) { bmin = 0.f; bmax = kInfF; }
        }
        float mint = fmaxf(rmint, bmin), mx = fminf(rmaxt, bmax);
        float sampled_t = __fadd_rn(mint, __fdiv_rn(-logf(__fsub_rn(1.f, sample)), majorant));
        bool valid_mi   = active && (sampled_t <= mx);
        float mi_t      = valid_mi ? sampled_t : kInfF;
        V3 mi_p         = ro + d * sampled_t;
        if (!(mi_t < kInfF)) break;                    // !mi.is_valid()
        if (si_t < mi_t) mi_t = kInfF;
        float tt = __fsub_rn(fminf(mi_t, si_t), mint);
        optical = __fadd_rn(optical, __fmul_rn(tt, majorant));
        t = __fadd_rn(t, mi_t);
        si_t = __fsub_rn(si_t, mi_t);
        ro    = mi_p;                                  // mi.spawn_ray(_ray.d), interaction.h:58-61
        rmint = __fmul_rn(__fadd_rn(1.f, fmaxf(fabsf(mi_p.x), fmaxf(fabsf(mi_p.y), fabsf(mi_p.z)))), kRayEpsilon);
        rmaxt = fmaxf(0.0f, __fsub_rn(maxt, t));
    }
    return expf(-optical);
}

// IsotropicPhaseFunction::eval (src/phase/isotropic.cpp:40-44) / HGPhaseFunction::eval (src/phase/hg.cpp:51-54,77-81)
__device__ inline float phase_eval(int phase, float g, V3 wi, V3 wo) {
    if (phase == 0) return kInvFourPi;
    float cos_theta = dot(wo, wi);
    float temp      = 1.0f + g * g + 2.0f * g * cos_theta;
    return kInvFourPi * (1 - g * g) / (temp * sqrtf(temp));
}

}  // namespace vrlb
