/*
 * vrl_oracle.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see vrl_oracle.h).
 *
 * Scalar fp32 CPU restatement of the reference's VRL gather.  Build:
 *   g++ -O2 -std=c++17 -ffp-contract=off -fno-fast-math -fPIC -shared -pthread
 * -ffp-contract=off matters: every '+', '-', '*', '/', sqrt below is exactly one IEEE binary32
 * operation in the written order, which is the arithmetic the CUDA kernels reproduce
 * (they are built with -fmad=false -prec-div=true -prec-sqrt=true).  Transcendentals go to the
 * platform libm, as the reference's do through Enoki's scalar fallbacks.
 *
 * Every function cites the reference lines it follows (paths relative to the reference root).
 */
#include "vrl_oracle.h"

#include <algorithm>
#include <array>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <queue>
#include <thread>
#include <vector>

namespace {

/* ------------------------------------------------------------------------------------------ */
/*  small vector type; every operation is componentwise scalar fp32                           */
/* ------------------------------------------------------------------------------------------ */
struct V3 {
    float x, y, z;
    V3() : x(0), y(0), z(0) {}
    V3(float a) : x(a), y(a), z(a) {}
    V3(float a, float b, float c) : x(a), y(b), z(c) {}
    explicit V3(const float *p) : x(p[0]), y(p[1]), z(p[2]) {}
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
inline V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
inline V3 operator*(V3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(float s, V3 a) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(V3 a, V3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 operator/(V3 a, float s) { return V3(a.x / s, a.y / s, a.z / s); }
/* dot = (x*x' + y*y') + z*z' : defined summation order of this oracle */
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(V3 a, V3 b) {
    return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
inline float squared_norm(V3 a) { return dot(a, a); }
inline float norm(V3 a) { return std::sqrt(dot(a, a)); }
inline V3 normalize(V3 a) { return a / norm(a); }
inline float safe_sqrt(float x) { return std::sqrt(std::max(x, 0.0f)); }
/* enoki::lerp = fmadd(b, t, fnmadd(a, t, a)) (array_math.h:1352); same order, uncontracted */
inline float lerp(float a, float b, float t) { return b * t + (a - a * t); }
inline float clampf(float v, float lo, float hi) { return std::min(std::max(v, lo), hi); }
inline float hmax_abs(V3 p) { return std::max(std::fabs(p.x), std::max(std::fabs(p.y), std::fabs(p.z))); }

constexpr float kEpsilon       = 5.9604644775390625e-08f;   /* math::Epsilon<float> = eps/2, math.h:36 */
constexpr float kRayEpsilon    = kEpsilon * 1500.0f;        /* math.h:37 */
constexpr float kShadowEpsilon = kRayEpsilon * 10.0f;       /* math.h:38 */
constexpr float kInf           = std::numeric_limits<float>::infinity();
constexpr float kInvFourPi     = 0.07957747154594766788f;

/* ------------------------------------------------------------------------------------------ */
/*  PCG32 -- ext/enoki/include/enoki/random.h:62-77 (seed), 69-77 (next_uint32),              */
/*  117-119 (next_float32), 265-286 (advance)                                                  */
/* ------------------------------------------------------------------------------------------ */
constexpr uint64_t PCG32_DEFAULT_STATE  = 0x853c49e6748fea9bULL;
constexpr uint64_t PCG32_DEFAULT_STREAM = 0xda3e39cb94b95bdbULL;
constexpr uint64_t PCG32_MULT           = 0x5851f42d4c957f2dULL;

struct Pcg32 {
    uint64_t state, inc;
    void seed(uint64_t initstate, uint64_t initseq) {
        state = 0;
        inc   = (initseq << 1) | 1u;
        next_u32();
        state += initstate;
        next_u32();
    }
    uint32_t next_u32() {
        uint64_t old = state;
        state        = old * PCG32_MULT + inc;
        uint32_t xorshifted = (uint32_t) (((old >> 18) ^ old) >> 27);
        uint32_t rot        = (uint32_t) (old >> 59);
        return (xorshifted >> rot) | (xorshifted << ((32 - rot) & 31));
    }
    float next_f32() {
        uint32_t bits = (next_u32() >> 9) | 0x3f800000u;
        float f;
        std::memcpy(&f, &bits, 4);
        return f - 1.0f;
    }
    void advance(int64_t delta_) {
        uint64_t cur_mult = PCG32_MULT, cur_plus = inc, acc_mult = 1, acc_plus = 0;
        uint64_t delta = (uint64_t) delta_;
        while (delta != 0) {
            if (delta & 1) {
                acc_mult *= cur_mult;
                acc_plus = acc_plus * cur_mult + cur_plus;
            }
            cur_plus = (cur_mult + 1) * cur_plus;
            cur_mult *= cur_mult;
            delta >>= 1;
        }
        state = acc_mult * state + acc_plus;
    }
};

/* IndependentSampler over PCG32Sampler::seed (sampler.cpp:83-96): scalar variant seeds
 * m_rng.seed(base_seed + seed_offset, PCG32_DEFAULT_STREAM). */
struct Sampler {
    Pcg32 rng;
    void seed(uint64_t base_seed, uint64_t seed_offset) { rng.seed(base_seed + seed_offset, PCG32_DEFAULT_STREAM); }
    float next_1d() { return rng.next_f32(); }
};

/* enoki::morton_decode<Point2u> (ext/enoki/include/enoki/morton.h:148-155): x = even bits, y = odd bits */
inline uint32_t compact_bits(uint32_t v) {
    v &= 0x55555555u;
    v = (v | (v >> 1)) & 0x33333333u;
    v = (v | (v >> 2)) & 0x0f0f0f0fu;
    v = (v | (v >> 4)) & 0x00ff00ffu;
    v = (v | (v >> 8)) & 0x0000ffffu;
    return v;
}

/* Spiral::Spiral/reset/next_block (src/librender/spiral.cpp:8-72): ids in spiral order from the centre */
std::vector<uint32_t> spiral_ids(int w, int h, int bs, int &nbx, int &nby) {
    nbx = (w + bs - 1) / bs;
    nby = (h + bs - 1) / bs;
    int count = nbx * nby;
    std::vector<uint32_t> ids((size_t) count, 0u);
    int px = nbx / 2, py = nby / 2, dir = 0 /* Right */, steps_left = 1, steps = 1;
    for (int counter = 0; counter < count; ++counter) {
        ids[(size_t) py * nbx + px] = (uint32_t) counter;
        if (counter + 1 != count) {
            do {
                switch (dir) {
                    case 0: ++px; break;   /* Right */
                    case 1: ++py; break;   /* Down  */
                    case 2: --px; break;   /* Left  */
                    case 3: --py; break;   /* Up    */
                }
                if (--steps_left == 0) {
                    dir = (dir + 1) % 4;
                    if (dir == 2 || dir == 0)
                        ++steps;
                    steps_left = steps;
                }
            } while (px < 0 || py < 0 || px >= nbx || py >= nby);
        }
    }
    return ids;
}

}  // namespace

extern "C" {

void vrlo_pcg32_seed(uint64_t initstate, uint64_t initseq, uint64_t *state, uint64_t *inc) {
    Pcg32 r; r.seed(initstate, initseq); *state = r.state; *inc = r.inc;
}
uint32_t vrlo_pcg32_next_u32(uint64_t *state, uint64_t inc) {
    Pcg32 r{ *state, inc }; uint32_t v = r.next_u32(); *state = r.state; return v;
}
float vrlo_pcg32_next_f32(uint64_t *state, uint64_t inc) {
    Pcg32 r{ *state, inc }; float v = r.next_f32(); *state = r.state; return v;
}
void vrlo_pcg32_advance(uint64_t *state, uint64_t inc, int64_t delta) {
    Pcg32 r{ *state, inc }; r.advance(delta); *state = r.state;
}
void vrlo_sampler_stream(uint64_t base_seed, uint64_t seed_offset, uint32_t n, float *out_f, uint32_t *out_u) {
    Sampler s; s.seed(base_seed, seed_offset);
    for (uint32_t i = 0; i < n; ++i) {
        Pcg32 copy = s.rng;
        if (out_u) out_u[i] = copy.next_u32();
        float f = s.next_1d();
        if (out_f) out_f[i] = f;
    }
}
uint32_t vrlo_spiral_block_ids(int32_t width, int32_t height, int32_t block_size, uint32_t *block_id, uint32_t cap) {
    int nbx, nby;
    std::vector<uint32_t> ids = spiral_ids(width, height, block_size, nbx, nby);
    for (size_t i = 0; i < ids.size() && i < cap; ++i) block_id[i] = ids[i];
    return (uint32_t) ids.size();
}
void vrlo_morton_decode(uint32_t i, uint32_t *x, uint32_t *y) {
    *x = compact_bits(i);
    *y = compact_bits(i >> 1);
}

}  // extern "C"

namespace {

/* ------------------------------------------------------------------------------------------ */
/*  Scene: triangle soup + a conservative BVH (stands in for ShapeKDTree, kdtree.h:2069-2172;  */
/*  only the set of triangle hits matters, and that is decided by the exact test below).      */
/* ------------------------------------------------------------------------------------------ */
struct Tri { V3 p0, p1, p2; uint32_t flags; float prm[3]; };   /* prm = the triangle's tri_reflectance entry (0 when absent) */

struct BvhNode { float lo[3], hi[3]; int left, right, first, count, axis; };

struct Hit { float t, u, v; int prim; bool valid() const { return prim >= 0; } };

}  // namespace

struct vrlo_scene {
    std::vector<Tri> tris;
    std::vector<vrlb_medium> media;
    std::vector<std::vector<float>> grids;
    std::vector<vrlb_emitter> emitters;
    std::vector<BvhNode> nodes;
    std::vector<int> order;
    float pad = 0.f;
};

namespace {

/* Mesh::ray_intersect_triangle, include/mitsuba/render/mesh.h:195-225 (Moeller-Trumbore) */
inline bool intersect_tri(const Tri &tr, V3 o, V3 d, float mint, float maxt, float &t, float &u, float &v) {
    V3 e1 = tr.p1 - tr.p0, e2 = tr.p2 - tr.p0;
    V3 pvec       = cross(d, e2);
    float inv_det = 1.0f / dot(e1, pvec);
    V3 tvec       = o - tr.p0;
    u             = dot(tvec, pvec) * inv_det;
    if (!(u >= 0.f && u <= 1.f)) return false;
    V3 qvec = cross(tvec, e1);
    v       = dot(d, qvec) * inv_det;
    if (!(v >= 0.f && u + v <= 1.f)) return false;
    t = dot(e2, qvec) * inv_det;
    if (!(t >= mint && t <= maxt)) return false;
    return true;
}

void build_bvh(vrlo_scene &s) {
    size_t n = s.tris.size();
    s.order.resize(n);
    for (size_t i = 0; i < n; ++i) s.order[i] = (int) i;
    s.nodes.clear();
    if (n == 0) return;
    float diag_lo[3] = { kInf, kInf, kInf }, diag_hi[3] = { -kInf, -kInf, -kInf };
    for (auto &t : s.tris)
        for (int k = 0; k < 3; ++k) {
            float a = std::min(t.p0[k], std::min(t.p1[k], t.p2[k])), b = std::max(t.p0[k], std::max(t.p1[k], t.p2[k]));
            diag_lo[k] = std::min(diag_lo[k], a); diag_hi[k] = std::max(diag_hi[k], b);
        }
    float ext = 0.f;
    for (int k = 0; k < 3; ++k) ext = std::max(ext, std::max(std::fabs(diag_lo[k]), std::fabs(diag_hi[k])));
    s.pad = 1e-4f * ext + 1e-6f;   /* conservative: boxes only ever admit more triangles to the exact test */

    struct Item { int begin, end, node; };
    std::vector<Item> stack;
    s.nodes.push_back(BvhNode());
    stack.push_back({ 0, (int) n, 0 });
    while (!stack.empty()) {
        Item it = stack.back(); stack.pop_back();
        BvhNode nd;
        float clo[3] = { kInf, kInf, kInf }, chi[3] = { -kInf, -kInf, -kInf };
        for (int k = 0; k < 3; ++k) { nd.lo[k] = kInf; nd.hi[k] = -kInf; }
        for (int i = it.begin; i < it.end; ++i) {
            const Tri &t = s.tris[s.order[i]];
            for (int k = 0; k < 3; ++k) {
                float a = std::min(t.p0[k], std::min(t.p1[k], t.p2[k])), b = std::max(t.p0[k], std::max(t.p1[k], t.p2[k]));
                nd.lo[k] = std::min(nd.lo[k], a); nd.hi[k] = std::max(nd.hi[k], b);
                float c = (t.p0[k] + t.p1[k] + t.p2[k]) / 3.0f;
                clo[k] = std::min(clo[k], c); chi[k] = std::max(chi[k], c);
            }
        }
        for (int k = 0; k < 3; ++k) { nd.lo[k] -= s.pad; nd.hi[k] += s.pad; }
        nd.left = nd.right = -1; nd.first = it.begin; nd.count = it.end - it.begin; nd.axis = 0;
        if (nd.count > 4) {
            /* binned SAH over the three axes (16 bins on centroids); falls back to the median */
            auto tri_lohi = [&](int idx, int k, float &a, float &b) {
                const Tri &t = s.tris[idx];
                a = std::min(t.p0[k], std::min(t.p1[k], t.p2[k])); b = std::max(t.p0[k], std::max(t.p1[k], t.p2[k]));
            };
            auto centroid = [&](int idx, int k) { const Tri &t = s.tris[idx]; return (t.p0[k] + t.p1[k] + t.p2[k]) / 3.0f; };
            const int NB = 16;
            int axis = -1, best_bin = -1; float best_cost = kInf;
            for (int k = 0; k < 3; ++k) {
                float ext = chi[k] - clo[k];
                if (!(ext > 0.f)) continue;
                int cnt[NB] = { 0 }; float blo[NB][3], bhi[NB][3];
                for (int b = 0; b < NB; ++b) for (int c = 0; c < 3; ++c) { blo[b][c] = kInf; bhi[b][c] = -kInf; }
                for (int i = it.begin; i < it.end; ++i) {
                    int idx = s.order[i];
                    int b = std::min(NB - 1, (int) ((centroid(idx, k) - clo[k]) / ext * NB));
                    cnt[b]++;
                    for (int c = 0; c < 3; ++c) { float a, bb; tri_lohi(idx, c, a, bb); blo[b][c] = std::min(blo[b][c], a); bhi[b][c] = std::max(bhi[b][c], bb); }
                }
                auto area = [](const float *lo, const float *hi) {
                    float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
                    return (dx < 0 || dy < 0 || dz < 0) ? 0.f : 2.f * (dx * dy + dy * dz + dz * dx);
                };
                float la[NB], ra[NB]; int lc[NB], rc[NB];
                float lo3[3] = { kInf, kInf, kInf }, hi3[3] = { -kInf, -kInf, -kInf }; int c0 = 0;
                for (int b = 0; b < NB; ++b) { for (int c = 0; c < 3; ++c) { lo3[c] = std::min(lo3[c], blo[b][c]); hi3[c] = std::max(hi3[c], bhi[b][c]); } c0 += cnt[b]; la[b] = area(lo3, hi3); lc[b] = c0; }
                for (int c = 0; c < 3; ++c) { lo3[c] = kInf; hi3[c] = -kInf; } c0 = 0;
                for (int b = NB - 1; b >= 0; --b) { for (int c = 0; c < 3; ++c) { lo3[c] = std::min(lo3[c], blo[b][c]); hi3[c] = std::max(hi3[c], bhi[b][c]); } c0 += cnt[b]; ra[b] = area(lo3, hi3); rc[b] = c0; }
                for (int b = 0; b + 1 < NB; ++b) {
                    if (lc[b] == 0 || rc[b + 1] == 0) continue;
                    float cost = la[b] * lc[b] + ra[b + 1] * rc[b + 1];
                    if (cost < best_cost) { best_cost = cost; axis = k; best_bin = b; }
                }
            }
            int mid;
            if (axis >= 0) {
                float ext = chi[axis] - clo[axis];
                auto bin_of = [&](int idx) { return std::min(NB - 1, (int) ((centroid(idx, axis) - clo[axis]) / ext * NB)); };
                auto itp = std::stable_partition(s.order.begin() + it.begin, s.order.begin() + it.end, [&](int idx) { return bin_of(idx) <= best_bin; });
                mid = (int) (itp - s.order.begin());
            } else {
                axis = 0; mid = (it.begin + it.end) / 2;
            }
            nd.count = 0; nd.axis = axis;
            nd.left  = (int) s.nodes.size(); s.nodes.push_back(BvhNode());
            nd.right = (int) s.nodes.size(); s.nodes.push_back(BvhNode());
            stack.push_back({ it.begin, mid, nd.left });
            stack.push_back({ mid, it.end, nd.right });
        }
        s.nodes[it.node] = nd;
    }
}

inline bool slab(const BvhNode &nd, V3 o, V3 inv, float mint, float maxt) {
    float t0 = mint, t1 = maxt;
    for (int k = 0; k < 3; ++k) {
        float a = (nd.lo[k] - o[k]) * inv[k], b = (nd.hi[k] - o[k]) * inv[k];
        float tn = std::fmin(a, b), tf = std::fmax(a, b);   /* fmin/fmax drop NaNs (0 * inf) */
        t0 = std::fmax(t0, tn); t1 = std::fmin(t1, tf);
    }
    return t0 <= t1 * 1.00001f + 1e-6f;
}

/* Scene::ray_intersect (scene.cpp:114-131): closest hit in [mint, maxt]; ties -> lowest prim index */
Hit ray_intersect(const vrlo_scene &s, V3 o, V3 d, float mint, float maxt) {
    Hit best{ kInf, 0.f, 0.f, -1 };
    if (s.nodes.empty()) return best;
    V3 inv(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    int stack[128], sp = 0;
    stack[sp++] = 0;
    while (sp) {
        const BvhNode &nd = s.nodes[stack[--sp]];
        float far = best.valid() ? best.t : maxt;
        if (!slab(nd, o, inv, mint, far)) continue;
        if (nd.left < 0) {
            for (int i = nd.first; i < nd.first + nd.count; ++i) {
                int prim = s.order[i];
                float t, u, v;
                if (intersect_tri(s.tris[prim], o, d, mint, maxt, t, u, v)) {
                    if (t < best.t || (t == best.t && prim < best.prim)) best = Hit{ t, u, v, prim };
                }
            }
        } else {
            /* near child last on the stack = visited first (left holds the smaller centroids) */
            if (d[nd.axis] >= 0.f) { stack[sp++] = nd.right; stack[sp++] = nd.left; }
            else                   { stack[sp++] = nd.left;  stack[sp++] = nd.right; }
        }
    }
    return best;
}

/* geometric normal of Mesh::compute_surface_interaction (mesh.cpp:477-487) */
inline V3 tri_normal(const Tri &t) { return normalize(cross(t.p1 - t.p0, t.p2 - t.p0)); }
/* si.p re-interpolated from barycentrics (mesh.cpp:467-484) */
inline V3 tri_point(const Tri &t, float b1, float b2) {
    float b0 = 1.f - b1 - b2;
    return t.p0 * b0 + t.p1 * b1 + t.p2 * b2;
}
/* SurfaceInteraction::target_medium (interaction.h:186-200): exterior when cos > 0 else interior */
inline int target_medium(const Tri &t, V3 d) {
    return dot(d, tri_normal(t)) > 0.f ? VRLB_SURF_GET_EXTERIOR(t.flags) : VRLB_SURF_GET_INTERIOR(t.flags);
}
inline bool is_medium_transition(const Tri &t) {
    return VRLB_SURF_GET_EXTERIOR(t.flags) >= 0 || VRLB_SURF_GET_INTERIOR(t.flags) >= 0;
}
inline bool is_null(const Tri &t) { return (t.flags & VRLB_SURF_BSDF_MASK) == VRLB_SURF_BSDF_NULL; }

/* coordinate_system, include/mitsuba/core/vector.h (Duff et al.) */
inline void coordinate_system(V3 n, V3 &s, V3 &t) {
    float sign = std::copysign(1.f, n.z), a = -1.f / (sign + n.z), b = n.x * n.y * a;
    s = V3(1.f + sign * n.x * n.x * a, sign * b, -sign * n.x);
    t = V3(b, sign + n.y * n.y * a, -n.y);
}
/* bsdf->sample(ctx, si, sample1, sample2) for the two delta BSDFs of the specular chains, default BSDFContext
 * (TransportMode::Radiance): SmoothConductor::sample (src/bsdfs/conductor.cpp: one-sided, wo = reflect(wi), weight =
 * specular_reflectance x Fresnel -- the descriptor carries the product, taken angle-independent) and
 * SmoothDielectric::sample (src/bsdfs/dielectric.cpp with fresnel() of include/mitsuba/core/fresnel.h: reflect when
 * sample1 <= r_i, else refract with the eta_ti^2 radiance scaling).  Returns false when the lobe is inactive. */
inline bool sample_delta(const Tri &tr, V3 n, V3 rd, float sample1, V3 &wo_world, V3 &weight, bool &transmitted) {
    const uint32_t kind = tr.flags & VRLB_SURF_BSDF_MASK;
    const bool twosided = (tr.flags & VRLB_SURF_TWOSIDED) != 0;
    transmitted = false;
    V3 s, t; coordinate_system(n, s, t);
    const V3 w = -rd;
    const V3 wi(dot(w, s), dot(w, t), dot(w, n));
    V3 wo;
    if (kind == VRLB_SURF_BSDF_MIRROR) {
        if (!(wi.z > 0.f) && !(twosided && wi.z < 0.f)) return false;
        wo = V3(-wi.x, -wi.y, wi.z);
        weight = V3(tr.prm[0], tr.prm[1], tr.prm[2]);
    } else {
        const float eta = tr.prm[0], cos_theta_i = wi.z;
        const bool outside_mask = cos_theta_i >= 0.f;
        const float rcp_eta = 1.f / eta, eta_it = outside_mask ? eta : rcp_eta, eta_ti = outside_mask ? rcp_eta : eta;
        const float cos_theta_t_sqr = 1.f - (1.f - cos_theta_i * cos_theta_i) * (eta_ti * eta_ti);
        const float cos_theta_i_abs = std::fabs(cos_theta_i), cos_theta_t_abs = std::sqrt(std::max(cos_theta_t_sqr, 0.f));
        const bool special_case = eta == 1.f || cos_theta_i_abs == 0.f;
        const float a_s = (cos_theta_i_abs - eta_it * cos_theta_t_abs) / (cos_theta_i_abs + eta_it * cos_theta_t_abs);
        const float a_p = (cos_theta_t_abs - eta_it * cos_theta_i_abs) / (cos_theta_t_abs + eta_it * cos_theta_i_abs);
        float r_i = 0.5f * (a_s * a_s + a_p * a_p);
        if (special_case) r_i = eta == 1.f ? 0.f : 1.f;
        const float cos_theta_t = cos_theta_i >= 0.f ? -cos_theta_t_abs : cos_theta_t_abs;
        if (sample1 <= r_i) { wo = V3(-wi.x, -wi.y, wi.z); weight = V3(tr.prm[1]); }
        else {
            wo = V3(-eta_ti * wi.x, -eta_ti * wi.y, cos_theta_t);
            weight = V3(tr.prm[2] * (eta_ti * eta_ti));
            transmitted = true;
        }
    }
    wo_world = s * wo.x + t * wo.y + n * wo.z;
    return true;
}

/* ------------------------------------------------------------------------------------------ */
/*  Media and phase functions                                                                 */
/* ------------------------------------------------------------------------------------------ */
/* ConstVolume::max() returns the MEAN of the colour (constant3d.cpp:37, srgb.h mean) and
 * m_max_density = scale * sigma_t->max() (homogeneous.cpp:26, nonlinear.cpp:53): grey extinction */
inline float max_density(const vrlb_medium &m) {
    if (m.kind == VRLB_MEDIUM_GRID) return m.scale * m.grid_max;             /* heterogeneous.cpp:29 */
    float mean = (m.sigma_t[0] + m.sigma_t[1] + m.sigma_t[2]) / 3.0f;        /* hmean */
    return m.scale * mean;
}
inline bool is_homogeneous(const vrlb_medium &m) { return m.kind != VRLB_MEDIUM_GRID; }

/* Medium::evalTransmittance, homogeneous branch (medium.h:112-117) and homoEvalTransmittance (103-107) */
inline float homo_transmittance(const vrlb_medium &m, float mint, float maxt) {
    float md        = max_density(m);
    float negLength = mint - maxt;
    float val       = std::exp(md * negLength);
    return md != 0.f ? val : 1.0f;
}

/* get_scattering_coefficients: sigma_s = scale*sigma_t*albedo (homogeneous.cpp:42-50, nonlinear.cpp:414-420) */
inline V3 sigma_s_const(const vrlb_medium &m) {
    V3 sigmat = V3(m.sigma_t) * m.scale;
    return sigmat * V3(m.albedo);
}

/* GridVolume::eval for a 1-channel grid: world -> [0,1]^3 by the `.vol` bbox (volume_data.h:25-33, 79-81; applied
 * here as (p - min) * rcp(extent)), then trilinear interpolation with clamp wrap (grid3d.cpp:236-340: fmadd(p, shape,
 * -.5), floor2int, the eight taps at ((z*ny + y)*nx + x), fmadd lerps in x, y, z order).  The reference does not test
 * the domain (eval_impl, grid3d.cpp:219-230): points outside see the clamped border values. */
inline float grid_eval(const vrlb_medium &m, V3 p) {
    const int n[3] = { m.grid_res[0], m.grid_res[1], m.grid_res[2] };
    float pw[3] = { p.x, p.y, p.z }, w1[3], w0[3];
    int i0[3], i1[3];
    for (int k = 0; k < 3; ++k) {
        float l  = (pw[k] - m.grid_bbox_min[k]) * (1.0f / (m.grid_bbox_max[k] - m.grid_bbox_min[k]));
        float q  = std::fma(l, (float) n[k], -.5f);
        float fl = std::floor(q);
        int   i  = (int) fl;
        w1[k] = q - fl; w0[k] = 1.f - w1[k];
        i0[k] = std::min(std::max(i, 0), n[k] - 1);
        i1[k] = std::min(std::max(i + 1, 0), n[k] - 1);
    }
    auto at = [&](int x, int y, int z) { return m.grid[((size_t) z * n[1] + y) * n[0] + x]; };
    float d000 = at(i0[0], i0[1], i0[2]), d100 = at(i1[0], i0[1], i0[2]), d010 = at(i0[0], i1[1], i0[2]),
          d110 = at(i1[0], i1[1], i0[2]), d001 = at(i0[0], i0[1], i1[2]), d101 = at(i1[0], i0[1], i1[2]),
          d011 = at(i0[0], i1[1], i1[2]), d111 = at(i1[0], i1[1], i1[2]);
    float v00 = std::fma(w0[0], d000, w1[0] * d100), v01 = std::fma(w0[0], d001, w1[0] * d101),
          v10 = std::fma(w0[0], d010, w1[0] * d110), v11 = std::fma(w0[0], d011, w1[0] * d111);
    float v0 = std::fma(w0[1], v00, w1[1] * v10), v1 = std::fma(w0[1], v01, w1[1] * v11);
    return std::fma(w0[2], v0, w1[2] * v1);
}
/* get_scattering_coefficients at a point: heterogeneous.cpp:43-51 for the grid medium (sigma_t = scale * grid(p),
 * sigma_s = sigma_t * albedo with a constant albedo volume), else the constant coefficients */
inline V3 sigma_s_at(const vrlb_medium &m, V3 p) {
    if (m.kind != VRLB_MEDIUM_GRID) return sigma_s_const(m);
    float sigmat = m.scale * grid_eval(m, p);
    return V3(m.albedo) * sigmat;
}

/* IsotropicPhaseFunction::eval = square_to_uniform_sphere_pdf = 1/(4 pi) (isotropic.cpp:40-44);
 * HGPhaseFunction::eval = eval_hg(dot(wo, mi.wi)) (hg.cpp:51-54, 77-81) */
inline float phase_eval(const vrlb_medium &m, V3 wi, V3 wo) {
    if (m.phase == VRLB_PHASE_ISOTROPIC) return kInvFourPi;
    float g = m.g, cos_theta = dot(wo, wi);
    float temp = 1.0f + g * g + 2.0f * g * cos_theta;
    return kInvFourPi * (1 - g * g) / (temp * std::sqrt(temp));
}

}  // namespace

namespace {

/* ------------------------------------------------------------------------------------------ */
/*  VRL (vrl_struct.h:60-86, 664-677) and camera ray part                                     */
/* ------------------------------------------------------------------------------------------ */
struct Vrl { V3 origin, direction; float length; V3 flux; int medium; };
struct Ray { V3 o, d; float maxt; };

struct ClosestPointInfo { float tCam, tVRL; };

/* VRL::findClosetPoint, vrl_struct.h:163-228 */
ClosestPointInfo find_closest_point(const Vrl &vrl, const Ray &r) {
    V3 u = vrl.direction * vrl.length;
    V3 v = r.d * r.maxt;
    V3 w = vrl.origin - r.o;

    float a = dot(u, u), b = dot(u, v), c = dot(v, v), d = dot(u, w), e = dot(v, w);
    float D = a * c - b * b;
    float sc, sN, sD = D;
    float tc, tN, tD = D;

    if (D < 0.0001f) {
        sN = 0.0f; sD = 1.0f; tN = e; tD = c;
    } else {
        sN = (b * e - c * d);
        tN = (a * e - b * d);
        if (sN < 0.0f) {
            sN = 0.0f; tN = e; tD = c;
        } else if (sN > sD) {
            sN = sD; tN = e + b; tD = c;
        }
    }

    if (tN < 0.0f) {
        tN = 0.0f;
        if (-d < 0.0f)      sN = 0.0f;
        else if (-d > a)    sN = sD;
        else { sN = -d; sD = a; }
    } else if (tN > tD) {
        tN = tD;
        if ((-d + b) < 0.0f)    sN = 0;
        else if ((-d + b) > a)  sN = sD;
        else { sN = (-d + b); sD = a; }
    }

    sc = (std::fabs(sN) < 0.0001f ? 0.0f : sN / sD);
    tc = (std::fabs(tN) < 0.0001f ? 0.0f : tN / tD);
    return ClosestPointInfo{ tc * r.maxt, sc * vrl.length };
}

struct SamplingInfo { V3 pCam; float tCam; V3 pVRL; float tVRL; float invPDF; bool valid; };
inline SamplingInfo sampling_invalid() { return SamplingInfo{ V3(0.f), 0.f, V3(0.f), 0.f, 0.f, false }; }

/* VRL::sampleMC, vrl_struct.h:239-243, with NLRay::at (44-56) for a single part */
SamplingInfo sample_mc(const Vrl &vrl, const Ray &ray, float u1, float u2) {
    float tCam = u1 * ray.maxt;
    float tVRL = u2 * vrl.length;
    return SamplingInfo{ ray.o + ray.d * tCam, tCam, vrl.origin + vrl.direction * tVRL, tVRL, vrl.length * ray.maxt, true };
}

/* Source of the two uniforms: draws lazily so that the sequential mode consumes exactly what the
 * reference consumes (no draw before the early-outs at vrl_struct.h:264, 275). */
struct DrawSrc {
    Pcg32 *rng;        /* non-null: draw from the stream */
    float u[2];        /* else: explicit uniforms        */
    int k;
    float next() { return rng ? rng->next_f32() : u[k++]; }
};

/* VRL::samplingVRL importance branch, vrl_struct.h:251-360 (the `if (true)` arm) */
SamplingInfo sampling_vrl(const Vrl &vrl, const Ray &ray, bool uniform, DrawSrc &draws) {
    if (uniform) {
        float u1 = draws.next(), u2 = draws.next();
        return sample_mc(vrl, ray, u1, u2);
    }
    ClosestPointInfo cp = find_closest_point(vrl, ray);
    float h;
    {
        V3 vh = vrl.origin + vrl.direction * cp.tVRL;
        V3 uh = ray.o + ray.d * cp.tCam;
        h     = squared_norm(uh - vh);
    }
    if (h == 0.0f) return sampling_invalid();
    h = safe_sqrt(h);

    float v0Hat = -cp.tVRL;
    float v1Hat = vrl.length + v0Hat;

    float sinTheta = squared_norm(cross(vrl.direction, ray.d));
    if (sinTheta == 0) return sampling_invalid();
    sinTheta = safe_sqrt(sinTheta);

    /* inverseCDF_A, vrl_struct.h:293-310 */
    auto A = [](float x, float h, float sinTheta) -> float {
        float value = (x / h) * sinTheta;
        return std::log(value + std::sqrt(value * value + 1));
    };
    float r  = draws.next();
    float a0 = A(v0Hat, h, sinTheta);
    float a1 = A(v1Hat, h, sinTheta);
    float vHat = (h * std::sinh(lerp(a0, a1, r))) / sinTheta;
    float invPDF_A = (a1 - a0) * std::sqrt(h * h + vHat * vHat * sinTheta * sinTheta);
    invPDF_A /= sinTheta;

    float tVRL = (vHat + cp.tVRL);
    V3 pVRL    = vrl.origin + vrl.direction * tVRL;

    /* Kulla-Fajardo equi-angular on the camera ray, vrl_struct.h:318-336 */
    float uHat  = dot(ray.d, (pVRL - ray.o));
    float u0Hat = -uHat;
    float u1Hat = ray.maxt + u0Hat;
    float hPoint = norm(ray.o + ray.d * uHat - pVRL);
    float eps    = draws.next();
    float thetaA = std::atan(u0Hat / hPoint);
    float thetaB = std::atan(u1Hat / hPoint);
    float uS     = hPoint * std::tan(lerp(thetaA, thetaB, eps));
    float invPDF_B = (thetaB - thetaA) * (hPoint * hPoint + uS * uS) / hPoint;
    float tCam = (uS - u0Hat);
    V3 pCam    = ray.o + ray.d * tCam;

    tCam = clampf(tCam, 0.0f, ray.maxt);      /* vrl_struct.h:340-341 */
    tVRL = clampf(tVRL, 0.0f, vrl.length);
    return SamplingInfo{ pCam, tCam, pVRL, tVRL, invPDF_A * invPDF_B, true };
}

/* Medium::evalTransmittance over [0, t] along a ray (medium.h:109-155).  Homogeneous / nonlinear
 * media: analytic grey exponential.  Grid media: see grid_transmittance() (part 5). */
float grid_transmittance(const vrlo_scene &s, const vrlb_medium &m, int medium_id, V3 o, V3 d, float maxt, Pcg32 *sub);

inline float medium_transmittance(const vrlo_scene &s, int medium_id, V3 o, V3 d, float maxt, Pcg32 *sub) {
    const vrlb_medium &m = s.media[(size_t) medium_id];
    if (is_homogeneous(m)) return homo_transmittance(m, 0.f, maxt);
    return grid_transmittance(s, m, medium_id, o, d, maxt, sub);
}

/* VRL::evalTransmittance (static), vrl_struct.h:509-572, called with p1OnSurface = p2OnSurface =
 * false and interactions = -1 (vrl_struct.h:630-632).  Grey result (all media return grey Tr).
 * Note Null::eval returns 0 (src/bsdfs/null.cpp:59-62): crossing ANY surface zeroes the result. */
float eval_transmittance(const vrlo_scene &s, bool trace, V3 p1, V3 p2, int medium_, Pcg32 *sub) {
    V3 d            = p2 - p1;
    float remaining = norm(d);
    d               = normalize(d);
    float lengthFactor = 1.f;
    V3 ro = p1; float mint = 0.f, maxt = remaining * lengthFactor;
    float transmittance = 1.0f;
    int interactions = 0;
    int medium = medium_;
    while (remaining > 0) {
        Hit si = trace ? ray_intersect(s, ro, d, mint, maxt) : Hit{ kInf, 0, 0, -1 };
        bool surface = si.valid();
        if (surface && !is_null(s.tris[(size_t) si.prim])) return 0.0f;
        if (medium >= 0)
            transmittance *= medium_transmittance(s, medium, ro, d, std::min(si.t, remaining), sub);
        if (!surface || transmittance == 0.0f) break;
        const Tri &tr = s.tris[(size_t) si.prim];
        transmittance *= 0.0f;                       /* bsdf->eval of the Null BSDF, null.cpp:59-62 */
        if (is_medium_transition(tr)) {
            if (medium != target_medium(tr, -d)) return 0.0f;
            medium = target_medium(tr, d);
        }
        if (++interactions > 100) break;
        ro = ro + d * si.t;
        remaining -= si.t;
        maxt = remaining * lengthFactor;
        mint = kEpsilon;
    }
    return transmittance;
}

/* VRL::getContrib, vrl_struct.h:584-662.  `wi` is NLRay::wi: the reference leaves it unset on the
 * importance path (quirk 1); defined here as -part.d, the commented-out intent (:611). */
V3 get_contrib_sampled(const vrlo_scene &s, bool trace, const Vrl &vrl, const Ray &ray, const SamplingInfo &sampling, Pcg32 *sub);
V3 get_contrib(const vrlo_scene &s, bool trace, const Vrl &vrl, const Ray &ray, bool uniform, DrawSrc &draws, Pcg32 *sub) {
    SamplingInfo sampling = sampling_vrl(vrl, ray, uniform, draws);
    return get_contrib_sampled(s, trace, vrl, ray, sampling, sub);
}
/* `ray` supplies wi = -ray.d (the part NLRay::at selected, vrl_struct.h:49) and the origin/direction of the
 * camera-side transmittance ray (ray.first(), vrl_struct.h:637-641; direction is irrelevant to the homogeneous Tr) */
V3 get_contrib_sampled(const vrlo_scene &s, bool trace, const Vrl &vrl, const Ray &ray, const SamplingInfo &sampling, Pcg32 *sub) {
    /* invalid() has pCam == pVRL == 0 -> lengthPtoP == 0 -> 0 */
    V3 dir           = sampling.pVRL - sampling.pCam;
    float lengthPtoP = norm(dir);
    if (lengthPtoP == 0) return V3(0.f);
    if (std::isnan(sampling.invPDF)) return V3(0.f);
    dir = dir / lengthPtoP;

    const vrlb_medium &m = s.media[(size_t) vrl.medium];
    V3 wi       = -ray.d;
    float rayPF = phase_eval(m, wi, dir);
    if (rayPF == 0) return V3(0.f);
    V3 sigmaSRay = sigma_s_at(m, sampling.pCam);

    float vrlPF = phase_eval(m, -vrl.direction, -dir);
    if (vrlPF == 0) return V3(0.f);
    V3 sigmaSVRL = sigma_s_at(m, sampling.pVRL);

    float vrlToRayTrans = eval_transmittance(s, trace, sampling.pCam, sampling.pVRL, vrl.medium, sub);
    if (vrlToRayTrans == 0.0f) return V3(0.f);
    float rayTrans = medium_transmittance(s, vrl.medium, ray.o, ray.d, sampling.tCam, sub);
    if (rayTrans == 0.0f) return V3(0.f);
    float vrlTrans = medium_transmittance(s, vrl.medium, vrl.origin, vrl.direction, sampling.tVRL, sub);
    if (vrlTrans == 0.0f) return V3(0.f);

    float fallOff = 1.0f / (lengthPtoP * lengthPtoP);
    V3 result = vrl.flux * fallOff * vrlPF * rayPF * V3(vrlTrans) * V3(rayTrans) * V3(vrlToRayTrans)
                * sigmaSRay * sigmaSVRL * sampling.invPDF;
    return result;
}

inline Vrl load_vrl(const vrlb_vrlset_desc &set, uint64_t j) {
    const float *c = set.soa10; uint64_t n = set.n;
    Vrl v;
    v.origin    = V3(c[0 * n + j], c[1 * n + j], c[2 * n + j]);
    v.direction = V3(c[3 * n + j], c[4 * n + j], c[5 * n + j]);
    v.length    = c[6 * n + j];
    v.flux      = V3(c[7 * n + j], c[8 * n + j], c[9 * n + j]);
    v.medium    = set.medium ? (int) set.medium[j] : 0;
    return v;
}

}  // namespace

namespace {

/* VRLMap::query, no-acceleration branch, vrl_map.h:133-211, for ONE camera part.
 * `rng` is the pixel's stream positioned at the first draw of this query.  In sequential mode it
 * is consumed exactly as the reference consumes it; in addressed mode eval (s, j) uses the draws at
 * offset ((s*total + first_index + j) * dpe) and the stream is left after q*total*dpe draws.
 * Li32: reference's sequential fp32 sum.  Li64: same fp32 contributions summed in fp64. */
uint64_t query(const vrlo_scene &s, const vrlb_vrlset_desc &set, const Ray &ray, const vrlb_options &opt,
               int rng_mode, Pcg32 &rng, V3 &Li32, double Li64[3]) {
    Li32 = V3(0.f); Li64[0] = Li64[1] = Li64[2] = 0.0;
    if (set.n == 0) return 0;
    const bool uniform = opt.use_uniform_sampling != 0, rr = opt.rr_vrl != 0, trace = opt.visibility != 0;
    const uint32_t q   = (uint32_t) std::max(opt.samples_per_query, 1);
    const uint64_t dpe = rr ? 3u : 2u;
    const uint64_t total = set.total ? set.total : set.n;
    const Pcg32 base = rng;
    uint64_t evals = 0;

    for (uint32_t i = 0; i < q; ++i) {
        Pcg32 cursor = base;
        if (rng_mode == VRLO_RNG_ADDRESSED)
            cursor.advance((int64_t) (((uint64_t) i * total + set.first_index) * dpe));
        for (uint64_t j = 0; j < set.n; ++j) {
            Vrl vrl = load_vrl(set, j);
            Pcg32 eval_rng = cursor;
            Pcg32 *r;
            if (rng_mode == VRLO_RNG_ADDRESSED) {
                eval_rng = cursor;
                for (uint64_t k = 0; k < dpe; ++k) cursor.next_u32();
                r = &eval_rng;
            } else {
                r = &rng;
            }
            /* grid-medium transmittance sub-stream (addressed mode only): keyed by the eval */
            Pcg32 sub; sub.seed(eval_rng.state, PCG32_DEFAULT_STREAM + 1);
            Pcg32 *subp = (rng_mode == VRLO_RNG_ADDRESSED) ? &sub : r;
            DrawSrc draws{ r, { 0.f, 0.f }, 0 };
            V3 contrib;
            if (!rr) {
                contrib = get_contrib(s, trace, vrl, ray, uniform, draws, subp) * set.scale;
            } else {
                /* EDistanceRoulette, vrl_map.h:171-190 */
                ClosestPointInfo res = find_closest_point(vrl, ray);
                float d2 = squared_norm((ray.o + ray.d * res.tCam) - (vrl.origin + vrl.direction * res.tVRL));
                float min_distance = std::sqrt(d2);
                float rrWeight     = std::min(1 / (min_distance * opt.scale_rr), 1.0f);
                if (rrWeight > r->next_f32())
                    contrib = (get_contrib(s, trace, vrl, ray, uniform, draws, subp) / rrWeight) * set.scale;
                else
                    continue;   /* pruned */
            }
            ++evals;
            if (std::isnan(contrib.x) || std::isnan(contrib.y) || std::isnan(contrib.z) ||
                std::isinf(contrib.x) || std::isinf(contrib.y) || std::isinf(contrib.z))
                continue;
            Li32 = Li32 + contrib;
            Li64[0] += (double) contrib.x; Li64[1] += (double) contrib.y; Li64[2] += (double) contrib.z;
        }
    }
    if (rng_mode == VRLO_RNG_ADDRESSED) {
        rng = base;
        rng.advance((int64_t) ((uint64_t) q * total * dpe));
    }
    if (q > 1) {
        Li32 = Li32 / (float) q;
        for (int k = 0; k < 3; ++k) Li64[k] /= (double) q;
    }
    return evals;
}

/* VRLMap::query for a whole multi-part NLRay, `use_nl_atomic_query` (vrl.cpp:613-617): uniform sampling over the total
 * length (VRL::sampleMC, vrl_struct.h:239-243, with NLRay::at walking the parts by sequential subtraction, :44-56),
 * distance roulette measured from the first part (vrl_map.h:141, 172-176).  Addressed draws as in query().  The
 * contributions are returned per part (the part `at` landed on), so that a caller can compare part-wise partial sums;
 * their sum over the parts is the reference's Li. */
uint64_t query_atomic(const vrlo_scene &s, const vrlb_vrlset_desc &set, const Ray *parts, int n_parts, const vrlb_options &opt,
                      int rng_mode, Pcg32 &rng, V3 *Li32, double (*Li64)[3]) {
    for (int k = 0; k < n_parts; ++k) { Li32[k] = V3(0.f); Li64[k][0] = Li64[k][1] = Li64[k][2] = 0.0; }
    if (set.n == 0 || n_parts == 0) return 0;
    const bool rr = opt.rr_vrl != 0, trace = opt.visibility != 0;
    const uint32_t q   = (uint32_t) std::max(opt.samples_per_query, 1);
    const uint64_t dpe = rr ? 3u : 2u;
    const uint64_t total = set.total ? set.total : set.n;
    float maxt = 0.f;
    for (int k = 0; k < n_parts; ++k) maxt += parts[k].maxt;                 /* NLRay::push_back, vrl_struct.h:23-26 */
    const Ray first = parts[0];
    const Pcg32 base = rng;
    uint64_t evals = 0;
    for (uint32_t i = 0; i < q; ++i) {
        Pcg32 cursor = base;
        if (rng_mode == VRLO_RNG_ADDRESSED) cursor.advance((int64_t) (((uint64_t) i * total + set.first_index) * dpe));
        for (uint64_t j = 0; j < set.n; ++j) {
            Vrl vrl = load_vrl(set, j);
            Pcg32 eval_rng = cursor;
            Pcg32 *r;
            if (rng_mode == VRLO_RNG_ADDRESSED) { for (uint64_t k = 0; k < dpe; ++k) cursor.next_u32(); r = &eval_rng; }
            else r = &rng;
            Pcg32 sub; sub.seed(eval_rng.state, PCG32_DEFAULT_STREAM + 1);
            Pcg32 *subp = (rng_mode == VRLO_RNG_ADDRESSED) ? &sub : r;
            float rrWeight = 1.f;
            if (rr) {
                ClosestPointInfo res = find_closest_point(vrl, first);
                float d2 = squared_norm((first.o + first.d * res.tCam) - (vrl.origin + vrl.direction * res.tVRL));
                rrWeight = std::min(1 / (std::sqrt(d2) * opt.scale_rr), 1.0f);
                if (!(rrWeight > r->next_f32())) continue;
            }
            /* sampleMC + NLRay::at */
            float u1 = r->next_f32(), u2 = r->next_f32();
            float tCam = u1 * maxt, tVRL = u2 * vrl.length;
            int hit = n_parts - 1; V3 pCam;
            {
                float dist = tCam; bool found = false;
                if (dist < maxt)
                    for (int k = 0; k < n_parts; ++k) {
                        if (dist < parts[k].maxt) { hit = k; pCam = parts[k].o + parts[k].d * dist; found = true; break; }
                        dist -= parts[k].maxt;
                    }
                if (!found) { hit = n_parts - 1; pCam = parts[hit].o + parts[hit].d * tCam; }
            }
            SamplingInfo si{ pCam, tCam, vrl.origin + vrl.direction * tVRL, tVRL, vrl.length * maxt, true };
            Ray wi_ray{ first.o, parts[hit].d, first.maxt };
            V3 contrib = get_contrib_sampled(s, trace, vrl, wi_ray, si, subp);
            if (rr) contrib = contrib / rrWeight;
            contrib = contrib * set.scale;
            ++evals;
            if (std::isnan(contrib.x) || std::isnan(contrib.y) || std::isnan(contrib.z) ||
                std::isinf(contrib.x) || std::isinf(contrib.y) || std::isinf(contrib.z))
                continue;
            Li32[hit] = Li32[hit] + contrib;
            Li64[hit][0] += (double) contrib.x; Li64[hit][1] += (double) contrib.y; Li64[hit][2] += (double) contrib.z;
        }
    }
    if (rng_mode == VRLO_RNG_ADDRESSED) { rng = base; rng.advance((int64_t) ((uint64_t) q * total * dpe)); }
    if (q > 1)
        for (int k = 0; k < n_parts; ++k) { Li32[k] = Li32[k] / (float) q; for (int c = 0; c < 3; ++c) Li64[k][c] /= (double) q; }
    return evals;
}

template <typename F> void parallel_for(uint64_t n, int n_threads, F &&fn) {
    if (n_threads <= 1 || n < 2) { for (uint64_t i = 0; i < n; ++i) fn(i); return; }
    std::atomic<uint64_t> next{ 0 };
    std::vector<std::thread> pool;
    for (int t = 0; t < n_threads; ++t)
        pool.emplace_back([&]() {
            for (;;) {
                uint64_t i = next.fetch_add(1);
                if (i >= n) break;
                fn(i);
            }
        });
    for (auto &th : pool) th.join();
}

}  // namespace

extern "C" {

void vrlo_find_closest_point(const float vrl_o[3], const float vrl_d[3], float vrl_len,
                             const float ray_o[3], const float ray_d[3], float ray_maxt, float *tCam, float *tVRL) {
    Vrl v{ V3(vrl_o), V3(vrl_d), vrl_len, V3(0.f), 0 };
    Ray r{ V3(ray_o), V3(ray_d), ray_maxt };
    ClosestPointInfo c = find_closest_point(v, r);
    *tCam = c.tCam; *tVRL = c.tVRL;
}

int vrlo_sampling_vrl(const float vrl_o[3], const float vrl_d[3], float vrl_len,
                      const float ray_o[3], const float ray_d[3], float ray_maxt,
                      int uniform, float u1, float u2, float *out9) {
    Vrl v{ V3(vrl_o), V3(vrl_d), vrl_len, V3(0.f), 0 };
    Ray r{ V3(ray_o), V3(ray_d), ray_maxt };
    DrawSrc draws{ nullptr, { u1, u2 }, 0 };
    SamplingInfo si = sampling_vrl(v, r, uniform != 0, draws);
    out9[0] = si.pCam.x; out9[1] = si.pCam.y; out9[2] = si.pCam.z; out9[3] = si.tCam;
    out9[4] = si.pVRL.x; out9[5] = si.pVRL.y; out9[6] = si.pVRL.z; out9[7] = si.tVRL; out9[8] = si.invPDF;
    return si.valid ? 1 : 0;
}

float vrlo_phase_eval(const vrlb_medium *m, const float wi[3], const float wo[3]) {
    return phase_eval(*m, V3(wi), V3(wo));
}

void vrlo_get_contrib(const vrlb_scene_desc *scene, const vrlb_medium *m, const float vrl10[10],
                      const float ray_o[3], const float ray_d[3], float ray_maxt,
                      int uniform, float u1, float u2, float out_rgb[3]) {
    vrlo_scene *s = nullptr;
    vrlo_scene local;
    if (scene) s = vrlo_scene_create(scene);
    else { local.media.push_back(*m); }
    const vrlo_scene &sc = s ? *s : local;
    Vrl v{ V3(vrl10), V3(vrl10 + 3), vrl10[6], V3(vrl10 + 7), 0 };
    Ray r{ V3(ray_o), V3(ray_d), ray_maxt };
    DrawSrc draws{ nullptr, { u1, u2 }, 0 };
    Pcg32 sub; sub.seed(0, PCG32_DEFAULT_STREAM + 1);
    V3 c = get_contrib(sc, s != nullptr, v, r, uniform != 0, draws, &sub);
    out_rgb[0] = c.x; out_rgb[1] = c.y; out_rgb[2] = c.z;
    if (s) vrlo_scene_destroy(s);
}

vrlo_scene *vrlo_scene_create(const vrlb_scene_desc *desc) {
    vrlo_scene *s = new vrlo_scene();
    s->tris.resize(desc->n_tris);
    for (uint32_t i = 0; i < desc->n_tris; ++i) {
        const float *p = desc->tri_verts + (size_t) i * 9;
        const float *r = desc->tri_reflectance ? desc->tri_reflectance + (size_t) i * 3 : nullptr;
        s->tris[i] = Tri{ V3(p), V3(p + 3), V3(p + 6), desc->tri_flags ? desc->tri_flags[i] : 0u, { r ? r[0] : 0.f, r ? r[1] : 0.f, r ? r[2] : 0.f } };
    }
    s->media.assign(desc->media, desc->media + desc->n_media);
    s->grids.resize(desc->n_media);
    for (uint32_t i = 0; i < desc->n_media; ++i) {
        vrlb_medium &m = s->media[i];
        if (m.kind == VRLB_MEDIUM_GRID && m.grid) {
            size_t n = (size_t) m.grid_res[0] * m.grid_res[1] * m.grid_res[2];
            s->grids[i].assign(m.grid, m.grid + n);
            m.grid = s->grids[i].data();
        }
    }
    if (desc->emitters) s->emitters.assign(desc->emitters, desc->emitters + desc->n_emitters);
    build_bvh(*s);
    return s;
}
void vrlo_scene_destroy(vrlo_scene *s) { delete s; }

int vrlo_ray_intersect(const vrlo_scene *s, const float o[3], const float d[3], float mint, float maxt,
                       float *t, float *u, float *v) {
    Hit h = ray_intersect(*s, V3(o), V3(d), mint, maxt);
    *t = h.t; *u = h.u; *v = h.v;
    return h.prim;
}

void vrlo_eval_transmittance(const vrlo_scene *s, const float p1[3], const float p2[3], int medium, float out_rgb[3]) {
    Pcg32 sub; sub.seed(0, PCG32_DEFAULT_STREAM + 1);
    float tr = eval_transmittance(*s, true, V3(p1), V3(p2), medium, &sub);
    out_rgb[0] = out_rgb[1] = out_rgb[2] = tr;
}

/* probe for the delta BSDFs of the specular chains (tests/test_oracle_cpu.py): returns 0 when the lobe is inactive */
int vrlo_sample_delta(uint32_t flags, const float prm[3], const float n[3], const float rd[3], float sample1, float wo[3], float weight[3], int *transmitted) {
    Tri tr{ V3(0.f), V3(0.f), V3(0.f), flags, { prm[0], prm[1], prm[2] } };
    V3 w, wt; bool tt = false;
    if (!sample_delta(tr, V3(n), V3(rd), sample1, w, wt, tt)) return 0;
    wo[0] = w.x; wo[1] = w.y; wo[2] = w.z; weight[0] = wt.x; weight[1] = wt.y; weight[2] = wt.z; *transmitted = tt ? 1 : 0;
    return 1;
}

/* probes for the grid medium (tests/test_oracle_cpu.py) */
float vrlo_grid_eval(const vrlo_scene *s, int medium, const float p[3]) { return grid_eval(s->media[(size_t) medium], V3(p)); }
float vrlo_grid_transmittance(const vrlo_scene *s, int medium, const float o[3], const float d[3], float maxt, uint64_t seed) {
    Pcg32 sub; sub.seed(seed, PCG32_DEFAULT_STREAM + 1);
    return grid_transmittance(*s, s->media[(size_t) medium], medium, V3(o), V3(d), maxt, &sub);
}

/* The direct in-scattering term of VRLIntegrator::sample for one in-medium camera segment (vrl.cpp:523-611): radius jitter
 * (the draw before the query position), gather points at r, 3r, 5r ... while gather_t <= maxt - step, grey transmittance
 * from the previous gather point (homoEvalTransmittance, medium.h:103-107), PhotonMap::estimateRadianceVolume with the
 * RadianceQueryVolume functor (photonmap.h:216-271) -- here a plain loop over the photon set, the reference's kd-tree
 * only changes the visiting order -- and the map's scale factor (vrl.cpp:610). */
/* PhotonMap::estimateRadianceVolume with the RadianceQueryVolume functor (photonmap.h:216-271): a plain loop over the
 * photon set, the reference's kd-tree range query only changes the visiting order */
static V3 estimate_radiance_volume(const vrlb_medium &m, const vrlb_photonset_desc &set, V3 gp, V3 wo, float radius) {
    V3 result(0.f);
    const float radiusSqr = radius * radius;
    for (uint64_t q = 0; q < set.n; ++q) {
        const V3 pp(set.pos + q * 3), pd(set.dir + q * 3), pw(set.power + q * 3);
        const float lengthSqr = squared_norm(gp - pp);
        if ((radiusSqr - lengthSqr) < 0) continue;
        const float photonPF = phase_eval(m, -pd, wo);
        const float q1 = 1.0f - lengthSqr / radiusSqr;
        result = result + pw * photonPF * (q1 * q1) / radiusSqr * 0.31830988618379067154f * 3.0f;
    }
    return result;
}
/* probe (tests/test_ref_golden_cpu.py): one estimateRadianceVolume call */
void vrlo_estimate_radiance_volume(const vrlb_medium *m, const vrlb_photonset_desc *set, const float gp[3], const float wo[3], float radius, float out[3]) {
    const V3 r = estimate_radiance_volume(*m, *set, V3(gp), V3(wo), radius);
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}

void vrlo_photon_gather_segments(const vrlo_scene *s, const vrlb_photonset_desc *set, const vrlb_segment *segs, uint64_t n_segs,
                                 const vrlb_options *opt, float *out_rgb) {
    (void) opt;
    for (uint64_t i = 0; i < n_segs; ++i) {
        const vrlb_segment &sg = segs[i];
        float *out = out_rgb + i * 3;
        out[0] = out[1] = out[2] = 0.f;
        const vrlb_medium &m = s->media[(size_t) sg.medium];
        if ((sg.part & 0xffffu) != 0 || set->n == 0) continue;
        /* a curved path = n_parts consecutive records (n_parts in the high half of `part`, or of `sample` for the atomic
         * query); the walk below is vrl.cpp:553-608 with the records standing for the NLRay's parts */
        uint64_t n_parts = std::max<uint32_t>(std::max(sg.part >> 16, sg.sample >> 16), 1u);
        n_parts = std::min(n_parts, n_segs - i);
        Pcg32 rng; rng.state = sg.rng_state; rng.inc = (PCG32_DEFAULT_STREAM << 1) | 1u;
        rng.advance(-1);
        const float radius = set->lookup_radius * lerp(0.75f, 1.25f, rng.next_f32());
        float stepSize = radius, leftOver = 0.0f;
        const V3 d0(sg.d);                                   /* mediumRay.d: the direction the path starts with */
        V3 last(sg.o), tr(sg.weight), direct(0.f);
        const float md = max_density(m);
        for (uint64_t k = 0; k < n_parts; ++k) {
            const V3 o(segs[i + k].o), d(segs[i + k].d);
            const float maxt = segs[i + k].maxt;
            float gather_t = 0.f;
            while (gather_t <= (maxt - (stepSize - leftOver))) {
                gather_t += (stepSize - leftOver);
                leftOver = 0.0f;
                const V3 gp = o + d * gather_t;
                stepSize = radius * 2;
                const float val = std::exp(md * -norm(gp - last));
                tr = tr * (md != 0.f ? val : 1.0f);
                const V3 result = estimate_radiance_volume(m, *set, gp, d0, radius);
                direct = direct + tr * result;
                last = gp;
            }
            if (k + 1 < n_parts) leftOver += maxt - gather_t;    /* vrl.cpp:570 */
        }
        direct = direct * set->scale;
        out[0] = direct.x; out[1] = direct.y; out[2] = direct.z;
    }
}

/* The surface radiance of VRLIntegrator::sample at the first smooth surface of the camera path (vrl.cpp:671-680):
 *   estimateCausticRadiance(si, r_c, k_c) + estimateRadiance(si, r_g, k_g)           (photonmap.h:131-211)
 * nnSearch (kdtree.h:319-400) = the k nearest photons with d^2 < R^2; when more than k lie inside, the search radius
 * shrinks to the k-th nearest distance, and THAT radius is what invSquaredRadius is computed from afterwards.  Here: all
 * in-range distances, sorted (the kd-tree's visiting order only permutes the sum).  bsdf->eval is SmoothDiffuse::eval
 * (diffuse.cpp): reflectance * INV_PI * cos_theta_o where cos_theta_i > 0 and cos_theta_o > 0, si.wi the camera side,
 * wo = to_local(-photon.direction); `twosided` flips the frame to the side si.wi lies on.  Either set may be NULL / empty. */
static V3 surface_estimate(const vrlb_photonset_desc *set, int32_t k, bool caustic, const vrlb_surface_query &q) {
    V3 result(0.f);
    if (!set || set->n == 0 || k < 1) return result;
    const V3 p(q.p);
    float squaredRadius = set->lookup_radius * set->lookup_radius;
    std::vector<std::pair<float, uint64_t>> found;
    for (uint64_t i = 0; i < set->n; ++i) {
        const float d2 = squared_norm(V3(set->pos + i * 3) - p);
        if (d2 < squaredRadius) found.push_back({ d2, i });
    }
    if (found.size() > (size_t) k) {
        std::sort(found.begin(), found.end());
        found.resize((size_t) k);
        squaredRadius = found.back().first;
    }
    std::sort(found.begin(), found.end(), [](auto &a, auto &b) { return a.second < b.second; });
    const float invSquaredRadius = 1.0f / squaredRadius;
    V3 n(q.n);
    float cos_i = dot(V3(q.wi), n);
    if (cos_i < 0.f && (q.flags & VRLB_SURF_TWOSIDED)) { n = -n; cos_i = -cos_i; }
    if (!(cos_i > 0.f)) return result;
    const V3 reflectance(q.reflectance);
    for (auto &f : found) {
        const V3 pd(set->dir + f.second * 3), pw(set->power + f.second * 3);
        const float cos_o = -dot(pd, n);
        if (!(cos_o > 0.f)) continue;
        float bs = 0.31830988618379067154f * cos_o;
        if (caustic) bs = bs * std::max(0.f, 1.0f - std::sqrt(f.first * invSquaredRadius));
        result = result + pw * (reflectance * bs);
    }
    return result * ((caustic ? 3.0f : 1.0f) * set->scale * 0.31830988618379067154f * invSquaredRadius);
}
void vrlo_surface_gather_queries(const vrlb_photonset_desc *global, int32_t k_global, const vrlb_photonset_desc *caustic, int32_t k_caustic,
                                 const vrlb_surface_query *queries, uint64_t n, float *out_rgb) {
    for (uint64_t i = 0; i < n; ++i) {
        const V3 c = surface_estimate(caustic, k_caustic, true, queries[i]), g = surface_estimate(global, k_global, false, queries[i]);
        const V3 r = c + g;
        out_rgb[i * 3] = r.x; out_rgb[i * 3 + 1] = r.y; out_rgb[i * 3 + 2] = r.z;
    }
}

/* VRLMap::toLong -> VRL::convertToLong (vrl_map.h:56-62, vrl_struct.h:146-154): Ray3f r(origin, direction, 0) keeps the
 * default [RayEpsilon, inf) interval; length = si.t of the closest surface; a miss is Log(Error) = throw in the reference */
uint64_t vrlo_vrls_to_long(const vrlo_scene *s, const float *soa10, uint64_t n, float *length_out) {
    uint64_t missed = 0;
    for (uint64_t j = 0; j < n; ++j) {
        V3 o(soa10[0 * n + j], soa10[1 * n + j], soa10[2 * n + j]), d(soa10[3 * n + j], soa10[4 * n + j], soa10[5 * n + j]);
        Hit h = ray_intersect(*s, o, d, kRayEpsilon, kInf);
        if (!h.valid()) { ++missed; length_out[j] = soa10[6 * n + j]; }
        else length_out[j] = h.t;
    }
    return missed;
}

/* VRLMap::dicingVRL -> VRL::dice (vrl_map.h:104-128, vrl_struct.h:88-128).  `avg_length` and `current_distance` are
 * doubles in the reference (`auto x = 0.0`), the dice length is passed as a Float; origin + direction * current_distance
 * is a float vector times a double, i.e. evaluated in fp64 and rounded once on assignment.  Homogeneous media only
 * (grid media would draw from the preprocess sampler).  Returns the number of diced VRLs; writes min(count, cap). */
uint64_t vrlo_vrls_dice(const vrlo_scene *s, const float *soa10, const uint16_t *medium, uint64_t n, uint32_t avg_number_dice,
                        float *out_soa10, uint16_t *out_medium, uint64_t cap) {
    double avg_length = 0.0;
    for (uint64_t j = 0; j < n; ++j) avg_length += soa10[6 * n + j];
    avg_length /= (double) n;
    const float avgLength = (float) (avg_length / (double) avg_number_dice);
    uint64_t w = 0;
    for (uint64_t j = 0; j < n; ++j) {
        const float length = soa10[6 * n + j];
        const int med = medium ? medium[j] : 0;
        const vrlb_medium &m = s->media[(size_t) med];
        double current_distance = 0.0;
        while (current_distance <= length) {
            double next_length = std::min((double) length - current_distance, (double) avgLength);
            float tr = 1.0f;
            if (current_distance != 0.0) tr = homo_transmittance(m, 0.0f, (float) current_distance);
            if (w < cap) {
                for (int c = 0; c < 3; ++c) {
                    out_soa10[(size_t) c * cap + w] = (float) ((double) soa10[c * n + j] + (double) soa10[(3 + c) * n + j] * current_distance);
                    out_soa10[(size_t) (3 + c) * cap + w] = soa10[(3 + c) * n + j];
                    out_soa10[(size_t) (7 + c) * cap + w] = soa10[(7 + c) * n + j] * tr;
                }
                out_soa10[(size_t) 6 * cap + w] = (float) next_length;
                if (out_medium) out_medium[w] = (uint16_t) med;
            }
            ++w;
            current_distance += avgLength;
        }
    }
    return w;
}

uint64_t vrlo_gather_segments(const vrlo_scene *s, const vrlb_vrlset_desc *vrls, const vrlb_segment *segs,
                              uint64_t n_segs, const vrlb_options *opt, int rng_mode, int n_threads,
                              float *out_rgb32, double *out_rgb64) {
    std::atomic<uint64_t> evals{ 0 };
    parallel_for(n_segs, n_threads, [&](uint64_t i) {
        const vrlb_segment &sg = segs[i];
        const uint32_t n_all = sg.sample >> 16;
        if (n_all > 0) {
            /* atomic query (use_nl_atomic_query): the record with part 0 evaluates the whole group */
            if (sg.part != 0) return;
            std::vector<Ray> parts(n_all);
            for (uint32_t k = 0; k < n_all; ++k) parts[k] = Ray{ V3(segs[i + k].o), V3(segs[i + k].d), segs[i + k].maxt };
            std::vector<V3> l32(n_all); std::vector<std::array<double, 3>> l64(n_all);
            Pcg32 rng{ sg.rng_state, (PCG32_DEFAULT_STREAM << 1) | 1u };
            evals += query_atomic(*s, *vrls, parts.data(), (int) n_all, *opt, rng_mode, rng, l32.data(), reinterpret_cast<double (*)[3]>(l64.data()));
            for (uint32_t k = 0; k < n_all; ++k) {
                if (out_rgb32) { out_rgb32[(i + k) * 3 + 0] = l32[k].x; out_rgb32[(i + k) * 3 + 1] = l32[k].y; out_rgb32[(i + k) * 3 + 2] = l32[k].z; }
                if (out_rgb64) { out_rgb64[(i + k) * 3 + 0] = l64[k][0]; out_rgb64[(i + k) * 3 + 1] = l64[k][1]; out_rgb64[(i + k) * 3 + 2] = l64[k][2]; }
            }
            return;
        }
        Ray ray{ V3(sg.o), V3(sg.d), sg.maxt };
        Pcg32 rng{ sg.rng_state, (PCG32_DEFAULT_STREAM << 1) | 1u };
        V3 li; double li64[3];
        evals += query(*s, *vrls, ray, *opt, rng_mode, rng, li, li64);
        if (out_rgb32) { out_rgb32[i * 3 + 0] = li.x; out_rgb32[i * 3 + 1] = li.y; out_rgb32[i * 3 + 2] = li.z; }
        if (out_rgb64) { out_rgb64[i * 3 + 0] = li64[0]; out_rgb64[i * 3 + 1] = li64[1]; out_rgb64[i * 3 + 2] = li64[2]; }
    });
    return evals.load();
}

}  // extern "C"

namespace {

/* ------------------------------------------------------------------------------------------ */
/*  Grid ("heterogeneous") medium                                                             */
/* ------------------------------------------------------------------------------------------ */
/* BoundingBox::ray_intersect, include/mitsuba/core/bbox.h:302-325 */
inline bool bbox_ray_intersect(const float lo[3], const float hi[3], V3 o, V3 d, float &mint, float &maxt) {
    bool active = true;
    float t1p[3], t2p[3];
    for (int k = 0; k < 3; ++k) {
        active = active && ((d[k] != 0.f) || ((o[k] > lo[k]) || (o[k] < hi[k])));
        float rcp = 1.0f / d[k];
        float t1 = (lo[k] - o[k]) * rcp, t2 = (hi[k] - o[k]) * rcp;
        t1p[k] = std::fmin(t1, t2); t2p[k] = std::fmax(t1, t2);
    }
    mint = std::max(t1p[0], std::max(t1p[1], t1p[2]));
    maxt = std::min(t2p[0], std::min(t2p[1], t2p[2]));
    return active && maxt >= mint;
}

/* Medium::evalTransmittance, non-homogeneous branch, include/mitsuba/render/medium.h:119-154,
 * with Medium::sample_interaction (src/librender/medium.cpp:36-75) and eval_tr_and_pdf (77-89)
 * inlined.  Stochastic product of majorant transmittances over accepted free-flight steps (a8).
 * Consumes 1 + (#steps + 1) draws from `sub`. */
float grid_transmittance(const vrlo_scene &, const vrlb_medium &m, int, V3 o_, V3 d, float maxt_, Pcg32 *sub) {
    float t = 0.f, maxt = maxt_;
    float majorant = max_density(m);
    (void) std::min(sub->next_f32() * 3.f, 2.f);   /* channel draw; the majorant is grey */
    float si_t = maxt_;
    float throughput = 1.0f;
    V3 ro = o_; float rmint = 0.f, rmaxt = maxt_;
    while (t < maxt) {
        float sample = sub->next_f32();
        float bmin, bmax;
        bool active = bbox_ray_intersect(m.grid_bbox_min, m.grid_bbox_max, ro, d, bmin, bmax);
        active = active && (std::isfinite(bmin) || std::isfinite(bmax));
        if (!active) { bmin = 0.f; bmax = kInf; }
        float mint = std::max(rmint, bmin), mx = std::min(rmaxt, bmax);
        float sampled_t = mint + (-std::log(1 - sample) / majorant);
        bool valid_mi   = active && (sampled_t <= mx);
        float mi_t      = valid_mi ? sampled_t : kInf;
        V3 mi_p         = ro + d * sampled_t;
        if (!(mi_t < kInf)) break;                       /* !mi.is_valid() */
        if (si_t < mi_t) mi_t = kInf;
        float tt = std::min(mi_t, si_t) - mint;
        throughput *= std::exp(-tt * majorant);
        t += mi_t;
        si_t -= mi_t;
        ro    = mi_p;                                    /* mi.spawn_ray(_ray.d), interaction.h:58-61 */
        rmint = (1.f + hmax_abs(mi_p)) * kRayEpsilon;
        rmaxt = std::max(0.0f, maxt - t);
    }
    return throughput;
}

/* ------------------------------------------------------------------------------------------ */
/*  Sensor: PerspectiveCamera::sample_ray_differential, src/sensors/perspective.cpp:183-216   */
/*  Transform * Point / * Vector: include/mitsuba/core/transform.h:105-127 (column accumulate) */
/* ------------------------------------------------------------------------------------------ */
inline void camera_ray(const vrlb_camera &c, float sx, float sy, V3 &o, V3 &d, float &mint, float &maxt) {
    const float *m = c.sample_to_camera;   /* row-major: m[r*4+c] */
    float r[4];
    for (int k = 0; k < 4; ++k) r[k] = m[k * 4 + 3];
    for (int k = 0; k < 4; ++k) r[k] = m[k * 4 + 0] * sx + r[k];
    for (int k = 0; k < 4; ++k) r[k] = m[k * 4 + 1] * sy + r[k];
    for (int k = 0; k < 4; ++k) r[k] = m[k * 4 + 2] * 0.f + r[k];
    V3 near_p(r[0] / r[3], r[1] / r[3], r[2] / r[3]);
    V3 dl = normalize(near_p);
    float inv_z = 1.0f / dl.z;
    mint = c.near_clip * inv_z;
    maxt = c.far_clip * inv_z;
    const float *w = c.to_world;
    o = V3(w[3], w[7], w[11]);
    float q[3];
    for (int k = 0; k < 3; ++k) q[k] = w[k * 4 + 0] * dl.x;
    for (int k = 0; k < 3; ++k) q[k] = w[k * 4 + 1] * dl.y + q[k];
    for (int k = 0; k < 3; ++k) q[k] = w[k * 4 + 2] * dl.z + q[k];
    d = V3(q[0], q[1], q[2]);
}

/* graded-IOR profile of the `nonlinear` medium (nonlinear.cpp:129-153, method 0) and one eikonal step */
constexpr float kNlPartCos = 0.99999048f;     /* cos(0.25 deg) */
inline float nl_step(const vrlb_medium &m) {
    if (m.kind != VRLB_MEDIUM_NONLINEAR || m.nl_res[1] < 1) return 0.f;
    return (m.nl_bbox_max[1] - m.nl_bbox_min[1]) / (float) m.nl_res[1];
}
/* index and vertical gradient at p; from_bottom = false (mirage.xml): norm = max(0, y / (height / 2)) with the WORLD y,
 * nonlinear.cpp:149-150 */
inline void nl_profile(const vrlb_medium &m, V3 p, float &n, float &g) {
    float h = m.nl_bbox_max[1] - m.nl_bbox_min[1];
    bool in = p.x >= m.nl_bbox_min[0] && p.x <= m.nl_bbox_max[0] && p.y >= m.nl_bbox_min[1] && p.y <= m.nl_bbox_max[1] &&
              p.z >= m.nl_bbox_min[2] && p.z <= m.nl_bbox_max[2];
    if (m.from_bottom) {
        float t = std::fmin(std::fmax((p.y - m.nl_bbox_min[1]) / h, 0.f), 1.f);
        n = m.bottom_ior + (m.top_ior - m.bottom_ior) * t;
        g = in ? (m.top_ior - m.bottom_ior) / h : 0.f;
    } else {
        float hh = h * 0.5f;
        float t = std::fmax(0.f, p.y / hh);
        n = m.bottom_ior + (m.top_ior - m.bottom_ior) * t;
        g = (in && p.y > 0.f) ? (m.top_ior - m.bottom_ior) / hh : 0.f;
    }
}
inline V3 nl_bend(const vrlb_medium &m, V3 p, V3 dir, float ds) {
    float n, g;
    nl_profile(m, p, n, g);
    V3 v = dir * n;
    v.y = v.y + ds * g;
    return normalize(v);
}

/* What one camera sample hands to the VRL query */
struct PartOut { Ray ray; V3 weight; int medium; uint32_t part; const Ray *all = nullptr; int n_all = 0; };   /* n_all > 0: one query over the whole NLRay (use_nl_atomic_query) */

struct SampleResult { V3 radiance; bool valid_ray; bool has_sq = false; vrlb_surface_query sq{}; };

/* test hook: where vrlo_build_surface_queries collects the surface queries of the paths render_pixel traces */
thread_local std::vector<vrlb_surface_query> *g_surface_queries = nullptr;

/* VRLIntegrator::sample, src/integrators/vrl/vrl.cpp:443-732, for the surface classes of the
 * ABI (null / smooth / emitter).  Photon-map terms (vrl.cpp:600-611, 671-680) are SURVEY 8f rank 3
 * and contribute 0 here.  `on_query(part, rng)` is called once per NLRay part in reference order
 * (vrl.cpp:613-629) with the sampler positioned at the query's first draw; it returns Li and must
 * leave the sampler after the query. */
template <typename OnQuery>
SampleResult trace_sample(const vrlo_scene &s, const vrlb_options &opt, Sampler &sampler, V3 ro, V3 rd,
                          float rmint, float rmaxt, int sensor_medium, OnQuery &&on_query) {
    V3 radiance(0.f), throughput(1.f);
    bool has_sq = false; vrlb_surface_query sq{};
    int medium = sensor_medium;
    Hit si{ kInf, 0, 0, -1 };
    bool valid_ray = false;                         /* no environment emitter in the ABI */
    bool specular_chain = !opt.hide_emitters;
    bool needs_intersection = true;
    uint32_t depth = 1;
    /* channel draw, vrl.cpp:476-479 */
    (void) std::min(sampler.next_1d() * 3.f, 2.f);

    for (int bounce = 0;; ++bounce) {
        bool active = !(throughput.x == 0.f && throughput.y == 0.f && throughput.z == 0.f);
        if (!active || depth >= (uint32_t) opt.max_depth) break;
        bool active_medium = medium >= 0, active_surface = !active_medium;
        if (bounce > 100000) break;

        if (active_medium) {
            if (!si.valid()) break;    /* reference spins to the 100k-bounce guard doing nothing (vrl.cpp:500-510,523) */
            valid_ray = true;
            /* radius jitter draw, vrl.cpp:527 (value only feeds the photon gather) */
            (void) sampler.next_1d();
            const vrlb_medium &m = s.media[(size_t) medium];
            /* linear camera ray: a single part [ray.o, si.t] (vrl.cpp:539-540, 595).  Non-linear
             * splitting (vrl.cpp:550-578) is added by trace_nl_parts() when the medium asks for it. */
            V3 indirect(0.f);
            V3 vrlThroughput = throughput;
            if (m.kind == VRLB_MEDIUM_NONLINEAR && nl_step(m) > 0.f && opt.use_non_linear_camera) {
                /* curved camera path (vrl.cpp:550-629).  The north star replaces the reference's cell-face refraction
                 * by a fixed-step eikonal marcher (step = one IOR layer); parts are cut when the direction has turned
                 * by more than kNlPartCos; one query per part with the preceding parts' transmittance (vrl.cpp:619-628) */
                V3 pstart = ro, pos = ro, dir = rd, dir0 = rd;
                float total = 0.f, step_mint = rmint, ds = nl_step(m); uint32_t part = 0; bool found = false;
                Hit hit{ kInf, 0, 0, -1 };
                const bool atomic = opt.use_nl_atomic_query != 0;      /* vrl.cpp:613-617: one query over all parts */
                std::vector<Ray> all;
                for (int step = 0; step < 100000; ++step) {
                    hit = ray_intersect(s, pos, dir, step_mint, ds);
                    step_mint = 0.f;
                    if (hit.valid()) { pos = pos + dir * hit.t; found = true; break; }
                    pos = pos + dir * ds;
                    V3 nd = nl_bend(m, pos, dir, ds);
                    if (dot(nd, dir0) < kNlPartCos) {
                        V3 chord = pos - pstart; float len = norm(chord);
                        if (atomic) all.push_back(Ray{ pstart, chord / len, len });
                        else {
                            V3 Li = on_query(PartOut{ Ray{ pstart, chord / len, len }, vrlThroughput, medium, part++ }, sampler);
                            indirect = indirect + Li * vrlThroughput;
                        }
                        vrlThroughput = vrlThroughput * homo_transmittance(m, 0.f, len);
                        total = total + len;
                        pstart = pos; dir0 = nd;
                    }
                    dir = nd;
                }
                if (!found) { si = Hit{ kInf, 0, 0, -1 }; radiance = radiance + indirect; break; }
                V3 chord = pos - pstart; float len = norm(chord);
                if (len > 0.f) {
                    if (atomic) all.push_back(Ray{ pstart, chord / len, len });
                    else {
                        V3 Li = on_query(PartOut{ Ray{ pstart, chord / len, len }, vrlThroughput, medium, part++ }, sampler);
                        indirect = indirect + Li * vrlThroughput;
                    }
                    total = total + len;
                }
                if (atomic && !all.empty()) {
                    V3 Li = on_query(PartOut{ all[0], throughput, medium, 0u, all.data(), (int) all.size() }, sampler);
                    indirect = indirect + Li * throughput;                                /* vrl.cpp:617 */
                }
                throughput = throughput * homo_transmittance(m, 0.f, total);             /* vrl.cpp:631 */
                si = hit; rd = dir; needs_intersection = false;
            } else {
                /* linear camera ray: a single part [ray.o, si.t] (vrl.cpp:539-540, 595) */
                Ray part{ ro, rd, si.t };
                V3 Li = on_query(PartOut{ part, vrlThroughput, medium, 0u }, sampler);
                indirect = indirect + Li * vrlThroughput;
                throughput = throughput * homo_transmittance(m, 0.f, part.maxt);         /* vrl.cpp:631 */
            }
            radiance = radiance + indirect;
            active_surface = true;
        }

        /* surface interactions, vrl.cpp:645-717 */
        if (active_surface && needs_intersection)
            si = ray_intersect(s, ro, rd, rmint, rmaxt);
        if (active_surface && si.valid() && specular_chain) {
            const Tri &tr = s.tris[(size_t) si.prim];
            if (tr.flags & VRLB_SURF_EMITTER) {
                /* AreaLight::eval: radiance where cos_theta(si.wi) > 0 (area.cpp:64-72) */
                if (dot(-rd, tri_normal(tr)) > 0.f) {
                    int e = VRLB_SURF_GET_EMITTER(tr.flags);
                    V3 Le = (size_t) e < s.emitters.size() ? V3(s.emitters[(size_t) e].radiance) : V3(0.f);
                    radiance = radiance + throughput * Le;
                }
            }
        }
        active_surface = active_surface && si.valid();
        if (active_surface) {
            const Tri &tr = s.tris[(size_t) si.prim];
            if (tr.flags & VRLB_SURF_EMITTER) break;            /* vrl.cpp:664-665 */
            valid_ray = true;
            if ((tr.flags & VRLB_SURF_BSDF_MASK) == VRLB_SURF_BSDF_SMOOTH) {        /* smooth BSDF: photon-map estimates, then break (vrl.cpp:671-680) */
                if (tr.prm[0] != 0.f || tr.prm[1] != 0.f || tr.prm[2] != 0.f) {     /* what the estimates read of si and the path */
                    const V3 p = tri_point(tr, si.u, si.v), n = tri_normal(tr);
                    sq.p[0] = p.x; sq.p[1] = p.y; sq.p[2] = p.z; sq.n[0] = n.x; sq.n[1] = n.y; sq.n[2] = n.z;
                    sq.flags = tr.flags & VRLB_SURF_TWOSIDED;
                    sq.wi[0] = -rd.x; sq.wi[1] = -rd.y; sq.wi[2] = -rd.z;
                    for (int k = 0; k < 3; ++k) sq.reflectance[k] = tr.prm[k];
                    sq.weight[0] = throughput.x; sq.weight[1] = throughput.y; sq.weight[2] = throughput.z;
                    has_sq = true;
                }
                break;
            }
            /* bsdf->sample consumes next_1d + next_2d (vrl.cpp:682).  Null::sample (null.cpp:38-52): weight 1, wo = -wi */
            const float sample1 = sampler.next_1d(); (void) sampler.next_1d(); (void) sampler.next_1d();
            if (!is_null(tr)) {
                /* specular chain, vrl.cpp:682-705 */
                V3 wo, weight; bool transmitted;
                if (!sample_delta(tr, tri_normal(tr), rd, sample1, wo, weight, transmitted)) break;
                throughput = throughput * weight;
                if (throughput.x == 0.f && throughput.y == 0.f && throughput.z == 0.f) break;
                rd = wo;
                depth += 1;
                specular_chain = true;
            }
            V3 p  = tri_point(tr, si.u, si.v);
            ro    = p;                                          /* si.spawn_ray, interaction.h:58-61 */
            rmint = (1.f + hmax_abs(p)) * kRayEpsilon;
            rmaxt = kInf;
            Hit si_new = ray_intersect(s, ro, rd, rmint, rmaxt);  /* vrl.cpp:707-711 */
            needs_intersection = false;
            if (is_medium_transition(tr)) medium = target_medium(tr, rd);   /* vrl.cpp:713-714 */
            si = si_new;
        }
        if (!(active_surface || active_medium)) break;          /* vrl.cpp:719 */
    }
    SampleResult res{ radiance, valid_ray };
    res.has_sq = has_sq; res.sq = sq;
    return res;
}

/* srgb_to_xyz, include/mitsuba/core/spectrum.h:221-227 */
inline V3 srgb_to_xyz(V3 c) {
    return V3(0.412453f * c.x + 0.357580f * c.y + 0.180423f * c.z,
              0.212671f * c.x + 0.715160f * c.y + 0.072169f * c.z,
              0.019334f * c.x + 0.119193f * c.y + 0.950227f * c.z);
}

/* SamplingIntegrator::render_block + render_sample (integrator.cpp:187-297) for one pixel:
 * seeds the sampler, runs spp samples, calls put(px, py, rgb, alpha) per sample. */
template <typename OnQuery, typename Put>
void render_pixel(const vrlo_scene &s, const vrlb_camera &cam, const vrlb_options &opt, uint32_t block_id,
                  uint32_t morton_i, int bx0, int by0, int bw, int bh, OnQuery &&on_query, Put &&put) {
    uint32_t lx = compact_bits(morton_i), ly = compact_bits(morton_i >> 1);
    if ((int) lx >= bw || (int) ly >= bh) return;
    Sampler sampler;
    uint32_t pixel_count = (uint32_t) (opt.block_size * opt.block_size);
    sampler.seed(opt.base_seed, (uint64_t) block_id * pixel_count + morton_i);   /* integrator.cpp:204 */
    int px = bx0 + (int) lx, py = by0 + (int) ly;
    for (int j = 0; j < opt.spp; ++j) {
        float ax = 0.5f, ay = 0.5f;
        if (opt.spp > 1) { ax = sampler.next_1d(); ay = sampler.next_1d(); }    /* integrator.cpp:249 */
        float posx = (float) px + ax, posy = (float) py + ay;
        (void) sampler.next_1d();                                                /* wavelength sample, :261 */
        float sx = posx / (float) cam.width, sy = posy / (float) cam.height;     /* :263-265, no crop */
        V3 o, d; float mint, maxt;
        camera_ray(cam, sx, sy, o, d, mint, maxt);
        auto q = [&](const PartOut &p, Sampler &smp) { return on_query(p, smp, px, py, j); };
        SampleResult r = trace_sample(s, opt, sampler, o, d, mint, maxt, cam.medium, q);
        /* ImageBlock::put, box filter branch (imageblock.cpp:162-169): pixel = ceil(pos - 1) inside the block */
        int fx = (int) std::ceil(posx - (float) bx0 - 1.0f), fy = (int) std::ceil(posy - (float) by0 - 1.0f);
        if (fx >= 0 && fy >= 0 && fx < bw && fy < bh) {
            put(bx0 + fx, by0 + fy, r.radiance, r.valid_ray ? 1.f : 0.f);
            if (r.has_sq && g_surface_queries) { r.sq.pixel = (uint32_t) ((by0 + fy) * cam.width + bx0 + fx); g_surface_queries->push_back(r.sq); }
        }
    }
}

}  // namespace

extern "C" {

/* graded-IOR probes: the profile (nonlinear.cpp:129-153) and the fixed-step eikonal march of a ray from o along d through
 * medium `medium` of the scene until the next surface (the marcher of trace_sample / the light tracer, without the part
 * cutting): out8 = exit point, exit direction, path length, steps.  Returns 1 if a surface was reached. */
float vrlo_nl_ior(const vrlb_medium *m, const float p[3]) {
    float n, g;
    nl_profile(*m, V3(p), n, g);
    return n;
}
int vrlo_nl_march(const vrlo_scene *s, int medium, const float o[3], const float d[3], float mint, float *out8) {
    const vrlb_medium &m = s->media[(size_t) medium];
    V3 pos(o), dir(d);
    float step_mint = mint, ds = nl_step(m), len = 0.f;
    int steps = 0; bool found = false;
    if (!(ds > 0.f)) return 0;
    for (; steps < 1000000; ++steps) {
        Hit hit = ray_intersect(*s, pos, dir, step_mint, ds);
        step_mint = 0.f;
        if (hit.valid()) { pos = pos + dir * hit.t; len += hit.t; found = true; break; }
        pos = pos + dir * ds; len += ds;
        dir = nl_bend(m, pos, dir, ds);
    }
    out8[0] = pos.x; out8[1] = pos.y; out8[2] = pos.z; out8[3] = dir.x; out8[4] = dir.y; out8[5] = dir.z; out8[6] = len; out8[7] = (float) steps;
    return found ? 1 : 0;
}

/* test hook: restrict vrlo_build_segments to the pixels x0 <= x < x1, y0 <= y < y1 (curved paths at 720p are minutes of
 * single-threaded marching for the whole film); x1 <= x0 resets to the whole film */
static int g_build_crop[4] = { 0, 0, 1 << 30, 1 << 30 };
void vrlo_set_build_crop(int x0, int y0, int x1, int y1) {
    if (x1 <= x0 || y1 <= y0) { g_build_crop[0] = g_build_crop[1] = 0; g_build_crop[2] = g_build_crop[3] = 1 << 30; }
    else { g_build_crop[0] = x0; g_build_crop[1] = y0; g_build_crop[2] = x1; g_build_crop[3] = y1; }
}
uint64_t vrlo_build_segments(const vrlo_scene *s, const vrlb_camera *cam, const vrlb_options *opt,
                             uint64_t n_vrls_total, int rng_mode, vrlb_segment *segs, uint64_t cap,
                             float *direct_rgb, float *alpha_w) {
    (void) rng_mode;   /* stream positions after the first query are only defined in addressed mode */
    int nbx, nby;
    std::vector<uint32_t> ids = spiral_ids(cam->width, cam->height, opt->block_size, nbx, nby);
    const uint64_t dpe = opt->rr_vrl ? 3u : 2u;
    const uint32_t q = (uint32_t) std::max(opt->samples_per_query, 1);
    uint64_t n = 0;
    if (direct_rgb) std::memset(direct_rgb, 0, sizeof(float) * 3 * (size_t) cam->width * cam->height);
    if (alpha_w) std::memset(alpha_w, 0, sizeof(float) * 2 * (size_t) cam->width * cam->height);
    /* emit in film order of blocks (row-major over blocks, Morton inside): the GPU builder uses the same order */
    for (int by = 0; by < nby; ++by)
        for (int bx = 0; bx < nbx; ++bx) {
            int bx0 = bx * opt->block_size, by0 = by * opt->block_size;
            int bw = std::min(opt->block_size, cam->width - bx0), bh = std::min(opt->block_size, cam->height - by0);
            uint32_t bid = ids[(size_t) by * nbx + bx];
            if (bx0 >= g_build_crop[2] || by0 >= g_build_crop[3] || bx0 + bw <= g_build_crop[0] || by0 + bh <= g_build_crop[1]) continue;
            for (uint32_t i = 0; i < (uint32_t) (opt->block_size * opt->block_size); ++i) {
                {   /* test hook: only the pixels of the crop window (vrlo_set_build_crop) */
                    const int cpx = bx0 + (int) compact_bits(i), cpy = by0 + (int) compact_bits(i >> 1);
                    if (cpx < g_build_crop[0] || cpy < g_build_crop[1] || cpx >= g_build_crop[2] || cpy >= g_build_crop[3]) continue;
                }
                auto on_query = [&](const PartOut &p, Sampler &smp, int px, int py, int j) {
                    if (p.n_all > 0) {
                        /* atomic query: the parts as consecutive records sharing the query's stream position;
                         * sample = j | n_parts << 16 (include/vrl_b200.h, vrlb_segment) */
                        for (int k = 0; k < p.n_all; ++k, ++n) {
                            if (n >= cap) continue;
                            vrlb_segment &g = segs[n]; const Ray &r = p.all[k];
                            g.o[0] = r.o.x; g.o[1] = r.o.y; g.o[2] = r.o.z; g.maxt = r.maxt;
                            g.d[0] = r.d.x; g.d[1] = r.d.y; g.d[2] = r.d.z;
                            g.pixel = (uint32_t) (py * cam->width + px);
                            g.weight[0] = p.weight.x; g.weight[1] = p.weight.y; g.weight[2] = p.weight.z;
                            g.medium = p.medium; g.rng_state = smp.rng.state;
                            g.sample = (uint32_t) j | ((uint32_t) p.n_all << 16); g.part = (uint32_t) k;
                        }
                        smp.rng.advance((int64_t) ((uint64_t) q * n_vrls_total * dpe));
                        return V3(0.f);
                    }
                    if (n < cap) {
                        vrlb_segment &g = segs[n];
                        g.o[0] = p.ray.o.x; g.o[1] = p.ray.o.y; g.o[2] = p.ray.o.z; g.maxt = p.ray.maxt;
                        g.d[0] = p.ray.d.x; g.d[1] = p.ray.d.y; g.d[2] = p.ray.d.z;
                        g.pixel = (uint32_t) (py * cam->width + px);
                        g.weight[0] = p.weight.x; g.weight[1] = p.weight.y; g.weight[2] = p.weight.z;
                        g.medium = p.medium; g.rng_state = smp.rng.state; g.sample = (uint32_t) j; g.part = p.part;
                    }
                    ++n;
                    smp.rng.advance((int64_t) ((uint64_t) q * n_vrls_total * dpe));
                    return V3(0.f);
                };
                auto put = [&](int x, int y, V3 rad, float alpha) {
                    size_t p = (size_t) y * cam->width + x;
                    if (direct_rgb) { direct_rgb[p * 3] += rad.x; direct_rgb[p * 3 + 1] += rad.y; direct_rgb[p * 3 + 2] += rad.z; }
                    if (alpha_w) { alpha_w[p * 2] += alpha; alpha_w[p * 2 + 1] += 1.f; }
                };
                render_pixel(*s, *cam, *opt, bid, i, bx0, by0, bw, bh, on_query, put);
            }
        }
    /* curved camera paths queried part by part: the records of one path are consecutive; each gets the path's part count
     * in the high half of `part` (include/vrl_b200.h, vrlb_segment) -- the direct in-scattering term walks the group */
    if (opt->use_non_linear_camera && !opt->use_nl_atomic_query) {
        const uint64_t m = std::min(n, cap);
        for (uint64_t a = 0; a < m;) {
            uint64_t b = a + 1;
            const vrlb_medium &med = s->media[(size_t) segs[a].medium];
            if (!(med.kind == VRLB_MEDIUM_NONLINEAR && nl_step(med) > 0.f)) { a = b; continue; }
            while (b < m && segs[b].part == segs[b - 1].part + 1 && segs[b].pixel == segs[a].pixel && segs[b].sample == segs[a].sample) ++b;
            const uint32_t enc = (uint32_t) std::min<uint64_t>(b - a, 0xffffu) << 16;
            for (uint64_t k = a; k < b; ++k) segs[k].part |= enc;
            a = b;
        }
    }
    return n;
}

/* The surface queries of the camera paths (one per path that ends on a smooth surface with a non-zero reflectance), in the
 * order vrlo_build_segments walks the film; returns the count (may exceed cap). */
uint64_t vrlo_build_surface_queries(const vrlo_scene *s, const vrlb_camera *cam, const vrlb_options *opt, uint64_t n_vrls_total,
                                    vrlb_surface_query *out, uint64_t cap) {
    std::vector<vrlb_surface_query> found;
    g_surface_queries = &found;
    vrlo_build_segments(s, cam, opt, n_vrls_total, VRLO_RNG_ADDRESSED, nullptr, 0, nullptr, nullptr);
    g_surface_queries = nullptr;
    for (uint64_t i = 0; i < std::min<uint64_t>(cap, found.size()); ++i) out[i] = found[i];
    return found.size();
}

uint64_t vrlo_render(const vrlo_scene *s, const vrlb_camera *cam, const vrlb_options *opt,
                     const vrlb_vrlset_desc *vrls, int rng_mode, int n_threads, const int32_t *crop,
                     float *xyzaw, float *rgb) {
    int nbx, nby;
    std::vector<uint32_t> ids = spiral_ids(cam->width, cam->height, opt->block_size, nbx, nby);
    size_t npix = (size_t) cam->width * cam->height;
    if (xyzaw) std::memset(xyzaw, 0, sizeof(float) * 5 * npix);
    if (rgb) std::memset(rgb, 0, sizeof(float) * 3 * npix);
    std::atomic<uint64_t> evals{ 0 };
    int cx0 = crop ? crop[0] : 0, cy0 = crop ? crop[1] : 0, cx1 = crop ? crop[2] : cam->width, cy1 = crop ? crop[3] : cam->height;
    /* one task per block, as tbb::parallel_for over blocks (integrator.cpp:111-142); blocks own
     * disjoint pixels with the box filter, so no film mutex is needed */
    parallel_for((uint64_t) nbx * nby, n_threads, [&](uint64_t b) {
        int bx = (int) (b % nbx), by = (int) (b / nbx);
        int bx0 = bx * opt->block_size, by0 = by * opt->block_size;
        int bw = std::min(opt->block_size, cam->width - bx0), bh = std::min(opt->block_size, cam->height - by0);
        if (bx0 >= cx1 || by0 >= cy1 || bx0 + bw <= cx0 || by0 + bh <= cy0) return;
        uint64_t local = 0;
        for (uint32_t i = 0; i < (uint32_t) (opt->block_size * opt->block_size); ++i) {
            uint32_t lx = compact_bits(i), ly = compact_bits(i >> 1);
            int px = bx0 + (int) lx, py = by0 + (int) ly;
            if (px < cx0 || py < cy0 || px >= cx1 || py >= cy1) continue;
            auto on_query = [&](const PartOut &p, Sampler &smp, int, int, int) {
                V3 li; double li64[3];
                if (p.n_all > 0) {
                    std::vector<V3> l32((size_t) p.n_all); std::vector<std::array<double, 3>> l64((size_t) p.n_all);
                    local += query_atomic(*s, *vrls, p.all, p.n_all, *opt, rng_mode, smp.rng, l32.data(),
                                          reinterpret_cast<double (*)[3]>(l64.data()));
                    li = V3(0.f);
                    for (const V3 &v : l32) li = li + v;
                    return li;
                }
                local += query(*s, *vrls, p.ray, *opt, rng_mode, smp.rng, li, li64);
                return li;
            };
            auto put = [&](int x, int y, V3 rad, float alpha) {
                size_t p = (size_t) y * cam->width + x;
                if (rgb) { rgb[p * 3] += rad.x; rgb[p * 3 + 1] += rad.y; rgb[p * 3 + 2] += rad.z; }
                if (xyzaw) {
                    V3 xyz = srgb_to_xyz(rad);
                    xyzaw[p * 5 + 0] += xyz.x; xyzaw[p * 5 + 1] += xyz.y; xyzaw[p * 5 + 2] += xyz.z;
                    xyzaw[p * 5 + 3] += alpha; xyzaw[p * 5 + 4] += 1.f;
                }
            };
            render_pixel(*s, *cam, *opt, ids[(size_t) by * nbx + bx], i, bx0, by0, bw, bh, on_query, put);
        }
        evals += local;
    });
    return evals.load();
}

}  // extern "C"

/* ========================================================================================== */
/*  Part 7: the VRL light cut -- src/integrators/vrl/vrl_lightcut.h.            */
/*  Pinned by tests/golden/ref_lightcut.json: that header compiled in place (ref_probe_lc.cpp) */
/* ========================================================================================== */
namespace {

/* ---- BoundingBox::getMinDistanceSqr(ray), include/mitsuba/core/bbox.h:335-816 ---------------------------------------
 * Squared distance between a box and a ray part.  The fork's routine is the published line / axis-aligned-box distance
 * query of Eberly ("Distance from a line to a box"): the line is reflected into the octant where its direction is
 * non-negative, the face (or edge, corner) of the box the line passes closest to is found from sign tests of cross
 * products, and the line parameter of the closest point decides between the line result and the two end points.
 * Quirks of the fork kept here because the cut's bound is computed from them:
 *   - the slab test in front (bbox.h:302-325, :806) is for the whole LINE, both directions: a part whose supporting line
 *     meets the box is at distance 0 however far the part itself is;
 *   - the box handed to the line query is given by its FULL extents (bbox.h:349, `box.extents()`) where the algorithm
 *     expects half extents: the line result is the distance to a box twice as large (the end-point results are not). */
struct LineBox { float sqr = 0.f, t = 0.f; };

/* the line leaves the reflected box through the face +e[i0] (or passes it beyond an edge / corner of that face) */
void lb_face(int i0, int i1, int i2, const float *p, const float *d, const float *pme, const float *e, LineBox &r) {
    float ppe[3] = { 0.f, 0.f, 0.f };
    ppe[i1] = p[i1] + e[i1];
    ppe[i2] = p[i2] + e[i2];
    /* closest box point at offsets (pme[i0], x1, x2) from the line's reference point: bbox.h:460-466 and its siblings */
    auto close = [&](float len_sqr, float x1, float x2) {
        float delta = d[i0] * pme[i0] + d[i1] * x1 + d[i2] * x2;
        float param = -delta / len_sqr;
        r.sqr += pme[i0] * pme[i0] + x1 * x1 + x2 * x2 + delta * param;
        r.t = param;
    };
    /* the edge of the face that runs along axis ia (ib = the face's other axis); with `only_if_ahead` the edge is taken
     * only when the projection falls on its positive side (bbox.h:538, :573) */
    auto edge = [&](int ia, int ib, bool only_if_ahead) -> bool {
        float len_sqr = d[i0] * d[i0] + d[ib] * d[ib];
        float tmp = len_sqr * ppe[ia] - d[ia] * (d[i0] * pme[i0] + d[ib] * ppe[ib]);
        if (only_if_ahead && !(tmp >= 0.f)) return false;
        float xa;
        if (tmp <= 2.f * len_sqr * e[ia]) { float t = tmp / len_sqr; len_sqr += d[ia] * d[ia]; xa = ppe[ia] - t; }   /* interior of the edge */
        else { len_sqr += d[ia] * d[ia]; xa = pme[ia]; }                                                              /* its far vertex       */
        if (ia == i1) close(len_sqr, xa, ppe[i2]); else close(len_sqr, ppe[i1], xa);
        return true;
    };
    const bool in1 = d[i0] * ppe[i1] >= d[i1] * pme[i0];
    if (in1) {
        if (d[i0] * ppe[i2] >= d[i2] * pme[i0]) { float inv = 1.f / d[i0]; r.t = -pme[i0] * inv; return; }   /* crosses the face: distance 0 */
        edge(i1, i2, false);
        return;
    }
    if (d[i0] * ppe[i2] >= d[i2] * pme[i0]) { edge(i2, i1, false); return; }
    if (edge(i1, i2, true)) return;
    if (edge(i2, i1, true)) return;
    float len_sqr = d[i0] * d[i0] + d[i1] * d[i1];                                                            /* the corner */
    len_sqr += d[i2] * d[i2];
    close(len_sqr, ppe[i1], ppe[i2]);
}

/* no zero direction component: bbox.h:626-657 */
void lb_general(const float *p, const float *d, const float *e, LineBox &r) {
    const float pme[3] = { p[0] - e[0], p[1] - e[1], p[2] - e[2] };
    if (d[1] * pme[0] >= d[0] * pme[1]) {
        if (d[2] * pme[0] >= d[0] * pme[2]) lb_face(0, 1, 2, p, d, pme, e, r);
        else lb_face(2, 0, 1, p, d, pme, e, r);
    } else {
        if (d[2] * pme[1] >= d[1] * pme[2]) lb_face(1, 2, 0, p, d, pme, e, r);
        else lb_face(2, 0, 1, p, d, pme, e, r);
    }
}
inline void lb_clamp_axis(int i, const float *p, const float *e, LineBox &r) {       /* the point-box term of one axis */
    if (p[i] < -e[i]) { float delta = p[i] + e[i]; r.sqr += delta * delta; }
    else if (p[i] > e[i]) { float delta = p[i] - e[i]; r.sqr += delta * delta; }
}
/* d[i2] == 0: a 2-D problem in the (i0, i1) plane plus the point-box term along i2, bbox.h:659-726 */
void lb_one_zero(int i0, int i1, int i2, const float *p, const float *d, const float *e, LineBox &r) {
    float pme0 = p[i0] - e[i0], pme1 = p[i1] - e[i1];
    float prod0 = d[i1] * pme0, prod1 = d[i0] * pme1;
    if (prod0 >= prod1) {
        float ppe1 = p[i1] + e[i1];
        float delta = prod0 - d[i0] * ppe1;
        if (delta >= 0.f) {
            float inv_l = 1.f / (d[i0] * d[i0] + d[i1] * d[i1]);
            r.sqr += delta * delta * inv_l;
            r.t = -(d[i0] * pme0 + d[i1] * ppe1) * inv_l;
        } else { float inv = 1.f / d[i0]; r.t = -pme0 * inv; }
    } else {
        float ppe0 = p[i0] + e[i0];
        float delta = prod1 - d[i1] * ppe0;
        if (delta >= 0.f) {
            float inv_l = 1.f / (d[i0] * d[i0] + d[i1] * d[i1]);
            r.sqr += delta * delta * inv_l;
            r.t = -(d[i0] * ppe0 + d[i1] * pme1) * inv_l;
        } else { float inv = 1.f / d[i1]; r.t = -pme1 * inv; }
    }
    lb_clamp_axis(i2, p, e, r);
}
/* only d[i0] != 0: bbox.h:728-761 */
void lb_two_zeros(int i0, int i1, int i2, const float *p, const float *d, const float *e, LineBox &r) {
    r.t = (e[i0] - p[i0]) / d[i0];
    lb_clamp_axis(i1, p, e, r);
    lb_clamp_axis(i2, p, e, r);
}

float box_point_sqr(const V3 &lo, const V3 &hi, const V3 &p) {                        /* bbox.h:221-224 */
    V3 a((p.x < lo.x) ? lo.x - p.x : 0.f, (p.y < lo.y) ? lo.y - p.y : 0.f, (p.z < lo.z) ? lo.z - p.z : 0.f);
    V3 b((p.x > hi.x) ? p.x - hi.x : 0.f, (p.y > hi.y) ? p.y - hi.y : 0.f, (p.z > hi.z) ? p.z - hi.z : 0.f);
    return squared_norm(a + b);
}

float box_ray_sqr(const V3 &lo, const V3 &hi, const Ray &ray) {
    /* bbox.h:806-808 after :302-325 -- the slab test on the line */
    {
        const float o[3] = { ray.o.x, ray.o.y, ray.o.z }, d[3] = { ray.d.x, ray.d.y, ray.d.z };
        const float l[3] = { lo.x, lo.y, lo.z }, h[3] = { hi.x, hi.y, hi.z };
        float mint = -kInf, maxt = kInf;
        bool first = true;
        for (int i = 0; i < 3; ++i) {
            float rcp = 1.0f / d[i];
            float t1 = (l[i] - o[i]) * rcp, t2 = (h[i] - o[i]) * rcp;
            float a = std::min(t1, t2), b = std::max(t1, t2);
            if (first) { mint = a; maxt = b; first = false; } else { mint = std::max(mint, a); maxt = std::min(maxt, b); }
        }
        if (maxt >= mint) return 0.0f;
    }
    /* bbox.h:345-365 */
    V3 c = (hi + lo) * 0.5f, ext = hi - lo;
    V3 pt = ray.o - c;
    float p[3] = { pt.x, pt.y, pt.z }, d[3] = { ray.d.x, ray.d.y, ray.d.z };
    const float e[3] = { ext.x, ext.y, ext.z };
    int positive = 0;
    for (int i = 0; i < 3; ++i) {
        if (d[i] < 0.f) { p[i] = -p[i]; d[i] = -d[i]; }
        if (d[i] > 0.f) positive |= 1 << i;
    }
    LineBox r;
    switch (positive) {                                                                 /* bbox.h:389-429 */
        case 7: lb_general(p, d, e, r); break;
        case 3: lb_one_zero(0, 1, 2, p, d, e, r); break;
        case 5: lb_one_zero(0, 2, 1, p, d, e, r); break;
        case 6: lb_one_zero(1, 2, 0, p, d, e, r); break;
        case 1: lb_two_zeros(0, 1, 2, p, d, e, r); break;
        case 2: lb_two_zeros(1, 0, 2, p, d, e, r); break;
        case 4: lb_two_zeros(2, 0, 1, p, d, e, r); break;
        default: lb_clamp_axis(0, p, e, r); lb_clamp_axis(1, p, e, r); lb_clamp_axis(2, p, e, r); break;
    }
    if (r.t < 0.f) return box_point_sqr(lo, hi, ray.o);
    if (r.t > ray.maxt) return box_point_sqr(lo, hi, ray.o + ray.d * ray.maxt);        /* line(maxt - mint), mint = 0 */
    return r.sqr;
}

/* VRLLightCut::sqrDistanceVRL(Ray3f, VRL), vrl_lightcut.h:772-838: the clamped closest points of two segments
 * (s on the VRL, t on the camera part), with the reference's 1e-4 absolute thresholds */
float seg_seg_sqr(const Ray &r, const Vrl &beam) {
    V3 u = beam.direction * beam.length, v = r.d * r.maxt, w = beam.origin - r.o;
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
    float sc = ((double) std::fabs(sN) < 0.0001) ? 0.f : sN / sD;
    float tc = ((double) std::fabs(tN) < 0.0001) ? 0.f : tN / tD;
    V3 dP = w + (u * sc) - (v * tc);
    return squared_norm(dP);
}

/* ---- the light tree ------------------------------------------------------------------------------------------------ */
struct LcNode {
    int parent = -1, child[2] = { -1, -1 };
    bool leaf = true;
    V3 lo, hi;                       /* aabb of every VRL of the cluster */
    Vrl rep;                         /* representative, flux = the cluster's */
    int rep_leaf = -1;               /* index of the VRL the representative's geometry comes from */
    V3 add_flux = V3(0.f);
    int same = -1;                   /* the child that shares the representative */
    float max_len = 0.f;
    std::vector<int> leaves;         /* Node::nodes: n1's then n2's */
};
inline V3 vmin(V3 a, V3 b) { return V3(std::min(a.x, b.x), std::min(a.y, b.y), std::min(a.z, b.z)); }
inline V3 vmax(V3 a, V3 b) { return V3(std::max(a.x, b.x), std::max(a.y, b.y), std::max(a.z, b.z)); }
inline float hmax3(V3 a) { return std::max(std::max(a.x, a.y), a.z); }

/* Node(const VRL &), vrl_lightcut.h:56-63 */
LcNode lc_leaf(const Vrl &v, int index) {
    LcNode n;
    V3 end = v.origin + v.direction * v.length;
    n.lo = vmin(v.origin, end); n.hi = vmax(v.origin, end);
    n.rep = v; n.rep_leaf = index; n.max_len = v.length;
    n.leaves.push_back(index);
    return n;
}
/* Node::dist, vrl_lightcut.h:101-109 */
inline float lc_dist(const LcNode &n1, const LcNode &n2) {
    float total = hmax3(n1.rep.flux + n2.rep.flux);
    V3 ext = vmax(n1.hi, n2.hi) - vmin(n1.lo, n2.lo);
    return squared_norm(ext) * total;
}
/* Node(Node *n1, Node *n2, Float v, bool), vrl_lightcut.h:111-151 (children referenced by index) */
LcNode lc_merge(const LcNode &n1, int i1, const LcNode &n2, int i2, float v) {
    LcNode m;
    m.lo = vmin(n1.lo, n2.lo); m.hi = vmax(n1.hi, n2.hi);
    m.leaf = false;
    m.max_len = std::max(n1.max_len, n2.max_len);
    m.child[0] = i1; m.child[1] = i2;
    m.leaves = n1.leaves;
    m.leaves.insert(m.leaves.end(), n2.leaves.begin(), n2.leaves.end());
    float leftWM = hmax3(n1.rep.flux) * n1.rep.length, rightWM = hmax3(n2.rep.flux) * n2.rep.length;
    float ratio = leftWM / (leftWM + rightWM);
    const int index = v < ratio ? 0 : 1;
    const LcNode &keep = index == 0 ? n1 : n2, &other = index == 0 ? n2 : n1;
    m.rep = keep.rep; m.rep_leaf = keep.rep_leaf;
    ratio = other.rep.length / keep.rep.length;
    m.add_flux = other.rep.flux * ratio;
    m.rep.flux = m.rep.flux + m.add_flux;
    m.same = index;
    return m;
}

}  // namespace

struct vrlo_lightcut {
    std::vector<LcNode> nodes;       /* leaves 0..n-1 in VRL order, then the merges in creation order */
    int root = -1;
    uint64_t n_vrls = 0, build_draws = 0;
    std::vector<int> preorder;       /* node ids in preorder (what the reference probe prints) */
};

namespace {

/* buildLightTree, vrl_lightcut.h:515-627: greedy bottom-up clustering.  Each live cluster keeps one queue entry
 * (itself, its cheapest partner at the time the entry was made); the cheapest entry is popped; it merges when both
 * ends are still live, else the live end(s) search again.  One sampler draw per merge picks the representative. */
void lc_build(vrlo_lightcut &lc, const vrlb_vrlset_desc &set, Sampler &sampler, int n_threads) {
    const size_t n = (size_t) set.n;
    lc.n_vrls = n;
    lc.nodes.reserve(2 * n);
    for (size_t j = 0; j < n; ++j) lc.nodes.push_back(lc_leaf(load_vrl(set, j), (int) j));
    if (n == 0) return;
    std::vector<int> slot(n);
    for (size_t j = 0; j < n; ++j) slot[j] = (int) j;
    struct QE { size_t id1; int n1; size_t id2; int n2; float distance; bool operator>(const QE &o) const { return distance > o.distance; } };
    std::priority_queue<QE, std::vector<QE>, std::greater<QE>> pairs;
    auto min_pair = [&](int node, size_t i) -> std::pair<size_t, float> {
        size_t o = 0; float best = std::numeric_limits<float>::max();
        for (size_t j = 0; j < n; ++j) {
            if (i == j || slot[j] < 0) continue;
            float d = lc_dist(lc.nodes[(size_t) node], lc.nodes[(size_t) slot[j]]);
            if (best > d) { o = j; best = d; }
        }
        return { o, best };
    };
    {
        std::vector<std::pair<size_t, float>> first(n);
        parallel_for(n, n_threads, [&](uint64_t i) { first[i] = min_pair(slot[i], (size_t) i); });
        for (size_t i = 0; i < n; ++i) pairs.push(QE{ i, slot[i], first[i].first, slot[first[i].first], first[i].second });
    }
    std::vector<char> has_parent(2 * n, 0);
    size_t live = n;
    while (live != 1) {
        QE p = pairs.top(); pairs.pop();
        const bool free1 = !has_parent[(size_t) p.n1], free2 = p.n2 >= 0 && !has_parent[(size_t) p.n2];
        if (free1 && free2) {
            float v = sampler.next_1d(); ++lc.build_draws;
            const int id = (int) lc.nodes.size();
            lc.nodes.push_back(lc_merge(lc.nodes[(size_t) p.n1], p.n1, lc.nodes[(size_t) p.n2], p.n2, v));
            lc.nodes[(size_t) p.n1].parent = id; lc.nodes[(size_t) p.n2].parent = id;
            has_parent[(size_t) p.n1] = has_parent[(size_t) p.n2] = 1;
            slot[p.id1] = id; slot[p.id2] = -1;
            auto b = min_pair(id, p.id1);
            pairs.push(QE{ p.id1, id, b.first, slot[b.first], b.second });
            --live;
        } else {
            if (free1) { auto b = min_pair(p.n1, p.id1); pairs.push(QE{ p.id1, p.n1, b.first, slot[b.first], b.second }); }
            if (free2) { auto b = min_pair(p.n2, p.id2); pairs.push(QE{ p.id2, p.n2, b.first, slot[b.first], b.second }); }
        }
    }
    for (size_t i = 0; i < n; ++i) if (slot[i] >= 0) lc.root = slot[i];
    /* preorder numbering */
    std::vector<int> stack{ lc.root };
    while (!stack.empty()) {
        int id = stack.back(); stack.pop_back();
        lc.preorder.push_back(id);
        const LcNode &nd = lc.nodes[(size_t) id];
        if (!nd.leaf) { stack.push_back(nd.child[1]); stack.push_back(nd.child[0]); }
    }
}

/* getClusterUpperBound, vrl_lightcut.h:429-512 */
float lc_bound(const vrlo_scene &s, const vrlo_lightcut &lc, const LcNode &cluster, const Ray *parts, int n_parts, float maxt_total, int better_dist, bool *unsupported) {
    const vrlb_medium &m = s.media[(size_t) cluster.rep.medium];
    float min_length;
    if ((int) cluster.leaves.size() > better_dist) {
        float best = std::numeric_limits<float>::max();
        for (int j = 0; j < n_parts; ++j) best = std::min(best, box_ray_sqr(cluster.lo, cluster.hi, parts[j]));
        min_length = safe_sqrt(best);
    } else {
        float best = std::numeric_limits<float>::max();
        for (int leaf : cluster.leaves) {
            float per = std::numeric_limits<float>::max();
            for (int j = 0; j < n_parts; ++j) per = std::min(per, seg_seg_sqr(parts[j], lc.nodes[(size_t) leaf].rep));
            best = std::min(best, per);
        }
        min_length = safe_sqrt(best);
    }
    if ((double) min_length < 0.0001) return std::numeric_limits<float>::max();
    if (!is_homogeneous(m)) { if (unsupported) *unsupported = true; return std::numeric_limits<float>::max(); }   /* Log(Error, ...), :485 */
    float tr = homo_transmittance(m, 0.f, min_length);
    V3 ss = sigma_s_const(m);
    V3 material = V3(1.f) * ss;
    material = material * ss;
    const double k = (m.phase == VRLB_PHASE_ISOTROPIC ? 1.0 : 2.0) / (16 * M_PI * M_PI);
    material = V3((float) ((double) material.x * k), (float) ((double) material.y * k), (float) ((double) material.z * k));
    float invRayPDF = cluster.max_len;
    float invVRLPDF = (float) (2.0 * std::atan((double) maxt_total * 0.5 / (double) min_length) / (double) min_length);
    float invPdf = invRayPDF * invVRLPDF;
    V3 tmp = cluster.rep.flux * invPdf * material * V3(tr);
    return hmax3(tmp);
}

/* getContrib on an NLRay (vrl_struct.h:584-662 with samplingVRL :247-360): importance sampling reads the first part
 * only; uniform sampling walks the parts (NLRay::at) */
V3 lc_contrib(const vrlo_scene &s, bool trace, const Vrl &vrl, const Ray *parts, int n_parts, float maxt_total, bool uniform, Pcg32 *r, Pcg32 *sub) {
    if (n_parts == 1 || !uniform) {
        DrawSrc draws{ r, { 0.f, 0.f }, 0 };
        return get_contrib(s, trace, vrl, parts[0], uniform, draws, sub);
    }
    float u1 = r->next_f32(), u2 = r->next_f32();
    float tCam = u1 * maxt_total, tVRL = u2 * vrl.length;
    int hit = n_parts - 1; V3 pCam; bool found = false;
    float dist = tCam;
    if (dist < maxt_total)
        for (int k = 0; k < n_parts; ++k) {
            if (dist < parts[k].maxt) { hit = k; pCam = parts[k].o + parts[k].d * dist; found = true; break; }
            dist -= parts[k].maxt;
        }
    if (!found) { hit = n_parts - 1; pCam = parts[hit].o + parts[hit].d * tCam; }
    SamplingInfo si{ pCam, tCam, vrl.origin + vrl.direction * tVRL, tVRL, vrl.length * maxt_total, true };
    Ray wi_ray{ parts[0].o, parts[hit].d, parts[0].maxt };
    return get_contrib_sampled(s, trace, vrl, wi_ray, si, sub);
}

struct LcStats { uint64_t evals = 0, pops = 0, max_heap = 0; float min_margin = std::numeric_limits<float>::max(); bool unsupported = false; };

/* VRLLightCut::query, vrl_lightcut.h:243-393 (stochastic = false), times m_scale and over samples_per_query as
 * VRLMap::query wraps it (vrl_map.h:143, 198-205).  Addressed draws: the fresh evaluation of a cluster whose
 * representative is VRL j, in sample i, reads the stream at ((i * total + j) * 2) -- every VRL is the representative of
 * exactly one chain of clusters and only the top of the chain is evaluated, so no position is read twice. */
V3 lc_query(const vrlo_scene &s, const vrlo_lightcut &lc, float scale, const Ray *parts, int n_parts, const vrlb_options &opt,
            int rng_mode, Pcg32 &rng, LcStats &st) {
    V3 total(0.f);
    if (lc.root < 0) return total;
    const bool uniform = opt.use_uniform_sampling != 0, trace = opt.visibility != 0;
    const uint32_t q = (uint32_t) std::max(opt.samples_per_query, 1);
    float maxt_total = 0.f;
    for (int k = 0; k < n_parts; ++k) maxt_total += parts[k].maxt;
    const Pcg32 base = rng;
    struct Cut { int node; float bound; V3 est; };
    auto cmp = [](const Cut &a, const Cut &b) { return a.bound < b.bound; };
    for (uint32_t i = 0; i < q; ++i) {
        auto eval = [&](const LcNode &nd) -> V3 {
            if (rng_mode == VRLO_RNG_ADDRESSED) {
                Pcg32 r = base;
                r.advance((int64_t) (((uint64_t) i * lc.n_vrls + (uint64_t) nd.rep_leaf) * 2u));
                Pcg32 sub; sub.seed(r.state, PCG32_DEFAULT_STREAM + 1);
                return lc_contrib(s, trace, nd.rep, parts, n_parts, maxt_total, uniform, &r, &sub);
            }
            return lc_contrib(s, trace, nd.rep, parts, n_parts, maxt_total, uniform, &rng, &rng);
        };
        auto bound = [&](const LcNode &nd) { return lc_bound(s, lc, nd, parts, n_parts, maxt_total, opt.threshold_better_dist, &st.unsupported); };
        std::priority_queue<Cut, std::vector<Cut>, decltype(cmp)> heap(cmp);
        const LcNode &root = lc.nodes[(size_t) lc.root];
        {
            float b = bound(root);
            V3 e = eval(root);
            heap.push(Cut{ lc.root, b, e });
        }
        V3 Li = heap.top().est;
        while (!heap.empty()) {
            Cut cur = heap.top(); heap.pop();
            const float thr = hmax3(Li * opt.threshold_error);
            {
                float margin = std::fabs(cur.bound - thr) / std::max(std::fabs(thr), 1e-30f);
                st.min_margin = std::min(st.min_margin, margin);
            }
            if (cur.bound < thr) break;
            ++st.pops;
            Li = Li - cur.est;
            const LcNode &cl = lc.nodes[(size_t) cur.node];
            for (int c = 0; c < 2; ++c) {
                if (cl.child[c] < 0) continue;
                const LcNode &ch = lc.nodes[(size_t) cl.child[c]];
                V3 est = ch.rep.flux;
                if (cl.same == c) {
                    V3 div = cl.rep.flux;
                    if (div.x > 0.0f) est.x /= div.x;
                    if (div.y > 0.0f) est.y /= div.y;
                    if (div.z > 0.0f) est.z /= div.z;
                    est = est * cur.est;
                    if (std::isnan(est.x) || std::isnan(est.y) || std::isnan(est.z)) est = V3(0.f);
                } else {
                    ++st.evals;
                    est = eval(ch);
                }
                Li = Li + est;
                if (!ch.leaf) heap.push(Cut{ cl.child[c], bound(ch), est });
            }
            st.max_heap = std::max<uint64_t>(st.max_heap, heap.size());
        }
        Li = V3(std::max(Li.x, 0.f), std::max(Li.y, 0.f), std::max(Li.z, 0.f));
        total = total + Li * scale;
    }
    if (rng_mode == VRLO_RNG_ADDRESSED) { rng = base; rng.advance((int64_t) ((uint64_t) q * lc.n_vrls * 2u)); }
    if (q > 1) total = total / (float) q;
    return total;
}

}  // namespace

extern "C" {

/* sampler.seed(0, seed_offset) supplies the representative draws, one per merge */
vrlo_lightcut *vrlo_lightcut_build(const vrlb_vrlset_desc *set, uint64_t base_seed, uint64_t seed_offset, int n_threads) {
    vrlo_lightcut *lc = new vrlo_lightcut();
    Sampler smp; smp.seed(base_seed, seed_offset);
    lc_build(*lc, *set, smp, n_threads);
    return lc;
}
void vrlo_lightcut_destroy(vrlo_lightcut *lc) { delete lc; }
uint64_t vrlo_lightcut_n_nodes(const vrlo_lightcut *lc) { return lc->preorder.size(); }
uint64_t vrlo_lightcut_build_draws(const vrlo_lightcut *lc) { return lc->build_draws; }
/* preorder dump.  f13: [n][13] = flux(3) add_flux(3) lo(3) hi(3) max_len; i6: [n][6] = leaf, representative VRL, the child sharing it, n_leaves, creation id, parent's creation id */
void vrlo_lightcut_nodes(const vrlo_lightcut *lc, float *f13, int32_t *i6) {
    for (size_t k = 0; k < lc->preorder.size(); ++k) {
        const LcNode &n = lc->nodes[(size_t) lc->preorder[k]];
        float *f = f13 + 13 * k; int32_t *i = i6 + 6 * k;
        f[0] = n.rep.flux.x; f[1] = n.rep.flux.y; f[2] = n.rep.flux.z;
        f[3] = n.add_flux.x; f[4] = n.add_flux.y; f[5] = n.add_flux.z;
        f[6] = n.lo.x; f[7] = n.lo.y; f[8] = n.lo.z; f[9] = n.hi.x; f[10] = n.hi.y; f[11] = n.hi.z; f[12] = n.max_len;
        i[0] = n.leaf ? 1 : 0; i[1] = n.rep_leaf; i[2] = n.same; i[3] = (int32_t) n.leaves.size(); i[4] = lc->preorder[k]; i[5] = n.parent;
    }
}
float vrlo_lc_node_dist(const float a10[10], const float b10[10]) {
    vrlb_vrlset_desc sa{ a10, nullptr, 1, 0, 1, 1.f, 0.f }, sb{ b10, nullptr, 1, 0, 1, 1.f, 0.f };
    return lc_dist(lc_leaf(load_vrl(sa, 0), 0), lc_leaf(load_vrl(sb, 0), 1));
}
/* out14 = pick, flux(3), add_flux(3), lo(3), hi(3), max_len */
void vrlo_lc_merge(const float a10[10], const float b10[10], float v, float *out14) {
    vrlb_vrlset_desc sa{ a10, nullptr, 1, 0, 1, 1.f, 0.f }, sb{ b10, nullptr, 1, 0, 1, 1.f, 0.f };
    LcNode m = lc_merge(lc_leaf(load_vrl(sa, 0), 0), 0, lc_leaf(load_vrl(sb, 0), 1), 1, v);
    out14[0] = (float) m.same; out14[1] = m.rep.flux.x; out14[2] = m.rep.flux.y; out14[3] = m.rep.flux.z;
    out14[4] = m.add_flux.x; out14[5] = m.add_flux.y; out14[6] = m.add_flux.z;
    out14[7] = m.lo.x; out14[8] = m.lo.y; out14[9] = m.lo.z; out14[10] = m.hi.x; out14[11] = m.hi.y; out14[12] = m.hi.z; out14[13] = m.max_len;
}
float vrlo_lc_box_dist(const float lo[3], const float hi[3], const float o[3], const float d[3], float maxt) {
    return box_ray_sqr(V3(lo), V3(hi), Ray{ V3(o), V3(d), maxt });
}
float vrlo_lc_seg_dist(const float vrl10[10], const float o[3], const float d[3], float maxt) {
    vrlb_vrlset_desc sa{ vrl10, nullptr, 1, 0, 1, 1.f, 0.f };
    return seg_seg_sqr(Ray{ V3(o), V3(d), maxt }, load_vrl(sa, 0));
}
/* parts7: n_parts x (o(3), d(3), maxt); node = preorder index */
float vrlo_lightcut_bound(const vrlo_scene *s, const vrlo_lightcut *lc, int32_t node, const float *parts7, int n_parts, int better_dist) {
    std::vector<Ray> parts((size_t) n_parts);
    float maxt = 0.f;
    for (int k = 0; k < n_parts; ++k) { parts[(size_t) k] = Ray{ V3(parts7 + 7 * k), V3(parts7 + 7 * k + 3), parts7[7 * k + 6] }; maxt += parts7[7 * k + 6]; }
    return lc_bound(*s, *lc, lc->nodes[(size_t) lc->preorder[(size_t) node]], parts.data(), n_parts, maxt, better_dist, nullptr);
}
/* one query on the stream sampler.seed(0, seed_offset) (sequential mode) / addressed from that position.
 * stats5 = evals (the reference's nb_evaluation), draws consumed (sequential mode), pops, largest heap, min decision margin */
void vrlo_lightcut_query(const vrlo_scene *s, const vrlo_lightcut *lc, float scale, const float *parts7, int n_parts, const vrlb_options *opt,
                         int rng_mode, uint64_t base_seed, uint64_t seed_offset, float out_rgb[3], double *stats5) {
    std::vector<Ray> parts((size_t) n_parts);
    for (int k = 0; k < n_parts; ++k) parts[(size_t) k] = Ray{ V3(parts7 + 7 * k), V3(parts7 + 7 * k + 3), parts7[7 * k + 6] };
    Sampler smp; smp.seed(base_seed, seed_offset);
    Pcg32 start = smp.rng;
    LcStats st;
    V3 li = lc_query(*s, *lc, scale, parts.data(), n_parts, *opt, rng_mode, smp.rng, st);
    out_rgb[0] = li.x; out_rgb[1] = li.y; out_rgb[2] = li.z;
    uint64_t draws = 0;
    if (rng_mode == VRLO_RNG_SEQUENTIAL) { while (start.state != smp.rng.state && draws < (1u << 24)) { start.next_u32(); ++draws; } }
    if (stats5) { stats5[0] = (double) st.evals; stats5[1] = (double) draws; stats5[2] = (double) st.pops; stats5[3] = (double) st.max_heap; stats5[4] = (double) st.min_margin; }
}
/* the gather through the cut over explicit segments (each record one query, or a group of n_parts records when
 * use_nl_atomic_query): out_rgb32 per segment (groups: on the first record), stats4 per segment = evals, pops, largest
 * heap, min decision margin.  Returns the fresh evaluations, or UINT64_MAX when a cluster sits in a non-homogeneous medium. */
uint64_t vrlo_lightcut_gather_segments(const vrlo_scene *s, const vrlo_lightcut *lc, float scale, const vrlb_segment *segs, uint64_t n_segs,
                                       const vrlb_options *opt, int rng_mode, int n_threads, float *out_rgb32, float *stats4) {
    std::atomic<uint64_t> evals{ 0 };
    std::atomic<int> bad{ 0 };
    parallel_for(n_segs, n_threads, [&](uint64_t i) {
        const vrlb_segment &sg = segs[i];
        const uint32_t n_all = sg.sample >> 16;
        if (n_all > 0 && (sg.part & 0xffffu) != 0) { if (out_rgb32) out_rgb32[i * 3] = out_rgb32[i * 3 + 1] = out_rgb32[i * 3 + 2] = 0.f; return; }
        const uint32_t n_parts = n_all > 0 ? n_all : 1;
        std::vector<Ray> parts(n_parts);
        for (uint32_t k = 0; k < n_parts; ++k) parts[k] = Ray{ V3(segs[i + k].o), V3(segs[i + k].d), segs[i + k].maxt };
        Pcg32 rng{ sg.rng_state, (PCG32_DEFAULT_STREAM << 1) | 1u };
        LcStats st;
        V3 li = lc_query(*s, *lc, scale, parts.data(), (int) n_parts, *opt, rng_mode, rng, st);
        if (st.unsupported) bad = 1;
        evals += st.evals;
        if (out_rgb32) { out_rgb32[i * 3 + 0] = li.x; out_rgb32[i * 3 + 1] = li.y; out_rgb32[i * 3 + 2] = li.z; }
        if (stats4) { stats4[i * 4 + 0] = (float) st.evals; stats4[i * 4 + 1] = (float) st.pops; stats4[i * 4 + 2] = (float) st.max_heap; stats4[i * 4 + 3] = st.min_margin; }
    });
    return bad.load() ? UINT64_MAX : evals.load();
}

}  // extern "C"
