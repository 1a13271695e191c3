/* TEST INFRASTRUCTURE.  A minimal stand-in for the Mitsuba 2 types that the reference's
 * src/integrators/vrl/vrl_struct.h names, so that THAT header (the reference's own VRL maths:
 * findClosetPoint, sampleMC, samplingVRL, evalTransmittance, getContrib) can be compiled where it
 * lies under /root/reference with plain g++ -- the full Mitsuba build is not possible here
 * (DESIGN.md section 2).  The arithmetic types are the reference's own vendored Enoki
 * (ext/enoki/include, header-only), so enoki::log/sinh/atan/tan/exp/safe_sqrt are the real ones.
 *
 * What is mocked (written here, NOT reference code): Ray (defaults as include/mitsuba/core/ray.h:34-48),
 * the epsilons of include/mitsuba/core/math.h:28-38, a homogeneous Medium whose transmittance follows
 * include/mitsuba/render/medium.h:109-118, isotropic/HG phase eval (src/phase/isotropic.cpp:142-146,
 * src/phase/hg.cpp:51-54,77-81), an IndependentSampler over the real enoki::PCG32
 * (src/librender/sampler.cpp:83-96), and a brute-force triangle Scene whose surfaces are all
 * non-null BSDFs.  Nothing under oracle/ is product code. */
#pragma once
#include <enoki/array.h>
#include <enoki/random.h>
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <iostream>
#include <limits>
#include <sstream>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#ifndef M_PI_2
#define M_PI_2 1.57079632679489661923
#endif

#define NAMESPACE_BEGIN(name) namespace name {
#define NAMESPACE_END(name) }

namespace mitsuba {
using namespace enoki; /* include/mitsuba/core/fwd.h:13 */

namespace math {
template <typename T> constexpr auto Infinity      = std::numeric_limits<scalar_t<T>>::infinity();
template <typename T> constexpr auto Epsilon       = std::numeric_limits<scalar_t<T>>::epsilon() / 2;
template <typename T> constexpr auto RayEpsilon    = Epsilon<T> * 1500;
template <typename T> constexpr auto ShadowEpsilon = RayEpsilon<T> * 10;
template <typename T> constexpr auto Max           = std::numeric_limits<scalar_t<T>>::max();   /* math.h:30 */
} // namespace math

enum LogLevel : int { Trace = 0, Debug = 100, Info = 200, Warn = 300, Error = 400 };   /* core/logger.h:11-17 (unscoped: the reference writes both Log(Error, ..) and Log(LogLevel::Error, ..)) */
struct MockLog { static inline int warnings = 0; };
template <typename... Args> inline void Log(LogLevel level, Args &&...) {
    if (level == LogLevel::Error) throw std::runtime_error("Log(Error)");
    if (level == LogLevel::Warn) MockLog::warnings++;
}
namespace string { template <typename T> const T &indent(const T &v) { return v; } }

template <typename T> constexpr bool is_rgb_v = true;

#ifdef REF_MOCK_VEC3
using MockVec3 = REF_MOCK_VEC3;     /* a probe's own 3-vector (ref_probe_lc.cpp: the real bbox.h asks Point::Vector) */
#else
using MockVec3 = enoki::Array<float, 3>;
#endif
using MockVec2 = enoki::Array<float, 2>;
#ifndef REF_MOCK_VEC3
using Vector   = MockVec3; /* names the dead anisotropic branch of samplingVRL uses unqualified */
#endif

struct MockRay {
    MockVec3 o, d, d_rcp;
    float mint = math::RayEpsilon<float>;
    float maxt = math::Infinity<float>;
    float time = 0.f;
    using Float = float; using Vector = MockVec3;      /* bbox.h:335-337 reads Ray::Float / Ray::Vector */
    MockRay() = default;
    MockRay(const MockVec3 &o_, const MockVec3 &d_, const float &t) : o(o_), d(d_), time(t) { d_rcp = rcp(d); }
    MockVec3 operator()(float t) const { return fmadd(d, MockVec3(t), o); }
};
inline std::ostream &operator<<(std::ostream &os, const MockRay &r) { return os << "Ray[" << r.o << r.d << "]"; }

struct MockFrame {
    MockVec3 s, t, n;
    MockFrame() : s(1, 0, 0), t(0, 1, 0), n(0, 0, 1) {}
    explicit MockFrame(const MockVec3 &n_) : n(n_) {
        /* any orthonormal completion: evalTransmittance only feeds wo to a BSDF eval that the
           mock answers with 0 (Null::eval) or never reaches */
        MockVec3 a = std::abs(n.x()) > 0.9f ? MockVec3(0, 1, 0) : MockVec3(1, 0, 0);
        s = normalize(cross(a, n)); t = cross(n, s);
    }
    MockVec3 to_local(const MockVec3 &v) const { return MockVec3(dot(v, s), dot(v, t), dot(v, n)); }
};

enum class BSDFFlags : uint32_t { None = 0, Null = 0x1, DiffuseReflection = 0x2 };
inline bool has_flag(uint32_t flags, BSDFFlags f) { return (flags & (uint32_t) f) != 0; }
enum class TransportMode { Radiance, Importance };
struct MockBSDFContext { MockBSDFContext(TransportMode, uint32_t, int) {} explicit MockBSDFContext(TransportMode) {} };

struct MockMedium;
struct MockSI;
struct MockBSDF {
    uint32_t m_flags = (uint32_t) BSDFFlags::DiffuseReflection;
    /* diffuse = false: what evalTransmittance meets (Null::eval = 0, or a value never read).  diffuse = true:
       SmoothDiffuse::eval (src/bsdfs/diffuse.cpp), for the photon-map probe */
    bool diffuse = false;
    MockVec3 reflectance = MockVec3(0.5f);
    uint32_t flags() const { return m_flags; }
    inline MockVec3 eval(const MockBSDFContext &, const MockSI &si, const MockVec3 &wo) const;
};
struct MockSI {
    float t = math::Infinity<float>;
    MockVec3 p, n;
    MockVec3 wi = MockVec3(0.f, 0.f, 1.f);      /* local */
    MockFrame sh_frame;
    MockVec2 duv_dx, duv_dy;
    const MockBSDF *m_bsdf = nullptr;
    MockVec3 to_local(const MockVec3 &v) const { return sh_frame.to_local(v); }
    MockVec3 to_world_mueller(const MockVec3 &value, const MockVec3 &, const MockVec3 &) const { return value; }
    bool is_valid() const { return t != math::Infinity<float>; }
    const MockBSDF *bsdf() const { return m_bsdf; }
    bool is_medium_transition() const { return false; }
    const MockMedium *target_medium(const MockVec3 &) const { return nullptr; }
};
inline MockVec3 MockBSDF::eval(const MockBSDFContext &, const MockSI &si, const MockVec3 &wo) const {
    if (!diffuse) return MockVec3(0.f);
    float cos_theta_i = si.wi.z(), cos_theta_o = wo.z();
    bool active = cos_theta_i > 0.f && cos_theta_o > 0.f;
    return active ? reflectance * (0.31830988618379067154f * cos_theta_o) : MockVec3(0.f);
}
struct MockMI {
    MockVec3 p, wi;
    float t = 0.f;
};

struct MockSampler {
    enoki::PCG32<float> rng;
    uint64_t draws = 0;
    void seed(uint64_t base_seed, uint64_t seed_offset) { rng.seed(base_seed + seed_offset, PCG32_DEFAULT_STREAM); draws = 0; }
    float next_1d(bool = true) { ++draws; return rng.next_float32(); }
    float next1D() { return next_1d(); }                  /* named by dead code only (vrl_lightcut.h:744) */
    MockSampler *clone() { return new MockSampler(*this); }
};
struct MockPhaseContext { explicit MockPhaseContext(MockSampler *) {} };

enum class PhaseFunctionFlags : uint32_t { None = 0x00, Isotropic = 0x01, Anisotropic = 0x02, Microflake = 0x04 };   /* render/phase.h:17-22 */
inline bool has_flag(uint32_t flags, PhaseFunctionFlags f) { return (flags & (uint32_t) f) != 0; }                   /* :34-36 */
struct MockPhase {
    bool hg = false;
    float g = 0.f;
    uint32_t flags(bool = true) const { return (uint32_t) (hg ? PhaseFunctionFlags::Anisotropic : PhaseFunctionFlags::Isotropic); }   /* isotropic.cpp / hg.cpp constructors */
    float eval(const MockPhaseContext &, const MockMI &mi, const MockVec3 &wo) const {
        const float inv4pi = 0.07957747154594766788f;
        if (!hg) return inv4pi;
        float cos_theta = dot(wo, mi.wi);
        float temp = 1.0f + g * g + 2.0f * g * cos_theta;
        return inv4pi * (1 - g * g) / (temp * enoki::sqrt(temp));
    }
};

struct MockMedium {
    MockVec3 sigma_t = MockVec3(0.f), albedo = MockVec3(0.f);
    float scale = 1.f;
    MockPhase phase;
    const MockPhase *phase_function() const { return &phase; }
    bool has_spectral_extinction() const { return true; }
    bool is_homogeneous() const { return true; }
    /* homogeneous.cpp:26 with ConstVolume::max() returning the mean (constant3d.cpp:37) */
    float max_density() const { return scale * hmean(sigma_t); }
    std::tuple<MockVec3, MockVec3, MockVec3> get_scattering_coefficients(const MockMI &, bool) const {
        MockVec3 st = sigma_t * scale, ss = st * albedo;
        return { ss, MockVec3(0.f), st };
    }
    MockVec3 evalTransmittance(const MockRay &ray, MockSampler *, bool, bool = false) const {
        float negLength = ray.mint - ray.maxt;
        float val = enoki::exp(max_density() * negLength);
        return MockVec3(max_density() != 0 ? val : 1.0f);
    }
};

struct MockScene {
    std::vector<MockVec3> tri; /* 3 vertices per triangle */
    struct BSphere { float radius = 1.f; };
    BSphere getBSphere() const { return BSphere(); }      /* named by dead code only (vrl_lightcut.h:726) */
    MockBSDF bsdf;
    MockSI ray_intersect(const MockRay &ray, bool = true) const {
        MockSI si;
        float best = ray.maxt;
        for (size_t k = 0; k + 2 < tri.size(); k += 3) {
            MockVec3 e1 = tri[k + 1] - tri[k], e2 = tri[k + 2] - tri[k];
            MockVec3 pv = cross(ray.d, e2);
            float det = dot(e1, pv);
            if (det == 0.f) continue;
            float inv = 1.f / det;
            MockVec3 tv = ray.o - tri[k];
            float u = dot(tv, pv) * inv;
            if (u < 0.f || u > 1.f) continue;
            MockVec3 qv = cross(tv, e1);
            float v = dot(ray.d, qv) * inv;
            if (v < 0.f || u + v > 1.f) continue;
            float t = dot(e2, qv) * inv;
            if (t < ray.mint || t > best) continue;
            best = t; si.t = t; si.n = normalize(cross(e1, e2)); si.m_bsdf = &bsdf;
        }
        return si;
    }
};
} // namespace mitsuba

/* REF_MOCK_MSVC_TYPENAME: photonmapper/kdtree.h writes `typedef typename Point3f PointType;` (accepted by MSVC only).
   With Point3f / Vector3f as macros for a qualified name the line is standard C++; the aliases are then left out here. */
#ifdef REF_MOCK_MSVC_TYPENAME
#define Point3f ::mitsuba::MockVec3
#define Vector3f ::mitsuba::MockVec3
#define REF_MOCK_POINT_ALIASES
#else
#define REF_MOCK_POINT_ALIASES using Point3f = mitsuba::MockVec3; using Vector3f = mitsuba::MockVec3;
#endif
#define MTS_IMPORT_TYPES(...)                                                                      \
    using Mask = bool; using UInt32 = uint32_t; using Int32 = int32_t;                             \
    REF_MOCK_POINT_ALIASES                                                                         \
    using Normal3f = mitsuba::MockVec3; using Ray3f = mitsuba::MockRay;                            \
    using Frame3f = mitsuba::MockFrame; using UnpolarizedSpectrum = mitsuba::MockVec3;             \
    using SurfaceInteraction3f = mitsuba::MockSI; using MediumInteraction3f = mitsuba::MockMI;     \
    using PhaseFunctionContext = mitsuba::MockPhaseContext; using BSDFContext = mitsuba::MockBSDFContext;
#define MTS_IMPORT_OBJECT_TYPES()                                                                  \
    using Scene = mitsuba::MockScene; using Sampler = mitsuba::MockSampler;                        \
    using Medium = mitsuba::MockMedium; using MediumPtr = const mitsuba::MockMedium *;             \
    using PhaseFunction = mitsuba::MockPhase; using BSDF = mitsuba::MockBSDF;
