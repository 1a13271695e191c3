// ref_probe_medium.cpp -- TEST INFRASTRUCTURE.  Compiles the reference's OWN Medium class -- include/mitsuba/render/medium.h
// and src/librender/medium.cpp, where they lie under /root/reference -- against a mock Mitsuba (ref_mock.h + the stand-ins
// below) on the reference's vendored Enoki, and prints inputs + outputs of
//     Medium::evalTransmittance      medium.h:109-155   (the stochastic transmittance of a non-homogeneous medium)
//     Medium::sample_interaction     medium.cpp:36-75
//     Medium::eval_tr_and_pdf        medium.cpp:77-89
//     Medium::homoEvalTransmittance  medium.h:103-107
// for a grid ("heterogeneous") medium as JSON -> tests/golden/ref_medium.json (oracle/make_golden.py); tests hold the
// oracle's grid_transmittance to it on the same PCG32 sub-streams.
//
// What is mocked (written here, NOT reference code): the medium subclass -- the three overrides of
// src/media/heterogeneous.cpp the transmittance reaches (max density = scale * grid max :28, get_combined_extinction
// :36-41, intersect_aabb = bbox.ray_intersect :54-56) --, MediumInteraction::spawn_ray (interaction.h:58-61) and
// is_valid, BoundingBox::ray_intersect (bbox.h:302-325), the sampler over the real enoki::PCG32, Properties / PluginManager
// for the constructor.  Shim for g++: none needed beyond the empty class-registration macros.
#include <cassert>
#include <cstdio>
#include <map>
#include <random>
#include "ref_mock.h"

namespace mitsuba {
template <typename T> struct ref {
    T *p = nullptr;
    ref() = default; ref(T *q) : p(q) {}
    T *get() const { return p; } T *operator->() const { return p; } T &operator*() const { return *p; }
    explicit operator bool() const { return p != nullptr; }
    ref &operator=(T *q) { p = q; return *this; }
};
struct Object { virtual ~Object() {} virtual std::string id() const { return ""; } virtual std::string to_string() const { return ""; } };
struct MedMockPhase : Object {};
struct MedMockRay : MockRay {                      // Ray3f incl. the six-argument constructor spawn_ray uses (ray.h:40-48)
    MockVec3 wavelengths;
    MedMockRay() = default;
    MedMockRay(const MockVec3 &o_, const MockVec3 &d_, float mint_, float maxt_, float time_, const MockVec3 &wl) : MockRay(o_, d_, time_) { mint = mint_; maxt = maxt_; wavelengths = wl; }
};
struct MedMockSI { float t = math::Infinity<float>; };
struct MedMockFrame { MedMockFrame() = default; explicit MedMockFrame(const MockVec3 &) {} };
template <typename MediumT> struct MedMockMI {     // MediumInteraction3f, include/mitsuba/render/interaction.h
    MockVec3 p, wi, wavelengths; float t = math::Infinity<float>, time = 0.f, mint = 0.f;
    MedMockFrame sh_frame; const MediumT *medium = nullptr;
    MockVec3 sigma_s, sigma_n, sigma_t, combined_extinction;
    bool is_valid() const { return t != math::Infinity<float>; }                                  // interaction.h:41-43
    MedMockRay spawn_ray(const MockVec3 &d) const {                                                // interaction.h:58-61
        return MedMockRay(p, d, (1.f + hmax(abs(p))) * math::RayEpsilon<float>, math::Infinity<float>, time, wavelengths);
    }
};
struct MedMockBBox {
    MockVec3 min, max;
    MedMockBBox() = default;
    MedMockBBox(const MockVec3 &a, const MockVec3 &b) : min(a), max(b) {}
    MockVec3 extents() const { return max - min; }                                                 // bbox.h:120
    std::tuple<bool, float, float> ray_intersect(const MockRay &ray) const {                       // bbox.h:302-325
        bool active = all(neq(ray.d, zero<MockVec3>()) || ((ray.o > min) || (ray.o < max)));
        MockVec3 t1 = (min - ray.o) * ray.d_rcp, t2 = (max - ray.o) * ray.d_rcp;
        MockVec3 t1p = enoki::min(t1, t2), t2p = enoki::max(t1, t2);
        float mint = hmax(t1p), maxt = hmin(t2p);
        active = active && maxt >= mint;
        return { active, mint, maxt };
    }
};
struct MedMockProps {                               // Properties: what Medium(const Properties &) asks
    std::string id() const { return "probe"; }
    std::vector<std::pair<std::string, ref<Object>>> objects(bool) const { return {}; }
    void mark_queried(const std::string &) const {}
    bool bool_(const char *, bool def) const { return def; }
    MedMockProps() = default; explicit MedMockProps(const char *) {}
};
struct PluginManager {
    static PluginManager *instance() { static PluginManager pm; return &pm; }
    template <typename T> T *create_object(const MedMockProps &) { static T obj; return &obj; }
};
template <typename... A> [[noreturn]] void Throw(A &&...) { throw std::runtime_error("Throw"); }
using enoki::max; using enoki::min;
inline float max(float a, float b) { return std::max(a, b); }
inline float min(float a, float b) { return std::min(a, b); }
inline uint32_t min(float a, uint32_t b) { return (uint32_t) std::min(a, (float) b); }              // `min(sample * n_channels, n_channels - 1)`, medium.h:126
}  // namespace mitsuba

#undef MTS_IMPORT_TYPES
#define MTS_IMPORT_TYPES(...)                                                                                          \
    using Mask = bool; using UInt32 = uint32_t; using Int32 = int32_t; using ScalarFloat = float;                       \
    using Point3f = mitsuba::MockVec3; using Vector3f = mitsuba::MockVec3;                                              \
    using Ray3f = mitsuba::MedMockRay; using Frame3f = mitsuba::MedMockFrame; using UnpolarizedSpectrum = mitsuba::MockVec3; \
    using SurfaceInteraction3f = mitsuba::MedMockSI; using MediumInteraction3f = mitsuba::MedMockMI<mitsuba::Medium<Float, Spectrum>>; \
    using ScalarBoundingBox3f = mitsuba::MedMockBBox; using Scene = mitsuba::MockScene; using Sampler = mitsuba::MockSampler; \
    using PhaseFunction = mitsuba::MedMockPhase; using Properties = mitsuba::MedMockProps; using Texture = int; using Shape = int;
#define MTS_EXPORT_RENDER
#define MTS_INLINE inline
#define MTS_DECLARE_CLASS()
#define MTS_EXTERN_CLASS_RENDER(...)
#define MTS_VARIANT template <typename Float, typename Spectrum>
#define MTS_MASKED_FUNCTION(...)
#define MTS_IMPLEMENT_CLASS_VARIANT(...)
#define MTS_INSTANTIATE_CLASS(...)
#define ENOKI_PINNED_OPERATOR_NEW(...)
#define ENOKI_MARK_USED(x) (void) x
#define ENOKI_CALL_SUPPORT_TEMPLATE_BEGIN(...)
#define ENOKI_CALL_SUPPORT_METHOD(...)
#define ENOKI_CALL_SUPPORT_TEMPLATE_END(...)
namespace mitsuba { using Properties = MedMockProps; }

#include "render/medium.h"             /* the reference's own header: -I$(REF)/include/mitsuba (its <mitsuba/...> includes resolve to the stand-ins in mock_include) */
#include "medium.cpp"                  /* the reference's own file, -I$(REF)/src/librender */

using namespace mitsuba;
using Base = Medium<float, MockVec3>;

// The grid medium as far as the transmittance sees it (src/media/heterogeneous.cpp:28-56)
struct GridMedium final : Base {
    MedMockBBox aabb;
    GridMedium(float scale, float grid_max, const MockVec3 &lo, const MockVec3 &hi) : Base(MedMockProps()) {
        m_is_homogeneous = false; m_has_spectral_extinction = false;
        m_scale = scale; m_max_density = scale * grid_max; m_inv_max_density = 1.0f / m_max_density;
        aabb.min = lo; aabb.max = hi;
    }
    std::tuple<bool, float, float> intersect_aabb(const MedMockRay &ray) const override { return aabb.ray_intersect(ray); }
    MockVec3 get_combined_extinction(const MediumInteraction3f &, bool) const override { return MockVec3(m_max_density); }
    std::tuple<MockVec3, MockVec3, MockVec3> get_scattering_coefficients(const MediumInteraction3f &, bool) const override {
        return { MockVec3(0.f), MockVec3(0.f), MockVec3(0.f) };               // not read by evalTransmittance
    }
    std::string to_string() const override { return "GridMedium"; }
};

static void p3(const MockVec3 &v) { std::printf("[%.9g, %.9g, %.9g]", v.x(), v.y(), v.z()); }

int main() {
    std::mt19937 gen(20261106u);
    auto uni = [&](float a, float b) { return a + (b - a) * (float) (gen() >> 8) * (1.0f / 16777216.0f); };
    const MockVec3 lo(-10.01f), hi(10.01f);
    std::printf("{\n\"grid_lo\": "); p3(lo); std::printf(", \"grid_hi\": "); p3(hi);
    std::printf(",\n\"cases\": [");
    bool first = true;
    const float majorants[3] = { 0.4f, 0.05f, 2.5f };
    for (int mj = 0; mj < 3; ++mj) {
        GridMedium med(majorants[mj] / 0.8f, 0.8f, lo, hi);                    // scale * grid max = the majorant
        for (int q = 0; q < 80; ++q) {
            // segments inside the box, some starting outside / leaving it, one axis-parallel
            MockVec3 o(uni(-9.f, 9.f), uni(-9.f, 9.f), uni(-9.f, 9.f)), e(uni(-9.f, 9.f), uni(-9.f, 9.f), uni(-9.f, 9.f));
            if (q % 10 == 7) o = MockVec3(uni(-14.f, -11.f), uni(-5.f, 5.f), uni(-5.f, 5.f));
            if (q % 10 == 8) e = MockVec3(uni(11.f, 14.f), uni(-5.f, 5.f), uni(-5.f, 5.f));
            if (q % 10 == 9) e = MockVec3(o.x(), o.y(), -o.z());
            MockVec3 d = e - o; float len = norm(d); d = d / len;
            if (q % 20 == 3) len = uni(0.001f, 0.05f);
            MedMockRay ray(o, d, 0.f, len, 0.f, MockVec3(0.f));
            const uint64_t seed = 0x9e3779b97f4a7c15ull * (uint64_t) (q + 1) + (uint64_t) mj;
            MockSampler sampler;
            sampler.rng.seed(seed, PCG32_DEFAULT_STREAM + 1); sampler.draws = 0;  // the oracle's per-eval sub-stream convention
            MockVec3 tr = med.evalTransmittance(ray, &sampler, true);
            std::printf("%s{\"majorant\": %.9g, \"o\": ", first ? "" : ",\n", med.max_density()); p3(o);
            std::printf(", \"d\": "); p3(d);
            std::printf(", \"maxt\": %.9g, \"seed\": %llu, \"tr\": %.9g, \"draws\": %llu}", len, (unsigned long long) seed, tr.x(), (unsigned long long) sampler.draws);
            first = false;
        }
    }
    std::printf("],\n\"homo\": [");
    {
        GridMedium med(1.f, 0.37f, lo, hi);
        for (int q = 0; q < 12; ++q) {
            float len = uni(0.f, 30.f);
            std::printf("%s[%.9g, %.9g]", q ? ", " : "", len, med.homoEvalTransmittance(len).x());
        }
    }
    std::printf("],\n\"homo_density\": 0.37,\n\"warnings\": %d\n}\n", MockLog::warnings);
    return 0;
}
