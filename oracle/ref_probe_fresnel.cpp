// ref_probe_fresnel.cpp -- TEST INFRASTRUCTURE.  Compiles the reference's OWN include/mitsuba/render/fresnel.h (where it
// lies under /root/reference) on top of the reference's vendored Enoki and prints inputs + outputs of fresnel(), reflect(),
// refract() (local-frame forms, the ones SmoothDielectric::sample / SmoothConductor::sample call) and fresnel_conductor()
// as JSON -> tests/golden/ref_fresnel.json (oracle/make_golden.py); tests/test_ref_golden_cpu.py holds the oracle's
// sample_delta to it.  <mitsuba/core/vector.h> resolves to the empty stand-in of oracle/mock_include; Vector / Normal are
// the Enoki arrays the real header aliases them to.
#include <cstdio>
#include <random>
#include <tuple>
#include <enoki/array.h>
#include <enoki/complex.h>
#define NAMESPACE_BEGIN(name) namespace name {
#define NAMESPACE_END(name) }
namespace mitsuba {
using namespace enoki;
template <typename F, size_t N> using Vector = enoki::Array<F, N>;
template <typename F, size_t N> using Normal = enoki::Array<F, N>;
}
#include <mitsuba/render/fresnel.h>

using namespace mitsuba;
static std::mt19937 gen(77u);
static float uni(float a, float b) { return a + (b - a) * (float) (gen() >> 8) * (1.0f / 16777216.0f); }

int main() {
    const float etas[5] = { 1.5046f / 1.000277f, 1.333f, 1.0f / 1.5f, 1.0f, 2.419f };
    std::printf("{\"dielectric\": [");
    for (int i = 0; i < 300; ++i) {
        const float eta = etas[i % 5];
        float cz = i < 10 ? (i % 2 ? 1.f : -1.f) : uni(-1.f, 1.f);
        if (i == 10) cz = 0.f;
        const float phi = uni(0.f, 6.2831853f), sn = std::sqrt(std::max(0.f, 1.f - cz * cz));
        Vector<float, 3> wi(sn * std::cos(phi), sn * std::sin(phi), cz);
        auto [r, cos_theta_t, eta_it, eta_ti] = fresnel<float>(wi.z(), eta);
        auto wr = reflect<float>(wi), wt = refract<float>(wi, cos_theta_t, eta_ti);
        std::printf("%s\n {\"eta\": %.9g, \"wi\": [%.9g, %.9g, %.9g], \"r\": %.9g, \"cos_theta_t\": %.9g, \"eta_it\": %.9g, \"eta_ti\": %.9g, "
                    "\"reflect\": [%.9g, %.9g, %.9g], \"refract\": [%.9g, %.9g, %.9g]}", i ? "," : "", eta, wi.x(), wi.y(), wi.z(), r, cos_theta_t, eta_it, eta_ti,
                    wr.x(), wr.y(), wr.z(), wt.x(), wt.y(), wt.z());
    }
    std::printf("],\n\"conductor_default\": [");      // SmoothConductor's default eta = 0, k = 1 (conductor.cpp): the Fresnel term the descriptor takes as 1
    for (int i = 0; i < 20; ++i) {
        const float c = (i + 1) / 20.f;
        std::printf("%s[%.9g, %.9g]", i ? ", " : "", c, fresnel_conductor<float>(c, Complex<float>(0.f, 1.f)));
    }
    std::printf("]}\n");
    return 0;
}
