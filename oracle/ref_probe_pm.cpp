// ref_probe_pm.cpp -- TEST INFRASTRUCTURE.  Compiles the reference's OWN photon map -- src/integrators/photonmapper/
// photonmap.h + kdtree.h, where they lie under /root/reference -- against the mock Mitsuba of ref_mock.h and the
// reference's vendored Enoki, and prints inputs + outputs of
//     PhotonMap::estimateRadiance          photonmap.h:131-173   (over PointKDTree::build / nnSearch, kdtree.h:266, 319-400)
//     PhotonMap::estimateCausticRadiance   photonmap.h:175-211
//     PhotonMap::estimateRadianceVolume    photonmap.h:216-271   (over PointKDTree::executeQuery, kdtree.h:678-745)
// as JSON; oracle/make_golden.py stores it as tests/golden/ref_photonmap.json and tests/test_ref_golden_cpu.py holds the
// oracle's restatements to it.  Shims for the fork's MSVC dialect (g++ rejects these constructs even in templates that
// are never instantiated), none of which touches the code under test:
//   * `typedef typename Point3f PointType;` (kdtree.h:148,151)  -> Point3f / Vector3f are macros for a qualified name
//     (REF_MOCK_MSVC_TYPENAME in ref_mock.h)
//   * `pos` undeclared in the unused beamQuery (kdtree.h:420)    -> `#define pos p` over both headers (renames the
//     RadianceQueryVolume member consistently)
//   * three-argument max / min in the same function              -> overloads below
//   * `: SimpleKDNode(photon)` naming a dependent base (photonmap.h:53) -> function-like macro spelling the base out
// The k of the nearest-neighbour estimates is even on purpose: estimateRadiance allocates k SearchResults where nnSearch
// needs k + 1 (photonmap.h:137 vs kdtree.h:316); with an even k the extra 8 bytes fall into malloc's slack.
#include <alloca.h>
#include <algorithm>
#include <atomic>
#include <cstdio>
#include <random>
#include <sstream>
#define REF_MOCK_MSVC_TYPENAME
#include "ref_mock.h"
namespace mitsuba {
template <typename Point> struct BoundingBox {                       // include/mitsuba/core/bbox.h, the members kdtree.h uses
    Point min, max;
    BoundingBox() { reset(); }
    void reset() { min = Point(std::numeric_limits<float>::infinity()); max = Point(-std::numeric_limits<float>::infinity()); }
    void expand(const Point &p) { min = enoki::min(min, p); max = enoki::max(max, p); }
    int major_axis() const { Point d = max - min; int a = 0; for (int i = 1; i < 3; ++i) if (d[i] > d[a]) a = i; return a; }
    float volume() const { return hprod(max - min); }
};
using enoki::max; using enoki::min;                                  // keep the two-argument forms visible next to the overloads
inline float max(float a, float b, float c) { return std::max(a, std::max(b, c)); }
inline float min(float a, float b, float c) { return std::min(a, std::min(b, c)); }
struct Timer { void reset() {} size_t value() const { return 0; } };
}  // namespace mitsuba
#define Assert(x) do { if (!(x)) { std::fprintf(stderr, "Assert(%s) failed\n", #x); std::abort(); } } while (0)
#define pos p
#include "kdtree.h"
#define SimpleKDNode(x) SimpleKDNode<Point, ::mitsuba::PhotonData<Float, Spectrum>>(x)
#include "photonmap.h"
#undef SimpleKDNode
#undef pos

using namespace mitsuba;
using PM = PhotonMap<float, MockVec3>;

static std::mt19937 gen(20260519u);
static float uni(float a = 0.f, float b = 1.f) { return a + (b - a) * (float) (gen() >> 8) * (1.0f / 16777216.0f); }
static MockVec3 unit() { for (;;) { MockVec3 v(uni(-1, 1), uni(-1, 1), uni(-1, 1)); float l = norm(v); if (l > 0.1f && l <= 1.f) return v / l; } }
static void p3(const MockVec3 &v) { std::printf("[%.9g, %.9g, %.9g]", v.x(), v.y(), v.z()); }

int main() {
    // ---- surface maps: photons in a thin slab around the plane z = 0 (so the kd-tree splits along all axes) ----
    const int N = 1500;
    std::vector<MockVec3> ppos(N), pdir(N), ppow(N);
    PM global(N), caustic(N);
    for (int i = 0; i < N; ++i) {
        ppos[i] = MockVec3(uni(-3, 3), uni(-3, 3), uni(-0.05f, 0.05f));
        pdir[i] = unit();
        ppow[i] = MockVec3(uni(0.5f, 1.5f), uni(0.4f, 1.2f), uni(0.3f, 0.9f));
        global.insert(ppos[i], PM::PhotonData(MockVec3(0, 0, 1), pdir[i], ppow[i], 1 + i % 7));
        if (i % 3 == 0) caustic.insert(ppos[i], PM::PhotonData(MockVec3(0, 0, 1), pdir[i], ppow[i], 1));
    }
    global.setScaleFactor(1.0f / 811); caustic.setScaleFactor(1.0f / 377);
    global.build(); caustic.build();
    std::printf("{\n\"surface_photons\": [");
    for (int i = 0; i < N; ++i) { std::printf("%s[", i ? ", " : ""); p3(ppos[i]); std::printf(", "); p3(pdir[i]); std::printf(", "); p3(ppow[i]); std::printf("]"); }
    std::printf("],\n\"global_scale\": %.9g, \"caustic_scale\": %.9g, \"caustic_every\": 3,\n\"surface\": [", 1.0f / 811, 1.0f / 377);
    const float radii[3] = { 0.6f, 0.25f, 1.2f }; const int ks[3] = { 20, 120, 2 };        // ~47 / 8 / 188 photons in range: shrunk, not shrunk, shrunk to the nearest two
    bool first = true;
    for (int c = 0; c < 3; ++c)
        for (int q = 0; q < 40; ++q) {
            MockBSDF bsdf; bsdf.diffuse = true; bsdf.reflectance = MockVec3(uni(0.2f, 0.9f), uni(0.2f, 0.9f), uni(0.2f, 0.9f));
            MockSI si; si.m_bsdf = &bsdf;
            si.p = MockVec3(uni(-2.4f, 2.4f), uni(-2.4f, 2.4f), uni(-0.02f, 0.02f));
            MockVec3 n = q % 4 == 0 ? unit() : MockVec3(0, 0, q % 8 < 4 ? 1.f : -1.f);
            si.n = n; si.sh_frame = MockFrame(n);
            MockVec3 wi_world = unit();
            si.wi = si.to_local(wi_world);
            MockVec3 g = global.estimateRadiance(si, radii[c], (size_t) ks[c]);
            MockVec3 ca = caustic.estimateCausticRadiance(si, 0.8f * radii[c], (size_t) ks[c]);
            std::printf("%s\n {\"radius\": %.9g, \"k\": %d, \"p\": ", first ? "" : ",", radii[c], ks[c]); first = false;
            p3(si.p); std::printf(", \"n\": "); p3(n); std::printf(", \"wi\": "); p3(wi_world); std::printf(", \"reflectance\": "); p3(bsdf.reflectance);
            std::printf(", \"global\": "); p3(g); std::printf(", \"caustic\": "); p3(ca); std::printf("}");
        }
    // ---- volume map ----
    const int NV = 1500;
    PM vol(NV);
    std::vector<MockVec3> vpos(NV), vdir(NV), vpow(NV);
    for (int i = 0; i < NV; ++i) {
        vpos[i] = MockVec3(uni(-3, 3), uni(-3, 3), uni(-3, 3)); vdir[i] = unit(); vpow[i] = MockVec3(uni(0.5f, 1.5f), uni(0.4f, 1.2f), uni(0.3f, 0.9f));
        vol.insert(vpos[i], PM::PhotonData(MockVec3(0.f), vdir[i], vpow[i], 1));
    }
    vol.build();
    std::printf("],\n\"volume_photons\": [");
    for (int i = 0; i < NV; ++i) { std::printf("%s[", i ? ", " : ""); p3(vpos[i]); std::printf(", "); p3(vdir[i]); std::printf(", "); p3(vpow[i]); std::printf("]"); }
    std::printf("],\n\"hg_g\": 0.6,\n\"volume\": [");
    MockMedium iso, hg; hg.phase.hg = true; hg.phase.g = 0.6f;
    MockSampler sampler;
    for (int q = 0; q < 60; ++q) {
        MockVec3 gp(uni(-2.5f, 2.5f), uni(-2.5f, 2.5f), uni(-2.5f, 2.5f)), wo = unit();
        const float r = q % 3 == 0 ? 0.9f : 0.45f;
        size_t M = 0;
        MockVec3 e = vol.estimateRadianceVolume(gp, wo, q % 2 ? &hg : &iso, &sampler, r, M);
        std::printf("%s\n {\"radius\": %.9g, \"hg\": %d, \"p\": ", q ? "," : "", r, q % 2); p3(gp); std::printf(", \"wo\": "); p3(wo);
        std::printf(", \"M\": %zu, \"out\": ", M); p3(e); std::printf("}");
    }
    std::printf("]\n}\n");
    return 0;
}
