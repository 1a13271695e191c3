// ref_probe_nl.cpp -- TEST INFRASTRUCTURE.  Compiles the reference's OWN non-linear medium -- src/media/nonlinear.cpp,
// where it lies under /root/reference -- against a mock Mitsuba (ref_mock.h + the stand-ins below) and the reference's
// vendored Enoki, and prints, for seeded rays through the media of the shipped fixtures, what the reference's cell walk
// does to them:
//     NonLinearMedia::build                        nonlinear.cpp:62-89    (cell grid, one IOR per cell)
//     NonLinearMedia::calculateIoR                 nonlinear.cpp:129-153  (incl. from_bottom = false, mirage.xml)
//     getNode / getNeighbour / getNormal           nonlinear.cpp:155-227
//     reflect / refract                            nonlinear.cpp:229-242
//     sampleNonLinearInteraction                   nonlinear.cpp:297-405  (Snell at the cell face, TIR -> reflect)
//     handleNonLinearInteraction                   nonlinear.cpp:244-295
// driven the way VRLIntegrator::sample drives them for a camera ray (vrl.cpp:550-577: sample, then while
// nli.t < si.t: handle, sample again).  Output: JSON -> tests/golden/ref_nonlinear.json (oracle/make_golden.py);
// tests hold the oracle's and the CUDA kernels' fixed-step eikonal marcher to these walks (exit point, exit
// direction, path length; convergence with the cell count).
//
// What is mocked (written here, NOT reference code): the Medium base class members nonlinear.cpp uses (medium.h:18-35
// NonLinearInteraction, :165-170 build, :210-218 members), Properties, a constant Volume, the bounding box functions of
// include/mitsuba/core/bbox.h the file calls (contains :169-176, center, extents, ray_intersect :302-325), and a scene
// that is the medium's container box.  Shims for the fork's MSVC dialect (g++ rejects them; none touches the code
// under test): unqualified base-class members (imported by MTS_IMPORT_BASE here), `Medium::build` naming the base
// template without arguments (a class-scope alias), `typedef NLNode<...> NLNode` (-fpermissive), unscoped LogLevel
// names, int / float mixes in max().
#include <cassert>
#include <cstdio>
#include <random>
#include "ref_mock.h"

namespace mitsuba {
template <typename T> struct ref {
    T *p = nullptr;
    ref() = default; ref(T *q) : p(q) {}
    T *get() const { return p; } T *operator->() const { return p; } T &operator*() const { return *p; }
};
struct NLMockMI {                                   // MediumInteraction3f, the members nonlinear.cpp touches
    MockVec3 p, wi; float t = 0.f; MockFrame sh_frame;
    MockVec3 sigma_s, sigma_n, sigma_t, combined_extinction;
};
struct NLMockVolume {                               // ConstVolume (src/textures/constant3d.cpp): eval = value, max = mean
    MockVec3 v;
    float max() const { return hmean(v); }
    MockVec3 eval(const NLMockMI &, bool = true) const { return v; }
};
template <typename T> std::ostream &operator<<(std::ostream &os, const ref<T> &) { return os << "ref"; }
struct NLMockProps {                                // Properties: every query answers with the scene file's value
    NLMockVolume sigma_t, albedo; float scale, top_ior, bottom_ior; int ior_method, res[3]; bool from_bottom;
    template <typename V> ref<V> volume(const char *name, float) const { return ref<V>(const_cast<V *>(std::string(name) == "albedo" ? &albedo : &sigma_t)); }
    float float_(const char *name, float def) const { std::string n(name); return n == "scale" ? scale : n == "top_ior" ? top_ior : n == "bottom_ior" ? bottom_ior : def; }
    int int_(const char *name, double def) const { std::string n(name); return n == "ior_method" ? ior_method : n == "res_x" ? res[0] : n == "res_y" ? res[1] : n == "res_z" ? res[2] : (int) def; }
    MockVec3 vector3f(const char *, const MockVec3 &def) const { return def; }
    bool bool_(const char *name, bool def) const { return std::string(name) == "from_bottom" ? from_bottom : def; }
};
struct NLMockBBox {                                 // BoundingBox<Point3f>, include/mitsuba/core/bbox.h
    MockVec3 min, max;
    NLMockBBox() : min(std::numeric_limits<float>::infinity()), max(-std::numeric_limits<float>::infinity()) {}
    NLMockBBox(const MockVec3 &a, const MockVec3 &b) : min(a), max(b) {}
    MockVec3 center() const { return (max + min) * 0.5f; }                        // bbox.h:112-114
    MockVec3 extents() const { return max - min; }                                // bbox.h:120
    bool contains(const MockVec3 &p) const { return all(p >= min) && all(p <= max); }   // bbox.h:169-176 (Strict = false)
    std::tuple<bool, float, float> ray_intersect(const MockRay &ray) const {      // bbox.h:302-325
        bool active = all(neq(ray.d, zero<MockVec3>()) || ((ray.o > min) || (ray.o < max)));
        MockVec3 t1 = (min - ray.o) * ray.d_rcp, t2 = (max - ray.o) * ray.d_rcp;
        MockVec3 t1p = enoki::min(t1, t2), t2p = enoki::max(t1, t2);
        float mint = hmax(t1p), maxt = hmin(t2p);
        active = active && maxt >= mint;
        return { active, mint, maxt };
    }
};
inline std::ostream &operator<<(std::ostream &os, const NLMockBBox &b) { return os << "BBox[" << b.min << b.max << "]"; }
struct NLMockRay : MockRay {                        // Ray3f with update() (ray.h:50-53) and mint = 0 spawn semantics
    using MockRay::MockRay;
    void update() { d_rcp = rcp(d); }
};
struct NLMockFrame : MockFrame {
    using MockFrame::MockFrame;
    static float cos_theta(const MockVec3 &v) { return v.z(); }                   // frame.h
};
struct NLMockScene {                                // the container box, seen from inside: closest face along the ray
    NLMockBBox box;
    MockSI ray_intersect(const NLMockRay &ray, bool = true) const {
        MockSI si;
        float best = ray.maxt;
        for (int k = 0; k < 3; ++k)
            for (int s = 0; s < 2; ++s) {
                if (ray.d[k] == 0.f) continue;
                float t = ((s ? box.max[k] : box.min[k]) - ray.o[k]) / ray.d[k];
                if (!(t >= ray.mint && t <= best)) continue;
                MockVec3 p = ray.o + ray.d * t;
                bool in = true;
                for (int j = 0; j < 3; ++j) if (j != k && !(p[j] >= box.min[j] - 1e-3f && p[j] <= box.max[j] + 1e-3f)) in = false;
                if (in) { best = t; si.t = t; si.p = p; }
            }
        return si;
    }
};
struct TraversalCallback { template <typename... A> void put_parameter(A &&...) {} template <typename... A> void put_object(A &&...) {} };

template <typename Float, typename Spectrum> class Medium {                      // the parts of include/mitsuba/render/medium.h in use
public:
    struct NonLinearInteraction {                                                // medium.h:18-35
        bool is_valid; MockVec3 wi, wo, n; float n1, n2; MockVec3 p; float t; float eta;
        inline NonLinearInteraction() { is_valid = false; n1 = n2 = 1.0f; t = math::Infinity<Float>; }
    };
    Medium(const NLMockProps &) {}
    virtual ~Medium() {}
    virtual void build(MockVec3 min, MockVec3 max) {                             // medium.h:165-170
        bbox = NLMockBBox(min, max);
        width = bbox.extents().x(); height = bbox.extents().y(); depth = bbox.extents().z();
    }
    virtual void traverse(TraversalCallback *) {}
    virtual std::string to_string() const = 0;
    virtual std::tuple<bool, float, float> intersect_aabb(const NLMockRay &) const = 0;
    virtual MockVec3 get_combined_extinction(const NLMockMI &, bool) const = 0;
    virtual std::tuple<MockVec3, MockVec3, MockVec3> get_scattering_coefficients(const NLMockMI &, bool) const = 0;
    virtual NonLinearInteraction sampleNonLinearInteraction(const NLMockRay &, uint32_t, bool) const { return NonLinearInteraction(); }
    virtual bool handleNonLinearInteraction(const NLMockScene *, MockSampler *, NonLinearInteraction &, MockSI &, NLMockMI &, NLMockRay &, MockVec3 &, uint32_t, bool) const { return false; }
protected:
    bool m_sample_emitters = false, m_is_homogeneous = false, m_is_nonlinear = false, m_has_spectral_extinction = false;
    float m_inv_max_density = 0.f, m_max_density = 0.f, m_scale = 1.f;
    NLMockBBox bbox;
    float width = 0.f, height = 0.f, depth = 0.f;
};
using enoki::max; using enoki::min;
inline float max(int a, float b) { return std::max((float) a, b); }              // `max(0, position.y() / ...)`, nonlinear.cpp:150
inline float max(double a, float b) { return std::max((float) a, b); }
inline float min(double a, float b) { return std::min((float) a, b); }
}  // namespace mitsuba

#undef MTS_IMPORT_TYPES
#undef MTS_IMPORT_OBJECT_TYPES
#define MTS_IMPORT_TYPES(...)                                                                                          \
    using Mask = bool; using UInt32 = uint32_t; using Int32 = int32_t;                                                 \
    using Point3f = mitsuba::MockVec3; using Vector3f = mitsuba::MockVec3; using Normal3f = mitsuba::MockVec3;         \
    using Ray3f = mitsuba::NLMockRay; using Frame3f = mitsuba::NLMockFrame; using UnpolarizedSpectrum = mitsuba::MockVec3; \
    using SurfaceInteraction3f = mitsuba::MockSI; using MediumInteraction3f = mitsuba::NLMockMI;                        \
    using ScalarBoundingBox3f = mitsuba::NLMockBBox; using Scene = mitsuba::NLMockScene; using Sampler = mitsuba::MockSampler; \
    using Volume = mitsuba::NLMockVolume; using Properties = mitsuba::NLMockProps;
#define MTS_IMPORT_OBJECT_TYPES()
// every base member nonlinear.cpp names without qualification, and `Medium::build` (the base, as MSVC resolves it)
#define MTS_IMPORT_BASE(Name, ...)                                                                                     \
    using Base = mitsuba::Name<Float, Spectrum>; using Medium = Base;                                                   \
    using typename Base::NonLinearInteraction;                                                                         \
    using Base::m_is_homogeneous; using Base::m_has_spectral_extinction; using Base::m_is_nonlinear; using Base::m_scale; \
    using Base::m_max_density; using Base::m_inv_max_density; using Base::bbox; using Base::width; using Base::height; using Base::depth;
#define MTS_MASKED_FUNCTION(...)
#define MTS_DECLARE_CLASS()
#define MTS_IMPLEMENT_CLASS_VARIANT(...)
#define MTS_EXPORT_PLUGIN(...)
#define MTS_INLINE inline
using mitsuba::NLMockProps;
namespace mitsuba { using Properties = NLMockProps; }

#include "nonlinear.cpp"   /* the reference's own file, -I$(REF)/src/media */

using namespace mitsuba;
using NLM = NonLinearMedia<float, MockVec3>;

static void p3(const MockVec3 &v) { std::printf("[%.9g, %.9g, %.9g]", v.x(), v.y(), v.z()); }

// stop: why the walk ended -- 0 the next cell face lies behind the container wall (the regular end), 1 sampleNonLinearInteraction
// returned an invalid interaction inside the container (the walk STALLS: see "stall" in DESIGN.md), 2 handleNonLinearInteraction
// refused, 3 event cap
// stall_layers: IOR layers the final straight stretch (from the last event to the container wall) passes through without
// being refracted -- 0 for a complete walk; > 0 when the walk gave up inside the medium
struct Walk { MockVec3 exit_p, exit_d, last_p; int crossings = 0, tir = 0, stop = 0, stall_layers = 0; float length = 0.f; std::vector<MockVec3> pts; };

// VRLIntegrator::sample, vrl.cpp:550-577, for a ray that starts inside the medium's container
static Walk walk(const NLM &med, const NLMockScene &scene, MockVec3 o, MockVec3 d, int max_events, bool keep_pts) {
    Walk w;
    NLMockRay ray(o, d, 0.f);
    ray.mint = 0.f;
    MockSampler sampler; MockSI si; NLMockMI mi; MockVec3 throughput(1.f);
    mi.t = math::Infinity<float>;
    si = scene.ray_intersect(ray);
    auto nli = med.sampleNonLinearInteraction(ray, 0u, true);
    while (nli.t < si.t && si.is_valid() && nli.is_valid && w.crossings < max_events) {
        const MockVec3 before = ray.d;
        const float seg = nli.t;
        bool valid = med.handleNonLinearInteraction(&scene, &sampler, nli, si, mi, ray, throughput, 0u, true);
        if (!valid) { w.stop = 2; break; }
        w.length += seg;
        ++w.crossings;
        if (before.y() * ray.d.y() < 0.f) ++w.tir;                               // the vertical component flipped: reflect(), nonlinear.cpp:367-372
        if (keep_pts) w.pts.push_back(ray.o);
        mi.t = math::Infinity<float>;
        nli = med.sampleNonLinearInteraction(ray, 0u, true);
    }
    if (w.stop == 0) w.stop = !nli.is_valid ? 1 : (w.crossings >= max_events ? 3 : 0);
    w.length += si.is_valid() ? si.t : 0.f;
    w.exit_p = si.is_valid() ? MockVec3(ray.o + ray.d * si.t) : ray.o;
    w.exit_d = ray.d;
    w.last_p = ray.o;
    return w;
}

struct MediumCase { const char *name; float half[3]; int res_y; float bottom, top; bool from_bottom; };

int main() {
    // the media of the shipped fixtures (container = unit cube scaled by 2 * half) and coarser / finer grids of the same
    const MediumCase cases[] = {
        { "cbox-nonlinear-homo.xml", { 400.f, 150.f, 400.f }, 640, 1.0f, 0.95f, true },
        { "cbox-nonlinear-homo.xml res_y 160", { 400.f, 150.f, 400.f }, 160, 1.0f, 0.95f, true },
        { "cbox-nonlinear-homo.xml res_y 2560", { 400.f, 150.f, 400.f }, 2560, 1.0f, 0.95f, true },
        { "mirage.xml", { 500.f, 250.f, 500.f }, 2400, 1.03f, 1.0f, false },
        { "mirage.xml res_y 600", { 500.f, 250.f, 500.f }, 600, 1.03f, 1.0f, false },
    };
    std::mt19937 gen(20261105u);
    auto uni = [&](float a, float b) { return a + (b - a) * (float) (gen() >> 8) * (1.0f / 16777216.0f); };
    std::printf("{\n\"media\": [");
    bool first_case = true;
    for (const MediumCase &c : cases) {
        NLMockProps props;
        props.sigma_t.v = MockVec3(0.05f); props.albedo.v = MockVec3(0.25f); props.scale = 0.05f;
        props.top_ior = c.top; props.bottom_ior = c.bottom; props.ior_method = 0;
        props.res[0] = 1; props.res[1] = c.res_y; props.res[2] = 1; props.from_bottom = c.from_bottom;
        NLM med(props);
        const MockVec3 lo(-c.half[0], -c.half[1], -c.half[2]), hi(c.half[0], c.half[1], c.half[2]);
        med.build(lo, hi);
        NLMockScene scene; scene.box = NLMockBBox(lo, hi);
        std::printf("%s{\"name\": \"%s\", \"lo\": ", first_case ? "" : ",\n", c.name); p3(lo);
        std::printf(", \"hi\": "); p3(hi);
        std::printf(", \"res_y\": %d, \"bottom_ior\": %.9g, \"top_ior\": %.9g, \"from_bottom\": %d,\n", c.res_y, c.bottom, c.top, c.from_bottom ? 1 : 0);
        // the IOR profile itself: calculateIoR at cell centres (what build() stores)
        std::printf("\"ior\": [");
        for (int k = 0; k < 16; ++k) {
            float y = lo.y() + (hi.y() - lo.y()) * (k + 0.5f) / 16.f;
            std::printf("%s[%.9g, %.9g]", k ? ", " : "", y, med.calculateIoR(MockVec3(0.f, y, 0.f)));
        }
        std::printf("],\n\"rays\": [");
        // the same seeded rays for every grid of a medium (re-seed per medium family)
        gen.seed(c.from_bottom ? 101u : 202u);
        for (int r = 0; r < 48; ++r) {
            // start inside the container (off the cell faces), mostly shallow directions: those are the ones that bend
            MockVec3 o(uni(-0.8f, 0.8f) * c.half[0], uni(-0.9f, 0.9f) * c.half[1], -c.half[2] + 0.37f);
            float elev = (r % 4 == 0 ? uni(-0.45f, 0.45f) : uni(-0.12f, 0.12f));          // radians from the horizontal
            float azim = uni(-0.3f, 0.3f);
            MockVec3 d(std::sin(azim) * std::cos(elev), std::sin(elev), std::cos(azim) * std::cos(elev));
            d = normalize(d);
            Walk w = walk(med, scene, o, d, 200000, r < 2 && c.res_y <= 640);
            {
                const float cell = (hi.y() - lo.y()) / (float) c.res_y;
                const int a = (int) std::floor((w.last_p.y() - lo.y()) / cell), b = (int) std::floor((w.exit_p.y() - lo.y()) / cell);
                w.stall_layers = std::abs(std::min(std::max(b, 0), c.res_y - 1) - std::min(std::max(a, 0), c.res_y - 1));
            }
            std::printf("%s{\"o\": ", r ? ",\n" : ""); p3(o); std::printf(", \"d\": "); p3(d);
            std::printf(", \"exit_p\": "); p3(w.exit_p); std::printf(", \"exit_d\": "); p3(w.exit_d);
            std::printf(", \"crossings\": %d, \"tir\": %d, \"stop\": %d, \"stall_layers\": %d, \"length\": %.9g", w.crossings, w.tir, w.stop, w.stall_layers, w.length);
            if (!w.pts.empty()) {
                std::printf(", \"pts\": [");
                for (size_t k = 0; k < w.pts.size(); ++k) { std::printf("%s", k ? ", " : ""); p3(w.pts[k]); }
                std::printf("]");
            }
            std::printf("}");
        }
        std::printf("]}");
        first_case = false;
    }
    // refract / reflect on their own (nonlinear.cpp:229-242): unit normal (0, 1, 0), several incidence angles and index pairs
    std::printf("],\n\"refract\": [");
    {
        NLMockProps props; props.sigma_t.v = MockVec3(1.f); props.albedo.v = MockVec3(0.5f); props.scale = 1.f;
        props.top_ior = 1.f; props.bottom_ior = 1.f; props.ior_method = 0; props.res[0] = props.res[1] = props.res[2] = 1; props.from_bottom = true;
        NLM med(props);
        const float pairs[4][2] = { { 1.0f, 0.95f }, { 0.95f, 1.0f }, { 1.03f, 1.0f }, { 1.0002f, 1.0f } };
        bool first = true;
        for (int k = 0; k < 4; ++k)
            for (int a = 0; a < 12; ++a) {
                float ang = 0.05f + 1.5f * a / 12.f;                                    // from the normal
                MockVec3 wi(std::sin(ang), -std::cos(ang), 0.f), n(0.f, 1.f, 0.f);
                MockVec3 wo = med.refract(wi, n, pairs[k][0], pairs[k][1]);
                MockVec3 rf = med.reflect(wi, n);
                std::printf("%s{\"wi\": ", first ? "" : ",\n"); p3(wi);
                std::printf(", \"n1\": %.9g, \"n2\": %.9g, \"wo\": ", pairs[k][0], pairs[k][1]); p3(wo);
                std::printf(", \"reflect\": "); p3(rf); std::printf("}");
                first = false;
            }
    }
    std::printf("],\n\"warnings\": %d\n}\n", MockLog::warnings);
    return 0;
}
