// vrl_integrator.h -- C++ host side of the drop-in: the `vrl` integrator plugin surface of
// VincentiusIV/mitsuba-NLVRL (class VRLIntegrator, src/integrators/vrl/vrl.cpp:20-859) over the C ABI of
// include/vrl_b200.h.  Same property names, defaults and error behaviour as the reference:
//   * Properties: typed getters with defaults, every property must be queried (xml.cpp:1050-1068 -> "unreferenced
//     property"), wrong type throws (properties.cpp);
//   * VRLIntegrator(const Properties &)  -- vrl.cpp:32-73 + SamplingIntegrator/Integrator base (integrator.cpp:26-38);
//   * preprocess(scene, sensor) / render(scene, sensor) / cancel() / postprocess() / to_string()
//     (include/mitsuba/render/integrator.h:41-54, 117-151); render() runs preprocess lazily, as Python callers
//     never call it;
//   * film contract: render() fills a full-frame XYZAW fp32 block (integrator.cpp:72-74, 158-174) that the caller
//     hands to Film::put.
// The Mitsuba-facing shim adapts Scene/Sensor/Film objects to HostScene/HostSensor; this file has
// no Mitsuba dependency so that it builds and is tested here, where the MSVC-only fork cannot be compiled.
#pragma once
#include <cstdint>
#include <map>
#include <set>
#include <stdexcept>
#include <string>
#include <variant>
#include <vector>

#include "../include/vrl_b200.h"

namespace vrlhost {

struct Vec3 { float x = 0, y = 0, z = 0; };

// Mirror of mitsuba::Properties for the value kinds the plugin reads (include/mitsuba/core/properties.h)
class Properties {
public:
    using Value = std::variant<bool, int64_t, double, std::string, Vec3>;
    explicit Properties(std::string plugin_name = "vrl") : m_plugin(std::move(plugin_name)) {}
    const std::string &plugin_name() const { return m_plugin; }
    void set_bool(const std::string &k, bool v) { m_values[k] = v; }
    void set_int(const std::string &k, int64_t v) { m_values[k] = v; }
    void set_float(const std::string &k, double v) { m_values[k] = v; }
    void set_string(const std::string &k, const std::string &v) { m_values[k] = v; }
    void set_vector3f(const std::string &k, Vec3 v) { m_values[k] = v; }
    bool has_property(const std::string &k) const { return m_values.count(k) != 0; }

    bool    bool_(const std::string &k, bool def) const;
    int64_t int_(const std::string &k, int64_t def) const;
    double  float_(const std::string &k, double def) const;      // accepts integers too, like Properties::float_
    Vec3    vector3f(const std::string &k, Vec3 def) const;
    // names never queried: the XML loader turns a non-empty list into a hard error (xml.cpp:1050-1068)
    std::vector<std::string> unqueried() const;

private:
    std::string m_plugin;
    std::map<std::string, Value> m_values;
    mutable std::set<std::string> m_queried;
};

// What the shim extracts from mitsuba::Scene (shapes -> triangles + surface words, media, emitters) ...
struct HostScene {
    std::vector<float>        tri_verts;   // 9 floats per triangle
    std::vector<uint32_t>     tri_flags;   // VRLB_SURF_*
    std::vector<float>        tri_reflectance;   // 3 per triangle (diffuse reflectance, read by the light tracer) or empty
    std::vector<vrlb_medium>  media;
    std::vector<vrlb_emitter> emitters;
    int32_t emission_medium = -1;              // medium at the emitters / laser origin (-1 = none)
    float   laser_power[3] = { 1.f, 1.f, 1.f };  // flux of the laser when `use_laser` (the reference takes it from the sampled emitter)
};
// ... and from mitsuba::Sensor / Film / Sampler
struct HostSensor {
    vrlb_camera camera{};
    int32_t  spp = 1;              // sampler sample_count
    uint64_t sampler_seed = 0;     // sampler `seed`
};
// The exported VRL set ("same exported VRL set" of the parity protocol); file format in INTEGRATION.md
struct VrlSet {
    std::vector<float>    soa10;   // 10 columns of n
    std::vector<uint16_t> medium;
    uint64_t n = 0;
    float scale = 1.f;             // VRLMap::m_scale
    void load(const std::string &path);
    void save(const std::string &path) const;
};

class VRLIntegrator {
public:
    explicit VRLIntegrator(const Properties &props);     // vrl.cpp:32-73
    ~VRLIntegrator();
    VRLIntegrator(const VRLIntegrator &) = delete;
    VRLIntegrator &operator=(const VRLIntegrator &) = delete;

    // VRL generation (vrl.cpp:85-441).  With the set supplied by the caller (parity protocol) this only uploads it.
    void set_vrls(const VrlSet &set);
    void preprocess(const HostScene &scene, const HostSensor &sensor);
    // Integrator::render: fills xyzaw (width*height*5 floats, row-major) and returns !stopped (integrator.cpp:177-182)
    bool render(const HostScene &scene, const HostSensor &sensor, std::vector<float> &xyzaw);
    void cancel();                                        // integrator.h:148
    void postprocess() const;                             // vrl.cpp:734-749: logs the counters
    std::string to_string() const;                        // vrl.cpp:790-810 style
    std::vector<std::string> aov_names() const { return {}; }
    [[noreturn]] void sample() const;                     // no CPU path: NotImplemented, as SURVEY 8b prescribes

    // property values, public for tests (names follow vrl.cpp's members)
    int64_t m_globalPhotons, m_causticPhotons, m_volumePhotons, m_targetVRLs, m_maxDepth, m_rrDepth;
    int64_t m_globalLookupSize, m_causticLookupSize, m_samplesPerQuery, m_diceVRL, m_lightcutSamples, m_thresholdBetterDist;
    double  m_volumeLookupRadiusRelative, m_globalLookupRadiusRelative, m_causticLookupRadiusRelative, minVRLLength, m_scaleRR, m_thresholdError;
    bool    m_useLaser, m_useNonLinear, m_useNonLinearCameraRays, m_useNLAtomicQuery, m_useDirectIllum, m_longVRL,
            m_useUniformSampling, m_useLightCut, m_stochasticLightcut, m_RRVRL, m_shootCenter, m_hideEmitters;
    Vec3    laserOrigin, laserDirection;
    int64_t m_blockSize, m_samplesPerPass; double m_timeout;    // base class (integrator.cpp:26-38)
    int64_t m_device;                                            // new: CUDA device index (default 0)
    int64_t m_mathMode;                                          // new: vrlb_options::math_mode (default 1)

    vrlb_timing last_timing() const { return m_timing; }

private:
    vrlb_options make_options(const HostSensor &sensor) const;
    void check(int rc, const char *what) const;
    vrlb_context *m_ctx = nullptr;
    bool m_preprocessed = false, m_have_vrls = false, m_lightcutBuilt = false;
    void build_lightcut(uint64_t base_seed);
    vrlb_timing m_timing{};
};

}  // namespace vrlhost
