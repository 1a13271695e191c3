// vrl_integrator.cpp -- see vrl_integrator.h.  Host logic only; all computation happens in libvrl_b200.so.
#include "vrl_integrator.h"
#include <algorithm>
#include <cmath>

#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>

namespace vrlhost {

namespace {
[[noreturn]] void Throw(const std::string &msg) { throw std::runtime_error(msg); }   // Throw(...) of mitsuba/core/logger.h
}

// ---- Properties ---------------------------------------------------------------------------------------------
bool Properties::bool_(const std::string &k, bool def) const {
    auto it = m_values.find(k);
    if (it == m_values.end()) return def;
    m_queried.insert(k);
    if (auto p = std::get_if<bool>(&it->second)) return *p;
    Throw("The property \"" + k + "\" has the wrong type (expected <boolean>).");
}
int64_t Properties::int_(const std::string &k, int64_t def) const {
    auto it = m_values.find(k);
    if (it == m_values.end()) return def;
    m_queried.insert(k);
    if (auto p = std::get_if<int64_t>(&it->second)) return *p;
    Throw("The property \"" + k + "\" has the wrong type (expected <integer>).");
}
double Properties::float_(const std::string &k, double def) const {
    auto it = m_values.find(k);
    if (it == m_values.end()) return def;
    m_queried.insert(k);
    if (auto p = std::get_if<double>(&it->second)) return *p;
    if (auto p = std::get_if<int64_t>(&it->second)) return (double) *p;
    Throw("The property \"" + k + "\" has the wrong type (expected <float>).");
}
Vec3 Properties::vector3f(const std::string &k, Vec3 def) const {
    auto it = m_values.find(k);
    if (it == m_values.end()) return def;
    m_queried.insert(k);
    if (auto p = std::get_if<Vec3>(&it->second)) return *p;
    Throw("The property \"" + k + "\" has the wrong type (expected <vector>).");
}
std::vector<std::string> Properties::unqueried() const {
    std::vector<std::string> out;
    for (auto &kv : m_values) if (!m_queried.count(kv.first)) out.push_back(kv.first);
    return out;
}

// ---- VRL set file: "VRLB" | u32 version | u64 n | f32 scale | u32 pad | 10*n f32 (SoA) | n u16 -----------------
void VrlSet::save(const std::string &path) const {
    std::ofstream f(path, std::ios::binary);
    if (!f) Throw("cannot write " + path);
    uint32_t ver = 1, pad = 0;
    f.write("VRLB", 4); f.write((const char *) &ver, 4); f.write((const char *) &n, 8);
    f.write((const char *) &scale, 4); f.write((const char *) &pad, 4);
    f.write((const char *) soa10.data(), (std::streamsize) (soa10.size() * 4));
    std::vector<uint16_t> m = medium; m.resize(n, 0);
    f.write((const char *) m.data(), (std::streamsize) (n * 2));
}
void VrlSet::load(const std::string &path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) Throw("cannot read " + path);
    char magic[4]; uint32_t ver = 0, pad = 0;
    f.read(magic, 4); f.read((char *) &ver, 4); f.read((char *) &n, 8); f.read((char *) &scale, 4); f.read((char *) &pad, 4);
    if (!f || std::memcmp(magic, "VRLB", 4) != 0 || ver != 1) Throw(path + ": not a VRLB version 1 file");
    soa10.resize((size_t) n * 10); medium.resize((size_t) n);
    f.read((char *) soa10.data(), (std::streamsize) (soa10.size() * 4));
    f.read((char *) medium.data(), (std::streamsize) (n * 2));
    if (!f) Throw(path + ": truncated VRLB file");
}

// ---- VRLIntegrator ------------------------------------------------------------------------------------------
VRLIntegrator::VRLIntegrator(const Properties &props) {
    // Integrator / SamplingIntegrator base, integrator.cpp:26-38
    m_blockSize      = props.int_("block_size", 0);
    if (m_blockSize < 0) Throw("\"block_size\" must be >= 0");                       // integrator.cpp:28-29
    m_samplesPerPass = props.int_("samples_per_pass", -1);
    m_timeout        = props.float_("timeout", -1.0);
    m_hideEmitters   = props.bool_("hide_emitters", false);
    // vrl.cpp:33-73 (the later global/caustic defaults win)
    m_globalPhotons  = props.int_("global_photons", 250000);
    m_causticPhotons = props.int_("caustic_photons", 250000);
    m_volumePhotons  = props.int_("volume_photons", 1000);
    m_maxDepth       = props.int_("max_depth", 512);
    m_targetVRLs     = props.int_("target_vrls", 1000);
    m_volumeLookupRadiusRelative  = props.float_("volume_lookup_radius_relative", 0.005);
    m_globalLookupRadiusRelative  = props.float_("global_lookup_radius_relative", 0.05);
    m_causticLookupRadiusRelative = props.float_("caustic_lookup_radius_relative", 0.0125);
    m_globalLookupSize  = props.int_("global_lookup_size", 120);
    m_causticLookupSize = props.int_("caustic_lookup_size", 120);
    m_samplesPerQuery   = props.int_("samples_per_query", 2);
    m_rrDepth           = props.int_("rr_depth", 5);
    m_useLaser          = props.bool_("use_laser", false);
    laserOrigin         = props.vector3f("laser_origin", Vec3());
    laserDirection      = props.vector3f("laser_direction", Vec3());
    m_useNonLinear            = props.bool_("use_non_linear", true);
    m_useNonLinearCameraRays  = props.bool_("use_non_linear_camera", true);
    m_useNLAtomicQuery        = props.bool_("use_nl_atomic_query", false);
    m_useDirectIllum          = props.bool_("use_direct_illum", true);
    m_diceVRL                 = props.int_("dice_vrl", 1);
    m_longVRL                 = props.bool_("long_vrl", false);
    m_useUniformSampling      = props.bool_("use_uniform_sampling", false) || (m_useNLAtomicQuery && m_useNonLinearCameraRays);
    minVRLLength              = props.float_("min_vrl_length", 5);
    m_useLightCut             = props.bool_("use_light_cut", false);
    m_stochasticLightcut      = props.bool_("stochastic_lightcut", false);
    m_lightcutSamples         = props.int_("lightcut_samples", 1);
    m_RRVRL                   = props.bool_("rr_vrl", false);
    m_scaleRR                 = props.float_("scale_rr", 0.5);
    m_thresholdBetterDist     = props.int_("threshold_better_dist", 8);
    m_thresholdError          = props.float_("threshold_error", 0.02);
    m_shootCenter             = props.bool_("shoot_center", true);
    // additions of this implementation
    m_device   = props.int_("device", 0);
    m_mathMode = props.int_("math_mode", 1);
    // the XML loader's check, xml.cpp:1050-1068
    auto rest = props.unqueried();
    if (!rest.empty()) {
        std::string msg = "Error while loading \"<integrator type=\\\"" + props.plugin_name() + "\\\">\": unreferenced propert" +
                          (rest.size() > 1 ? "ies " : "y ");
        for (size_t i = 0; i < rest.size(); ++i) msg += (i ? ", \"" : "\"") + rest[i] + "\"";
        Throw(msg);
    }
    // The light cut (vrl_lightcut.h): the library builds the reference's tree over the final VRL set (vrlb_lightcut_build in
    // preprocess / set_vrls) and every query walks its cut (vrl_map.h:198-201).  stochastic_lightcut re-draws the
    // representatives per query on a per-thread copy of the tree (vrl_lightcut.h:153-193, 257-261): not built.
    if (m_useLightCut && m_stochasticLightcut)
        Throw("\"stochastic_lightcut\" is not supported by the B200 light cut (deterministic representatives only)");
    if (m_useLightCut && m_useNLAtomicQuery && m_useNonLinearCameraRays)
        Throw("\"use_light_cut\" with \"use_nl_atomic_query\": multi-part queries through the cut are not supported");
    if (m_samplesPerQuery < 1) Throw("\"samples_per_query\" must be >= 1");
}

VRLIntegrator::~VRLIntegrator() { if (m_ctx) vrlb_destroy(m_ctx); }

void VRLIntegrator::check(int rc, const char *what) const {
    if (rc == VRLB_OK) return;
    std::string msg = std::string(what) + ": " + (m_ctx ? vrlb_last_error(m_ctx) : vrlb_last_error(nullptr));
    Throw(msg);
}

void VRLIntegrator::sample() const { Throw("VRLIntegrator::sample(): not implemented on the host; the B200 path has no CPU fallback"); }

vrlb_options VRLIntegrator::make_options(const HostSensor &sensor) const {
    vrlb_options o{};
    o.spp = sensor.spp;
    // integrator.cpp:91-99: 32 unless `block_size` is given (the thread-count dependent halving is a CPU detail;
    // parity runs pin block_size = 32)
    o.block_size = m_blockSize > 0 ? (int32_t) m_blockSize : 32;
    o.base_seed = sensor.sampler_seed;
    o.samples_per_query = (int32_t) m_samplesPerQuery;
    o.use_uniform_sampling = m_useUniformSampling;
    o.rr_vrl = m_RRVRL; o.scale_rr = (float) m_scaleRR;
    o.max_depth = (int32_t) m_maxDepth;
    o.use_non_linear_camera = m_useNonLinearCameraRays; o.use_nl_atomic_query = m_useNLAtomicQuery;
    o.visibility = 1; o.hide_emitters = m_hideEmitters; o.math_mode = (int32_t) m_mathMode;
    o.use_light_cut = m_useLightCut; o.threshold_error = (float) m_thresholdError; o.threshold_better_dist = (int32_t) m_thresholdBetterDist;
    return o;
}

void VRLIntegrator::set_vrls(const VrlSet &set) {
    if (!m_ctx) check(vrlb_create((int) m_device, &m_ctx), "vrlb_create");
    vrlb_vrlset_desc d{};
    d.soa10 = set.soa10.data(); d.medium = set.medium.empty() ? nullptr : set.medium.data();
    d.n = set.n; d.first_index = 0; d.total = set.n; d.scale = set.scale;
    check(vrlb_vrls_upload(m_ctx, &d), "vrlb_vrls_upload");
    m_have_vrls = true;
    if (m_useLightCut) build_lightcut(0);
}

void VRLIntegrator::preprocess(const HostScene &scene, const HostSensor &sensor) {
    if (!m_ctx) check(vrlb_create((int) m_device, &m_ctx), "vrlb_create");
    vrlb_scene_desc sd{};
    sd.tri_verts = scene.tri_verts.data(); sd.tri_flags = scene.tri_flags.data(); sd.n_tris = (uint32_t) scene.tri_flags.size();
    sd.media = scene.media.data(); sd.n_media = (uint32_t) scene.media.size();
    sd.emitters = scene.emitters.data(); sd.n_emitters = (uint32_t) scene.emitters.size();
    if (scene.tri_reflectance.size() == scene.tri_flags.size() * 3 && !scene.tri_reflectance.empty()) sd.tri_reflectance = scene.tri_reflectance.data();
    check(vrlb_scene_set(m_ctx, &sd), "vrlb_scene_set");
    if (!m_have_vrls) {
        // light tracing on the device (vrl.cpp:117-374 -> generate.cu); scale = 1 / particles traced (vrl.cpp:404,423)
        vrlb_generate_desc g{};
        g.target_vrls = (uint64_t) m_targetVRLs; g.seed = 0;                       // sampler->seed(0), vrl.cpp:98
        g.max_depth = (int32_t) m_maxDepth; g.rr_depth = (int32_t) m_rrDepth; g.min_vrl_length = (float) minVRLLength;
        g.use_non_linear = m_useNonLinear; g.use_laser = m_useLaser;
        const float lo[3] = { laserOrigin.x, laserOrigin.y, laserOrigin.z }, ld[3] = { laserDirection.x, laserDirection.y, laserDirection.z };
        for (int k = 0; k < 3; ++k) { g.laser_origin[k] = lo[k]; g.laser_direction[k] = ld[k]; g.laser_power[k] = scene.laser_power[k]; }
        g.start_medium = scene.emission_medium;
        uint64_t n = 0, particles = 0;
        check(vrlb_vrls_generate(m_ctx, &g, &n, &particles), "vrlb_vrls_generate");
        std::fprintf(stderr, "[VRLIntegrator] VRL Map Created: %llu VRLs from %llu particles\n", (unsigned long long) n, (unsigned long long) particles);
        m_have_vrls = true;
        // the volume photon map of the direct in-scattering term (vrl.cpp:79: capacity volume_photons iff use_direct_illum;
        // lookup radius = volume_lookup_radius_relative * |bbox.center - bbox.max|, vrl.cpp:449-451).  Same particles as above.
        double radius = 0.0;                                                       // sceneRadius = |bbox.center - bbox.max|, vrl.cpp:447
        if (!scene.tri_verts.empty()) {
            double lo[3] = { 1e300, 1e300, 1e300 }, hi[3] = { -1e300, -1e300, -1e300 };
            for (size_t i = 0; i < scene.tri_verts.size(); ++i) { lo[i % 3] = std::min(lo[i % 3], (double) scene.tri_verts[i]); hi[i % 3] = std::max(hi[i % 3], (double) scene.tri_verts[i]); }
            radius = 0.5 * std::sqrt((hi[0] - lo[0]) * (hi[0] - lo[0]) + (hi[1] - lo[1]) * (hi[1] - lo[1]) + (hi[2] - lo[2]) * (hi[2] - lo[2]));
        }
        // the global surface map (vrl.cpp:76, 379-390): photons at every smooth surface the same particles meet; the camera
        // path collects it at its first smooth surface (vrl.cpp:671-680); the caustic map likewise when the scene has a
        // dielectric (a caustic photon needs `wasTransmitted`, vrl.cpp:345, 776-780).
        if (m_globalPhotons > 0 && radius > 0.0 && sd.tri_reflectance) {
            uint64_t np = 0, paths = 0;
            check(vrlb_surface_photons_generate(m_ctx, VRLB_PHOTONMAP_GLOBAL, &g, (uint64_t) m_globalPhotons, (float) (m_globalLookupRadiusRelative * radius),
                                                (int32_t) m_globalLookupSize, &np, &paths), "vrlb_surface_photons_generate");
            if (np) std::fprintf(stderr, "[VRLIntegrator] Building global PM, size: %llu (%llu particles), lookup radius %g\n",
                                 (unsigned long long) np, (unsigned long long) paths, m_globalLookupRadiusRelative * radius);
            else std::fprintf(stderr, "[VRLIntegrator] No global photons\n");
            bool any_dielectric = false;
            for (uint32_t f : scene.tri_flags) any_dielectric = any_dielectric || (f & VRLB_SURF_BSDF_MASK) == VRLB_SURF_BSDF_DIELECTRIC;
            if (m_causticPhotons > 0 && any_dielectric) {                   // only a transmission makes a caustic photon (vrl.cpp:345, 776-780)
                check(vrlb_surface_photons_generate(m_ctx, VRLB_PHOTONMAP_CAUSTIC, &g, (uint64_t) m_causticPhotons, (float) (m_causticLookupRadiusRelative * radius),
                                                    (int32_t) m_causticLookupSize, &np, &paths), "vrlb_surface_photons_generate");
                if (np) std::fprintf(stderr, "[VRLIntegrator] Building caustic PM, size: %llu (%llu particles), lookup radius %g\n",
                                     (unsigned long long) np, (unsigned long long) paths, m_causticLookupRadiusRelative * radius);
                else std::fprintf(stderr, "[VRLIntegrator] No caustic photons\n");
            }
        }
        if (m_useDirectIllum && m_volumePhotons > 0 && radius > 0.0) {
            uint64_t np = 0, paths = 0;
            check(vrlb_photons_generate(m_ctx, &g, (uint64_t) m_volumePhotons, (float) (m_volumeLookupRadiusRelative * radius), &np, &paths), "vrlb_photons_generate");
            std::fprintf(stderr, "[VRLIntegrator] Building volume PM, size: %llu (%llu particles), lookup radius %g\n",
                         (unsigned long long) np, (unsigned long long) paths, m_volumeLookupRadiusRelative * radius);
        }
    }
    // vrl.cpp:408-415: the map may be turned into long beams and / or diced before it is used
    if (m_longVRL) check(vrlb_vrls_to_long(m_ctx), "vrlb_vrls_to_long");
    if (m_diceVRL > 1) {
        uint64_t n = 0;
        check(vrlb_vrls_dice(m_ctx, (uint32_t) m_diceVRL, &n), "vrlb_vrls_dice");
        std::fprintf(stderr, "[VRLIntegrator] Dicing the VRL... Number of VRLs: %llu\n", (unsigned long long) n);
    }
    if (m_useLightCut && (!m_lightcutBuilt || m_longVRL || m_diceVRL > 1)) build_lightcut(sensor.sampler_seed);
    m_preprocessed = true;
}

// VRLMap::build's ELightCutAcceleration branch (vrl_map.h:89-97, called from vrl.cpp:424 with the preprocess sampler)
void VRLIntegrator::build_lightcut(uint64_t base_seed) {
    check(vrlb_lightcut_build(m_ctx, base_seed, 0), "vrlb_lightcut_build");
    vrlb_lightcut_stats st{};
    vrlb_lightcut_get_stats(m_ctx, &st);
    std::fprintf(stderr, "[VRLIntegrator] Building VRL lightcut acceleration: %llu nodes, %llu partner searches, %.0f ms\n",
                 (unsigned long long) st.n_nodes, (unsigned long long) st.build_searches, st.build_ms);
    m_lightcutBuilt = true;
}

bool VRLIntegrator::render(const HostScene &scene, const HostSensor &sensor, std::vector<float> &xyzaw) {
    if (!m_preprocessed) preprocess(scene, sensor);          // Python callers never call preprocess
    vrlb_options o = make_options(sensor);
    xyzaw.assign((size_t) sensor.camera.width * sensor.camera.height * 5, 0.f);
    int rc = vrlb_render(m_ctx, &sensor.camera, &o, nullptr, xyzaw.data());
    if (rc == VRLB_ERR_CANCELLED) return false;              // render() returns !m_stop, integrator.cpp:182
    check(rc, "vrlb_render");
    vrlb_get_timing(m_ctx, &m_timing);
    postprocess();
    return true;
}

void VRLIntegrator::cancel() { if (m_ctx) vrlb_cancel(m_ctx); }

void VRLIntegrator::postprocess() const {
    // vrl.cpp:734-749 logs query counters; same information from the device timings
    std::fprintf(stderr, "[VRLIntegrator] %llu camera segments x %llu VRLs: %llu evals, gather %.3f ms (%.3g evals/s), %u launches\n",
                 (unsigned long long) m_timing.n_segments, (unsigned long long) (m_ctx ? vrlb_vrls_count(m_ctx) : 0),
                 (unsigned long long) m_timing.n_evals, m_timing.ms_gather,
                 m_timing.ms_gather > 0 ? m_timing.n_evals / (m_timing.ms_gather * 1e-3) : 0.0, m_timing.total_launches);
}

std::string VRLIntegrator::to_string() const {
    std::ostringstream o;
    o << "VRLIntegrator[\n  max_depth = " << m_maxDepth << ",\n  rr_depth = " << m_rrDepth << ",\n  target_vrls = " << m_targetVRLs
      << ",\n  samples_per_query = " << m_samplesPerQuery << ",\n  use_uniform_sampling = " << m_useUniformSampling
      << ",\n  rr_vrl = " << m_RRVRL << ",\n  use_light_cut = " << m_useLightCut << " (threshold_error = " << m_thresholdError
      << ", threshold_better_dist = " << m_thresholdBetterDist << ")"
      << ",\n  backend = libvrl_b200 (sm_100a), math_mode = " << m_mathMode << "\n]";
    return o.str();
}

}  // namespace vrlhost

// ---- plugin discovery symbols (include/mitsuba/core/class.h:206-211, plugin.cpp:84-118) -------------------------
extern "C" {
const char *plugin_name() { return "VRLIntegrator"; }
const char *plugin_descr() { return "Virtual ray lights, B200 gather (libvrl_b200)"; }
}
