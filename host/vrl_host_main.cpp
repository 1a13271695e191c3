// vrl_host_main.cpp -- headless driver over the C++ host mirror (the role src/mitsuba/mitsuba.cpp:79-112 plays for
// the reference: load -> preprocess -> render -> film).  Usage:
//   vrl_host render <scene.vrls> <set.vrlb> <out.f32> [key=value ...]     integrator properties as key=value
//   vrl_host selftest                                                     property / error behaviour, no GPU needed
//   vrl_host vol-info <file.vol> | obj-info <file.obj>                     the loaders of vrl_io.h (no GPU needed)
//   vrl_host xml-info <scene.xml> [-Dname=value ...]                       what the scene loader of vrl_xml.h sees (no GPU needed)
//   vrl_host convert-film <in.f32> <width> <height> <out.exr|out.pfm> [float16]   raw XYZAW film -> developed image
//   vrl_host render-xml <scene.xml> <out.exr|out.pfm|out.f32> [-Dname=value ...] [key=value ...]
//                                                                         the role of `mitsuba scene.xml`: load -> VRLIntegrator ->
//                                                                         preprocess (light tracing on the device) -> render -> film;
//                                                                         key=value overrides integrator properties of the file
// An output name ending in .pfm receives the developed film (XYZ / W -> linear sRGB, hdrfilm.cpp:259-262) instead of
// the raw XYZAW floats.
// scene.vrls (test harness dump, written by tests/): u32 n_tris | verts | flags | u32 n_media | vrlb_medium[] |
// u32 n_emitters | vrlb_emitter[] | vrlb_camera | i32 spp | u64 seed.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>

#include "vrl_integrator.h"
#include "vrl_io.h"
#include "vrl_xml.h"
using namespace vrlhost;

static void set_kv(Properties &p, const std::string &kv) {
    auto eq = kv.find('=');
    if (eq == std::string::npos) throw std::runtime_error("expected key=value: " + kv);
    std::string k = kv.substr(0, eq), v = kv.substr(eq + 1);
    if (v == "true" || v == "false") p.set_bool(k, v == "true");
    else if (v.find_first_of(".eE") != std::string::npos) p.set_float(k, std::stod(v));
    else p.set_int(k, std::stoll(v));
}

static int selftest() {
    int fails = 0;
    auto expect_throw = [&](const char *name, auto &&fn, const char *needle) {
        try { fn(); std::printf("FAIL %s: no exception\n", name); ++fails; }
        catch (const std::exception &e) {
            if (std::strstr(e.what(), needle)) std::printf("ok   %s: %s\n", name, e.what());
            else { std::printf("FAIL %s: unexpected message: %s\n", name, e.what()); ++fails; }
        }
    };
    { Properties p; VRLIntegrator it(p);
      bool ok = it.m_targetVRLs == 1000 && it.m_samplesPerQuery == 2 && it.m_maxDepth == 512 && it.m_rrDepth == 5 && it.m_globalPhotons == 250000 &&
                it.m_volumePhotons == 1000 && !it.m_useUniformSampling && !it.m_RRVRL && it.m_scaleRR == 0.5 && it.minVRLLength == 5 &&
                it.m_useNonLinearCameraRays && !it.m_useNLAtomicQuery && it.m_blockSize == 0;
      std::printf("%s defaults (vrl.cpp:33-73)\n", ok ? "ok  " : "FAIL"); fails += !ok; }
    { Properties p; p.set_bool("use_nl_atomic_query", true); VRLIntegrator it(p);
      bool ok = it.m_useUniformSampling;   // forced, vrl.cpp:62
      std::printf("%s use_nl_atomic_query forces uniform sampling\n", ok ? "ok  " : "FAIL"); fails += !ok; }
    expect_throw("unreferenced property", [] { Properties p; p.set_int("vrlTargetNum", 5); VRLIntegrator it(p); }, "unreferenced property");
    expect_throw("wrong type", [] { Properties p; p.set_string("target_vrls", "many"); VRLIntegrator it(p); }, "wrong type");
    expect_throw("sample() has no CPU path", [] { Properties p; VRLIntegrator it(p); it.sample(); }, "no CPU fallback");
    { Properties p; p.set_bool("use_light_cut", true); p.set_float("threshold_error", 0.05); p.set_int("threshold_better_dist", 4); VRLIntegrator it(p);
      std::printf("ok   use_light_cut is accepted (the cut is walked on the device)\n"); }
    expect_throw("stochastic_lightcut", [] { Properties p; p.set_bool("use_light_cut", true); p.set_bool("stochastic_lightcut", true); VRLIntegrator it(p); }, "not supported");
    { VrlSet s; s.n = 3; s.scale = 0.25f; s.soa10.resize(30); for (int i = 0; i < 30; ++i) s.soa10[i] = (float) i; s.medium = { 0, 1, 0 };
      s.save("/tmp/_vrlb_selftest.vrlb"); VrlSet t; t.load("/tmp/_vrlb_selftest.vrlb");
      bool ok = t.n == 3 && t.scale == 0.25f && t.soa10 == s.soa10 && t.medium == s.medium;
      std::printf("%s VRLB file round trip\n", ok ? "ok  " : "FAIL"); fails += !ok; }
    { VolumeGrid g; g.nx = 2; g.ny = 3; g.nz = 2; g.data.resize(12); for (int i = 0; i < 12; ++i) g.data[i] = 0.5f * (float) i; g.bbox_max[1] = 2.f;
      write_vol("/tmp/_vol_selftest.vol", g); VolumeGrid h = read_vol("/tmp/_vol_selftest.vol");
      bool ok = h.nx == 2 && h.ny == 3 && h.nz == 2 && h.data == g.data && h.max == 5.5f && h.bbox_max[1] == 2.f;
      std::printf("%s .vol round trip\n", ok ? "ok  " : "FAIL"); fails += !ok; }
    expect_throw("bad .vol header", [] { std::ofstream("/tmp/_vol_bad.vol") << "VOX"; read_vol("/tmp/_vol_bad.vol"); }, "Invalid volume file");
    { std::ofstream("/tmp/_obj_selftest.obj") << "v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvn 0 0 1\nf 1//1 2//1 3//1 4//1\n";
      std::vector<float> t = read_obj("/tmp/_obj_selftest.obj");
      bool ok = t.size() == 18 && t[3] == 1.f && t[9] == 0.f && t[12] == 1.f && t[13] == 1.f && t[15] == 0.f && t[16] == 1.f;   // quad fanned: (1,2,3) (1,3,4)
      std::printf("%s OBJ quad is fanned into two triangles\n", ok ? "ok  " : "FAIL"); fails += !ok; }
    expect_throw("OBJ invalid index", [] { std::ofstream("/tmp/_obj_bad.obj") << "v 0 0 0\nf 1 2 3\n"; read_obj("/tmp/_obj_bad.obj"); }, "reference to invalid vertex");
    { std::vector<float> x = { 0.412453f, 0.212671f, 0.019334f, 1.f, 2.f, 0, 0, 0, 0, 0 };      // pixel 0 = XYZ of pure red, twice weighted
      x[0] *= 2; x[1] *= 2; x[2] *= 2;
      std::vector<float> rgb = develop_rgb(x, 2, 1);
      bool ok = std::fabs(rgb[0] - 1.f) < 1e-4f && std::fabs(rgb[1]) < 1e-4f && std::fabs(rgb[2]) < 1e-4f && rgb[3] == 0.f;
      write_pfm("/tmp/_pfm_selftest.pfm", rgb, 2, 1);
      std::printf("%s develop XYZ/W -> sRGB + PFM\n", ok ? "ok  " : "FAIL"); fails += !ok; }
    std::printf("%s\n", fails ? "SELFTEST FAILED" : "SELFTEST PASSED");
    return fails ? 1 : 0;
}

int main(int argc, char **argv) {
    try {
        if (argc >= 2 && std::string(argv[1]) == "selftest") return selftest();
        if (argc == 3 && std::string(argv[1]) == "vol-info") {
            VolumeGrid g = read_vol(argv[2]);
            std::printf("shape %d %d %d channels %d bbox %.9g %.9g %.9g %.9g %.9g %.9g max %.9g mean %.9g\n", g.nx, g.ny, g.nz, g.channels,
                        g.bbox_min[0], g.bbox_min[1], g.bbox_min[2], g.bbox_max[0], g.bbox_max[1], g.bbox_max[2], g.max, g.mean);
            return 0;
        }
        if ((argc == 6 || (argc == 7 && std::string(argv[6]) == "float16")) && std::string(argv[1]) == "convert-film") {   // raw XYZAW floats -> .exr / .pfm (HDRFilm::develop)
            const int w = std::atoi(argv[3]), h = std::atoi(argv[4]);
            std::ifstream f(argv[2], std::ios::binary);
            std::vector<float> film((size_t) std::max(w, 0) * (size_t) std::max(h, 0) * 5);
            if (!f || film.empty() || !f.read((char *) film.data(), (std::streamsize) (film.size() * 4)))
                throw std::runtime_error(std::string("cannot read ") + argv[3] + " x " + argv[4] + " XYZAW pixels from " + argv[2]);
            write_film(argv[5], film, w, h, argc == 7);
            return 0;
        }
        if (argc == 3 && std::string(argv[1]) == "obj-info") {
            std::vector<float> t = read_obj(argv[2]);
            double lo[3] = { 1e300, 1e300, 1e300 }, hi[3] = { -1e300, -1e300, -1e300 }, sum = 0;
            for (size_t i = 0; i < t.size(); ++i) { lo[i % 3] = std::min(lo[i % 3], (double) t[i]); hi[i % 3] = std::max(hi[i % 3], (double) t[i]); sum += t[i]; }
            std::printf("triangles %zu bbox %.9g %.9g %.9g %.9g %.9g %.9g checksum %.9g\n", t.size() / 9, lo[0], lo[1], lo[2], hi[0], hi[1], hi[2], sum);
            return 0;
        }
        if (argc >= 4 && std::string(argv[1]) == "dump-xml") {
            // scene.xml -> the binary scene dump the Python harness reads (scenes.load_scene_dump): what the XML loader hands the
            // plugin, so that tests can run the SAME loaded fixture through the oracle and through the C ABI
            std::map<std::string, std::string> defs;
            for (int i = 4; i < argc; ++i) {
                std::string a = argv[i];
                auto eq = a.find('=');
                if (a.rfind("-D", 0) != 0 || eq == std::string::npos) throw std::runtime_error("expected -Dname=value: " + a);
                defs[a.substr(2, eq - 2)] = a.substr(eq + 1);
            }
            LoadedScene ls = load_scene_xml(argv[2], defs);
            for (auto &w : ls.warnings) std::fprintf(stderr, "[xml] warning: %s\n", w.c_str());
            std::ofstream o(argv[3], std::ios::binary);
            if (!o) throw std::runtime_error(std::string("cannot write ") + argv[3]);
            auto put = [&](const void *p, size_t n) { o.write((const char *) p, (std::streamsize) n); };
            auto put_u32 = [&](uint32_t v) { put(&v, 4); };
            put("VRLD2\0\0\0", 8);
            put_u32((uint32_t) ls.scene.tri_flags.size());
            put(ls.scene.tri_verts.data(), ls.scene.tri_verts.size() * 4); put(ls.scene.tri_flags.data(), ls.scene.tri_flags.size() * 4);
            put_u32((uint32_t) ls.scene.media.size()); put(ls.scene.media.data(), ls.scene.media.size() * sizeof(vrlb_medium));
            put_u32((uint32_t) ls.scene.emitters.size()); put(ls.scene.emitters.data(), ls.scene.emitters.size() * sizeof(vrlb_emitter));
            put(&ls.sensor.camera, sizeof(vrlb_camera)); put(&ls.sensor.spp, 4); put(&ls.sensor.sampler_seed, 8);
            put_u32(ls.scene.tri_reflectance.empty() ? 0u : 1u);
            put(ls.scene.tri_reflectance.data(), ls.scene.tri_reflectance.size() * 4);
            for (const vrlb_medium &m : ls.scene.media)
                if (m.kind == VRLB_MEDIUM_GRID) put(m.grid, (size_t) m.grid_res[0] * m.grid_res[1] * m.grid_res[2] * 4);
            put(&ls.scene.emission_medium, 4); put(ls.scene.laser_power, 12);
            if (!o) throw std::runtime_error(std::string("write failed: ") + argv[3]);
            std::printf("{\"triangles\": %zu, \"media\": %zu, \"emitters\": %zu, \"integrator\": \"%s\"}\n", ls.scene.tri_flags.size(), ls.scene.media.size(),
                        ls.scene.emitters.size(), ls.integrator_type.c_str());
            return 0;
        }
        if (argc >= 3 && (std::string(argv[1]) == "xml-info" || std::string(argv[1]) == "render-xml")) {
            const bool render = std::string(argv[1]) == "render-xml";
            if (render && argc < 4) { std::fprintf(stderr, "usage: vrl_host render-xml scene.xml out.pfm [-Dname=value ...] [key=value ...]\n"); return 2; }
            std::map<std::string, std::string> defs; std::vector<std::string> overrides;
            for (int i = render ? 4 : 3; i < argc; ++i) {
                std::string a = argv[i];
                if (a.rfind("-D", 0) == 0) { auto eq = a.find('='); if (eq == std::string::npos) throw std::runtime_error("expected -Dname=value: " + a); defs[a.substr(2, eq - 2)] = a.substr(eq + 1); }
                else overrides.push_back(a);
            }
            LoadedScene ls = load_scene_xml(argv[2], defs);
            for (auto &w : ls.warnings) std::fprintf(stderr, "[xml] warning: %s\n", w.c_str());
            if (!render) {
                const vrlb_camera &c = ls.sensor.camera;
                double lo[3] = { 1e300, 1e300, 1e300 }, hi[3] = { -1e300, -1e300, -1e300 };
                for (size_t i = 0; i < ls.scene.tri_verts.size(); ++i) { lo[i % 3] = std::min(lo[i % 3], (double) ls.scene.tri_verts[i]); hi[i % 3] = std::max(hi[i % 3], (double) ls.scene.tri_verts[i]); }
                size_t n_null = 0, n_emit = 0, n_int = 0;
                for (uint32_t f : ls.scene.tri_flags) { n_null += (f & VRLB_SURF_BSDF_MASK) == VRLB_SURF_BSDF_NULL; n_emit += (f & VRLB_SURF_EMITTER) != 0; n_int += VRLB_SURF_GET_INTERIOR(f) >= 0; }
                std::printf("{\"triangles\": %zu, \"null_triangles\": %zu, \"emitter_triangles\": %zu, \"interior_triangles\": %zu,\n", ls.scene.tri_flags.size(), n_null, n_emit, n_int);
                std::printf(" \"bbox\": [%.9g, %.9g, %.9g, %.9g, %.9g, %.9g],\n", lo[0], lo[1], lo[2], hi[0], hi[1], hi[2]);
                {   // what the light tracer sees of the BSDFs: triangles that reflect particles, their mean reflectance
                    size_t n_refl = 0; double mean[3] = { 0, 0, 0 };
                    const auto &rf = ls.scene.tri_reflectance;
                    for (size_t i = 0; i + 2 < rf.size(); i += 3) if ((ls.scene.tri_flags[i / 3] & VRLB_SURF_BSDF_MASK) == VRLB_SURF_BSDF_SMOOTH && (rf[i] > 0 || rf[i + 1] > 0 || rf[i + 2] > 0)) { ++n_refl; for (int k = 0; k < 3; ++k) mean[k] += rf[i + (size_t) k]; }
                    for (int k = 0; k < 3; ++k) mean[k] = n_refl ? mean[k] / (double) n_refl : 0.0;
                    size_t n_two = 0, n_mirror = 0, n_glass = 0;
                    for (uint32_t f : ls.scene.tri_flags) { n_two += (f & VRLB_SURF_TWOSIDED) != 0; n_mirror += (f & VRLB_SURF_BSDF_MASK) == VRLB_SURF_BSDF_MIRROR; n_glass += (f & VRLB_SURF_BSDF_MASK) == VRLB_SURF_BSDF_DIELECTRIC; }
                    std::printf(" \"reflecting_triangles\": %zu, \"twosided_triangles\": %zu, \"mirror_triangles\": %zu, \"dielectric_triangles\": %zu, \"mean_reflectance\": [%.9g, %.9g, %.9g],\n", n_refl, n_two, n_mirror, n_glass, mean[0], mean[1], mean[2]);
                }
                std::printf(" \"media\": [");
                for (size_t i = 0; i < ls.scene.media.size(); ++i) {
                    const vrlb_medium &m = ls.scene.media[i];
                    std::printf("%s{\"id\": \"%s\", \"kind\": %d, \"phase\": %d, \"g\": %.9g, \"scale\": %.9g, \"sigma_t\": [%.9g, %.9g, %.9g], \"albedo\": [%.9g, %.9g, %.9g], "
                                "\"grid_res\": [%d, %d, %d], \"grid_bbox\": [%.9g, %.9g, %.9g, %.9g, %.9g, %.9g], \"grid_max\": %.9g, "
                                "\"nl_res\": [%d, %d, %d], \"nl_bbox\": [%.9g, %.9g, %.9g, %.9g, %.9g, %.9g], \"top_ior\": %.9g, \"bottom_ior\": %.9g}",
                                i ? ", " : "", ls.media_ids[i].c_str(), m.kind, m.phase, m.g, m.scale, m.sigma_t[0], m.sigma_t[1], m.sigma_t[2], m.albedo[0], m.albedo[1], m.albedo[2],
                                m.grid_res[0], m.grid_res[1], m.grid_res[2], m.grid_bbox_min[0], m.grid_bbox_min[1], m.grid_bbox_min[2], m.grid_bbox_max[0], m.grid_bbox_max[1], m.grid_bbox_max[2], m.grid_max,
                                m.nl_res[0], m.nl_res[1], m.nl_res[2], m.nl_bbox_min[0], m.nl_bbox_min[1], m.nl_bbox_min[2], m.nl_bbox_max[0], m.nl_bbox_max[1], m.nl_bbox_max[2], m.top_ior, m.bottom_ior);
                }
                std::printf("],\n \"emitters\": [");
                for (size_t i = 0; i < ls.scene.emitters.size(); ++i) std::printf("%s[%.9g, %.9g, %.9g]", i ? ", " : "", ls.scene.emitters[i].radiance[0], ls.scene.emitters[i].radiance[1], ls.scene.emitters[i].radiance[2]);
                std::printf("],\n \"sensor\": {\"width\": %d, \"height\": %d, \"spp\": %d, \"seed\": %llu, \"medium\": %d, \"near_clip\": %.9g, \"far_clip\": %.9g,\n  \"sample_to_camera\": [",
                            c.width, c.height, ls.sensor.spp, (unsigned long long) ls.sensor.sampler_seed, c.medium, c.near_clip, c.far_clip);
                for (int i = 0; i < 16; ++i) std::printf("%s%.9g", i ? ", " : "", c.sample_to_camera[i]);
                std::printf("],\n  \"to_world\": [");
                for (int i = 0; i < 16; ++i) std::printf("%s%.9g", i ? ", " : "", c.to_world[i]);
                std::printf("]},\n \"integrator\": \"%s\", \"warnings\": %zu}\n", ls.integrator_type.c_str(), ls.warnings.size());
                if (ls.integrator_type == "vrl") {                         // the constructor is where property errors surface (xml.cpp:1050-1068)
                    Properties props = ls.integrator;
                    for (auto &kv : overrides) set_kv(props, kv);
                    VRLIntegrator integrator(props);
                    std::cerr << integrator.to_string() << std::endl;
                }
                return 0;
            }
            if (ls.integrator_type != "vrl") throw std::runtime_error("the scene's integrator is \"" + ls.integrator_type + "\"; this plugin is <integrator type=\"vrl\">");
            Properties props = ls.integrator;
            for (auto &kv : overrides) set_kv(props, kv);
            VRLIntegrator integrator(props);
            std::cerr << integrator.to_string() << std::endl;
            integrator.preprocess(ls.scene, ls.sensor);
            std::vector<float> film;
            bool ok = integrator.render(ls.scene, ls.sensor, film);
            std::string out = argv[3];
            write_film(out, film, ls.sensor.camera.width, ls.sensor.camera.height, ls.film_half);
            return ok ? 0 : 3;
        }
        if (argc < 5 || std::string(argv[1]) != "render") { std::fprintf(stderr, "usage: vrl_host render scene.vrls set.vrlb out.f32 [k=v ...] | selftest\n"); return 2; }
        HostScene sc; HostSensor se;
        std::ifstream f(argv[2], std::ios::binary);
        if (!f) throw std::runtime_error(std::string("cannot read ") + argv[2]);
        uint32_t n = 0;
        f.read((char *) &n, 4); sc.tri_verts.resize((size_t) n * 9); sc.tri_flags.resize(n);
        f.read((char *) sc.tri_verts.data(), (std::streamsize) (sc.tri_verts.size() * 4)); f.read((char *) sc.tri_flags.data(), (std::streamsize) (n * 4));
        f.read((char *) &n, 4); sc.media.resize(n); f.read((char *) sc.media.data(), (std::streamsize) (n * sizeof(vrlb_medium)));
        for (auto &m : sc.media) m.grid = nullptr;
        f.read((char *) &n, 4); sc.emitters.resize(n); f.read((char *) sc.emitters.data(), (std::streamsize) (n * sizeof(vrlb_emitter)));
        f.read((char *) &se.camera, sizeof(vrlb_camera)); f.read((char *) &se.spp, 4); f.read((char *) &se.sampler_seed, 8);
        if (!f) throw std::runtime_error("truncated scene dump");
        VrlSet set; set.load(argv[3]);
        Properties props("vrl");
        for (int i = 5; i < argc; ++i) set_kv(props, argv[i]);
        VRLIntegrator integrator(props);
        std::cerr << integrator.to_string() << std::endl;
        integrator.set_vrls(set);
        integrator.preprocess(sc, se);
        std::vector<float> film;
        bool ok = integrator.render(sc, se, film);
        std::string out = argv[4];
        write_film(out, film, se.camera.width, se.camera.height);
        return ok ? 0 : 3;
    } catch (const std::exception &e) {                    // mitsuba.cpp:251-255: exceptions end the run with a message
        std::fprintf(stderr, "Caught a critical exception: %s\n", e.what());
        return 1;
    }
}
