// vrl_host_main.cpp -- headless driver over the C++ host mirror (the role src/mitsuba/mitsuba.cpp:79-112 plays for
// the reference: load -> preprocess -> render -> film).  Usage:
//   vrl_host render <scene.vrls> <set.vrlb> <out.f32> [key=value ...]     integrator properties as key=value
//   vrl_host selftest                                                     property / error behaviour, no GPU needed
// scene.vrls (test harness dump, written by tests/): u32 n_tris | verts | flags | u32 n_media | vrlb_medium[] |
// u32 n_emitters | vrlb_emitter[] | vrlb_camera | i32 spp | u64 seed.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>

#include "vrl_integrator.h"
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
    expect_throw("lightcut not yet", [] { Properties p; p.set_bool("use_light_cut", true); VRLIntegrator it(p); }, "lightcut");
    { VrlSet s; s.n = 3; s.scale = 0.25f; s.soa10.resize(30); for (int i = 0; i < 30; ++i) s.soa10[i] = (float) i; s.medium = { 0, 1, 0 };
      s.save("/tmp/_vrlb_selftest.vrlb"); VrlSet t; t.load("/tmp/_vrlb_selftest.vrlb");
      bool ok = t.n == 3 && t.scale == 0.25f && t.soa10 == s.soa10 && t.medium == s.medium;
      std::printf("%s VRLB file round trip\n", ok ? "ok  " : "FAIL"); fails += !ok; }
    std::printf("%s\n", fails ? "SELFTEST FAILED" : "SELFTEST PASSED");
    return fails ? 1 : 0;
}

int main(int argc, char **argv) {
    try {
        if (argc >= 2 && std::string(argv[1]) == "selftest") return selftest();
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
        std::ofstream o(argv[4], std::ios::binary);
        o.write((const char *) film.data(), (std::streamsize) (film.size() * 4));
        return ok ? 0 : 3;
    } catch (const std::exception &e) {                    // mitsuba.cpp:251-255: exceptions end the run with a message
        std::fprintf(stderr, "Caught a critical exception: %s\n", e.what());
        return 1;
    }
}
