/* TEST INFRASTRUCTURE.  Compiles the reference's OWN src/integrators/vrl/vrl_struct.h (included where it
 * lies under /root/reference, never copied) against the mock Mitsuba of ref_mock.h and the reference's
 * vendored Enoki, instantiates VRL<float, Color3f-like> and prints golden vectors for the functions the
 * hot path is made of:
 *   closest        VRL::findClosetPoint                        vrl_struct.h:163-228
 *   sampling       VRL::samplingVRL (importance and uniform)   vrl_struct.h:239-360
 *   transmittance  VRL::evalTransmittance                      vrl_struct.h:509-572
 *   contrib        VRL::getContrib                             vrl_struct.h:584-662
 *   query          getContrib over a VRL list on ONE sequential sampler stream, summed with the
 *                  NaN/Inf rejection, m_scale and 1/samples of VRLMap::query (vrl_map.h:151-169, 203-204;
 *                  that loop is restated here because vrl_map.h pulls in the whole light-cut/kd-tree stack)
 * Inputs are generated here from a seeded enoki::PCG32 and written next to the outputs, so the fixture
 * (tests/golden/ref_vrl_struct.json, written by oracle/make_golden.py) is self-contained. */
#include "ref_mock.h"
#include "vrl_struct.h"
#include <cstdio>

using namespace mitsuba;
using V3  = MockVec3;
using VRLf = VRL<float, MockVec3>;
using NLf  = NLRay<float, MockVec3>;

static enoki::PCG32<float> g_in(20261104u, 7u);
static float uni(float a, float b) { return a + (b - a) * g_in.next_float32(); }
static V3 rand_dir() {
    for (;;) {
        V3 v(uni(-1, 1), uni(-1, 1), uni(-1, 1));
        float n = squared_norm(v);
        if (n > 1e-3f && n <= 1.f) return normalize(v);
    }
}
static void p3(const V3 &v) { printf("%.9g, %.9g, %.9g", (double) v.x(), (double) v.y(), (double) v.z()); }
static void pf(float f);
static void p3n(const V3 &v) { pf(v.x()); printf(", "); pf(v.y()); printf(", "); pf(v.z()); }
static void pf(float f) {
    if (std::isnan(f)) printf("\"nan\""); else if (std::isinf(f)) printf(f > 0 ? "\"inf\"" : "\"-inf\""); else printf("%.9g", (double) f);
}

struct Pair { V3 vo, vd; float vlen; V3 ro, rd; float rmaxt; V3 flux; };

static Pair make_pair(int kind) {
    Pair p;
    p.vo = V3(uni(-8, 8), uni(-8, 8), uni(-8, 8));
    p.vd = rand_dir();
    p.vlen = uni(0.2f, 12.f);
    p.ro = V3(uni(-8, 8), uni(-8, 8), uni(-8, 8));
    p.rd = rand_dir();
    p.rmaxt = uni(0.5f, 25.f);
    p.flux = V3(uni(0.1f, 2.f), uni(0.1f, 2.f), uni(0.1f, 2.f));
    if (kind == 1) p.rd = p.vd;                                   /* parallel: sinTheta == 0 early-out */
    if (kind == 2) p.rd = -p.vd;
    if (kind == 3) { p.ro = p.vo + p.vd * (0.3f * p.vlen) - p.rd * (0.4f * p.rmaxt); } /* crossing: h ~ 0 */
    if (kind == 4) { p.ro = p.vo + p.vd * (p.vlen + 3.f) + rand_dir() * 0.5f; }        /* beyond the VRL end */
    if (kind == 5) { p.vlen = 0.01f; }                                                 /* tiny VRL */
    if (kind == 6) { p.rd = normalize(p.vd + rand_dir() * 1e-3f); }                    /* nearly parallel */
    return p;
}
static VRLf make_vrl(const Pair &p, const MockMedium *m) {
    VRLf v(p.vo, m, p.flux, 1, 0, false);
    v.direction = p.vd; v.length = p.vlen;
    return v;
}
static NLf make_ray(const Pair &p) {
    NLf nl;
    MockRay r(p.ro, p.rd, 0.f);
    r.mint = 0.f; r.maxt = p.rmaxt;
    nl.push_back(r);
    nl.wi = -p.rd; /* quirk 1: the reference leaves wi unset on the importance path; the defined value is -d */
    return nl;
}
static void print_pair(const Pair &p) {
    printf("\"vrl\": ["); p3(p.vo); printf(", "); p3(p.vd); printf(", %.9g, ", (double) p.vlen); p3(p.flux);
    printf("], \"ray\": ["); p3(p.ro); printf(", "); p3(p.rd); printf(", %.9g]", (double) p.rmaxt);
}

int main() {
    MockMedium med;
    med.sigma_t = V3(0.08f, 0.11f, 0.14f); med.albedo = V3(0.9f, 0.8f, 0.7f); med.scale = 1.5f;
    MockMedium med_hg = med; med_hg.phase.hg = true; med_hg.phase.g = 0.6f;
    MockScene empty, occl;
    /* occluders: a quad in the plane z = 0.5 and a tilted triangle */
    const float q[][3] = { {-3,-3,0.5f},{3,-3,0.5f},{3,3,0.5f}, {-3,-3,0.5f},{3,3,0.5f},{-3,3,0.5f}, {-6,1,-4},{-1,5,2},{-5,6,-3} };
    for (auto &v : q) occl.tri.push_back(V3(v[0], v[1], v[2]));

    printf("{\n \"medium\": {\"sigma_t\": [0.08, 0.11, 0.14], \"albedo\": [0.9, 0.8, 0.7], \"scale\": 1.5, \"g\": 0.6},\n");
    printf(" \"triangles\": [");
    for (size_t i = 0; i < occl.tri.size(); ++i) { if (i) printf(", "); p3(occl.tri[i]); }
    printf("],\n");

    /* ---- closest ---- */
    printf(" \"closest\": [\n");
    const int NC = 160;
    for (int i = 0; i < NC; ++i) {
        Pair p = make_pair(i < 100 ? 0 : 1 + (i % 6));
        VRLf v = make_vrl(p, &med);
        MockRay r(p.ro, p.rd, 0.f); r.mint = 0; r.maxt = p.rmaxt;
        auto c = v.findClosetPoint(r);
        printf("  {"); print_pair(p); printf(", \"out\": [%.9g, %.9g]}%s\n", (double) c.tCam, (double) c.tVRL, i + 1 < NC ? "," : "");
    }
    printf(" ],\n");

    /* ---- sampling ---- */
    printf(" \"sampling\": [\n");
    const int NS = 200;
    for (int i = 0; i < NS; ++i) {
        Pair p = make_pair(i < 140 ? 0 : 1 + (i % 6));
        int uniform = (i % 4) == 3;
        VRLf v = make_vrl(p, &med);
        NLf nl = make_ray(p);
        MockSampler s, s2; s.seed(0, 1000 + i); s2.seed(0, 1000 + i);
        float u1 = s2.next_1d(), u2 = s2.next_1d();
        auto r = v.samplingVRL(&empty, nl, &s, uniform != 0, 0);
        printf("  {"); print_pair(p);
        printf(", \"uniform\": %d, \"seed\": %d, \"u\": [%.9g, %.9g], \"draws\": %d, \"out\": [", uniform, 1000 + i, (double) u1, (double) u2, (int) s.draws);
        p3n(r.pCam); printf(", "); pf(r.tCam); printf(", "); p3n(r.pVRL); printf(", "); pf(r.tVRL); printf(", "); pf(r.invPDF);
        printf("]}%s\n", i + 1 < NS ? "," : "");
    }
    printf(" ],\n");

    /* ---- transmittance ---- */
    printf(" \"transmittance\": [\n");
    const int NT = 80;
    for (int i = 0; i < NT; ++i) {
        V3 a(uni(-6, 6), uni(-6, 6), uni(-5, 5)), b(uni(-6, 6), uni(-6, 6), uni(-5, 5));
        MockSampler s; s.seed(0, 5000 + i);
        int interactions = -1;
        V3 tr = VRLf::evalTransmittance(&occl, a, false, b, false, &med, interactions, &s, 0, true);
        printf("  {\"p1\": ["); p3(a); printf("], \"p2\": ["); p3(b); printf("], \"out\": ["); p3n(tr); printf("]}%s\n", i + 1 < NT ? "," : "");
    }
    printf(" ],\n");

    /* ---- contrib ---- */
    printf(" \"contrib\": [\n");
    const int NG = 240;
    for (int i = 0; i < NG; ++i) {
        Pair p = make_pair(i < 180 ? 0 : 1 + (i % 6));
        int uniform = (i % 5) == 4, hg = (i % 3) == 1, vis = (i % 2) == 1;
        const MockMedium *m = hg ? &med_hg : &med;
        VRLf v = make_vrl(p, m);
        NLf nl = make_ray(p);
        MockSampler s, s2; s.seed(0, 9000 + i); s2.seed(0, 9000 + i);
        float u1 = s2.next_1d(), u2 = s2.next_1d();
        V3 c = v.getContrib(vis ? &occl : &empty, uniform != 0, nl, &s, 0);
        printf("  {"); print_pair(p);
        printf(", \"uniform\": %d, \"hg\": %d, \"vis\": %d, \"u\": [%.9g, %.9g], \"draws\": %d, \"out\": [", uniform, hg, vis, (double) u1, (double) u2, (int) s.draws);
        pf(c.x()); printf(", "); pf(c.y()); printf(", "); pf(c.z());
        printf("]}%s\n", i + 1 < NG ? "," : "");
    }
    printf(" ],\n");

    /* ---- query: one sequential stream over a VRL list (vrl_map.h:151-169) ---- */
    const int NV = 96, NQ = 12;
    std::vector<Pair> vp(NV);
    std::vector<VRLf> vrls;
    for (int j = 0; j < NV; ++j) { vp[j] = make_pair(j % 16 == 15 ? 1 : 0); }
    printf(" \"query_vrls\": [");
    for (int j = 0; j < NV; ++j) {
        printf("%s[", j ? ", " : ""); p3(vp[j].vo); printf(", "); p3(vp[j].vd); printf(", %.9g, ", (double) vp[j].vlen); p3(vp[j].flux); printf("]");
    }
    printf("],\n \"query_scale\": 0.015625,\n \"query\": [\n");
    const float m_scale = 0.015625f;
    for (int k = 0; k < NQ; ++k) {
        Pair p = make_pair(0);
        /* rays parallel to VRL 15, 31, ... exercise the early-out that consumes no draws */
        if (k % 4 == 3) p.rd = vp[15].vd;
        int uniform = (k % 3) == 2, vis = (k % 2), samples = 1 + (k % 2);
        NLf nl = make_ray(p);
        MockSampler s; s.seed(0, 20000 + k);
        V3 Li(0.f); int nb = 0;
        for (int smp = 0; smp < samples; ++smp)
            for (int j = 0; j < NV; ++j) {
                VRLf v = make_vrl(vp[j], &med);
                V3 c = v.getContrib(vis ? &occl : &empty, uniform != 0, nl, &s, 0) * m_scale;
                if (std::isnan(c[0]) || std::isnan(c[1]) || std::isnan(c[2]) || std::isinf(c[0]) || std::isinf(c[1]) || std::isinf(c[2])) continue;
                Li += c; nb++;
            }
        if (samples > 1) Li /= (float) samples;
        printf("  {\"ray\": ["); p3(p.ro); printf(", "); p3(p.rd); printf(", %.9g], \"uniform\": %d, \"vis\": %d, \"samples\": %d, \"seed\": %d, \"draws\": %d, \"nb\": %d, \"out\": [",
               (double) p.rmaxt, uniform, vis, samples, 20000 + k, (int) s.draws, nb);
        p3n(Li); printf("]}%s\n", k + 1 < NQ ? "," : "");
    }
    printf(" ],\n");

    /* ---- query_rr: the same loop with the distance Russian roulette of vrl_map.h:171-190 (EDistanceRoulette): the
     *      roulette draw comes first, a survivor's contribution is divided by its survival weight ---- */
    printf(" \"query_rr\": [\n");
    for (int k = 0; k < 8; ++k) {
        Pair p = make_pair(0);
        int uniform = (k % 3) == 2, vis = (k % 2), samples = 1 + (k % 2);
        const float scaleRR = k < 4 ? 0.5f : 2.0f;
        NLf nl = make_ray(p);
        auto ray = nl.first();
        MockSampler s; s.seed(0, 30000 + k);
        V3 Li(0.f); int nb = 0, pruned = 0;
        for (int smp = 0; smp < samples; ++smp)
            for (int j = 0; j < NV; ++j) {
                VRLf v = make_vrl(vp[j], &med);
                auto res = v.findClosetPoint(ray);
                float min_distance_sqr = sqrt(enoki::squared_norm((ray.o + ray.d * res.tCam) - (v.origin + v.direction * res.tVRL)));
                float rrWeight = min(1 / (min_distance_sqr * scaleRR), 1.0);
                if (rrWeight > s.next_1d()) {
                    V3 c = (v.getContrib(vis ? &occl : &empty, uniform != 0, nl, &s, 0) / rrWeight) * m_scale;
                    if (std::isnan(c[0]) || std::isnan(c[1]) || std::isnan(c[2]) || std::isinf(c[0]) || std::isinf(c[1]) || std::isinf(c[2])) continue;
                    Li += c; nb++;
                } else pruned++;
            }
        if (samples > 1) Li /= (float) samples;
        printf("  {\"ray\": ["); p3(p.ro); printf(", "); p3(p.rd); printf(", %.9g], \"uniform\": %d, \"vis\": %d, \"samples\": %d, \"seed\": %d, \"scale_rr\": %.9g, \"draws\": %d, \"nb\": %d, \"pruned\": %d, \"out\": [",
               (double) p.rmaxt, uniform, vis, samples, 30000 + k, (double) scaleRR, (int) s.draws, nb, pruned);
        p3n(Li); printf("]}%s\n", k + 1 < 8 ? "," : "");
    }
    printf(" ],\n");

    /* ---- dice: VRL::dice (vrl_struct.h:88-128) with the dice length VRLMap::dicingVRL derives (vrl_map.h:104-118,
     *      double mean of the lengths / avgNumberDice, passed as a Float); long: VRL::convertToLong (146-154) ---- */
    {
        const int ND = 24, avgNumberDice = 3;
        std::vector<Pair> dp(ND);
        for (int j = 0; j < ND; ++j) { dp[j] = make_pair(j % 8 == 7 ? 5 : 0); }
        dp[3].vlen = 0.f;                       /* zero-length VRL: one zero-length piece */
        auto avg_length = 0.0;
        for (int j = 0; j < ND; ++j) avg_length += dp[j].vlen;
        avg_length /= ND;
        printf(" \"dice_in\": [");
        for (int j = 0; j < ND; ++j) { printf("%s[", j ? ", " : ""); p3(dp[j].vo); printf(", "); p3(dp[j].vd); printf(", %.9g, ", (double) dp[j].vlen); p3(dp[j].flux); printf("]"); }
        printf("],\n \"dice_n\": %d,\n \"dice_out\": [\n", avgNumberDice);
        MockSampler s; s.seed(0, 777);
        bool first = true;
        for (int j = 0; j < ND; ++j) {
            VRLf v = make_vrl(dp[j], &med);
            auto pieces = v.dice(&empty, &s, avg_length / avgNumberDice);
            for (auto &q2 : pieces) {
                printf("%s  [%d, ", first ? "" : ",\n", j); p3(q2.origin); printf(", "); p3(q2.direction); printf(", %.9g, ", (double) q2.length); p3(q2.flux); printf("]");
                first = false;
            }
        }
        printf("\n ],\n \"long\": [\n");
        /* a closed room around the occluders so that every VRL sees a surface */
        MockScene room = occl;
        const float R = 12.f;
        const float c8[8][3] = { {-R,-R,-R},{R,-R,-R},{R,R,-R},{-R,R,-R},{-R,-R,R},{R,-R,R},{R,R,R},{-R,R,R} };
        const int f12[12][3] = { {0,1,2},{0,2,3},{4,6,5},{4,7,6},{0,4,5},{0,5,1},{3,2,6},{3,6,7},{0,3,7},{0,7,4},{1,5,6},{1,6,2} };
        for (auto &f : f12) for (int k = 0; k < 3; ++k) room.tri.push_back(V3(c8[f[k]][0], c8[f[k]][1], c8[f[k]][2]));
        for (int j = 0; j < ND; ++j) {
            VRLf v = make_vrl(dp[j], &med);
            v.convertToLong(&room);
            printf("  %.9g%s\n", (double) v.length, j + 1 < ND ? "," : "");
        }
        printf(" ],\n \"room_triangles\": [");
        for (size_t i = 0; i < room.tri.size(); ++i) { if (i) printf(", "); p3(room.tri[i]); }
        printf("],\n");
    }
    printf(" \"warnings\": %d\n}\n", MockLog::warnings);
    return 0;
}
