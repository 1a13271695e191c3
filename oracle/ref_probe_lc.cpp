/* TEST INFRASTRUCTURE.  Compiles the reference's OWN VRL light cut -- src/integrators/vrl/vrl_lightcut.h with
 * vrl_struct.h, the fork's BoundingBox::getMinDistanceSqr (include/mitsuba/core/bbox.h:335-600) and the Point/Vector
 * types of include/mitsuba/core/vector.h, all where they lie under /root/reference, never copied -- against the mock
 * Mitsuba of ref_mock.h on the reference's vendored Enoki, and prints golden vectors for SURVEY 8f-2:
 *   dist          Node::dist                                   vrl_lightcut.h:101-109   (the merge metric)
 *   merge         Node(Node*, Node*, v, n1_merge)              vrl_lightcut.h:111-151   (representative, addFlux, isSame)
 *   tree          VRLLightCut(...) -> buildLightTree           vrl_lightcut.h:196-212, 515-627, in preorder
 *   box_dist      BoundingBox::getMinDistanceSqr(ray)          bbox.h:335-600
 *   seg_dist      VRLLightCut::sqrDistanceVRL(ray, vrl)        vrl_lightcut.h:772-838
 *   bound         VRLLightCut::getClusterUpperBound            vrl_lightcut.h:429-512
 *   query         VRLLightCut::query                           vrl_lightcut.h:243-393   (Li, nb_evaluation, sampler draws)
 * -> tests/golden/ref_lightcut.json (oracle/make_golden.py); tests hold the oracle's restatement to it.
 *
 * How it is made to compile with g++ (oracle/Makefile, _ref/refprobe_lc):
 *   - vrl_lightcut.h is reached through a directory of symbolic links (oracle/_ref/lc_view/) whose kd_tree.h is a
 *     four-declaration stand-in: the reference's kd_tree.h is dead code (only the never-called buildTreeStable names
 *     it) that no conforming compiler accepts (`new KdNode<Node>` on a typedef, levComp used before its declaration);
 *   - one missing `typename` (vrl_lightcut.h:540, `Node::DistanceResult{..}` inside a lambda; MSVC accepts it) is
 *     inserted in the preprocessed stream between two g++ invocations -- a pipe, nothing is written to disk;
 *   - `private`/`protected` read as `public` so that the bound, the distances and the root can be printed.
 * What is mocked (written here or in ref_mock.h, NOT reference code): see ref_mock.h; BoundingSphere and type_suffix
 * (named by parts of bbox.h this probe never instantiates). */
#include <enoki/array.h>
#include <mutex>
#include <queue>
#include <vector>
#define NAMESPACE_BEGIN(name) namespace name {
#define NAMESPACE_END(name) }
namespace mitsuba {
using namespace enoki;
template <typename V, size_t S> struct Vector; template <typename V, size_t S> struct Point; template <typename V, size_t S> struct Normal;
}
#include "core/vector.h"                 /* the reference's own: -I$(REF)/include/mitsuba */
namespace mitsuba { template <typename V, size_t S> Vector(const Point<V, S> &) -> Vector<V, S>; }   /* `Vector wo = ...`, vrl_struct.h:548 */
#define REF_MOCK_VEC3 mitsuba::Point<float, 3>
#include "ref_mock.h"
#define MTS_INLINE inline
namespace mitsuba { template <typename P> struct BoundingSphere { P c; float r; }; template <typename T> const char *type_suffix() { return ""; } }
#include "core/bbox.h"                   /* the reference's own */
#define private public
#define protected public
#include "lc_view/vrl_lightcut.h"        /* the reference's own, through oracle/_ref/lc_view */
#undef private
#undef protected
#include <cstdio>

using namespace mitsuba;
using V3   = MockVec3;
using VRLf = VRL<float, MockVec3>;
using NLf  = NLRay<float, MockVec3>;
using LC   = VRLLightCut<float, MockVec3>;
using Box  = BoundingBox<MockVec3>;

static enoki::PCG32<float> g_in(20261106u, 11u);
static float uni(float a, float b) { return a + (b - a) * g_in.next_float32(); }
static V3 rand_dir() {
    for (;;) {
        V3 v(uni(-1, 1), uni(-1, 1), uni(-1, 1));
        float n = squared_norm(v);
        if (n > 1e-3f && n <= 1.f) return normalize(v);
    }
}
static void pf(float f) {
    if (std::isnan(f)) printf("\"nan\""); else if (std::isinf(f)) printf(f > 0 ? "\"inf\"" : "\"-inf\""); else printf("%.9g", (double) f);
}
static void p3(const V3 &v) { pf(v.x()); printf(", "); pf(v.y()); printf(", "); pf(v.z()); }

struct In { V3 o, d; float len; V3 flux; };
static In make_vrl_in(float spread, float max_len) {
    In v;
    v.o = V3(uni(-spread, spread), uni(-spread, spread), uni(-spread, spread));
    v.d = rand_dir();
    v.len = uni(0.05f, max_len);
    float w = uni(0.05f, 1.f);
    v.flux = V3(uni(0.1f, 2.f), uni(0.1f, 2.f), uni(0.1f, 2.f)) * w;
    return v;
}
static VRLf to_vrl(const In &p, const MockMedium *m) {
    VRLf v(p.o, m, p.flux, 1, 0, false);
    v.direction = p.d; v.length = p.len;
    return v;
}
static void print_in(const In &p) { printf("["); p3(p.o); printf(", "); p3(p.d); printf(", %.9g, ", (double) p.len); p3(p.flux); printf("]"); }

static MockRay make_part(const V3 &o, const V3 &d, float maxt) { MockRay r(o, d, 0.f); r.mint = 0.f; r.maxt = maxt; return r; }
static void print_ray(const MockRay &r) { printf("["); p3(r.o); printf(", "); p3(r.d); printf(", %.9g]", (double) r.maxt); }

static int find_leaf(const std::vector<In> &in, const VRLf &rep) {
    for (size_t j = 0; j < in.size(); ++j)
        if (in[j].o.x() == rep.origin.x() && in[j].o.y() == rep.origin.y() && in[j].o.z() == rep.origin.z() && in[j].len == rep.length) return (int) j;
    return -1;
}
static void print_node(const LC::Node *n, const std::vector<In> &in, bool &first) {
    printf("%s  {\"leaf\": %d, \"rep\": %d, \"flux\": [", first ? "" : ",\n", n->isLeaf ? 1 : 0, find_leaf(in, n->represent)); first = false;
    p3(n->represent.flux); printf("], \"add_flux\": ["); if (n->isLeaf) printf("0, 0, 0"); else p3(n->addFlux);
    printf("], \"same\": %d, \"lo\": [", n->isLeaf ? -1 : (n->isSame[0] ? 0 : 1)); p3(n->aabb.min); printf("], \"hi\": ["); p3(n->aabb.max);
    printf("], \"max_len\": %.9g, \"n_leaves\": %d}", (double) n->maxVRLLenght, (int) n->nodes.size());
    if (!n->isLeaf) { print_node(n->children[0], in, first); print_node(n->children[1], in, first); }
}
static void collect(const LC::Node *n, std::vector<const LC::Node *> &out) {
    out.push_back(n);
    if (!n->isLeaf) { collect(n->children[0], out); collect(n->children[1], out); }
}

int main() {
    MockMedium med;
    med.sigma_t = V3(0.08f, 0.11f, 0.14f); med.albedo = V3(0.9f, 0.8f, 0.7f); med.scale = 1.5f;
    MockMedium med_hg = med; med_hg.phase.hg = true; med_hg.phase.g = 0.6f;
    MockScene empty, occl;
    const float q[][3] = { {-3,-3,0.5f},{3,-3,0.5f},{3,3,0.5f}, {-3,-3,0.5f},{3,3,0.5f},{-3,3,0.5f}, {-6,1,-4},{-1,5,2},{-5,6,-3} };
    for (auto &v : q) occl.tri.push_back(V3(v[0], v[1], v[2]));

    printf("{\n \"medium\": {\"sigma_t\": [0.08, 0.11, 0.14], \"albedo\": [0.9, 0.8, 0.7], \"scale\": 1.5, \"g\": 0.6},\n \"triangles\": [");
    for (size_t i = 0; i < occl.tri.size(); ++i) { if (i) printf(", "); p3(occl.tri[i]); }
    printf("],\n");

    /* ---- dist + merge ---- */
    printf(" \"merge\": [\n");
    const int NM = 60;
    for (int i = 0; i < NM; ++i) {
        In a = make_vrl_in(8.f, 6.f), b = make_vrl_in(8.f, 6.f);
        if (i % 10 == 9) b.flux = V3(0.f);
        LC::Node *n1 = new LC::Node(to_vrl(a, &med)), *n2 = new LC::Node(to_vrl(b, &med));
        float d12 = n1->dist(n2).distance, d21 = n2->dist(n1).distance;
        float v = uni(0.f, 1.f);
        LC::Node *m = new LC::Node(n1, n2, v, true);
        printf("  {\"a\": "); print_in(a); printf(", \"b\": "); print_in(b);
        printf(", \"dist\": [%.9g, %.9g], \"v\": %.9g, \"pick\": %d, \"flux\": [", (double) d12, (double) d21, (double) v, m->isSame[0] ? 0 : 1);
        p3(m->represent.flux); printf("], \"add_flux\": ["); p3(m->addFlux); printf("], \"lo\": ["); p3(m->aabb.min); printf("], \"hi\": ["); p3(m->aabb.max);
        printf("], \"max_len\": %.9g}%s\n", (double) m->maxVRLLenght, i + 1 < NM ? "," : "");
        delete m;
    }
    printf(" ],\n");

    /* ---- box_dist: every direction-sign / zero-component case of the line-box query, rays starting inside, beyond either end ---- */
    printf(" \"box_dist\": [\n");
    const int NB = 240;
    for (int i = 0; i < NB; ++i) {
        V3 c(uni(-5, 5), uni(-5, 5), uni(-5, 5)), h(uni(0.1f, 4.f), uni(0.1f, 4.f), uni(0.1f, 4.f));
        if (i % 24 == 23) h = V3(h.x(), 0.f, h.z());                                       /* flat box: a VRL parallel to an axis plane */
        Box box(c - h, c + h);
        V3 o(uni(-12, 12), uni(-12, 12), uni(-12, 12)), d = rand_dir();
        const int zk = i % 8;                                                               /* zero out direction components */
        if (zk == 1) d = normalize(V3(d.x(), d.y(), 0.f));
        if (zk == 2) d = normalize(V3(d.x(), 0.f, d.z()));
        if (zk == 3) d = normalize(V3(0.f, d.y(), d.z()));
        if (zk == 4) d = V3(d.x() < 0 ? -1.f : 1.f, 0.f, 0.f);
        if (zk == 5) d = V3(0.f, d.y() < 0 ? -1.f : 1.f, 0.f);
        if (zk == 6) d = V3(0.f, 0.f, d.z() < 0 ? -1.f : 1.f);
        if (i % 16 == 7) o = c + h * V3(uni(-0.9f, 0.9f), uni(-0.9f, 0.9f), uni(-0.9f, 0.9f)); /* starts inside */
        if (i % 16 == 15) o = c - d * uni(6.f, 14.f) + rand_dir() * 0.3f;                   /* aims at the box */
        float maxt = (i % 5 == 4) ? uni(0.05f, 2.f) : uni(2.f, 30.f);
        MockRay r = make_part(o, d, maxt);
        float v = box.getMinDistanceSqr(r);
        printf("  {\"lo\": ["); p3(box.min); printf("], \"hi\": ["); p3(box.max); printf("], \"ray\": "); print_ray(r); printf(", \"out\": "); pf(v);
        printf("}%s\n", i + 1 < NB ? "," : "");
    }
    printf(" ],\n");

    /* ---- the tree over N VRLs ---- */
    const int N = 192;
    std::vector<In> in(N);
    for (int j = 0; j < N; ++j) in[j] = make_vrl_in(j % 3 == 0 ? 2.5f : 8.f, j % 7 == 0 ? 9.f : 3.f);
    std::vector<VRLf> vrls;
    for (int j = 0; j < N; ++j) vrls.push_back(to_vrl(in[j], &med));
    printf(" \"vrls\": [");
    for (int j = 0; j < N; ++j) { printf("%s", j ? ", " : ""); print_in(in[j]); }
    printf("],\n \"build_seed\": 4242,\n");
    MockSampler bs; bs.seed(0, 4242);
    LC lc(&empty, vrls, &bs, 8, 0.02f, false, true, false, 1);
    printf(" \"build_draws\": %d,\n \"tree\": [\n", (int) bs.draws);
    { bool first = true; print_node(lc.m_root, in, first); }
    printf("\n ],\n");
    std::vector<const LC::Node *> all;
    collect(lc.m_root, all);

    /* ---- seg_dist ---- */
    printf(" \"seg_dist\": [\n");
    const int ND = 120;
    for (int i = 0; i < ND; ++i) {
        In v = make_vrl_in(8.f, 6.f);
        V3 o(uni(-8, 8), uni(-8, 8), uni(-8, 8)), d = rand_dir();
        if (i % 6 == 5) d = v.d;                                                             /* parallel */
        if (i % 12 == 3) o = v.o + v.d * (0.4f * v.len) - d * 1.5f;                          /* crossing */
        MockRay r = make_part(o, d, uni(0.3f, 20.f));
        float s = lc.sqrDistanceVRL(r, to_vrl(v, &med));
        printf("  {\"vrl\": "); print_in(v); printf(", \"ray\": "); print_ray(r); printf(", \"out\": "); pf(s); printf("}%s\n", i + 1 < ND ? "," : "");
    }
    printf(" ],\n");

    /* ---- rays: single parts and curved (3-part) paths ---- */
    const int NR = 40;
    std::vector<NLf> rays(NR);
    printf(" \"rays\": [\n");
    for (int k = 0; k < NR; ++k) {
        V3 o(uni(-9, 9), uni(-9, 9), uni(-9, 9)), d = rand_dir();
        if (k % 8 == 7) { const In &v = in[(k * 5) % N]; o = v.o + v.d * (0.5f * v.len) - d * 2.f; }   /* through a VRL: bound = max */
        const int n_parts = (k % 5 == 4) ? 3 : 1;
        printf("  [");
        for (int p = 0; p < n_parts; ++p) {
            float maxt = uni(1.f, 18.f) / n_parts;
            MockRay r = make_part(o, d, maxt);
            rays[k].push_back(r);
            printf("%s", p ? ", " : ""); print_ray(r);
            o = o + d * maxt; d = normalize(d + rand_dir() * 0.15f);
        }
        rays[k].wi = -rays[k].parts[0].d;
        printf("]%s\n", k + 1 < NR ? "," : "");
    }
    printf(" ],\n");

    /* ---- bound: every 3rd node of the tree x every 4th ray, for isotropic and HG media ---- */
    printf(" \"bound\": [\n");
    {
        bool first = true;
        for (int hg = 0; hg < 2; ++hg) {
            std::vector<VRLf> vv;
            for (int j = 0; j < N; ++j) vv.push_back(to_vrl(in[j], hg ? &med_hg : &med));
            MockSampler s2; s2.seed(0, 4242);
            LC lc2(&empty, vv, &s2, 8, 0.02f, false, true, false, 1);
            std::vector<const LC::Node *> nodes2; collect(lc2.m_root, nodes2);
            for (size_t a = hg; a < nodes2.size(); a += (hg ? 9 : 3))
                for (int k = (int) (a % 4); k < NR; k += 4) {
                    size_t nb = 0; MockSampler s; s.seed(0, 1);
                    float b = lc2.getClusterUpperBound(&empty, &s, *nodes2[a], rays[k], 0, nb);
                    printf("%s  [%d, %d, %d, ", first ? "" : ",\n", hg, (int) a, k); pf(b); printf("]"); first = false;
                }
        }
    }
    printf("\n ],\n");

    /* ---- query ---- */
    printf(" \"query\": [\n");
    {
        bool first = true;
        const float errs[3] = { 0.02f, 0.1f, 0.5f };
        for (int cfg = 0; cfg < 6; ++cfg) {
            const int uniform = cfg % 2, vis = (cfg / 2) % 2, hg = cfg == 4, better = cfg == 5 ? 0 : 8;
            const float err = errs[cfg % 3];
            std::vector<VRLf> vv;
            for (int j = 0; j < N; ++j) vv.push_back(to_vrl(in[j], hg ? &med_hg : &med));
            MockSampler s2; s2.seed(0, 4242);
            LC lcq(&empty, vv, &s2, better, err, uniform != 0, true, false, 1);
            for (int k = 0; k < NR; ++k) {
                MockSampler s; s.seed(0, 60000 + 100 * cfg + k);
                NLf nl = rays[k];
                LC::LCQuery qq{ nl, &s, 0 };
                size_t nbb = 0;
                V3 Li = lcq.query(vis ? &occl : &empty, qq, nbb, 0);
                printf("%s  {\"cfg\": %d, \"uniform\": %d, \"vis\": %d, \"hg\": %d, \"better\": %d, \"err\": %.9g, \"ray\": %d, \"seed\": %d, \"draws\": %d, \"nb\": %d, \"out\": [",
                       first ? "" : ",\n", cfg, uniform, vis, hg, better, (double) err, k, 60000 + 100 * cfg + k, (int) s.draws, (int) qq.nb_evaluation);
                p3(Li); printf("]}"); first = false;
            }
        }
    }
    printf("\n ],\n \"warnings\": %d\n}\n", MockLog::warnings);
    return 0;
}
