// occluder_proxies.cpp -- conservative proxies of a connected mesh component for the shadow (any-hit) query.
//
// The reference's visibility test (VRL::evalTransmittance, src/integrators/vrl/vrl_struct.h:509-572) returns 0 as
// soon as the connection segment meets ANY surface.  For a *closed* component that lets a segment be classified
// without touching its triangles:
//   * outer proxies (bounding sphere, oriented bounding box): a segment that misses one of them misses the mesh;
//   * inner proxies (inscribed sphere, inscribed oriented box; both strictly inside the solid): a segment that has an
//     end point outside the component's AABB and meets an inner proxy passes from outside the closed surface to a
//     point enclosed by it, so it crosses the surface -- occluded.
// Only segments in the thin layer between the inner and the outer proxies go on to the BVH.  All proxies are built
// here from the triangles alone (no primitive-type knowledge), in double precision, and are pulled in / pushed out
// by `margin` so that fp32 evaluation on the device stays conservative.
#include <algorithm>
#include <array>
#include <cmath>
#include <cstring>
#include <map>
#include <vector>

#include "bvh_build.h"

namespace vrlb {
namespace {

struct D3 { double x, y, z; };
inline D3 operator+(D3 a, D3 b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
inline D3 operator-(D3 a, D3 b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
inline D3 operator*(D3 a, double s) { return { a.x * s, a.y * s, a.z * s }; }
inline double dot(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline D3 cross(D3 a, D3 b) { return { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }
inline double comp(D3 a, int k) { return k == 0 ? a.x : (k == 1 ? a.y : a.z); }

struct Tri { D3 a, b, c; };

// every undirected edge (vertex positions compared bit-exactly) is shared by exactly two triangles
bool is_closed(const float *verts, const std::vector<uint32_t> &sub) {
    using Key = std::array<uint32_t, 6>;
    std::map<Key, int> edges;
    auto bits = [](float f) { if (f == 0.f) f = 0.f; uint32_t u; std::memcpy(&u, &f, 4); return u; };
    for (uint32_t t : sub) {
        const float *p = verts + (size_t) t * 9;
        for (int e = 0; e < 3; ++e) {
            const float *u = p + 3 * e, *v = p + 3 * ((e + 1) % 3);
            std::array<uint32_t, 3> ku = { bits(u[0]), bits(u[1]), bits(u[2]) }, kv = { bits(v[0]), bits(v[1]), bits(v[2]) };
            if (ku == kv) return false;                       // degenerate edge
            if (kv < ku) std::swap(ku, kv);
            edges[Key{ ku[0], ku[1], ku[2], kv[0], kv[1], kv[2] }]++;
        }
    }
    for (auto &kv : edges) if (kv.second != 2) return false;
    return !edges.empty();
}

// parity of the crossings of a ray from p along dir (double-precision Moeller-Trumbore)
int crossings(const std::vector<Tri> &tris, D3 p, D3 dir) {
    int n = 0;
    for (const Tri &t : tris) {
        D3 e1 = t.b - t.a, e2 = t.c - t.a, pv = cross(dir, e2);
        double det = dot(e1, pv);
        if (std::fabs(det) < 1e-300) continue;
        double inv = 1.0 / det;
        D3 tv = p - t.a;
        double u = dot(tv, pv) * inv;
        if (u < 0.0 || u > 1.0) continue;
        D3 qv = cross(tv, e1);
        double v = dot(dir, qv) * inv;
        if (v < 0.0 || u + v > 1.0) continue;
        if (dot(e2, qv) * inv > 0.0) ++n;
    }
    return n;
}
bool inside(const std::vector<Tri> &tris, D3 p) {
    // three generic directions must agree (a ray through an edge could be counted twice)
    const D3 dirs[3] = { { 0.5773, 0.3015, 0.7589 }, { -0.4242, 0.8111, 0.4028 }, { 0.1237, -0.6419, -0.7567 } };
    int votes = 0;
    for (const D3 &d : dirs) votes += crossings(tris, p, d) & 1;
    return votes == 3;
}

// distance from p to triangle (closest point on triangle, Ericson 5.1.5)
double dist_point_tri(D3 p, const Tri &t) {
    D3 ab = t.b - t.a, ac = t.c - t.a, ap = p - t.a;
    double d1 = dot(ab, ap), d2 = dot(ac, ap);
    D3 q;
    if (d1 <= 0 && d2 <= 0) q = t.a;
    else {
        D3 bp = p - t.b; double d3 = dot(ab, bp), d4 = dot(ac, bp);
        if (d3 >= 0 && d4 <= d3) q = t.b;
        else {
            double vc = d1 * d4 - d3 * d2;
            if (vc <= 0 && d1 >= 0 && d3 <= 0) q = t.a + ab * (d1 / (d1 - d3));
            else {
                D3 cp = p - t.c; double d5 = dot(ab, cp), d6 = dot(ac, cp);
                if (d6 >= 0 && d5 <= d6) q = t.c;
                else {
                    double vb = d5 * d2 - d1 * d6;
                    if (vb <= 0 && d2 >= 0 && d6 <= 0) q = t.a + ac * (d2 / (d2 - d6));
                    else {
                        double va = d3 * d6 - d5 * d4;
                        if (va <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0) q = t.b + (t.c - t.b) * ((d4 - d3) / ((d4 - d3) + (d5 - d6)));
                        else { double den = 1.0 / (va + vb + vc); q = t.a + ab * (vb * den) + ac * (vc * den); }
                    }
                }
            }
        }
    }
    D3 d = p - q;
    return std::sqrt(dot(d, d));
}

// triangle (in box-local coordinates, box centred at the origin) vs box of half extents h: separating-axis test
bool tri_box_overlap(const Tri &t, const double h[3]) {
    const D3 v[3] = { t.a, t.b, t.c };
    for (int k = 0; k < 3; ++k) {
        double mn = std::min({ comp(v[0], k), comp(v[1], k), comp(v[2], k) }), mx = std::max({ comp(v[0], k), comp(v[1], k), comp(v[2], k) });
        if (mn > h[k] || mx < -h[k]) return false;
    }
    const D3 e[3] = { v[1] - v[0], v[2] - v[1], v[0] - v[2] };
    D3 n = cross(e[0], e[1]);
    {
        double r = h[0] * std::fabs(n.x) + h[1] * std::fabs(n.y) + h[2] * std::fabs(n.z);
        double s = dot(n, v[0]);
        if (std::fabs(s) > r) return false;
    }
    const D3 axes[3] = { { 1, 0, 0 }, { 0, 1, 0 }, { 0, 0, 1 } };
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            D3 a = cross(axes[i], e[j]);
            double p0 = dot(a, v[0]), p1 = dot(a, v[1]), p2 = dot(a, v[2]);
            double r = h[0] * std::fabs(a.x) + h[1] * std::fabs(a.y) + h[2] * std::fabs(a.z);
            if (std::min({ p0, p1, p2 }) > r || std::max({ p0, p1, p2 }) < -r) return false;
        }
    return true;
}

// symmetric 3x3 eigenvectors by cyclic Jacobi; columns of V
void jacobi3(double A[3][3], double V[3][3]) {
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) V[i][j] = i == j;
    for (int sweep = 0; sweep < 32; ++sweep) {
        double off = std::fabs(A[0][1]) + std::fabs(A[0][2]) + std::fabs(A[1][2]);
        if (off < 1e-14 * (std::fabs(A[0][0]) + std::fabs(A[1][1]) + std::fabs(A[2][2]) + 1e-300)) break;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                if (std::fabs(A[p][q]) < 1e-300) continue;
                double theta = (A[q][q] - A[p][p]) / (2.0 * A[p][q]);
                double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
                double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < 3; ++k) { double akp = A[k][p], akq = A[k][q]; A[k][p] = c * akp - s * akq; A[k][q] = s * akp + c * akq; }
                for (int k = 0; k < 3; ++k) { double apk = A[p][k], aqk = A[q][k]; A[p][k] = c * apk - s * aqk; A[q][k] = s * apk + c * aqk; }
                for (int k = 0; k < 3; ++k) { double vkp = V[k][p], vkq = V[k][q]; V[k][p] = c * vkp - s * vkq; V[k][q] = s * vkp + c * vkq; }
            }
    }
}

}  // namespace

void build_solid_proxies(const float *verts, const std::vector<uint32_t> &sub, float margin, DSolid &out) {
    std::memset(&out, 0, sizeof(out));
    std::vector<Tri> tris; tris.reserve(sub.size());
    D3 lo = { 1e300, 1e300, 1e300 }, hi = { -1e300, -1e300, -1e300 }, mean = { 0, 0, 0 };
    for (uint32_t t : sub) {
        const float *p = verts + (size_t) t * 9;
        Tri tr = { { p[0], p[1], p[2] }, { p[3], p[4], p[5] }, { p[6], p[7], p[8] } };
        tris.push_back(tr);
        for (const D3 &v : { tr.a, tr.b, tr.c }) {
            lo = { std::min(lo.x, v.x), std::min(lo.y, v.y), std::min(lo.z, v.z) };
            hi = { std::max(hi.x, v.x), std::max(hi.y, v.y), std::max(hi.z, v.z) };
            mean = mean + v;
        }
    }
    const double nv = 3.0 * (double) tris.size();
    mean = mean * (1.0 / nv);
    const double m = margin;
    out.lo[0] = (float) (lo.x - m); out.lo[1] = (float) (lo.y - m); out.lo[2] = (float) (lo.z - m);
    out.hi[0] = (float) (hi.x + m); out.hi[1] = (float) (hi.y + m); out.hi[2] = (float) (hi.z + m);
    // bounding sphere: the better of AABB centre / vertex mean
    D3 cands[2] = { (lo + hi) * 0.5, mean };
    double best_r = 1e300; D3 best_c = cands[0];
    for (const D3 &c : cands) {
        double r = 0;
        for (const Tri &t : tris) for (const D3 &v : { t.a, t.b, t.c }) r = std::max(r, std::sqrt(dot(v - c, v - c)));
        if (r < best_r) { best_r = r; best_c = c; }
    }
    out.c[0] = (float) best_c.x; out.c[1] = (float) best_c.y; out.c[2] = (float) best_c.z;
    out.r_out2 = (float) ((best_r + m) * (best_r + m) * (1.0 + 1e-5));
    // oriented box: the tightest of a few candidate frames -- PCA of the vertices, PCA of the area-weighted normals, and
    // (normal, edge) frames of the largest triangles (exact for box-like meshes, whatever their rotation)
    std::vector<std::array<D3, 3>> frames;
    auto add_frame = [&](D3 a0, D3 a1) {
        double l0 = std::sqrt(dot(a0, a0)); if (l0 < 1e-30) return;
        a0 = a0 * (1.0 / l0);
        a1 = a1 - a0 * dot(a0, a1);
        double l1 = std::sqrt(dot(a1, a1)); if (l1 < 1e-12) return;
        a1 = a1 * (1.0 / l1);
        frames.push_back({ a0, a1, cross(a0, a1) });
    };
    {
        double C[3][3] = { { 0 } }, N[3][3] = { { 0 } }, V[3][3];
        for (const Tri &t : tris) {
            for (const D3 &v : { t.a, t.b, t.c }) {
                D3 d = v - mean; const double dd[3] = { d.x, d.y, d.z };
                for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) C[i][j] += dd[i] * dd[j];
            }
            D3 nn = cross(t.b - t.a, t.c - t.a);                      // length = 2 * area
            double l = std::sqrt(dot(nn, nn)); if (l < 1e-300) continue;
            const double nd[3] = { nn.x, nn.y, nn.z };
            for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) N[i][j] += nd[i] * nd[j] / l;
        }
        jacobi3(C, V); add_frame({ V[0][0], V[1][0], V[2][0] }, { V[0][1], V[1][1], V[2][1] });
        jacobi3(N, V); add_frame({ V[0][0], V[1][0], V[2][0] }, { V[0][1], V[1][1], V[2][1] });
        std::vector<std::pair<double, size_t>> by_area;
        for (size_t i = 0; i < tris.size(); ++i) { D3 nn = cross(tris[i].b - tris[i].a, tris[i].c - tris[i].a); by_area.push_back({ -dot(nn, nn), i }); }
        std::sort(by_area.begin(), by_area.end());
        for (size_t k = 0; k < by_area.size() && k < 4; ++k) {
            const Tri &t = tris[by_area[k].second];
            D3 nn = cross(t.b - t.a, t.c - t.a);
            add_frame(nn, t.b - t.a); add_frame(nn, t.c - t.b); add_frame(nn, t.a - t.c);
        }
        frames.push_back({ D3{ 1, 0, 0 }, D3{ 0, 1, 0 }, D3{ 0, 0, 1 } });
    }
    D3 ax[3] = { { 1, 0, 0 }, { 0, 1, 0 }, { 0, 0, 1 } };
    double pmin[3], pmax[3], best_vol = 1e300;
    for (const auto &fr : frames) {
        double mn[3] = { 1e300, 1e300, 1e300 }, mx[3] = { -1e300, -1e300, -1e300 };
        for (const Tri &t : tris) for (const D3 &v : { t.a, t.b, t.c })
            for (int k = 0; k < 3; ++k) { double s = dot(v, fr[k]); mn[k] = std::min(mn[k], s); mx[k] = std::max(mx[k], s); }
        double vol = (mx[0] - mn[0]) * (mx[1] - mn[1]) * (mx[2] - mn[2]);
        if (vol < best_vol * (1.0 - 1e-9)) { best_vol = vol; for (int k = 0; k < 3; ++k) { ax[k] = fr[k]; pmin[k] = mn[k]; pmax[k] = mx[k]; } }
    }
    D3 oc = { 0, 0, 0 }; double h[3];
    for (int k = 0; k < 3; ++k) { oc = oc + ax[k] * (0.5 * (pmin[k] + pmax[k])); h[k] = 0.5 * (pmax[k] - pmin[k]); }
    for (int k = 0; k < 3; ++k) {
        out.ax[3 * k + 0] = (float) ax[k].x; out.ax[3 * k + 1] = (float) ax[k].y; out.ax[3 * k + 2] = (float) ax[k].z;
        out.h_out[k] = (float) ((h[k] + m) * (1.0 + 1e-5));
    }
    out.oc[0] = (float) oc.x; out.oc[1] = (float) oc.y; out.oc[2] = (float) oc.z;
    const double vol_obb = h[0] * h[1] * h[2] * 8, vol_aabb = (hi.x - lo.x) * (hi.y - lo.y) * (hi.z - lo.z);
    out.flags = 0;
    if (vol_obb < 0.8 * vol_aabb) out.flags |= VRLB_SOLID_OUTER_OBB;   // worth a transform only when it is tighter than the AABB
    // inner proxies need a closed surface
    if (!is_closed(verts, sub)) return;
    {   // inscribed sphere: best of a few interior candidates
        const D3 ic[3] = { best_c, oc, mean };
        double rin = 0; D3 cin = ic[0];
        for (const D3 &c : ic) {
            if (!inside(tris, c)) continue;
            double r = 1e300;
            for (const Tri &t : tris) r = std::min(r, dist_point_tri(c, t));
            if (r > rin) { rin = r; cin = c; }
        }
        rin -= m;
        if (rin > 0) {
            out.ci[0] = (float) cin.x; out.ci[1] = (float) cin.y; out.ci[2] = (float) cin.z;
            out.r_in2 = (float) (rin * rin * (1.0 - 1e-5));
            out.flags |= VRLB_SOLID_INNER_SPHERE;
        }
    }
    if (inside(tris, oc)) {   // inscribed oriented box: the outer box scaled about its centre until it is free of triangles
        std::vector<Tri> local(tris.size());
        for (size_t i = 0; i < tris.size(); ++i) {
            auto tl = [&](D3 v) { D3 d = v - oc; return D3{ dot(d, ax[0]), dot(d, ax[1]), dot(d, ax[2]) }; };
            local[i] = { tl(tris[i].a), tl(tris[i].b), tl(tris[i].c) };
        }
        auto free_of_tris = [&](double s) {
            const double hs[3] = { h[0] * s, h[1] * s, h[2] * s };
            for (const Tri &t : local) if (tri_box_overlap(t, hs)) return false;
            return true;
        };
        double a = 0.0, b = 1.0;
        if (free_of_tris(1e-3)) {
            a = 1e-3;
            for (int it = 0; it < 30; ++it) { double mid = 0.5 * (a + b); if (free_of_tris(mid)) a = mid; else b = mid; }
            double hin[3]; bool ok = true;
            for (int k = 0; k < 3; ++k) { hin[k] = h[k] * a - m; ok = ok && hin[k] > 0; }
            // keep it only when it adds to the inscribed sphere (volume comparison)
            double rin = std::sqrt((double) out.r_in2);
            if (ok && hin[0] * hin[1] * hin[2] * 8 > 0.5 * (4.18879 * rin * rin * rin)) {
                for (int k = 0; k < 3; ++k) out.h_in[k] = (float) (hin[k] * (1.0 - 1e-5));
                out.flags |= VRLB_SOLID_INNER_OBB;
            }
        }
    }
    // Every proxy is conservative on its own, so dropping one only sends a few more segments to the traversal.  Keep the
    // per-eval test short where a proxy adds nothing (a warp pays for the longest path any of its 32 lanes takes):
    //  * sphere-like solid: the layer between the inscribed and the bounding sphere is thin -> no oriented boxes;
    //  * box-like solid: the inscribed sphere lies (nearly) inside the inscribed box -> no inscribed sphere.
    if ((out.flags & VRLB_SOLID_INNER_SPHERE) && out.r_in2 >= 0.85f * out.r_out2)
        out.flags &= ~(VRLB_SOLID_OUTER_OBB | VRLB_SOLID_INNER_OBB);
    if ((out.flags & VRLB_SOLID_INNER_SPHERE) && (out.flags & VRLB_SOLID_INNER_OBB)) {
        const double rin = std::sqrt((double) out.r_in2);
        bool inside_box = true;
        for (int k = 0; k < 3; ++k) {
            double off = 0;
            for (int c = 0; c < 3; ++c) off += ((double) out.ci[c] - (double) out.oc[c]) * (double) out.ax[3 * k + c];
            inside_box = inside_box && std::fabs(off) + 0.9 * rin <= (double) out.h_in[k];
        }
        if (inside_box) out.flags &= ~VRLB_SOLID_INNER_SPHERE;
    }
}

}  // namespace vrlb
