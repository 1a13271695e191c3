// gather_exact.cuh -- the reference's getContrib in its own IEEE operation order (one rounding per written
// operation; translation units including this file are compiled with -fmad=false).  Used by the exact gather
// (gather.cu, math_mode 0) for every eval and by the fast gather (gather_fast.cu, math_mode 1) for the evals
// its conditioning guard routes here.
#pragma once
#include "gather_core.cuh"

namespace vrlb {

// ---- VRL::findClosetPoint, vrl_struct.h:163-228 -----------------------------------------------------
__device__ inline void find_closest_point(V3 vo, V3 vd, float vlen, V3 ro, V3 rd, float rmaxt, float &tCam, float &tVRL) {
    V3 u = vd * vlen;
    V3 v = rd * rmaxt;
    V3 w = vo - ro;
    float a = dot(u, u), b = dot(u, v), c = dot(v, v), d = dot(u, w), e = dot(v, w);
    float D = a * c - b * b;
    float sN, sD = D, tN, tD = D;
    if (D < 0.0001f) {
        sN = 0.0f; sD = 1.0f; tN = e; tD = c;
    } else {
        sN = (b * e - c * d);
        tN = (a * e - b * d);
        if (sN < 0.0f) { sN = 0.0f; tN = e; tD = c; }
        else if (sN > sD) { sN = sD; tN = e + b; tD = c; }
    }
    if (tN < 0.0f) {
        tN = 0.0f;
        if (-d < 0.0f) sN = 0.0f;
        else if (-d > a) sN = sD;
        else { sN = -d; sD = a; }
    } else if (tN > tD) {
        tN = tD;
        if ((-d + b) < 0.0f) sN = 0;
        else if ((-d + b) > a) sN = sD;
        else { sN = (-d + b); sD = a; }
    }
    float sc = (fabsf(sN) < 0.0001f ? 0.0f : sN / sD);
    float tc = (fabsf(tN) < 0.0001f ? 0.0f : tN / tD);
    tCam = tc * rmaxt;
    tVRL = sc * vlen;
}

struct SegRegs { V3 o, d; float maxt; };

// ---- VRL::getContrib (vrl_struct.h:584-662) with samplingVRL (239-360) inlined.  Returns the scalar
//      factor that multiplies flux * sigma_s(cam) * sigma_s(vrl); 0 when the reference returns 0. ----
//      GRID: the VRL's medium may be a grid medium; then `sub_seed` keys the eval's transmittance sub-stream and
//      `st2` receives sigma_t(pCam) * sigma_t(pVRL) (the caller multiplies albedo^2; heterogeneous.cpp:43-51). ----
//      `seg_out` (3 + 3 + 1 floats), when given, receives the connection segment pCam, direction, length and the
//      shadow query is left to the caller (the fast kernel parks it in its warp queue).  PROXIED: the shadow query
//      goes through the solids' conservative proxies first (any_hit_proxied: same answer, fewer BVH walks). ----
template <bool UNIFORM, bool VIS, bool GRID = false, bool PROXIED = false>
__device__ inline float eval_contrib(const DScene &scene, const SegRegs &ray, V3 vo, V3 vd, float vlen,
                                     float max_density, int phase, float g, float u1, float u2,
                                     const DMedium *gm = nullptr, uint64_t sub_seed = 0, float *st2 = nullptr, float *seg_out = nullptr,
                                     float nla_off = 0.f, float nla_total = -1.f) {
    V3 pCam, pVRL; float tCam, tVRL, invPDF;
    if (UNIFORM) {
        // VRL::sampleMC, vrl_struct.h:239-243 (single-part NLRay::at, :44-56)
        if (nla_total >= 0.f) {
            // use_nl_atomic_query: `ray` is one part of a longer NLRay of length nla_total that starts nla_off before it;
            // the eval belongs to the part NLRay::at lands on, every other part's thread contributes 0
            tCam = u1 * nla_total;
            float local = tCam - nla_off;
            if (!(local >= 0.f && local < ray.maxt)) return 0.f;
            pCam = ray.o + ray.d * local;
            invPDF = vlen * nla_total;
        } else {
            tCam = u1 * ray.maxt;
            pCam = ray.o + ray.d * tCam;
            invPDF = vlen * ray.maxt;
        }
        tVRL = u2 * vlen;
        pVRL = vo + vd * tVRL;
    } else {
        float cpCam, cpVRL;
        find_closest_point(vo, vd, vlen, ray.o, ray.d, ray.maxt, cpCam, cpVRL);
        V3 vh = vo + vd * cpVRL;
        V3 uh = ray.o + ray.d * cpCam;
        float h = squared_norm(uh - vh);
        if (h == 0.0f) return 0.f;                               // vrl_struct.h:264
        h = safe_sqrt(h);
        float v0Hat = -cpVRL;
        float v1Hat = vlen + v0Hat;
        float sinTheta = squared_norm(cross(vd, ray.d));
        if (sinTheta == 0) return 0.f;                           // vrl_struct.h:275
        sinTheta = safe_sqrt(sinTheta);
        // inverseCDF_A, vrl_struct.h:293-310 (Novak et al. 2012, eq. 10-13)
        float z0 = (v0Hat / h) * sinTheta, z1 = (v1Hat / h) * sinTheta;
        float a0 = logf(z0 + sqrtf(z0 * z0 + 1));
        float a1 = logf(z1 + sqrtf(z1 * z1 + 1));
        float vHat = (h * sinhf(lerpf(a0, a1, u1))) / sinTheta;
        float invPDF_A = (a1 - a0) * sqrtf(h * h + vHat * vHat * sinTheta * sinTheta);
        invPDF_A /= sinTheta;
        tVRL = (vHat + cpVRL);
        pVRL = vo + vd * tVRL;
        // equi-angular sampling of the camera part, vrl_struct.h:318-336 (Kulla & Fajardo 2012)
        float uHat  = dot(ray.d, (pVRL - ray.o));
        float u0Hat = -uHat;
        float u1Hat = ray.maxt + u0Hat;
        float hPoint = norm(ray.o + ray.d * uHat - pVRL);
        float thetaA = atanf(u0Hat / hPoint);
        float thetaB = atanf(u1Hat / hPoint);
        float uS     = hPoint * tanf(lerpf(thetaA, thetaB, u2));
        float invPDF_B = (thetaB - thetaA) * (hPoint * hPoint + uS * uS) / hPoint;
        tCam = (uS - u0Hat);
        pCam = ray.o + ray.d * tCam;
        tCam = fminf(fmaxf(tCam, 0.0f), ray.maxt);               // vrl_struct.h:340-341
        tVRL = fminf(fmaxf(tVRL, 0.0f), vlen);
        invPDF = invPDF_A * invPDF_B;
    }
    V3 dir = pVRL - pCam;
    float lengthPtoP = norm(dir);
    if (lengthPtoP == 0) return 0.f;                             // vrl_struct.h:589
    if (isnan(invPDF)) return 0.f;                               // vrl_struct.h:594
    dir = dir / lengthPtoP;
    float rayPF = phase_eval(phase, g, -ray.d, dir);             // wi = -part.d (quirk 1)
    if (rayPF == 0) return 0.f;
    float vrlPF = phase_eval(phase, g, -vd, -dir);
    if (vrlPF == 0) return 0.f;
    // VRL::evalTransmittance pCam -> pVRL, vrl_struct.h:509-572
    float vrlToRayTrans, rayTrans, vrlTrans;
    V3 dd = pVRL - pCam;
    float remaining = norm(dd);
    dd = normalize(dd);
    if (seg_out) { seg_out[0] = pCam.x; seg_out[1] = pCam.y; seg_out[2] = pCam.z; seg_out[3] = dd.x; seg_out[4] = dd.y; seg_out[5] = dd.z; seg_out[6] = remaining; }
    if (VIS && remaining > 0 && (PROXIED ? any_hit_proxied(scene, pCam, dd, 0.f, remaining) : any_hit(scene, pCam, dd, 0.f, remaining))) return 0.f;
    if (GRID && gm->kind == VRLB_MEDIUM_GRID) {
        // get_scattering_coefficients at both points (vrl_struct.h:616, 624), then the three stochastic
        // transmittances in the reference's call order (:632, :641, :647) on the eval's sub-stream
        *st2 = (gm->scale * grid_eval(*gm, pCam)) * (gm->scale * grid_eval(*gm, pVRL));
        Pcg32s sub; sub.seed(sub_seed, PCG32_DEFAULT_STREAM + 1u);
        vrlToRayTrans = remaining > 0 ? grid_transmittance(*gm, pCam, dd, remaining, sub) : 1.0f;
        if (vrlToRayTrans == 0.0f) return 0.f;
        rayTrans = grid_transmittance(*gm, ray.o, ray.d, tCam, sub);
        if (rayTrans == 0.0f) return 0.f;
        vrlTrans = grid_transmittance(*gm, vo, vd, tVRL, sub);
        if (vrlTrans == 0.0f) return 0.f;
    } else {
        vrlToRayTrans = remaining > 0 ? homo_transmittance(max_density, remaining) : 1.0f;
        if (vrlToRayTrans == 0.0f) return 0.f;
        rayTrans = homo_transmittance(max_density, tCam);
        if (rayTrans == 0.0f) return 0.f;
        vrlTrans = homo_transmittance(max_density, tVRL);
        if (vrlTrans == 0.0f) return 0.f;
    }
    float fallOff = 1.0f / (lengthPtoP * lengthPtoP);
    return fallOff * vrlPF * rayPF * vrlTrans * rayTrans * vrlToRayTrans * invPDF;
}

}  // namespace vrlb
