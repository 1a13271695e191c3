// generate.cu -- VRL generation by GPU light tracing (north-star subsystem 1).
//
// Replaces the single-threaded light tracer of VRLIntegrator::preprocess (src/integrators/vrl/vrl.cpp:117-374)
// and the cell-by-cell Snell refraction of NonLinearMedia (src/media/nonlinear.cpp:244-405): one thread per emitted
// particle, each with its own PCG32 stream seed(base_seed + particle index) (the reference runs one serial stream;
// generation is compared statistically).  Inside a graded-IOR ("nonlinear") medium the particle is
// marched with a FIXED-STEP EIKONAL integrator
//        v = n * dir,   dv/ds = grad n,   dx/ds = v / n          (semi-implicit Euler, step = one IOR layer)
// through n(y) = lerp(bottom_ior, top_ior, (y - y0) / height) (nonlinear.cpp:129-153, method 0) and the curved path is
// emitted as piecewise-linear VRLs: a segment is cut when the direction has changed and the segment is longer than
// min_vrl_length (vrl.cpp:215-223), at every real scattering event (vrl.cpp:278-284) and at every surface
// (vrl.cpp:313-314).  Homogeneous media take one straight flight per event.
//
// Events (vrl.cpp:180-357): Russian roulette after rr_depth; free-flight distance t = -log(1 - u) / sigma_maj with the
// grey majorant (medium.cpp:36-75; tr/pdf cancels for a homogeneous extinction); real scatter: throughput *= albedo,
// new direction from the phase function (isotropic.cpp / hg.cpp); null-BSDF surfaces are crossed and switch the
// medium; a smooth surface reflects the particle diffusely (SmoothDiffuse::sample: cosine lobe, weight = the
// triangle's reflectance from vrlb_scene_desc::tri_reflectance; without that table particles end there); an emitter
// surface ends it before the VRL in flight is stored (vrl.cpp:309-310).  A VRL carries the flux and depth it was
// created with, so a roulette survival boosts the VRLs created afterwards, not the one in flight -- and a roulette
// death drops the VRL in flight (vrl.cpp:181-185 runs before :283/:313): the reference's set is reproduced, energy loss
// included.
// Flux: laser = power; area emitter = Le * pi * area * pi (the second pi is vrl.cpp:151-152).
//
// Output is deterministic for a given (seed, particle count): pass 1 counts the VRLs of every particle, an exclusive
// scan gives each particle its slot range, pass 2 re-traces and writes.  Records are written as 48 B DVrl.
#include "vrlb_device.cuh"
#include "vrlb_kernels.h"

namespace vrlb {

constexpr float kPi = 3.14159265358979323846f;

// one eikonal step: nl_bend (vrlb_device.cuh)
__device__ inline V3 eikonal_bend(const DMedium &m, V3 p, V3 dir, float ds) { return nl_bend(m, p, dir, ds); }

__device__ inline V3 sample_uniform_sphere(float u1, float u2) {         // warp::square_to_uniform_sphere
    float z = 1.f - 2.f * u2, r = sqrtf(fmaxf(1.f - z * z, 0.f));
    float s, c; sincosf(2.f * kPi * u1, &s, &c);
    return V3(r * c, r * s, z);
}
__device__ inline V3 sample_phase(const DMedium &m, V3 wi /* direction of travel */, float u1, float u2) {
    if (m.phase == VRLB_PHASE_ISOTROPIC) return sample_uniform_sphere(u1, u2);
    // Henyey-Greenstein about the direction of travel (hg.cpp:56-75: cos_theta relative to -wi = travel direction)
    float g = m.g, cos_theta;
    if (fabsf(g) < 1e-3f) cos_theta = 1.f - 2.f * u1;
    else { float sq = (1.f - g * g) / (1.f - g + 2.f * g * u1); cos_theta = (1.f + g * g - sq * sq) / (2.f * g); }
    float sin_theta = sqrtf(fmaxf(1.f - cos_theta * cos_theta, 0.f));
    float s, c; sincosf(2.f * kPi * u2, &s, &c);
    V3 a, b; coordinate_system(wi, a, b);
    return normalize(a * (sin_theta * c) + b * (sin_theta * s) + wi * cos_theta);
}

struct Emit {
    V3 o, d, flux; int medium;
};

__device__ inline Emit emit_particle(const GenParams &p, Pcg32 &rng) {
    Emit e; e.medium = p.start_medium;
    float u0 = rng.next_f32(), u1 = rng.next_f32(), u2 = rng.next_f32(), u3 = rng.next_f32(), u4 = rng.next_f32();
    if (p.use_laser || p.n_emit_tris == 0) {                               // vrl.cpp:156-162
        e.o = V3(p.laser_o[0], p.laser_o[1], p.laser_o[2]);
        e.d = normalize(V3(p.laser_d[0], p.laser_d[1], p.laser_d[2]));
        e.flux = V3(p.laser_power[0], p.laser_power[1], p.laser_power[2]);
        return e;
    }
    // area emitter (area.cpp:74-117): triangle proportional to area (cdf), uniform point, cosine direction
    int lo = 0, hi = p.n_emit_tris - 1;
    float target = u0 * p.emit_area;
    while (lo < hi) { int mid = (lo + hi) >> 1; if (p.emit_cdf[mid] < target) lo = mid + 1; else hi = mid; }
    const DTri &t = p.scene.tris[p.emit_tris[lo]];
    float su = sqrtf(u1), b0 = 1.f - su, b1 = u2 * su;                      // warp::square_to_uniform_triangle
    V3 p0(t.p0[0], t.p0[1], t.p0[2]), e1(t.e1[0], t.e1[1], t.e1[2]), e2(t.e2[0], t.e2[1], t.e2[2]);
    V3 n = tri_normal(t);
    e.o = p0 + e1 * b1 + e2 * (1.f - b0 - b1);
    float r = sqrtf(u3), s, c; sincosf(2.f * kPi * u4, &s, &c);           // cosine hemisphere
    V3 a, b; coordinate_system(n, a, b);
    e.d = normalize(a * (r * c) + b * (r * s) + n * sqrtf(fmaxf(1.f - u3, 0.f)));
    int id = VRLB_SURF_GET_EMITTER(t.flags);
    V3 Le(p.scene.emitters[id * 4], p.scene.emitters[id * 4 + 1], p.scene.emitters[id * 4 + 2]);
    e.flux = Le * (kPi * p.emit_area * kPi);                               // Le * pi / pdf_pos, then * pi (vrl.cpp:151-152)
    e.o = e.o + e.d * ((1.f + hmax_abs(e.o)) * kRayEpsilon);
    return e;
}

template <bool WRITE>
__global__ void __launch_bounds__(128) generate_kernel(GenParams p) {
    uint64_t pi = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (pi >= p.n_particles) return;
    const DScene &s = p.scene;
    Pcg32 rng; rng.seed(p.base_seed + p.first_particle + pi);
    uint32_t count = 0;
    uint64_t slot = WRITE ? p.offsets[pi] : 0;
    const uint64_t slot_end = WRITE ? p.offsets[pi + 1] : 0;

    Emit e = emit_particle(p, rng);
    (void) rng.next_f32();                                                 // channel draw, once per particle (vrl.cpp:172-176)
    V3 pos = e.o, dir = e.d, throughput(1.f);
    // depth counts real events (medium scatterings + reflections) = the reference's `depth` - 1 (it starts at 1,
    // vrl.cpp:171); medium_depth is its mediumDepth (vrl.cpp:164)
    int medium = e.medium, depth = 0, medium_depth = 0;
    bool was_transmitted = false;                                          // vrl.cpp:165, 345: the last surface event was a transmission
    // a VRL keeps the flux and depth it was created with (VRL(ray.o, medium, throughput * flux, depth ...), vrl.cpp:178,
    // 283, 354): the roulette's 1/q of a later iteration does not reach the VRL already in flight
    V3 vrl_o = pos, vrl_d0 = dir, vrl_flux = throughput * e.flux; int vrl_depth = 0; bool vrl_open = medium >= 0;
    auto open_vrl = [&](V3 at, V3 d0) { vrl_o = at; vrl_d0 = d0; vrl_flux = throughput * e.flux; vrl_depth = depth; vrl_open = medium >= 0; };
    auto push = [&](V3 end) {                                               // VRL::setEndPoint + VRLMap::push_back
        if (p.collect_surface) return;
        V3 dd = end - vrl_o; float len = norm(dd);
        if (!(len > 0.f) || medium < 0 || (vrl_flux.x == 0.f && vrl_flux.y == 0.f && vrl_flux.z == 0.f)) return;
        if (WRITE && slot < slot_end) {
            DVrl v; V3 dn = dd / len;
            v.ox = vrl_o.x; v.oy = vrl_o.y; v.oz = vrl_o.z; v.length = len;
            v.dx = dn.x; v.dy = dn.y; v.dz = dn.z; v.medium = (uint32_t) medium;
            v.fr = vrl_flux.x; v.fg = vrl_flux.y; v.fb = vrl_flux.z; v.pad = __int_as_float(vrl_depth);
            p.out[slot++] = v;
        }
        ++count;
    };

    for (int bounce = 0; bounce < 100000 && count < p.max_per_particle; ++bounce) {
        if (throughput.x == 0.f && throughput.y == 0.f && throughput.z == 0.f) break;
        {                                                                  // Russian roulette, vrl.cpp:181-185: the draw is
            float q = fminf(fmaxf(throughput.x, fmaxf(throughput.y, throughput.z)), 0.95f);   // taken every iteration
            const float u = rng.next_f32();
            if (depth + 1 > p.rr_depth) { if (!(u < q)) break; throughput = throughput * (1.f / q); }
        }
        if (depth + 1 >= p.max_depth) break;                               // vrl.cpp:187-189
        // distance to the next medium event
        float t_event = VRLB_INF;
        const DMedium *m = medium >= 0 ? &s.media[medium] : nullptr;
        if (m) {
            float u = rng.next_f32();                                      // sample_interaction(ray, next_1d, channel), vrl.cpp:206
            t_event = m->max_density > 0.f ? -logf(1.f - u) / m->max_density : VRLB_INF;
        }
        const bool curved = m && m->kind == VRLB_MEDIUM_NONLINEAR && m->nl_step > 0.f && p.use_non_linear;
        // flight: straight (one closest-hit) or marched in eikonal steps
        bool surface = false; Hit si; si.tri = -1; si.t = VRLB_INF;
        float travelled = 0.f;
        if (!curved) {
            si = closest_hit(s, pos, dir, 0.f, t_event);
            surface = si.tri >= 0;
            float t = surface ? si.t : t_event;
            if (!(t < VRLB_INF)) break;                                    // escaped
            pos = pos + dir * t; travelled = t;
        } else {
            for (int step = 0; step < p.max_steps; ++step) {
                float ds = fminf(m->nl_step, t_event - travelled);
                si = closest_hit(s, pos, dir, 0.f, ds);
                if (si.tri >= 0) { pos = pos + dir * si.t; travelled += si.t; surface = true; break; }
                pos = pos + dir * ds; travelled += ds;
                if (travelled >= t_event) break;
                V3 nd = eikonal_bend(*m, pos, dir, ds);
                // cut a VRL when the direction has changed and the piece is long enough (vrl.cpp:215-223)
                if (vrl_open && dot(nd, vrl_d0) < 1.f - 1e-7f && norm(pos - vrl_o) > p.min_vrl_length) {
                    push(pos);
                    open_vrl(pos, nd);
                }
                dir = nd;
            }
            if (!surface && travelled < t_event) break;                     // step budget exhausted
        }
        if (surface) {
            const DTri &tr = s.tris[si.tri];
            if (tr.flags & VRLB_SURF_EMITTER) break;                        // si.shape->is_emitter(): vrl.cpp:309-310, before the VRL is stored
            if (vrl_open) push(pos);                                        // vrl.cpp:312-313
            // handleSurfaceInteraction (vrl.cpp:315, 772-782): a smooth BSDF stores a photon (position, travel direction,
            // flux * throughput before the reflection) whichever side it is met from; !wasTransmitted -> the global map
            const uint32_t bsdf_kind = tr.flags & VRLB_SURF_BSDF_MASK;
            // collect_surface 1 = the global map (!wasTransmitted), 2 = the caustic map (wasTransmitted), vrl.cpp:776-780
            if (p.collect_surface && p.tri_refl && bsdf_kind == VRLB_SURF_BSDF_SMOOTH && was_transmitted == (p.collect_surface == 2)) {
                const float *rf0 = p.tri_refl + (size_t) tr.prim * 3;
                if (__ldg(rf0) != 0.f || __ldg(rf0 + 1) != 0.f || __ldg(rf0 + 2) != 0.f) {   // reflectance 0 stands for a BSDF without a smooth lobe
                    if (WRITE && slot < slot_end) {
                        const V3 w = throughput * e.flux;
                        p.surf_out[slot++] = DPhoton{ pos.x, pos.y, pos.z, 0.f, dir.x, dir.y, dir.z, 0.f, w.x, w.y, w.z, 0.f };
                    }
                    ++count;
                }
            }
            const float u1 = rng.next_f32(), u2 = rng.next_f32(), u3 = rng.next_f32();     // bsdf->sample(ctx, si, next_1d, next_2d), vrl.cpp:320
            if (bsdf_kind == VRLB_SURF_BSDF_NULL) {
                was_transmitted = false;                                    // non_null_bsdf && Transmission, vrl.cpp:345
                if (VRLB_SURF_GET_EXTERIOR(tr.flags) >= 0 || VRLB_SURF_GET_INTERIOR(tr.flags) >= 0) medium = target_medium(tr, dir);
                pos = pos + dir * ((1.f + hmax_abs(pos)) * kRayEpsilon);    // spawn_ray offset
                open_vrl(pos, dir);
                continue;
            }
            if (bsdf_kind != VRLB_SURF_BSDF_SMOOTH) {
                // specular chain on the light side (vrl.cpp:318-358): SmoothConductor / SmoothDielectric::sample
                if (!p.tri_refl) break;
                const float *rf = p.tri_refl + (size_t) tr.prim * 3;
                const DeltaSample ds = sample_delta(bsdf_kind, __ldg(rf), __ldg(rf + 1), __ldg(rf + 2), (tr.flags & VRLB_SURF_TWOSIDED) != 0u,
                                                    tri_normal(tr), dir, u1);
                if (!ds.valid) break;
                throughput = throughput * ds.weight;
                if (throughput.x == 0.f && throughput.y == 0.f && throughput.z == 0.f) break;
                dir = ds.wo;
                ++depth;
                was_transmitted = ds.transmitted;
                if (was_transmitted) medium_depth = 0;                      // vrl.cpp:346-347
                if (VRLB_SURF_GET_EXTERIOR(tr.flags) >= 0 || VRLB_SURF_GET_INTERIOR(tr.flags) >= 0) medium = target_medium(tr, dir);
                pos = pos + dir * ((1.f + hmax_abs(pos)) * kRayEpsilon);
                open_vrl(pos, dir);
                continue;
            }
            was_transmitted = false;
            // SmoothDiffuse::sample (src/bsdfs/diffuse.cpp): nothing behind a one-sided surface, else a cosine-distributed
            // direction about the normal (warp::square_to_cosine_hemisphere) with weight = reflectance
            V3 n = tri_normal(tr);
            float cos_i = -dot(dir, n);
            if (cos_i < 0.f && (tr.flags & VRLB_SURF_TWOSIDED)) { n = -n; cos_i = -cos_i; }
            if (!(cos_i > 0.f) || !p.tri_refl) break;
            const float *rf = p.tri_refl + (size_t) tr.prim * 3;
            throughput = throughput * V3(__ldg(rf), __ldg(rf + 1), __ldg(rf + 2));
            if (throughput.x == 0.f && throughput.y == 0.f && throughput.z == 0.f) break;
            float dx, dy;                                                   // warp::square_to_uniform_disk_concentric
            {
                const float x = 2.f * u2 - 1.f, y = 2.f * u3 - 1.f;
                const bool quad = fabsf(x) < fabsf(y);
                const float r = quad ? y : x, rp = quad ? x : y;
                float phi = 0.25f * kPi * (r != 0.f ? rp / r : 0.f);
                if (quad) phi = 0.5f * kPi - phi;
                float sn, cs; sincosf(phi, &sn, &cs);
                dx = r * cs; dy = r * sn;
            }
            const float dz = sqrtf(fmaxf(1.f - dx * dx - dy * dy, 0.f));
            V3 a, b; coordinate_system(n, a, b);
            dir = normalize(a * dx + b * dy + n * dz);
            ++depth;                                                        // non-null BSDF, vrl.cpp:336-337
            if (VRLB_SURF_GET_EXTERIOR(tr.flags) >= 0 || VRLB_SURF_GET_INTERIOR(tr.flags) >= 0) medium = target_medium(tr, dir);   // vrl.cpp:351-353
            pos = pos + dir * ((1.f + hmax_abs(pos)) * kRayEpsilon);
            open_vrl(pos, dir);                                             // vrl.cpp:354
            continue;
        }
        // tentative collision: null or real scattering (delta tracking), vrl.cpp:246-294.  The free flight was sampled with the
        // majorant; in a grid medium the collision is real with probability sigma_t(x) / majorant, otherwise the particle
        // carries on from here -- same direction, VRL in flight not cut, depth untouched (vrl.cpp:259-264; the throughput
        // factors of the two branches, sigma_n / sigma_n and sigma_s / sigma_t, leave 1 and the albedo).  Constant
        // extinction: sigma_t = majorant, every collision is real.
        const float u_null = rng.next_f32();                               // the null-scatter decision's draw, vrl.cpp:246
        if (m->kind == VRLB_MEDIUM_GRID) {
            const float sigma_t = m->scale * grid_eval(*m, pos);
            if (u_null >= sigma_t / m->max_density) continue;
        }
        ++depth;                                                           // vrl.cpp:254
        if (depth + 1 >= p.max_depth) break;                               // vrl.cpp:257-258: the scattering is dropped
        throughput = throughput * V3(m->albedo[0], m->albedo[1], m->albedo[2]);
        if (!WRITE && p.photon_out && medium_depth == 0 && !p.photon_flag[pi]) {   // handleMediumInteraction, vrl.cpp:276-281, 785 (one slot per
                                                                           // particle: a second direct scattering after a transmission is not stored)
            const V3 w = throughput * e.flux;                              // flux * throughput after the albedo
            p.photon_out[pi] = DPhoton{ pos.x, pos.y, pos.z, 0.f, dir.x, dir.y, dir.z, 0.f, w.x, w.y, w.z, 0.f };
            p.photon_flag[pi] = 1u;
        }
        ++medium_depth;
        push(pos);                                                         // vrl.cpp:283-288
        float u1 = rng.next_f32(), u2 = rng.next_f32();
        dir = sample_phase(*m, dir, u1, u2);
        open_vrl(pos, dir);
    }
    if (!WRITE) p.counts[pi] = count;

}

cudaError_t launch_generate(const GenParams &p, bool write, cudaStream_t stream) {
    if (p.n_particles == 0) return cudaSuccess;
    uint32_t grid = (uint32_t) ((p.n_particles + 127) / 128);
    if (write) generate_kernel<true><<<grid, 128, 0, stream>>>(p);
    else generate_kernel<false><<<grid, 128, 0, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace vrlb
