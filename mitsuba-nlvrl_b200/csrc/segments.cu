// segments.cu -- camera-segment generation: the part of VRLIntegrator::sample
// (src/integrators/vrl/vrl.cpp:443-732) that precedes and surrounds the VRL query, driven per pixel
// like SamplingIntegrator::render_block / render_sample (src/librender/integrator.cpp:187-297).
//
// One thread per film pixel in (block, Morton) order, so that consecutive segments -- and therefore the
// 32 segments a gather warp holds -- are spatial neighbours.  The pixel's PCG32 stream is consumed in
// reference order; the q*N*dpe draws of each query are skipped with PCG32::advance (random.h:265-286).
#include "vrlb_device.cuh"
#include "vrlb_kernels.h"

namespace vrlb {

__device__ inline uint32_t compact_bits(uint32_t v) {   // enoki::morton_decode, morton.h:148-155
    v &= 0x55555555u;
    v = (v | (v >> 1)) & 0x33333333u;
    v = (v | (v >> 2)) & 0x0f0f0f0fu;
    v = (v | (v >> 4)) & 0x00ff00ffu;
    v = (v | (v >> 8)) & 0x0000ffffu;
    return v;
}

// PerspectiveCamera::sample_ray_differential (src/sensors/perspective.cpp:183-216); Transform * Point /
// Vector accumulate column by column (include/mitsuba/core/transform.h:105-127)
__device__ inline void camera_ray(const vrlb_camera &c, float sx, float sy, V3 &o, V3 &d, float &mint, float &maxt) {
    const float *m = c.sample_to_camera;
    float r[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) r[k] = m[k * 4 + 3];
#pragma unroll
    for (int k = 0; k < 4; ++k) r[k] = m[k * 4 + 0] * sx + r[k];
#pragma unroll
    for (int k = 0; k < 4; ++k) r[k] = m[k * 4 + 1] * sy + r[k];
#pragma unroll
    for (int k = 0; k < 4; ++k) r[k] = m[k * 4 + 2] * 0.f + r[k];
    V3 near_p(r[0] / r[3], r[1] / r[3], r[2] / r[3]);
    V3 dl = normalize(near_p);
    float inv_z = 1.0f / dl.z;
    mint = c.near_clip * inv_z;
    maxt = c.far_clip * inv_z;
    const float *w = c.to_world;
    o = V3(w[3], w[7], w[11]);
    float q[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) q[k] = w[k * 4 + 0] * dl.x;
#pragma unroll
    for (int k = 0; k < 3; ++k) q[k] = w[k * 4 + 1] * dl.y + q[k];
#pragma unroll
    for (int k = 0; k < 3; ++k) q[k] = w[k * 4 + 2] * dl.z + q[k];
    d = V3(q[0], q[1], q[2]);
}

// cosine of the turning angle after which a curved camera path starts a new straight part (0.25 degrees)
constexpr float NL_PART_COS = 0.99999048f;
__global__ void __launch_bounds__(128) segments_kernel(SegmentParams p) {
    const int bs = p.opt.block_size;
    const uint32_t pixel_count = (uint32_t) (bs * bs);
    uint64_t tglobal = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t blk = tglobal / pixel_count;
    if (blk >= (uint64_t) p.nbx * p.nby) return;
    uint32_t mi = (uint32_t) (tglobal % pixel_count);
    int bx = (int) (blk % p.nbx), by = (int) (blk / p.nbx);
    int bx0 = bx * bs, by0 = by * bs;
    int bw = min(bs, p.cam.width - bx0), bh = min(bs, p.cam.height - by0);
    uint32_t lx = compact_bits(mi), ly = compact_bits(mi >> 1);
    if ((int) lx >= bw || (int) ly >= bh) return;
    const int px = bx0 + (int) lx, py = by0 + (int) ly;
    const DScene &s = p.scene;

    Pcg32 rng;
    rng.seed(p.opt.base_seed + (uint64_t) p.block_ids[blk] * pixel_count + mi);   // integrator.cpp:204, sampler.cpp:87-94
    const uint64_t dpe = p.opt.rr_vrl ? 3u : 2u;
    const uint64_t query_draws = (uint64_t) max(p.opt.samples_per_query, 1) * p.n_vrls_total * dpe;

    for (int j = 0; j < p.opt.spp; ++j) {
        float ax = 0.5f, ay = 0.5f;
        if (p.opt.spp > 1) { ax = rng.next_f32(); ay = rng.next_f32(); }       // integrator.cpp:249
        float posx = (float) px + ax, posy = (float) py + ay;
        (void) rng.next_f32();                                                  // wavelength sample, :261
        float sx = posx / (float) p.cam.width, sy = posy / (float) p.cam.height;
        V3 ro, rd; float rmint, rmaxt;
        camera_ray(p.cam, sx, sy, ro, rd, rmint, rmaxt);
        // ImageBlock::put, box-filter branch (imageblock.cpp:162-169)
        int fx = (int) ceilf(posx - (float) bx0 - 1.0f), fy = (int) ceilf(posy - (float) by0 - 1.0f);
        const bool on_film = fx >= 0 && fy >= 0 && fx < bw && fy < bh;
        const uint32_t film_pixel = (uint32_t) ((by0 + fy) * p.cam.width + (bx0 + fx));

        // ---- VRLIntegrator::sample, vrl.cpp:443-732 ----
        V3 radiance(0.f), throughput(1.f);
        int medium = p.cam.medium;
        Hit si; si.t = VRLB_INF; si.tri = -1;
        bool valid_ray = false, specular_chain = !p.opt.hide_emitters, needs_intersection = true;
        uint32_t depth = 1;
        (void) rng.next_f32();                                                  // channel draw, vrl.cpp:476-479
        for (int bounce = 0; bounce <= 100000; ++bounce) {
            bool active = !(throughput.x == 0.f && throughput.y == 0.f && throughput.z == 0.f);
            if (!active || depth >= (uint32_t) p.opt.max_depth) break;
            bool active_medium = medium >= 0, active_surface = !active_medium;
            if (active_medium) {
                if (si.tri < 0) break;             // reference idles to its 100k-bounce guard here (vrl.cpp:500-510, 523)
                valid_ray = true;
                (void) rng.next_f32();             // photon-gather radius jitter, vrl.cpp:527
                const DMedium &m = s.media[medium];
                auto emit = [&](V3 po, V3 pd, float pmaxt, V3 weight, uint32_t part) {
                    if (on_film) {
                        unsigned long long slot = atomicAdd(p.counter, 1ull);
                        if (slot < p.cap) {
                            vrlb_segment g;
                            g.o[0] = po.x; g.o[1] = po.y; g.o[2] = po.z; g.maxt = pmaxt;
                            g.d[0] = pd.x; g.d[1] = pd.y; g.d[2] = pd.z; g.pixel = film_pixel;
                            g.weight[0] = weight.x; g.weight[1] = weight.y; g.weight[2] = weight.z;
                            g.medium = medium; g.rng_state = rng.state; g.sample = (uint32_t) j; g.part = part;
                            p.segs[slot] = g;
                        }
                    }
                    rng.advance(query_draws);
                };
                if (m.kind == VRLB_MEDIUM_NONLINEAR && m.nl_step > 0.f && p.opt.use_non_linear_camera) {
                    // curved camera path (vrl.cpp:550-629): marched with the eikonal integrator of generate.cu, cut into
                    // straight parts when the direction has turned by more than NL_PART_COS; one query per part, the
                    // preceding parts' transmittance carried in the weight (vrl.cpp:619-628)
                    V3 pos = ro, dir = rd; float total = 0.f; bool found = false; Hit hit; hit.tri = -1; hit.t = VRLB_INF; hit.u = hit.v = 0.f;
                    // the march; on_part(origin, direction, length, index) is called for every straight part in order
                    auto march = [&](auto &&on_part) -> uint32_t {
                        V3 pstart = ro, dir0 = rd; pos = ro; dir = rd; total = 0.f; found = false;
                        float step_mint = rmint; uint32_t part = 0;
                        for (int step = 0; step < 100000; ++step) {
                            hit = closest_hit(s, pos, dir, step_mint, m.nl_step);
                            step_mint = 0.f;
                            if (hit.tri >= 0) { pos = pos + dir * hit.t; found = true; break; }
                            pos = pos + dir * m.nl_step;
                            V3 nd = nl_bend(m, pos, dir, m.nl_step);
                            if (dot(nd, dir0) < NL_PART_COS) {
                                V3 chord = pos - pstart; float len = norm(chord);
                                on_part(pstart, chord / len, len, part++);
                                total = total + len;
                                pstart = pos; dir0 = nd;
                            }
                            dir = nd;
                        }
                        if (!found) return part;
                        V3 chord = pos - pstart; float len = norm(chord);
                        if (len > 0.f) { on_part(pstart, chord / len, len, part++); total = total + len; }
                        return part;
                    };
                    if (p.opt.use_nl_atomic_query) {
                        // one query over the whole curved path (vrl.cpp:613-617): count the parts, reserve consecutive
                        // records, write them with the query's stream position and sample = j | n_parts << 16
                        const uint32_t n_parts = march([&](V3, V3, float, uint32_t) {});
                        if (!found) { si.tri = -1; break; }
                        if (n_parts > 0) {
                            unsigned long long slot0 = on_film ? atomicAdd(p.counter, (unsigned long long) n_parts) : 0ull;
                            if (on_film)
                                march([&](V3 po, V3 pd, float pmaxt, uint32_t part) {
                                    unsigned long long slot = slot0 + part;
                                    if (slot < p.cap) {
                                        vrlb_segment g;
                                        g.o[0] = po.x; g.o[1] = po.y; g.o[2] = po.z; g.maxt = pmaxt;
                                        g.d[0] = pd.x; g.d[1] = pd.y; g.d[2] = pd.z; g.pixel = film_pixel;
                                        g.weight[0] = throughput.x; g.weight[1] = throughput.y; g.weight[2] = throughput.z;
                                        g.medium = medium; g.rng_state = rng.state; g.sample = (uint32_t) j | (n_parts << 16); g.part = part;
                                        p.segs[slot] = g;
                                    }
                                });
                            rng.advance(query_draws);
                        }
                    } else {
                        // one query per part (vrl.cpp:619-628).  The parts of one path are written as consecutive records
                        // (count, reserve, write) with part = index | n_parts << 16, so that the direct in-scattering term
                        // (photons.cu) can walk the path in order with the step phase it carries across parts
                        const uint32_t n_parts = march([&](V3, V3, float, uint32_t) {});
                        const uint32_t n_enc = min(n_parts, 0xffffu) << 16;
                        unsigned long long slot0 = (on_film && n_parts > 0) ? atomicAdd(p.counter, (unsigned long long) n_parts) : 0ull;
                        V3 vrlThroughput = throughput;
                        march([&](V3 po, V3 pd, float pmaxt, uint32_t part) {
                            unsigned long long slot = slot0 + part;
                            if (on_film && slot < p.cap) {
                                vrlb_segment g;
                                g.o[0] = po.x; g.o[1] = po.y; g.o[2] = po.z; g.maxt = pmaxt;
                                g.d[0] = pd.x; g.d[1] = pd.y; g.d[2] = pd.z; g.pixel = film_pixel;
                                g.weight[0] = vrlThroughput.x; g.weight[1] = vrlThroughput.y; g.weight[2] = vrlThroughput.z;
                                g.medium = medium; g.rng_state = rng.state; g.sample = (uint32_t) j; g.part = part | n_enc;
                                p.segs[slot] = g;
                            }
                            rng.advance(query_draws);
                            vrlThroughput = vrlThroughput * homo_transmittance(m.max_density, pmaxt);
                        });
                        if (!found) { si.tri = -1; break; }
                    }
                    throughput = throughput * homo_transmittance(m.max_density, total);        // vrl.cpp:631
                    si = hit; rd = dir; needs_intersection = false;
                } else {
                    emit(ro, rd, si.t, throughput, 0u);
                    throughput = throughput * homo_transmittance(m.max_density, si.t);      // vrl.cpp:631
                }
                active_surface = true;
            }
            if (active_surface && needs_intersection) si = closest_hit(s, ro, rd, rmint, rmaxt);
            if (active_surface && si.tri >= 0 && specular_chain) {
                const DTri &tr = s.tris[si.tri];
                if ((tr.flags & VRLB_SURF_EMITTER) && dot(-rd, tri_normal(tr)) > 0.f) {   // area.cpp:64-72
                    int e = VRLB_SURF_GET_EMITTER(tr.flags);
                    if (e < s.n_emitters) {
                        V3 Le(s.emitters[e * 4 + 0], s.emitters[e * 4 + 1], s.emitters[e * 4 + 2]);
                        radiance = radiance + throughput * Le;
                    }
                }
            }
            active_surface = active_surface && si.tri >= 0;
            if (active_surface) {
                const DTri &tr = s.tris[si.tri];
                if (tr.flags & VRLB_SURF_EMITTER) break;                              // vrl.cpp:664-665
                valid_ray = true;
                const uint32_t bsdf_kind = tr.flags & VRLB_SURF_BSDF_MASK;
                if (bsdf_kind == VRLB_SURF_BSDF_SMOOTH) {                              // smooth BSDF: vrl.cpp:671-680
                    // radiance += (caustic + global photon estimate at si) * throughput: handed to surface_photons.cu
                    if (p.squeries && on_film) {
                        const float *rf = p.tri_refl + (size_t) tr.prim * 3;
                        const float r0 = __ldg(rf), r1 = __ldg(rf + 1), r2 = __ldg(rf + 2);
                        if (r0 != 0.f || r1 != 0.f || r2 != 0.f) {
                            unsigned long long slot = atomicAdd(p.sq_counter, 1ull);
                            if (slot < p.sq_cap) {
                                const float *vp = s.verts + (size_t) tr.prim * 9;             // si.p, mesh.cpp:467-484
                                const float b1 = si.u, b2 = si.v, b0 = 1.f - b1 - b2;
                                const V3 pnt = V3(vp[0], vp[1], vp[2]) * b0 + V3(vp[3], vp[4], vp[5]) * b1 + V3(vp[6], vp[7], vp[8]) * b2;
                                const V3 n = tri_normal(tr);
                                vrlb_surface_query q;
                                q.p[0] = pnt.x; q.p[1] = pnt.y; q.p[2] = pnt.z; q.pixel = film_pixel;
                                q.n[0] = n.x; q.n[1] = n.y; q.n[2] = n.z; q.flags = tr.flags & VRLB_SURF_TWOSIDED;
                                q.wi[0] = -rd.x; q.wi[1] = -rd.y; q.wi[2] = -rd.z; q.pad0 = 0.f;
                                q.reflectance[0] = r0; q.reflectance[1] = r1; q.reflectance[2] = r2; q.pad1 = 0.f;
                                q.weight[0] = throughput.x; q.weight[1] = throughput.y; q.weight[2] = throughput.z; q.pad2 = 0.f;
                                p.squeries[slot] = q;
                            }
                        }
                    }
                    break;
                }
                const float sample1 = rng.next_f32(); (void) rng.next_f32(); (void) rng.next_f32();  // bsdf->sample(ctx, si, next_1d, next_2d), vrl.cpp:682
                if (bsdf_kind != VRLB_SURF_BSDF_NULL) {
                    // specular chain (vrl.cpp:682-705): delta reflection / refraction, throughput *= weight, depth += 1
                    if (!p.tri_refl) break;
                    const float *rf = p.tri_refl + (size_t) tr.prim * 3;
                    const DeltaSample ds = sample_delta(bsdf_kind, __ldg(rf), __ldg(rf + 1), __ldg(rf + 2), (tr.flags & VRLB_SURF_TWOSIDED) != 0u,
                                                        tri_normal(tr), rd, sample1);
                    if (!ds.valid) break;
                    throughput = throughput * ds.weight;
                    if (throughput.x == 0.f && throughput.y == 0.f && throughput.z == 0.f) break;
                    rd = ds.wo;
                    depth += 1u;
                    specular_chain = true;                                            // vrl.cpp:700-701
                }
                const float *vp = s.verts + (size_t) tr.prim * 9;                     // si.p, mesh.cpp:467-484
                float b1 = si.u, b2 = si.v, b0 = 1.f - b1 - b2;
                V3 pnt = V3(vp[0], vp[1], vp[2]) * b0 + V3(vp[3], vp[4], vp[5]) * b1 + V3(vp[6], vp[7], vp[8]) * b2;
                ro = pnt;                                                             // si.spawn_ray, interaction.h:58-61
                rmint = (1.f + hmax_abs(pnt)) * kRayEpsilon;
                rmaxt = VRLB_INF;
                Hit si_new = closest_hit(s, ro, rd, rmint, rmaxt);                    // vrl.cpp:707-711
                needs_intersection = false;
                if (VRLB_SURF_GET_EXTERIOR(tr.flags) >= 0 || VRLB_SURF_GET_INTERIOR(tr.flags) >= 0)
                    medium = target_medium(tr, rd);                                   // vrl.cpp:713-714
                si = si_new;
            }
            if (!(active_surface || active_medium)) break;                            // vrl.cpp:719
        }
        if (on_film) {
            if (p.add_direct && (radiance.x != 0.f || radiance.y != 0.f || radiance.z != 0.f)) {
                atomicAdd(p.film_rgb + (size_t) film_pixel * 3 + 0, radiance.x);
                atomicAdd(p.film_rgb + (size_t) film_pixel * 3 + 1, radiance.y);
                atomicAdd(p.film_rgb + (size_t) film_pixel * 3 + 2, radiance.z);
            }
            atomicAdd(p.film_aw + (size_t) film_pixel * 2 + 0, valid_ray ? 1.f : 0.f);
            atomicAdd(p.film_aw + (size_t) film_pixel * 2 + 1, 1.f);
        }
    }
}

cudaError_t launch_segments(const SegmentParams &p, cudaStream_t stream) {
    uint64_t threads = (uint64_t) p.nbx * p.nby * (uint64_t) (p.opt.block_size * p.opt.block_size);
    uint32_t grid = (uint32_t) ((threads + 127) / 128);
    if (grid == 0) return cudaSuccess;
    segments_kernel<<<grid, 128, 0, stream>>>(p);
    return cudaGetLastError();
}

// ---- film + VRL packing helpers ---------------------------------------------------------------------
__global__ void film_xyzaw_kernel(const float *rgb, const float *aw, float *out, uint64_t n) {
    uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float r = rgb[i * 3], g = rgb[i * 3 + 1], b = rgb[i * 3 + 2];
    // srgb_to_xyz, include/mitsuba/core/spectrum.h:221-227
    out[i * 5 + 0] = 0.412453f * r + 0.357580f * g + 0.180423f * b;
    out[i * 5 + 1] = 0.212671f * r + 0.715160f * g + 0.072169f * b;
    out[i * 5 + 2] = 0.019334f * r + 0.119193f * g + 0.950227f * b;
    out[i * 5 + 3] = aw[i * 2];
    out[i * 5 + 4] = aw[i * 2 + 1];
}
cudaError_t launch_film_xyzaw(const float *film_rgb, const float *film_aw, float *xyzaw, uint64_t n, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    film_xyzaw_kernel<<<(uint32_t) ((n + 255) / 256), 256, 0, stream>>>(film_rgb, film_aw, xyzaw, n);
    return cudaGetLastError();
}

__global__ void pack_vrls_kernel(const float *soa, const uint16_t *medium, uint64_t n, DVrl *out) {
    uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    DVrl v;
    v.ox = soa[0 * n + j]; v.oy = soa[1 * n + j]; v.oz = soa[2 * n + j]; v.length = soa[6 * n + j];
    v.dx = soa[3 * n + j]; v.dy = soa[4 * n + j]; v.dz = soa[5 * n + j]; v.medium = medium ? medium[j] : 0u;
    v.fr = soa[7 * n + j]; v.fg = soa[8 * n + j]; v.fb = soa[9 * n + j]; v.pad = 0.f;
    out[j] = v;
}
cudaError_t launch_pack_vrls(const float *soa10, const uint16_t *medium, uint64_t n, DVrl *out, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    pack_vrls_kernel<<<(uint32_t) ((n + 255) / 256), 256, 0, stream>>>(soa10, medium, n, out);
    return cudaGetLastError();
}
__global__ void unpack_vrls_kernel(const DVrl *in, uint64_t n, float *soa, uint16_t *medium) {
    uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    DVrl v = in[j];
    soa[0 * n + j] = v.ox; soa[1 * n + j] = v.oy; soa[2 * n + j] = v.oz;
    soa[3 * n + j] = v.dx; soa[4 * n + j] = v.dy; soa[5 * n + j] = v.dz; soa[6 * n + j] = v.length;
    soa[7 * n + j] = v.fr; soa[8 * n + j] = v.fg; soa[9 * n + j] = v.fb;
    if (medium) medium[j] = (uint16_t) v.medium;
}
cudaError_t launch_unpack_vrls(const DVrl *in, uint64_t n, float *soa10, uint16_t *medium, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    unpack_vrls_kernel<<<(uint32_t) ((n + 255) / 256), 256, 0, stream>>>(in, n, soa10, medium);
    return cudaGetLastError();
}

}  // namespace vrlb
