/*
 * vrl_oracle.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (scalar fp32, one IEEE operation per written operation, built with
 * -ffp-contract=off) of the VRL gather hot path of VincentiusIV/mitsuba-NLVRL.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
 *
 * PARITY STATUS: pinned for the homogeneous-medium gather.  The reference's own vrl_struct.h is compiled in
 * place against a mock Mitsuba (oracle/ref_probe.cpp + ref_mock.h -> oracle/_ref/refprobe) and its outputs
 * for findClosetPoint / samplingVRL / evalTransmittance / getContrib / sequential-stream query sums are the
 * fixture tests/golden/ref_vrl_struct.json, which tests/test_ref_golden_cpu.py holds this file to (closest
 * point bit-exact, getContrib within 3.5e-7).  The PCG32 stream is pinned by tests/golden/enoki_pcg32.json
 * (Enoki's random.h compiled into oracle/_ref/enoki_probe).  The photon-map estimates (estimateRadiance,
 * estimateCausticRadiance, estimateRadianceVolume) are pinned the same way: photonmapper/photonmap.h + kdtree.h
 * compiled in place (oracle/ref_probe_pm.cpp -> tests/golden/ref_photonmap.json), agreement 2e-7.  "Parity unpinned"
 * still applies to the grid-medium stochastic transmittance, the camera-segment construction and VRL generation: no
 * reference code could be run for those (the delta BSDF samples inside them are pinned bit-exact through the reference's
 * fresnel.h, oracle/ref_probe_fresnel.cpp).
 *
 * Input structs are the public ABI PODs of include/vrl_b200.h (interface only).
 */
#ifndef VRL_ORACLE_H
#define VRL_ORACLE_H
#include "../include/vrl_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* rng addressing modes (SURVEY 8d "PCG32 stream addressing") */
enum { VRLO_RNG_ADDRESSED = 0,   /* fixed draws per (part, sample, VRL); what the GPU implements */
       VRLO_RNG_SEQUENTIAL = 1 };/* bit-faithful to the reference: early-outs consume fewer draws */

/* -- PCG32 / sampler (ext/enoki/include/enoki/random.h:62-119,265-286; sampler.cpp:83-96) -- */
void     vrlo_pcg32_seed(uint64_t initstate, uint64_t initseq, uint64_t *state, uint64_t *inc);
uint32_t vrlo_pcg32_next_u32(uint64_t *state, uint64_t inc);
float    vrlo_pcg32_next_f32(uint64_t *state, uint64_t inc);
void     vrlo_pcg32_advance(uint64_t *state, uint64_t inc, int64_t delta);
/* n floats of the stream of IndependentSampler::seed(seed_offset) with base seed `base_seed` */
void     vrlo_sampler_stream(uint64_t base_seed, uint64_t seed_offset, uint32_t n, float *out_f, uint32_t *out_u);

/* -- block / pixel order (spiral.cpp:27-72, integrator.cpp:203-206) -- */
/* fills block_id[by*nbx+bx] with the spiral id of every block; returns number of blocks */
uint32_t vrlo_spiral_block_ids(int32_t width, int32_t height, int32_t block_size, uint32_t *block_id, uint32_t cap);
void     vrlo_morton_decode(uint32_t i, uint32_t *x, uint32_t *y);

/* -- single-eval probes (vrl_struct.h:163-228, 239-360, 584-662) -- */
void  vrlo_find_closest_point(const float vrl_o[3], const float vrl_d[3], float vrl_len,
                              const float ray_o[3], const float ray_d[3], float ray_maxt,
                              float *tCam, float *tVRL);
/* returns 1 if valid; out9 = pCam(3) tCam pVRL(3) tVRL invPDF */
int   vrlo_sampling_vrl(const float vrl_o[3], const float vrl_d[3], float vrl_len,
                        const float ray_o[3], const float ray_d[3], float ray_maxt,
                        int uniform, float u1, float u2, float *out9);
float vrlo_phase_eval(const vrlb_medium *m, const float wi[3], const float wo[3]);
/* one getContrib with explicit uniforms, homogeneous media; visibility through `scene` if non-NULL */
void  vrlo_get_contrib(const vrlb_scene_desc *scene, const vrlb_medium *m,
                       const float vrl10[10], const float ray_o[3], const float ray_d[3], float ray_maxt,
                       int uniform, float u1, float u2, float out_rgb[3]);

/* -- scene queries (scene.cpp:114-131, mesh.h:195-225) -- */
typedef struct vrlo_scene vrlo_scene;
vrlo_scene *vrlo_scene_create(const vrlb_scene_desc *desc);
void        vrlo_scene_destroy(vrlo_scene *s);
/* closest hit; returns prim index or -1; t/u/v written */
int   vrlo_ray_intersect(const vrlo_scene *s, const float o[3], const float d[3], float mint, float maxt,
                         float *t, float *u, float *v);
/* VRL::evalTransmittance (vrl_struct.h:509-572) for p1->p2 starting in medium `medium` */
void  vrlo_eval_transmittance(const vrlo_scene *s, const float p1[3], const float p2[3], int medium, float out_rgb[3]);

/* -- direct in-scattering term from the volume photon map (vrl.cpp:523-611, photonmap.h:216-271), per segment, throughput included -- */
void vrlo_photon_gather_segments(const vrlo_scene *s, const vrlb_photonset_desc *set, const vrlb_segment *segs, uint64_t n_segs,
                                 const vrlb_options *opt, float *out_rgb);
/* caustic + global surface estimate per query (vrl.cpp:671-680, photonmap.h:131-211); either set may be NULL */
void vrlo_surface_gather_queries(const vrlb_photonset_desc *global, int32_t k_global, const vrlb_photonset_desc *caustic, int32_t k_caustic,
                                 const vrlb_surface_query *queries, uint64_t n, float *out_rgb);
/* SmoothConductor / SmoothDielectric::sample as the specular chains call it (vrl.cpp:320, 682); 0 = inactive lobe */
int vrlo_sample_delta(uint32_t flags, const float prm[3], const float n[3], const float rd[3], float sample1, float wo[3], float weight[3], int *transmitted);
/* PhotonMap::estimateRadianceVolume (photonmap.h:216-271) at one gather point */
void vrlo_estimate_radiance_volume(const vrlb_medium *m, const vrlb_photonset_desc *set, const float gp[3], const float wo[3], float radius, float out[3]);
/* surface queries of the camera paths (vrl.cpp:671-680), film order of vrlo_build_segments; returns the count */
uint64_t vrlo_build_surface_queries(const vrlo_scene *s, const vrlb_camera *cam, const vrlb_options *opt, uint64_t n_vrls_total,
                                    vrlb_surface_query *out, uint64_t cap);

/* -- VRL-set transforms of VRLIntegrator::preprocess (vrl.cpp:408-415) -- */
/* VRLMap::toLong (vrl_map.h:56-62): new lengths; returns the number of VRLs that see no surface (the reference throws) */
uint64_t vrlo_vrls_to_long(const vrlo_scene *s, const float *soa10, uint64_t n, float *length_out);
/* VRLMap::dicingVRL (vrl_map.h:104-128): out_soa10 is [10][cap]; returns the diced count (may exceed cap: nothing lost but the tail) */
void vrlo_set_build_crop(int x0, int y0, int x1, int y1);
float vrlo_nl_ior(const vrlb_medium *m, const float p[3]);
int vrlo_nl_march(const vrlo_scene *s, int medium, const float o[3], const float d[3], float mint, float *out8);
uint64_t vrlo_vrls_dice(const vrlo_scene *s, const float *soa10, const uint16_t *medium, uint64_t n, uint32_t avg_number_dice,
                        float *out_soa10, uint16_t *out_medium, uint64_t cap);

/* -- camera segments: VRLIntegrator::sample up to the query (vrl.cpp:443-631) -- */
/* returns number of segments written (<= cap); n_vrls_total fixes the draws a query consumes in
 * addressed mode.  direct_rgb (width*height*3, may be NULL) receives emitter radiance seen on the
 * specular chain (vrl.cpp:651-657); alpha_w (width*height*2, may be NULL) receives sum(alpha), sum(w). */
uint64_t vrlo_build_segments(const vrlo_scene *s, const vrlb_camera *cam, const vrlb_options *opt,
                             uint64_t n_vrls_total, int rng_mode,
                             vrlb_segment *segs, uint64_t cap, float *direct_rgb, float *alpha_w);

/* -- gather: VRLMap::query over explicit segments (vrl_map.h:133-211) -- */
/* out_rgb32: per segment Li with the reference's sequential fp32 sum; out_rgb64: same fp32
 * per-eval contributions summed in fp64 (either may be NULL). Returns the number of evals.
 * Multi-threaded over segments when n_threads > 1. */
uint64_t vrlo_gather_segments(const vrlo_scene *s, const vrlb_vrlset_desc *vrls,
                              const vrlb_segment *segs, uint64_t n_segs, const vrlb_options *opt,
                              int rng_mode, int n_threads, float *out_rgb32, double *out_rgb64);

/* -- whole frame: SamplingIntegrator::render (integrator.cpp:51-183) in 32x32 blocks over
 *    n_threads std::threads; film = XYZAW fp32 (imageblock.cpp:162-169).  crop = {x0,y0,x1,y1}
 *    restricts the rendered pixels (NULL = full frame) for bounded CPU-baseline samples.
 *    Returns evals performed. -- */
uint64_t vrlo_render(const vrlo_scene *s, const vrlb_camera *cam, const vrlb_options *opt,
                     const vrlb_vrlset_desc *vrls, int rng_mode, int n_threads,
                     const int32_t *crop, float *xyzaw, float *rgb);

/* -- the VRL light cut (vrl_lightcut.h), pinned by tests/golden/ref_lightcut.json -- */
typedef struct vrlo_lightcut vrlo_lightcut;
vrlo_lightcut *vrlo_lightcut_build(const vrlb_vrlset_desc *set, uint64_t base_seed, uint64_t seed_offset, int n_threads);
void     vrlo_lightcut_destroy(vrlo_lightcut *lc);
uint64_t vrlo_lightcut_n_nodes(const vrlo_lightcut *lc);
uint64_t vrlo_lightcut_build_draws(const vrlo_lightcut *lc);
void     vrlo_lightcut_nodes(const vrlo_lightcut *lc, float *f13, int32_t *i6);
float    vrlo_lc_node_dist(const float a10[10], const float b10[10]);
void     vrlo_lc_merge(const float a10[10], const float b10[10], float v, float *out14);
float    vrlo_lc_box_dist(const float lo[3], const float hi[3], const float o[3], const float d[3], float maxt);
float    vrlo_lc_seg_dist(const float vrl10[10], const float o[3], const float d[3], float maxt);
float    vrlo_lightcut_bound(const vrlo_scene *s, const vrlo_lightcut *lc, int32_t node, const float *parts7, int n_parts, int better_dist);
void     vrlo_lightcut_query(const vrlo_scene *s, const vrlo_lightcut *lc, float scale, const float *parts7, int n_parts, const vrlb_options *opt,
                             int rng_mode, uint64_t base_seed, uint64_t seed_offset, float out_rgb[3], double *stats5);
uint64_t vrlo_lightcut_gather_segments(const vrlo_scene *s, const vrlo_lightcut *lc, float scale, const vrlb_segment *segs, uint64_t n_segs,
                                       const vrlb_options *opt, int rng_mode, int n_threads, float *out_rgb32, float *stats4);

#ifdef __cplusplus
}
#endif
#endif
