/*
 * vrl_b200.h -- C ABI of the B200-native virtual-ray-light (VRL) volumetric gather.
 *
 * This is the drop-in boundary for the hot path of the `vrl` integrator of
 * VincentiusIV/mitsuba-NLVRL (reference paths below are relative to the
 * reference checkout).  The reference has no FFI on this path: everything is
 * C++ virtual calls inside one process.  The entry
 * points declared here are what a Mitsuba-side plugin shim `plugins/vrl.so`
 * (class VRLIntegrator, src/integrators/vrl/vrl.cpp:20-859) binds instead of
 * running the scalar CPU code; INTEGRATION.md shows that shim.
 *
 * Plain C: opaque handle, POD structs, fp32 arrays, explicit sizes.  No C++
 * or torch types cross this boundary.  All functions return VRLB_OK (0) or a
 * negative error code; vrlb_last_error() gives the message.  There is no CPU
 * fallback: every compute entry point fails with VRLB_ERR_CUDA when no
 * sm_100 device is usable.
 */
#ifndef VRL_B200_H
#define VRL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VRLB_ABI_VERSION 3

/* ---- status codes ------------------------------------------------------ */
enum {
    VRLB_OK            = 0,
    VRLB_ERR_INVALID   = -1,  /* bad argument (mirrors Throw(...) in Properties / xml.cpp:1050-1068) */
    VRLB_ERR_CUDA      = -2,  /* CUDA runtime failure or no device                                 */
    VRLB_ERR_STATE     = -3,  /* call order violated (e.g. gather before a VRL set is resident)   */
    VRLB_ERR_CANCELLED = -4,  /* vrlb_cancel() observed; mirrors Integrator::cancel, integrator.h:148 */
    VRLB_ERR_NOMEM     = -5,
    VRLB_ERR_UNSUPPORTED = -6 /* a combination the reference itself rejects, or one this library does not build (message says which) */
};

/* ---- surfaces ---------------------------------------------------------- */
/* One uint32 per triangle.  Restates what the hot path asks of a surface:
 * BSDFFlags::Null (vrl_struct.h:529), BSDFFlags::Smooth (vrl.cpp:671),
 * Shape::is_emitter (vrl.cpp:664), is_medium_transition/target_medium
 * (vrl_struct.h:552-557, vrl.cpp:713-714). */
#define VRLB_SURF_BSDF_MASK      0x3u
#define VRLB_SURF_BSDF_SMOOTH    0x0u   /* diffuse & co: ends the camera path, occludes          */
/* The delta classes below continue the camera path and the light-tracing particle (specular chains, vrl.cpp:318-358,
 * 682-716); towards shadow segments they occlude like every other surface (vrl_struct.h:529). */
#define VRLB_SURF_BSDF_NULL      0x1u   /* index-matched boundary: passed through                */
#define VRLB_SURF_BSDF_MIRROR    0x2u   /* smooth conductor (src/bsdfs/conductor.cpp): delta reflection, weight =
                                         * tri_reflectance (specular_reflectance x Fresnel, taken angle-independent) */
#define VRLB_SURF_BSDF_DIELECTRIC 0x3u  /* smooth dielectric (src/bsdfs/dielectric.cpp): tri_reflectance holds
                                         * { eta = int_ior / ext_ior, specular_reflectance, specular_transmittance } */
#define VRLB_SURF_EMITTER        0x4u   /* area emitter attached                                 */
#define VRLB_SURF_TWOSIDED       0x8u   /* `twosided` adapter: the diffuse lobe reflects on both sides (light tracing) */
#define VRLB_SURF_INTERIOR(m)    ((((uint32_t)(m)) + 1u) << 8)   /* medium id m on the -n side  */
#define VRLB_SURF_EXTERIOR(m)    ((((uint32_t)(m)) + 1u) << 16)  /* medium id m on the +n side  */
#define VRLB_SURF_EMITTER_ID(e)  (((uint32_t)(e)) << 24)
#define VRLB_SURF_GET_INTERIOR(f) ((int)(((f) >> 8) & 0xffu) - 1)   /* -1 = none */
#define VRLB_SURF_GET_EXTERIOR(f) ((int)(((f) >> 16) & 0xffu) - 1)
#define VRLB_SURF_GET_EMITTER(f)  ((int)(((f) >> 24) & 0xffu))

/* ---- media ------------------------------------------------------------- */
enum { VRLB_MEDIUM_HOMOGENEOUS = 0,   /* src/media/homogeneous.cpp   */
       VRLB_MEDIUM_GRID        = 1,   /* src/media/heterogeneous.cpp + src/textures/grid3d.cpp */
       VRLB_MEDIUM_NONLINEAR   = 2 }; /* src/media/nonlinear.cpp (homogeneous coefficients + IOR cell grid) */
enum { VRLB_PHASE_ISOTROPIC = 0,      /* src/phase/isotropic.cpp:40-44 */
       VRLB_PHASE_HG        = 1 };    /* src/phase/hg.cpp:51-54,77-81  */

typedef struct vrlb_medium {
    int32_t kind;
    int32_t phase;
    float   g;                 /* HG asymmetry                                         */
    float   scale;             /* `scale` property                                     */
    float   sigma_t[3];        /* constant sigma_t (homogeneous / nonlinear)            */
    float   albedo[3];         /* constant albedo                                      */
    /* grid medium (kind == VRLB_MEDIUM_GRID): scalar sigma_t grid, x fastest
     * ((z*ny + y)*nx + x, grid3d.cpp:291), `.vol` bbox (volume_data.h:42-107)   */
    const float *grid;         /* host pointer, nx*ny*nz floats; copied on scene_set    */
    int32_t grid_res[3];
    float   grid_bbox_min[3];
    float   grid_bbox_max[3];
    float   grid_max;          /* `m_sigmat->max()` of the grid (majorant before scale) */
    /* nonlinear medium (kind == VRLB_MEDIUM_NONLINEAR), nonlinear.cpp:40-89,129-153   */
    int32_t nl_res[3];
    float   nl_bbox_min[3];
    float   nl_bbox_max[3];
    float   top_ior, bottom_ior;
    int32_t ior_method;        /* only 0 (linear in height) is on the path             */
    int32_t from_bottom;
} vrlb_medium;

typedef struct vrlb_emitter {
    float radiance[3];         /* area emitter radiance (src/emitters/area.cpp)         */
    float pad;
} vrlb_emitter;

typedef struct vrlb_scene_desc {
    const float       *tri_verts;   /* n_tris * 9 floats: p0 p1 p2                      */
    const uint32_t    *tri_flags;   /* n_tris surface words (VRLB_SURF_*)               */
    uint32_t           n_tris;
    const vrlb_medium *media;
    uint32_t           n_media;
    const vrlb_emitter*emitters;
    uint32_t           n_emitters;
    /* ABI 3.  Diffuse reflectance of the SMOOTH surfaces, n_tris * 3 floats, or NULL.  Only the light tracer reads it
     * (vrlb_vrls_generate / vrlb_photons_generate): a particle that meets a smooth surface is reflected with
     * SmoothDiffuse::sample (src/bsdfs/diffuse.cpp: cosine lobe on the side of the normal the particle arrived on,
     * weight = reflectance) and carries on, as vrl.cpp:307-358 does.  NULL = reflectance 0: particles end there. */
    const float       *tri_reflectance;
} vrlb_scene_desc;

/* ---- sensor ------------------------------------------------------------ */
/* What PerspectiveCamera::sample_ray_differential (src/sensors/perspective.cpp:183-216)
 * reads: m_sample_to_camera, the to_world transform, clip planes; the film size. Row-major 4x4. */
typedef struct vrlb_camera {
    float   sample_to_camera[16];
    float   to_world[16];
    float   near_clip, far_clip;
    int32_t width, height;
    int32_t medium;            /* sensor->medium(): medium id the camera sits in, -1 = none */
    int32_t pad;
} vrlb_camera;

/* ---- integrator options ------------------------------------------------ */
/* Subset of VRLIntegrator's Properties that drives the gather (vrl.cpp:33-73) plus the
 * SamplingIntegrator / sampler state that addresses the PCG32 streams
 * (integrator.cpp:187-297, sampler.cpp:83-96). */
typedef struct vrlb_options {
    int32_t  spp;                   /* sampler sample_count                               */
    int32_t  block_size;            /* image block size (pin to 32 for parity, quirk 4)   */
    uint64_t base_seed;             /* sampler `seed` property                            */
    int32_t  samples_per_query;     /* q                                                  */
    int32_t  use_uniform_sampling;  /* sampleMC instead of importance sampling            */
    int32_t  rr_vrl;                /* distance Russian roulette (vrl_map.h:171-190)      */
    float    scale_rr;
    int32_t  max_depth;
    int32_t  use_non_linear_camera; /* split camera rays at IOR cell faces (vrl.cpp:550)  */
    int32_t  use_nl_atomic_query;   /* one query over the whole NLRay (vrl.cpp:613-617)   */
    int32_t  visibility;            /* 1: trace shadow segments through the BVH; 0: scene has no
                                       occluder between medium points (Tr_vis = medium Tr) */
    int32_t  hide_emitters;
    int32_t  math_mode;             /* 0: exact (the reference's IEEE operation order, every eval);
                                       1: production: fast arithmetic (FMA, MUFU) with a conditioning guard that
                                          sends near-singular evals (VRL closer than 1 % of the scene's extent to the camera part, or
                                          asinh argument < -8) through the exact path; parity <= 1e-4 per pixel;
                                       2: fast arithmetic without the guard (diagnostic: perf ceiling) */
    int32_t  reserved[2];           /* 0.  (development: override guard thresholds, [0] = h in 1e-3 units, [1] = z) */
    /* the VRL light cut (vrl_lightcut.h; properties vrl.cpp:66-72) */
    int32_t  use_light_cut;         /* 1: every query walks the cut of the light tree built by vrlb_lightcut_build
                                       (VRLMap::query's ELightCutAcceleration branch, vrl_map.h:198-201) */
    float    threshold_error;       /* m_errorRatio: a cluster is refined while its bound >= threshold_error * hmax(Li) (0.02) */
    int32_t  threshold_better_dist; /* clusters of at most this many VRLs bound the distance by exact segment-segment
                                       distances instead of the box (8) */
    int32_t  lightcut_heap;         /* entries of a query's refinement heap; 0 = default max(4096, 4 N^(2/3)), four times the
                                       largest heap measured on the bench frame.  A query that needs more is counted in
                                       vrlb_lightcut_stats and finishes with the exact sum over the VRLs (slow: N evals on one lane) */
} vrlb_options;

/* ---- VRL set ----------------------------------------------------------- */
/* SoA of 10 fp32 columns, each n long, in this order:
 *   ox oy oz  dx dy dz  length  flux_r flux_g flux_b        (struct VRL, vrl_struct.h:664-677)
 * i.e. column c of VRL j is soa10[c*n + j].  `medium` (may be NULL = all 0) is the medium id.
 * `scale` is VRLMap::m_scale = 1/volumePathCount (vrl.cpp:404,423).
 * `first_index`/`total` place this set inside a larger exported set: PCG32 stream addressing
 * uses the GLOBAL VRL index, so a rank holding a shard passes its offset here. */
typedef struct vrlb_vrlset_desc {
    const float    *soa10;
    const uint16_t *medium;
    uint64_t        n;
    uint64_t        first_index;
    uint64_t        total;
    float           scale;
    float           pad;
} vrlb_vrlset_desc;

/* ---- camera segments --------------------------------------------------- */
/* One straight part of a camera path inside a medium: an NLRay part (vrl_struct.h:10-57)
 * together with what VRLIntegrator::sample carries around its query (vrl.cpp:613-631). 64 bytes. */
typedef struct vrlb_segment {
    float    o[3];  float    maxt;
    float    d[3];  uint32_t pixel;       /* y*width + x of the film pixel the sample lands in */
    float    weight[3];                   /* throughput outside the query (vrlThroughput)      */
    int32_t  medium;
    uint64_t rng_state;                   /* PCG32 state at the first draw of this part's query */
    uint32_t sample;                      /* sample index within the pixel; with use_nl_atomic_query the high
                                           * 16 bits hold n_parts of the NLRay and its parts are consecutive */
    uint32_t part;                        /* low 16 bits: part index within the NLRay.  Curved paths queried part by
                                           * part: high 16 bits = n_parts, the parts are consecutive records (the direct
                                           * in-scattering term walks them in order, vrl.cpp:553-608) */
} vrlb_segment;

/* ---- timings of the last render, CUDA events on the library's stream --- */
typedef struct vrlb_timing {
    float    ms_upload;      /* host->device copies issued by the call         */
    float    ms_segments;    /* camera-segment kernel(s)                       */
    float    ms_gather;      /* gather kernel(s) incl. visibility              */
    float    ms_film;        /* film conversion                                */
    float    ms_download;    /* device->host copy                              */
    uint32_t gather_launches;
    uint32_t total_launches;
    uint64_t n_evals;        /* P*S*q*N getContrib calls (vrl_struct.h:584)    */
    uint64_t n_segments;
    uint64_t launches_lifetime; /* kernels of this library launched since vrlb_create (never reset) */
    float    ms_reduce;      /* NCCL film all-reduce (0 when single GPU)       */
    float    pad;
} vrlb_timing;

typedef struct vrlb_context vrlb_context;

/* ---- lifetime ---------------------------------------------------------- */
int  vrlb_abi_version(void);
/* Creates a context on CUDA device `device` (one context per GPU; one process per GPU under
 * torchrun).  Replaces the VRLIntegrator constructor's allocations, vrl.cpp:75-79. */
int  vrlb_create(int device, vrlb_context **out);
void vrlb_destroy(vrlb_context *ctx);
const char *vrlb_last_error(const vrlb_context *ctx);   /* ctx may be NULL: last create error */
/* Run the library's work on an existing CUDA stream (cudaStream_t as void*); NULL = own stream. */
int  vrlb_set_stream(vrlb_context *ctx, void *cuda_stream);

/* ---- scene: replaces Scene::ray_intersect + ShapeKDTree (scene.cpp:114-131, kdtree.h:2069-2172),
 *      the media plugins and phase functions -------------------------------------------------- */
int  vrlb_scene_set(vrlb_context *ctx, const vrlb_scene_desc *scene);

/* ---- VRL set: replaces VRLMap::push_back/build/setScaleFactor (vrl_map.h:43-53,80-102,220) --- */
int  vrlb_vrls_upload(vrlb_context *ctx, const vrlb_vrlset_desc *set);
int  vrlb_vrls_download(vrlb_context *ctx, float *soa10, uint16_t *medium, uint64_t cap, uint64_t *n_out);
uint64_t vrlb_vrls_count(const vrlb_context *ctx);

/* ---- VRL generation on the device: replaces the light tracer of VRLIntegrator::preprocess (vrl.cpp:117-374) and
 *      NonLinearMedia::sampleNonLinearInteraction / handleNonLinearInteraction (nonlinear.cpp:244-405).  Particles are
 *      marched through graded-IOR media with a fixed-step eikonal integrator (step = one IOR layer, height / res_y) and
 *      emit piecewise-linear VRLs.  The resident set is replaced; m_scale = 1 / (particles traced) (vrl.cpp:404,423). */
typedef struct vrlb_generate_desc {
    uint64_t target_vrls;          /* `target_vrls`: generation stops at the particle that reaches it            */
    uint64_t seed;                 /* particle i uses PCG32 seed(seed + i)                                       */
    int32_t  max_depth, rr_depth;  /* vrl.cpp:36,46                                                              */
    float    min_vrl_length;       /* vrl.cpp:64                                                                 */
    int32_t  use_non_linear;       /* vrl.cpp:53: 0 = straight flights even in a nonlinear medium                */
    int32_t  use_laser;            /* vrl.cpp:50,156-162; otherwise the scene's area emitters (VRLB_SURF_EMITTER) */
    float    laser_origin[3], laser_direction[3], laser_power[3];
    int32_t  start_medium;         /* medium id at the emission point, -1 = outside all media                    */
    int32_t  max_steps;            /* eikonal step budget per flight (0 = 100000)                                */
    int32_t  reserved[2];
} vrlb_generate_desc;
/* m_scale of the resident set, and the scatter depth of every resident VRL (VRL::m_scatterDepth, vrl_struct.h:675;
 * 0 for uploaded sets) */
float vrlb_vrls_scale(const vrlb_context *ctx);
int  vrlb_vrls_download_depth(vrlb_context *ctx, int32_t *depth, uint64_t cap);
int  vrlb_vrls_generate(vrlb_context *ctx, const vrlb_generate_desc *gen, uint64_t *n_vrls_out, uint64_t *n_particles_out);
/* The two transforms VRLIntegrator::preprocess applies to the map before the gather (vrl.cpp:408-415), on the resident set:
 *   vrlb_vrls_to_long: `long_vrl`  -- VRLMap::toLong / VRL::convertToLong (vrl_map.h:56-62, vrl_struct.h:146-154): every VRL is
 *                      extended to the closest surface along its direction; VRLB_ERR_STATE if one sees none (the reference throws);
 *   vrlb_vrls_dice:    `dice_vrl` > 1 -- VRLMap::dicingVRL / VRL::dice (vrl_map.h:104-128, vrl_struct.h:88-128): every VRL is cut
 *                      into pieces of (mean length / avg_number_dice), flux attenuated by the medium transmittance up to the piece;
 *                      the set keeps VRL-major order and its m_scale.  Homogeneous / nonlinear media only. */
int  vrlb_vrls_to_long(vrlb_context *ctx);
int  vrlb_vrls_dice(vrlb_context *ctx, uint32_t avg_number_dice, uint64_t *n_vrls_out);
/* vrlb_vrls_dice for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  Same pieces as
 * vrlb_vrls_dice (VRL::dice arithmetic, VRL-major order) for one homogeneous medium of grey extinction `max_density`
 * (= scale * mean(sigma_t)), but every buffer is the caller's and lives on `device`: d_soa10 = the 10 columns of the
 * input set (column c of VRL j at d_soa10[c*n + j]); d_soa10_out = the diced set with column stride cap_out (column c
 * of piece i at d_soa10_out[c*cap_out + i]); d_n_out = two uint64 on the device: [0] the diced size, [1] a status
 * (0 ok; 1 the mean length is not positive; 2 the diced size exceeds cap_out -- in both cases nothing is written).
 * The work is enqueued on `cuda_stream` (cudaStream_t as void*, NULL = the default stream) and the call returns without
 * synchronising.  It calls no cudaMalloc / cudaFree: the piece counts and offsets live in a static __device__ array
 * sized for VRLB_NOALLOC_MAX_VRLS input VRLs (~0.75 MB held by every GPU that loads the library, until exit); a larger
 * n is VRLB_ERR_INVALID (message through vrlb_last_error(NULL)).  All calls on a device share that array: calls that
 * overlap on the GPU (two streams, two threads) race -- serialise them. */
#define VRLB_NOALLOC_MAX_VRLS 65536u
int  vrlb_vrls_dice_noalloc(int device, void *cuda_stream, const float *d_soa10, uint64_t n, float max_density,
                            uint32_t avg_number_dice, float *d_soa10_out, uint64_t cap_out, uint64_t *d_n_out);

/* ---- volume photon map: the direct in-scattering term of VRLIntegrator::sample (vrl.cpp:523-611) ----------------------
 * A resident photon set switches the term on (`use_direct_illum`): vrlb_gather / vrlb_render then add, for every in-medium
 * camera segment, sum over the gather points (every 2 r along the segment, r = lookup_radius * lerp(0.75, 1.25, u)) of
 * throughput * Tr * PhotonMap::estimateRadianceVolume (photonmap.h:216-271) * scale.  Photons are what the light tracer
 * deposits at a particle's first medium scattering (vrl.cpp:278-281, 785: position, propagation direction, power after the
 * albedo).  pos / dir / power are n x 3 floats each; `scale` = 1 / directVolumePathCount (vrl.cpp:431); `lookup_radius` =
 * volume_lookup_radius_relative * scene radius (vrl.cpp:451).  On a curved (nonlinear-medium) camera path the gather
 * points run over all its parts with the step phase carried from part to part (vrl.cpp:553-577), the transmittance taken
 * over the straight distance between consecutive points and the phase function against the first part's direction. */
typedef struct vrlb_photonset_desc {
    const float *pos, *dir, *power;
    uint64_t     n;
    float        scale;
    float        lookup_radius;
} vrlb_photonset_desc;
int  vrlb_photons_upload(vrlb_context *ctx, const vrlb_photonset_desc *set);   /* n = 0 clears the set (term off) */
uint64_t vrlb_photons_count(const vrlb_context *ctx);
/* Light tracing on the device for the photon set (generate.cu, the same particles as vrlb_vrls_generate): one photon at each
 * particle's first medium scattering, in particle order, until `capacity` (`volume_photons`) are stored; scale as the reference
 * counts its paths (vrl.cpp:124, 372-373, 431: every particle traced while the map was not full, except the one that fills
 * it).  capacity = 0 or a scene where nothing scatters leaves no set (term off). */
int  vrlb_photons_generate(vrlb_context *ctx, const vrlb_generate_desc *gen, uint64_t capacity, float lookup_radius,
                           uint64_t *n_photons_out, uint64_t *n_particles_out);
/* pos / dir / power (n x 3 each, any may be NULL) of the resident set in grid order, and its scale */
int  vrlb_photons_download(vrlb_context *ctx, float *pos, float *dir, float *power, uint64_t cap, float *scale_out);
/* Test/diagnostic entry: the direct term per resident camera segment (throughput included), host buffer n_segs*3. */
int  vrlb_photon_gather_segments(vrlb_context *ctx, const vrlb_options *opt, float *direct_out);

/* ---- surface photon maps: the radiance the camera path collects at its first smooth surface (vrl.cpp:671-680) --------
 * radiance += throughput * (causticMap.estimateCausticRadiance(si, r_c, k_c) + globalMap.estimateRadiance(si, r_g, k_g)),
 * photonmap.h:131-211: the k nearest photons within the lookup radius (kd-tree nnSearch, kdtree.h:319-400: when more than
 * k lie inside, the radius shrinks to the k-th nearest), each weighted with bsdf->eval(si, wo = -photon direction)
 * (SmoothDiffuse::eval = reflectance / pi * cos_theta_o where both cosines are positive), times scale / (pi r^2); the
 * caustic estimate adds the cone filter max(0, 1 - d / r) and a factor 3.  Photons are what the light tracer stores at
 * every smooth surface a particle meets (handleSurfaceInteraction, vrl.cpp:315, 772-782): position, travel direction,
 * flux * throughput before the reflection; caustic = the last surface event before it was a transmission.
 * scale = 1 / surfacePathCount for both maps (vrl.cpp:379-399). */
enum { VRLB_PHOTONMAP_GLOBAL = 0, VRLB_PHOTONMAP_CAUSTIC = 1 };
/* set->lookup_radius = {global,caustic}_lookup_radius_relative * scene radius (vrl.cpp:448-450); lookup_size =
 * {global,caustic}_lookup_size (1 .. 1024).  n = 0 clears the map. */
int  vrlb_surface_photons_upload(vrlb_context *ctx, int which, const vrlb_photonset_desc *set, int32_t lookup_size);
uint64_t vrlb_surface_photons_count(const vrlb_context *ctx, int which);
/* Light tracing for the global or the caustic map (the same particles as vrlb_vrls_generate): one photon per smooth-surface
 * hit -- caustic when the particle's last surface event was a transmission through a dielectric, global otherwise
 * (wasTransmitted, vrl.cpp:345, 776-780) -- (needs
 * vrlb_scene_desc::tri_reflectance; a zero reflectance stands for a BSDF without a smooth lobe and stores nothing), in
 * particle order, until `capacity` (`global_photons`) are stored; scale = 1 / (particles traced while the map was not
 * full, the one that fills it not counted) as vrl.cpp:124, 368-369, 379 count surfacePathCount. */
int  vrlb_surface_photons_generate(vrlb_context *ctx, int which, const vrlb_generate_desc *gen, uint64_t capacity, float lookup_radius,
                                   int32_t lookup_size, uint64_t *n_photons_out, uint64_t *n_particles_out);
int  vrlb_surface_photons_download(vrlb_context *ctx, int which, float *pos, float *dir, float *power, uint64_t cap, float *scale_out);
/* One surface query = the end of a camera path on a smooth surface; vrlb_gather adds weight * estimate to the film. */
typedef struct vrlb_surface_query {
    float    p[3];   uint32_t pixel;      /* si.p; y*width + x                                                   */
    float    n[3];   uint32_t flags;      /* geometric normal (= shading frame here); VRLB_SURF_TWOSIDED or 0     */
    float    wi[3];  float    pad0;       /* world direction towards the camera, -ray.d                           */
    float    reflectance[3]; float pad1;  /* of the surface's diffuse BSDF                                        */
    float    weight[3]; float pad2;       /* path throughput                                                      */
} vrlb_surface_query;
/* (vrlb_segments_build replaces the resident queries with those of the paths it traces; vrlb_segments_upload drops them.) */
int  vrlb_surface_queries_upload(vrlb_context *ctx, const vrlb_surface_query *q, uint64_t n, int32_t width, int32_t height);
int  vrlb_surface_queries_download(vrlb_context *ctx, vrlb_surface_query *q, uint64_t cap, uint64_t *n_out);
/* Test/diagnostic entry: caustic + global estimate per resident query BEFORE the weight, host buffer n*3. */
int  vrlb_surface_gather_queries(vrlb_context *ctx, float *radiance_out);

/* ---- camera segments: replaces the medium branch of VRLIntegrator::sample up to the query
 *      (vrl.cpp:443-631) driven by SamplingIntegrator::render_block/render_sample
 *      (integrator.cpp:187-297) ----------------------------------------------------------------- */
int  vrlb_segments_build(vrlb_context *ctx, const vrlb_camera *cam, const vrlb_options *opt);
int  vrlb_segments_upload(vrlb_context *ctx, const vrlb_segment *segs, uint64_t n, int32_t width, int32_t height, int32_t spp);
int  vrlb_segments_download(vrlb_context *ctx, vrlb_segment *segs, uint64_t cap, uint64_t *n_out);
uint64_t vrlb_segments_count(const vrlb_context *ctx);

/* ---- gather: replaces VRLMap::query -> VRL::getContrib (vrl_map.h:133-211, vrl_struct.h:584-662)
 *      for every resident camera segment against every resident VRL.  Accumulates
 *      weight * Li into the device radiance buffer (fp32 RGB, width*height*3).
 *      A long gather is issued as launches of about a quarter of a second over consecutive VRL ranges (multiples of 512,
 *      so the film receives the values one launch would add); the call then paces them -- it returns when all but the
 *      last are done -- and looks at the cancel flag between two launches.  With opt->use_light_cut the queries walk
 *      the cut of the resident light tree instead (vrlb_lightcut_build). ----------------- */
int  vrlb_film_clear(vrlb_context *ctx);
int  vrlb_gather(vrlb_context *ctx, const vrlb_options *opt);
/* Same, and additionally returns Li per resident segment BEFORE the weight (host, n_segments*3 fp32):
 * exactly the `color` VRLMap::query returns (vrl_map.h:210), for per-segment parity checks. */
int  vrlb_gather_segments(vrlb_context *ctx, const vrlb_options *opt, float *li_out);

/* ---- the VRL light cut: replaces VRLLightCut (src/integrators/vrl/vrl_lightcut.h), which VRLMap::build creates when
 *      use_light_cut is set (vrl_map.h:89-97) and VRLMap::query walks instead of the whole set (vrl_map.h:198-201). ---- */
/* Builds the reference's light tree (vrl_lightcut.h:196-212 -> buildLightTree :515-627: greedy bottom-up clustering by
 * Node::dist :101-109, one sampler draw per merge for the representative :137-150) over the resident VRL set -- the same
 * tree as the reference builds from the same VRLs and the same sampler (sampler.seed(base_seed, seed_offset), the sampler
 * VRLIntegrator::preprocess hands to VRLMap::build, vrl.cpp:424).  The partner searches run on the device.  A context
 * holding a shard builds the tree of its shard.  Any call that changes the VRL set drops the tree.
 * Not built: stochastic_lightcut (shuffleRepresentativeLight re-draws representatives per query on a per-thread copy of
 * the tree, vrl_lightcut.h:153-193) -> the host mirror reports it as unsupported. */
int  vrlb_lightcut_build(vrlb_context *ctx, uint64_t base_seed, uint64_t seed_offset);
/* The tree in preorder, for parity checks: f13[k] = cluster flux(3), addFlux(3), aabb lo(3) hi(3), maxVRLLenght;
 * i6[k] = is leaf, VRL of the representative, child sharing it (isSame; -1 for a leaf), VRLs in the cluster, id, parent id.
 * Returns the node count through n_out (2 n - 1); writes at most cap nodes. */
int  vrlb_lightcut_download(vrlb_context *ctx, float *f13, int32_t *i6, uint64_t cap, uint64_t *n_out);
typedef struct vrlb_lightcut_stats {
    uint64_t n_nodes;        /* 2 n - 1                                                              */
    uint64_t build_searches; /* partner searches of the build (n in one launch + one launch each)     */
    float    build_ms;       /* wall time of vrlb_lightcut_build                                      */
    float    pad;
    uint64_t queries;        /* of the last gather through the cut: queries (segments x samples_per_query ...) */
    uint64_t evaluations;    /* ... fresh getContrib calls (LCQuery::nb_evaluation summed)            */
    uint64_t refinements;    /* ... clusters replaced by their children                               */
    uint64_t overflows;      /* ... queries whose heap outgrew lightcut_heap and were summed exactly   */
    uint64_t max_heap;       /* ... largest refinement heap                                           */
} vrlb_lightcut_stats;
int  vrlb_lightcut_get_stats(const vrlb_context *ctx, vrlb_lightcut_stats *out);
/* vrlb_gather_segments through the cut (opt->use_light_cut is implied), additionally returning per segment
 * stats4 = fresh evaluations, refinements, largest heap, 1 if the query overflowed (host, n_segments*4 fp32; may be NULL). */
int  vrlb_lightcut_gather_segments(vrlb_context *ctx, const vrlb_options *opt, float *li_out, float *stats4_out);
/* Device pointer of the fp32 RGB radiance accumulation buffer (width*height*3 floats) so that a
 * multi-GPU host can sum VRL shards with NCCL without a host round trip. */
int  vrlb_film_device_ptr(vrlb_context *ctx, void **dev_ptr, uint64_t *n_floats);
/* Film contract of SamplingIntegrator::render_sample / ImageBlock::put (integrator.cpp:282-294,
 * imageblock.cpp:162-169): 5 fp32 channels X,Y,Z,A,W per pixel, weighted sums (box filter). */
int  vrlb_film_download_xyzaw(vrlb_context *ctx, float *xyzaw);
int  vrlb_film_download_rgb(vrlb_context *ctx, float *rgb);

/* ---- whole frame from host buffers: Integrator::render (integrator.cpp:51-183) --------------- */
int  vrlb_render(vrlb_context *ctx, const vrlb_camera *cam, const vrlb_options *opt,
                 const vrlb_vrlset_desc *vrls /* NULL: keep resident set */, float *xyzaw_out);

/* ---- multi-GPU: VRLs are sharded across ranks (one context per GPU / process), every rank gathers all
 *      pixels against its shard, and the fp32 RGB film buffers are summed with one NCCL all-reduce over
 *      NVLink.  Replaces the mutex-protected HDRFilm::put merge (hdrfilm.cpp:210-214).
 *      NCCL is bound at run time (dlopen of libnccl.so.2), so single-GPU users need no NCCL at all. ---- */
#define VRLB_COMM_ID_BYTES 128
int  vrlb_comm_unique_id(void *id /* VRLB_COMM_ID_BYTES, filled on the calling rank */);
int  vrlb_comm_init(vrlb_context *ctx, const void *id, int rank, int n_ranks);
int  vrlb_comm_destroy(vrlb_context *ctx);
/* Sum the radiance film over all ranks in place, on the context's stream.  Emitter radiance seen
 * directly (vrl.cpp:651-657) and the A/W channels are written by rank 0 only / replicated, so the
 * sum counts them once. */
int  vrlb_film_allreduce(vrlb_context *ctx);
/* XYZAW conversion on the device only (no download); device pointer to width*height*5 floats. */
int  vrlb_film_resolve(vrlb_context *ctx, void **dev_xyzaw);

/* ---- measurement helper: FP32 FMA issue peak of this device, measured by an FFMA-saturating
 *      microkernel (roofline denominator of the gather).  tflops = 2 flops per FMA. ---- */
int  vrlb_measure_fp32_peak(vrlb_context *ctx, float *tflops);

/* ---- diagnostic, host only (no device needed): the shadow-query structure the library builds for a mesh.
 *      Solids are the connected components that get conservative proxies (mitsuba-nlvrl_b200/csrc/
 *      occluder_proxies.cpp); each is written as VRLB_SOLID_WORDS 32-bit words:
 *      aabb lo[3] hi[3] | u32 root | u32 flags (1 outer box, 2 inscribed sphere, 4 inscribed box) |
 *      bounding sphere c[3] r^2 | inscribed sphere c[3] r^2 | box axes[9] (rows) | box centre[3] |
 *      outer half extents[3] | inscribed half extents[3].  counts[0] = solids, counts[1] = shells,
 *      counts[2] = triangles left in the main BVH. ---- */
#define VRLB_SOLID_WORDS 34
int  vrlb_shadow_proxies(const float *tri_verts, uint32_t n_tris, float *solids_out, uint32_t cap_solids, uint32_t counts[3]);

/* Depth (root = 0) of the closest-hit BVH and of the deepest shadow BVH the library builds for a mesh: the device
 * traversals use fixed per-lane stacks, so the host build bounds its depth (28; balanced median splits once a binned-SAH
 * branch would go deeper).  Host only.  Stands where the reference's kd-tree has its own depth limit
 * (include/mitsuba/render/kdtree.h: max_depth). */
int  vrlb_bvh_depths(const float *tri_verts, uint32_t n_tris, uint32_t depths[2]);
/* Which builder the NEXT vrlb_scene_set uses for the closest-hit tree (the one the camera paths and the light tracer
 * walk, Scene::ray_intersect -> ShapeKDTree, scene.cpp:114-131, kdtree.h:2069-2172): 0 = binned SAH on the host (default),
 * 1 = LBVH on the device (Morton sort + parallel hierarchy + bottom-up fit, bvh_lbvh.cu).  Hits do not depend on the tree.
 * A device tree deeper than the traversal stack allows falls back to the host build.  With builder 1 the trees of the
 * shadow structure (main geometry, each shell, each solid) are built on the device as well; its components, shell boxes
 * and the solids' proxies are found on the host either way. */
int  vrlb_set_bvh_builder(vrlb_context *ctx, int builder);
/* info = closest-hit tree of the last vrlb_scene_set: builder used (0 / 1), depth (root = 0), nodes, build time in
 * microseconds; shadow structure: trees built on the device, depth of its deepest tree */
int  vrlb_bvh_info(const vrlb_context *ctx, uint32_t info[6]);

/* Integrator::cancel (integrator.h:148-151).  Thread-safe against a running vrlb_gather / vrlb_render of the same context:
 * the gather stops between two of its launches (< 1 s) and returns VRLB_ERR_CANCELLED; the film then holds a partial sum. */
int  vrlb_cancel(vrlb_context *ctx);
int  vrlb_get_timing(const vrlb_context *ctx, vrlb_timing *out);
int  vrlb_synchronize(vrlb_context *ctx);

#ifdef __cplusplus
}
#endif
#endif /* VRL_B200_H */
