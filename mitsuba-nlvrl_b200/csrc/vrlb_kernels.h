// vrlb_kernels.h -- host-visible launch interface of the CUDA kernels (internal to libvrl_b200.so).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/vrl_b200.h"
#include "vrlb_device.cuh"

namespace vrlb {

struct GatherParams {
    DScene              scene;
    const vrlb_segment *segs;        // device, 64 B records
    uint64_t            n_segs;
    const DVrl         *vrls;        // device, 48 B records (this rank's shard)
    const DVrl         *vrls_fast;   // same records pre-multiplied for the fast kernel (launch_prep_fast_vrls)
    const DVrlV        *vrls_vis;    // 64-byte records of the fast visibility kernel (launch_prep_vis_vrls)
    uint64_t            n_vrls;
    uint64_t            first_index; // global index of vrls[0] (PCG32 stream addressing)
    uint64_t            total;       // size of the whole exported set
    float               scale;       // VRLMap::m_scale
    float               scale_rr;
    float               guard_h, guard_z; // fast gather: evals with closest distance < guard_h or asinh argument < -guard_z take the exact path
    uint32_t            q;           // samples_per_query
    int32_t             nla;         // use_nl_atomic_query: records with sample >> 16 = n_parts form one NLRay (exact kernel only)
    float              *film_rgb;    // device, width*height*3 fp32, accumulated with RED.ADD (may be null)
    float              *seg_out;     // device, n_segs*3: Li per segment before the weight (may be null)
    uint32_t           *wd;          // pinned host memory, 32 words: record left by the in-kernel watchdog before it traps
    float               guard_h2;    // guard_h^2, filled by launch_gather_fast
    float               safe[6];     // intersection of the shell boxes (lo, hi), filled by launch_gather_fast: compare operands from the constant bank
};

cudaError_t launch_gather(const GatherParams &p, bool uniform, bool rr, bool vis, bool grid, cudaStream_t stream);
// math_mode 1 / 2 (gather_fast.cu): iso = every medium isotropic, guard = route ill-conditioned evals to the exact path
cudaError_t launch_gather_fast(const GatherParams &p, bool uniform, bool rr, bool vis, bool iso, bool guard, cudaStream_t stream);
cudaError_t launch_prep_fast_vrls(const DVrl *in, uint64_t n, const DMedium *media, float scale, DVrl *out, cudaStream_t stream);
cudaError_t launch_prep_vis_vrls(const DVrl *in, uint64_t n, const DScene &scene, float scale, DVrlV *out, cudaStream_t stream);
void shell_safe_box(const DScene &scene, float safe[6]);   // intersection of the (shrunk) shell boxes, +-inf without shells

struct SegmentParams {
    DScene        scene;
    vrlb_camera   cam;
    vrlb_options  opt;
    const uint32_t *block_ids;   // device: spiral id per block, row-major over blocks
    int32_t       nbx, nby;
    uint64_t      n_vrls_total;
    vrlb_segment *segs;          // device output, capacity `cap`
    uint64_t      cap;
    unsigned long long *counter; // device: number of segments appended (may exceed cap -> overflow)
    float        *film_rgb;      // device: emitter radiance seen directly is added here
    float        *film_aw;       // device: width*height*2: sum(alpha), sum(weight)
    int32_t       add_direct;    // 0 on ranks > 0 of a VRL-sharded render: emitter radiance is counted once
    // camera paths that end on a smooth surface (vrl.cpp:671-680): one query each, when the scene carries reflectances
    vrlb_surface_query *squeries; uint64_t sq_cap; unsigned long long *sq_counter;   // null = not collected
    const float  *tri_refl;      // n_tris * 3, original triangle order
};
cudaError_t launch_segments(const SegmentParams &p, cudaStream_t stream);

// XYZAW film from the RGB accumulation buffer (integrator.cpp:282-294)
cudaError_t launch_film_xyzaw(const float *film_rgb, const float *film_aw, float *xyzaw, uint64_t n_pixels, cudaStream_t stream);
// SoA10 (+ medium ids) -> 48 B records, and back
cudaError_t launch_pack_vrls(const float *soa10, const uint16_t *medium, uint64_t n, DVrl *out, cudaStream_t stream);
cudaError_t launch_unpack_vrls(const DVrl *in, uint64_t n, float *soa10, uint16_t *medium, cudaStream_t stream);

// VRL generation by light tracing (generate.cu)
struct DPhoton;
struct GenParams {
    DScene    scene;
    uint64_t  n_particles, first_particle, base_seed;
    int32_t   max_depth, rr_depth, max_steps, use_non_linear, use_laser, start_medium;
    uint32_t  max_per_particle;
    float     min_vrl_length;
    float     laser_o[3], laser_d[3], laser_power[3];
    const int32_t *emit_tris;     // device: indices (BVH order) of emitter triangles
    const float   *emit_cdf;      // device: inclusive area cdf over emit_tris
    int32_t   n_emit_tris;
    float     emit_area;
    uint32_t *counts;             // device, per particle (count pass)
    const uint64_t *offsets;      // device, n_particles + 1 (write pass)
    DVrl     *out;                // device, VRL records
    DPhoton  *photon_out;     // device or null (count pass only): one slot per particle of the wave, the volume photon of
    uint32_t *photon_flag;        //   its FIRST medium scattering (vrl.cpp:278-281) and whether it deposited one
    const float *tri_refl;        // device or null: n_tris * 3 diffuse reflectance, ORIGINAL triangle order (DTri::prim)
    // surface photon collection (handleSurfaceInteraction, vrl.cpp:315, 772-782): counts / offsets / the two passes then
    // refer to surface photons instead of VRLs, and no VRL is written
    int32_t   collect_surface;
    DPhoton  *surf_out;
};
cudaError_t launch_generate(const GenParams &p, bool write, cudaStream_t stream);

// VRL-set transforms between light tracing and the gather (vrl_post.cu): long beams, dicing (count pass when out == nullptr)
cudaError_t launch_vrls_to_long(const DScene &s, DVrl *v, uint64_t n, unsigned long long *missed, cudaStream_t stream);
cudaError_t launch_vrls_dice(const DVrl *in, uint64_t n, const DMedium *media, float dice_len, uint32_t *counts,
                             const uint64_t *offsets, DVrl *out, cudaStream_t stream);

// Volume photon map and the direct in-scattering term (photons.cu; vrl.cpp:523-611, photonmap.h:216-271)
struct DPhoton {             // 48 B: position, propagation direction, power (PhotonData, photonmapper/photon.h)
    float px, py, pz, pad0;
    float dx, dy, dz, pad1;
    float wr, wg, wb, pad2;
};
struct PhotonGatherParams {
    DScene              scene;
    const vrlb_segment *segs; uint64_t n_segs;
    const DPhoton      *photons;      // sorted by grid cell
    const uint32_t     *cell_start;   // n_cells + 1
    uint64_t            n_photons;
    float               grid_lo[3], inv_cell; int32_t grid_res[3];
    float               lookup_radius, scale;
    int32_t             nl_camera;    // use_non_linear_camera (curved paths arrive as groups of consecutive parts)
    float              *film_rgb, *seg_out;
};
cudaError_t launch_photon_gather(const PhotonGatherParams &p, cudaStream_t stream);

// Surface photon maps (surface_photons.cu; vrl.cpp:671-680, photonmap.h:131-211)
struct PhotonGridDev {                // one resident photon map, binned like the volume map
    const DPhoton  *photons; const uint32_t *cell_start; uint64_t n;
    float grid_lo[3], inv_cell; int32_t grid_res[3];
    float lookup_radius, scale; int32_t lookup_size;
};
constexpr int VRLB_MAX_LOOKUP = 1024;
struct SurfaceGatherParams {
    const vrlb_surface_query *queries; uint64_t n_queries;
    PhotonGridDev global, caustic;
    float *film_rgb, *query_out;      // film: weight * estimate by atomicAdd; query_out: estimate per query (n * 3)
    float *heap;                      // n_queries * heap_stride floats of scratch (k-nearest squared distances)
    int32_t heap_stride;
};
cudaError_t launch_surface_gather(const SurfaceGatherParams &p, cudaStream_t stream);

// message returned by vrlb_last_error(NULL), for entry points that take no context (noalloc.cu)
void set_global_error(const char *msg);

// FFMA-saturating microkernel (peak.cu)
cudaError_t launch_fp32_peak(float *scratch, int blocks, int iters, cudaStream_t stream);

}  // namespace vrlb
