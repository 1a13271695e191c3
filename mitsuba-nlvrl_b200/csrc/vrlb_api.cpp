// vrlb_api.cpp -- implementation of the C ABI declared in include/vrl_b200.h.
//
// Host side of libvrl_b200.so: owns device memory, the CUDA stream and events, builds the BVH,
// mirrors the control flow of Integrator::render (src/librender/integrator.cpp:51-183) and
// VRLIntegrator::preprocess/render (src/integrators/vrl/vrl.cpp) around the kernels of gather.cu and
// segments.cu.  No CPU compute fallback: without a usable CUDA device every entry point fails.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/vrl_b200.h"
#include "bvh_build.h"
#include "vrlb_kernels.h"
#include "lightcut.h"

using namespace vrlb;

namespace {

std::string g_create_error;
int (*g_nccl_destroy)(void *) = nullptr;   // set once NCCL is bound (vrlb_comm_*)

template <typename T> struct DevBuf {
    T *ptr = nullptr; size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (ptr) cudaFree(ptr);
        ptr = nullptr; cap = 0;
        cudaError_t e = cudaMalloc((void **) &ptr, std::max<size_t>(n, 1) * sizeof(T));
        if (e == cudaSuccess) cap = n;
        return e;
    }
    void release() { if (ptr) cudaFree(ptr); ptr = nullptr; cap = 0; }
};

}  // namespace

struct vrlb_context {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    std::string error;
    std::atomic<int> cancel{ 0 };

    // scene
    DevBuf<DTri> tris; DevBuf<float> verts; DevBuf<DNode> nodes; DevBuf<DWNode> wnodes; DevBuf<DTri> stris; DevBuf<DMedium> media; DevBuf<float> emitters; DevBuf<float> tri_refl; bool have_refl = false; DevBuf<float> grids; bool any_grid = false; DevBuf<DSolid> solids;
    DScene scene{};
    bool have_scene = false;
    float scene_extent = 20.f;        // largest side of the geometry's bounding box (the conditioning guard scales with it)
    int bvh_builder = 0;              // vrlb_set_bvh_builder: 0 = host binned SAH, 1 = device LBVH (closest-hit tree)
    uint32_t bvh_info[6] = { 0, 0, 0, 0, 0, 0 };   // closest-hit tree: builder used, depth, nodes, build time in microseconds; shadow trees built on the device, their largest depth
    DevBuf<uint32_t> flags_dev;
    std::vector<vrlb_medium> host_media;
    std::vector<DTri> host_tris;      // BVH order, for the emitter table of vrlb_vrls_generate

    // VRL set
    DevBuf<DVrl> vrls, vrls_fast; DevBuf<DVrlV> vrls_vis; bool fast_dirty = true; bool all_iso = true;
    DevBuf<float> stage_soa; DevBuf<uint16_t> stage_med;
    uint64_t n_vrls = 0, vrl_first = 0, vrl_total = 0; float vrl_scale = 1.f;
    // light cut over the resident set (lightcut.cu); dropped whenever the set changes
    LightcutTree *lc = nullptr;
    DevBuf<unsigned long long> lc_heap; DevBuf<float4> lc_slots; DevBuf<unsigned long long> lc_stats; DevBuf<float> lc_seg_stats;
    vrlb_lightcut_stats lc_last{};
    void vrls_changed() { fast_dirty = true; if (lc) { lightcut_free(lc); lc = nullptr; } }

    // volume photon map (photons.cu): sorted by grid cell
    struct PhotonMapRes {
        DevBuf<DPhoton> photons; DevBuf<uint32_t> cells; uint64_t n = 0;
        float scale = 0.f, radius = 0.f, lo[3] = { 0, 0, 0 }, inv_cell = 0.f; int32_t res[3] = { 1, 1, 1 }; int32_t lookup_size = 0;
    };
    PhotonMapRes vol;                 // volume photon map (photons.cu)
    PhotonMapRes surf[2];             // global / caustic surface maps (surface_photons.cu), VRLB_PHOTONMAP_*
    DevBuf<vrlb_surface_query> squeries; uint64_t n_squeries = 0;   // camera paths that end on a reflecting surface

    // segments
    DevBuf<vrlb_segment> segs; uint64_t n_segs = 0;
    DevBuf<unsigned long long> counter;
    DevBuf<uint32_t> block_ids; std::vector<uint32_t> host_block_ids; int bid_w = 0, bid_h = 0, bid_bs = 0;

    // film
    DevBuf<float> film_rgb, film_aw, film_xyzaw;
    int width = 0, height = 0;

    // pinned staging for host<->device copies of the e2e path
    void *pinned = nullptr; size_t pinned_cap = 0;
    uint32_t *wd = nullptr;           // pinned, 32 words: in-kernel watchdog record (gather_fast.cu)

    cudaEvent_t ev[10] = { nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr };   // [8], [9]: the gather's launch pacing
    vrlb_timing timing{};
    uint64_t launches = 0;       // lifetime kernel-launch count of this library

    // multi-GPU film reduce
    void *nccl_comm = nullptr; int rank = 0, n_ranks = 1;
};

namespace {

int fail(vrlb_context *ctx, int code, const char *fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    if (ctx) ctx->error = buf; else g_create_error = buf;
    return code;
}
// CUDA failure: message + the in-kernel watchdog's record when it is the one that trapped (gather_fast.cu)
int fail_cuda(vrlb_context *ctx, const char *what, cudaError_t e) {
    const uint32_t *w = ctx->wd;
    if (w && w[0]) {
        char rec[512]; int n = 0;
        for (int i = 0; i < 25 && n < 480; ++i) n += std::snprintf(rec + n, sizeof(rec) - (size_t) n, "%s%x", i ? " " : "", w[i]);
        return fail(ctx, VRLB_ERR_CUDA, "%s: %s [gather watchdog code %u: %s]", what, cudaGetErrorString(e), w[0], rec);
    }
    return fail(ctx, VRLB_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
}
#define CU(ctx, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail_cuda(ctx, #call, e_); } while (0)

int ensure_pinned(vrlb_context *ctx, size_t bytes) {
    if (bytes <= ctx->pinned_cap) return VRLB_OK;
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    ctx->pinned = nullptr; ctx->pinned_cap = 0;
    CU(ctx, cudaMallocHost(&ctx->pinned, bytes));
    ctx->pinned_cap = bytes;
    return VRLB_OK;
}

// Spiral block ids, src/librender/spiral.cpp:8-72 (ids in spiral order from the film centre)
std::vector<uint32_t> spiral_ids(int w, int h, int bs, int &nbx, int &nby) {
    nbx = (w + bs - 1) / bs; nby = (h + bs - 1) / bs;
    int count = nbx * nby;
    std::vector<uint32_t> ids((size_t) count, 0u);
    int px = nbx / 2, py = nby / 2, dir = 0, steps_left = 1, steps = 1;
    for (int c = 0; c < count; ++c) {
        ids[(size_t) py * nbx + px] = (uint32_t) c;
        if (c + 1 != count) {
            do {
                switch (dir) { case 0: ++px; break; case 1: ++py; break; case 2: --px; break; default: --py; break; }
                if (--steps_left == 0) {
                    dir = (dir + 1) % 4;
                    if (dir == 0 || dir == 2) ++steps;
                    steps_left = steps;
                }
            } while (px < 0 || py < 0 || px >= nbx || py >= nby);
        }
    }
    return ids;
}

int ensure_film(vrlb_context *ctx, int w, int h) {
    if (w <= 0 || h <= 0) return fail(ctx, VRLB_ERR_INVALID, "film size %dx%d", w, h);
    size_t np = (size_t) w * h;
    bool fresh = (w != ctx->width || h != ctx->height);
    CU(ctx, ctx->film_rgb.reserve(np * 3));
    CU(ctx, ctx->film_aw.reserve(np * 2));
    CU(ctx, ctx->film_xyzaw.reserve(np * 5));
    ctx->width = w; ctx->height = h;
    if (fresh) {
        CU(ctx, cudaMemsetAsync(ctx->film_rgb.ptr, 0, np * 3 * sizeof(float), ctx->stream));
        CU(ctx, cudaMemsetAsync(ctx->film_aw.ptr, 0, np * 2 * sizeof(float), ctx->stream));
    }
    return VRLB_OK;
}

float elapsed(cudaEvent_t a, cudaEvent_t b) { float ms = 0.f; cudaEventElapsedTime(&ms, a, b); return ms; }

}  // namespace

namespace vrlb { void set_global_error(const char *msg) { g_create_error = msg ? msg : ""; } }

extern "C" {

int vrlb_abi_version(void) { return VRLB_ABI_VERSION; }

int vrlb_create(int device, vrlb_context **out) {
    if (!out) return fail(nullptr, VRLB_ERR_INVALID, "out is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(nullptr, VRLB_ERR_CUDA, "no CUDA device (%s); libvrl_b200 has no CPU fallback", cudaGetErrorString(e));
    if (device < 0 || device >= n) return fail(nullptr, VRLB_ERR_INVALID, "device %d out of range [0,%d)", device, n);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return fail(nullptr, VRLB_ERR_CUDA, "%s", cudaGetErrorString(e));
    if (prop.major < 10)
        return fail(nullptr, VRLB_ERR_CUDA, "device %d is sm_%d%d; this library carries sm_100a code only", device, prop.major, prop.minor);
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail(nullptr, VRLB_ERR_CUDA, "%s", cudaGetErrorString(e));
    vrlb_context *ctx = new vrlb_context();
    ctx->device = device;
    if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) {
        delete ctx; return fail(nullptr, VRLB_ERR_CUDA, "%s", cudaGetErrorString(e));
    }
    for (auto &ev : ctx->ev) cudaEventCreate(&ev);
    cudaMalloc((void **) &ctx->counter.ptr, 2 * sizeof(unsigned long long)); ctx->counter.cap = 2;   // [0] segments, [1] surface queries
    *out = ctx;
    return VRLB_OK;
}

void vrlb_destroy(vrlb_context *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    ctx->tris.release(); ctx->verts.release(); ctx->nodes.release(); ctx->wnodes.release(); ctx->stris.release(); ctx->media.release(); ctx->emitters.release(); ctx->grids.release(); ctx->solids.release();
    ctx->tri_refl.release(); ctx->squeries.release(); ctx->flags_dev.release();
    ctx->vol.photons.release(); ctx->vol.cells.release();
    for (auto &m : ctx->surf) { m.photons.release(); m.cells.release(); }
    if (ctx->wd) cudaFreeHost(ctx->wd);
    if (ctx->lc) lightcut_free(ctx->lc);
    ctx->lc_heap.release(); ctx->lc_slots.release(); ctx->lc_stats.release(); ctx->lc_seg_stats.release();
    ctx->vrls.release(); ctx->vrls_fast.release(); ctx->vrls_vis.release(); ctx->stage_soa.release(); ctx->stage_med.release(); ctx->segs.release(); ctx->counter.release();
    if (ctx->nccl_comm && g_nccl_destroy) g_nccl_destroy(ctx->nccl_comm);
    ctx->block_ids.release(); ctx->film_rgb.release(); ctx->film_aw.release(); ctx->film_xyzaw.release();
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    for (auto &ev : ctx->ev) if (ev) cudaEventDestroy(ev);
    if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char *vrlb_last_error(const vrlb_context *ctx) { return ctx ? ctx->error.c_str() : g_create_error.c_str(); }

int vrlb_set_stream(vrlb_context *ctx, void *cuda_stream) {
    if (!ctx) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    if (cuda_stream) { ctx->stream = (cudaStream_t) cuda_stream; ctx->own_stream = false; }
    else { CU(ctx, cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)); ctx->own_stream = true; }
    return VRLB_OK;
}

int vrlb_synchronize(vrlb_context *ctx) {
    if (!ctx) return VRLB_ERR_INVALID;
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return VRLB_OK;
}

int vrlb_shadow_proxies(const float *tri_verts, uint32_t n_tris, float *solids_out, uint32_t cap_solids, uint32_t counts[3]) {
    static_assert(sizeof(DSolid) == VRLB_SOLID_WORDS * 4, "DSolid layout is part of the diagnostic ABI");
    if ((n_tris && !tri_verts) || !counts) return VRLB_ERR_INVALID;
    ShadowAccel sa; build_shadow_accel(tri_verts, nullptr, n_tris, sa);
    counts[0] = (uint32_t) sa.solids.size(); counts[1] = (uint32_t) sa.shells.size();
    counts[2] = n_tris;
    // triangles not in the main BVH: recount from the leaf references is overkill; shells + solids own the rest
    if (solids_out)
        for (size_t i = 0; i < sa.solids.size() && i < cap_solids; ++i) std::memcpy(solids_out + i * VRLB_SOLID_WORDS, &sa.solids[i], sizeof(DSolid));
    counts[2] = sa.n_main_tris;
    return VRLB_OK;
}

int vrlb_bvh_depths(const float *tri_verts, uint32_t n_tris, uint32_t depths[2]) {
    if ((n_tris && !tri_verts) || !depths) return VRLB_ERR_INVALID;
    std::vector<DNode> nodes; std::vector<DTri> tris;
    build_bvh(tri_verts, nullptr, n_tris, nodes, tris);
    depths[0] = (uint32_t) last_bvh_depth();
    ShadowAccel sa; build_shadow_accel(tri_verts, nullptr, n_tris, sa);
    depths[1] = (uint32_t) sa.max_depth;
    return VRLB_OK;
}

int vrlb_set_bvh_builder(vrlb_context *ctx, int builder) {
    if (!ctx) return VRLB_ERR_INVALID;
    if (builder != 0 && builder != 1) return fail(ctx, VRLB_ERR_INVALID, "bvh builder %d unknown (0 host binned SAH, 1 device LBVH)", builder);
    ctx->bvh_builder = builder;
    return VRLB_OK;
}
int vrlb_bvh_info(const vrlb_context *ctx, uint32_t info[6]) {
    if (!ctx || !info) return VRLB_ERR_INVALID;
    for (int k = 0; k < 6; ++k) info[k] = ctx->bvh_info[k];
    return VRLB_OK;
}

int vrlb_cancel(vrlb_context *ctx) { if (!ctx) return VRLB_ERR_INVALID; ctx->cancel.store(1); return VRLB_OK; }

int vrlb_get_timing(const vrlb_context *ctx, vrlb_timing *out) {
    if (!ctx || !out) return VRLB_ERR_INVALID;
    *out = ctx->timing;
    out->launches_lifetime = ctx->launches;
    // kernel durations from the CUDA events recorded around them on the library's stream, once complete
    if (ctx->timing.gather_launches && cudaEventQuery(ctx->ev[3]) == cudaSuccess) out->ms_gather = elapsed(ctx->ev[2], ctx->ev[3]);
    return VRLB_OK;
}

// ---- scene -----------------------------------------------------------------------------------------
int vrlb_scene_set(vrlb_context *ctx, const vrlb_scene_desc *sc) {
    if (!ctx || !sc) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (sc->n_tris && !sc->tri_verts) return fail(ctx, VRLB_ERR_INVALID, "tri_verts is NULL");
    if (sc->n_media == 0 || !sc->media) return fail(ctx, VRLB_ERR_INVALID, "a scene needs at least one medium");
    if (sc->n_tris >= (1u << 19)) return fail(ctx, VRLB_ERR_INVALID, "at most 2^19 - 1 triangles on this path (shadow leaf-queue packing)");
    if (sc->n_media > 255 || sc->n_emitters > 255) return fail(ctx, VRLB_ERR_INVALID, "at most 255 media / emitters");
    for (uint32_t i = 0; i < sc->n_tris; ++i) {
        uint32_t f = sc->tri_flags ? sc->tri_flags[i] : 0u;
        if (VRLB_SURF_GET_INTERIOR(f) >= (int) sc->n_media || VRLB_SURF_GET_EXTERIOR(f) >= (int) sc->n_media)
            return fail(ctx, VRLB_ERR_INVALID, "triangle %u references medium out of range", i);
        if ((f & VRLB_SURF_BSDF_MASK) > VRLB_SURF_BSDF_NULL && !sc->tri_reflectance)
            return fail(ctx, VRLB_ERR_INVALID, "triangle %u: BSDF class %u needs its parameters in tri_reflectance", i, f & VRLB_SURF_BSDF_MASK);
        if ((f & VRLB_SURF_BSDF_MASK) == VRLB_SURF_BSDF_DIELECTRIC && !(sc->tri_reflectance[(size_t) i * 3] > 0.f))
            return fail(ctx, VRLB_ERR_INVALID, "triangle %u: dielectric eta must be > 0", i);
    }
    std::vector<DMedium> dm(sc->n_media);
    std::vector<float> grid_host; std::vector<std::pair<uint32_t, size_t>> grid_off;
    for (uint32_t i = 0; i < sc->n_media; ++i) {
        const vrlb_medium &m = sc->media[i];
        if (m.kind != VRLB_MEDIUM_HOMOGENEOUS && m.kind != VRLB_MEDIUM_NONLINEAR && m.kind != VRLB_MEDIUM_GRID)
            return fail(ctx, VRLB_ERR_INVALID, "medium %u: unknown kind %d", i, m.kind);
        if (m.phase != VRLB_PHASE_ISOTROPIC && m.phase != VRLB_PHASE_HG) return fail(ctx, VRLB_ERR_INVALID, "medium %u: unknown phase %d", i, m.phase);
        if (m.phase == VRLB_PHASE_HG && !(m.g > -1.f && m.g < 1.f))
            return fail(ctx, VRLB_ERR_INVALID, "The asymmetry parameter must lie in the interval (-1, 1)!");   // hg.cpp:45-46
        DMedium d{};
        float mean = (m.sigma_t[0] + m.sigma_t[1] + m.sigma_t[2]) / 3.0f;      // ConstVolume::max() = mean, constant3d.cpp:37
        d.max_density = m.scale * mean;                                          // homogeneous.cpp:26
        for (int k = 0; k < 3; ++k) d.sigma_s[k] = (m.sigma_t[k] * m.scale) * m.albedo[k];   // homogeneous.cpp:42-50
        d.g = m.g; d.phase = m.phase; d.kind = m.kind; d.scale = m.scale;
        if (m.kind == VRLB_MEDIUM_GRID) {
            if (!m.grid || m.grid_res[0] < 1 || m.grid_res[1] < 1 || m.grid_res[2] < 1) return fail(ctx, VRLB_ERR_INVALID, "medium %u: grid medium needs grid and grid_res", i);
            for (int k = 0; k < 3; ++k)
                if (!(m.grid_bbox_max[k] > m.grid_bbox_min[k])) return fail(ctx, VRLB_ERR_INVALID, "medium %u: empty grid_bbox", i);
            if (!(m.grid_max > 0.f)) return fail(ctx, VRLB_ERR_INVALID, "medium %u: grid_max (the majorant before scale) must be positive", i);
            d.max_density = m.scale * m.grid_max;                                // heterogeneous.cpp:29
            for (int k = 0; k < 3; ++k) { d.sigma_s[k] = 0.f; d.g_res[k] = m.grid_res[k]; d.g_lo[k] = m.grid_bbox_min[k]; d.g_hi[k] = m.grid_bbox_max[k]; }
            grid_off.push_back({ i, grid_host.size() });
            grid_host.insert(grid_host.end(), m.grid, m.grid + (size_t) m.grid_res[0] * m.grid_res[1] * m.grid_res[2]);
        }
        for (int k = 0; k < 3; ++k) { d.albedo[k] = m.albedo[k]; d.nl_lo[k] = m.nl_bbox_min[k]; d.nl_hi[k] = m.nl_bbox_max[k]; }
        d.ior_bottom = m.bottom_ior; d.ior_top = m.top_ior; d.nl_step = 0.f; d.from_bottom = m.from_bottom ? 1 : 0;
        if (m.kind == VRLB_MEDIUM_NONLINEAR) {
            if (m.ior_method != 0) return fail(ctx, VRLB_ERR_INVALID, "medium %u: only ior_method 0 (linear in height) is on the path", i);
            if (!(m.nl_bbox_max[1] > m.nl_bbox_min[1]) || m.nl_res[1] < 1) return fail(ctx, VRLB_ERR_INVALID, "medium %u: nonlinear medium needs nl_bbox and nl_res", i);
            if (!(m.bottom_ior > 0.f && m.top_ior > 0.f)) return fail(ctx, VRLB_ERR_INVALID, "medium %u: IOR must be positive", i);
            d.nl_step = (m.nl_bbox_max[1] - m.nl_bbox_min[1]) / (float) m.nl_res[1];
        }
        dm[i] = d;
    }
    if (sc->n_tris) {
        float lo[3] = { INFINITY, INFINITY, INFINITY }, hi[3] = { -INFINITY, -INFINITY, -INFINITY };
        for (size_t i = 0; i < (size_t) sc->n_tris * 9; ++i) { lo[i % 3] = std::min(lo[i % 3], sc->tri_verts[i]); hi[i % 3] = std::max(hi[i % 3], sc->tri_verts[i]); }
        const float ext = std::max(hi[0] - lo[0], std::max(hi[1] - lo[1], hi[2] - lo[2]));
        ctx->scene_extent = (ext > 0.f && std::isfinite(ext)) ? ext : 20.f;
    } else ctx->scene_extent = 20.f;
    std::vector<DNode> nodes; std::vector<DTri> tris;
    bool device_tree = false;
    const auto t_bvh = std::chrono::steady_clock::now();
    if (ctx->bvh_builder == 1 && sc->n_tris > 0) {
        // closest-hit tree on the device (bvh_lbvh.cu): same layout, same triangle arithmetic; the host build's margin
        float lo[3] = { INFINITY, INFINITY, INFINITY }, hi[3] = { -INFINITY, -INFINITY, -INFINITY }, mx = 0.f;
        for (size_t i = 0; i < (size_t) sc->n_tris * 9; ++i) {
            lo[i % 3] = std::min(lo[i % 3], sc->tri_verts[i]); hi[i % 3] = std::max(hi[i % 3], sc->tri_verts[i]);
            mx = std::max(mx, std::fabs(sc->tri_verts[i]));
        }
        const float pad = 1e-4f * mx + 1e-6f;
        const uint32_t n = sc->n_tris;
        CU(ctx, ctx->verts.reserve((size_t) n * 9)); CU(ctx, ctx->flags_dev.reserve(n));
        CU(ctx, ctx->tris.reserve(n)); CU(ctx, ctx->nodes.reserve((size_t) 2 * n));
        CU(ctx, cudaStreamSynchronize(ctx->stream));
        CU(ctx, cudaMemcpy(ctx->verts.ptr, sc->tri_verts, (size_t) n * 9 * sizeof(float), cudaMemcpyHostToDevice));
        CU(ctx, cudaMemcpy(ctx->flags_dev.ptr, sc->tri_flags, (size_t) n * sizeof(uint32_t), cudaMemcpyHostToDevice));
        int depth = 0;
        CU(ctx, build_bvh_device(ctx->verts.ptr, ctx->flags_dev.ptr, n, pad, lo, hi, ctx->nodes.ptr, ctx->tris.ptr, &depth, &ctx->launches, ctx->stream));
        if (depth <= 44) {                       // closest_hit's stack holds depth + 1 entries of its 48
            device_tree = true;
            tris.resize(n); nodes.resize((size_t) 2 * n - 1);
            CU(ctx, cudaMemcpy(tris.data(), ctx->tris.ptr, (size_t) n * sizeof(DTri), cudaMemcpyDeviceToHost));      // the emitter table reads them
            CU(ctx, cudaMemcpy(nodes.data(), ctx->nodes.ptr, nodes.size() * sizeof(DNode), cudaMemcpyDeviceToHost));
            ctx->bvh_info[0] = 1; ctx->bvh_info[1] = (uint32_t) depth;
        }
    }
    if (!device_tree) {
        build_bvh(sc->tri_verts, sc->tri_flags, sc->n_tris, nodes, tris);
        ctx->bvh_info[0] = 0; ctx->bvh_info[1] = (uint32_t) last_bvh_depth();
    }
    ctx->bvh_info[2] = (uint32_t) nodes.size();
    ctx->bvh_info[3] = (uint32_t) std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_bvh).count();
    // the shadow structure's per-component trees through the device builder as well, when it is selected (components,
    // shells and proxies are found on the host either way)
    SubsetBuilder dev_builder = [&](const float *v, const uint32_t *f, uint32_t m, std::vector<DNode> &nd, std::vector<DTri> &tr, int &depth) -> bool {
        if (m == 0) return false;
        float lo[3] = { INFINITY, INFINITY, INFINITY }, hi[3] = { -INFINITY, -INFINITY, -INFINITY }, mx = 0.f;
        for (size_t i = 0; i < (size_t) m * 9; ++i) { lo[i % 3] = std::min(lo[i % 3], v[i]); hi[i % 3] = std::max(hi[i % 3], v[i]); mx = std::max(mx, std::fabs(v[i])); }
        DevBuf<float> dv; DevBuf<uint32_t> df; DevBuf<DNode> dn; DevBuf<DTri> dt;
        bool ok = dv.reserve((size_t) m * 9) == cudaSuccess && df.reserve(m) == cudaSuccess && dn.reserve((size_t) 2 * m) == cudaSuccess && dt.reserve(m) == cudaSuccess;
        ok = ok && cudaMemcpy(dv.ptr, v, (size_t) m * 9 * sizeof(float), cudaMemcpyHostToDevice) == cudaSuccess;
        ok = ok && cudaMemcpy(df.ptr, f, (size_t) m * sizeof(uint32_t), cudaMemcpyHostToDevice) == cudaSuccess;
        ok = ok && build_bvh_device(dv.ptr, df.ptr, m, 1e-4f * mx + 1e-6f, lo, hi, dn.ptr, dt.ptr, &depth, &ctx->launches, ctx->stream) == cudaSuccess;
        if (ok) {
            nd.resize((size_t) 2 * m - 1); tr.resize(m);
            ok = cudaMemcpy(nd.data(), dn.ptr, nd.size() * sizeof(DNode), cudaMemcpyDeviceToHost) == cudaSuccess &&
                 cudaMemcpy(tr.data(), dt.ptr, (size_t) m * sizeof(DTri), cudaMemcpyDeviceToHost) == cudaSuccess;
        }
        dv.release(); df.release(); dn.release(); dt.release();
        return ok;
    };
    uint32_t n_dev_components = 0;
    ShadowAccel sa; build_shadow_accel(sc->tri_verts, sc->tri_flags, sc->n_tris, sa, ctx->bvh_builder == 1 ? &dev_builder : nullptr, &n_dev_components);
    ctx->bvh_info[4] = n_dev_components; ctx->bvh_info[5] = (uint32_t) sa.max_depth;
    CU(ctx, ctx->wnodes.reserve(sa.wnodes.size())); CU(ctx, ctx->stris.reserve(sa.tris.size()));
    CU(ctx, ctx->tris.reserve(tris.size())); CU(ctx, ctx->nodes.reserve(nodes.size()));
    CU(ctx, ctx->verts.reserve((size_t) sc->n_tris * 9)); CU(ctx, ctx->media.reserve(dm.size()));
    CU(ctx, ctx->emitters.reserve((size_t) sc->n_emitters * 4));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    if (!tris.empty()) {
        if (!device_tree) {
            CU(ctx, cudaMemcpy(ctx->tris.ptr, tris.data(), tris.size() * sizeof(DTri), cudaMemcpyHostToDevice));
            CU(ctx, cudaMemcpy(ctx->nodes.ptr, nodes.data(), nodes.size() * sizeof(DNode), cudaMemcpyHostToDevice));
        }
        if (!sa.wnodes.empty()) CU(ctx, cudaMemcpy(ctx->wnodes.ptr, sa.wnodes.data(), sa.wnodes.size() * sizeof(DWNode), cudaMemcpyHostToDevice));
        CU(ctx, cudaMemcpy(ctx->stris.ptr, sa.tris.data(), sa.tris.size() * sizeof(DTri), cudaMemcpyHostToDevice));
        CU(ctx, cudaMemcpy(ctx->verts.ptr, sc->tri_verts, (size_t) sc->n_tris * 9 * sizeof(float), cudaMemcpyHostToDevice));
    }
    if (!grid_host.empty()) {
        CU(ctx, ctx->grids.reserve(grid_host.size()));
        CU(ctx, cudaMemcpy(ctx->grids.ptr, grid_host.data(), grid_host.size() * sizeof(float), cudaMemcpyHostToDevice));
        for (auto &go : grid_off) dm[go.first].grid = ctx->grids.ptr + go.second;
    }
    ctx->any_grid = !grid_off.empty();
    CU(ctx, cudaMemcpy(ctx->media.ptr, dm.data(), dm.size() * sizeof(DMedium), cudaMemcpyHostToDevice));
    if (sc->n_emitters) CU(ctx, cudaMemcpy(ctx->emitters.ptr, sc->emitters, (size_t) sc->n_emitters * sizeof(vrlb_emitter), cudaMemcpyHostToDevice));
    ctx->have_refl = sc->tri_reflectance != nullptr && sc->n_tris > 0;      // read by the light tracer only (generate.cu)
    if (ctx->have_refl) {
        for (size_t i = 0; i < (size_t) sc->n_tris * 3; ++i)
            if (!(sc->tri_reflectance[i] >= 0.f)) return fail(ctx, VRLB_ERR_INVALID, "tri_reflectance[%zu] is negative or NaN", i);
        CU(ctx, ctx->tri_refl.reserve((size_t) sc->n_tris * 3));
        CU(ctx, cudaMemcpy(ctx->tri_refl.ptr, sc->tri_reflectance, (size_t) sc->n_tris * 3 * sizeof(float), cudaMemcpyHostToDevice));
    }
    ctx->scene.tris = ctx->tris.ptr; ctx->scene.verts = ctx->verts.ptr; ctx->scene.nodes = ctx->nodes.ptr;
    ctx->scene.wnodes = ctx->wnodes.ptr; ctx->scene.stris = ctx->stris.ptr; ctx->scene.main_root = sa.main_root;
    for (int k = 0; k < 3; ++k) { ctx->scene.main_lo[k] = sa.main_lo[k]; ctx->scene.main_hi[k] = sa.main_hi[k]; }
    ctx->scene.n_shells = (int32_t) sa.shells.size();
    for (size_t i = 0; i < sa.shells.size(); ++i) ctx->scene.shells[i] = sa.shells[i];
    ctx->scene.n_solids = (int32_t) sa.solids.size();
    CU(ctx, ctx->solids.reserve(std::max<size_t>(sa.solids.size(), 1)));
    if (!sa.solids.empty()) CU(ctx, cudaMemcpy(ctx->solids.ptr, sa.solids.data(), sa.solids.size() * sizeof(DSolid), cudaMemcpyHostToDevice));
    ctx->scene.solids = ctx->solids.ptr;
    ctx->scene.n_tris = (int32_t) tris.size(); ctx->scene.n_nodes = (int32_t) nodes.size();
    ctx->scene.media = ctx->media.ptr; ctx->scene.n_media = (int32_t) dm.size();
    ctx->scene.emitters = ctx->emitters.ptr; ctx->scene.n_emitters = (int32_t) sc->n_emitters;
    ctx->host_media.assign(sc->media, sc->media + sc->n_media);
    ctx->host_tris = tris;
    ctx->have_scene = true; ctx->fast_dirty = true;
    ctx->all_iso = true;
    for (const DMedium &d : dm) if (d.phase != VRLB_PHASE_ISOTROPIC) ctx->all_iso = false;
    return VRLB_OK;
}

// ---- VRL set ---------------------------------------------------------------------------------------
int vrlb_vrls_upload(vrlb_context *ctx, const vrlb_vrlset_desc *set) {
    if (!ctx || !set) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (set->n && !set->soa10) return fail(ctx, VRLB_ERR_INVALID, "soa10 is NULL");
    uint64_t total = set->total ? set->total : set->n;
    if (set->first_index + set->n > total) return fail(ctx, VRLB_ERR_INVALID, "VRL shard [%llu,+%llu) exceeds total %llu",
                                                        (unsigned long long) set->first_index, (unsigned long long) set->n, (unsigned long long) total);
    uint64_t n = set->n;
    CU(ctx, ctx->vrls.reserve(n)); CU(ctx, ctx->stage_soa.reserve(n * 10));
    if (set->medium) CU(ctx, ctx->stage_med.reserve(n));
    size_t soa_bytes = n * 10 * sizeof(float), med_bytes = set->medium ? n * sizeof(uint16_t) : 0;
    int rc = ensure_pinned(ctx, soa_bytes + med_bytes + 64);
    if (rc) return rc;
    cudaEventRecord(ctx->ev[0], ctx->stream);
    if (n) {
        // pageable -> pinned on the host, then one async H2D per array (no batched memcpy API)
        std::memcpy(ctx->pinned, set->soa10, soa_bytes);
        CU(ctx, cudaMemcpyAsync(ctx->stage_soa.ptr, ctx->pinned, soa_bytes, cudaMemcpyHostToDevice, ctx->stream));
        if (set->medium) {
            std::memcpy((char *) ctx->pinned + soa_bytes, set->medium, med_bytes);
            CU(ctx, cudaMemcpyAsync(ctx->stage_med.ptr, (char *) ctx->pinned + soa_bytes, med_bytes, cudaMemcpyHostToDevice, ctx->stream));
        }
        CU(ctx, launch_pack_vrls(ctx->stage_soa.ptr, set->medium ? ctx->stage_med.ptr : nullptr, n, ctx->vrls.ptr, ctx->stream));
        ctx->timing.total_launches += 1; ctx->launches += 1;
    }
    cudaEventRecord(ctx->ev[1], ctx->stream);
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->timing.ms_upload = elapsed(ctx->ev[0], ctx->ev[1]);
    ctx->vrls_changed();
    ctx->n_vrls = n; ctx->vrl_first = set->first_index; ctx->vrl_total = total; ctx->vrl_scale = set->scale;
    return VRLB_OK;
}

uint64_t vrlb_vrls_count(const vrlb_context *ctx) { return ctx ? ctx->n_vrls : 0; }

int vrlb_vrls_download(vrlb_context *ctx, float *soa10, uint16_t *medium, uint64_t cap, uint64_t *n_out) {
    if (!ctx || !soa10) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    uint64_t n = ctx->n_vrls;
    if (n_out) *n_out = n;
    if (cap < n) return fail(ctx, VRLB_ERR_INVALID, "capacity %llu < %llu VRLs", (unsigned long long) cap, (unsigned long long) n);
    if (n == 0) return VRLB_OK;
    CU(ctx, ctx->stage_soa.reserve(n * 10)); CU(ctx, ctx->stage_med.reserve(n));
    CU(ctx, launch_unpack_vrls(ctx->vrls.ptr, n, ctx->stage_soa.ptr, ctx->stage_med.ptr, ctx->stream));
    ctx->launches += 1;
    CU(ctx, cudaMemcpyAsync(soa10, ctx->stage_soa.ptr, n * 10 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    if (medium) CU(ctx, cudaMemcpyAsync(medium, ctx->stage_med.ptr, n * sizeof(uint16_t), cudaMemcpyDeviceToHost, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return VRLB_OK;
}

float vrlb_vrls_scale(const vrlb_context *ctx) { return ctx ? ctx->vrl_scale : 0.f; }

int vrlb_vrls_download_depth(vrlb_context *ctx, int32_t *depth, uint64_t cap) {
    if (!ctx || !depth) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (cap < ctx->n_vrls) return fail(ctx, VRLB_ERR_INVALID, "capacity too small");
    std::vector<DVrl> tmp((size_t) ctx->n_vrls);
    if (ctx->n_vrls) CU(ctx, cudaMemcpy(tmp.data(), ctx->vrls.ptr, tmp.size() * sizeof(DVrl), cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < tmp.size(); ++i) std::memcpy(&depth[i], &tmp[i].pad, 4);
    return VRLB_OK;
}

// light-tracer parameters shared by VRL and volume-photon generation: emitter table (triangles flagged VRLB_SURF_EMITTER,
// area cdf) + the integrator properties of vrlb_generate_desc
static int setup_generate(vrlb_context *ctx, const vrlb_generate_desc *gen, GenParams &p, DevBuf<int32_t> &d_etris, DevBuf<float> &d_ecdf) {
    // emitter table: triangles flagged VRLB_SURF_EMITTER, area cdf
    std::vector<int32_t> etris; std::vector<float> ecdf; double area = 0.0;
    if (!gen->use_laser) {
        for (size_t i = 0; i < ctx->host_tris.size(); ++i) {
            const DTri &t = ctx->host_tris[i];
            if (!(t.flags & VRLB_SURF_EMITTER)) continue;
            double cx = (double) t.e1[1] * t.e2[2] - (double) t.e1[2] * t.e2[1], cy = (double) t.e1[2] * t.e2[0] - (double) t.e1[0] * t.e2[2],
                   cz = (double) t.e1[0] * t.e2[1] - (double) t.e1[1] * t.e2[0];
            area += 0.5 * std::sqrt(cx * cx + cy * cy + cz * cz);
            etris.push_back((int32_t) i); ecdf.push_back((float) area);
        }
        if (etris.empty()) return fail(ctx, VRLB_ERR_INVALID, "no emitter: the scene has no VRLB_SURF_EMITTER triangle and use_laser is 0");
    }
    if (!etris.empty()) {
        CU(ctx, d_etris.reserve(etris.size())); CU(ctx, d_ecdf.reserve(ecdf.size()));
        CU(ctx, cudaMemcpy(d_etris.ptr, etris.data(), etris.size() * 4, cudaMemcpyHostToDevice));
        CU(ctx, cudaMemcpy(d_ecdf.ptr, ecdf.data(), ecdf.size() * 4, cudaMemcpyHostToDevice));
    }
    p.scene = ctx->scene; p.base_seed = gen->seed;
    p.max_depth = gen->max_depth > 0 ? gen->max_depth : 512; p.rr_depth = gen->rr_depth; p.max_steps = gen->max_steps > 0 ? gen->max_steps : 100000;
    p.use_non_linear = gen->use_non_linear; p.use_laser = gen->use_laser; p.start_medium = gen->start_medium;
    p.max_per_particle = 4096; p.min_vrl_length = gen->min_vrl_length;
    for (int k = 0; k < 3; ++k) { p.laser_o[k] = gen->laser_origin[k]; p.laser_d[k] = gen->laser_direction[k]; p.laser_power[k] = gen->laser_power[k]; }
    p.emit_tris = d_etris.ptr; p.emit_cdf = d_ecdf.ptr; p.n_emit_tris = (int32_t) etris.size(); p.emit_area = (float) area;
    p.tri_refl = ctx->have_refl ? ctx->tri_refl.ptr : nullptr;

    return VRLB_OK;
}

int vrlb_vrls_generate(vrlb_context *ctx, const vrlb_generate_desc *gen, uint64_t *n_vrls_out, uint64_t *n_particles_out) {
    if (!ctx || !gen) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (!ctx->have_scene) return fail(ctx, VRLB_ERR_STATE, "vrlb_scene_set must precede vrlb_vrls_generate");
    if (gen->target_vrls == 0) return fail(ctx, VRLB_ERR_INVALID, "target_vrls must be > 0");
    if (gen->start_medium >= ctx->scene.n_media) return fail(ctx, VRLB_ERR_INVALID, "start_medium out of range");
    DevBuf<int32_t> d_etris; DevBuf<float> d_ecdf; DevBuf<uint32_t> d_counts; DevBuf<uint64_t> d_offsets;
    GenParams p{};
    { int rc = setup_generate(ctx, gen, p, d_etris, d_ecdf); if (rc) return rc; }
    const uint64_t target = gen->target_vrls;
    uint64_t have = 0, particles = 0, producing = 0;              // producing: particles that stored at least one VRL (volumePath, vrl.cpp:370)
    uint64_t wave = std::min<uint64_t>(std::max<uint64_t>(target / 4, 4096), 1u << 22);
    std::vector<uint32_t> counts; std::vector<uint64_t> offsets;
    bool done = false;
    for (int iter = 0; iter < 64 && !done; ++iter) {
        if (ctx->cancel.exchange(0)) return fail(ctx, VRLB_ERR_CANCELLED, "cancelled");
        CU(ctx, d_counts.reserve(wave)); CU(ctx, d_offsets.reserve(wave + 1));
        p.n_particles = wave; p.first_particle = particles; p.counts = d_counts.ptr;
        CU(ctx, launch_generate(p, false, ctx->stream)); ctx->launches += 1;
        counts.resize(wave);
        CU(ctx, cudaMemcpyAsync(counts.data(), d_counts.ptr, wave * 4, cudaMemcpyDeviceToHost, ctx->stream));
        CU(ctx, cudaStreamSynchronize(ctx->stream));
        offsets.assign(wave + 1, 0);
        uint64_t used = wave;
        for (uint64_t i = 0; i < wave; ++i) {
            offsets[i + 1] = offsets[i] + counts[i];
            if (counts[i]) ++producing;
            if (have + offsets[i + 1] >= target) { used = i + 1; done = true; break; }       // the particle that fills the map is the last one
        }
        const uint64_t add = offsets[used];
        if (add) {
            // grow the record buffer, keeping what earlier waves wrote
            if (have + add > ctx->vrls.cap) {
                DevBuf<DVrl> bigger; CU(ctx, bigger.reserve(std::max<uint64_t>(have + add, target + target / 8 + 4096)));
                if (have) CU(ctx, cudaMemcpy(bigger.ptr, ctx->vrls.ptr, have * sizeof(DVrl), cudaMemcpyDeviceToDevice));
                ctx->vrls.release(); ctx->vrls = bigger;
            }
            CU(ctx, cudaMemcpyAsync(d_offsets.ptr, offsets.data(), (used + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
            p.n_particles = used; p.offsets = d_offsets.ptr; p.out = ctx->vrls.ptr + have;
            CU(ctx, launch_generate(p, true, ctx->stream)); ctx->launches += 1;
            CU(ctx, cudaStreamSynchronize(ctx->stream));
        }
        have += add; particles += used;
        if (!done) {
            if (have == 0 && iter >= 3) break;                                  // nothing scatters: give up below
            double per = have ? (double) have / (double) particles : 0.05;
            wave = (uint64_t) std::min<double>(std::max<double>(1.2 * (double) (target - have) / per, 1024.0), (double) (1u << 22));
        }
    }
    d_etris.release(); d_ecdf.release(); d_counts.release(); d_offsets.release();
    if (have == 0) return fail(ctx, VRLB_ERR_STATE, "light tracing produced no VRL (no particle reached a medium)");
    // VRLMap::push_back refuses once the map holds target_vrls (vrl_map.h:43-45): the VRLs of the filling particle beyond
    // the target are dropped.  m_scale = 1 / volumePathCount (vrl.cpp:404, 423), and a particle counts when it stored a VRL
    // and the map is not full after it (vrl.cpp:370-371): the filling particle does not.
    if (have > target) have = target;
    const uint64_t counted = done && producing > 1 ? producing - 1 : producing;
    ctx->n_vrls = have; ctx->vrl_first = 0; ctx->vrl_total = have; ctx->vrl_scale = 1.0f / (float) std::max<uint64_t>(counted, 1);
    ctx->vrls_changed();
    if (n_vrls_out) *n_vrls_out = have;
    if (n_particles_out) *n_particles_out = particles;
    return VRLB_OK;
}

// ---- VRL-set transforms (vrl_post.cu) ----------------------------------------------------------------
int vrlb_vrls_to_long(vrlb_context *ctx) {
    if (!ctx) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (!ctx->have_scene) return fail(ctx, VRLB_ERR_STATE, "vrlb_scene_set must precede vrlb_vrls_to_long");
    if (ctx->n_vrls == 0) return VRLB_OK;
    CU(ctx, ctx->counter.reserve(1));
    CU(ctx, cudaMemsetAsync(ctx->counter.ptr, 0, sizeof(unsigned long long), ctx->stream));
    CU(ctx, launch_vrls_to_long(ctx->scene, ctx->vrls.ptr, ctx->n_vrls, ctx->counter.ptr, ctx->stream)); ctx->launches += 1;
    unsigned long long missed = 0;
    CU(ctx, cudaMemcpyAsync(&missed, ctx->counter.ptr, sizeof missed, cudaMemcpyDeviceToHost, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->vrls_changed();
    // VRL::convertToLong: Log(Error, ...) throws in the reference
    if (missed) return fail(ctx, VRLB_ERR_STATE, "Impossible to convert to long vrl... (%llu VRLs see no surface)", missed);
    return VRLB_OK;
}

int vrlb_vrls_dice(vrlb_context *ctx, uint32_t avg_number_dice, uint64_t *n_out) {
    if (!ctx || avg_number_dice == 0) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (!ctx->have_scene) return fail(ctx, VRLB_ERR_STATE, "vrlb_scene_set must precede vrlb_vrls_dice");
    if (ctx->any_grid) return fail(ctx, VRLB_ERR_STATE, "dicing inside a grid medium draws the reference's stochastic transmittance from the "
                                                        "preprocess sampler; only homogeneous / nonlinear media are diced on this path");
    const uint64_t n = ctx->n_vrls;
    if (n_out) *n_out = n;
    if (n == 0) return VRLB_OK;
    // average length: the reference's serial double sum over the map (vrl_map.h:106-110)
    CU(ctx, ctx->stage_soa.reserve(n * 10)); CU(ctx, ctx->stage_med.reserve(n));
    CU(ctx, launch_unpack_vrls(ctx->vrls.ptr, n, ctx->stage_soa.ptr, ctx->stage_med.ptr, ctx->stream)); ctx->launches += 1;
    std::vector<float> len((size_t) n);
    CU(ctx, cudaMemcpyAsync(len.data(), ctx->stage_soa.ptr + 6 * n, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    double avg = 0.0;
    for (float l : len) avg += l;
    avg /= (double) n;
    const float dice_len = (float) (avg / (double) avg_number_dice);      // VRL::dice(scene, sampler, const Float avgLength)
    if (!(dice_len > 0.f)) return fail(ctx, VRLB_ERR_STATE, "dicing needs a positive mean VRL length");
    DevBuf<uint32_t> d_counts; DevBuf<uint64_t> d_offsets; DevBuf<DVrl> diced;
    CU(ctx, d_counts.reserve(n)); CU(ctx, d_offsets.reserve(n + 1));
    CU(ctx, launch_vrls_dice(ctx->vrls.ptr, n, ctx->media.ptr, dice_len, d_counts.ptr, nullptr, nullptr, ctx->stream)); ctx->launches += 1;
    std::vector<uint32_t> counts((size_t) n); std::vector<uint64_t> offsets((size_t) n + 1, 0);
    CU(ctx, cudaMemcpyAsync(counts.data(), d_counts.ptr, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    for (uint64_t i = 0; i < n; ++i) offsets[i + 1] = offsets[i] + counts[i];
    const uint64_t total = offsets[n];
    if (total == 0) return fail(ctx, VRLB_ERR_STATE, "dicing produced no VRL");
    CU(ctx, diced.reserve(total));
    CU(ctx, cudaMemcpyAsync(d_offsets.ptr, offsets.data(), (n + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
    CU(ctx, launch_vrls_dice(ctx->vrls.ptr, n, ctx->media.ptr, dice_len, d_counts.ptr, d_offsets.ptr, diced.ptr, ctx->stream)); ctx->launches += 1;
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->vrls.release(); ctx->vrls = diced; diced.ptr = nullptr; diced.cap = 0;
    d_counts.release(); d_offsets.release();
    ctx->n_vrls = total; ctx->vrl_first = 0; ctx->vrl_total = total;      // m_scale is untouched (vrl.cpp:412-423)
    ctx->vrls_changed();
    if (n_out) *n_out = total;
    return VRLB_OK;
}

// ---- camera segments -------------------------------------------------------------------------------
int vrlb_segments_build(vrlb_context *ctx, const vrlb_camera *cam, const vrlb_options *opt) {
    if (!ctx || !cam || !opt) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (!ctx->have_scene) return fail(ctx, VRLB_ERR_STATE, "vrlb_scene_set must precede vrlb_segments_build");
    if (opt->spp < 1 || opt->block_size < 1 || opt->block_size > 1024) return fail(ctx, VRLB_ERR_INVALID, "spp/block_size out of range");
    if (cam->medium >= ctx->scene.n_media) return fail(ctx, VRLB_ERR_INVALID, "camera medium out of range");
    int rc = ensure_film(ctx, cam->width, cam->height);
    if (rc) return rc;
    if (ctx->bid_w != cam->width || ctx->bid_h != cam->height || ctx->bid_bs != opt->block_size) {
        int nbx, nby;
        ctx->host_block_ids = spiral_ids(cam->width, cam->height, opt->block_size, nbx, nby);
        CU(ctx, ctx->block_ids.reserve(ctx->host_block_ids.size()));
        CU(ctx, cudaMemcpyAsync(ctx->block_ids.ptr, ctx->host_block_ids.data(), ctx->host_block_ids.size() * sizeof(uint32_t),
                                cudaMemcpyHostToDevice, ctx->stream));
        CU(ctx, cudaStreamSynchronize(ctx->stream));
        ctx->bid_w = cam->width; ctx->bid_h = cam->height; ctx->bid_bs = opt->block_size;
    }
    // linear camera rays: at most one in-medium part per null-bounded medium region crossed; start with
    // 2 per sample and grow on overflow (the kernel counts what it needed)
    uint64_t samples = (uint64_t) cam->width * cam->height * (uint64_t) opt->spp;
    uint64_t cap = std::max<uint64_t>(ctx->segs.cap, samples + samples / 8 + 1024);
    for (int attempt = 0; attempt < 2; ++attempt) {
        CU(ctx, ctx->segs.reserve(cap));
        SegmentParams p{};
        p.scene = ctx->scene; p.cam = *cam; p.opt = *opt;
        p.block_ids = ctx->block_ids.ptr;
        p.nbx = (cam->width + opt->block_size - 1) / opt->block_size; p.nby = (cam->height + opt->block_size - 1) / opt->block_size;
        p.n_vrls_total = ctx->vrl_total; p.segs = ctx->segs.ptr; p.cap = ctx->segs.cap; p.counter = ctx->counter.ptr;
        p.film_rgb = ctx->film_rgb.ptr; p.film_aw = ctx->film_aw.ptr;
        p.add_direct = ctx->rank == 0 ? 1 : 0;   // emitter radiance is counted once across VRL shards
        if (ctx->have_refl) {                    // at most one surface query per camera sample
            CU(ctx, ctx->squeries.reserve(samples));
            p.squeries = ctx->squeries.ptr; p.sq_cap = samples; p.sq_counter = ctx->counter.ptr + 1; p.tri_refl = ctx->tri_refl.ptr;   // (also the delta BSDFs' parameters)
        }
        size_t np = (size_t) cam->width * cam->height;
        cudaEventRecord(ctx->ev[0], ctx->stream);
        CU(ctx, cudaMemsetAsync(ctx->counter.ptr, 0, 2 * sizeof(unsigned long long), ctx->stream));
        CU(ctx, cudaMemsetAsync(ctx->film_rgb.ptr, 0, np * 3 * sizeof(float), ctx->stream));
        CU(ctx, cudaMemsetAsync(ctx->film_aw.ptr, 0, np * 2 * sizeof(float), ctx->stream));
        CU(ctx, launch_segments(p, ctx->stream));
        cudaEventRecord(ctx->ev[1], ctx->stream);
        unsigned long long n2[2] = { 0, 0 };
        CU(ctx, cudaMemcpyAsync(n2, ctx->counter.ptr, sizeof n2, cudaMemcpyDeviceToHost, ctx->stream));
        CU(ctx, cudaStreamSynchronize(ctx->stream));
        const unsigned long long n = n2[0];
        ctx->n_squeries = ctx->have_refl ? std::min<uint64_t>(n2[1], samples) : 0;
        ctx->timing.ms_segments = elapsed(ctx->ev[0], ctx->ev[1]);
        ctx->timing.total_launches += 1; ctx->launches += 1;
        if (n <= ctx->segs.cap) { ctx->n_segs = n; ctx->timing.n_segments = n; return VRLB_OK; }
        cap = n;
    }
    return fail(ctx, VRLB_ERR_NOMEM, "segment buffer overflow");
}

int vrlb_segments_upload(vrlb_context *ctx, const vrlb_segment *segs, uint64_t n, int32_t width, int32_t height, int32_t spp) {
    if (!ctx || (n && !segs)) return VRLB_ERR_INVALID;
    (void) spp;
    cudaSetDevice(ctx->device);
    int rc = ensure_film(ctx, width, height);
    if (rc) return rc;
    for (uint64_t i = 0; i < n; ++i)
        if (segs[i].pixel >= (uint64_t) width * height) return fail(ctx, VRLB_ERR_INVALID, "segment %llu: pixel out of range", (unsigned long long) i);
    CU(ctx, ctx->segs.reserve(n));
    if (n) CU(ctx, cudaMemcpyAsync(ctx->segs.ptr, segs, n * sizeof(vrlb_segment), cudaMemcpyHostToDevice, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->n_segs = n; ctx->timing.n_segments = n;
    ctx->n_squeries = 0;                         // the surface queries belonged to the camera paths just replaced
    return VRLB_OK;
}

int vrlb_segments_download(vrlb_context *ctx, vrlb_segment *segs, uint64_t cap, uint64_t *n_out) {
    if (!ctx) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (n_out) *n_out = ctx->n_segs;
    if (!segs) return VRLB_OK;
    if (cap < ctx->n_segs) return fail(ctx, VRLB_ERR_INVALID, "capacity too small");
    if (ctx->n_segs) CU(ctx, cudaMemcpyAsync(segs, ctx->segs.ptr, ctx->n_segs * sizeof(vrlb_segment), cudaMemcpyDeviceToHost, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return VRLB_OK;
}

uint64_t vrlb_segments_count(const vrlb_context *ctx) { return ctx ? ctx->n_segs : 0; }

// ---- gather ----------------------------------------------------------------------------------------
int vrlb_film_clear(vrlb_context *ctx) {
    if (!ctx) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (ctx->width == 0) return fail(ctx, VRLB_ERR_STATE, "no film yet");
    size_t np = (size_t) ctx->width * ctx->height;
    CU(ctx, cudaMemsetAsync(ctx->film_rgb.ptr, 0, np * 3 * sizeof(float), ctx->stream));
    CU(ctx, cudaMemsetAsync(ctx->film_aw.ptr, 0, np * 2 * sizeof(float), ctx->stream));
    return VRLB_OK;
}

// VRLMap::query's ELightCutAcceleration branch (vrl_map.h:198-201) for every resident segment
static int lightcut_gather_impl(vrlb_context *ctx, const vrlb_options *opt, float *seg_out_dev, float *seg_stats_dev) {
    if (!ctx->lc) return fail(ctx, VRLB_ERR_STATE, "use_light_cut: no light tree resident (vrlb_lightcut_build after the VRL set is final)");
    if (ctx->any_grid) return fail(ctx, VRLB_ERR_UNSUPPORTED, "use_light_cut with a grid medium: the reference raises \"Heterogeneous Lightcut not implemented\" (vrl_lightcut.h:485)");
    if (opt->use_nl_atomic_query && opt->use_non_linear_camera) return fail(ctx, VRLB_ERR_UNSUPPORTED, "use_light_cut with use_nl_atomic_query: multi-part queries through the cut are not built (queries are per part, vrl.cpp:619-628)");
    if (!(opt->threshold_error >= 0.f)) return fail(ctx, VRLB_ERR_INVALID, "threshold_error %g", (double) opt->threshold_error);
    if (lightcut_n_vrls(ctx->lc) != ctx->n_vrls) return fail(ctx, VRLB_ERR_STATE, "light tree is stale");
    LightcutParams p{};
    p.scene = ctx->scene; p.segs = ctx->segs.ptr; p.n_segs = ctx->n_segs;
    p.nodes = lightcut_nodes(ctx->lc); p.leaf_order = lightcut_leaf_order(ctx->lc); p.jump = lightcut_jump(ctx->lc); p.leaf_geom = lightcut_leaf_geom(ctx->lc);
    p.vrls = ctx->vrls.ptr; p.n_vrls = ctx->n_vrls; p.first_index = ctx->vrl_first; p.total = ctx->vrl_total;
    p.root = lightcut_root(ctx->lc); p.scale = ctx->vrl_scale; p.err = opt->threshold_error; p.better_dist = opt->threshold_better_dist;
    p.q = (uint32_t) opt->samples_per_query;
    // a heap never holds more than the tree's inner nodes
    // measured largest heaps at threshold_error 0.02 on the bench frame: 700 / 1852 / 5847 at 4096 / 16384 / 100000 VRLs
    // (~0.75 N^(2/3)); the default leaves a factor 4 and never goes below 4096
    const uint64_t want = opt->lightcut_heap > 0 ? (uint64_t) opt->lightcut_heap
                                                 : std::max<uint64_t>(4096u, (uint64_t) (4.0 * std::cbrt((double) ctx->n_vrls) * std::cbrt((double) ctx->n_vrls)));
    p.cap = (uint32_t) ((std::max<uint64_t>(2, std::min<uint64_t>(want, ctx->n_vrls)) + 3) & ~3ull);      // runs of whole 32-byte sectors
    int n_sms = 148; cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, ctx->device);
    const uint32_t grid = lightcut_grid(p.n_segs, p.cap, n_sms, (size_t) 64 << 30);
    const size_t entries = (size_t) grid * 128 * p.cap;
    CU(ctx, ctx->lc_heap.reserve(entries)); CU(ctx, ctx->lc_slots.reserve(entries)); CU(ctx, ctx->lc_stats.reserve(8));
    CU(ctx, cudaMemsetAsync(ctx->lc_stats.ptr, 0, 8 * sizeof(unsigned long long), ctx->stream));
    p.heap_key = ctx->lc_heap.ptr; p.slot_est = ctx->lc_slots.ptr; p.stats = ctx->lc_stats.ptr;
    p.film_rgb = seg_out_dev ? nullptr : ctx->film_rgb.ptr; p.seg_out = seg_out_dev; p.seg_stats = seg_stats_dev;
    const bool vis = opt->visibility != 0 && ctx->scene.n_tris > 0;
    cudaEventRecord(ctx->ev[2], ctx->stream);
    CU(ctx, launch_lightcut_gather(p, grid, opt->use_uniform_sampling != 0, vis, ctx->stream));
    cudaEventRecord(ctx->ev[3], ctx->stream);
    unsigned long long st[5] = { 0, 0, 0, 0, 0 };
    CU(ctx, cudaMemcpyAsync(st, ctx->lc_stats.ptr, sizeof st, cudaMemcpyDeviceToHost, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->lc_last.queries = st[0] * p.q; ctx->lc_last.evaluations = st[1]; ctx->lc_last.refinements = st[2]; ctx->lc_last.overflows = st[3]; ctx->lc_last.max_heap = st[4];
    const bool launched = p.n_segs && p.n_vrls;
    ctx->timing.gather_launches = launched ? 1 : 0;
    ctx->timing.total_launches += launched ? 1 : 0; ctx->launches += launched ? 1 : 0;
    ctx->timing.n_evals = st[1] + st[0] * p.q;           // fresh evaluations + one root evaluation per query sample
    ctx->timing.ms_gather = elapsed(ctx->ev[2], ctx->ev[3]);
    return VRLB_OK;
}

static int gather_impl(vrlb_context *ctx, const vrlb_options *opt, float *seg_out_dev) {
    if (!ctx->have_scene) return fail(ctx, VRLB_ERR_STATE, "no scene resident");
    if (ctx->width == 0) return fail(ctx, VRLB_ERR_STATE, "no camera segments resident");
    if (opt->samples_per_query < 1) return fail(ctx, VRLB_ERR_INVALID, "samples_per_query must be >= 1");
    if (opt->math_mode < 0 || opt->math_mode > 2) return fail(ctx, VRLB_ERR_INVALID, "math_mode %d unknown (0 exact, 1 fast + conditioning guard, 2 fast unguarded)", opt->math_mode);
    if (ctx->cancel.exchange(0)) return fail(ctx, VRLB_ERR_CANCELLED, "cancelled");
    if (opt->use_light_cut) return lightcut_gather_impl(ctx, opt, seg_out_dev, nullptr);
    GatherParams p{};
    p.scene = ctx->scene; p.segs = ctx->segs.ptr; p.n_segs = ctx->n_segs;
    p.vrls = ctx->vrls.ptr; p.n_vrls = ctx->n_vrls; p.first_index = ctx->vrl_first; p.total = ctx->vrl_total;
    p.scale = ctx->vrl_scale; p.scale_rr = opt->scale_rr; p.q = (uint32_t) opt->samples_per_query;
    p.film_rgb = ctx->film_rgb.ptr; p.seg_out = seg_out_dev;
    if (!ctx->wd) { CU(ctx, cudaHostAlloc((void **) &ctx->wd, 32 * sizeof(uint32_t), cudaHostAllocMapped)); std::memset(ctx->wd, 0, 32 * sizeof(uint32_t)); }
    p.wd = ctx->wd;
    const bool vis = opt->visibility != 0 && ctx->scene.n_tris > 0;
    // grid media run on the exactly-rounded kernel in every math mode (their stochastic transmittance is a chain of
    // accept/reject decisions on logf results; a fast variant would flip decisions against the oracle)
    // use_nl_atomic_query && use_non_linear_camera forces uniform sampling (vrl.cpp:62) and runs on the exact kernel
    const bool nla = opt->use_nl_atomic_query != 0 && opt->use_non_linear_camera != 0;
    const bool uniform = opt->use_uniform_sampling != 0 || nla;
    p.nla = nla ? 1 : 0;
    const bool fast = opt->math_mode != 0 && !ctx->any_grid && !nla;
    if (fast && ctx->fast_dirty && ctx->n_vrls) {
        CU(ctx, ctx->vrls_fast.reserve(ctx->n_vrls));
        CU(ctx, launch_prep_fast_vrls(ctx->vrls.ptr, ctx->n_vrls, ctx->media.ptr, ctx->vrl_scale, ctx->vrls_fast.ptr, ctx->stream));
        ctx->timing.total_launches += 1; ctx->launches += 1;
        if (ctx->scene.n_tris > 0) {         // the visibility kernel's 64-byte records (safe-box intervals of this scene's shells)
            CU(ctx, ctx->vrls_vis.reserve(ctx->n_vrls));
            CU(ctx, launch_prep_vis_vrls(ctx->vrls.ptr, ctx->n_vrls, ctx->scene, ctx->vrl_scale, ctx->vrls_vis.ptr, ctx->stream));
            ctx->timing.total_launches += 1; ctx->launches += 1;
        }
        ctx->fast_dirty = false;
    }
    p.vrls_fast = ctx->vrls_fast.ptr; p.vrls_vis = ctx->vrls_vis.ptr;
    // Conditioning guard: evals whose VRL passes closer than 1 % of the scene's extent to the camera part take the exactly
    // rounded path.  The fast path's position error is ~1e-6 of the coordinates' magnitude and the equi-angular pdf amplifies
    // it by 1 / h, so the threshold has to scale with the scene (0.2 on the +-10 box-fog scene the kernel was tuned on).
    p.guard_h = opt->reserved[0] ? (float) opt->reserved[0] * 1e-3f : 0.01f * ctx->scene_extent / 1.001f;   // dev knob (milli-units); 0 = shipped default
    p.guard_z = opt->reserved[1] ? (float) opt->reserved[1] : 8.0f;
    // The film path cuts the sum over the VRLs into launches of ~1/4 s so that vrlb_cancel has the latency of the
    // reference's per-block should_stop() (integrator.cpp:124): the host keeps one launch queued ahead of the running one
    // and looks at the flag before every launch.  Cuts sit on multiples of 512 VRLs, the visibility kernel's trace-batch
    // window, so a chunked gather adds the same per-eval values as one launch would (the film sums them by fp32 RED
    // either way).  Per-segment output (tests) stays one launch: seg_out is assigned, not accumulated.
    const double rate = !fast ? (ctx->any_grid ? 2e9 : 2e10) : 4e10;                 // evals / s, low estimates per kernel
    uint64_t chunk = p.n_vrls;
    if (!seg_out_dev && p.n_segs) {
        const double per_launch = 0.25 * rate / ((double) p.n_segs * p.q);
        chunk = std::max<uint64_t>(512, (uint64_t) per_launch / 512 * 512);
        if (chunk * 5 / 4 >= p.n_vrls) chunk = p.n_vrls;                                 // a short tail is not worth a launch
    }
    const DVrl *v0 = p.vrls, *vf0 = p.vrls_fast; const DVrlV *vv0 = p.vrls_vis;
    const uint64_t n_all = p.n_vrls, first_all = p.first_index;
    uint32_t n_launch = 0;
    cudaEventRecord(ctx->ev[2], ctx->stream);
    for (uint64_t lo = 0; lo < n_all; lo += chunk) {
        if (n_launch && ctx->cancel.exchange(0)) {
            cudaStreamSynchronize(ctx->stream);
            return fail(ctx, VRLB_ERR_CANCELLED, "cancelled after %llu of %llu VRLs", (unsigned long long) lo, (unsigned long long) n_all);
        }
        p.n_vrls = std::min<uint64_t>(chunk, n_all - lo); p.first_index = first_all + lo;
        p.vrls = v0 + lo; p.vrls_fast = vf0 ? vf0 + lo : nullptr; p.vrls_vis = vv0 ? vv0 + lo : nullptr;
        if (!fast) CU(ctx, launch_gather(p, uniform, opt->rr_vrl != 0, vis, ctx->any_grid, ctx->stream));
        else CU(ctx, launch_gather_fast(p, uniform, opt->rr_vrl != 0, vis, ctx->all_iso, opt->math_mode == 1, ctx->stream));
        if (chunk < n_all) {                     // wait for the launch BEFORE this one: one stays queued behind the running one
            if (n_launch) CU(ctx, cudaEventSynchronize(ctx->ev[8 + ((n_launch - 1) & 1)]));
            cudaEventRecord(ctx->ev[8 + (n_launch & 1)], ctx->stream);
        }
        ++n_launch;
    }
    cudaEventRecord(ctx->ev[3], ctx->stream);
    if (!p.n_segs) n_launch = 0;
    ctx->timing.gather_launches = n_launch;
    ctx->timing.total_launches += n_launch; ctx->launches += n_launch;
    ctx->timing.n_evals = p.n_segs * n_all * p.q;
    return VRLB_OK;
}

static int surface_gather_impl(vrlb_context *ctx, float *out_dev);

// direct term from the resident photon set (no-op without one); film path when seg_out_dev is null
static int photon_gather_impl(vrlb_context *ctx, const vrlb_options *opt, float *seg_out_dev) {
    if (ctx->vol.n == 0 && !seg_out_dev) return VRLB_OK;
    if (!seg_out_dev && ctx->n_ranks > 1 && ctx->rank != 0) return VRLB_OK;   // VRL shards: the films are summed, the direct term is rank 0's
    if (!ctx->have_scene) return fail(ctx, VRLB_ERR_STATE, "no scene resident");
    if (ctx->width == 0) return fail(ctx, VRLB_ERR_STATE, "no camera segments resident");
    PhotonGatherParams p{};
    p.scene = ctx->scene; p.segs = ctx->segs.ptr; p.n_segs = ctx->n_segs;
    p.photons = ctx->vol.photons.ptr; p.cell_start = ctx->vol.cells.ptr; p.n_photons = ctx->vol.n;
    for (int k = 0; k < 3; ++k) { p.grid_lo[k] = ctx->vol.lo[k]; p.grid_res[k] = ctx->vol.res[k]; }
    p.inv_cell = ctx->vol.inv_cell; p.lookup_radius = ctx->vol.radius; p.scale = ctx->vol.scale;
    p.nl_camera = opt->use_non_linear_camera;
    p.film_rgb = seg_out_dev ? nullptr : ctx->film_rgb.ptr; p.seg_out = seg_out_dev;
    CU(ctx, launch_photon_gather(p, ctx->stream));
    if (p.n_segs) { ctx->timing.total_launches += 1; ctx->launches += 1; }
    return VRLB_OK;
}

int vrlb_gather(vrlb_context *ctx, const vrlb_options *opt) {
    if (!ctx || !opt) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    int rc = gather_impl(ctx, opt, nullptr);
    if (rc == VRLB_OK) rc = photon_gather_impl(ctx, opt, nullptr);
    if (rc == VRLB_OK) rc = surface_gather_impl(ctx, nullptr);
    return rc;
}

// Photon set -> resident map: uniform grid over the photons' bounds, counting sort by cell (set order inside a cell, so
// the sums over a cell are deterministic)
static int bin_photons(vrlb_context *ctx, const vrlb_photonset_desc *set, vrlb_context::PhotonMapRes &m) {
    if (set->n == 0) { m.n = 0; return VRLB_OK; }
    if (!set->pos || !set->dir || !set->power) return fail(ctx, VRLB_ERR_INVALID, "photon arrays missing");
    if (!(set->lookup_radius > 0.f)) return fail(ctx, VRLB_ERR_INVALID, "lookup_radius must be > 0");
    if (set->n > 0xfffffff0ull) return fail(ctx, VRLB_ERR_INVALID, "too many photons");
    const uint64_t n = set->n;
    // uniform grid over the photons' bounds, cell = 2 r_max (r_max = 1.25 lookup_radius): a gather sphere meets <= 8 cells
    float lo[3] = { set->pos[0], set->pos[1], set->pos[2] }, hi[3] = { lo[0], lo[1], lo[2] };
    for (uint64_t i = 0; i < n; ++i) for (int k = 0; k < 3; ++k) { float v = set->pos[i * 3 + k]; if (!(v == v)) return fail(ctx, VRLB_ERR_INVALID, "NaN photon position"); lo[k] = std::min(lo[k], v); hi[k] = std::max(hi[k], v); }
    float cell = 2.5f * set->lookup_radius;
    for (;;) {
        double cells = 1;
        for (int k = 0; k < 3; ++k) { m.res[k] = std::max(1, (int) std::floor((hi[k] - lo[k]) / cell) + 1); cells *= m.res[k]; }
        if (cells <= 16.0e6) break;
        cell *= 1.5f;
    }
    const size_t n_cells = (size_t) m.res[0] * m.res[1] * m.res[2];
    const float inv_cell = 1.0f / cell;
    std::vector<uint32_t> cell_of((size_t) n), order((size_t) n), start(n_cells + 1, 0);
    for (uint64_t i = 0; i < n; ++i) {
        uint32_t c[3];
        for (int k = 0; k < 3; ++k) c[k] = (uint32_t) std::min(m.res[k] - 1, std::max(0, (int) std::floor((set->pos[i * 3 + k] - lo[k]) * inv_cell)));
        cell_of[i] = (c[2] * (uint32_t) m.res[1] + c[1]) * (uint32_t) m.res[0] + c[0];
        start[cell_of[i] + 1] += 1;
    }
    for (size_t c = 0; c < n_cells; ++c) start[c + 1] += start[c];
    { std::vector<uint32_t> cur(start.begin(), start.end() - 1);
      for (uint64_t i = 0; i < n; ++i) order[cur[cell_of[i]]++] = (uint32_t) i; }           // counting sort: set order inside a cell
    std::vector<DPhoton> sorted((size_t) n);
    for (uint64_t k = 0; k < n; ++k) {
        const uint64_t i = order[k];
        sorted[k] = DPhoton{ set->pos[i * 3], set->pos[i * 3 + 1], set->pos[i * 3 + 2], 0.f, set->dir[i * 3], set->dir[i * 3 + 1], set->dir[i * 3 + 2], 0.f,
                             set->power[i * 3], set->power[i * 3 + 1], set->power[i * 3 + 2], 0.f };
    }
    CU(ctx, m.photons.reserve(n)); CU(ctx, m.cells.reserve(n_cells + 1));
    CU(ctx, cudaMemcpyAsync(m.photons.ptr, sorted.data(), n * sizeof(DPhoton), cudaMemcpyHostToDevice, ctx->stream));
    CU(ctx, cudaMemcpyAsync(m.cells.ptr, start.data(), (n_cells + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    for (int k = 0; k < 3; ++k) m.lo[k] = lo[k];
    m.inv_cell = inv_cell; m.scale = set->scale; m.radius = set->lookup_radius; m.n = n;
    return VRLB_OK;
}

int vrlb_photons_upload(vrlb_context *ctx, const vrlb_photonset_desc *set) {
    if (!ctx || !set) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    return bin_photons(ctx, set, ctx->vol);
}

// Volume photons by light tracing: the same particles as vrlb_vrls_generate (particle i = PCG32 seed(seed + i)), one
// photon at a particle's first medium scattering, in particle order until `capacity` photons are stored.  Scale factor
// as the reference counts it (vrl.cpp:372-373, 431): a path is counted when it deposited AND the map is not full after
// it, i.e. capacity - 1 paths for a filled map.
int vrlb_photons_generate(vrlb_context *ctx, const vrlb_generate_desc *gen, uint64_t capacity, float lookup_radius,
                          uint64_t *n_photons_out, uint64_t *n_particles_out) {
    if (!ctx || !gen) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (!ctx->have_scene) return fail(ctx, VRLB_ERR_STATE, "vrlb_scene_set must precede vrlb_photons_generate");
    if (gen->start_medium >= ctx->scene.n_media) return fail(ctx, VRLB_ERR_INVALID, "start_medium out of range");
    if (n_photons_out) *n_photons_out = 0;
    if (n_particles_out) *n_particles_out = 0;
    if (capacity == 0) { ctx->vol.n = 0; return VRLB_OK; }            // PhotonMap(m_useDirectIllum ? m_volumePhotons : 0), vrl.cpp:79
    DevBuf<int32_t> d_etris; DevBuf<float> d_ecdf; DevBuf<uint32_t> d_counts, d_flags; DevBuf<DPhoton> d_ph;
    GenParams p{};
    { int rc = setup_generate(ctx, gen, p, d_etris, d_ecdf); if (rc) return rc; }
    std::vector<float> pos, dir, power;
    std::vector<uint32_t> flags; std::vector<DPhoton> ph;
    uint64_t have = 0, particles = 0;
    uint64_t wave = std::min<uint64_t>(std::max<uint64_t>(capacity + capacity / 4, 4096), 1u << 20);
    bool done = false;
    for (int iter = 0; iter < 256 && !done; ++iter) {
        if (ctx->cancel.exchange(0)) return fail(ctx, VRLB_ERR_CANCELLED, "cancelled");
        CU(ctx, d_counts.reserve(wave)); CU(ctx, d_flags.reserve(wave)); CU(ctx, d_ph.reserve(wave));
        CU(ctx, cudaMemsetAsync(d_flags.ptr, 0, wave * 4, ctx->stream));
        p.n_particles = wave; p.first_particle = particles; p.counts = d_counts.ptr; p.photon_out = d_ph.ptr; p.photon_flag = d_flags.ptr;
        CU(ctx, launch_generate(p, false, ctx->stream)); ctx->launches += 1;
        flags.resize(wave); ph.resize(wave);
        CU(ctx, cudaMemcpyAsync(flags.data(), d_flags.ptr, wave * 4, cudaMemcpyDeviceToHost, ctx->stream));
        CU(ctx, cudaMemcpyAsync(ph.data(), d_ph.ptr, wave * sizeof(DPhoton), cudaMemcpyDeviceToHost, ctx->stream));
        CU(ctx, cudaStreamSynchronize(ctx->stream));
        uint64_t used = wave;
        for (uint64_t i = 0; i < wave; ++i) {
            if (!flags[i]) continue;
            const DPhoton &q = ph[i];
            pos.insert(pos.end(), { q.px, q.py, q.pz }); dir.insert(dir.end(), { q.dx, q.dy, q.dz }); power.insert(power.end(), { q.wr, q.wg, q.wb });
            if (++have == capacity) { used = i + 1; done = true; break; }
        }
        particles += used;
        if (!done && have == 0 && iter >= 3) break;                       // nothing scatters in a medium
    }
    d_etris.release(); d_ecdf.release(); d_counts.release(); d_flags.release(); d_ph.release();
    if (n_photons_out) *n_photons_out = have;
    if (n_particles_out) *n_particles_out = particles;
    if (have == 0) { ctx->vol.n = 0; return VRLB_OK; }
    // directVolumePath starts out true for every particle traced while the map is not full and is counted when the map is
    // still not full after it (vrl.cpp:124, 372-373): every such particle counts, scattered or not, but not the one that fills the map
    const uint64_t counted = done ? particles - 1 : particles;
    vrlb_photonset_desc d{};
    d.pos = pos.data(); d.dir = dir.data(); d.power = power.data(); d.n = have;
    d.scale = counted ? 1.0f / (float) counted : 1.0f; d.lookup_radius = lookup_radius;
    return vrlb_photons_upload(ctx, &d);
}

// Global / caustic surface photons by light tracing: the same particles again, one photon per smooth-surface hit that
// was (caustic) or was not (global) preceded by a transmission as the last surface event (vrl.cpp:776-780) (several per
// particle: count pass, host scan, write pass like the VRLs), in particle order until `capacity` are stored.
// Scale = 1 / surfacePathCount: every particle traced while the map was not full counts, except the one that fills it
// (vrl.cpp:124, 368-369, 379).
int vrlb_surface_photons_generate(vrlb_context *ctx, int which, const vrlb_generate_desc *gen, uint64_t capacity, float lookup_radius,
                                  int32_t lookup_size, uint64_t *n_photons_out, uint64_t *n_particles_out) {
    if (!ctx || !gen) return VRLB_ERR_INVALID;
    if (which != VRLB_PHOTONMAP_GLOBAL && which != VRLB_PHOTONMAP_CAUSTIC) return fail(ctx, VRLB_ERR_INVALID, "unknown photon map %d", which);
    cudaSetDevice(ctx->device);
    if (!ctx->have_scene) return fail(ctx, VRLB_ERR_STATE, "vrlb_scene_set must precede vrlb_surface_photons_generate");
    if (gen->start_medium >= ctx->scene.n_media) return fail(ctx, VRLB_ERR_INVALID, "start_medium out of range");
    if (n_photons_out) *n_photons_out = 0;
    if (n_particles_out) *n_particles_out = 0;
    vrlb_context::PhotonMapRes &map = ctx->surf[which];
    if (capacity == 0 || !ctx->have_refl) { map.n = 0; return VRLB_OK; }      // no reflecting surface: nothing to store
    if (lookup_size < 1 || lookup_size > VRLB_MAX_LOOKUP) return fail(ctx, VRLB_ERR_INVALID, "lookup_size must lie in 1 .. %d", VRLB_MAX_LOOKUP);
    DevBuf<int32_t> d_etris; DevBuf<float> d_ecdf; DevBuf<uint32_t> d_counts; DevBuf<uint64_t> d_offsets; DevBuf<DPhoton> d_ph;
    GenParams p{};
    { int rc = setup_generate(ctx, gen, p, d_etris, d_ecdf); if (rc) return rc; }
    p.collect_surface = which == VRLB_PHOTONMAP_CAUSTIC ? 2 : 1;
    std::vector<DPhoton> all; std::vector<DPhoton> ph;
    std::vector<uint32_t> counts; std::vector<uint64_t> offsets;
    uint64_t particles = 0;
    const uint64_t wave = std::min<uint64_t>(std::max<uint64_t>(capacity / 2, 4096), 1u << 20);
    bool done = false;
    for (int iter = 0; iter < 256 && !done; ++iter) {
        if (ctx->cancel.exchange(0)) return fail(ctx, VRLB_ERR_CANCELLED, "cancelled");
        CU(ctx, d_counts.reserve(wave)); CU(ctx, d_offsets.reserve(wave + 1));
        p.n_particles = wave; p.first_particle = particles; p.counts = d_counts.ptr; p.offsets = nullptr; p.surf_out = nullptr;
        CU(ctx, launch_generate(p, false, ctx->stream)); ctx->launches += 1;
        counts.resize(wave);
        CU(ctx, cudaMemcpyAsync(counts.data(), d_counts.ptr, wave * 4, cudaMemcpyDeviceToHost, ctx->stream));
        CU(ctx, cudaStreamSynchronize(ctx->stream));
        offsets.assign(wave + 1, 0);
        uint64_t used = wave;
        for (uint64_t i = 0; i < wave; ++i) {
            offsets[i + 1] = offsets[i] + counts[i];
            if (all.size() + offsets[i + 1] >= capacity) { used = i + 1; done = true; break; }   // this particle fills the map
        }
        const uint64_t add = offsets[used];
        if (add) {
            CU(ctx, d_ph.reserve(add));
            CU(ctx, cudaMemcpyAsync(d_offsets.ptr, offsets.data(), (used + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
            p.n_particles = used; p.offsets = d_offsets.ptr; p.surf_out = d_ph.ptr;
            CU(ctx, launch_generate(p, true, ctx->stream)); ctx->launches += 1;
            ph.resize(add);
            CU(ctx, cudaMemcpyAsync(ph.data(), d_ph.ptr, add * sizeof(DPhoton), cudaMemcpyDeviceToHost, ctx->stream));
            CU(ctx, cudaStreamSynchronize(ctx->stream));
            all.insert(all.end(), ph.begin(), ph.end());
        }
        particles += used;
        if (!done && all.empty() && iter >= 3) break;                       // no particle ever meets a reflecting surface
    }
    d_etris.release(); d_ecdf.release(); d_counts.release(); d_offsets.release(); d_ph.release();
    if (all.size() > capacity) all.resize(capacity);                        // PhotonMap::insert refuses once full
    if (n_photons_out) *n_photons_out = all.size();
    if (n_particles_out) *n_particles_out = particles;
    if (all.empty()) { map.n = 0; return VRLB_OK; }
    const uint64_t counted = done ? particles - 1 : particles;
    std::vector<float> pos(all.size() * 3), dir(all.size() * 3), power(all.size() * 3);
    for (size_t i = 0; i < all.size(); ++i) {
        pos[i * 3] = all[i].px; pos[i * 3 + 1] = all[i].py; pos[i * 3 + 2] = all[i].pz;
        dir[i * 3] = all[i].dx; dir[i * 3 + 1] = all[i].dy; dir[i * 3 + 2] = all[i].dz;
        power[i * 3] = all[i].wr; power[i * 3 + 1] = all[i].wg; power[i * 3 + 2] = all[i].wb;
    }
    vrlb_photonset_desc d{};
    d.pos = pos.data(); d.dir = dir.data(); d.power = power.data(); d.n = all.size();
    d.scale = counted ? 1.0f / (float) counted : 1.0f; d.lookup_radius = lookup_radius;
    int rc = bin_photons(ctx, &d, map);
    if (rc == VRLB_OK) map.lookup_size = lookup_size;
    return rc;
}

uint64_t vrlb_photons_count(const vrlb_context *ctx) { return ctx ? ctx->vol.n : 0; }

static int download_photons(vrlb_context *ctx, const vrlb_context::PhotonMapRes &m, float *pos, float *dir, float *power, uint64_t cap, float *scale_out) {
    if (scale_out) *scale_out = m.scale;
    const uint64_t n = m.n;
    if (cap < n) return fail(ctx, VRLB_ERR_INVALID, "capacity %llu < %llu photons", (unsigned long long) cap, (unsigned long long) n);
    if (n == 0) return VRLB_OK;
    std::vector<DPhoton> tmp((size_t) n);
    CU(ctx, cudaMemcpy(tmp.data(), m.photons.ptr, n * sizeof(DPhoton), cudaMemcpyDeviceToHost));
    for (uint64_t i = 0; i < n; ++i) {
        if (pos)   { pos[i * 3] = tmp[i].px; pos[i * 3 + 1] = tmp[i].py; pos[i * 3 + 2] = tmp[i].pz; }
        if (dir)   { dir[i * 3] = tmp[i].dx; dir[i * 3 + 1] = tmp[i].dy; dir[i * 3 + 2] = tmp[i].dz; }
        if (power) { power[i * 3] = tmp[i].wr; power[i * 3 + 1] = tmp[i].wg; power[i * 3 + 2] = tmp[i].wb; }
    }
    return VRLB_OK;
}

int vrlb_photons_download(vrlb_context *ctx, float *pos, float *dir, float *power, uint64_t cap, float *scale_out) {
    if (!ctx) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    return download_photons(ctx, ctx->vol, pos, dir, power, cap, scale_out);
}

int vrlb_photon_gather_segments(vrlb_context *ctx, const vrlb_options *opt, float *direct_out) {
    if (!ctx || !opt || !direct_out) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    DevBuf<float> tmp;
    CU(ctx, tmp.reserve(ctx->n_segs * 3 + 1));
    CU(ctx, cudaMemsetAsync(tmp.ptr, 0, (ctx->n_segs * 3 + 1) * sizeof(float), ctx->stream));
    int rc = photon_gather_impl(ctx, opt, tmp.ptr);
    if (rc == VRLB_OK) {
        cudaError_t e = cudaMemcpyAsync(direct_out, tmp.ptr, ctx->n_segs * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) rc = fail(ctx, VRLB_ERR_CUDA, "%s", cudaGetErrorString(e));
    }
    tmp.release();
    return rc;
}

// ---- surface photon maps (surface_photons.cu) ------------------------------------------------------
int vrlb_surface_photons_upload(vrlb_context *ctx, int which, const vrlb_photonset_desc *set, int32_t lookup_size) {
    if (!ctx || !set) return VRLB_ERR_INVALID;
    if (which != VRLB_PHOTONMAP_GLOBAL && which != VRLB_PHOTONMAP_CAUSTIC) return fail(ctx, VRLB_ERR_INVALID, "unknown photon map %d", which);
    if (set->n && (lookup_size < 1 || lookup_size > VRLB_MAX_LOOKUP)) return fail(ctx, VRLB_ERR_INVALID, "lookup_size must lie in 1 .. %d", VRLB_MAX_LOOKUP);
    cudaSetDevice(ctx->device);
    int rc = bin_photons(ctx, set, ctx->surf[which]);
    if (rc == VRLB_OK) ctx->surf[which].lookup_size = lookup_size;
    return rc;
}
uint64_t vrlb_surface_photons_count(const vrlb_context *ctx, int which) {
    return (ctx && (which == VRLB_PHOTONMAP_GLOBAL || which == VRLB_PHOTONMAP_CAUSTIC)) ? ctx->surf[which].n : 0;
}
int vrlb_surface_photons_download(vrlb_context *ctx, int which, float *pos, float *dir, float *power, uint64_t cap, float *scale_out) {
    if (!ctx || (which != VRLB_PHOTONMAP_GLOBAL && which != VRLB_PHOTONMAP_CAUSTIC)) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    return download_photons(ctx, ctx->surf[which], pos, dir, power, cap, scale_out);
}
int vrlb_surface_queries_upload(vrlb_context *ctx, const vrlb_surface_query *q, uint64_t n, int32_t width, int32_t height) {
    if (!ctx || (n && !q) || width <= 0 || height <= 0) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    for (uint64_t i = 0; i < n; ++i) if (q[i].pixel >= (uint64_t) width * (uint64_t) height) return fail(ctx, VRLB_ERR_INVALID, "surface query %llu: pixel out of range", (unsigned long long) i);
    { int rc = ensure_film(ctx, width, height); if (rc) return rc; }
    CU(ctx, ctx->squeries.reserve(std::max<uint64_t>(n, 1)));
    if (n) CU(ctx, cudaMemcpyAsync(ctx->squeries.ptr, q, n * sizeof(vrlb_surface_query), cudaMemcpyHostToDevice, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->n_squeries = n;
    return VRLB_OK;
}
int vrlb_surface_queries_download(vrlb_context *ctx, vrlb_surface_query *q, uint64_t cap, uint64_t *n_out) {
    if (!ctx) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (n_out) *n_out = ctx->n_squeries;
    if (!q) return VRLB_OK;
    if (cap < ctx->n_squeries) return fail(ctx, VRLB_ERR_INVALID, "capacity %llu < %llu surface queries", (unsigned long long) cap, (unsigned long long) ctx->n_squeries);
    if (ctx->n_squeries) CU(ctx, cudaMemcpy(q, ctx->squeries.ptr, ctx->n_squeries * sizeof(vrlb_surface_query), cudaMemcpyDeviceToHost));
    return VRLB_OK;
}
static void fill_grid(const vrlb_context::PhotonMapRes &m, PhotonGridDev &g) {
    g.photons = m.photons.ptr; g.cell_start = m.cells.ptr; g.n = m.n;
    for (int k = 0; k < 3; ++k) { g.grid_lo[k] = m.lo[k]; g.grid_res[k] = m.res[k]; }
    g.inv_cell = m.inv_cell; g.lookup_radius = m.radius; g.scale = m.scale; g.lookup_size = m.lookup_size;
}
// surface term from the resident maps (no-op without photons or queries); film path when out_dev is null
static int surface_gather_impl(vrlb_context *ctx, float *out_dev) {
    const bool have_maps = ctx->surf[0].n || ctx->surf[1].n;
    if (ctx->n_squeries == 0 || (!have_maps && !out_dev)) return VRLB_OK;
    if (!out_dev && ctx->n_ranks > 1 && ctx->rank != 0) return VRLB_OK;       // VRL shards: the films are summed, this term is rank 0's
    SurfaceGatherParams p{};
    fill_grid(ctx->surf[VRLB_PHOTONMAP_GLOBAL], p.global); fill_grid(ctx->surf[VRLB_PHOTONMAP_CAUSTIC], p.caustic);
    p.heap_stride = std::max(1, std::max(ctx->surf[0].n ? ctx->surf[0].lookup_size : 0, ctx->surf[1].n ? ctx->surf[1].lookup_size : 0));
    // queries in batches so that the k-nearest scratch stays small (batch * k floats)
    const uint64_t batch = 1u << 18;
    DevBuf<float> heap; CU(ctx, heap.reserve(std::min<uint64_t>(batch, ctx->n_squeries) * (uint64_t) p.heap_stride));
    p.heap = heap.ptr;
    for (uint64_t first = 0; first < ctx->n_squeries; first += batch) {
        p.queries = ctx->squeries.ptr + first; p.n_queries = std::min<uint64_t>(batch, ctx->n_squeries - first);
        p.film_rgb = out_dev ? nullptr : ctx->film_rgb.ptr; p.query_out = out_dev ? out_dev + first * 3 : nullptr;
        CU(ctx, launch_surface_gather(p, ctx->stream));
        ctx->timing.total_launches += 1; ctx->launches += 1;
    }
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    heap.release();
    return VRLB_OK;
}
int vrlb_surface_gather_queries(vrlb_context *ctx, float *radiance_out) {
    if (!ctx || !radiance_out) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    DevBuf<float> tmp;
    CU(ctx, tmp.reserve(ctx->n_squeries * 3 + 1));
    CU(ctx, cudaMemsetAsync(tmp.ptr, 0, (ctx->n_squeries * 3 + 1) * sizeof(float), ctx->stream));
    int rc = surface_gather_impl(ctx, tmp.ptr);
    if (rc == VRLB_OK && ctx->n_squeries) {
        cudaError_t e = cudaMemcpy(radiance_out, tmp.ptr, ctx->n_squeries * 3 * sizeof(float), cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) rc = fail(ctx, VRLB_ERR_CUDA, "%s", cudaGetErrorString(e));
    }
    tmp.release();
    return rc;
}

/* Test/diagnostic entry: gather and also return Li per resident segment (before the weight), host buffer n_segs*3. */
int vrlb_gather_segments(vrlb_context *ctx, const vrlb_options *opt, float *li_out) {
    if (!ctx || !opt || !li_out) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    DevBuf<float> tmp;
    CU(ctx, tmp.reserve(ctx->n_segs * 3 + 1));
    CU(ctx, cudaMemsetAsync(tmp.ptr, 0, (ctx->n_segs * 3 + 1) * sizeof(float), ctx->stream));
    int rc = gather_impl(ctx, opt, tmp.ptr);
    if (rc == VRLB_OK) {
        cudaError_t e = cudaMemcpyAsync(li_out, tmp.ptr, ctx->n_segs * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) rc = fail(ctx, VRLB_ERR_CUDA, "%s", cudaGetErrorString(e));
        else ctx->timing.ms_gather = elapsed(ctx->ev[2], ctx->ev[3]);
    }
    tmp.release();
    return rc;
}

int vrlb_lightcut_build(vrlb_context *ctx, uint64_t base_seed, uint64_t seed_offset) {
    if (!ctx) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (ctx->lc) { lightcut_free(ctx->lc); ctx->lc = nullptr; }
    if (ctx->n_vrls == 0) return fail(ctx, VRLB_ERR_STATE, "no VRL set resident");
    cudaError_t e = lightcut_build(ctx->vrls.ptr, ctx->n_vrls, ctx->vrl_first, base_seed, seed_offset, ctx->stream, &ctx->lc, &ctx->launches);
    if (e != cudaSuccess) return fail_cuda(ctx, "lightcut_build", e);
    lightcut_build_stats(ctx->lc, &ctx->lc_last.n_nodes, &ctx->lc_last.build_searches, &ctx->lc_last.build_ms);
    return VRLB_OK;
}

int vrlb_lightcut_download(vrlb_context *ctx, float *f13, int32_t *i6, uint64_t cap, uint64_t *n_out) {
    if (!ctx || !n_out || (cap && (!f13 || !i6))) return VRLB_ERR_INVALID;
    if (!ctx->lc) return fail(ctx, VRLB_ERR_STATE, "no light tree resident");
    *n_out = lightcut_dump(ctx->lc, f13, i6, cap);
    return VRLB_OK;
}

int vrlb_lightcut_get_stats(const vrlb_context *ctx, vrlb_lightcut_stats *out) {
    if (!ctx || !out) return VRLB_ERR_INVALID;
    *out = ctx->lc_last;
    return VRLB_OK;
}

int vrlb_lightcut_gather_segments(vrlb_context *ctx, const vrlb_options *opt, float *li_out, float *stats4_out) {
    if (!ctx || !opt || !li_out) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (!ctx->have_scene) return fail(ctx, VRLB_ERR_STATE, "no scene resident");
    if (ctx->width == 0) return fail(ctx, VRLB_ERR_STATE, "no camera segments resident");
    if (opt->samples_per_query < 1) return fail(ctx, VRLB_ERR_INVALID, "samples_per_query must be >= 1");
    DevBuf<float> tmp;
    CU(ctx, tmp.reserve(ctx->n_segs * 3 + 1)); CU(ctx, ctx->lc_seg_stats.reserve(ctx->n_segs * 4 + 1));
    CU(ctx, cudaMemsetAsync(tmp.ptr, 0, (ctx->n_segs * 3 + 1) * sizeof(float), ctx->stream));
    CU(ctx, cudaMemsetAsync(ctx->lc_seg_stats.ptr, 0, (ctx->n_segs * 4 + 1) * sizeof(float), ctx->stream));
    int rc = lightcut_gather_impl(ctx, opt, tmp.ptr, ctx->lc_seg_stats.ptr);
    if (rc == VRLB_OK) {
        cudaError_t e = cudaMemcpyAsync(li_out, tmp.ptr, ctx->n_segs * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess && stats4_out) e = cudaMemcpyAsync(stats4_out, ctx->lc_seg_stats.ptr, ctx->n_segs * 4 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) rc = fail(ctx, VRLB_ERR_CUDA, "%s", cudaGetErrorString(e));
    }
    tmp.release();
    return rc;
}

int vrlb_film_device_ptr(vrlb_context *ctx, void **dev_ptr, uint64_t *n_floats) {
    if (!ctx || !dev_ptr) return VRLB_ERR_INVALID;
    if (ctx->width == 0) return fail(ctx, VRLB_ERR_STATE, "no film yet");
    *dev_ptr = ctx->film_rgb.ptr;
    if (n_floats) *n_floats = (uint64_t) ctx->width * ctx->height * 3;
    return VRLB_OK;
}

int vrlb_film_resolve(vrlb_context *ctx, void **dev_xyzaw) {
    if (!ctx) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (ctx->width == 0) return fail(ctx, VRLB_ERR_STATE, "no film yet");
    size_t np = (size_t) ctx->width * ctx->height;
    cudaEventRecord(ctx->ev[4], ctx->stream);
    CU(ctx, launch_film_xyzaw(ctx->film_rgb.ptr, ctx->film_aw.ptr, ctx->film_xyzaw.ptr, np, ctx->stream));
    cudaEventRecord(ctx->ev[5], ctx->stream);
    ctx->timing.total_launches += 1; ctx->launches += 1;
    if (dev_xyzaw) *dev_xyzaw = ctx->film_xyzaw.ptr;
    return VRLB_OK;
}

int vrlb_film_download_xyzaw(vrlb_context *ctx, float *xyzaw) {
    if (!ctx || !xyzaw) return VRLB_ERR_INVALID;
    int rc = vrlb_film_resolve(ctx, nullptr);
    if (rc) return rc;
    size_t np = (size_t) ctx->width * ctx->height;
    rc = ensure_pinned(ctx, np * 5 * sizeof(float));
    if (rc) return rc;
    cudaEventRecord(ctx->ev[6], ctx->stream);
    CU(ctx, cudaMemcpyAsync(ctx->pinned, ctx->film_xyzaw.ptr, np * 5 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    cudaEventRecord(ctx->ev[7], ctx->stream);
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    std::memcpy(xyzaw, ctx->pinned, np * 5 * sizeof(float));
    ctx->timing.ms_film = elapsed(ctx->ev[4], ctx->ev[5]);
    ctx->timing.ms_download = elapsed(ctx->ev[6], ctx->ev[7]);
    if (ctx->timing.gather_launches) ctx->timing.ms_gather = elapsed(ctx->ev[2], ctx->ev[3]);
    return VRLB_OK;
}

int vrlb_film_download_rgb(vrlb_context *ctx, float *rgb) {
    if (!ctx || !rgb) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (ctx->width == 0) return fail(ctx, VRLB_ERR_STATE, "no film yet");
    size_t np = (size_t) ctx->width * ctx->height;
    CU(ctx, cudaMemcpyAsync(rgb, ctx->film_rgb.ptr, np * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->timing.ms_gather = elapsed(ctx->ev[2], ctx->ev[3]);
    return VRLB_OK;
}

// ---- whole frame (Integrator::render, integrator.cpp:51-183): upload -> segments -> gather -> film ----
int vrlb_render(vrlb_context *ctx, const vrlb_camera *cam, const vrlb_options *opt, const vrlb_vrlset_desc *vrls, float *xyzaw_out) {
    if (!ctx || !cam || !opt) return VRLB_ERR_INVALID;
    if (!xyzaw_out && ctx->rank == 0) return fail(ctx, VRLB_ERR_INVALID, "xyzaw_out is NULL (only ranks > 0 may skip the film download)");
    ctx->timing = vrlb_timing{};
    int rc;
    if (vrls && (rc = vrlb_vrls_upload(ctx, vrls)) != VRLB_OK) return rc;
    if ((rc = vrlb_segments_build(ctx, cam, opt)) != VRLB_OK) return rc;
    /* an empty map is fine: VRLMap::query returns 0 (vrl_map.h:134-137) */
    if ((rc = vrlb_gather(ctx, opt)) != VRLB_OK) return rc;
    if (ctx->n_ranks > 1 && (rc = vrlb_film_allreduce(ctx)) != VRLB_OK) return rc;
    if (xyzaw_out) return vrlb_film_download_xyzaw(ctx, xyzaw_out);
    return vrlb_synchronize(ctx);
}

// ---- multi-GPU film reduce: NCCL bound at run time ---------------------------------------------------
struct Id128 { char bytes[VRLB_COMM_ID_BYTES]; };   // ncclUniqueId is passed by value
namespace {
struct NcclApi {
    void *lib = nullptr;
    int (*GetUniqueId)(void *) = nullptr;
    int (*CommInitRank)(void **, int, Id128, int) = nullptr;
    int (*CommDestroy)(void *) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, void *, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
};
NcclApi g_nccl;
const char *load_nccl() {
    if (g_nccl.lib) return nullptr;
    // the soname first: inside a process that already loaded NCCL (e.g. torch) this binds to that copy
    const char *names[] = { "libnccl.so.2", "libnccl.so" };
    void *h = nullptr;
    for (const char *n : names) if ((h = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
    if (!h) return "libnccl.so.2 not found (dlopen)";
    g_nccl.GetUniqueId   = (decltype(g_nccl.GetUniqueId)) dlsym(h, "ncclGetUniqueId");
    g_nccl.CommInitRank  = (decltype(g_nccl.CommInitRank)) dlsym(h, "ncclCommInitRank");
    g_nccl.CommDestroy   = (decltype(g_nccl.CommDestroy)) dlsym(h, "ncclCommDestroy");
    g_nccl.AllReduce     = (decltype(g_nccl.AllReduce)) dlsym(h, "ncclAllReduce");
    g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString)) dlsym(h, "ncclGetErrorString");
    if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.CommDestroy || !g_nccl.AllReduce) return "NCCL symbols missing";
    g_nccl.lib = h; g_nccl_destroy = g_nccl.CommDestroy;
    return nullptr;
}
const char *nccl_err(int rc) { return g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "nccl error"; }
}  // namespace

int vrlb_comm_unique_id(void *id) {
    if (!id) return VRLB_ERR_INVALID;
    if (const char *e = load_nccl()) return fail(nullptr, VRLB_ERR_STATE, "%s", e);
    int rc = g_nccl.GetUniqueId(id);
    if (rc != 0) return fail(nullptr, VRLB_ERR_CUDA, "ncclGetUniqueId: %s", nccl_err(rc));
    return VRLB_OK;
}

int vrlb_comm_init(vrlb_context *ctx, const void *id, int rank, int n_ranks) {
    if (!ctx || !id || n_ranks < 1 || rank < 0 || rank >= n_ranks) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (ctx->nccl_comm) return fail(ctx, VRLB_ERR_STATE, "communicator already initialised");
    if (const char *e = load_nccl()) return fail(ctx, VRLB_ERR_STATE, "%s", e);
    Id128 uid; std::memcpy(uid.bytes, id, sizeof uid.bytes);
    int rc = g_nccl.CommInitRank(&ctx->nccl_comm, n_ranks, uid, rank);
    if (rc != 0) { ctx->nccl_comm = nullptr; return fail(ctx, VRLB_ERR_CUDA, "ncclCommInitRank: %s", nccl_err(rc)); }
    ctx->rank = rank; ctx->n_ranks = n_ranks;
    return VRLB_OK;
}

int vrlb_comm_destroy(vrlb_context *ctx) {
    if (!ctx) return VRLB_ERR_INVALID;
    if (ctx->nccl_comm) { cudaSetDevice(ctx->device); cudaStreamSynchronize(ctx->stream); g_nccl.CommDestroy(ctx->nccl_comm); }
    ctx->nccl_comm = nullptr; ctx->rank = 0; ctx->n_ranks = 1;
    return VRLB_OK;
}

int vrlb_film_allreduce(vrlb_context *ctx) {
    if (!ctx) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (ctx->width == 0) return fail(ctx, VRLB_ERR_STATE, "no film yet");
    if (ctx->n_ranks == 1) return VRLB_OK;
    if (!ctx->nccl_comm) return fail(ctx, VRLB_ERR_STATE, "vrlb_comm_init must precede vrlb_film_allreduce");
    size_t n = (size_t) ctx->width * ctx->height * 3;
    int rc = g_nccl.AllReduce(ctx->film_rgb.ptr, ctx->film_rgb.ptr, n, /* ncclFloat32 */ 7, /* ncclSum */ 0, ctx->nccl_comm, ctx->stream);
    if (rc != 0) return fail(ctx, VRLB_ERR_CUDA, "ncclAllReduce: %s", nccl_err(rc));   // NCCL errors abort the render
    return VRLB_OK;
}

int vrlb_measure_fp32_peak(vrlb_context *ctx, float *tflops) {
    if (!ctx || !tflops) return VRLB_ERR_INVALID;
    cudaSetDevice(ctx->device);
    cudaDeviceProp prop; CU(ctx, cudaGetDeviceProperties(&prop, ctx->device));
    DevBuf<float> scratch; CU(ctx, scratch.reserve(4));
    const int blocks = prop.multiProcessorCount * 8, iters = 4096;
    float best = 0.f;
    for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(ctx->ev[0], ctx->stream);
        CU(ctx, launch_fp32_peak(scratch.ptr, blocks, iters, ctx->stream));
        cudaEventRecord(ctx->ev[1], ctx->stream);
        CU(ctx, cudaStreamSynchronize(ctx->stream));
        ctx->launches += 1;
        double flops = (double) blocks * 256.0 * iters * 64.0 * 2.0;
        float t = (float) (flops / (elapsed(ctx->ev[0], ctx->ev[1]) * 1e-3) * 1e-12);
        if (rep > 0 && t > best) best = t;
    }
    scratch.release();
    *tflops = best;
    return VRLB_OK;
}

}  // extern "C"
