// noalloc.cu -- vrlb_vrls_dice_noalloc: VRL dicing (VRLMap::dicingVRL -> VRL::dice, vrl_map.h:104-128,
// vrl_struct.h:88-128; the arithmetic of vrl_post.cu) for callers that may not allocate: CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
// vrlb_vrls_dice keeps its per-VRL piece counts and write offsets in cudaMalloc'ed scratch and scans them on the host.
// Here that scratch is a file-scope __device__ array of a fixed, modest size (VRLB_NOALLOC_MAX_VRLS input VRLs), which
// every GPU that loads the library holds until the process exits; the scan runs on the device.  The entry point calls
// no cudaMalloc / cudaFree / allocating library routine.  Every call shares the array: nothing in it has to start at
// zero (each call overwrites the part it reads), but two calls that overlap on the GPU -- different streams, or
// different host threads -- race on it.  Serialise them.
#include <cstdio>
#include <mutex>

#include "vrlb_device.cuh"
#include "vrlb_kernels.h"

namespace vrlb {

struct DiceScratch {
    uint64_t offsets[VRLB_NOALLOC_MAX_VRLS + 1];    // exclusive prefix of counts; offsets[n] = diced size
    uint32_t counts[VRLB_NOALLOC_MAX_VRLS];
    float    dice_len;                              // mean length / avg_number_dice, computed on the device
    uint32_t status;                                // 0 ok, 1 non-positive mean length, 2 output capacity too small
};
// In host code this name is only a stand-in for the symbol; the device address comes from cudaGetSymbolAddress.
static __device__ __align__(256) unsigned char g_dice_scratch[sizeof(DiceScratch)];

// dice_count / the piece arithmetic restate vrl_post.cu (static there); see the comments in that file.
__device__ __forceinline__ uint32_t na_dice_count(float length, float dice_len) {
    const double L = (double) dice_len, len = (double) length;
    if (!(L > 0.0) || !(len >= 0.0)) return 0u;
    double kf = floor(len / L);
    if (kf > 4.0e6) kf = 4.0e6;
    uint32_t k = (uint32_t) kf;
    while (__dmul_rn((double) (k + 1), L) <= len) ++k;
    while (k > 0 && __dmul_rn((double) k, L) > len) --k;
    return k + 1;
}

// the reference's serial double sum over the map (vrl_map.h:106-110), one thread: 65 536 additions at most
__global__ void na_mean_kernel(const float *len, uint64_t n, uint32_t avg_number_dice, DiceScratch *s) {
    double avg = 0.0;
    for (uint64_t i = 0; i < n; ++i) avg = __dadd_rn(avg, (double) len[i]);
    avg = __ddiv_rn(avg, (double) n);
    const float d = (float) __ddiv_rn(avg, (double) avg_number_dice);
    s->dice_len = d;
    s->status = d > 0.f ? 0u : 1u;
}
__global__ void na_count_kernel(const float *len, uint64_t n, DiceScratch *s) {
    uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) s->counts[j] = s->status ? 0u : na_dice_count(len[j], s->dice_len);
}
// exclusive scan of <= 65 536 counts by one CTA of 1024 threads: each thread owns a contiguous run
__global__ void __launch_bounds__(1024) na_scan_kernel(uint64_t n, uint64_t cap_out, uint64_t *n_out, DiceScratch *s) {
    __shared__ uint64_t part[1024];
    const uint32_t t = threadIdx.x;
    const uint64_t per = (n + 1023) / 1024, a = min((uint64_t) t * per, n), b = min(a + per, n);
    uint64_t sum = 0;
    for (uint64_t i = a; i < b; ++i) sum += s->counts[i];
    part[t] = sum;
    __syncthreads();
    if (t == 0) {
        uint64_t run = 0;
        for (int i = 0; i < 1024; ++i) { uint64_t v = part[i]; part[i] = run; run += v; }
        s->offsets[n] = run;
        if (s->status == 0 && run > cap_out) s->status = 2u;
        n_out[0] = run;
        n_out[1] = s->status;
    }
    __syncthreads();
    uint64_t run = part[t];
    for (uint64_t i = a; i < b; ++i) { s->offsets[i] = run; run += s->counts[i]; }
}
__global__ void na_write_kernel(const float *in, uint64_t n, float max_density, float *out, uint64_t stride, const DiceScratch *s) {
    uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n || s->status) return;
    float v[10];
#pragma unroll
    for (int c = 0; c < 10; ++c) v[c] = in[(uint64_t) c * n + j];
    const uint32_t cnt = s->counts[j];
    const uint64_t dst = s->offsets[j];
    const double L = (double) s->dice_len;
    for (uint32_t k = 0; k < cnt; ++k) {
        const double cur = __dmul_rn((double) k, L);
        const float next = (float) fmin(__dsub_rn((double) v[6], cur), L);
        float tr = 1.0f;
        if (cur != 0.0) {
            const float neg = __fsub_rn(0.0f, (float) cur);
            tr = max_density != 0.f ? expf(__fmul_rn(max_density, neg)) : 1.0f;    // medium.h:112-117
        }
        const uint64_t o = dst + k;
        out[0 * stride + o] = (float) __dadd_rn((double) v[0], __dmul_rn((double) v[3], cur));
        out[1 * stride + o] = (float) __dadd_rn((double) v[1], __dmul_rn((double) v[4], cur));
        out[2 * stride + o] = (float) __dadd_rn((double) v[2], __dmul_rn((double) v[5], cur));
        out[3 * stride + o] = v[3]; out[4 * stride + o] = v[4]; out[5 * stride + o] = v[5];
        out[6 * stride + o] = next;
        out[7 * stride + o] = __fmul_rn(v[7], tr); out[8 * stride + o] = __fmul_rn(v[8], tr); out[9 * stride + o] = __fmul_rn(v[9], tr);
    }
}

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_na_mutex;
DiceScratch *g_na_ptr[kMaxDevices] = {};     // device address of g_dice_scratch, per device

// once per device, with that device current
cudaError_t scratch_for_current_device(int device, DiceScratch **out) {
    std::lock_guard<std::mutex> lock(g_na_mutex);
    if (!g_na_ptr[device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_dice_scratch);
        if (e != cudaSuccess) return e;
        g_na_ptr[device] = static_cast<DiceScratch *>(p);
    }
    *out = g_na_ptr[device];
    return cudaSuccess;
}
}  // namespace

}  // namespace vrlb

using namespace vrlb;

extern "C" int vrlb_vrls_dice_noalloc(int device, void *cuda_stream, const float *d_soa10, uint64_t n, float max_density,
                                      uint32_t avg_number_dice, float *d_soa10_out, uint64_t cap_out, uint64_t *d_n_out) {
    char msg[256];
    if (device < 0 || device >= kMaxDevices || !d_soa10 || !d_soa10_out || !d_n_out || avg_number_dice == 0 || n == 0 || cap_out == 0) {
        set_global_error("vrlb_vrls_dice_noalloc: bad argument (device, NULL buffer, n = 0, cap_out = 0 or avg_number_dice = 0)");
        return VRLB_ERR_INVALID;
    }
    if (n > VRLB_NOALLOC_MAX_VRLS) {
        snprintf(msg, sizeof msg, "vrlb_vrls_dice_noalloc: %llu VRLs exceed the static scratch (VRLB_NOALLOC_MAX_VRLS = %u); "
                                  "use vrlb_vrls_dice, which allocates", (unsigned long long) n, (unsigned) VRLB_NOALLOC_MAX_VRLS);
        set_global_error(msg);
        return VRLB_ERR_INVALID;
    }
    cudaError_t e = cudaSetDevice(device);
    DiceScratch *s = nullptr;
    if (e == cudaSuccess) e = scratch_for_current_device(device, &s);
    if (e != cudaSuccess) {
        snprintf(msg, sizeof msg, "vrlb_vrls_dice_noalloc: %s", cudaGetErrorString(e));
        set_global_error(msg);
        return VRLB_ERR_CUDA;
    }
    cudaStream_t st = (cudaStream_t) cuda_stream;
    const uint32_t grid = (uint32_t) ((n + 127) / 128);
    const float *len = d_soa10 + 6 * n;
    na_mean_kernel<<<1, 1, 0, st>>>(len, n, avg_number_dice, s);
    na_count_kernel<<<grid, 128, 0, st>>>(len, n, s);
    na_scan_kernel<<<1, 1024, 0, st>>>(n, cap_out, d_n_out, s);
    na_write_kernel<<<grid, 128, 0, st>>>(d_soa10, n, max_density, d_soa10_out, cap_out, s);
    e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(msg, sizeof msg, "vrlb_vrls_dice_noalloc: launch failed: %s", cudaGetErrorString(e));
        set_global_error(msg);
        return VRLB_ERR_CUDA;
    }
    return VRLB_OK;
}
