// sba_noalloc.cu -- batched multi-view DLT triangulation for callers that may not allocate device memory (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Same kernels as sba_triangulate (K0 pose cache + K5 k_triangulate); what differs is where their scratch
// lives: sba_triangulate keeps the camera cache in the handle's arena (cudaMalloc at finalize) and stages its arguments in
// cudaMalloc'ed buffers, here every intermediate sits in ONE file-scope device array of fixed size:
//     camera cache (24 doubles per camera), Q block (12 per camera), DLT record (8 per camera), the sanitised camera-model index per camera,
//     the clamped view offsets, the sanitised camera index of every view, one "bad input" flag per point.
// Every GPU that loads the library holds the array (SBA_NOALLOC_SCRATCH_BYTES = 6.75 MiB) until the process exits.
//
// The array's name is a host-side stand-in in host code: the device address comes from cudaGetSymbolAddress, once per device,
// with that device current (noalloc_scratch below).  All calls on a device share the array and it keeps the previous call's
// contents: every region a call reads is rewritten by that call first (nothing relies on zero-initialisation; the bad-input
// flags are stored by every point's thread, not accumulated), and two calls that overlap on the GPU (different streams) race.
#include <mutex>
#include <string>

#include "sba_b200.h"
#include "sba_kernel_util.cuh"

namespace sba {

namespace {

constexpr size_t align256(size_t b) { return (b + 255) / 256 * 256; }
constexpr size_t NA_OFF_CACHE = 0;
constexpr size_t NA_OFF_Q = NA_OFF_CACHE + align256(sizeof(double) * CAM_CACHE_STRIDE * SBA_NOALLOC_MAX_CAMS);
constexpr size_t NA_OFF_REC = NA_OFF_Q + align256(sizeof(double) * 12 * SBA_NOALLOC_MAX_CAMS);
constexpr size_t NA_OFF_MODEL = NA_OFF_REC + align256(sizeof(double) * 8 * SBA_NOALLOC_MAX_CAMS);
constexpr size_t NA_OFF_VOFF = NA_OFF_MODEL + align256(sizeof(int32_t) * SBA_NOALLOC_MAX_CAMS);
constexpr size_t NA_OFF_VCAM = NA_OFF_VOFF + align256(sizeof(int32_t) * ((size_t)SBA_NOALLOC_MAX_POINTS + 1));
constexpr size_t NA_OFF_BAD = NA_OFF_VCAM + align256(sizeof(int32_t) * (size_t)SBA_NOALLOC_MAX_VIEWS);
constexpr size_t NA_BYTES = NA_OFF_BAD + align256((size_t)SBA_NOALLOC_MAX_POINTS);
static_assert(NA_BYTES <= SBA_NOALLOC_SCRATCH_BYTES, "include/sba_b200.h states a smaller scratch than the layout needs");

static __device__ __align__(256) unsigned char g_noalloc_scratch[SBA_NOALLOC_SCRATCH_BYTES];

// device address of the scratch array on the CURRENT device (looked up once per device)
std::mutex g_na_mutex;
unsigned char* g_na_ptr[64] = {};
int noalloc_scratch(unsigned char** out, std::string& err) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { err = std::string("sba_triangulate_noalloc: cudaGetDevice: ") + cudaGetErrorString(e); return e == cudaErrorNoDevice ? SBA_ERR_NO_DEVICE : SBA_ERR_CUDA; }
    if (dev < 0 || dev >= 64) { err = "sba_triangulate_noalloc: device ordinal beyond 63"; return SBA_ERR_INVALID; }
    std::lock_guard<std::mutex> lock(g_na_mutex);
    if (!g_na_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            err = std::string("sba_triangulate_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e);
            return SBA_ERR_CUDA;
        }
        g_na_ptr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_na_ptr[dev];
    return SBA_OK;
}

// camera-model index per camera, sanitised (NULL input: every camera uses model 0)
__global__ void k_na_models(int n_cams, int n_models, const int32_t* __restrict__ cam_model, int32_t* model_s) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_cams) return;
    const int m = cam_model ? __ldg(cam_model + c) : 0;
    model_s[c] = (m >= 0 && m < n_models) ? m : 0;
}
// offsets clamped into [0, n_views] (so that K5 reads in bounds whatever the caller passed); thread per offset
__global__ void k_na_offsets(int n_points, int n_views, const int32_t* __restrict__ view_off, int32_t* off_s) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n_points) return;
    off_s[i] = min(max(__ldg(view_off + i), 0), n_views);
}
// thread per point: camera indices of its views sanitised, bad = some offset, camera or camera-model index was out of range
__global__ void k_na_views(int n_points, int n_views, int n_cams, int n_models, const int32_t* __restrict__ view_off,
                           const int32_t* __restrict__ view_cam, const int32_t* __restrict__ cam_model, int32_t* cam_s, uint8_t* bad) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_points) return;
    const int r0 = __ldg(view_off + i), r1 = __ldg(view_off + i + 1);
    bool b = r0 < 0 || r1 < r0 || r1 > n_views;
    const int k0 = min(max(r0, 0), n_views), k1 = min(max(r1, 0), n_views);
    for (int k = k0; k < k1; k++) {
        const int c = __ldg(view_cam + k);
        const bool in = c >= 0 && c < n_cams;
        if (in && cam_model) { const int m = __ldg(cam_model + c); b = b || m < 0 || m >= n_models; }
        b = b || !in;
        cam_s[k] = in ? c : 0;
    }
    bad[i] = b ? 1 : 0;
}
__global__ void k_na_fixup(int n_points, const uint8_t* __restrict__ bad, double* xyz_out, uint8_t* ok_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_points || !bad[i]) return;
    xyz_out[3 * (size_t)i] = 0.0; xyz_out[3 * (size_t)i + 1] = 0.0; xyz_out[3 * (size_t)i + 2] = 0.0;
    if (ok_out) ok_out[i] = 0;
}

}  // namespace

int triangulate_noalloc(int32_t n_cams, const double* raxis, const double* t, const int32_t* cam_model, int32_t n_models,
                        const double* intr10, int32_t n_points, int32_t n_views, const int32_t* view_off, const int32_t* view_cam,
                        const double* view_uv, double* xyz_out, uint8_t* ok_out, void* stream, std::string& err) {
    if (n_cams < 0 || n_models < 0 || n_points < 0 || n_views < 0) { err = "sba_triangulate_noalloc: negative size"; return SBA_ERR_INVALID; }
    if (n_cams > SBA_NOALLOC_MAX_CAMS || n_points > SBA_NOALLOC_MAX_POINTS || n_views > SBA_NOALLOC_MAX_VIEWS) {
        err = "sba_triangulate_noalloc: " + std::to_string(n_cams) + " cameras / " + std::to_string(n_points) + " points / " +
              std::to_string(n_views) + " views exceed the static scratch (at most " + std::to_string(SBA_NOALLOC_MAX_CAMS) + " / " +
              std::to_string(SBA_NOALLOC_MAX_POINTS) + " / " + std::to_string(SBA_NOALLOC_MAX_VIEWS) + "): split the batch or use sba_triangulate";
        return SBA_ERR_INVALID;
    }
    unsigned char* base = nullptr;
    const int rc = noalloc_scratch(&base, err);   // also with n_points == 0: the once-per-device set-up (do it outside a capture)
    if (rc) return rc;
    if (n_points == 0) return SBA_OK;
    if (!view_off || !xyz_out || (n_views > 0 && (!view_cam || !view_uv))) { err = "sba_triangulate_noalloc: null view / output arrays"; return SBA_ERR_INVALID; }
    if (n_cams > 0 && (n_models == 0 || !raxis || !t || !intr10)) { err = "sba_triangulate_noalloc: cameras without poses / intrinsics models"; return SBA_ERR_INVALID; }
    if (n_views > 0 && n_cams == 0) { err = "sba_triangulate_noalloc: views without cameras"; return SBA_ERR_INVALID; }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int32_t* model_s = reinterpret_cast<int32_t*>(base + NA_OFF_MODEL);
    int32_t* off_s = reinterpret_cast<int32_t*>(base + NA_OFF_VOFF);
    int32_t* cam_s = reinterpret_cast<int32_t*>(base + NA_OFF_VCAM);
    uint8_t* bad = base + NA_OFF_BAD;
    DevProblem P{};   // only what K0 and K5 read: no offsets, no mountings
    P.Nc = n_cams;
    P.cam_r = const_cast<double*>(raxis); P.cam_t = const_cast<double*>(t);
    P.intr = const_cast<double*>(intr10);
    P.cam_model = model_s;
    P.cam_cache = reinterpret_cast<double*>(base + NA_OFF_CACHE);
    P.cam_Q = reinterpret_cast<double*>(base + NA_OFF_Q);
    P.dlt_rec = reinterpret_cast<double*>(base + NA_OFF_REC);
    if (n_cams) {
        k_na_models<<<cdiv(n_cams, 256), 256, 0, st>>>(n_cams, n_models, cam_model, model_s);
        launch_pose_cache(P, st);
    }
    k_na_offsets<<<cdiv(n_points + 1, 256), 256, 0, st>>>(n_points, n_views, view_off, off_s);
    k_na_views<<<cdiv(n_points, 256), 256, 0, st>>>(n_points, n_views, n_cams, n_models, view_off, view_cam, cam_model, cam_s, bad);
    launch_triangulate(P, n_points, off_s, cam_s, view_uv, xyz_out, ok_out, st);
    k_na_fixup<<<cdiv(n_points, 256), 256, 0, st>>>(n_points, bad, xyz_out, ok_out);
    const cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) { err = std::string("sba_triangulate_noalloc: launch: ") + cudaGetErrorString(e); return SBA_ERR_CUDA; }
    return SBA_OK;
}

}  // namespace sba
