// sba_kernel_util.cuh -- warp/block reduction helpers shared by the kernel files.
// Every reduction is a fixed-shape tree: results are bitwise reproducible run to run.
#pragma once
#include "sba_device.h"
#include "sba_math.cuh"

namespace sba {

constexpr int SBA_INTR_PAD = 10;

#define FULL 0xffffffffu

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(FULL, v, o));
    return v;
}

// deterministic block-wide sum (blockDim.x multiple of 32, <= 1024); result valid in every thread
__device__ __forceinline__ double block_sum(double v, double* sh /*>=33*/) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    if (w == 0) {
        double t = lane < nw ? sh[lane] : 0.0;
        t = warp_sum(t);
        if (lane == 0) sh[32] = t;
    }
    __syncthreads();
    return sh[32];
}
__device__ __forceinline__ double block_max(double v, double* sh) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    if (w == 0) {
        double t = lane < nw ? sh[lane] : 0.0;
        t = warp_max(t);
        if (lane == 0) sh[32] = t;
    }
    __syncthreads();
    return sh[32];
}

__device__ __forceinline__ double2 ldg2(const double* p) { return __ldg(reinterpret_cast<const double2*>(p)); }
// streaming (evict-first) variants for the big planes that are read/written once per pass, so that they do not
// push the gathered per-point arrays (y_p, points, Hinv) out of the 126 MB L2
__device__ __forceinline__ double2 ldcs2(const double* p) { return __ldcs(reinterpret_cast<const double2*>(p)); }
__device__ __forceinline__ void stcs(double* p, double v) { __stcs(p, v); }
__device__ __forceinline__ void st2(double* p, double a, double b) { *reinterpret_cast<double2*>(p) = make_double2(a, b); }

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }

// Sum of SIX values over the warp with 8 double shuffles instead of 30: at every step the two partner lanes split the
// values between them (6 -> 3 -> 2|1 -> 1) and only then run the plain butterfly.  Returns the total of value *j in
// the lanes with (lane & 3) == 0 and *j < 6 (lanes 0, 4, 8, 16, 20, 24 -> values 0, 1, 2, 3, 4, 5).  Fixed shape ->
// deterministic.  Shuffles share the MIO pipe with the L1 gathers, which is what bounds the camera pass.
__device__ __forceinline__ double warp_sum6(const double (&v)[6], int lane, int* j) {
    const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
    const double r0 = (b4 ? v[3] : v[0]) + __shfl_xor_sync(FULL, b4 ? v[0] : v[3], 16);
    const double r1 = (b4 ? v[4] : v[1]) + __shfl_xor_sync(FULL, b4 ? v[1] : v[4], 16);
    const double r2 = (b4 ? v[5] : v[2]) + __shfl_xor_sync(FULL, b4 ? v[2] : v[5], 16);
    const double gA = __shfl_xor_sync(FULL, b3 ? r0 : r2, 8);   // b3 = 0 receives the partner's r0, b3 = 1 its r2
    const double gB = __shfl_xor_sync(FULL, r1, 8);
    const double p0 = (b3 ? r2 : r0) + gA;
    const double p1 = b3 ? 0.0 : r1 + gB;
    double q = (b2 ? p1 : p0) + __shfl_xor_sync(FULL, b2 ? p0 : p1, 4);
    q += __shfl_xor_sync(FULL, q, 2);
    q += __shfl_xor_sync(FULL, q, 1);
    *j = (b3 && b2) ? 6 : (b4 ? 3 : 0) + (b3 ? 2 : (b2 ? 1 : 0));
    return q;
}

// sum of component a over the camera-pass tile partials w0..w1-1 of one camera: four interleaved chains (shorter dependent
// chain than a serial sum), combined in a fixed order -> deterministic
__device__ __forceinline__ double sum_partials(const double* __restrict__ wp, int w0, int w1, int a) {
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    int w = w0;
    for (; w + 3 < w1; w += 4) {
        a0 += wp[6 * (size_t)w + a]; a1 += wp[6 * (size_t)(w + 1) + a];
        a2 += wp[6 * (size_t)(w + 2) + a]; a3 += wp[6 * (size_t)(w + 3) + a];
    }
    for (; w < w1; w++) a0 += wp[6 * (size_t)w + a];
    return (a0 + a1) + (a2 + a3);
}

// 256-bit read-only load (sm_100: LDG.E.256): one instruction -- one L1 wavefront per distinct line -- for a whole
// 32-byte record (a point, a y_p entry, half a camera operand).  The gather passes are bound by L1 wavefronts +
// shuffles (both go through the MIO pipe), not by bytes, so halving the instruction count of a gather matters.
struct d4 { double x, y, z, w; };
__device__ __forceinline__ d4 ld256(const double* p) {
    d4 v;
    asm("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w) : "l"(p));
    return v;
}

__device__ __forceinline__ void st256(double* p, double a, double b, double c, double d) {
    asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" :: "l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
}

// per-camera operand of the point pass (sba_kernels_points.cu) for the input vector x6 = (x_r, x_t) of camera c:
//   (a_c = R Jr x_r, g_c = x_t + (t - c0) x a_c, pad 2)   -- 64 B, two 256-bit loads --
// so that  C x_r - A x_t = A ((X - t) x a_c - x_t) = A ((X - c0) x a_c - g_c).  c0 = P.center (fixed at finalize, the
// mean camera position) keeps |X - c0|, |t - c0| of the order of the scene extent whatever the world origin is.
constexpr int CAM_VEC_STRIDE = 8;
__device__ __forceinline__ void cam_vec_store(const DevProblem& P, int c, const double* x6) {
    const double* cc = P.cam_cache + (size_t)CAM_CACHE_STRIDE * c;
    const double* Jr = cc + 12;
    const double w0 = Jr[0] * x6[0] + Jr[1] * x6[1] + Jr[2] * x6[2];
    const double w1 = Jr[3] * x6[0] + Jr[4] * x6[1] + Jr[5] * x6[2];
    const double w2 = Jr[6] * x6[0] + Jr[7] * x6[1] + Jr[8] * x6[2];
    // R = (Rt)^T
    const double a0 = cc[0] * w0 + cc[3] * w1 + cc[6] * w2;
    const double a1 = cc[1] * w0 + cc[4] * w1 + cc[7] * w2;
    const double a2 = cc[2] * w0 + cc[5] * w1 + cc[8] * w2;
    const double t0 = cc[9] - P.center[0], t1 = cc[10] - P.center[1], t2 = cc[11] - P.center[2];
    double* o = P.cam_vec + (size_t)CAM_VEC_STRIDE * c;
    st2(o, a0, a1);
    st2(o + 2, a2, x6[3] + (t1 * a2 - t2 * a1));
    st2(o + 4, x6[4] + (t2 * a0 - t0 * a2), x6[5] + (t0 * a1 - t1 * a0));
    st2(o + 6, 0.0, 0.0);
}

// ---- explicit reduced camera system (sba_kernels_explicit.cu): tile geometry and the fold of the mat-vec partials
constexpr int EX_TB = 64;            // cameras per tile side
constexpr int EX_TD = 6 * EX_TB;     // 384 scalars per tile side
constexpr int EX_CHUNK = 64;         // rows per mat-vec CTA (128 measured slower: 0.093 vs 0.085 ms) (EXH_NCHUNK in sba_solver_explicit.inc = EX_TD / EX_CHUNK)
constexpr int EX_MVW = EX_CHUNK / 16; // warps per mat-vec CTA, 16 rows each
constexpr int EX_NCHUNK = EX_TD / EX_CHUNK;

// y = S x from the partials, element i of camera tile K: the static list ex_fold_off[ex_fold_ptr[K] .. ex_fold_ptr[K+1])
// holds the offsets (relative to ex_row) of the partial vectors that contribute to tile K -- row partials of the tiles
// (K, J >= K) in J order, then the column partials of the tiles (I <= K, K) in (I, chunk) order.  Block-cooperative:
// all threads of the block belong to ONE camera tile (EX_TD % blockDim.x == 0); the offsets are staged in shared memory
// so the ~200 loads of a thread are independent and pipeline.  Four interleaved chains combined in a fixed order ->
// deterministic.  Every thread of the block must call it (barriers inside); `valid` = the thread owns element i.
__device__ __forceinline__ double explicit_fold(const DevProblem& P, int i, bool valid, int* s_off /* blockDim.x ints */) {
    const int K = (blockIdx.x * blockDim.x) / EX_TD, e = i - K * EX_TD;
    const int m0 = __ldg(P.ex_fold_ptr + K), m1 = __ldg(P.ex_fold_ptr + K + 1);
    const double* base = P.ex_row + e;
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    for (int mb = m0; mb < m1; mb += blockDim.x) {
        __syncthreads();
        if (mb + (int)threadIdx.x < m1) s_off[threadIdx.x] = __ldg(P.ex_fold_off + mb + threadIdx.x);
        __syncthreads();
        const int n = min((int)blockDim.x, m1 - mb);
        if (valid) {
            int m = 0;
            // 16 independent loads in flight per thread (the fold is latency-bound: ~100-200 L2 reads per element on few
            // SMs); entry m still goes to chain m % 4, so the sum is bitwise the one of the plain 4-chain loop below
            for (; m + 15 < n; m += 16) {
                double v[16];
#pragma unroll
                for (int k = 0; k < 16; k++) v[k] = __ldcg(base + s_off[m + k]);
#pragma unroll
                for (int k = 0; k < 16; k += 4) { a0 += v[k]; a1 += v[k + 1]; a2 += v[k + 2]; a3 += v[k + 3]; }
            }
            for (; m + 3 < n; m += 4) {
                a0 += base[s_off[m]]; a1 += base[s_off[m + 1]]; a2 += base[s_off[m + 2]]; a3 += base[s_off[m + 3]];
            }
            for (; m < n; m++) a0 += base[s_off[m]];
        }
    }
    return (a0 + a1) + (a2 + a3);
}

}  // namespace sba
