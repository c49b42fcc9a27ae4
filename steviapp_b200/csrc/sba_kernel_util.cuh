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

// C = d res / d rAxis is NOT stored in camera order: with d = X - t = R Pbar and Q = (R Jr)^T (cam_Q, written by K0)
//   C = W Pi [Pbar]x Jr = A [d]x Q^T,   row k:  C_k = (A_k x d) Q^T
// for ANY stiffness W, so the passes that need C rebuild it from the A planes, the Pbar planes (Pc) and two per-camera
// 3x3 constants: 24 B/obs read instead of 48, and K1 writes 48 B/obs less.
struct CamRQ { double Rt[9], Q[9], dt[3]; };   // Rt of the SENSOR pose, Q of the body pose, dt = ts - t (zero without offsets / mountings)
__device__ __forceinline__ CamRQ load_cam_rq(const DevProblem& P, int cam) {
    CamRQ c;
    const double* cc = P.cam_cache + (size_t)CAM_CACHE_STRIDE * cam;
    const d4 a0 = ld256(cc), a1 = ld256(cc + 4), a2 = ld256(cc + 8);
    c.Rt[0] = a0.x; c.Rt[1] = a0.y; c.Rt[2] = a0.z; c.Rt[3] = a0.w; c.Rt[4] = a1.x; c.Rt[5] = a1.y; c.Rt[6] = a1.z; c.Rt[7] = a1.w; c.Rt[8] = a2.x;
    const double* q = P.cam_Q + 12 * (size_t)cam;
    const d4 q0 = ld256(q), q1 = ld256(q + 4), q2 = ld256(q + 8);
    c.Q[0] = q0.x; c.Q[1] = q0.y; c.Q[2] = q0.z; c.Q[3] = q0.w; c.Q[4] = q1.x; c.Q[5] = q1.y; c.Q[6] = q1.z; c.Q[7] = q1.w; c.Q[8] = q2.x;
    c.dt[0] = q2.y; c.dt[1] = q2.z; c.dt[2] = q2.w;
    return c;
}
// A: rows (A[0..2]), (A[3..5]); Pb = Pbar of the slot; C out in the same layout
__device__ __forceinline__ void rebuild_C(const double* A, double p0, double p1, double p2, const CamRQ& c, double* C) {
    const double d0 = c.Rt[0] * p0 + c.Rt[3] * p1 + c.Rt[6] * p2 + c.dt[0];   // d = X - t = Rs Pbar + (ts - t), Rs = Rt^T
    const double d1 = c.Rt[1] * p0 + c.Rt[4] * p1 + c.Rt[7] * p2 + c.dt[1];
    const double d2 = c.Rt[2] * p0 + c.Rt[5] * p1 + c.Rt[8] * p2 + c.dt[2];
#pragma unroll
    for (int r = 0; r < 2; r++) {
        const double a0 = A[3 * r], a1 = A[3 * r + 1], a2 = A[3 * r + 2];
        const double w0 = a1 * d2 - a2 * d1, w1 = a2 * d0 - a0 * d2, w2 = a0 * d1 - a1 * d0;
#pragma unroll
        for (int k = 0; k < 3; k++) C[3 * r + k] = w0 * c.Q[3 * k] + w1 * c.Q[3 * k + 1] + w2 * c.Q[3 * k + 2];
    }
}

// Cheaper form for the passes that SUM over the observations of one camera: C_k = w_k Q^T with w_k = A_k x d, and Q is a
// per-camera constant, so they accumulate with the "pre-Q" rows J' = [w | -A] (21 FMA per observation instead of 39) and
// apply T = diag(Q, I) once per camera:  sum Jc^T M Jc = T (sum J'^T M J') T^T,  sum Jc^T r = T sum J'^T r.
// Rt = 12 doubles: Rt (9) of the sensor pose followed by dt = ts - t (load_rt)
__device__ __forceinline__ void preq_rows(const double* A, double p0, double p1, double p2, const double* Rt, double* w) {
    const double d0 = Rt[0] * p0 + Rt[3] * p1 + Rt[6] * p2 + Rt[9];   // d = X - t = Rs Pbar + (ts - t), Rs = Rt^T
    const double d1 = Rt[1] * p0 + Rt[4] * p1 + Rt[7] * p2 + Rt[10];
    const double d2 = Rt[2] * p0 + Rt[5] * p1 + Rt[8] * p2 + Rt[11];
#pragma unroll
    for (int r = 0; r < 2; r++) {
        const double a0 = A[3 * r], a1 = A[3 * r + 1], a2 = A[3 * r + 2];
        w[3 * r] = a1 * d2 - a2 * d1; w[3 * r + 1] = a2 * d0 - a0 * d2; w[3 * r + 2] = a0 * d1 - a1 * d0;
    }
}
__device__ __forceinline__ void load_rt(const DevProblem& P, int cam, double* Rt) {
    const double* cc = P.cam_cache + (size_t)CAM_CACHE_STRIDE * cam;
    const d4 a0 = ld256(cc), a1 = ld256(cc + 4), a2 = ld256(cc + 8);
    Rt[0] = a0.x; Rt[1] = a0.y; Rt[2] = a0.z; Rt[3] = a0.w; Rt[4] = a1.x; Rt[5] = a1.y; Rt[6] = a1.z; Rt[7] = a1.w; Rt[8] = a2.x;
    const d4 q2 = ld256(P.cam_Q + 12 * (size_t)cam + 8);   // (Q8, dt0, dt1, dt2)
    Rt[9] = q2.y; Rt[10] = q2.z; Rt[11] = q2.w;
}
// in place: packed symmetric 6x6 H' (21, row-major upper) -> T H' T^T, and g' (6) -> T g' (either may be null), T = diag(Q, I)
__device__ __forceinline__ void preq_finish(const double* Q, double* H21, double* g6) {
    if (H21) {
        double M[36];
        int q = 0;
#pragma unroll
        for (int a = 0; a < 6; a++)
#pragma unroll
            for (int b = a; b < 6; b++) { M[6 * a + b] = H21[q]; M[6 * b + a] = H21[q]; q++; }
        double L[36];   // L = T M: rows 0..2 mixed by Q
#pragma unroll
        for (int c = 0; c < 6; c++) {
#pragma unroll
            for (int r = 0; r < 3; r++) L[6 * r + c] = Q[3 * r] * M[c] + Q[3 * r + 1] * M[6 + c] + Q[3 * r + 2] * M[12 + c];
#pragma unroll
            for (int r = 3; r < 6; r++) L[6 * r + c] = M[6 * r + c];
        }
        q = 0;   // (L T^T)[a][b] for b >= a: columns 0..2 mixed by Q
#pragma unroll
        for (int a = 0; a < 6; a++)
#pragma unroll
            for (int b = a; b < 6; b++) {
                H21[q++] = b < 3 ? L[6 * a] * Q[3 * b] + L[6 * a + 1] * Q[3 * b + 1] + L[6 * a + 2] * Q[3 * b + 2] : L[6 * a + b];
            }
    }
    if (g6) {
        const double g0 = g6[0], g1 = g6[1], g2 = g6[2];
        g6[0] = Q[0] * g0 + Q[1] * g1 + Q[2] * g2; g6[1] = Q[3] * g0 + Q[4] * g1 + Q[5] * g2; g6[2] = Q[6] * g0 + Q[7] * g1 + Q[8] * g2;
    }
}

// per-camera operand of the point pass (sba_kernels_points.cu) for the input vector x6 = (x_r, x_t) of camera c:
//   (a_c = R Jr x_r, g_c = x_t + (t - c0) x a_c, pad 2)   -- 64 B, two 256-bit loads --
// so that  C x_r - A x_t = A ((X - t) x a_c - x_t) = A ((X - c0) x a_c - g_c).  c0 = P.center (fixed at finalize, the
// mean camera position) keeps |X - c0|, |t - c0| of the order of the scene extent whatever the world origin is.
constexpr int CAM_VEC_STRIDE = 8;
__device__ __forceinline__ void cam_vec_store(const DevProblem& P, int c, const double* x6) {
    const double* cc = P.cam_cache + (size_t)CAM_CACHE_STRIDE * c;
    const double* Q = P.cam_Q + 12 * (size_t)c;   // Q = (R Jr)^T of the BODY pose, Q[9..11] = ts - t
    // a = R Jr x_r = Q^T x_r
    const double a0 = Q[0] * x6[0] + Q[3] * x6[1] + Q[6] * x6[2];
    const double a1 = Q[1] * x6[0] + Q[4] * x6[1] + Q[7] * x6[2];
    const double a2 = Q[2] * x6[0] + Q[5] * x6[1] + Q[8] * x6[2];
    const double t0 = cc[9] - Q[9] - P.center[0], t1 = cc[10] - Q[10] - P.center[1], t2 = cc[11] - Q[11] - P.center[2];   // body t - c0
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
