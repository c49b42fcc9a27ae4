// sba_kernels_points.cu -- the point-order ("q slot") kernels: K1p linearisation, K2a point blocks and
// pass 1 of the implicit Schur operator / back-substitution.
//
// One thread per q slot; the planes qA, qC, qres are read as coalesced 256-byte runs per warp.  The sum
// over the observations of a point is a segmented warp reduction: points with k <= 32 observations never
// straddle a 32-slot window (layout builder), so a fixed-shape shuffle-down tree keyed on the point index
// gives every point's sum in its first lane -- deterministic, no atomics.  Points with k > 32 ("heavy")
// own whole windows; their per-window partials go to heavy_part and a second tiny kernel folds them in
// window order.
#include "sba_kernel_util.cuh"

namespace sba {

// after the call the first lane of every run of equal keys holds the run's sum (fixed order)
template <int N>
__device__ __forceinline__ void seg_reduce(double (&v)[N], int key, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int k2 = __shfl_down_sync(FULL, key, o);
        const bool take = (lane + o < 32) && (k2 == key);
#pragma unroll
        for (int i = 0; i < N; i++) {
            const double t = __shfl_down_sync(FULL, v[i], o);
            if (take) v[i] += t;
        }
    }
}
__device__ __forceinline__ bool seg_head(int key, int lane) {
    const int prev = __shfl_up_sync(FULL, key, 1);
    return lane == 0 || prev != key;
}

// ---------------------------------------------------------------------------
// K1p: linearisation in point order.  Same expressions as K1 (sba_math.cuh), so both orders hold
// bitwise identical A, C, res.  In: 16 (uv) + 8 (pt, cam idx) per slot, camera record gathered from
// L2 (168 B), point from L1/L2; out: 112 B per slot as 14 coalesced planes.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_linearize_q(DevProblem P) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= P.Nq) return;
    const int pt = __ldcs(P.q_pt + q);
    if (pt < 0) return;   // padding: planes stay zero
    const int cam = __ldcs(P.q_cam + q);
    const double2 x01 = ldg2(P.pts + 4 * (size_t)pt), x23 = ldg2(P.pts + 4 * (size_t)pt + 2);
    const double X[3] = {x01.x, x01.y, x23.x};
    const double* cc = P.cam_cache + (size_t)CAM_CACHE_STRIDE * cam;
    double ccr[22];
#pragma unroll
    for (int i = 0; i < 11; i++) { const double2 v = ldg2(cc + 2 * i); ccr[2 * i] = v.x; ccr[2 * i + 1] = v.y; }
    const double* in = P.intr + 10 * __ldg(P.cam_model + cam);
    double intr[10];
#pragma unroll
    for (int i = 0; i < 5; i++) { const double2 v = ldg2(in + 2 * i); intr[2 * i] = v.x; intr[2 * i + 1] = v.y; }
    const double2 uv = ldcs2(P.q_uv + 2 * (size_t)q);
    double d[3];
    dir_from_uv(uv.x, uv.y, intr, d);
    double W[4];
    const double* Wp = nullptr;
    if (P.q_info) {
        const double2 w01 = ldcs2(P.q_info + 4 * (size_t)q), w23 = ldcs2(P.q_info + 4 * (size_t)q + 2);
        W[0] = w01.x; W[1] = w01.y; W[2] = w23.x; W[3] = w23.y;
        Wp = W;
    }
    double r[2] = {0.0, 0.0}, Jc[12], Jp[6];
    const bool ok = obs_eval<true>(ccr, X, d, Wp, P.cam_kind[cam], r, Jc, Jp);
    if (!ok) {
        r[0] = r[1] = 0.0;
#pragma unroll
        for (int i = 0; i < 12; i++) Jc[i] = 0.0;
#pragma unroll
        for (int i = 0; i < 6; i++) Jp[i] = 0.0;
    }
    const size_t Nq = (size_t)P.Nq;
    stcs(P.qres + (q), r[0]);
    stcs(P.qres + (Nq + q), r[1]);
#pragma unroll
    for (int j = 0; j < 3; j++) {
        stcs(P.qC + ((size_t)j * Nq + q), Jc[j]);
        stcs(P.qC + ((size_t)(3 + j) * Nq + q), Jc[6 + j]);
        stcs(P.qA + ((size_t)j * Nq + q), Jp[j]);
        stcs(P.qA + ((size_t)(3 + j) * Nq + q), Jp[3 + j]);
    }
}
void launch_linearize_points(const DevProblem& P, cudaStream_t st) {
    if (P.Nq == 0) return;
    k_linearize_q<<<cdiv(P.Nq, 256), 256, 0, st>>>(P);
}

// ---------------------------------------------------------------------------
// K2a point blocks: Hpp = sum A^T A (packed sym), gp = sum A^T res.  64 B/slot in (A, res planes),
// 96 B/point out.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void store_point_blocks(const DevProblem& P, int pt, const double* v) {
    const bool fixed = P.pt_fixed[pt];
    double* Hp = P.Hpp + 8 * (size_t)pt;
    double* gp = P.gp + 4 * (size_t)pt;
    if (fixed) {
        st2(Hp, 0.0, 0.0); st2(Hp + 2, 0.0, 0.0); st2(Hp + 4, 0.0, 0.0); st2(Hp + 6, 0.0, 0.0);
        st2(gp, 0.0, 0.0); st2(gp + 2, 0.0, 0.0);
    } else {
        st2(Hp, v[0], v[1]); st2(Hp + 2, v[2], v[3]); st2(Hp + 4, v[4], v[5]); st2(Hp + 6, 0.0, 0.0);
        st2(gp, v[6], v[7]); st2(gp + 2, v[8], 0.0);
    }
}
__global__ void __launch_bounds__(256) k_point_blocks_q(DevProblem P) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;   // Nq is a multiple of 32: whole warps in range
    if (q >= P.Nq) return;
    const int lane = threadIdx.x & 31;
    const int pt = __ldcs(P.q_pt + q);
    const size_t Nq = (size_t)P.Nq;
    double v[9];
    {
        double A[6];
#pragma unroll
        for (int a = 0; a < 6; a++) A[a] = __ldcs(P.qA + (size_t)a * Nq + q);
        const double r0 = __ldcs(P.qres + q), r1 = __ldcs(P.qres + Nq + q);
        v[0] = A[0] * A[0] + A[3] * A[3]; v[1] = A[0] * A[1] + A[3] * A[4]; v[2] = A[0] * A[2] + A[3] * A[5];
        v[3] = A[1] * A[1] + A[4] * A[4]; v[4] = A[1] * A[2] + A[4] * A[5]; v[5] = A[2] * A[2] + A[5] * A[5];
        v[6] = A[0] * r0 + A[3] * r1; v[7] = A[1] * r0 + A[4] * r1; v[8] = A[2] * r0 + A[5] * r1;
    }
    seg_reduce<9>(v, pt, lane);
    if (pt >= 0 && seg_head(pt, lane)) {
        const int h = __ldg(P.pt_heavy + pt);
        if (h < 0) {
            store_point_blocks(P, pt, v);
        } else {
            double* o = P.heavy_part + (size_t)HP * (__ldg(P.heavy_w0 + h) + ((q - __ldg(P.pt_q0 + pt)) >> 5));
#pragma unroll
            for (int i = 0; i < 9; i++) o[i] = v[i];
        }
    }
}
__global__ void k_point_blocks_heavy(DevProblem P) {
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= P.Nh) return;
    double v[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int w = P.heavy_w0[h]; w < P.heavy_w0[h + 1]; w++)
#pragma unroll
        for (int i = 0; i < 9; i++) v[i] += P.heavy_part[(size_t)HP * w + i];
    store_point_blocks(P, P.heavy_pt[h], v);
}
void launch_point_blocks(const DevProblem& P, cudaStream_t st) {
    if (P.Nq == 0) return;
    k_point_blocks_q<<<cdiv(P.Nq, 256), 256, 0, st>>>(P);
    if (P.Nh) k_point_blocks_heavy<<<cdiv(P.Nh, 128), 128, 0, st>>>(P);
}

// ---------------------------------------------------------------------------
// operator pass 1 / back-substitution.  x must be zero at fixed cameras.
//   MODE 0: y_p = Hinv * sum_o A^T (C x_r - A x_t)
//   MODE 2: dp  = -Hinv * (gp + sum_o A^T (C x_r - A x_t))   (block_solver.hpp:455-477 analogue)
// 104 B/slot in (A, C planes + pt, cam idx), camera vector gathered (48 B, L1-resident: 85 KB for
// 1778 cameras); 96 B/point in (Hinv [+ gp]) and 32 B/point out, by the head lanes only.
// ---------------------------------------------------------------------------
template <int MODE>
__device__ __forceinline__ void finish_point_pass(const DevProblem& P, int pt, double u0, double u1, double u2) {
    if (MODE == 2) {
        const double* g = P.gp + 4 * (size_t)pt;
        const double2 g01 = ldg2(g), g23 = ldg2(g + 2);
        u0 += g01.x; u1 += g01.y; u2 += g23.x;
    }
    const double* Hp = P.Hinv + 8 * (size_t)pt;
    const double2 h01 = ldg2(Hp), h23 = ldg2(Hp + 2), h45 = ldg2(Hp + 4);
    double y0 = h01.x * u0 + h01.y * u1 + h23.x * u2;
    double y1 = h01.y * u0 + h23.y * u1 + h45.x * u2;
    double y2 = h23.x * u0 + h45.x * u1 + h45.y * u2;
    double* o = (MODE == 2 ? P.dp : P.y_p) + 4 * (size_t)pt;
    if (MODE == 2) { y0 = -y0; y1 = -y1; y2 = -y2; }
    st2(o, y0, y1); st2(o + 2, y2, 0.0);
}
template <int MODE>
__global__ void __launch_bounds__(256) k_point_pass_q(DevProblem P, const double* __restrict__ x) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= P.Nq) return;
    const int lane = threadIdx.x & 31;
    const int pt = __ldcs(P.q_pt + q);
    const size_t Nq = (size_t)P.Nq;
    double u[3] = {0.0, 0.0, 0.0};
    if (pt >= 0) {
        const double* xc = x + 6 * (size_t)__ldcs(P.q_cam + q);
        const double2 x01 = ldg2(xc), x23 = ldg2(xc + 2), x45 = ldg2(xc + 4);
        double A[6], C[6];
#pragma unroll
        for (int a = 0; a < 6; a++) {
            A[a] = __ldcs(P.qA + (size_t)a * Nq + q);
            C[a] = __ldcs(P.qC + (size_t)a * Nq + q);
        }
        // rows of Jc x = C x_r - A x_t
        const double m0 = C[0] * x01.x + C[1] * x01.y + C[2] * x23.x - (A[0] * x23.y + A[1] * x45.x + A[2] * x45.y);
        const double m1 = C[3] * x01.x + C[4] * x01.y + C[5] * x23.x - (A[3] * x23.y + A[4] * x45.x + A[5] * x45.y);
        u[0] = A[0] * m0 + A[3] * m1;
        u[1] = A[1] * m0 + A[4] * m1;
        u[2] = A[2] * m0 + A[5] * m1;
    }
    seg_reduce<3>(u, pt, lane);
    if (pt >= 0 && seg_head(pt, lane)) {
        const int h = __ldg(P.pt_heavy + pt);
        if (h < 0) {
            finish_point_pass<MODE>(P, pt, u[0], u[1], u[2]);
        } else {
            double* o = P.heavy_part + (size_t)HP * (__ldg(P.heavy_w0 + h) + ((q - __ldg(P.pt_q0 + pt)) >> 5));
            o[0] = u[0]; o[1] = u[1]; o[2] = u[2];
        }
    }
}
template <int MODE>
__global__ void k_point_pass_heavy(DevProblem P) {
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= P.Nh) return;
    double u0 = 0.0, u1 = 0.0, u2 = 0.0;
    for (int w = P.heavy_w0[h]; w < P.heavy_w0[h + 1]; w++) {
        u0 += P.heavy_part[(size_t)HP * w]; u1 += P.heavy_part[(size_t)HP * w + 1]; u2 += P.heavy_part[(size_t)HP * w + 2];
    }
    finish_point_pass<MODE>(P, P.heavy_pt[h], u0, u1, u2);
}
// MODE 1: y_p = Hinv * gp, thread per point (no observation data)
__global__ void __launch_bounds__(256) k_point_hinv_g(DevProblem P) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P.Np) return;
    const double* g = P.gp + 4 * (size_t)p;
    const double2 g01 = ldg2(g), g23 = ldg2(g + 2);
    const double* Hp = P.Hinv + 8 * (size_t)p;
    const double2 h01 = ldg2(Hp), h23 = ldg2(Hp + 2), h45 = ldg2(Hp + 4);
    double* o = P.y_p + 4 * (size_t)p;
    st2(o, h01.x * g01.x + h01.y * g01.y + h23.x * g23.x, h01.y * g01.x + h23.y * g01.y + h45.x * g23.x);
    st2(o + 2, h23.x * g01.x + h45.x * g01.y + h45.y * g23.x, 0.0);
}
void launch_point_pass(const DevProblem& P, const double* x, int mode, cudaStream_t st) {
    if (P.Np == 0) return;
    if (mode == 1) { k_point_hinv_g<<<cdiv(P.Np, 256), 256, 0, st>>>(P); return; }
    const int nb = cdiv(P.Nq, 256);
    if (mode == 0) {
        k_point_pass_q<0><<<nb, 256, 0, st>>>(P, x);
        if (P.Nh) k_point_pass_heavy<0><<<cdiv(P.Nh, 128), 128, 0, st>>>(P);
    } else {
        k_point_pass_q<2><<<nb, 256, 0, st>>>(P, x);
        if (P.Nh) k_point_pass_heavy<2><<<cdiv(P.Nh, 128), 128, 0, st>>>(P);
    }
}

}  // namespace sba
