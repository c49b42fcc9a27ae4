// sba_kernels_op.cu -- LM damping, point-block inversion, SCHUR_JACOBI
// preconditioner and the two passes of the implicit Schur operator
//   S x = (Hcc + Dc) x - sum_p W_p (Hpp + Dp)^-1 W_p^T x,   W_o = Jc_o^T Jp_o
// (Ceres ImplicitSchurComplement behind modularsbasolver.cpp:938; g2o analogue
// modules/g2o/g2o/core/block_solver.hpp:371-447).
#include <cstdlib>

#include "sba_kernel_util.cuh"

namespace sba {

// Ceres jacobi_scaling: scale_j = 1 / (1 + sqrt(diag(J^T J)_j)), computed once.
__global__ void k_scaling(DevProblem P, int enabled) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P.Nc * 6) {
        const int c = i / 6, a = i - 6 * c;
        P.scale_c[i] = enabled ? 1.0 / (1.0 + sqrt(P.Hcc[21 * (size_t)c + sym6(a, a)])) : 1.0;
    }
    if (i < P.Np) {
        const double* H = P.Hpp + 8 * (size_t)i;
        double* s = P.scale_p + 4 * (size_t)i;
        s[0] = enabled ? 1.0 / (1.0 + sqrt(H[0])) : 1.0;
        s[1] = enabled ? 1.0 / (1.0 + sqrt(H[3])) : 1.0;
        s[2] = enabled ? 1.0 / (1.0 + sqrt(H[5])) : 1.0;
        s[3] = 0.0;
    }
}
void launch_scaling(const DevProblem& P, int enabled, cudaStream_t st) {
    const int n = max(P.Nc * 6, P.Np);
    k_scaling<<<cdiv(n, 256), 256, 0, st>>>(P, enabled);
}

// Ceres LevenbergMarquardtStrategy: D^2 = clamp(diag(J~^T J~), min, max) / radius in the
// scaled variables; in the unscaled variables the damping is D^2 / scale^2.
__device__ __forceinline__ double lm_diag(double h, double s, double radius, double dmin, double dmax) {
    const double d = fmin(fmax(h * s * s, dmin), dmax);
    return d / radius / (s * s);
}
__global__ void k_damping(DevProblem P, double radius, double dmin, double dmax) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P.Nc * 6) {
        const int c = i / 6, a = i - 6 * c;
        P.lam_c[i] = lm_diag(P.Hcc[21 * (size_t)c + sym6(a, a)], P.scale_c[i], radius, dmin, dmax);
    }
    if (i < P.Np) {
        const double* H = P.Hpp + 8 * (size_t)i;
        const double* s = P.scale_p + 4 * (size_t)i;
        double* l = P.lam_p + 4 * (size_t)i;
        l[0] = lm_diag(H[0], s[0], radius, dmin, dmax);
        l[1] = lm_diag(H[3], s[1], radius, dmin, dmax);
        l[2] = lm_diag(H[5], s[2], radius, dmin, dmax);
        l[3] = 0.0;
    }
}
void launch_damping(const DevProblem& P, double radius, double dmin, double dmax, cudaStream_t st) {
    const int n = max(P.Nc * 6, P.Np);
    k_damping<<<cdiv(n, 256), 256, 0, st>>>(P, radius, dmin, dmax);
}

// per-point damped inverse in registers (SchurEliminator e-block step)
__global__ void __launch_bounds__(256) k_invert_points(DevProblem P) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P.Np) return;
    double Hi[6] = {0, 0, 0, 0, 0, 0};
    if (!P.pt_fixed[p]) {
        const double* Hp = P.Hpp + 8 * (size_t)p;
        const double* l = P.lam_p + 4 * (size_t)p;
        const double2 h01 = ldg2(Hp), h23 = ldg2(Hp + 2), h45 = ldg2(Hp + 4), l01 = ldg2(l), l23 = ldg2(l + 2);
        const double H[6] = {h01.x + l01.x, h01.y, h23.x, h23.y + l01.y, h45.x, h45.y + l23.x};
        if (!inv3_sym(H, Hi)) {
            P.flags[FL_SINGULAR_PT] = 1;
#pragma unroll
            for (int i = 0; i < 6; i++) Hi[i] = 0.0;
        }
    }
    double* o = P.Hinv + 8 * (size_t)p;
    st2(o, Hi[0], Hi[1]); st2(o + 2, Hi[2], Hi[3]); st2(o + 4, Hi[4], Hi[5]); st2(o + 6, 0.0, 0.0);
}
void launch_invert_points(const DevProblem& P, cudaStream_t st) {
    if (P.Np == 0) return;
    k_invert_points<<<cdiv(P.Np, 256), 256, 0, st>>>(P);
}

// ---------------------------------------------------------------------------
// diagonal blocks of S: warp per tile accumulates sum_o Jc^T (A Hinv A^T) Jc
// (21 unique entries), gathers Hinv_p (64 B) per observation.
// ---------------------------------------------------------------------------
// PIPE: Rt in shared memory, the lane's next slot in flight during the arithmetic of the current one (point index two slots
// ahead, planes and the Hinv gather one slot ahead): the kernel is latency-bound at 16 warps / SM (ncu r2f: DRAM 38 %, 22 % warps active).
struct SdIn { double A[6], p[3]; d4 h0, h1; };   // h0 = (h00 h01 h02 h11), h1 = (h12 h22 - -)
__device__ __forceinline__ void sd_load(const DevProblem& P, size_t Ns, int s, int pt, SdIn& x) {
#pragma unroll
    for (int a = 0; a < 6; a++) x.A[a] = __ldcs(P.Ac + (size_t)a * Ns + s);
    x.p[0] = __ldcs(P.Pc + s); x.p[1] = __ldcs(P.Pc + Ns + s); x.p[2] = __ldcs(P.Pc + 2 * Ns + s);
    const double* Hp = P.Hinv + 8 * (size_t)max(pt, 0);
    x.h0 = ld256(Hp); x.h1 = ld256(Hp + 4);
}
__device__ __forceinline__ void sd_accumulate(const SdIn& x, const double* Rt, double (&D)[21]) {
    const double* A = x.A;
    double w[6], j0[6], j1[6];   // pre-Q rows J' = [A x d | -A]; diag(Q, I) is applied per camera after the tile sum
    preq_rows(A, x.p[0], x.p[1], x.p[2], Rt, w);
#pragma unroll
    for (int a = 0; a < 3; a++) { j0[a] = w[a]; j1[a] = w[3 + a]; j0[3 + a] = -A[a]; j1[3 + a] = -A[3 + a]; }
    const double h00 = x.h0.x, h01 = x.h0.y, h02 = x.h0.z, h11 = x.h0.w, h12 = x.h1.x, h22 = x.h1.y;
    // BH = A * Hinv (2x3), T = BH * A^T (2x2)
    const double b00 = A[0] * h00 + A[1] * h01 + A[2] * h02;
    const double b01 = A[0] * h01 + A[1] * h11 + A[2] * h12;
    const double b02 = A[0] * h02 + A[1] * h12 + A[2] * h22;
    const double b10 = A[3] * h00 + A[4] * h01 + A[5] * h02;
    const double b11 = A[3] * h01 + A[4] * h11 + A[5] * h12;
    const double b12 = A[3] * h02 + A[4] * h12 + A[5] * h22;
    const double T00 = b00 * A[0] + b01 * A[1] + b02 * A[2];
    const double T01 = b00 * A[3] + b01 * A[4] + b02 * A[5];
    const double T10 = b10 * A[0] + b11 * A[1] + b12 * A[2];
    const double T11 = b10 * A[3] + b11 * A[4] + b12 * A[5];
    int q = 0;
#pragma unroll
    for (int a = 0; a < 6; a++) {
        const double t0 = j0[a] * T00 + j1[a] * T10, t1 = j0[a] * T01 + j1[a] * T11;
#pragma unroll
        for (int b = a; b < 6; b++) D[q++] += t0 * j0[b] + t1 * j1[b];
    }
}
template <bool PIPE>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32, 2) k_schur_diag(DevProblem P) {
    __shared__ double s_rt[PIPE ? WARPS_PER_BLOCK : 1][12];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * WARPS_PER_BLOCK + warp;
    if (tile >= P.Nt) return;
    const int start = __ldg(P.tile_start + tile), end = start + __ldg(P.tile_len + tile);
    const size_t Ns = (size_t)P.Ns;
    double D[21];
#pragma unroll
    for (int i = 0; i < 21; i++) D[i] = 0.0;
    if (!PIPE) {
        double Rt[12];
        load_rt(P, __ldg(P.tile_cam + tile), Rt);
        for (int s = start + lane; s < end; s += 32) {
            const int pt = __ldcs(P.obs_pt + s);
            if (pt < 0) continue;
            SdIn x;
            sd_load(P, Ns, s, pt, x);
            sd_accumulate(x, Rt, D);
        }
    } else {
        double* Rt = s_rt[warp];
        {
            const int cam = __ldg(P.tile_cam + tile);
            if (lane < 9) Rt[lane] = __ldg(P.cam_cache + (size_t)CAM_CACHE_STRIDE * cam + lane);
            else if (lane < 12) Rt[lane] = __ldg(P.cam_Q + 12 * (size_t)cam + lane);   // dt = ts - t behind Q (load_rt)
        }
        __syncwarp();
        int s = start + lane;
        int pt1 = s < end ? __ldcs(P.obs_pt + s) : -1;
        int pt2 = s + 32 < end ? __ldcs(P.obs_pt + s + 32) : -1;
        SdIn cur, nxt;
        if (s < end) sd_load(P, Ns, s, pt1, cur);
        for (; s < end; s += 32) {
            const int pt = pt1;
            pt1 = pt2;
            pt2 = s + 64 < end ? __ldcs(P.obs_pt + s + 64) : -1;
            const bool more = s + 32 < end;
            if (more) sd_load(P, Ns, s + 32, pt1, nxt);
            if (pt >= 0) sd_accumulate(cur, Rt, D);   // padding slot: nothing to add
            if (more) cur = nxt;
        }
    }
    double* out = P.tile_part + (size_t)TP * tile;
#pragma unroll
    for (int i = 0; i < 21; i++) { const double v = warp_sum(D[i]); if (lane == 0) out[i] = v; }
}
__global__ void k_tile_reduce21(DevProblem P, double* dst) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.Nc * 21) return;
    const int c = i / 21, comp = i - c * 21;
    double acc = 0.0;
    if (!P.cam_fixed[c]) {
        const int t0 = P.cam_tile_off[c], t1 = P.cam_tile_off[c + 1];
        for (int t = t0; t < t1; t++) acc += P.tile_part[(size_t)TP * t + comp];
    }
    dst[i] = acc;
}
void launch_schur_diag(const DevProblem& P, cudaStream_t st) {
    const char* e = std::getenv("SBA_SD_PIPE");   // measurement knob
    // measured slower than the plain loop on the Venice shape (0.1864 vs 0.1828 ms incl. reduce + inversion, scripts/ab_tiles.py:
    // 128 registers with spills): kept as a knob, off by default
    const bool pipe = e ? std::atoi(e) != 0 : false;
    if (P.Nt) {
        if (pipe) k_schur_diag<true><<<cdiv(P.Nt, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, 0, st>>>(P);
        else k_schur_diag<false><<<cdiv(P.Nt, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, 0, st>>>(P);
    }
    k_tile_reduce21<<<cdiv(P.Nc * 21, 256), 256, 0, st>>>(P, P.Sd);
    launch_preq_finish(P, P.Sd, nullptr, st);
}

// Minv = (Hcc + lam_c - Sd)^-1 per camera (6x6 Cholesky in registers)
__global__ void __launch_bounds__(128) k_schur_diag_invert(DevProblem P) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= P.Nc) return;
    double D[36], Di[36];
    const double* H = P.Hcc + 21 * (size_t)c;
    const double* S = P.Sd + 21 * (size_t)c;
    int q = 0;
#pragma unroll
    for (int a = 0; a < 6; a++)
#pragma unroll
        for (int b = a; b < 6; b++) {
            double v = H[q] - S[q];
            if (a == b) v += P.lam_c[6 * (size_t)c + a];
            D[6 * a + b] = v; D[6 * b + a] = v;
            q++;
        }
    if (!inv6_spd(D, Di)) {
        P.flags[FL_NOT_PD] = 1;
#pragma unroll
        for (int i = 0; i < 36; i++) Di[i] = 0.0;
    }
#pragma unroll
    for (int i = 0; i < 36; i++) P.Minv[36 * (size_t)c + i] = Di[i];
}
void launch_schur_diag_invert(const DevProblem& P, cudaStream_t st) {
    k_schur_diag_invert<<<cdiv(P.Nc, 128), 128, 0, st>>>(P);
}

}  // namespace sba
