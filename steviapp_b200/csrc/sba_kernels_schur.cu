// sba_kernels_schur.cu -- K2: normal-equation blocks, point-block inversion,
// SCHUR_JACOBI preconditioner, and the implicit Schur operator S x.
//
// What ceres::Solve does for the reference behind modularsbasolver.cpp:938
// (SchurEliminator / ImplicitSchurComplement / SchurJacobiPreconditioner); the
// in-tree statement of the same algebra is g2o's BlockSolver::solve
// (modules/g2o/g2o/core/block_solver.hpp:329-479).
#include <cstdlib>

#include "sba_kernel_util.cuh"

namespace sba {

// ---------------------------------------------------------------------------
// priors: FixedSizeNormalPrior<3,3>  r = A (x - b), J = A
// (costfunctors/fixedsizenormalprior.h:89-108).  One thread per target group.
// ---------------------------------------------------------------------------
__device__ __forceinline__ const double* prior_target(const DevProblem& P, int kind, int idx) {
    return kind == 0 ? P.pts + 4 * (size_t)idx : (kind == 1 ? P.cam_r + 3 * (size_t)idx : P.cam_t + 3 * (size_t)idx);
}

// prior residuals; a single block so the cost sum is a fixed tree
__global__ void __launch_bounds__(RED_BLOCK) k_prior_eval(DevProblem P, int store) {
    __shared__ double sh[33];
    double acc = 0.0, acc_cam = 0.0;   // point priors / camera priors (the latter are replicated on every rank)
    for (int g = threadIdx.x; g < P.Ngroup; g += blockDim.x) {
        const int is_cam = P.pg_kind[g] != 0;
        const double* x = prior_target(P, P.pg_kind[g], P.pg_index[g]);
        for (int i = P.pg_off[g]; i < P.pg_off[g + 1]; i++) {
            const double* A = P.prior_A + 9 * (size_t)i;
            const double* b = P.prior_b + 3 * (size_t)i;
            const double e0 = x[0] - b[0], e1 = x[1] - b[1], e2 = x[2] - b[2];
#pragma unroll
            for (int k = 0; k < 3; k++) {
                const double r = A[3 * k] * e0 + A[3 * k + 1] * e1 + A[3 * k + 2] * e2;
                if (store) P.prior_res[3 * (size_t)i + k] = r;
                if (is_cam) acc_cam += r * r; else acc += r * r;
            }
        }
    }
    acc = block_sum(acc, sh);
    acc_cam = block_sum(acc_cam, sh);
    if (threadIdx.x == 0) { P.scalars[SC_COST_PRIOR] = acc; P.scalars[SC_COST_PRIOR_CAM] = acc_cam; }
}
void launch_prior_eval(const DevProblem& P, int store, cudaStream_t st) { k_prior_eval<<<1, RED_BLOCK, 0, st>>>(P, store); }

// multi-GPU: this rank's (sum r^2, cheirality flag) appended to the Hcc|gc all-reduce payload
__global__ void k_pack_tail(DevProblem P, double* tail) {
    tail[0] = P.scalars[SC_COST_OBS] + P.scalars[SC_COST_PRIOR];
    tail[1] = P.flags[FL_CHEIRALITY] ? 1.0 : 0.0;
}
void launch_pack_tail(const DevProblem& P, double* tail, cudaStream_t st) { k_pack_tail<<<1, 1, 0, st>>>(P, tail); }

// sum of the per-tile observation costs (fixed tree) -> SC_COST_OBS (sum r^2, not halved)
__global__ void __launch_bounds__(RED_BLOCK) k_cost_reduce(DevProblem P) {
    __shared__ double sh[33];
    double acc = 0.0;
    for (int t = threadIdx.x; t < P.Nt; t += blockDim.x) acc += P.tile_cost[t];
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) P.scalars[SC_COST_OBS] = acc;
}
void launch_cost_reduce(const DevProblem& P, cudaStream_t st) { k_cost_reduce<<<1, RED_BLOCK, 0, st>>>(P); }

// adds A^T A and A^T r of the priors to the diagonal blocks (after the observation sums)
__global__ void k_prior_blocks(DevProblem P, int cameras) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= P.Ngroup) return;
    const int kind = P.pg_kind[g], idx = P.pg_index[g];
    if ((kind != 0) != (cameras != 0)) return;
    if (kind == 0 ? P.pt_fixed[idx] : P.cam_fixed[idx]) return;
    double H[6] = {0, 0, 0, 0, 0, 0}, gg[3] = {0, 0, 0};
    for (int i = P.pg_off[g]; i < P.pg_off[g + 1]; i++) {
        const double* A = P.prior_A + 9 * (size_t)i;
        const double* r = P.prior_res + 3 * (size_t)i;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const double a0 = A[3 * k], a1 = A[3 * k + 1], a2 = A[3 * k + 2];
            gg[0] += a0 * r[k]; gg[1] += a1 * r[k]; gg[2] += a2 * r[k];
            H[0] += a0 * a0; H[1] += a0 * a1; H[2] += a0 * a2; H[3] += a1 * a1; H[4] += a1 * a2; H[5] += a2 * a2;
        }
    }
    if (kind == 0) {
        double* Hp = P.Hpp + 8 * (size_t)idx;
        double* gp = P.gp + 4 * (size_t)idx;
        // a point without observations heads no q slot, so k_linearize_q never rewrites its blocks: the priors ARE its
        // blocks (assigned, not accumulated -- otherwise they would grow by A^T A with every linearisation)
        const bool unobserved = P.pt_off[idx] == P.pt_off[idx + 1];
#pragma unroll
        for (int i = 0; i < 6; i++) Hp[i] = (unobserved ? 0.0 : Hp[i]) + H[i];
#pragma unroll
        for (int i = 0; i < 3; i++) gp[i] = (unobserved ? 0.0 : gp[i]) + gg[i];
    } else {
        const int off = kind == 1 ? 0 : 3;
        double* Hc = P.Hcc + 21 * (size_t)idx;
        double* gc = P.gc + 6 * (size_t)idx + off;
        Hc[sym6(off, off)] += H[0]; Hc[sym6(off, off + 1)] += H[1]; Hc[sym6(off, off + 2)] += H[2];
        Hc[sym6(off + 1, off + 1)] += H[3]; Hc[sym6(off + 1, off + 2)] += H[4]; Hc[sym6(off + 2, off + 2)] += H[5];
        gc[0] += gg[0]; gc[1] += gg[1]; gc[2] += gg[2];
    }
}
void launch_prior_blocks(const DevProblem& P, int cameras, cudaStream_t st) {
    if (P.Ngroup == 0) return;
    k_prior_blocks<<<cdiv(P.Ngroup, 128), 128, 0, st>>>(P, cameras);
}

// ---------------------------------------------------------------------------
// K2b camera blocks: warp per tile.  Every lane accumulates its slots'
// Jc^T Jc (21 unique) and Jc^T res (6) with Jc = [C | -A]; a fixed xor-shuffle
// tree folds the 32 lanes; lane 0 writes the tile partial.  k_tile_reduce then
// sums the tiles of each camera in tile order.
// Reads 88 B/obs (res, A, Pbar planes, coalesced); C rebuilt in registers.
// ---------------------------------------------------------------------------
// PIPE: Rt lives in shared memory (12 doubles per warp) and the 11 plane values of the lane's NEXT slot are loaded before the
// arithmetic of the current one: a pure stream at 16 warps / SM (120 registers) did not keep enough bytes in flight (0.58 of HBM).
struct K2bIn { double A[6], p[3], r[2]; };
__device__ __forceinline__ void k2b_load(const DevProblem& P, size_t Ns, int s, K2bIn& x) {
#pragma unroll
    for (int a = 0; a < 6; a++) x.A[a] = __ldcs(P.Ac + (size_t)a * Ns + s);
    x.p[0] = __ldcs(P.Pc + s); x.p[1] = __ldcs(P.Pc + Ns + s); x.p[2] = __ldcs(P.Pc + 2 * Ns + s);
    x.r[0] = __ldcs(P.res + s); x.r[1] = __ldcs(P.res + Ns + s);
}
__device__ __forceinline__ void k2b_accumulate(const K2bIn& x, const double* Rt, double (&H)[21], double (&g)[6]) {
    double j0[6], j1[6], w[6];   // the two pre-Q rows J' = [A x d | -A] (Jc = J' diag(Q, I)^T, applied per camera by k_preq_finish)
    preq_rows(x.A, x.p[0], x.p[1], x.p[2], Rt, w);
#pragma unroll
    for (int a = 0; a < 3; a++) { j0[a] = w[a]; j1[a] = w[3 + a]; j0[3 + a] = -x.A[a]; j1[3 + a] = -x.A[3 + a]; }
    int q = 0;
#pragma unroll
    for (int a = 0; a < 6; a++) {
        g[a] += j0[a] * x.r[0] + j1[a] * x.r[1];
#pragma unroll
        for (int b = a; b < 6; b++) H[q++] += j0[a] * j0[b] + j1[a] * j1[b];
    }
}
template <int PIPE>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32, 2) k_camera_blocks(DevProblem P) {
    __shared__ double s_rt[PIPE ? WARPS_PER_BLOCK : 1][12];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * WARPS_PER_BLOCK + warp;
    if (tile >= P.Nt) return;
    const int start = __ldg(P.tile_start + tile), end = start + __ldg(P.tile_len + tile);
    const size_t Ns = (size_t)P.Ns;
    double H[21], g[6];
#pragma unroll
    for (int i = 0; i < 21; i++) H[i] = 0.0;
#pragma unroll
    for (int i = 0; i < 6; i++) g[i] = 0.0;
    if (!PIPE) {
        double Rt[12];
        load_rt(P, __ldg(P.tile_cam + tile), Rt);
        for (int s = start + lane; s < end; s += 32) {
            K2bIn x;
            k2b_load(P, Ns, s, x);
            k2b_accumulate(x, Rt, H, g);
        }
    } else if (PIPE == 2) {
        // two consecutive slots per lane, 128-bit loads: twice the bytes in flight per load instruction (tiles start on even slots
        // and are padded to a multiple of 32: the pair never leaves the tile's padded range)
        double* Rt = s_rt[warp];
        {
            const int cam = __ldg(P.tile_cam + tile);
            if (lane < 9) Rt[lane] = __ldg(P.cam_cache + (size_t)CAM_CACHE_STRIDE * cam + lane);
            else if (lane < 12) Rt[lane] = __ldg(P.cam_Q + 12 * (size_t)cam + lane);
        }
        __syncwarp();
        for (int s = start + 2 * lane; s < end; s += 64) {
            K2bIn x0, x1;
#pragma unroll
            for (int a = 0; a < 6; a++) { const double2 v = ldcs2(P.Ac + (size_t)a * Ns + s); x0.A[a] = v.x; x1.A[a] = v.y; }
#pragma unroll
            for (int a = 0; a < 3; a++) { const double2 v = ldcs2(P.Pc + (size_t)a * Ns + s); x0.p[a] = v.x; x1.p[a] = v.y; }
#pragma unroll
            for (int a = 0; a < 2; a++) { const double2 v = ldcs2(P.res + (size_t)a * Ns + s); x0.r[a] = v.x; x1.r[a] = v.y; }
            k2b_accumulate(x0, Rt, H, g);
            k2b_accumulate(x1, Rt, H, g);
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
        K2bIn cur, nxt;
        if (s < end) k2b_load(P, Ns, s, cur);
        for (; s < end; s += 32) {
            const bool more = s + 32 < end;
            if (more) k2b_load(P, Ns, s + 32, nxt);
            k2b_accumulate(cur, Rt, H, g);
            if (more) cur = nxt;
        }
    }
    double* out = P.tile_part + (size_t)TP * tile;
#pragma unroll
    for (int i = 0; i < 21; i++) { const double v = warp_sum(H[i]); if (lane == 0) out[i] = v; }
#pragma unroll
    for (int i = 0; i < 6; i++) { const double v = warp_sum(g[i]); if (lane == 0) out[21 + i] = v; }
}

// thread per (camera, component): fixed-order sum over the camera's tiles.
// n components starting at tile-partial offset `src`; dst AoS with stride n.
__global__ void k_tile_reduce(DevProblem P, int src, int n, double* dst, int zero_fixed) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.Nc * n) return;
    const int c = i / n, comp = i - c * n;
    double acc = 0.0;
    if (!(zero_fixed && P.cam_fixed[c])) {
        const int t0 = P.cam_tile_off[c], t1 = P.cam_tile_off[c + 1];
        for (int t = t0; t < t1; t++) acc += P.tile_part[(size_t)TP * t + src + comp];
    }
    dst[i] = acc;
}
// thread per camera: pre-Q sums -> T (.) T^T / T (.), T = diag(Q, I) of the camera (see preq_finish)
__global__ void k_preq_finish(DevProblem P, double* H21, double* g6) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= P.Nc) return;
    double Q[9], H[21], g[6];
    const double* q = P.cam_Q + 12 * (size_t)c;
#pragma unroll
    for (int i = 0; i < 9; i++) Q[i] = q[i];
#pragma unroll
    for (int i = 0; i < 21; i++) H[i] = H21[21 * (size_t)c + i];
    if (g6) {
#pragma unroll
        for (int i = 0; i < 6; i++) g[i] = g6[6 * (size_t)c + i];
    }
    preq_finish(Q, H, g6 ? g : nullptr);
#pragma unroll
    for (int i = 0; i < 21; i++) H21[21 * (size_t)c + i] = H[i];
    if (g6) {
#pragma unroll
        for (int i = 0; i < 6; i++) g6[6 * (size_t)c + i] = g[i];
    }
}
void launch_preq_finish(const DevProblem& P, double* H21, double* g6, cudaStream_t st) {
    k_preq_finish<<<cdiv(P.Nc, 128), 128, 0, st>>>(P, H21, g6);
}
void launch_camera_blocks(const DevProblem& P, cudaStream_t st) {
    const char* e = std::getenv("SBA_K2B_PIPE");   // measurement knob
    // Venice shape, stage 9 (scripts/ab_tiles.py): 0.1104 ms with two slots per lane, 0.1134 plain, 0.1238 with the prefetch
    // (126 registers and the register moves of the rotation cost more than the overlap gives); a chunk-blocked plane layout
    // ([256 slots][plane]) read through the same kernel was no faster (0.1135): the plane stride is not what limits it
    const int pipe = e ? std::atoi(e) : 2;
    if (P.Nt) {
        if (pipe == 2 && P.Ns % 2 == 0) k_camera_blocks<2><<<cdiv(P.Nt, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, 0, st>>>(P);
        else if (pipe == 1) k_camera_blocks<1><<<cdiv(P.Nt, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, 0, st>>>(P);
        else k_camera_blocks<0><<<cdiv(P.Nt, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, 0, st>>>(P);
    }
    k_tile_reduce<<<cdiv(P.Nc * 21, 256), 256, 0, st>>>(P, 0, 21, P.Hcc, 1);
    k_tile_reduce<<<cdiv(P.Nc * 6, 256), 256, 0, st>>>(P, 21, 6, P.gc, 1);
    launch_preq_finish(P, P.Hcc, P.gc, st);
}

}  // namespace sba
