// sba_kernels_pcg.cu -- K3 block-Jacobi (SCHUR_JACOBI) preconditioned CG vector
// updates, K4 step application / model cost change / norms, output gathers and
// the batched DLT triangulation (K5).
//
// PCG follows the loop ceres::Solve runs for ITERATIVE_SCHUR behind
// modularsbasolver.cpp:938; in-tree analogue modules/g2o/g2o/solvers/pcg/linear_solver_pcg.hpp:84-165.
// The 6*Nc vectors are tiny (85 KB for 1778 cameras), so each update is ONE
// single-block kernel: every dot product is a fixed-shape tree, bitwise
// reproducible, and identical on every rank of a multi-GPU run.
#include <cooperative_groups.h>

#include <algorithm>

#include <cstdlib>

#include "sba_kernel_util.cuh"

namespace sba {

__device__ __forceinline__ double minv_row(const DevProblem& P, const double* r, int i) {
    const int c = i / 6, a = i - 6 * c;
    const double* M = P.Minv + 36 * (size_t)c + 6 * a;
    const double* rc = r + 6 * (size_t)c;
    double v = 0.0;
#pragma unroll
    for (int b = 0; b < 6; b++) v += M[b] * rc[b];
    return v;
}

// x = 0, r = rhs, z = Minv r, p = z, rz = r.z, rz0 = rz, iteration counter = 0
__global__ void __launch_bounds__(RED_BLOCK) k_pcg_init(DevProblem P) {
    __shared__ double sh[33];
    const int n = 6 * P.Nc;
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const double z = minv_row(P, P.rhs, i);
        const double r = P.rhs[i];
        P.pcg_x[i] = 0.0; P.pcg_r[i] = r; P.pcg_z[i] = z; P.pcg_p[i] = z;
        acc += r * z;
    }
    acc = block_sum(acc, sh);   // (contains __syncthreads: every p of this block is visible below)
    if (threadIdx.x == 0) {
        P.scalars[SC_RZ] = acc; P.scalars[SC_RZ_NEXT] = acc; P.scalars[SC_RZ0] = acc;
        P.flags[FL_PCG_ITERS] = 0; P.flags[FL_PCG_ITERS_NEXT] = 0;
    }
    for (int cam = threadIdx.x; cam < P.Nc; cam += blockDim.x) cam_vec_store(P, cam, P.pcg_p + 6 * (size_t)cam);   // operand of the first point pass
}
void launch_pcg_init(const DevProblem& P, cudaStream_t st) { k_pcg_init<<<1, RED_BLOCK, 0, st>>>(P); }

// ---------------------------------------------------------------------------
// One CG iteration after the camera pass, as three small multi-block kernels (a single block is
// bound by one SM's L2 bandwidth: 31 us for 1778 cameras).  192 threads = 32 cameras per block,
// thread per vector element.  Dot products: per-block partial -> every block folds the partial
// array with the same fixed tree, so alpha / beta are bitwise identical in all blocks and ranks.
//   A: q = -sum(tile partials) [+ Hcc p + lam p],  partial p.q        (replaces k_camera_reduce)
//   B: alpha = rz / p.q;  x += alpha p;  r -= alpha q;  z = Minv r;  partial r.z
//   C: beta = rz_new / rz;  p = z + beta p;  scalars for the next iteration
// B and C freeze (no state change) once rz <= tol^2 rz0 or the iteration cap is reached, so the host
// may launch iterations in batches and still stop at exactly the iteration a sequential loop would.
// Scalars are double-buffered (SC_RZ / SC_RZ_NEXT, FL_PCG_ITERS / FL_PCG_ITERS_NEXT): C writes the
// NEXT copies, A promotes them, so no block ever reads a scalar another block of the same launch writes.
// ---------------------------------------------------------------------------
constexpr int PCG_BLOCK = 192;

__device__ __forceinline__ double fold_partials(const double* part, int nb, double* sh) {
    double v = 0.0;
    for (int j = threadIdx.x; j < nb; j += blockDim.x) v += part[j];
    return block_sum(v, sh);
}
__device__ __forceinline__ bool pcg_frozen(const DevProblem& P, double tol2, int max_iters) {
    const double rz = P.scalars[SC_RZ], rz0 = P.scalars[SC_RZ0];
    return !(rz > tol2 * rz0) || !(rz0 > 0.0) || P.flags[FL_PCG_ITERS] >= max_iters;
}

// ---- peer-memory all-reduce of one vector slice per block (multi-GPU, PM_XCHG) ----------------------------------
// Every rank pushes its block's PCG_BLOCK partial values straight into the mailbox of every peer (plain stores to
// IPC-mapped memory: NVLink / NVSwitch writes), publishes a per-(source, block) flag with release semantics, waits for
// the same block of every peer, and sums the G slices in RANK ORDER -- the same order and the same operands on every
// rank, so q (hence alpha, beta and every later iterate) stays bitwise identical across ranks, as with NCCL's
// replicated result but without a kernel boundary and a collective launch per PCG iteration.  Buffers alternate with
// the parity of the exchange number: a rank can be at most one exchange ahead of a peer (it needs the peer's flag of
// exchange k+1, which the peer sets only after its kernel of exchange k has completed).
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t* xchg_flags(const DevProblem& P, int g) {
    return reinterpret_cast<uint32_t*>(P.xchg_peer[g] + 2 * (size_t)P.xchg_n * P.xchg_stride);
}
// every thread of the block calls it with its own value; i = global element index (< xchg_stride)
__device__ __forceinline__ double peer_allreduce(const DevProblem& P, double v, int i, uint32_t seq) {
    const int G = P.xchg_n, me = P.xchg_rank, nb = P.xchg_nflag;
    const int par = (int)(seq & 1u);
    const size_t stride = (size_t)P.xchg_stride;
    for (int g = 0; g < G; g++)
        if (g != me) __stcg(P.xchg_peer[g] + (size_t)(par * G + me) * stride + i, v);
    __threadfence_system();
    __syncthreads();
    if ((int)threadIdx.x < G && (int)threadIdx.x != me) {   // one thread per peer: publish, then wait for that peer
        const int g = threadIdx.x;
        st_release_sys(xchg_flags(P, g) + (size_t)(par * G + me) * nb + blockIdx.x, seq);
        const uint32_t* mine = xchg_flags(P, me) + (size_t)(par * G + g) * nb + blockIdx.x;
        const long long t0 = clock64();
        while (ld_acquire_sys(mine) != seq) {
            if (clock64() - t0 > 6000000000LL) { P.flags[FL_XCHG_TIMEOUT] = 1; break; }   // ~3 s: a peer died; the host reports it
        }
    }
    __syncthreads();
    double acc = 0.0;
    for (int g = 0; g < G; g++)
        acc += g == me ? v : __ldcg(P.xchg_peer[me] + (size_t)(par * G + g) * stride + i);
    return acc;
}

// local part of q_i = (S p)_i according to the mode (see PM_* in sba_device.h); block-uniform barriers inside
__device__ __forceinline__ double pcg_local_q(const DevProblem& P, int mode, int i, int n, int* s_off) {
    const int src = mode & PM_SRC_MASK;
    if (src == PM_SRC_FOLD) return explicit_fold(P, i, i < n, s_off);
    if (i >= n) return 0.0;
    if (src == PM_SRC_Q) return P.pcg_q[i];
    const int c = i / 6, a = i - 6 * c;
    double acc = 0.0;
    if (!P.cam_fixed[c]) acc = sum_partials(P.mv_part, P.cam_tile_off[c], P.cam_tile_off[c + 1], a);
    return -acc;
}
__device__ __forceinline__ double pcg_hcc_p(const DevProblem& P, int i) {   // ((Hcc + lam) p)_i
    const int c = i / 6, a = i - 6 * c;
    const double* p = P.pcg_p + 6 * (size_t)c;
    const double* H = P.Hcc + 21 * (size_t)c;
    double hx = 0.0;
#pragma unroll
    for (int b = 0; b < 6; b++) hx += H[sym6(a, b)] * p[b];
    return hx + P.lam_c[i] * p[a];
}

__global__ void __launch_bounds__(PCG_BLOCK) k_pcg_A(DevProblem P, int mode) {
    __shared__ double sh[33];
    __shared__ int s_off[PCG_BLOCK];
    const int i = blockIdx.x * blockDim.x + threadIdx.x, n = 6 * P.Nc;
    double q_i = pcg_local_q(P, mode, i, n, s_off);
    if (i == 0) { P.scalars[SC_RZ] = P.scalars[SC_RZ_NEXT]; P.flags[FL_PCG_ITERS] = P.flags[FL_PCG_ITERS_NEXT]; }
    double pq = 0.0;
    if (i < n) {
        if (mode & PM_HCC) q_i += pcg_hcc_p(P, i);
        P.pcg_q[i] = q_i;
        pq = P.pcg_p[i] * q_i;
    }
    pq = block_sum(pq, sh);
    if (threadIdx.x == 0) P.pcg_part[blockIdx.x] = pq;
}
__global__ void __launch_bounds__(PCG_BLOCK) k_pcg_B(DevProblem P, double tol2, int max_iters, int nb) {
    __shared__ double sh[33];
    __shared__ double sr[PCG_BLOCK];
    if (pcg_frozen(P, tol2, max_iters)) return;   // uniform over the grid
    const int i = blockIdx.x * blockDim.x + threadIdx.x, n = 6 * P.Nc;
    const double pq = fold_partials(P.pcg_part, nb, sh);
    const double alpha = P.scalars[SC_RZ] / pq;
    double r = 0.0;
    if (i < n) {
        P.pcg_x[i] += alpha * P.pcg_p[i];
        r = P.pcg_r[i] - alpha * P.pcg_q[i];
        P.pcg_r[i] = r;
    }
    sr[threadIdx.x] = r;
    __syncthreads();
    double rz = 0.0;
    if (i < n) {
        const int c = i / 6, a = i - 6 * c;
        const double* M = P.Minv + 36 * (size_t)c + 6 * a;
        const double* rc = sr + (threadIdx.x - a);
        double z = 0.0;
#pragma unroll
        for (int b = 0; b < 6; b++) z += M[b] * rc[b];
        P.pcg_z[i] = z;
        rz = r * z;
    }
    rz = block_sum(rz, sh);
    if (threadIdx.x == 0) P.pcg_part[nb + blockIdx.x] = rz;
}
__global__ void __launch_bounds__(PCG_BLOCK) k_pcg_C(DevProblem P, double tol2, int max_iters, int nb) {
    __shared__ double sh[33];
    __shared__ double sp[PCG_BLOCK];
    if (pcg_frozen(P, tol2, max_iters)) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x, n = 6 * P.Nc;
    const double rz = P.scalars[SC_RZ];
    const double rz_new = fold_partials(P.pcg_part + nb, nb, sh);
    const double beta = rz_new / rz;
    double pn = 0.0;
    if (i < n) { pn = P.pcg_z[i] + beta * P.pcg_p[i]; P.pcg_p[i] = pn; }
    sp[threadIdx.x] = pn;
    __syncthreads();
    if (i < n && i % 6 == 0) cam_vec_store(P, i / 6, sp + threadIdx.x);   // operand of the next point pass (PCG_BLOCK % 6 == 0)
    if (i == 0) { P.scalars[SC_RZ_NEXT] = rz_new; P.flags[FL_PCG_ITERS_NEXT] = P.flags[FL_PCG_ITERS] + 1; }
}
// A, B and C in ONE cooperative launch (two grid-wide barriers instead of two kernel boundaries): same arithmetic,
// same trees, same freeze rule; used when the device can keep all pcg_blocks co-resident.
__global__ void __launch_bounds__(PCG_BLOCK) k_pcg_ABC(DevProblem P, int mode, double tol2, int max_iters, int nb) {
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    __shared__ double sh[33];
    __shared__ double sr[PCG_BLOCK];
    __shared__ int s_off[PCG_BLOCK];
    const int i = blockIdx.x * blockDim.x + threadIdx.x, n = 6 * P.Nc;
    // ---- A (reads the scalars of the previous iteration through the NEXT copies: nobody writes them before the last barrier)
    const double rz = P.scalars[SC_RZ_NEXT], rz0 = P.scalars[SC_RZ0];
    const int it = P.flags[FL_PCG_ITERS_NEXT];
    const bool frozen = !(rz > tol2 * rz0) || !(rz0 > 0.0) || it >= max_iters;
    if (frozen) return;   // uniform over the grid AND over the ranks (replicated scalars): no barrier, no exchange
    if ((mode & PM_XCHG) && P.flags[FL_XCHG_DEAD]) return;   // an earlier launch timed out on a peer (set after its last barrier: uniform)
    const uint32_t seq = (mode & PM_XCHG) ? *P.xchg_ctr + 1u : 0u;   // (written only after the last barrier)
    double q_i = pcg_local_q(P, mode, i, n, s_off);
    if (mode & PM_XCHG) q_i = peer_allreduce(P, q_i, i, seq);
    double pq = 0.0, p_i = 0.0;
    if (i < n) {
        p_i = P.pcg_p[i];
        if (mode & PM_HCC) q_i += pcg_hcc_p(P, i);
        P.pcg_q[i] = q_i;
        pq = p_i * q_i;
    }
    pq = block_sum(pq, sh);
    if (threadIdx.x == 0) P.pcg_part[blockIdx.x] = pq;
    grid.sync();
    // ---- B
    const double alpha = rz / fold_partials(P.pcg_part, nb, sh);
    double r = 0.0;
    if (i < n) {
        P.pcg_x[i] += alpha * p_i;
        r = P.pcg_r[i] - alpha * q_i;
        P.pcg_r[i] = r;
    }
    sr[threadIdx.x] = r;
    __syncthreads();
    double rzp = 0.0, z = 0.0;
    if (i < n) {
        const int c = i / 6, a = i - 6 * c;
        const double* M = P.Minv + 36 * (size_t)c + 6 * a;
        const double* rc = sr + (threadIdx.x - a);
#pragma unroll
        for (int b = 0; b < 6; b++) z += M[b] * rc[b];
        P.pcg_z[i] = z;
        rzp = r * z;
    }
    rzp = block_sum(rzp, sh);
    if (threadIdx.x == 0) P.pcg_part[nb + blockIdx.x] = rzp;
    grid.sync();
    // ---- C
    const double rz_new = fold_partials(P.pcg_part + nb, nb, sh);
    const double beta = rz_new / rz;
    const double pn = z + beta * p_i;
    if (i < n) P.pcg_p[i] = pn;
    sr[threadIdx.x] = pn;   // (sr was last read before the block_sum barriers of phase B)
    __syncthreads();
    if (i < n && i % 6 == 0) cam_vec_store(P, i / 6, sr + threadIdx.x);   // operand of the next point pass (PCG_BLOCK % 6 == 0)
    grid.sync();   // every block has read SC_RZ_NEXT / FL_PCG_ITERS_NEXT before they are overwritten
    if (i == 0) {
        P.scalars[SC_RZ] = rz; P.flags[FL_PCG_ITERS] = it;
        P.scalars[SC_RZ_NEXT] = rz_new; P.flags[FL_PCG_ITERS_NEXT] = it + 1;
        if (mode & PM_XCHG) { *P.xchg_ctr = seq; if (P.flags[FL_XCHG_TIMEOUT]) P.flags[FL_XCHG_DEAD] = 1; }
    }
}

int pcg_blocks(const DevProblem& P) { return cdiv(6 * P.Nc, PCG_BLOCK); }
int pcg_xchg_nflag(const DevProblem& P) { return pcg_blocks(P); }
int pcg_xchg_stride(const DevProblem& P) { return pcg_blocks(P) * PCG_BLOCK; }
size_t pcg_xchg_ll_bytes(const DevProblem& P, int nranks) { return 2 * (size_t)nranks * pcg_solve_items(P.Nc) * pcg_solve_slot() * 16; }
size_t pcg_xchg_bytes(const DevProblem& P, int nranks) {
    return 2 * (size_t)nranks * pcg_xchg_stride(P) * sizeof(double) + 2 * (size_t)nranks * pcg_xchg_nflag(P) * sizeof(uint32_t);
}
static int pcg_coop_blocks() {
    static int ok_blocks = -1;
    if (ok_blocks < 0) {
        int dev = 0, coop = 0, sms = 0, per_sm = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_pcg_ABC, PCG_BLOCK, 0);
        ok_blocks = coop ? sms * per_sm : 0;
    }
    return ok_blocks;
}
bool pcg_fused_available(const DevProblem& P) { return pcg_blocks(P) <= pcg_coop_blocks(); }
// returns false when the cooperative path is unavailable (caller falls back to A + B + C; PM_XCHG needs the fused kernel)
bool launch_pcg_fused(const DevProblem& P, int mode, double tol2, int max_iters, cudaStream_t st) {
    int nb = pcg_blocks(P);
    if (nb > pcg_coop_blocks()) return false;
    DevProblem Pc = P;
    void* args[] = {&Pc, &mode, &tol2, &max_iters, &nb};
    return cudaLaunchCooperativeKernel((void*)k_pcg_ABC, dim3(nb), dim3(PCG_BLOCK), args, 0, st) == cudaSuccess;
}
void launch_pcg_A(const DevProblem& P, int mode, cudaStream_t st) { k_pcg_A<<<pcg_blocks(P), PCG_BLOCK, 0, st>>>(P, mode); }
void launch_pcg_BC(const DevProblem& P, double tol2, int max_iters, cudaStream_t st) {
    const int nb = pcg_blocks(P);
    k_pcg_B<<<nb, PCG_BLOCK, 0, st>>>(P, tol2, max_iters, nb);
    k_pcg_C<<<nb, PCG_BLOCK, 0, st>>>(P, tol2, max_iters, nb);
}

// ---------------------------------------------------------------------------
// K4: model cost change  -sum_o m.(res + m/2),  m = Jc dc + Jp dp
// (Ceres TrustRegionMinimizer model_cost_change).  Warp per tile, tile partial
// into tile_cost (free between the linear solve and the trial evaluation).
// ---------------------------------------------------------------------------
// PIPE: the point index two slots ahead and the dp gather one slot ahead are in flight during the arithmetic (same latency
// chain as K1: index -> gather).
template <bool PIPE>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32, PIPE ? 3 : 4) k_model_tiles(DevProblem P, const double* __restrict__ dc) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * WARPS_PER_BLOCK + warp;
    if (tile >= P.Nt) return;
    const int cam = __ldg(P.tile_cam + tile);
    const int start = __ldg(P.tile_start + tile), end = start + __ldg(P.tile_len + tile);
    const size_t Ns = (size_t)P.Ns;
    // m = Jc dc + Jp dp = A ((X - t) x a_c - x_t + dp) = A ((X - c0) x a_c - g_c + dp) with the per-camera operand (a_c, g_c) of dc
    // (cam_vec, written by launch_model_change), X - c0 = R Pbar + (t - c0): no C planes, no per-tile 3x3 products
    const double* cc = P.cam_cache + (size_t)CAM_CACHE_STRIDE * cam;
    const d4 r0 = ld256(cc), r1 = ld256(cc + 4), r2 = ld256(cc + 8);   // Rt (9), t (3)
    const d4 o0 = ld256(P.cam_vec + (size_t)CAM_VEC_STRIDE * cam), o1 = ld256(P.cam_vec + (size_t)CAM_VEC_STRIDE * cam + 4);
    const double ac0 = o0.x, ac1 = o0.y, ac2 = o0.z, gc0 = o0.w, gc1 = o1.x, gc2 = o1.y;
    const double tc0 = r2.y - P.center[0], tc1 = r2.z - P.center[1], tc2 = r2.w - P.center[2];
    double acc = 0.0;
    auto slot = [&](int s, const d4& dpv) {
        double m0 = 0.0, m1 = 0.0, A[6];
#pragma unroll
        for (int a = 0; a < 6; a++) A[a] = __ldcs(P.Ac + (size_t)a * Ns + s);
        const double p0 = __ldcs(P.Pc + s), p1 = __ldcs(P.Pc + Ns + s), p2 = __ldcs(P.Pc + 2 * Ns + s);
        const double d0 = r0.x * p0 + r0.w * p1 + r1.z * p2 + tc0;   // X - c0 = R Pbar + (t - c0), R = Rt^T
        const double d1 = r0.y * p0 + r1.x * p1 + r1.w * p2 + tc1;
        const double d2 = r0.z * p0 + r1.y * p1 + r2.x * p2 + tc2;
        const double v[3] = {d1 * ac2 - d2 * ac1 - gc0 + dpv.x, d2 * ac0 - d0 * ac2 - gc1 + dpv.y, d0 * ac1 - d1 * ac0 - gc2 + dpv.z};
#pragma unroll
        for (int a = 0; a < 3; a++) { m0 += A[a] * v[a]; m1 += A[3 + a] * v[a]; }
        if (P.sh_gd) { m0 += __ldcs(P.sh_gd + s); m1 += __ldcs(P.sh_gd + Ns + s); }   // + G_o delta of the shared blocks (sba_kernels_shared.cu)
        const double e0 = __ldcs(P.res + s), e1 = __ldcs(P.res + Ns + s);
        acc += m0 * (e0 + 0.5 * m0) + m1 * (e1 + 0.5 * m1);
    };
    if (!PIPE) {
        for (int s = start + lane; s < end; s += 32) {
            const int pt = __ldcs(P.obs_pt + s);
            if (pt < 0) continue;
            slot(s, ld256(P.dp + 4 * (size_t)pt));
        }
    } else {
        int s = start + lane;
        int pt1 = s < end ? __ldcs(P.obs_pt + s) : -1;
        int pt2 = s + 32 < end ? __ldcs(P.obs_pt + s + 32) : -1;
        d4 dp1 = ld256(P.dp + 4 * (size_t)max(pt1, 0));
        for (; s < end; s += 32) {
            const int pt = pt1;
            const d4 dpv = dp1;
            pt1 = pt2;
            pt2 = s + 64 < end ? __ldcs(P.obs_pt + s + 64) : -1;
            if (s + 32 < end) dp1 = ld256(P.dp + 4 * (size_t)max(pt1, 0));
            if (pt >= 0) slot(s, dpv);
        }
    }
    acc = warp_sum(acc);
    if (lane == 0) P.tile_cost[tile] = acc;
}

// per-block partials over points: [0] sum dp^2, [1] sum X^2 (free points), [2] max |gp|
__global__ void __launch_bounds__(256) k_point_stats(DevProblem P) {
    __shared__ double sh[33];
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    double s2 = 0.0, x2 = 0.0, gm = 0.0;
    if (p < P.Np && !P.pt_fixed[p]) {
        const double* d = P.dp + 4 * (size_t)p;
        const double* X = P.pts + 4 * (size_t)p;
        const double* g = P.gp + 4 * (size_t)p;
        s2 = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
        x2 = X[0] * X[0] + X[1] * X[1] + X[2] * X[2];
        gm = fmax(fabs(g[0]), fmax(fabs(g[1]), fabs(g[2])));
    }
    s2 = block_sum(s2, sh);
    x2 = block_sum(x2, sh);
    gm = block_max(gm, sh);
    if (threadIdx.x == 0) {
        P.block_part[4 * (size_t)blockIdx.x] = s2;
        P.block_part[4 * (size_t)blockIdx.x + 1] = x2;
        P.block_part[4 * (size_t)blockIdx.x + 2] = gm;
    }
}

// single block: folds the tile / block partials and the camera + prior parts into scalars.
//  SC_MODEL      = -(obs part + point-prior part)       (local to this rank)
//  SC_MODEL_PRIOR_CAM = -(camera-prior part)            (replicated)
//  SC_STEP2 / SC_X2   = point parts (local);  SC_STEP2_CAM / SC_X2_CAM = camera parts (replicated)
__global__ void __launch_bounds__(RED_BLOCK) k_model_reduce(DevProblem P, const double* __restrict__ dc) {
    __shared__ double sh[33];
    double acc = 0.0;
    for (int t = threadIdx.x; t < P.Nt; t += blockDim.x) acc += P.tile_cost[t];
    double mp = 0.0, mc = 0.0;
    for (int g = threadIdx.x; g < P.Ngroup; g += blockDim.x) {
        const int kind = P.pg_kind[g], idx = P.pg_index[g];
        if (kind == 0 ? P.pt_fixed[idx] : P.cam_fixed[idx]) continue;
        const double* d = kind == 0 ? P.dp + 4 * (size_t)idx : (kind == 1 ? dc + 6 * (size_t)idx : dc + 6 * (size_t)idx + 3);
        double v = 0.0;
        for (int i = P.pg_off[g]; i < P.pg_off[g + 1]; i++) {
            const double* A = P.prior_A + 9 * (size_t)i;
#pragma unroll
            for (int k = 0; k < 3; k++) {
                const double m = A[3 * k] * d[0] + A[3 * k + 1] * d[1] + A[3 * k + 2] * d[2];
                v += m * (P.prior_res[3 * (size_t)i + k] + 0.5 * m);
            }
        }
        if (kind == 0) mp += v; else mc += v;
    }
    double s2 = 0.0, x2 = 0.0;
    for (int b = threadIdx.x; b < P.NB; b += blockDim.x) { s2 += P.block_part[4 * (size_t)b]; x2 += P.block_part[4 * (size_t)b + 1]; }
    double cs2 = 0.0, cx2 = 0.0;
    for (int c = threadIdx.x; c < P.Nc; c += blockDim.x) {
        if (P.cam_fixed[c]) continue;
#pragma unroll
        for (int a = 0; a < 6; a++) cs2 += dc[6 * (size_t)c + a] * dc[6 * (size_t)c + a];
#pragma unroll
        for (int a = 0; a < 3; a++) cx2 += P.cam_r[3 * (size_t)c + a] * P.cam_r[3 * (size_t)c + a] + P.cam_t[3 * (size_t)c + a] * P.cam_t[3 * (size_t)c + a];
    }
    acc = block_sum(acc, sh); mp = block_sum(mp, sh); mc = block_sum(mc, sh);
    s2 = block_sum(s2, sh); x2 = block_sum(x2, sh); cs2 = block_sum(cs2, sh); cx2 = block_sum(cx2, sh);
    if (threadIdx.x == 0) {
        P.scalars[SC_MODEL] = -(acc + mp);
        P.scalars[SC_MODEL_PRIOR_CAM] = -mc;
        P.scalars[SC_STEP2] = s2; P.scalars[SC_X2] = x2;
        P.scalars[SC_STEP2_CAM] = cs2; P.scalars[SC_X2_CAM] = cx2;
    }
}
void launch_model_change(const DevProblem& P, const double* dc, cudaStream_t st) {
    launch_cam_vec(P, dc, st);   // per-camera operand (a_c, g_c) of dc for k_model_tiles
    if (P.Nt) {
        const char* e = std::getenv("SBA_MODEL_PIPE");   // measurement knob
        if (e ? std::atoi(e) != 0 : true) k_model_tiles<true><<<cdiv(P.Nt, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, 0, st>>>(P, dc);
        else k_model_tiles<false><<<cdiv(P.Nt, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, 0, st>>>(P, dc);
    }
    if (P.NB) k_point_stats<<<P.NB, 256, 0, st>>>(P);
    k_model_reduce<<<1, RED_BLOCK, 0, st>>>(P, dc);
}

// max-norm of the gradient over free blocks: SC_GMAX (points, local), SC_GMAX_CAM (replicated)
__global__ void __launch_bounds__(RED_BLOCK) k_grad_max(DevProblem P) {
    __shared__ double sh[33];
    double gm = 0.0, gcm = 0.0;
    for (int b = threadIdx.x; b < P.NB; b += blockDim.x) gm = fmax(gm, P.block_part[4 * (size_t)b + 2]);
    for (int i = threadIdx.x; i < 6 * P.Nc; i += blockDim.x)
        if (!P.cam_fixed[i / 6]) gcm = fmax(gcm, fabs(P.gc[i]));
    gm = block_max(gm, sh); gcm = block_max(gcm, sh);
    if (threadIdx.x == 0) { P.scalars[SC_GMAX] = gm; P.scalars[SC_GMAX_CAM] = gcm; }
}
void launch_grad_max(const DevProblem& P, cudaStream_t st) {
    if (P.NB) k_point_stats<<<P.NB, 256, 0, st>>>(P);
    k_grad_max<<<1, RED_BLOCK, 0, st>>>(P);
}

}  // namespace sba
