// sba_kernels_shared.cu -- SHARED parameter blocks in the reduced camera system:
//   * the five intrinsic blocks f[1], pp[2], k[3], t[2], B[2] of a camera model, free unless the Camera is fixed or
//     FixedParameter::CameraInternal is set (PinholeCamProjModule::init, sbamodules/pinholecameraprojectormodule.cpp:302-533;
//     the 8-block cost AutoDiffCostFunction<UVCost, 2, 3,3,3, 1,2,3,2,2> at :105-122);
//   * the boresight / lever-arm blocks of a mounting (LeverArmCost / PoseTransformLeverArmCost, blocks 2,3,3,3,3,3,1,2,3,2,2,
//     :60-99,147-166; costfunctors/posedecoratorfunctors.h:148-264).
// Ceres keeps such blocks in the reduced camera system next to the poses (they are not e-blocks).  Here a GROUP of shared
// columns (8 per free model: f ppx ppy k1 k2 k3 t1 t2 -- d res / d B == 0, pinholecameraprojectormodule.h:46-47 --, 6 per
// free mounting) becomes one or two PSEUDO-CAMERAS of 6 columns behind the real cameras, so the explicit tiles, the
// block-Jacobi preconditioner and the persistent PCG kernel run unchanged on Nce = Nc + pseudo-cameras.  With
//   G_o = d res_o / d (group columns)   (2 x 8, closed forms obs_intr_jac_pbar / obs_lever_jac_pbar, sba_math.cuh)
// the bordered Schur complement needs, per group g,
//   W_g[p]  = sum_{o in p} A_o^T G_o                      (3 x 8 per point; point order, segmented warp reduction)
//   H_gg'   = sum_o G_g^T G_g',  g_g = sum_o G_g^T res    (fixed-shape block trees)
//   Y_g[p]  = (Hpp + D)^-1 W_g[p]
//   S_cg    = sum_{o in c} Jc_o^T (G_o [c in g] - A_o Y_g[p(o)])          (6 x 8 per camera; camera order, warp per tile)
//   S_gg'   = H_gg' - sum_p W_g^T Y_g',  rhs_g = -(g_g - sum_p W_g^T Hinv gp)
// and the back-substitution / model change get the extra terms -Y_g delta_g and G_o delta_g.
// Every sum is a fixed tree (no floating-point atomics): bitwise reproducible.
#include <algorithm>

#include "sba_kernel_util.cuh"

namespace sba {

constexpr int SHW = 8;        // internal columns per group
constexpr int SH_RED = 72;    // doubles per block partial of the group reductions

// segmented warp reduction keyed on the point index (same scheme as sba_kernels_points.cu)
__device__ __forceinline__ int sh_seg_setup(int key, int lane, bool* head) {
    const int prev = __shfl_up_sync(FULL, key, 1);
    *head = lane == 0 || prev != key;
    const unsigned heads = __ballot_sync(FULL, *head);
    const unsigned above = lane == 31 ? 0u : heads >> (lane + 1);
    return above ? __ffs(above) : 32 - lane;
}

// G (2 x 8 row-major) of one observation for a group of kind `kind` (0 camera model, 1 mounting); dead columns zeroed
__device__ __forceinline__ void shared_jac(const DevProblem& P, int g, int kind, int cam, double Px, double Py, double Pz, double u, double v,
                                           const double* W, double* G) {
    const double* in = P.intr + 10 * (size_t)__ldg(P.cam_model + cam);
    const int sph = P.cam_kind[cam];
#pragma unroll
    for (int i = 0; i < 16; i++) G[i] = 0.0;
    if (Px == 0.0 && Py == 0.0 && Pz == 0.0) return;   // failed row (K1 marks it with Pbar = 0)
    if (kind == 0) {
        double J[20];
        if (!obs_intr_jac_pbar(Px, Py, Pz, u, v, in, W, sph, J)) return;
#pragma unroll
        for (int c = 0; c < 8; c++) { G[c] = J[c]; G[8 + c] = J[10 + c]; }
    } else {
        double d[3], J[12];
        dir_from_uv(u, v, in, d);
        if (!obs_lever_jac_pbar(Px, Py, Pz, d, P.mount_cache + (size_t)CAM_CACHE_STRIDE * P.cam_mount[cam], W, sph, J)) return;
#pragma unroll
        for (int c = 0; c < 6; c++) { G[c] = J[c]; G[8 + c] = J[6 + c]; }
    }
    const uint8_t* live = P.grp_live + 8 * g;
#pragma unroll
    for (int c = 0; c < 8; c++) if (!live[c]) { G[c] = 0.0; G[8 + c] = 0.0; }
}

// deterministic block sum of N values per thread -> part[blockIdx][0..N) ; 256 threads
template <int N>
__device__ __forceinline__ void block_reduce_store(double (&v)[N], double* part, double (*sm)[SH_RED]) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < N; i++) { const double s = warp_sum(v[i]); if (lane == 0) sm[w][i] = s; }
    __syncthreads();
    if ((int)threadIdx.x < N) {
        double a = 0.0;
        for (int k = 0; k < (int)(blockDim.x >> 5); k++) a += sm[k][threadIdx.x];   // warp order: fixed
        part[(size_t)blockIdx.x * SH_RED + threadIdx.x] = a;
    }
}

// ---------------------------------------------------------------------------
// per linearisation, point order: W_g per point, partial sums of G^T G (36 packed) and G^T res (8)
// ---------------------------------------------------------------------------
// Persistent grid (round 2): a thread walks its slots with a grid stride and keeps the 44 sums of G^T G | G^T res in registers
// across them; the block tree runs ONCE per block at the end (it was once per 256 slots: 220 double shuffles per warp and 21 000
// block partials for a single-block fold on the Venice shape).  Fixed slot -> thread assignment: deterministic.
__global__ void __launch_bounds__(256) k_shared_points(DevProblem P, int g, int kind) {
    __shared__ double sm[8][SH_RED];
    const int lane = threadIdx.x & 31;
    double hg[44];
#pragma unroll
    for (int i = 0; i < 44; i++) hg[i] = 0.0;
    // Nq % 32 == 0 and the stride is a multiple of 32: the lanes of a warp stay together (the shuffles below need that)
    for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < P.Nq; q += gridDim.x * blockDim.x) {
    const int pt = __ldcs(P.q_pt + q);
    double v[24];
#pragma unroll
    for (int i = 0; i < 24; i++) v[i] = 0.0;
    if (pt >= 0) {
        const int cam = __ldcs(P.q_cam + q);
        if (__ldg(P.cam_grp + 2 * cam + kind) == g) {
            const d4 Xp = ld256(P.pts + 4 * (size_t)pt);
            const double X[3] = {Xp.x, Xp.y, Xp.z};
            const double* cc = P.cam_cache + (size_t)CAM_CACHE_STRIDE * cam;
            double ccr[12];
            {
                const d4 c0 = ld256(cc), c1 = ld256(cc + 4), c2 = ld256(cc + 8);
                ccr[0] = c0.x; ccr[1] = c0.y; ccr[2] = c0.z; ccr[3] = c0.w; ccr[4] = c1.x; ccr[5] = c1.y; ccr[6] = c1.z; ccr[7] = c1.w;
                ccr[8] = c2.x; ccr[9] = c2.y; ccr[10] = c2.z; ccr[11] = c2.w;
            }
            const double2 uv = ldcs2(P.q_uv + 2 * (size_t)q);
            double W[4];
            const double* Wp = nullptr;
            if (P.q_info) {
                const double2 w01 = ldcs2(P.q_info + 4 * (size_t)q), w23 = ldcs2(P.q_info + 4 * (size_t)q + 2);
                W[0] = w01.x; W[1] = w01.y; W[2] = w23.x; W[3] = w23.y;
                Wp = W;
            }
            double d[3], r[2] = {0.0, 0.0};
            dir_from_uv(uv.x, uv.y, P.intr + 10 * (size_t)__ldg(P.cam_model + cam), d);
            const bool ok = obs_eval<false>(ccr, X, d, Wp, P.cam_kind[cam], r, nullptr, nullptr);
            if (ok) {
                const double vx = X[0] - ccr[9], vy = X[1] - ccr[10], vz = X[2] - ccr[11];
                const double Px = ccr[0] * vx + ccr[1] * vy + ccr[2] * vz, Py = ccr[3] * vx + ccr[4] * vy + ccr[5] * vz,
                             Pz = ccr[6] * vx + ccr[7] * vy + ccr[8] * vz;
                double G[16], A[6];
                shared_jac(P, g, kind, cam, Px, Py, Pz, uv.x, uv.y, Wp, G);
                const size_t Nq = (size_t)P.Nq;
#pragma unroll
                for (int i = 0; i < 6; i++) A[i] = __ldcs(P.qA + (size_t)i * Nq + q);
#pragma unroll
                for (int a = 0; a < 3; a++)
#pragma unroll
                    for (int c = 0; c < 8; c++) v[8 * a + c] = A[a] * G[c] + A[3 + a] * G[8 + c];
                int k = 0;
#pragma unroll
                for (int c1 = 0; c1 < 8; c1++)
#pragma unroll
                    for (int c2 = c1; c2 < 8; c2++) hg[k++] += G[c1] * G[c2] + G[8 + c1] * G[8 + c2];
#pragma unroll
                for (int c = 0; c < 8; c++) hg[36 + c] += G[c] * r[0] + G[8 + c] * r[1];
            }
        }
    }
    // W: segmented reduction over the point's slots
    bool head;
    const int rem = sh_seg_setup(pt, lane, &head);
    const int longest = __reduce_max_sync(FULL, rem);
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        if (o >= longest) break;
        const bool take = o < rem;
#pragma unroll
        for (int i = 0; i < 24; i++) { const double t = __shfl_down_sync(FULL, v[i], o); if (take) v[i] += t; }
    }
    if (pt >= 0 && head) {
        const int h = __ldg(P.pt_heavy + pt);
        double* o = h < 0 ? P.sh_W + ((size_t)g * P.Np + pt) * 24
                          : P.sh_heavy + 24 * (size_t)(__ldg(P.heavy_w0 + h) + ((q - __ldg(P.pt_q0 + pt)) >> 5));
#pragma unroll
        for (int i = 0; i < 24; i += 4) st256(o + i, v[i], v[i + 1], v[i + 2], v[i + 3]);
    }
    }   // slots of this thread
    block_reduce_store<44>(hg, P.sh_part, sm);
}
__global__ void k_shared_points_heavy(DevProblem P, int g) {
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= P.Nh) return;
    double v[24];
#pragma unroll
    for (int i = 0; i < 24; i++) v[i] = 0.0;
    for (int w = P.heavy_w0[h]; w < P.heavy_w0[h + 1]; w++)
#pragma unroll
        for (int i = 0; i < 24; i++) v[i] += P.sh_heavy[24 * (size_t)w + i];
    double* o = P.sh_W + ((size_t)g * P.Np + P.heavy_pt[h]) * 24;
#pragma unroll
    for (int i = 0; i < 24; i++) o[i] = v[i];
}
// cross term of a camera-model group and a mounting group on the SAME observation: sum_o G_m^T G_l (8 x 8)
__global__ void __launch_bounds__(256) k_shared_cross(DevProblem P, int gm, int gl) {
    __shared__ double sm[8][SH_RED];
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    const int pt = q < P.Nq ? __ldcs(P.q_pt + q) : -1;
    double x[64];
#pragma unroll
    for (int i = 0; i < 64; i++) x[i] = 0.0;
    if (pt >= 0) {
        const int cam = __ldcs(P.q_cam + q);
        if (__ldg(P.cam_grp + 2 * cam) == gm && __ldg(P.cam_grp + 2 * cam + 1) == gl) {
            const d4 Xp = ld256(P.pts + 4 * (size_t)pt);
            const double* cc = P.cam_cache + (size_t)CAM_CACHE_STRIDE * cam;
            const double vx = Xp.x - cc[9], vy = Xp.y - cc[10], vz = Xp.z - cc[11];
            const double Px = cc[0] * vx + cc[1] * vy + cc[2] * vz, Py = cc[3] * vx + cc[4] * vy + cc[5] * vz, Pz = cc[6] * vx + cc[7] * vy + cc[8] * vz;
            const double2 uv = ldcs2(P.q_uv + 2 * (size_t)q);
            double W[4];
            const double* Wp = nullptr;
            if (P.q_info) { for (int i = 0; i < 4; i++) W[i] = P.q_info[4 * (size_t)q + i]; Wp = W; }
            const bool ok = P.cam_kind[cam] || Pz >= 0.0;
            if (ok) {
                double Gm[16], Gl[16];
                shared_jac(P, gm, 0, cam, Px, Py, Pz, uv.x, uv.y, Wp, Gm);
                shared_jac(P, gl, 1, cam, Px, Py, Pz, uv.x, uv.y, Wp, Gl);
#pragma unroll
                for (int a = 0; a < 8; a++)
#pragma unroll
                    for (int b = 0; b < 8; b++) x[8 * a + b] = Gm[a] * Gl[b] + Gm[8 + a] * Gl[8 + b];
            }
        }
    }
    block_reduce_store<64>(x, P.sh_part, sm);
}
// single block: folds n block partials of `cnt` values each (fixed tree) and hands value k to `put(k, value)`
template <class Put>
__device__ __forceinline__ void fold_block_partials(const double* part, int n, int cnt, Put put) {
    __shared__ double sh[33];
    for (int k = 0; k < cnt; k++) {
        double a = 0.0;
        for (int b = threadIdx.x; b < n; b += blockDim.x) a += part[(size_t)b * SH_RED + k];
        a = block_sum(a, sh);
        if (threadIdx.x == 0) put(k, a);
        __syncthreads();
    }
}
__global__ void __launch_bounds__(RED_BLOCK) k_shared_fold_hg(DevProblem P, int g, int nblk) {
    double* H = P.sh_H + ((size_t)g * P.Ng + g) * 64;
    double* gg = P.sh_g + 8 * (size_t)g;
    fold_block_partials(P.sh_part, nblk, 44, [&](int k, double a) {
        if (k >= 36) { gg[k - 36] = a; return; }
        int c1 = 0, rest = k;
        while (rest >= 8 - c1) { rest -= 8 - c1; c1++; }
        const int c2 = c1 + rest;
        H[8 * c1 + c2] = a; H[8 * c2 + c1] = a;
    });
}
__global__ void __launch_bounds__(RED_BLOCK) k_shared_fold_cross(DevProblem P, int gm, int gl, int nblk) {
    double* A = P.sh_H + ((size_t)gm * P.Ng + gl) * 64;
    double* B = P.sh_H + ((size_t)gl * P.Ng + gm) * 64;
    fold_block_partials(P.sh_part, nblk, 64, [&](int k, double a) { A[k] = a; B[8 * (k & 7) + (k >> 3)] = a; });
}
void launch_shared_linearize(const DevProblem& P, cudaStream_t st) {
    if (!P.Ng) return;
    cudaMemsetAsync(P.sh_H, 0, ((size_t)P.Ng * P.Ng * 64 + 8 * (size_t)P.Ng) * sizeof(double), st);   // sh_g follows sh_H
    if (!P.Nq) return;
    const int nb = cdiv(P.Nq, 256);
    for (int g = 0; g < P.Ng; g++) {
        const int nbp = std::min(nb, 148 * 4);   // persistent: a few blocks per SM, every thread walks its slots
        k_shared_points<<<nbp, 256, 0, st>>>(P, g, P.grp_kind[g]);
        if (P.Nh) k_shared_points_heavy<<<cdiv(P.Nh, 128), 128, 0, st>>>(P, g);
        k_shared_fold_hg<<<1, RED_BLOCK, 0, st>>>(P, g, nbp);
    }
    if (P.cam_mount) {
        for (int gm = 0; gm < P.Ng; gm++)
            for (int gl = 0; gl < P.Ng; gl++) {
                if (P.grp_kind[gm] != 0 || P.grp_kind[gl] != 1) continue;
                k_shared_cross<<<nb, 256, 0, st>>>(P, gm, gl);
                k_shared_fold_cross<<<1, RED_BLOCK, 0, st>>>(P, gm, gl, nb);
            }
    }
}

// column c (0..7) of group g -> (pseudo-camera, column inside it): a model spreads over two pseudo-cameras (6 + 2), a mounting has one
__device__ __forceinline__ int grp_col_index(const DevProblem& P, int g, int c) { return 6 * P.grp_blk[g] + c; }
__device__ __forceinline__ int grp_nblk(const DevProblem& P, int g) { return P.grp_kind[g] == 0 ? 2 : 1; }

// priors on the intrinsic blocks: residuals r = A (x - b) (dim x dim), cost into the replicated scalar
__global__ void __launch_bounds__(128) k_shared_prior_eval(DevProblem P, int store) {
    __shared__ double sh[33];
    double acc = 0.0;
    for (int i = threadIdx.x; i < P.Nipr; i += blockDim.x) {
        const int blk = P.ipr_block[i];
        const int first = blk == 0 ? 0 : (blk == 1 ? 1 : (blk == 2 ? 3 : (blk == 3 ? 6 : 8))), dim = blk == 0 ? 1 : (blk == 2 ? 3 : 2);
        const double* x = P.intr + 10 * (size_t)P.ipr_model[i] + first;
        const double* A = P.ipr_A + 9 * (size_t)i;
        const double* b = P.ipr_b + 3 * (size_t)i;
        for (int k = 0; k < 3; k++) {
            double r = 0.0;
            if (k < dim) for (int j = 0; j < dim; j++) r += A[dim * k + j] * (x[j] - b[j]);
            if (store) P.ipr_res[3 * (size_t)i + k] = r;
            acc += r * r;
        }
    }
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) P.scalars[SC_COST_PRIOR_CAM] += acc;   // after k_prior_eval on the same stream
}
void launch_shared_prior_eval(const DevProblem& P, int store, cudaStream_t st) {
    if (P.Nipr) k_shared_prior_eval<<<1, 128, 0, st>>>(P, store);
}

// after the all-reduce of sh_H | sh_g: priors on the intrinsic blocks, then Hcc / gc of the pseudo-cameras (dead columns: unit diagonal)
__global__ void k_shared_lin_finish(DevProblem P) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        for (int i = 0; i < P.Nipr; i++) {
            const int blk = P.ipr_block[i], m = P.ipr_model[i];
            if (blk == 4) continue;
            int g = -1;
            for (int k = 0; k < P.Ng; k++) if (P.grp_kind[k] == 0 && P.grp_index[k] == m) g = k;
            if (g < 0) continue;
            const int first = blk == 0 ? 0 : (blk == 1 ? 1 : (blk == 2 ? 3 : 6)), dim = blk == 0 ? 1 : (blk == 2 ? 3 : 2);
            if (!P.grp_live[8 * g + first]) continue;
            const double* A = P.ipr_A + 9 * (size_t)i;
            double* H = P.sh_H + ((size_t)g * P.Ng + g) * 64;
            for (int a = 0; a < dim; a++) {
                for (int r = 0; r < dim; r++) P.sh_g[8 * g + first + a] += A[dim * r + a] * P.ipr_res[3 * (size_t)i + r];
                for (int b = 0; b < dim; b++) for (int r = 0; r < dim; r++) H[8 * (first + a) + first + b] += A[dim * r + a] * A[dim * r + b];
            }
        }
    }
    __syncthreads();
    if (blockIdx.x != 0) return;
    for (int g = threadIdx.x; g < P.Ng; g += blockDim.x) {
        const double* H = P.sh_H + ((size_t)g * P.Ng + g) * 64;
        const int nb = grp_nblk(P, g);
        for (int b = 0; b < nb; b++) {
            const int blk = P.grp_blk[g] + b;
            for (int i = 0; i < 6; i++) {
                const int ci = 6 * b + i;
                const bool li = ci < 8 && P.grp_live[8 * g + ci];
                P.gc[6 * (size_t)blk + i] = li ? P.sh_g[8 * g + ci] : 0.0;
                for (int j = i; j < 6; j++) {
                    const int cj = 6 * b + j;
                    const bool lj = cj < 8 && P.grp_live[8 * g + cj];
                    P.Hcc[21 * (size_t)blk + sym6(i, j)] = (li && lj) ? H[8 * ci + cj] : (i == j ? 1.0 : 0.0);
                }
            }
        }
    }
}
void launch_shared_lin_finish(const DevProblem& P, cudaStream_t st) {
    if (P.Ng) k_shared_lin_finish<<<1, 128, 0, st>>>(P);
}

// ---------------------------------------------------------------------------
// per LM iteration
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_shared_Y(DevProblem P) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;   // (group, point)
    if (i >= (size_t)P.Ng * P.Np) return;
    const int p = (int)(i % (size_t)P.Np);
    const double* Hp = P.Hinv + 8 * (size_t)p;
    const d4 h0 = ld256(Hp), h1 = ld256(Hp + 4);   // (h00 h01 h02 h11) (h12 h22 - -)
    const double* W = P.sh_W + 24 * i;
    double* Y = P.sh_Y + 24 * i;
#pragma unroll
    for (int c = 0; c < 8; c += 4) {
        const d4 w0 = ld256(W + c), w1 = ld256(W + 8 + c), w2 = ld256(W + 16 + c);
        st256(Y + c, h0.x * w0.x + h0.y * w1.x + h0.z * w2.x, h0.x * w0.y + h0.y * w1.y + h0.z * w2.y,
              h0.x * w0.z + h0.y * w1.z + h0.z * w2.z, h0.x * w0.w + h0.y * w1.w + h0.z * w2.w);
        st256(Y + 8 + c, h0.y * w0.x + h0.w * w1.x + h1.x * w2.x, h0.y * w0.y + h0.w * w1.y + h1.x * w2.y,
              h0.y * w0.z + h0.w * w1.z + h1.x * w2.z, h0.y * w0.w + h0.w * w1.w + h1.x * w2.w);
        st256(Y + 16 + c, h0.z * w0.x + h1.x * w1.x + h1.y * w2.x, h0.z * w0.y + h1.x * w1.y + h1.y * w2.y,
              h0.z * w0.z + h1.x * w1.z + h1.y * w2.z, h0.z * w0.w + h1.x * w1.w + h1.y * w2.w);
    }
}
// sum_p W_g1^T Y_g2 (64) [+ sum_p W_g1^T (Hinv gp) (8) when g1 == g2]: persistent blocks, fixed trees
__global__ void __launch_bounds__(256) k_shared_schur(DevProblem P, int g1, int g2) {
    __shared__ double sm[8][SH_RED];
    double acc[72];
#pragma unroll
    for (int i = 0; i < 72; i++) acc[i] = 0.0;
    const double* W1 = P.sh_W + (size_t)g1 * P.Np * 24;
    const double* Y2 = P.sh_Y + (size_t)g2 * P.Np * 24;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < P.Np; p += gridDim.x * blockDim.x) {
        double w[24], y[24];
#pragma unroll
        for (int i = 0; i < 24; i += 4) {
            const d4 a = ld256(W1 + 24 * (size_t)p + i), b = ld256(Y2 + 24 * (size_t)p + i);
            w[i] = a.x; w[i + 1] = a.y; w[i + 2] = a.z; w[i + 3] = a.w; y[i] = b.x; y[i + 1] = b.y; y[i + 2] = b.z; y[i + 3] = b.w;
        }
#pragma unroll
        for (int c1 = 0; c1 < 8; c1++)
#pragma unroll
            for (int c2 = 0; c2 < 8; c2++) acc[8 * c1 + c2] += w[c1] * y[c2] + w[8 + c1] * y[8 + c2] + w[16 + c1] * y[16 + c2];
        if (g1 == g2) {
            const d4 yh = ld256(P.y_p + 4 * (size_t)p);   // Hinv gp (written by the rhs pass of this LM iteration)
#pragma unroll
            for (int c = 0; c < 8; c++) acc[64 + c] += w[c] * yh.x + w[8 + c] * yh.y + w[16 + c] * yh.z;
        }
    }
    block_reduce_store<72>(acc, P.sh_part, sm);
}
__global__ void __launch_bounds__(RED_BLOCK) k_shared_fold_schur(DevProblem P, int g1, int g2, int nblk) {
    double* acc = P.sh_acc + (size_t)P.Ng * P.Nc * 48;
    double* A = acc + ((size_t)g1 * P.Ng + g2) * 64;
    double* B = acc + ((size_t)g2 * P.Ng + g1) * 64;
    double* rS = acc + (size_t)P.Ng * P.Ng * 64 + 8 * (size_t)g1;
    fold_block_partials(P.sh_part, nblk, g1 == g2 ? 72 : 64, [&](int k, double a) {
        if (k >= 64) { rS[k - 64] = a; return; }
        A[k] = a;
        if (g1 != g2) B[8 * (k & 7) + (k >> 3)] = a;
    });
}
// border, camera order: warp per tile, 48 accumulators J'^T E with the pre-Q rows J' = [A x d | -A] and
// E = G [camera in g] - A Y_g[p]; tile partials in sh_tile (48 per tile)
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32) k_shared_border(DevProblem P, int g, int kind, double* __restrict__ tile_out) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * WARPS_PER_BLOCK + warp;
    if (tile >= P.Nt) return;
    const int cam = __ldg(P.tile_cam + tile);
    const int start = __ldg(P.tile_start + tile), len = __ldg(P.tile_len + tile);
    const size_t Ns = (size_t)P.Ns;
    const bool own = __ldg(P.cam_grp + 2 * cam + kind) == g;
    double Rt[12];
    load_rt(P, cam, Rt);
    const double* Yg = P.sh_Y + (size_t)g * P.Np * 24;
    double acc[48];
#pragma unroll
    for (int i = 0; i < 48; i++) acc[i] = 0.0;
    for (int s = start + lane; s < start + len; s += 32) {
        const int pt = __ldcs(P.obs_pt + s);
        if (pt < 0) continue;
        double A[6], w[6];
#pragma unroll
        for (int a = 0; a < 6; a++) A[a] = __ldcs(P.Ac + (size_t)a * Ns + s);
        const double p0 = __ldcs(P.Pc + s), p1 = __ldcs(P.Pc + Ns + s), p2 = __ldcs(P.Pc + 2 * Ns + s);
        preq_rows(A, p0, p1, p2, Rt, w);
        double E[16];
        if (own) {
            const double2 uv = ldcs2(P.obs_uv + 2 * (size_t)s);
            double W[4];
            const double* Wp = nullptr;
            if (P.obs_info) { for (int i = 0; i < 4; i++) W[i] = __ldcs(P.obs_info + i * Ns + s); Wp = W; }
            shared_jac(P, g, kind, cam, p0, p1, p2, uv.x, uv.y, Wp, E);
        } else {
#pragma unroll
            for (int i = 0; i < 16; i++) E[i] = 0.0;
        }
        const double* Y = Yg + 24 * (size_t)pt;
#pragma unroll
        for (int c = 0; c < 8; c += 4) {
            const d4 y0 = ld256(Y + c), y1 = ld256(Y + 8 + c), y2 = ld256(Y + 16 + c);
            const double ya[4] = {y0.x, y0.y, y0.z, y0.w}, yb[4] = {y1.x, y1.y, y1.z, y1.w}, yc[4] = {y2.x, y2.y, y2.z, y2.w};
#pragma unroll
            for (int k = 0; k < 4; k++) {
                E[c + k] -= A[0] * ya[k] + A[1] * yb[k] + A[2] * yc[k];
                E[8 + c + k] -= A[3] * ya[k] + A[4] * yb[k] + A[5] * yc[k];
            }
        }
#pragma unroll
        for (int a = 0; a < 3; a++)
#pragma unroll
            for (int c = 0; c < 8; c++) {
                acc[8 * a + c] += w[a] * E[c] + w[3 + a] * E[8 + c];
                acc[8 * (3 + a) + c] -= A[a] * E[c] + A[3 + a] * E[8 + c];
            }
    }
    double* out = tile_out + 48 * (size_t)tile;
#pragma unroll
    for (int i = 0; i < 48; i++) { const double v = warp_sum(acc[i]); if (lane == 0) out[i] = v; }
}
// thread per (camera, column): sums the camera's tiles in tile order, applies T = diag(Q, I) to the six rows
__global__ void k_shared_border_reduce(DevProblem P, int g, const double* __restrict__ tile_in) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.Nc * 8) return;
    const int cam = i >> 3, c = i & 7;
    double v[6] = {0, 0, 0, 0, 0, 0};
    if (!P.cam_fixed[cam]) {
        for (int t = P.cam_tile_off[cam]; t < P.cam_tile_off[cam + 1]; t++)
#pragma unroll
            for (int a = 0; a < 6; a++) v[a] += tile_in[48 * (size_t)t + 8 * a + c];
        const double* Q = P.cam_Q + 12 * (size_t)cam;
        const double v0 = v[0], v1 = v[1], v2 = v[2];
        v[0] = Q[0] * v0 + Q[1] * v1 + Q[2] * v2; v[1] = Q[3] * v0 + Q[4] * v1 + Q[5] * v2; v[2] = Q[6] * v0 + Q[7] * v1 + Q[8] * v2;
    }
    double* out = P.sh_acc + ((size_t)g * P.Nc + cam) * 48;
    const double* pb = P.pair_border ? P.pair_border + ((size_t)g * P.Nc + cam) * 48 : nullptr;   // UV <-> UV factors: + Jc^T Ji (replicated: rank 0 only, see launch)
#pragma unroll
    for (int a = 0; a < 6; a++) out[8 * a + c] = v[a] + (pb ? pb[8 * a + c] : 0.0);
}
int shared_reduce_blocks() {
    static int n = 0;
    if (!n) {
        int dev = 0, sms = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        n = std::max(1, sms) * 4;
    }
    return n;
}
void launch_shared_build(const DevProblem& P, cudaStream_t st) {
    if (!P.Ng) return;
    const size_t npg = (size_t)P.Ng * P.Np;
    if (npg) k_shared_Y<<<(unsigned)((npg + 255) / 256), 256, 0, st>>>(P);
    const int nb = std::min(P.sh_nblk, std::max(1, cdiv(P.Np, 256)));
    for (int g1 = 0; g1 < P.Ng; g1++)
        for (int g2 = g1; g2 < P.Ng; g2++) {
            k_shared_schur<<<nb, 256, 0, st>>>(P, g1, g2);
            k_shared_fold_schur<<<1, RED_BLOCK, 0, st>>>(P, g1, g2, nb);
        }
    double* tile_buf = P.sh_part + (size_t)SH_RED * std::max(cdiv(P.Nq, 256), P.sh_nblk);   // 48 per tile, behind the block partials
    for (int g = 0; g < P.Ng; g++) {
        if (P.Nt) k_shared_border<<<cdiv(P.Nt, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, 0, st>>>(P, g, P.grp_kind[g], tile_buf);
        k_shared_border_reduce<<<cdiv(P.Nc * 8, 256), 256, 0, st>>>(P, g, tile_buf);
    }
}

// after the all-reduce of sh_acc: Sd and rhs of the pseudo-cameras, and (write_tiles) the border blocks S_cg and the blocks
// between pseudo-cameras into the tiles.  Thread per 6x6 block (e, s), e < s or e == s (diagonal: Sd / rhs only).
__device__ __forceinline__ size_t sh_ex_elem(const DevProblem& P, int i, int j) {
    const int I = i / EX_TB, J = j / EX_TB;
    const int tile = P.ex_tile_of[I * P.ex_NT + J];
    return (size_t)tile * EX_TD * EX_TD + (size_t)(6 * (i - I * EX_TB)) * EX_TD + 6 * (j - J * EX_TB);
}
__global__ void k_shared_build_finish(DevProblem P, int write_tiles) {
    const int nsh = P.Nce - P.Nc;
    const size_t total = (size_t)P.Nce * nsh;
    const size_t id = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= total) return;
    const int e = (int)(id / nsh), s = P.Nc + (int)(id % nsh);
    if (e > s) return;
    // group / block of the column pseudo-camera s
    int g2 = 0;
    while (g2 + 1 < P.Ng && P.grp_blk[g2 + 1] <= s) g2++;
    const int b2 = s - P.grp_blk[g2];
    const double* accS = P.sh_acc + (size_t)P.Ng * P.Nc * 48;
    const double* rS = accS + (size_t)P.Ng * P.Ng * 64;
    auto live2 = [&](int j) { const int c = 6 * b2 + j; return c < 8 && P.grp_live[8 * g2 + c]; };
    if (e == s) {
        const double* SS = accS + ((size_t)g2 * P.Ng + g2) * 64;
        for (int i = 0; i < 6; i++) {
            const int ci = 6 * b2 + i;
            P.rhs[6 * (size_t)s + i] = live2(i) ? -(P.sh_g[8 * g2 + ci] - rS[8 * g2 + ci]) : 0.0;
            for (int j = i; j < 6; j++) P.Sd[21 * (size_t)s + sym6(i, j)] = (live2(i) && live2(j)) ? SS[8 * ci + 6 * b2 + j] : 0.0;
        }
        return;
    }
    if (!write_tiles) return;
    double* out = P.ex_S + sh_ex_elem(P, e, s);
    if (e < P.Nc) {   // border block of a real camera
        const double* B = P.sh_acc + ((size_t)g2 * P.Nc + e) * 48;
        for (int i = 0; i < 6; i++)
            for (int j = 0; j < 6; j++) out[(size_t)i * EX_TD + j] = live2(j) ? B[8 * i + 6 * b2 + j] : 0.0;
        return;
    }
    int g1 = 0;
    while (g1 + 1 < P.Ng && P.grp_blk[g1 + 1] <= e) g1++;
    const int b1 = e - P.grp_blk[g1];
    const double* H = P.sh_H + ((size_t)g1 * P.Ng + g2) * 64;
    const double* SS = accS + ((size_t)g1 * P.Ng + g2) * 64;
    for (int i = 0; i < 6; i++) {
        const int ci = 6 * b1 + i;
        const bool li = ci < 8 && P.grp_live[8 * g1 + ci];
        for (int j = 0; j < 6; j++) out[(size_t)i * EX_TD + j] = (li && live2(j)) ? H[8 * ci + 6 * b2 + j] - SS[8 * ci + 6 * b2 + j] : 0.0;
    }
}
void launch_shared_build_finish(const DevProblem& P, int write_tiles, cudaStream_t st) {
    if (!P.Ng) return;
    const size_t total = (size_t)P.Nce * (P.Nce - P.Nc);
    k_shared_build_finish<<<(unsigned)((total + 127) / 128), 128, 0, st>>>(P, write_tiles);
}

// ---------------------------------------------------------------------------
// back-substitution / model change / update
// ---------------------------------------------------------------------------
__device__ __forceinline__ void grp_delta(const DevProblem& P, const double* x, int g, double* d8) {
    const int base = 6 * P.grp_blk[g];
#pragma unroll
    for (int c = 0; c < 8; c++) d8[c] = P.grp_live[8 * g + c] ? x[base + c] : 0.0;   // a model's columns 6, 7 are the first two of its second pseudo-camera
}
// dp -= sum_g Y_g[p] delta_g
__global__ void __launch_bounds__(256) k_shared_backsub(DevProblem P, const double* __restrict__ x) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P.Np) return;
    double u0 = 0.0, u1 = 0.0, u2 = 0.0;
    for (int g = 0; g < P.Ng; g++) {
        double d8[8];
        grp_delta(P, x, g, d8);
        const double* Y = P.sh_Y + ((size_t)g * P.Np + p) * 24;
#pragma unroll
        for (int c = 0; c < 8; c++) { u0 += Y[c] * d8[c]; u1 += Y[8 + c] * d8[c]; u2 += Y[16 + c] * d8[c]; }
    }
    double* o = P.dp + 4 * (size_t)p;
    o[0] -= u0; o[1] -= u1; o[2] -= u2;
}
// G_o delta of the camera's own groups, two planes over the slots (read by k_model_tiles)
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32) k_shared_gd(DevProblem P, const double* __restrict__ x) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * WARPS_PER_BLOCK + warp;
    if (tile >= P.Nt) return;
    const int cam = __ldg(P.tile_cam + tile);
    const int start = __ldg(P.tile_start + tile), len = __ldg(P.tile_len + tile);
    const size_t Ns = (size_t)P.Ns;
    const int gk[2] = {__ldg(P.cam_grp + 2 * cam), __ldg(P.cam_grp + 2 * cam + 1)};
    double d8[2][8];
#pragma unroll
    for (int k = 0; k < 2; k++) {
#pragma unroll
        for (int c = 0; c < 8; c++) d8[k][c] = 0.0;
        if (gk[k] >= 0) grp_delta(P, x, gk[k], d8[k]);
    }
    for (int s = start + lane; s < start + len; s += 32) {
        double m0 = 0.0, m1 = 0.0;
        if (__ldcs(P.obs_pt + s) >= 0) {
            const double p0 = __ldcs(P.Pc + s), p1 = __ldcs(P.Pc + Ns + s), p2 = __ldcs(P.Pc + 2 * Ns + s);
            const double2 uv = ldcs2(P.obs_uv + 2 * (size_t)s);
            double W[4];
            const double* Wp = nullptr;
            if (P.obs_info) { for (int i = 0; i < 4; i++) W[i] = P.obs_info[i * Ns + s]; Wp = W; }
#pragma unroll
            for (int k = 0; k < 2; k++) {   // unrolled: the group kind is a compile-time constant inside, d8 stays in registers
                if (gk[k] < 0) continue;
                double G[16];
                shared_jac(P, gk[k], k, cam, p0, p1, p2, uv.x, uv.y, Wp, G);
#pragma unroll
                for (int c = 0; c < 8; c++) { m0 += G[c] * d8[k][c]; m1 += G[8 + c] * d8[k][c]; }
            }
        }
        P.sh_gd[s] = m0; P.sh_gd[Ns + s] = m1;
    }
}
void launch_shared_backsub(const DevProblem& P, const double* x, cudaStream_t st) {
    if (!P.Ng) return;
    if (P.Np) k_shared_backsub<<<cdiv(P.Np, 256), 256, 0, st>>>(P, x);
    if (P.Nt) k_shared_gd<<<cdiv(P.Nt, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, 0, st>>>(P, x);
}
// value of column c of group g
__device__ __forceinline__ double* grp_param(const DevProblem& P, int g, int c) {
    return P.grp_kind[g] == 0 ? P.intr + 10 * (size_t)P.grp_index[g] + c : P.mount + 6 * (size_t)P.grp_index[g] + c;
}
// step / parameter norms of the live shared columns and the model change of the intrinsics priors: ADDED to the replicated
// scalars k_model_reduce has just written (same stream)
__global__ void k_shared_stats(DevProblem P, const double* __restrict__ x) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double s2 = 0.0, x2 = 0.0, mc = 0.0;
    for (int g = 0; g < P.Ng; g++) {
        double d8[8];
        grp_delta(P, x, g, d8);
        const int w = P.grp_kind[g] == 0 ? 8 : 6;
        for (int c = 0; c < w; c++) if (P.grp_live[8 * g + c]) { s2 += d8[c] * d8[c]; const double v = *grp_param(P, g, c); x2 += v * v; }
    }
    for (int i = 0; i < P.Nipr; i++) {
        const int blk = P.ipr_block[i], m = P.ipr_model[i];
        if (blk == 4) continue;
        int g = -1;
        for (int k = 0; k < P.Ng; k++) if (P.grp_kind[k] == 0 && P.grp_index[k] == m) g = k;
        if (g < 0) continue;
        const int first = blk == 0 ? 0 : (blk == 1 ? 1 : (blk == 2 ? 3 : 6)), dim = blk == 0 ? 1 : (blk == 2 ? 3 : 2);
        if (!P.grp_live[8 * g + first]) continue;
        double d8[8];
        grp_delta(P, x, g, d8);
        const double* A = P.ipr_A + 9 * (size_t)i;
        for (int a = 0; a < dim; a++) {
            double mm = 0.0;
            for (int j = 0; j < dim; j++) mm += A[dim * a + j] * d8[first + j];
            mc += mm * (P.ipr_res[3 * (size_t)i + a] + 0.5 * mm);
        }
    }
    P.scalars[SC_STEP2_CAM] += s2; P.scalars[SC_X2_CAM] += x2; P.scalars[SC_MODEL_PRIOR_CAM] -= mc;
}
void launch_shared_stats(const DevProblem& P, const double* x, cudaStream_t st) {
    if (P.Ng) k_shared_stats<<<1, 32, 0, st>>>(P, x);
}
// one block: the save of every entry precedes the update behind the barrier
__global__ void __launch_bounds__(1024) k_shared_apply(DevProblem P, const double* __restrict__ x, int restore) {
    if (restore) {
        for (int i = threadIdx.x; i < 10 * P.Nm; i += blockDim.x) P.intr[i] = P.intr_saved[i];
        for (int i = threadIdx.x; i < 6 * P.Nmt; i += blockDim.x) P.mount[i] = P.mount_saved[i];
        return;
    }
    for (int i = threadIdx.x; i < 10 * P.Nm; i += blockDim.x) P.intr_saved[i] = P.intr[i];
    for (int i = threadIdx.x; i < 6 * P.Nmt; i += blockDim.x) P.mount_saved[i] = P.mount[i];
    __syncthreads();
    if (threadIdx.x == 0) {   // a handful of columns
        for (int g = 0; g < P.Ng; g++) {
            double d8[8];
            grp_delta(P, x, g, d8);
            const int w = P.grp_kind[g] == 0 ? 8 : 6;
            for (int c = 0; c < w; c++) *grp_param(P, g, c) += d8[c];
        }
    }
}
void launch_shared_apply(const DevProblem& P, const double* x, cudaStream_t st) {
    if (P.Ng) k_shared_apply<<<1, 1024, 0, st>>>(P, x, 0);
}
void launch_shared_restore(const DevProblem& P, cudaStream_t st) {
    if (!P.Ng) return;
    k_shared_apply<<<1, 1024, 0, st>>>(P, nullptr, 1);
}

}  // namespace sba
