// sba_kernels_dense.cu -- dense FP64 Cholesky of the explicit reduced camera system on the device: the DIRECT reduced solve
// (the reference's default solver type is ceres::SPARSE_SCHUR, i.e. a Cholesky factorisation of S, not CG:
// modularsbasolver.h:672, modularsbasolver.cpp:898) and the selected inverse behind the point covariances of
// ModularSBASolver::std_step (modularsbasolver.cpp:952-1004; pairs requested by landmarkssbamodule.cpp:264-297).
//
// S (6 Nce scalars per side, 10 752 for the Venice shape) is already explicit in HBM as 384x384 tiles; it is copied into a
// dense row-major matrix and factorised S = L L^T by a right-looking blocked algorithm with 64x64 blocks:
//   per block column:  one CTA factorises the diagonal block in shared memory and inverts the triangular factor;
//                      panel  L21 = A21 inv(L11)^T  and trailing update  A22 -= L21 L21^T  are tile GEMMs.
// The same tile GEMM (64x64 tile per CTA, 4x4 micro-tile per thread, K staged through shared memory in steps of 16) carries the
// triangular inverse Z = L^-1 and S^-1 = Z^T Z.  This is the one place on the path with a real contraction (n^3 / 3 flops);
// FP64 on the CUDA cores.  No atomics: deterministic.
#include <algorithm>

#include "sba_kernel_util.cuh"

namespace sba {

constexpr int DB = 64;        // block size

// tiles (upper block triangle, diagonal tiles carry S_ii / 2 on their block diagonal and zeros below) -> dense symmetric
// row-major n x n; rows / columns whose diagonal is zero (constant cameras without damping) get a unit diagonal
__global__ void __launch_bounds__(256) k_dense_from_tiles(DevProblem P, double* __restrict__ D, int n) {
    const int t = blockIdx.x;
    const int I = P.ex_tile_I[t], J = P.ex_tile_J[t];
    const double* T = P.ex_Sown + (size_t)t * EX_TD * EX_TD;
    for (int e = threadIdx.x; e < EX_TD * EX_TD; e += blockDim.x) {
        const int r = e / EX_TD, c = e - r * EX_TD;
        const double v = T[e];
        const size_t gr = (size_t)EX_TD * I + r, gc = (size_t)EX_TD * J + c;
        if (I != J) { D[gr * n + gc] = v; D[gc * n + gr] = v; }
        else {
            // upper part of the diagonal tile holds S, its 6x6 diagonal blocks hold S_ii / 2 in full (both triangles)
            const bool same_block = r / 6 == c / 6;
            if (same_block) D[gr * n + gc] = 2.0 * v;
            else if (c > r) { D[gr * n + gc] = v; D[gc * n + gr] = v; }
        }
    }
}
__global__ void k_dense_fix_diag(double* D, int n, int n_live) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (i >= n_live || D[(size_t)i * n + i] == 0.0) D[(size_t)i * n + i] = 1.0;   // padding rows / constant blocks: decoupled unit rows
}

// C (M x N, ldc) = beta C + alpha op(A) op(B); 64x64 tile per CTA.  K range per tile by kmode:
//  0 full; 1 k >= max(bi, bj) blocks (Z^T Z with lower-triangular Z); 2 lower tiles only (bj <= bi), full K; 3 k >= bj blocks
template <bool AT, bool BT>
__global__ void __launch_bounds__(256) k_dgemm(const double* __restrict__ A, int lda, const double* __restrict__ B, int ldb,
                                               double* __restrict__ C, int ldc, int K, double alpha, double beta, int kmode) {
    __shared__ double As[16][DB + 4], Bs[16][DB + 4];
    const int bi = blockIdx.y, bj = blockIdx.x;
    if (kmode == 2 && bj > bi) return;
    int k0 = 0;
    if (kmode == 1) k0 = DB * max(bi, bj);
    if (kmode == 3) k0 = DB * bj;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    double acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.0;
    const int m0 = DB * bi, n0 = DB * bj;
    for (int k = k0; k < K; k += 16) {
        // stage A (64 x 16) and B (16 x 64): 1024 elements each, 4 per thread
#pragma unroll
        for (int e = threadIdx.x; e < 1024; e += 256) {
            if (AT) { const int kk = e >> 6, m = e & 63; As[kk][m] = A[(size_t)(k + kk) * lda + m0 + m]; }
            else { const int m = e >> 4, kk = e & 15; As[kk][m] = A[(size_t)(m0 + m) * lda + k + kk]; }
            if (BT) { const int nn = e >> 4, kk = e & 15; Bs[kk][nn] = B[(size_t)(n0 + nn) * ldb + k + kk]; }
            else { const int kk = e >> 6, nn = e & 63; Bs[kk][nn] = B[(size_t)(k + kk) * ldb + n0 + nn]; }
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; kk++) {
            double a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; i++) { a[i] = As[kk][ty + 16 * i]; b[i] = Bs[kk][tx + 16 * i]; }
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int j = 0; j < 4; j++) acc[i][j] += a[i] * b[j];
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
            double* c = C + (size_t)(m0 + ty + 16 * i) * ldc + n0 + tx + 16 * j;
            *c = (beta == 0.0 ? 0.0 : beta * *c) + alpha * acc[i][j];
        }
}

// one CTA (256 threads): Cholesky of the 64x64 diagonal block at D[k0, k0] in shared memory (lower factor written back, upper part
// zeroed) and the inverse of the factor into Linv (64x64 row-major, lower triangular).  flag set when a pivot is not positive.
__global__ void __launch_bounds__(256) k_potrf_block(double* __restrict__ D, int n, int k0, double* __restrict__ Linv, int* flag) {
    extern __shared__ double potrf_smem[];   // two 64 x 65 arrays (66.5 KB: dynamic)
    double (*L)[DB + 1] = reinterpret_cast<double (*)[DB + 1]>(potrf_smem);
    double (*X)[DB + 1] = reinterpret_cast<double (*)[DB + 1]>(potrf_smem + DB * (DB + 1));
    __shared__ int bad;
    const int tid = threadIdx.x;
    if (tid == 0) bad = 0;
    for (int e = tid; e < DB * DB; e += 256) { const int r = e >> 6, c = e & 63; L[r][c] = D[(size_t)(k0 + r) * n + k0 + c]; }
    __syncthreads();
    for (int j = 0; j < DB; j++) {
        if (tid == 0) {
            const double d = L[j][j];
            if (!(d > 0.0)) { bad = 1; L[j][j] = 1.0; } else L[j][j] = sqrt(d);
        }
        __syncthreads();
        if (tid > j && tid < DB) L[tid][j] /= L[j][j];
        __syncthreads();
        // trailing lower triangle (rows r > j, columns j < c <= r), elements dealt round-robin to the 256 threads
        const int m = DB - 1 - j;
        for (int e = tid; e < m * m; e += 256) {
            const int r = j + 1 + e / m, c = j + 1 + e % m;
            if (c <= r) L[r][c] -= L[r][j] * L[c][j];
        }
        __syncthreads();
    }
    for (int e = tid; e < DB * DB; e += 256) { const int r = e >> 6, c = e & 63; D[(size_t)(k0 + r) * n + k0 + c] = c <= r ? L[r][c] : 0.0; }
    // inverse: thread t < 64 solves L x = e_t (column t of L^-1) by forward substitution, X in shared memory
    if (tid < DB) {
        const int t = tid;
        for (int i = 0; i < DB; i++) {
            double sacc = i == t ? 1.0 : 0.0;
            for (int k = t; k < i; k++) sacc -= L[i][k] * X[k][t];
            X[i][t] = i < t ? 0.0 : sacc / L[i][i];
        }
    }
    __syncthreads();
    for (int e = tid; e < DB * DB; e += 256) { const int r = e >> 6, c = e & 63; Linv[(size_t)r * DB + c] = X[r][c]; }
    if (tid == 0 && bad) *flag = 1;
}

// in-place dense Cholesky, lower factor; Linv_all: (n / 64) inverse diagonal blocks; work: n x 64 panel scratch.  Returns launches.
// Two levels of blocking: 64-wide block columns are factorised and applied INSIDE a 256-wide panel only; the matrix behind the
// panel is updated once per panel with K = 256 (a quarter of the big-GEMM launches, four times the arithmetic per C tile).
int dense_cholesky(double* D, int n, double* Linv_all, double* work, int* flag, cudaStream_t st) {
    const int NB = n / DB;
    constexpr int PW = 4;   // blocks per panel
    constexpr size_t POTRF_SMEM = 2 * DB * (DB + 1) * sizeof(double);
    static bool attr_set = false;
    if (!attr_set) { cudaFuncSetAttribute(k_potrf_block, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)POTRF_SMEM); attr_set = true; }
    int launches = 0;
    for (int pb = 0; pb < NB; pb += PW) {
        const int pe = std::min(NB, pb + PW);   // panel = block columns pb .. pe-1
        for (int kb = pb; kb < pe; kb++) {
            const int k0 = DB * kb, rem = NB - kb - 1;
            double* Li = Linv_all + (size_t)kb * DB * DB;
            k_potrf_block<<<1, 256, POTRF_SMEM, st>>>(D, n, k0, Li, flag);
            launches++;
            if (!rem) break;
            double* A21 = D + (size_t)(k0 + DB) * n + k0;
            // L21 = A21 Linv^T for every row below (through the scratch panel: the GEMM is not in place)
            k_dgemm<false, true><<<dim3(1, rem), 256, 0, st>>>(A21, n, Li, DB, work, DB, DB, 1.0, 0.0, 0);
            cudaMemcpy2DAsync(A21, (size_t)n * sizeof(double), work, DB * sizeof(double), DB * sizeof(double), (size_t)rem * DB, cudaMemcpyDeviceToDevice, st);
            launches += 2;
            // update of the remaining block columns of THIS panel: A[k+1.., k+1..pe) -= L21 L21[0 .. pe-k-1)^T  (rows: all below;
            // tiles above the diagonal are written too -- never read)
            const int pc = pe - kb - 1;
            if (pc > 0) {
                k_dgemm<false, true><<<dim3(pc, rem), 256, 0, st>>>(A21, n, A21, n, D + (size_t)(k0 + DB) * n + k0 + DB, n, DB, -1.0, 1.0, 0);
                launches++;
            }
        }
        // everything behind the panel: A22 -= L_panel L_panel^T, K = panel width, lower tiles only
        const int rem = NB - pe;
        if (rem > 0) {
            const double* Lp = D + (size_t)(DB * pe) * n + DB * pb;
            k_dgemm<false, true><<<dim3(rem, rem), 256, 0, st>>>(Lp, n, Lp, n, D + (size_t)(DB * pe) * n + DB * pe, n, DB * (pe - pb), -1.0, 1.0, 2);
            launches++;
        }
    }
    return launches;
}

// x = S^-1 b with the factor, right-looking by 64-row blocks.  One launch per block step: EVERY CTA recomputes the block's
// solution x_k = inv(L_kk) b_k (forward) / inv(L_kk)^T b_k (backward) from the by-then final b_k (a 64x64 product), CTA 0 stores it,
// and the other CTAs subtract its contribution from their share of the rest of the right-hand side:
//   forward:  b[r] -= L[r, k-block] . x_k   for rows r behind the block (each CTA 64 rows, row segments of 512 B);
//   backward: b[c] -= L[k-block, c]^T x_k   for columns c in front of the block (each CTA 256 columns, coalesced along c).
// n^2 / 2 matrix elements are read once, by many CTAs, instead of by one.
__global__ void __launch_bounds__(256) k_trsv_step(const double* __restrict__ L, int n, const double* __restrict__ Linv_all, double* x, int kb, int backward) {
    __shared__ double bk[DB], xk[DB];
    const int tid = threadIdx.x, k0 = DB * kb;
    if (tid < DB) bk[tid] = x[k0 + tid];
    __syncthreads();
    const double* Li = Linv_all + (size_t)kb * DB * DB;
    if (tid < DB) {
        double v = 0.0;
        if (!backward) { for (int j = 0; j <= tid; j++) v += Li[(size_t)tid * DB + j] * bk[j]; }
        else { for (int j = tid; j < DB; j++) v += Li[(size_t)j * DB + tid] * bk[j]; }
        xk[tid] = v;
    }
    __syncthreads();
    if (blockIdx.x == 0) { if (tid < DB) x[k0 + tid] = xk[tid]; return; }
    const int w = blockIdx.x - 1;
    if (!backward) {
        // 64 rows per CTA, four threads per row (16 columns each), folded in fixed order
        __shared__ double part[4][DB];
        const int r = k0 + DB + DB * w + (tid & 63), h = tid >> 6;
        double acc = 0.0;
        if (r < n) { const double* row = L + (size_t)r * n + k0 + 16 * h; for (int c = 0; c < 16; c++) acc += row[c] * xk[16 * h + c]; }
        part[h][tid & 63] = acc;
        __syncthreads();
        if (h == 0 && r < n) x[r] -= (part[0][tid] + part[1][tid]) + (part[2][tid] + part[3][tid]);
    } else {
        const int c = 256 * w + tid;
        if (c < k0) {
            double acc = 0.0;
            for (int r = 0; r < DB; r++) acc += L[(size_t)(k0 + r) * n + c] * xk[r];
            x[c] -= acc;
        }
    }
}
int dense_solve(const double* L, int n, const double* Linv_all, double* x, cudaStream_t st) {
    const int NB = n / DB;
    for (int kb = 0; kb < NB; kb++) k_trsv_step<<<1 + (NB - 1 - kb), 256, 0, st>>>(L, n, Linv_all, x, kb, 0);
    for (int kb = NB - 1; kb >= 0; kb--) k_trsv_step<<<1 + (DB * kb + 255) / 256, 256, 0, st>>>(L, n, Linv_all, x, kb, 1);
    return 2 * NB;
}

// Z = L^-1 (lower triangular, n x n, zero above the diagonal blocks) and Sinv = Z^T Z.  Z and Sinv are caller buffers; work: 64 x n.
__global__ void k_copy_block(const double* __restrict__ src, int lds, double* __restrict__ dst, int ldd) {
    for (int e = threadIdx.x; e < DB * DB; e += blockDim.x) { const int r = e / DB, c = e - r * DB; dst[(size_t)r * ldd + c] = src[(size_t)r * lds + c]; }
}
int dense_inverse(const double* L, int n, const double* Linv_all, double* Z, double* Sinv, double* work, cudaStream_t st) {
    const int NB = n / DB;
    int launches = 0;
    cudaMemsetAsync(Z, 0, (size_t)n * n * sizeof(double), st);
    for (int ib = 0; ib < NB; ib++) {
        const double* Li = Linv_all + (size_t)ib * DB * DB;
        k_copy_block<<<1, 256, 0, st>>>(Li, DB, Z + (size_t)(DB * ib) * n + DB * ib, n);   // Z_ii = inv(L_ii)
        launches++;
        if (!ib) continue;
        // T = L_{i, 0:i} Z_{0:i, 0:i} (64 x 64 i), K range of column block j starts at j (Z lower triangular); Z_{i, 0:i} = -inv(L_ii) T
        k_dgemm<false, false><<<dim3(ib, 1), 256, 0, st>>>(L + (size_t)(DB * ib) * n, n, Z, n, work, n, DB * ib, 1.0, 0.0, 3);
        k_dgemm<false, false><<<dim3(ib, 1), 256, 0, st>>>(Li, DB, work, n, Z + (size_t)(DB * ib) * n, n, DB, -1.0, 0.0, 0);
        launches += 2;
    }
    k_dgemm<true, false><<<dim3(NB, NB), 256, 0, st>>>(Z, n, Z, n, Sinv, n, n, 1.0, 0.0, 1);   // Sinv = Z^T Z
    return launches + 1;
}

// ---------------------------------------------------------------------------
// point covariances: Cov_p = Hinv + sum_{e, f} (W_e Hinv)^T Sinv[e, f] (W_f Hinv) over the reduced-side blocks e, f the point
// touches -- its observations' cameras (W = Jc^T A, 6 x 3) and the shared groups (W_g^T, 8 x 3).  One WARP per point: the
// lanes split the (e, f) pairs, a fixed shuffle tree sums the nine entries.  Undamped blocks (the caller linearised with
// zero damping): (J^T J)^-1 restricted to the point, what ceres::Covariance returns for the pair (block, block).
// ---------------------------------------------------------------------------
__device__ __forceinline__ void cov_entry(const DevProblem& P, int p, int q0, int k, int e, const double* Hi, int* ext, double* M /*up to 8 x 3*/, int* rows) {
    if (e < k) {   // observation e of the point: M = Jc^T (A Hinv), Jc = [C | -A], C = A [X - t]x Q^T
        const int q = q0 + e;
        const int cam = P.q_cam[q];
        const size_t Nq = (size_t)P.Nq;
        double A[6];
        for (int i = 0; i < 6; i++) A[i] = P.qA[(size_t)i * Nq + q];
        const double* cc = P.cam_cache + (size_t)CAM_CACHE_STRIDE * cam;
        const double* Q = P.cam_Q + 12 * (size_t)cam;
        const double* X = P.pts + 4 * (size_t)p;
        const double d[3] = {X[0] - (cc[9] - Q[9]), X[1] - (cc[10] - Q[10]), X[2] - (cc[11] - Q[11])};   // X - t (body)
        double Jc[12];
        for (int r = 0; r < 2; r++) {
            const double a0 = A[3 * r], a1 = A[3 * r + 1], a2 = A[3 * r + 2];
            const double w0 = a1 * d[2] - a2 * d[1], w1 = a2 * d[0] - a0 * d[2], w2 = a0 * d[1] - a1 * d[0];
            for (int c = 0; c < 3; c++) { Jc[6 * r + c] = w0 * Q[3 * c] + w1 * Q[3 * c + 1] + w2 * Q[3 * c + 2]; Jc[6 * r + 3 + c] = -A[3 * r + c]; }
        }
        double AH[6];   // A Hinv (2 x 3)
        for (int r = 0; r < 2; r++)
            for (int c = 0; c < 3; c++) AH[3 * r + c] = A[3 * r] * Hi[c] + A[3 * r + 1] * Hi[3 + c] + A[3 * r + 2] * Hi[6 + c];
        const bool fixed = P.cam_fixed[cam];
        for (int a = 0; a < 6; a++)
            for (int c = 0; c < 3; c++) M[3 * a + c] = fixed ? 0.0 : Jc[a] * AH[c] + Jc[6 + a] * AH[3 + c];
        *ext = 6 * cam; *rows = 6;
    } else {       // shared group: M = W_g^T Hinv, rows = its 8 columns (contiguous in the extended layout)
        const int g = e - k;
        const double* W = P.sh_W + ((size_t)g * P.Np + p) * 24;
        for (int c8 = 0; c8 < 8; c8++)
            for (int c = 0; c < 3; c++) M[3 * c8 + c] = P.grp_live[8 * g + c8] ? W[c8] * Hi[c] + W[8 + c8] * Hi[3 + c] + W[16 + c8] * Hi[6 + c] : 0.0;
        *ext = 6 * P.grp_blk[g]; *rows = 8;
    }
}
__global__ void __launch_bounds__(256) k_point_covariance(DevProblem P, const double* __restrict__ Sinv, int n, double* __restrict__ cov9) {
    const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (p >= P.Np) return;
    const double* Hp = P.Hinv + 8 * (size_t)p;
    const double Hi[9] = {Hp[0], Hp[1], Hp[2], Hp[1], Hp[3], Hp[4], Hp[2], Hp[4], Hp[5]};
    const int k = P.pt_off[p + 1] - P.pt_off[p], q0 = P.pt_q0[p];
    const int ne = k + P.Ng;
    double acc[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    if (!P.pt_fixed[p])
        for (long long pr = lane; pr < (long long)ne * ne; pr += 32) {
            const int e = (int)(pr / ne), f = (int)(pr - (long long)e * ne);
            double Me[24], Mf[24];
            int xe, xf, re, rf;
            cov_entry(P, p, q0, k, e, Hi, &xe, Me, &re);
            cov_entry(P, p, q0, k, f, Hi, &xf, Mf, &rf);
            for (int a = 0; a < re; a++) {
                double t[3] = {0, 0, 0};   // row a of Sinv[e, f] times Mf
                for (int b = 0; b < rf; b++) {
                    const double s = Sinv[(size_t)(xe + a) * n + xf + b];
                    t[0] += s * Mf[3 * b]; t[1] += s * Mf[3 * b + 1]; t[2] += s * Mf[3 * b + 2];
                }
                for (int c = 0; c < 3; c++) { acc[3 * c] += Me[3 * a + c] * t[0]; acc[3 * c + 1] += Me[3 * a + c] * t[1]; acc[3 * c + 2] += Me[3 * a + c] * t[2]; }
            }
        }
    for (int i = 0; i < 9; i++) acc[i] = warp_sum(acc[i]);
    if (lane == 0) for (int i = 0; i < 9; i++) cov9[9 * (size_t)p + i] = (P.pt_fixed[p] ? 0.0 : Hi[i]) + acc[i];
}

void launch_dense_from_tiles(const DevProblem& P, double* D, int n, cudaStream_t st) {
    cudaMemsetAsync(D, 0, (size_t)n * n * sizeof(double), st);
    if (P.ex_nown) k_dense_from_tiles<<<P.ex_nown, 256, 0, st>>>(P, D, n);
    k_dense_fix_diag<<<cdiv(n, 256), 256, 0, st>>>(D, n, 6 * P.Nc);
}
void launch_point_covariance(const DevProblem& P, const double* Sinv, int n, double* cov9, cudaStream_t st) {
    if (P.Np) k_point_covariance<<<cdiv(P.Np * 32, 256), 256, 0, st>>>(P, Sinv, n, cov9);
}

}  // namespace sba
