// sba_kernels_explicit.cu -- the EXPLICIT reduced camera system
//   S = (Hcc + Dc) - sum_p W_p (Hpp + Dp)^-1 W_p^T,   W_o = Jc_o^T A_o  (6x3)
// (Ceres SchurEliminator behind modularsbasolver.cpp:938; g2o analogue block_solver.hpp:371-447),
// kept as dense 384x384 tiles over the upper block triangle, and its mat-vec for the PCG.
//
// Why: one application of the implicit operator streams every observation twice and is bound by gathers
// (L1 wavefronts), 150 times per LM iteration.  When the co-visibility graph is dense (Venice: 69 % of the
// camera pairs share a point) the tiles are SMALLER than the observation planes and their mat-vec is a pure
// coalesced stream that reads every stored byte once (symmetric storage).
//
// Determinism (no floating-point atomics): every (point, observation pair) contribution has a STATIC slot in a
// block-sorted buffer (finalize builds the list); the assembly kernel sums the run of a block in buffer order.
//
// Algebra: with d = X - t, Q = (R Jr)^T and C = A [d]x Q^T (see sba_kernels_points.cu),
//   W_o = T_c D_o (A_o^T A_o),  T_c = diag(Q_c, I),  D_o = [ [d_o]x^T ; -I ]  (6x3)
//   W_i Hinv W_j^T = T_i ( D_i K D_j^T ) T_j^T,   K = A_i^T (A_i Hinv A_j^T) A_j   (3x3)
// so a contribution needs only the point-order A planes, the point, Hinv and the two camera centres; the
// per-camera factors T are applied once per block by the assembly kernel.
#include <algorithm>
#include <cstdlib>

#include "sba_kernel_util.cuh"

namespace sba {

// ---------------------------------------------------------------------------
// producer: one thread per contribution, point order (neighbouring threads share the point, Hinv and most A rows:
// the gathers hit L1).  Writes the COMPACT contribution (K 3x3, X) = 12 doubles = 96 B = three whole sectors to its
// static slot; the 6x6 block D_i K D_j^T = [ -U [dj]x , U ; -K [dj]x , K ], U = [di]x K, d = X - t, is expanded by the
// assembly kernel, where t_i and t_j are per-block constants (3x less traffic than storing the 36 entries).
// ---------------------------------------------------------------------------
constexpr int EX_CW = 12;   // doubles per stored contribution
__global__ void __launch_bounds__(256) k_explicit_contrib(DevProblem P) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= P.ex_ncontrib) return;
    const int4 rec = __ldcs(P.ex_rec + c);   // {qa, qb, dest, point}
    const int qa = rec.x, qb = rec.y;
    const size_t dest = (size_t)rec.z;
    const int pt = rec.w;
    const size_t Nq = (size_t)P.Nq;
    double Aa[6], Ab[6];
#pragma unroll
    for (int i = 0; i < 6; i++) { Aa[i] = __ldg(P.qA + (size_t)i * Nq + qa); Ab[i] = __ldg(P.qA + (size_t)i * Nq + qb); }
    const d4 X = ld256(P.pts + 4 * (size_t)pt);
    const d4 h0 = ld256(P.Hinv + 8 * (size_t)pt), h1 = ld256(P.Hinv + 8 * (size_t)pt + 4);   // (h00 h01 h02 h11) (h12 h22 - -)
    // E = Aa Hinv (2x3)
    const double E0 = Aa[0] * h0.x + Aa[1] * h0.y + Aa[2] * h0.z, E1 = Aa[0] * h0.y + Aa[1] * h0.w + Aa[2] * h1.x,
                 E2 = Aa[0] * h0.z + Aa[1] * h1.x + Aa[2] * h1.y;
    const double E3 = Aa[3] * h0.x + Aa[4] * h0.y + Aa[5] * h0.z, E4 = Aa[3] * h0.y + Aa[4] * h0.w + Aa[5] * h1.x,
                 E5 = Aa[3] * h0.z + Aa[4] * h1.x + Aa[5] * h1.y;
    // G = E Ab^T (2x2)
    const double G00 = E0 * Ab[0] + E1 * Ab[1] + E2 * Ab[2], G01 = E0 * Ab[3] + E1 * Ab[4] + E2 * Ab[5];
    const double G10 = E3 * Ab[0] + E4 * Ab[1] + E5 * Ab[2], G11 = E3 * Ab[3] + E4 * Ab[4] + E5 * Ab[5];
    // F = G Ab (2x3);  K = Aa^T F (3x3)
    double F[6];
#pragma unroll
    for (int j = 0; j < 3; j++) { F[j] = G00 * Ab[j] + G01 * Ab[3 + j]; F[3 + j] = G10 * Ab[j] + G11 * Ab[3 + j]; }
    double K[9];
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int j = 0; j < 3; j++) K[3 * r + j] = Aa[r] * F[j] + Aa[3 + r] * F[3 + j];
    double* o = P.ex_contrib + EX_CW * dest;   // 96 B, 32-byte aligned
    st256(o, K[0], K[1], K[2], K[3]);
    st256(o + 4, K[4], K[5], K[6], K[7]);
    st256(o + 8, K[8], X.x, X.y, X.z);
}

// ---------------------------------------------------------------------------
// assembly: one THREAD per off-diagonal block (i < j).  The run of the block is contiguous in the buffer; the thread
// streams it with 256-bit loads, expands every contribution with its block-constant camera centres and keeps the 36
// sums in registers (run order -> deterministic, no shuffles).  Then S_ij = -T_i (sum B) T_j^T, T = diag(Q, I), is
// written into its tile.  Blocks touching a fixed camera are zero (its Jc is zero in the reference's linearisation).
// Neighbouring threads hold neighbouring j of the same i: similar run lengths, adjacent 48-byte row segments.
// ---------------------------------------------------------------------------
__device__ __forceinline__ size_t ex_elem(const DevProblem& P, int i, int j) {   // offset of element (0,0) of block (i,j), i <= j
    const int I = i / EX_TB, J = j / EX_TB;
    const int tile = P.ex_tile_of[I * P.ex_NT + J];
    return (size_t)tile * EX_TD * EX_TD + (size_t)(6 * (i - I * EX_TB)) * EX_TD + 6 * (j - J * EX_TB);
}
__global__ void __launch_bounds__(128) k_explicit_assemble(DevProblem P) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= P.ex_nblocks) return;
    const int i = __ldg(P.ex_blk_i + b), j = __ldg(P.ex_blk_j + b);
    const int r0 = __ldg(P.ex_blk_off + b), r1 = r0 + __ldg(P.ex_blk_len + b);
    d4 ti = ld256(P.cam_cache + (size_t)CAM_CACHE_STRIDE * i + 8);   // (Rt8, t0, t1, t2) of the sensor pose
    d4 tj = ld256(P.cam_cache + (size_t)CAM_CACHE_STRIDE * j + 8);
    {   // body t = ts - (ts - t): the second term sits behind Q (zero without offsets / mountings)
        const d4 qi = ld256(P.cam_Q + 12 * (size_t)i + 8), qj = ld256(P.cam_Q + 12 * (size_t)j + 8);
        ti.y -= qi.y; ti.z -= qi.z; ti.w -= qi.w; tj.y -= qj.y; tj.z -= qj.z; tj.w -= qj.w;
    }
    double B[36];
#pragma unroll
    for (int e = 0; e < 36; e++) B[e] = 0.0;
    const double* src = P.ex_contrib + EX_CW * (size_t)r0;
    d4 n0 = ld256(src), n1 = ld256(src + 4), n2 = ld256(src + 8);   // r1 > r0 always
    for (int r = r0; r < r1; r++) {
        const d4 k0 = n0, k1 = n1, k2 = n2;
        src += EX_CW;
        if (r + 1 < r1) { n0 = ld256(src); n1 = ld256(src + 4); n2 = ld256(src + 8); }   // next contribution in flight during the math
        const double K[9] = {k0.x, k0.y, k0.z, k0.w, k1.x, k1.y, k1.z, k1.w, k2.x};
        const double da[3] = {k2.y - ti.y, k2.z - ti.z, k2.w - ti.w};
        const double db[3] = {k2.y - tj.y, k2.z - tj.z, k2.w - tj.w};
        // U = [da]x K : column c of U = da x K[:,c]
        double U[9];
#pragma unroll
        for (int c = 0; c < 3; c++) {
            U[c] = da[1] * K[6 + c] - da[2] * K[3 + c];
            U[3 + c] = da[2] * K[c] - da[0] * K[6 + c];
            U[6 + c] = da[0] * K[3 + c] - da[1] * K[c];
        }
        // M [db]x : row r = M[r,:] x db
#pragma unroll
        for (int q = 0; q < 3; q++) {
            const double u0 = U[3 * q], u1 = U[3 * q + 1], u2 = U[3 * q + 2];
            B[6 * q + 0] -= u1 * db[2] - u2 * db[1]; B[6 * q + 1] -= u2 * db[0] - u0 * db[2]; B[6 * q + 2] -= u0 * db[1] - u1 * db[0];
            B[6 * q + 3] += u0; B[6 * q + 4] += u1; B[6 * q + 5] += u2;
            const double v0 = K[3 * q], v1 = K[3 * q + 1], v2 = K[3 * q + 2];
            B[6 * (3 + q) + 0] -= v1 * db[2] - v2 * db[1]; B[6 * (3 + q) + 1] -= v2 * db[0] - v0 * db[2]; B[6 * (3 + q) + 2] -= v0 * db[1] - v1 * db[0];
            B[6 * (3 + q) + 3] += v0; B[6 * (3 + q) + 4] += v1; B[6 * (3 + q) + 5] += v2;
        }
    }
    double* out = P.ex_S + ex_elem(P, i, j);
    if (P.cam_fixed[i] || P.cam_fixed[j]) {
#pragma unroll
        for (int r = 0; r < 6; r++)
#pragma unroll
            for (int c = 0; c < 6; c += 2) st2(out + (size_t)r * EX_TD + c, 0.0, 0.0);
        return;
    }
    // S_ij = -T_i B T_j^T, T = diag(Q, I): first the left factor on the top three rows, then the right factor on the left
    // three columns
    double Qm[9];
    {
        const double* Qi = P.cam_Q + 12 * (size_t)i;
        const d4 q0 = ld256(Qi), q1 = ld256(Qi + 4), q2 = ld256(Qi + 8);
        Qm[0] = q0.x; Qm[1] = q0.y; Qm[2] = q0.z; Qm[3] = q0.w; Qm[4] = q1.x; Qm[5] = q1.y; Qm[6] = q1.z; Qm[7] = q1.w; Qm[8] = q2.x;
    }
#pragma unroll
    for (int c = 0; c < 6; c++) {
        const double b0 = B[c], b1 = B[6 + c], b2 = B[12 + c];
        B[c] = Qm[0] * b0 + Qm[1] * b1 + Qm[2] * b2;
        B[6 + c] = Qm[3] * b0 + Qm[4] * b1 + Qm[5] * b2;
        B[12 + c] = Qm[6] * b0 + Qm[7] * b1 + Qm[8] * b2;
    }
    {
        const double* Qj = P.cam_Q + 12 * (size_t)j;
        const d4 q0 = ld256(Qj), q1 = ld256(Qj + 4), q2 = ld256(Qj + 8);
        Qm[0] = q0.x; Qm[1] = q0.y; Qm[2] = q0.z; Qm[3] = q0.w; Qm[4] = q1.x; Qm[5] = q1.y; Qm[6] = q1.z; Qm[7] = q1.w; Qm[8] = q2.x;
    }
#pragma unroll
    for (int r = 0; r < 6; r++) {
        const double b0 = B[6 * r], b1 = B[6 * r + 1], b2 = B[6 * r + 2];
        // (B Tj^T)[r][c] = sum_b B[r][b] Qj[c][b]
        B[6 * r] = b0 * Qm[0] + b1 * Qm[1] + b2 * Qm[2];
        B[6 * r + 1] = b0 * Qm[3] + b1 * Qm[4] + b2 * Qm[5];
        B[6 * r + 2] = b0 * Qm[6] + b1 * Qm[7] + b2 * Qm[8];
    }
#pragma unroll
    for (int r = 0; r < 6; r++)
#pragma unroll
        for (int c = 0; c < 6; c += 2) st2(out + (size_t)r * EX_TD + c, -B[6 * r + c], -B[6 * r + c + 1]);
}

// Variant with TWO lanes per block: the even lane sums rows 0..2 of the block (the [di]x K rows), the odd lane rows 3..5
// (the K rows).  Both lanes stream the same run (identical addresses in one warp: one request), each keeps 18 sums instead
// of 36 -- half the registers per thread and twice the threads, i.e. four times the loads in flight of the latency-bound
// one-thread-per-block walk (ncu r2f: 164 registers, 17.5 % warps active).  Every entry is summed in the same order as
// in k_explicit_assemble: bitwise identical tiles.
__device__ __forceinline__ void asm2_add(const d4& k0, const d4& k1, const d4& k2, const d4& ti, const d4& tj, int half, double (&B)[18]) {
    const double K[9] = {k0.x, k0.y, k0.z, k0.w, k1.x, k1.y, k1.z, k1.w, k2.x};
    const double db[3] = {k2.y - tj.y, k2.z - tj.z, k2.w - tj.w};
    double M[9];   // the three rows this lane owns: U = [da]x K (even lane) or K itself (odd lane)
    if (half == 0) {
        const double da[3] = {k2.y - ti.y, k2.z - ti.z, k2.w - ti.w};
#pragma unroll
        for (int c = 0; c < 3; c++) {
            M[c] = da[1] * K[6 + c] - da[2] * K[3 + c];
            M[3 + c] = da[2] * K[c] - da[0] * K[6 + c];
            M[6 + c] = da[0] * K[3 + c] - da[1] * K[c];
        }
    } else {
#pragma unroll
        for (int e = 0; e < 9; e++) M[e] = K[e];
    }
#pragma unroll
    for (int q = 0; q < 3; q++) {
        const double u0 = M[3 * q], u1 = M[3 * q + 1], u2 = M[3 * q + 2];
        B[6 * q + 0] -= u1 * db[2] - u2 * db[1]; B[6 * q + 1] -= u2 * db[0] - u0 * db[2]; B[6 * q + 2] -= u0 * db[1] - u1 * db[0];
        B[6 * q + 3] += u0; B[6 * q + 4] += u1; B[6 * q + 5] += u2;
    }
}
__device__ __forceinline__ d4 shfl_xor1(const d4& v) {
    d4 r;
    r.x = __shfl_xor_sync(FULL, v.x, 1); r.y = __shfl_xor_sync(FULL, v.y, 1); r.z = __shfl_xor_sync(FULL, v.z, 1); r.w = __shfl_xor_sync(FULL, v.w, 1);
    return r;
}
// SPLIT: the two lanes load ALTERNATE contributions of the run (even lane r, r + 2, ..; odd lane r + 1, r + 3, ..) and hand
// them to each other by shuffle: twice the distinct bytes in flight per block for the same registers
template <bool SPLIT, bool DIRECT>
__global__ void __launch_bounds__(128) k_explicit_assemble2(DevProblem P) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int half = t & 1;
    int b = t >> 1;
    const bool live = b < P.ex_nblocks;
    if (!SPLIT && !live) return;
    if (!live) b = P.ex_nblocks - 1;   // SPLIT: whole warps stay for the shuffles (the pair of a dead block is dead too)
    const int i = __ldg(P.ex_blk_i + b), j = __ldg(P.ex_blk_j + b);
    const int r0 = __ldg(P.ex_blk_off + b), r1 = r0 + __ldg(P.ex_blk_len + b);
    d4 ti = ld256(P.cam_cache + (size_t)CAM_CACHE_STRIDE * i + 8);   // (Rt8, t0, t1, t2) of the sensor pose
    d4 tj = ld256(P.cam_cache + (size_t)CAM_CACHE_STRIDE * j + 8);
    {   // body t = ts - (ts - t)
        const d4 qi = ld256(P.cam_Q + 12 * (size_t)i + 8), qj = ld256(P.cam_Q + 12 * (size_t)j + 8);
        ti.y -= qi.y; ti.z -= qi.z; ti.w -= qi.w; tj.y -= qj.y; tj.z -= qj.z; tj.w -= qj.w;
    }
    double B[18];   // rows 3 * half .. 3 * half + 2 of the block
#pragma unroll
    for (int e = 0; e < 18; e++) B[e] = 0.0;
    if (!SPLIT) {
        const double* src = P.ex_contrib + EX_CW * (size_t)r0;
        d4 n0 = ld256(src), n1 = ld256(src + 4), n2 = ld256(src + 8);   // r1 > r0 always
        for (int r = r0; r < r1; r++) {
            const d4 k0 = n0, k1 = n1, k2 = n2;
            src += EX_CW;
            if (r + 1 < r1) { n0 = ld256(src); n1 = ld256(src + 4); n2 = ld256(src + 8); }
            asm2_add(k0, k1, k2, ti, tj, half, B);
        }
    } else {
        // pairs of contributions (r, r + 1): this lane loads r + half; trip count made warp-uniform for the shuffles
        const int npair = (r1 - r0 + 1) >> 1;
        const int nmax = __reduce_max_sync(FULL, npair);
        const double* src = P.ex_contrib + EX_CW * (size_t)(r0 + half);
        d4 n0 = {0, 0, 0, 0}, n1 = n0, n2 = n0;
        if (r0 + half < r1) { n0 = ld256(src); n1 = ld256(src + 4); n2 = ld256(src + 8); }
        for (int k = 0; k < nmax; k++) {
            const int r = r0 + 2 * k;
            const d4 m0 = n0, m1 = n1, m2 = n2;   // contribution r + half
            src += 2 * EX_CW;
            if (r + 2 + half < r1) { n0 = ld256(src); n1 = ld256(src + 4); n2 = ld256(src + 8); }
            const d4 o0 = shfl_xor1(m0), o1 = shfl_xor1(m1), o2 = shfl_xor1(m2);   // the partner's: r + 1 - half
            if (half == 0) {
                if (r < r1) asm2_add(m0, m1, m2, ti, tj, 0, B);
                if (r + 1 < r1) asm2_add(o0, o1, o2, ti, tj, 0, B);
            } else {
                if (r < r1) asm2_add(o0, o1, o2, ti, tj, 1, B);
                if (r + 1 < r1) asm2_add(m0, m1, m2, ti, tj, 1, B);
            }
        }
        if (!live) return;
    }
    // DIRECT: straight into the block's place in its panel (row stride m6) -- no dense tile, no compaction pass
    const size_t ld = DIRECT ? (size_t)__ldg(P.ex_blk_m6 + b) : (size_t)EX_TD;
    double* out = DIRECT ? P.sp_data + __ldg(P.ex_blk_dst + b) + (size_t)(3 * half) * ld : P.ex_S + ex_elem(P, i, j) + (size_t)(3 * half) * EX_TD;
    if (P.cam_fixed[i] || P.cam_fixed[j]) {
#pragma unroll
        for (int r = 0; r < 3; r++)
#pragma unroll
            for (int c = 0; c < 6; c += 2) st2(out + (size_t)r * ld + c, 0.0, 0.0);
        return;
    }
    double Qm[9];
    if (half == 0) {   // left factor Q_i on rows 0..2 (rows 3..5 keep the identity)
        const double* Qi = P.cam_Q + 12 * (size_t)i;
        const d4 q0 = ld256(Qi), q1 = ld256(Qi + 4), q2 = ld256(Qi + 8);
        Qm[0] = q0.x; Qm[1] = q0.y; Qm[2] = q0.z; Qm[3] = q0.w; Qm[4] = q1.x; Qm[5] = q1.y; Qm[6] = q1.z; Qm[7] = q1.w; Qm[8] = q2.x;
#pragma unroll
        for (int c = 0; c < 6; c++) {
            const double b0 = B[c], b1 = B[6 + c], b2 = B[12 + c];
            B[c] = Qm[0] * b0 + Qm[1] * b1 + Qm[2] * b2;
            B[6 + c] = Qm[3] * b0 + Qm[4] * b1 + Qm[5] * b2;
            B[12 + c] = Qm[6] * b0 + Qm[7] * b1 + Qm[8] * b2;
        }
    }
    {
        const double* Qj = P.cam_Q + 12 * (size_t)j;
        const d4 q0 = ld256(Qj), q1 = ld256(Qj + 4), q2 = ld256(Qj + 8);
        Qm[0] = q0.x; Qm[1] = q0.y; Qm[2] = q0.z; Qm[3] = q0.w; Qm[4] = q1.x; Qm[5] = q1.y; Qm[6] = q1.z; Qm[7] = q1.w; Qm[8] = q2.x;
    }
#pragma unroll
    for (int r = 0; r < 3; r++) {
        const double b0 = B[6 * r], b1 = B[6 * r + 1], b2 = B[6 * r + 2];
        B[6 * r] = b0 * Qm[0] + b1 * Qm[1] + b2 * Qm[2];
        B[6 * r + 1] = b0 * Qm[3] + b1 * Qm[4] + b2 * Qm[5];
        B[6 * r + 2] = b0 * Qm[6] + b1 * Qm[7] + b2 * Qm[8];
    }
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 6; c += 2) st2(out + (size_t)r * ld + c, -B[6 * r + c], -B[6 * r + c + 1]);
}

// diagonal blocks: S_ii = Hcc + lam - Sd, stored HALVED in the diagonal tiles (T x + T^T x counts them twice).  Written
// into the OWNED copy of the tile (multi-GPU: after the reduce-scatter of the partial tiles, by the owner only; Hcc and
// Sd are already all-reduced there)
__global__ void k_explicit_diag(DevProblem P) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= P.Nc * 36) return;
    const int c = t / 36, e = t - 36 * c, r = e / 6, cc = e - 6 * r;
    const int I = c / EX_TB;
    const int tile = P.ex_tile_of[I * P.ex_NT + I] - P.ex_own0;
    if (tile < 0 || tile >= P.ex_nown) return;
    const int q = sym6(r, cc);
    double v = P.Hcc[21 * (size_t)c + q] - P.Sd[21 * (size_t)c + q];
    if (r == cc) v += P.lam_c[6 * (size_t)c + r];
    const int l = 6 * (c - I * EX_TB);
    P.ex_Sown[(size_t)tile * EX_TD * EX_TD + (size_t)(l + r) * EX_TD + l + cc] = 0.5 * v;
}

// direct mode: the (halved) diagonal block of every camera into its panel
__global__ void k_explicit_diag_direct(DevProblem P) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= P.Nc * 36) return;
    const int c = t / 36, e = t - 36 * c, r = e / 6, cc = e - 6 * r;
    const int q = sym6(r, cc);
    double v = P.Hcc[21 * (size_t)c + q] - P.Sd[21 * (size_t)c + q];
    if (r == cc) v += P.lam_c[6 * (size_t)c + r];
    P.sp_data[P.ex_dg_dst[c] + (size_t)r * P.ex_dg_m6[c] + cc] = 0.5 * v;
}

void launch_explicit_build(const DevProblem& P, cudaStream_t st) {
    if (!P.ex_on) return;
    if (P.ex_ncontrib) k_explicit_contrib<<<cdiv(P.ex_ncontrib, 256), 256, 0, st>>>(P);
    const char* e = std::getenv("SBA_EX_ASM2");   // measurement knob: two lanes per block
    const int two = e ? std::atoi(e) : 2;   // Venice build (scripts/ab_build.py): 1.155 ms one thread per block, 1.099 two lanes, 1.073 two lanes sharing the loads
    if (P.ex_nblocks) {
        // (the shared-load variant capped at 128 registers for a fourth block per SM spills and is slower: 1.072 ms)
        if (P.ex_direct) k_explicit_assemble2<true, true><<<cdiv(2 * P.ex_nblocks, 128), 128, 0, st>>>(P);
        else if (two == 2) k_explicit_assemble2<true, false><<<cdiv(2 * P.ex_nblocks, 128), 128, 0, st>>>(P);
        else if (two) k_explicit_assemble2<false, false><<<cdiv(2 * P.ex_nblocks, 128), 128, 0, st>>>(P);
        else k_explicit_assemble<<<cdiv(P.ex_nblocks, 128), 128, 0, st>>>(P);
    }
}
void launch_explicit_diag(const DevProblem& P, cudaStream_t st) {
    if (!P.ex_on) return;
    if (P.ex_direct) { k_explicit_diag_direct<<<cdiv(P.Nc * 36, 256), 256, 0, st>>>(P); return; }
    k_explicit_diag<<<cdiv(P.Nc * 36, 256), 256, 0, st>>>(P);
}

// dense tiles -> block-sparse panels of the persistent PCG kernel (sp_on): one warp per panel copies the 6 rows of the
// block row through the panel's column map (48-byte runs of a tile row -> contiguous panel rows)
__global__ void __launch_bounds__(256) k_explicit_compact(DevProblem P) {
    const int q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (q >= P.sp_npanel) return;
    const int4 pn = __ldg(P.sp_panel + q);
    const int m6 = pn.z;
    const uint16_t* stat = P.sp_idx + 8 * (size_t)pn.y;   // header + column map
    const uint16_t* cmap = stat + 8;
    const double* src = P.ex_Sown + (size_t)(pn.w >> 6) * EX_TD * EX_TD + (size_t)(6 * (pn.w & 63)) * EX_TD;
    double* out = P.sp_data + 2 * (size_t)pn.x;
    double* dst = out + 2;
    for (int c = lane; c < m6; c += 32) {
        const int cj = __ldg(cmap + c);
#pragma unroll
        for (int r = 0; r < 6; r++) __stcs(dst + r * m6 + c, __ldcs(src + (size_t)r * EX_TD + cj));
    }
    // 16-byte units of the static part: the header in front, the map behind the values
    const uint4* st4 = reinterpret_cast<const uint4*>(stat);
    if (lane == 0) reinterpret_cast<uint4*>(out)[0] = __ldg(st4);
    uint4* map_out = reinterpret_cast<uint4*>(dst + 6 * m6);
    for (int k = lane; k < (2 * m6 + 15) / 16; k += 32) map_out[k] = __ldg(st4 + 1 + k);
}
// multi-GPU: this rank's slice of the summed tiles, pulled from the partial tiles of all ranks (peer memory over NVLink /
// NVSwitch, own partial from HBM), summed in RANK ORDER -- same operands, same order on every run: deterministic.  Replaces
// ncclReduceScatter on the 479 MB partial tiles; G loads of 32 bytes per thread in flight, two rounds unrolled.
template <int G>
__global__ void __launch_bounds__(256) k_explicit_pull_reduce(DevProblem P, size_t first, size_t n4) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += 2 * stride) {
        const size_t j = i + stride;
        d4 a[G], b[G];
#pragma unroll
        for (int g = 0; g < G; g++) a[g] = ld256(P.ex_peer_S[g] + first + 4 * i);
        if (j < n4) {
#pragma unroll
            for (int g = 0; g < G; g++) b[g] = ld256(P.ex_peer_S[g] + first + 4 * j);
        }
        d4 s = a[0];
#pragma unroll
        for (int g = 1; g < G; g++) { s.x += a[g].x; s.y += a[g].y; s.z += a[g].z; s.w += a[g].w; }
        st256(P.ex_Sown + 4 * i, s.x, s.y, s.z, s.w);
        if (j < n4) {
            d4 u = b[0];
#pragma unroll
            for (int g = 1; g < G; g++) { u.x += b[g].x; u.y += b[g].y; u.z += b[g].z; u.w += b[g].w; }
            st256(P.ex_Sown + 4 * j, u.x, u.y, u.z, u.w);
        }
    }
}
void launch_explicit_pull_reduce(const DevProblem& P, int nranks, int rank, size_t slice_doubles, cudaStream_t st) {
    const size_t n4 = slice_doubles / 4, first = (size_t)rank * slice_doubles;
    const int grid = 148 * 8;
    switch (nranks) {
        case 2: k_explicit_pull_reduce<2><<<grid, 256, 0, st>>>(P, first, n4); break;
        case 3: k_explicit_pull_reduce<3><<<grid, 256, 0, st>>>(P, first, n4); break;
        case 4: k_explicit_pull_reduce<4><<<grid, 256, 0, st>>>(P, first, n4); break;
        case 5: k_explicit_pull_reduce<5><<<grid, 256, 0, st>>>(P, first, n4); break;
        case 6: k_explicit_pull_reduce<6><<<grid, 256, 0, st>>>(P, first, n4); break;
        case 7: k_explicit_pull_reduce<7><<<grid, 256, 0, st>>>(P, first, n4); break;
        case 8: k_explicit_pull_reduce<8><<<grid, 256, 0, st>>>(P, first, n4); break;
        default: break;
    }
}

// the inverse copy (direct mode keeps only the panels up to date): panel values back into their dense tile
__global__ void __launch_bounds__(256) k_explicit_decompact(DevProblem P) {
    const int q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (q >= P.sp_npanel) return;
    const int4 pn = __ldg(P.sp_panel + q);
    const int m6 = pn.z;
    const uint16_t* cmap = P.sp_idx + 8 * (size_t)pn.y + 8;
    double* tile = P.ex_Sown + (size_t)(pn.w >> 6) * EX_TD * EX_TD + (size_t)(6 * (pn.w & 63)) * EX_TD;
    const double* src = P.sp_data + 2 * (size_t)pn.x + 2;
    for (int c = lane; c < m6; c += 32) {
        const int cj = __ldg(cmap + c);
#pragma unroll
        for (int r = 0; r < 6; r++) tile[(size_t)r * EX_TD + cj] = src[r * m6 + c];
    }
}
void launch_explicit_decompact(const DevProblem& P, cudaStream_t st) {
    if (!P.ex_on || !P.sp_on || !P.sp_npanel) return;
    k_explicit_decompact<<<cdiv(P.sp_npanel * 32, 256), 256, 0, st>>>(P);
}
__global__ void __launch_bounds__(256) k_explicit_sparse_dirs(DevProblem P) {
    const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (c >= P.sp_nchunk) return;
    const int4 ch = __ldg(P.sp_chunk + c);
    const uint4* src = reinterpret_cast<const uint4*>(P.sp_idx + 8 * (size_t)ch.z);
    uint4* dst = reinterpret_cast<uint4*>(P.sp_data + 2 * (size_t)ch.x);
    for (int k = lane; k < ch.w / 16; k += 32) dst[k] = __ldg(src + k);
}
void launch_explicit_sparse_dirs(const DevProblem& P, cudaStream_t st) {
    if (!P.sp_on || !P.sp_nchunk) return;
    k_explicit_sparse_dirs<<<cdiv(P.sp_nchunk * 32, 256), 256, 0, st>>>(P);
}
void launch_explicit_compact(const DevProblem& P, cudaStream_t st) {
    if (!P.ex_on || !P.ps_on || !P.sp_on || P.ex_direct) return;   // direct mode: the panels are written by the assembly itself
    k_explicit_compact<<<cdiv(P.sp_npanel * 32, 256), 256, 0, st>>>(P);
}

// ---------------------------------------------------------------------------
// mat-vec on the tiles.  CTA = (tile, EX_CHUNK-row chunk), EX_MVW warps x 16 rows.  Lane l owns the columns l + 32 k
// (k < 12) of the tile: every load instruction of a warp is one coalesced 256-byte run of a tile row, read once
// and used for BOTH products
//   row partial  y_I[r] += sum_c T[r][c] x_J[c]   (8 rows at a time in registers, then ONE split butterfly:
//                                                  9 double shuffles for the 8 sums instead of 40)
//   col partial  y_J[c] += sum_r T[r][c] x_I[r]   (12 register accumulators per lane, no shuffles at all;
//                                                  the four warps are folded through shared memory in warp order).
// Algorithmic bytes: 8 B per stored scalar, i.e. 384*384*8 per tile -- the whole symmetric product from one read.
// ---------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum8(const double (&v)[8], int lane) {
    const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
    double r[4];
#pragma unroll
    for (int i = 0; i < 4; i++) r[i] = (b4 ? v[4 + i] : v[i]) + __shfl_xor_sync(FULL, b4 ? v[i] : v[4 + i], 16);
    double p[2];
#pragma unroll
    for (int i = 0; i < 2; i++) p[i] = (b3 ? r[2 + i] : r[i]) + __shfl_xor_sync(FULL, b3 ? r[i] : r[2 + i], 8);
    double q = (b2 ? p[1] : p[0]) + __shfl_xor_sync(FULL, b2 ? p[0] : p[1], 4);
    q += __shfl_xor_sync(FULL, q, 2);
    q += __shfl_xor_sync(FULL, q, 1);
    return q;   // total of value (lane >> 2) in every lane
}
// pcg != 0: launched inside the PCG batch -- return at once when the solve has already converged / hit its cap (the
// host reads the convergence scalar only every pcg_check_every iterations; same freeze rule as k_pcg_ABC)
__global__ void __launch_bounds__(32 * EX_MVW) k_explicit_matvec(DevProblem P, const double* __restrict__ x, int pcg, double tol2, int max_iters) {
    __shared__ double s_col[EX_MVW][EX_TD];
    if (pcg) {
        const double rz = P.scalars[SC_RZ_NEXT], rz0 = P.scalars[SC_RZ0];
        if (!(rz > tol2 * rz0) || !(rz0 > 0.0) || P.flags[FL_PCG_ITERS_NEXT] >= max_iters) return;   // uniform over the grid
    }
    const int tile = blockIdx.x / EX_NCHUNK, h = blockIdx.x - tile * EX_NCHUNK;   // local tile of this rank
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int I = __ldg(P.ex_tile_I + P.ex_own0 + tile), J = __ldg(P.ex_tile_J + P.ex_own0 + tile);
    const int n = 6 * P.Nc;
    const double* T = P.ex_Sown + (size_t)tile * EX_TD * EX_TD;
    double xJ[12], col[12];
#pragma unroll
    for (int k = 0; k < 12; k++) {
        const int g = EX_TD * J + lane + 32 * k;
        xJ[k] = g < n ? __ldg(x + g) : 0.0;
        col[k] = 0.0;
    }
    const int row0 = EX_CHUNK * h + 16 * warp;
#pragma unroll 1
    for (int g8 = 0; g8 < 2; g8++) {
        double racc[8];
#pragma unroll
        for (int rr = 0; rr < 8; rr++) {
            const int r = row0 + 8 * g8 + rr;
            const int gi = EX_TD * I + r;
            const double xi = gi < n ? __ldg(x + gi) : 0.0;   // warp-uniform
            const double* Tr = T + (size_t)r * EX_TD + lane;
            double acc = 0.0;
#pragma unroll
            for (int k = 0; k < 12; k++) {
                const double v = __ldcs(Tr + 32 * k);
                acc += v * xJ[k];
                col[k] += v * xi;
            }
            racc[rr] = acc;
        }
        const double s = warp_sum8(racc, lane);
        if ((lane & 3) == 0) P.ex_row[(size_t)tile * EX_TD + row0 + 8 * g8 + (lane >> 2)] = s;
    }
#pragma unroll
    for (int k = 0; k < 12; k++) s_col[warp][lane + 32 * k] = col[k];
    __syncthreads();
    for (int c = threadIdx.x; c < EX_TD; c += 32 * EX_MVW) {
        double a = s_col[0][c];
#pragma unroll
        for (int w = 1; w < EX_MVW; w++) a += s_col[w][c];   // warp order: fixed
        P.ex_col[((size_t)tile * EX_NCHUNK + h) * EX_TD + c] = a;
    }
}
void launch_explicit_matvec(const DevProblem& P, const double* x, cudaStream_t st, int pcg, double tol2, int max_iters) {
    if (P.ex_nown) k_explicit_matvec<<<P.ex_nown * EX_NCHUNK, 32 * EX_MVW, 0, st>>>(P, x, pcg, tol2, max_iters);
}

__global__ void __launch_bounds__(128) k_explicit_gather(DevProblem P, double* y) {
    __shared__ int s_off[128];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < 6 * P.Nc;
    const double v = explicit_fold(P, i, valid, s_off);
    if (valid) y[i] = v;
}
void launch_explicit_gather(const DevProblem& P, double* y, cudaStream_t st) {
    k_explicit_gather<<<cdiv(6 * P.Nc, 128), 128, 0, st>>>(P, y);
}

}  // namespace sba
