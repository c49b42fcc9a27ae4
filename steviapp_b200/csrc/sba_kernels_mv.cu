// sba_kernels_mv.cu -- the two passes of the implicit Schur operator and the
// back-substitution (K2c / K4 of SURVEY Appendix C).
#include "sba_kernel_util.cuh"

namespace sba {

// ---------------------------------------------------------------------------
// point pass: thread per point over its 128-byte records (A: bytes 0-47,
// C: bytes 48-95 -> 3 sectors per observation), camera vector gathered (48 B,
// L1/L2 resident).  x must be zero at fixed cameras.
//   mode 0: y_p = Hinv * sum_o A^T (C x_r - A x_t)
//   mode 1: y_p = Hinv * gp
//   mode 2: dp  = -Hinv * (gp + sum_o A^T (C x_r - A x_t))     (back-substitution,
//           block_solver.hpp:455-477 analogue)
// ---------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(256) k_point_pass(DevProblem P, const double* __restrict__ x) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P.Np) return;
    double u0 = 0.0, u1 = 0.0, u2 = 0.0;
    if (MODE != 0) {
        const double* g = P.gp + 4 * (size_t)p;
        const double2 g01 = ldg2(g), g23 = ldg2(g + 2);
        u0 = g01.x; u1 = g01.y; u2 = g23.x;
    }
    if (MODE != 1) {
        const int k0 = __ldg(P.pt_off + p), k1 = __ldg(P.pt_off + p + 1);
        for (int k = k0; k < k1; k++) {
            const double* rec = P.prec + (size_t)PREC * k;
            const double* xc = x + 6 * (size_t)__ldg(P.pos_cam + k);
            const double2 a01 = ldg2(rec), a23 = ldg2(rec + 2), a45 = ldg2(rec + 4);
            const double2 c01 = ldg2(rec + 6), c23 = ldg2(rec + 8), c45 = ldg2(rec + 10);
            const double2 x01 = ldg2(xc), x23 = ldg2(xc + 2), x45 = ldg2(xc + 4);
            // rows of Jc x = C x_r - A x_t
            const double m0 = c01.x * x01.x + c01.y * x01.y + c23.x * x23.x - (a01.x * x23.y + a01.y * x45.x + a23.x * x45.y);
            const double m1 = c23.y * x01.x + c45.x * x01.y + c45.y * x23.x - (a23.y * x23.y + a45.x * x45.x + a45.y * x45.y);
            u0 += a01.x * m0 + a23.y * m1;
            u1 += a01.y * m0 + a45.x * m1;
            u2 += a23.x * m0 + a45.y * m1;
        }
    }
    const double* Hp = P.Hinv + 8 * (size_t)p;
    const double2 h01 = ldg2(Hp), h23 = ldg2(Hp + 2), h45 = ldg2(Hp + 4);
    double y0 = h01.x * u0 + h01.y * u1 + h23.x * u2;
    double y1 = h01.y * u0 + h23.y * u1 + h45.x * u2;
    double y2 = h23.x * u0 + h45.x * u1 + h45.y * u2;
    double* o = (MODE == 2 ? P.dp : P.y_p) + 4 * (size_t)p;
    if (MODE == 2) { y0 = -y0; y1 = -y1; y2 = -y2; }
    st2(o, y0, y1); st2(o + 2, y2, 0.0);
}
void launch_point_pass(const DevProblem& P, const double* x, int mode, cudaStream_t st) {
    if (P.Np == 0) return;
    const int nb = cdiv(P.Np, 256);
    if (mode == 0) k_point_pass<0><<<nb, 256, 0, st>>>(P, x);
    else if (mode == 1) k_point_pass<1><<<nb, 256, 0, st>>>(P, x);
    else k_point_pass<2><<<nb, 256, 0, st>>>(P, x);
}

// ---------------------------------------------------------------------------
// camera pass: warp per tile, sum_o Jc^T (A y_p) with Jc = [C | -A]; reads the
// C and A planes coalesced (96 B/obs) and gathers y_p (one 32 B sector).
// Lane sums folded by a fixed shuffle tree -> tile partial (6 doubles).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32) k_camera_pass(DevProblem P) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * WARPS_PER_BLOCK + warp;
    if (tile >= P.Nt) return;
    const int start = __ldg(P.tile_start + tile), len = __ldg(P.tile_len + tile);
    const size_t Ns = (size_t)P.Ns;
    double w[6] = {0, 0, 0, 0, 0, 0};
    for (int s = start + lane; s < start + len; s += 32) {
        const int pt = __ldg(P.obs_pt + s);
        if (pt < 0) continue;
        const double2 y01 = ldg2(P.y_p + 4 * (size_t)pt), y23 = ldg2(P.y_p + 4 * (size_t)pt + 2);
        double A[6], C[6];
#pragma unroll
        for (int a = 0; a < 6; a++) {
            A[a] = __ldg(P.Ac + (size_t)a * Ns + s);
            C[a] = __ldg(P.Cc + (size_t)a * Ns + s);
        }
        const double b0 = A[0] * y01.x + A[1] * y01.y + A[2] * y23.x;
        const double b1 = A[3] * y01.x + A[4] * y01.y + A[5] * y23.x;
#pragma unroll
        for (int a = 0; a < 3; a++) {
            w[a] += C[a] * b0 + C[3 + a] * b1;
            w[3 + a] -= A[a] * b0 + A[3 + a] * b1;
        }
    }
    double* out = P.tile_part + (size_t)TP * tile;
#pragma unroll
    for (int i = 0; i < 6; i++) { const double v = warp_sum(w[i]); if (lane == 0) out[i] = v; }
}

// thread per (camera, row):  out = -sum_tiles  [+ finish terms].  Fixed cameras: the sum is 0
// (their Jc is zero in the reference's linearisation).
// finish: out += hsign * (Hcc x)_row + damp * lam_c x + base
__device__ __forceinline__ double finish_terms(const DevProblem& P, int c, int a, const double* x, const double* base,
                                               double hsign, int damp) {
    double v = 0.0;
    const size_t i = 6 * (size_t)c + a;
    if (hsign != 0.0) {
        const double* H = P.Hcc + 21 * (size_t)c;
        double hx = 0.0;
#pragma unroll
        for (int b = 0; b < 6; b++) hx += H[sym6(a, b)] * x[6 * (size_t)c + b];
        v += hsign * hx;
    }
    if (damp == 1 || (damp == 2 && P.cam_fixed[c])) v += P.lam_c[i] * x[i];   // 2: rows of fixed cameras only (test hook)
    if (base) v += base[i];
    return v;
}
__global__ void k_camera_reduce(DevProblem P, const double* x, const double* base, double hsign, int damp,
                                int finish, double* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.Nc * 6) return;
    const int c = i / 6, a = i - 6 * c;
    double acc = 0.0;
    if (!P.cam_fixed[c]) {
        const int t0 = P.cam_tile_off[c], t1 = P.cam_tile_off[c + 1];
        for (int t = t0; t < t1; t++) acc += P.tile_part[(size_t)TP * t + a];
    }
    acc = -acc;
    if (finish) acc += finish_terms(P, c, a, x, base, hsign, damp);
    out[i] = acc;
}
__global__ void k_camera_finish(DevProblem P, const double* x, const double* base, double hsign, int damp, double* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.Nc * 6) return;
    const int c = i / 6, a = i - 6 * c;
    out[i] += finish_terms(P, c, a, x, base, hsign, damp);
}
void launch_camera_pass(const DevProblem& P, const double* x, const double* base, double hsign, int damp, int finish,
                        double* out, cudaStream_t st) {
    if (P.Nt) k_camera_pass<<<cdiv(P.Nt, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, 0, st>>>(P);
    k_camera_reduce<<<cdiv(P.Nc * 6, 256), 256, 0, st>>>(P, x, base, hsign, damp, finish, out);
}
void launch_camera_finish(const DevProblem& P, const double* x, const double* base, double hsign, int damp, double* out,
                          cudaStream_t st) {
    k_camera_finish<<<cdiv(P.Nc * 6, 256), 256, 0, st>>>(P, x, base, hsign, damp, out);
}

__global__ void k_zero_fixed(DevProblem P, const double* x, double* xz) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.Nc * 6) return;
    xz[i] = P.cam_fixed[i / 6] ? 0.0 : x[i];
}
void launch_zero_fixed(const DevProblem& P, const double* x, double* xz, cudaStream_t st) {
    k_zero_fixed<<<cdiv(P.Nc * 6, 256), 256, 0, st>>>(P, x, xz);
}

__global__ void k_negate(double* x, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) x[i] = -x[i];
}
void launch_negate(double* x, int n, cudaStream_t st) { k_negate<<<cdiv(n, 256), 256, 0, st>>>(x, n); }

}  // namespace sba
