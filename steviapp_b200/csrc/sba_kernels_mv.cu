// sba_kernels_mv.cu -- pass 2 (camera order) of the implicit Schur operator; pass 1 and the
// back-substitution live in sba_kernels_points.cu.
#include "sba_kernel_util.cuh"

namespace sba {

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
    const CamRQ crq = load_cam_rq(P, __ldg(P.tile_cam + tile));
    double w[6] = {0, 0, 0, 0, 0, 0};
    for (int s = start + lane; s < start + len; s += 32) {
        const int pt = __ldcs(P.obs_pt + s);
        if (pt < 0) continue;
        const d4 y = ld256(P.y_p + 4 * (size_t)pt);
        double A[6], C[6];
#pragma unroll
        for (int a = 0; a < 6; a++) A[a] = __ldcs(P.Ac + (size_t)a * Ns + s);
        rebuild_C(A, __ldcs(P.Pc + s), __ldcs(P.Pc + Ns + s), __ldcs(P.Pc + 2 * Ns + s), crq, C);
        const double b0 = A[0] * y.x + A[1] * y.y + A[2] * y.z;
        const double b1 = A[3] * y.x + A[4] * y.y + A[5] * y.z;
#pragma unroll
        for (int a = 0; a < 3; a++) {
            w[a] += C[a] * b0 + C[3 + a] * b1;
            w[3 + a] -= A[a] * b0 + A[3 + a] * b1;
        }
    }
    double* out = P.mv_part + 6 * (size_t)tile;
#pragma unroll
    for (int i = 0; i < 6; i++) { const double v = warp_sum(w[i]); if (lane == 0) out[i] = v; }
}

// ---------------------------------------------------------------------------
// camera pass, FACTORED form (identity stiffness): with A = Pi R^T and C = Pi [Pbar]x Jr,
//   Jc^T (A y) = [ -Jr^T (Pbar x n) ; -R n ],   n = Pi^T Pi (R^T y),
// so a lane only needs Pbar (24 B, coalesced planes) [+ 16 B scale constants on spherical cameras],
// the point index and the gathered y_p; Pi is rebuilt in registers (pi_from_pbar) and the per-camera
// factors Jr^T, R are applied once per tile by lane 0.  60 B/obs instead of 132 B/obs.
// ---------------------------------------------------------------------------
// The pass is latency-bound (dependent index -> gather chain), not byte-bound, so every lane first issues the
// index loads of a batch of CPF_BATCH slots, then all their y_p gathers and Pbar loads, and only then does the math:
// CPF_BATCH independent gathers in flight per lane instead of one.
template <int CPF_BATCH>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32, CPF_BATCH >= 4 ? 3 : (CPF_BATCH == 2 ? 4 : 6)) k_camera_pass_f(DevProblem P) {
    __shared__ double s_rt[WARPS_PER_BLOCK][10];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * WARPS_PER_BLOCK + warp;
    if (tile >= P.Nt) return;
    const int cam = __ldg(P.tile_cam + tile);
    const int start = __ldg(P.tile_start + tile), len = __ldg(P.tile_len + tile);
    const size_t Ns = (size_t)P.Ns;
    const double* cc = P.cam_cache + (size_t)CAM_CACHE_STRIDE * cam;
    if (lane < 9) s_rt[warp][lane] = __ldg(cc + lane);
    if (lane == 9) s_rt[warp][9] = __ldg(cc + 21);   // f of the camera's model (written by K0): d_z of the plane cost
    __syncwarp();
    const double* Rt = s_rt[warp];
    const int spherical = P.cam_kind[cam];
    const double f = Rt[9];
    double S1[3] = {0, 0, 0}, S2[3] = {0, 0, 0};
    for (int base = start + lane; base < start + len; base += 32 * CPF_BATCH) {
        int pt[CPF_BATCH];
#pragma unroll
        for (int i = 0; i < CPF_BATCH; i++) {
            const int s = base + 32 * i;
            pt[i] = s < start + len ? __ldcs(P.obs_pt + s) : -1;
        }
        d4 y[CPF_BATCH];
        double Pb[CPF_BATCH][3], kk[CPF_BATCH][2];
#pragma unroll
        for (int i = 0; i < CPF_BATCH; i++) {
            const int s = base + 32 * i;
            if (pt[i] >= 0) {
                y[i] = ld256(P.y_p + 4 * (size_t)pt[i]);   // one 256-bit gather per observation
                Pb[i][0] = __ldcs(P.Pc + s); Pb[i][1] = __ldcs(P.Pc + Ns + s); Pb[i][2] = __ldcs(P.Pc + 2 * Ns + s);
                kk[i][0] = f; kk[i][1] = f;
                if (spherical) { kk[i][0] = __ldcs(P.Kc + s); kk[i][1] = __ldcs(P.Kc + Ns + s); }
            }
        }
#pragma unroll
        for (int i = 0; i < CPF_BATCH; i++) {
            if (pt[i] < 0) continue;
            const double Px = Pb[i][0], Py = Pb[i][1], Pz = Pb[i][2];
            if (Px == 0.0 && Py == 0.0 && Pz == 0.0) continue;   // failed row (zero Jacobian)
            double p00, p02, p11, p12;
            pi_from_pbar(Px, Py, Pz, kk[i][0], kk[i][1], spherical, &p00, &p02, &p11, &p12);
            const double r0 = Rt[0] * y[i].x + Rt[1] * y[i].y + Rt[2] * y[i].z;
            const double r1 = Rt[3] * y[i].x + Rt[4] * y[i].y + Rt[5] * y[i].z;
            const double r2 = Rt[6] * y[i].x + Rt[7] * y[i].y + Rt[8] * y[i].z;
            const double b0 = p00 * r0 + p02 * r2, b1 = p11 * r1 + p12 * r2;      // b = A y
            const double n0 = p00 * b0, n1 = p11 * b1, n2 = p02 * b0 + p12 * b1;  // n = Pi^T b
            S2[0] += n0; S2[1] += n1; S2[2] += n2;
            S1[0] += Py * n2 - Pz * n1; S1[1] += Pz * n0 - Px * n2; S1[2] += Px * n1 - Py * n0;   // Pbar x n
        }
    }
#pragma unroll
    for (int i = 0; i < 3; i++) { S1[i] = warp_sum(S1[i]); S2[i] = warp_sum(S2[i]); }
    if (lane == 0) {
        double* out = P.mv_part + 6 * (size_t)tile;
        // with Rs = Rt^T (sensor pose), Q = (R Jr)^T (body pose) and dt = ts - t:
        //   C^T b summed = -Q (Rs S1 + dt x (Rs S2)) ;  -A^T b summed = -Rs S2
        // (without offsets / mountings dt = 0 and Q Rs = Jr^T: the form -Jr^T S1 of the plain projector)
        const double* Q = P.cam_Q + 12 * (size_t)cam;
        double u[3], v[3];
#pragma unroll
        for (int a = 0; a < 3; a++) {
            u[a] = Rt[a] * S1[0] + Rt[3 + a] * S1[1] + Rt[6 + a] * S1[2];
            v[a] = Rt[a] * S2[0] + Rt[3 + a] * S2[1] + Rt[6 + a] * S2[2];
        }
        const double d0 = __ldg(Q + 9), d1 = __ldg(Q + 10), d2 = __ldg(Q + 11);
        u[0] += d1 * v[2] - d2 * v[1]; u[1] += d2 * v[0] - d0 * v[2]; u[2] += d0 * v[1] - d1 * v[0];
#pragma unroll
        for (int a = 0; a < 3; a++) {
            out[a] = -(__ldg(Q + 3 * a) * u[0] + __ldg(Q + 3 * a + 1) * u[1] + __ldg(Q + 3 * a + 2) * u[2]);
            out[3 + a] = -v[a];
        }
    }
}

// thread per (camera, row):  out = -sum of the camera's tile partials  [+ finish terms].  Fixed cameras: the sum is 0
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
        acc = sum_partials(P.mv_part, P.cam_tile_off[c], P.cam_tile_off[c + 1], a);
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
void launch_camera_pass_tiles(const DevProblem& P, cudaStream_t st) {
    if (!P.Nt) return;
    if (P.cp_factored) {
        const int nb = cdiv(P.Nt, WARPS_PER_BLOCK);
        if (P.cp_variant == 1) k_camera_pass_f<4><<<nb, WARPS_PER_BLOCK * 32, 0, st>>>(P);
        else if (P.cp_variant == 2) k_camera_pass_f<1><<<nb, WARPS_PER_BLOCK * 32, 0, st>>>(P);
        else k_camera_pass_f<2><<<nb, WARPS_PER_BLOCK * 32, 0, st>>>(P);   // measured best on B200 (Venice shape)
    }
    else k_camera_pass<<<cdiv(P.Nt, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, 0, st>>>(P);
}
void launch_camera_pass(const DevProblem& P, const double* x, const double* base, double hsign, int damp, int finish,
                        double* out, cudaStream_t st) {
    launch_camera_pass_tiles(P, st);
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
