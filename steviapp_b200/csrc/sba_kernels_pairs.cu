// sba_kernels_pairs.cu -- UV <-> UV cross-projection factors: landmark-free two-view constraints between
// two poses, 4 residuals each (UV2UVCost, costfunctors/modularuvprojection.h:660-795; built by
// ProjectorModule::addCrossProjectionCostFunction, modularsbasolver.cpp:1616-1739; caller
// sbamodules/correspondencessetsbamodule.cpp:1173-1214) under ceres::HuberLoss / CauchyLoss / ArctanLoss(1)
// (correspondencessetsbamodule.cpp:2459-2471).  A factor touches no point, so nothing is eliminated: its Gauss-Newton blocks
// J1^T J1, J2^T J2 go to the camera diagonal blocks, J1^T J2 straight into the (c1, c2) block of the reduced camera system,
// and -- when the intrinsic blocks of the cameras' models are free -- J^T Ji into the border / pseudo-camera blocks of
// sba_kernels_shared.cu.  Jacobians by forward-mode duals in registers (sba_jet.cuh: the functor solves a 3x2 least-squares
// problem per evaluation); the loss enters as Ceres' Corrector does it (residual and Jacobian rescaling).
// Deterministic: every sum runs over a static, ordered incidence list.
#include "sba_jet.cuh"
#include "sba_kernel_util.cuh"

namespace sba {

typedef DJet<18> J18;   // pose1 (6), pose2 (6), view ray 1 (3), view ray 2 (3)

__device__ __forceinline__ int pair_group(const DevProblem& P, int cam) { return P.Ng ? P.cam_grp[2 * cam] : -1; }

// thread per factor.  store: corrected residuals / Jacobians for the linear algebra; always: rho(s) per factor
__global__ void __launch_bounds__(128) k_pair_eval(DevProblem P, int store, double* __restrict__ rho_out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= P.Npair) return;
    const int c1 = P.pair_cam[2 * k], c2 = P.pair_cam[2 * k + 1];
    const double* uv = P.pair_uv + 4 * (size_t)k;
    const double* in1 = P.intr + 10 * (size_t)P.cam_model[c1];
    const double* in2 = P.intr + 10 * (size_t)P.cam_model[c2];
    const double* w1 = P.pair_info ? P.pair_info + 8 * (size_t)k : nullptr;
    const double* w2 = P.pair_info ? w1 + 4 : nullptr;
    double r[4], rho[3];
    if (!store) {
        double d1[3], d2[3];
        dir_from_uv(uv[0], uv[1], in1, d1);
        dir_from_uv(uv[2], uv[3], in2, d2);
        uv2uv_residual<double>(P.cam_r + 3 * c1, P.cam_t + 3 * c1, P.cam_r + 3 * c2, P.cam_t + 3 * c2, d1, d2, w1, w2, r);
        loss_rho(P.pair_loss, P.pair_loss_a, r[0] * r[0] + r[1] * r[1] + r[2] * r[2] + r[3] * r[3], rho);
        rho_out[k] = rho[0];
        return;
    }
    double d1[3], d2[3], dd1[24], dd2[24];
    dir_from_uv_jac(uv[0], uv[1], in1, d1, dd1);
    dir_from_uv_jac(uv[2], uv[3], in2, d2, dd2);
    J18 rp1[3], tp1[3], rp2[3], tp2[3], jd1[3], jd2[3], out[4];
    for (int a = 0; a < 3; a++) {
        rp1[a] = J18(P.cam_r[3 * c1 + a], a); tp1[a] = J18(P.cam_t[3 * c1 + a], 3 + a);
        rp2[a] = J18(P.cam_r[3 * c2 + a], 6 + a); tp2[a] = J18(P.cam_t[3 * c2 + a], 9 + a);
        jd1[a] = J18(d1[a], 12 + a); jd2[a] = J18(d2[a], 15 + a);
    }
    uv2uv_residual<J18>(rp1, tp1, rp2, tp2, jd1, jd2, w1, w2, out);
    for (int i = 0; i < 4; i++) r[i] = out[i].a;
    const double s = r[0] * r[0] + r[1] * r[1] + r[2] * r[2] + r[3] * r[3];
    loss_rho(P.pair_loss, P.pair_loss_a, s, rho);
    rho_out[k] = rho[0];
    double sq1, rs, asn;
    loss_corrector(s, rho, &sq1, &rs, &asn);
    // raw Jacobian columns: 12 pose columns, then 8 + 8 intrinsic columns through the rays (chain rule)
    const int g1 = pair_group(P, c1), g2 = pair_group(P, c2);
    const bool f1 = P.cam_fixed[c1], f2 = P.cam_fixed[c2];
    double J[4][28];
    for (int i = 0; i < 4; i++) {
        for (int a = 0; a < 6; a++) { J[i][a] = f1 ? 0.0 : out[i].v[a]; J[i][6 + a] = f2 ? 0.0 : out[i].v[6 + a]; }
        for (int c = 0; c < 8; c++) {
            J[i][12 + c] = (g1 >= 0 && P.grp_live[8 * g1 + c]) ? out[i].v[12] * dd1[c] + out[i].v[13] * dd1[8 + c] + out[i].v[14] * dd1[16 + c] : 0.0;
            J[i][20 + c] = (g2 >= 0 && P.grp_live[8 * g2 + c]) ? out[i].v[15] * dd2[c] + out[i].v[16] * dd2[8 + c] + out[i].v[17] * dd2[16 + c] : 0.0;
        }
    }
    // Corrector: J' = sqrt(rho') (J - alpha_sq r (r^T J)), r' = scale r
    for (int a = 0; a < 28; a++) {
        const double rj = r[0] * J[0][a] + r[1] * J[1][a] + r[2] * J[2][a] + r[3] * J[3][a];
        for (int i = 0; i < 4; i++) J[i][a] = sq1 * (J[i][a] - asn * r[i] * rj);
    }
    for (int i = 0; i < 4; i++) {
        P.pair_r[4 * (size_t)k + i] = rs * r[i];
        for (int a = 0; a < 12; a++) P.pair_J[(4 * (size_t)k + i) * 12 + a] = J[i][a];
        if (P.pair_Ji) for (int a = 0; a < 16; a++) P.pair_Ji[(4 * (size_t)k + i) * 16 + a] = J[i][12 + a];
    }
}
// sum of rho over the factors (fixed tree), ADDED to the replicated cost scalar (after k_prior_eval on the same stream)
__global__ void __launch_bounds__(RED_BLOCK) k_pair_cost(DevProblem P, const double* __restrict__ rho) {
    __shared__ double sh[33];
    double acc = 0.0;
    for (int k = threadIdx.x; k < P.Npair; k += blockDim.x) acc += rho[k];
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) P.scalars[SC_COST_PRIOR_CAM] += acc;
}
void launch_pair_eval(const DevProblem& P, int store, cudaStream_t st) {
    if (!P.Npair) return;
    k_pair_eval<<<cdiv(P.Npair, 128), 128, 0, st>>>(P, store, P.pair_rho);
    k_pair_cost<<<1, RED_BLOCK, 0, st>>>(P, P.pair_rho);
}

// thread per camera: its incident (factor, side) entries in list order -> Hcc, gc
__global__ void k_pair_camera(DevProblem P) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= P.Nc || P.cam_fixed[c]) return;
    const int e0 = P.pair_inc_off[c], e1 = P.pair_inc_off[c + 1];
    if (e0 == e1) return;
    double H[21], g[6];
    for (int i = 0; i < 21; i++) H[i] = 0.0;
    for (int i = 0; i < 6; i++) g[i] = 0.0;
    for (int e = e0; e < e1; e++) {
        const int k = P.pair_inc[e] >> 1, side = P.pair_inc[e] & 1;
        const double* J = P.pair_J + 48 * (size_t)k + 6 * side;
        const double* r = P.pair_r + 4 * (size_t)k;
        int q = 0;
        for (int a = 0; a < 6; a++) {
            for (int i = 0; i < 4; i++) g[a] += J[12 * i + a] * r[i];
            for (int b = a; b < 6; b++) { double v = 0.0; for (int i = 0; i < 4; i++) v += J[12 * i + a] * J[12 * i + b]; H[q++] += v; }
        }
    }
    for (int i = 0; i < 21; i++) P.Hcc[21 * (size_t)c + i] += H[i];
    for (int i = 0; i < 6; i++) P.gc[6 * (size_t)c + i] += g[i];
}
// free intrinsics: sum_k Ji^T Ji into sh_H, Ji^T r into sh_g (thread per entry, factors in order), Jc^T Ji per camera into pair_border
__global__ void k_pair_shared(DevProblem P) {
    const int id = blockIdx.x * blockDim.x + threadIdx.x;
    const int nH = P.Ng * P.Ng * 64, ng8 = P.Ng * 8;
    if (id < nH + ng8) {
        const bool isg = id >= nH;
        const int ga = isg ? (id - nH) / 8 : id / (P.Ng * 64), gb = isg ? -1 : (id / 64) % P.Ng;
        const int ca = isg ? (id - nH) % 8 : (id % 64) / 8, cb = isg ? -1 : id % 8;
        double acc = 0.0;
        for (int k = 0; k < P.Npair; k++) {
            const int g1 = pair_group(P, P.pair_cam[2 * k]), g2 = pair_group(P, P.pair_cam[2 * k + 1]);
            const double* Ji = P.pair_Ji + 64 * (size_t)k;
            const double* r = P.pair_r + 4 * (size_t)k;
            for (int sa = 0; sa < 2; sa++) {
                if ((sa ? g2 : g1) != ga) continue;
                if (isg) { for (int i = 0; i < 4; i++) acc += Ji[16 * i + 8 * sa + ca] * r[i]; continue; }
                for (int sb = 0; sb < 2; sb++) {
                    if ((sb ? g2 : g1) != gb) continue;
                    for (int i = 0; i < 4; i++) acc += Ji[16 * i + 8 * sa + ca] * Ji[16 * i + 8 * sb + cb];
                }
            }
        }
        if (isg) P.sh_g[id - nH] += acc; else P.sh_H[id] += acc;
    }
}
__global__ void k_pair_border(DevProblem P) {
    const int id = blockIdx.x * blockDim.x + threadIdx.x;   // (group, camera)
    if (id >= P.Ng * P.Nc) return;
    const int g = id / P.Nc, c = id - g * P.Nc;
    double B[48];
    for (int i = 0; i < 48; i++) B[i] = 0.0;
    if (!P.cam_fixed[c])
        for (int e = P.pair_inc_off[c]; e < P.pair_inc_off[c + 1]; e++) {
            const int k = P.pair_inc[e] >> 1, side = P.pair_inc[e] & 1;
            const double* J = P.pair_J + 48 * (size_t)k + 6 * side;
            const double* Ji = P.pair_Ji + 64 * (size_t)k;
            for (int sb = 0; sb < 2; sb++) {
                if (pair_group(P, P.pair_cam[2 * k + sb]) != g) continue;
                for (int a = 0; a < 6; a++)
                    for (int cc = 0; cc < 8; cc++) { double v = 0.0; for (int i = 0; i < 4; i++) v += J[12 * i + a] * Ji[16 * i + 8 * sb + cc]; B[8 * a + cc] += v; }
            }
        }
    for (int i = 0; i < 48; i++) P.pair_border[48 * (size_t)id + i] = B[i];
}
void launch_pair_camera_blocks(const DevProblem& P, cudaStream_t st) {
    if (!P.Npair) return;
    k_pair_camera<<<cdiv(P.Nc, 128), 128, 0, st>>>(P);
    if (P.Ng && P.pair_Ji) {
        k_pair_shared<<<cdiv(P.Ng * P.Ng * 64 + P.Ng * 8, 128), 128, 0, st>>>(P);
        k_pair_border<<<cdiv(P.Ng * P.Nc, 64), 64, 0, st>>>(P);
    }
}

// thread per distinct camera pair (i < j): sum of J_i^T J_j over its factors in list order -> block (i, j) of the tiles
// (added to the assembled point contributions when the block has some, assigned otherwise)
__global__ void k_pair_blocks(DevProblem P) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= P.Npblk) return;
    const int i = P.pblk_i[b], j = P.pblk_j[b];
    double B[36];
    for (int e = 0; e < 36; e++) B[e] = 0.0;
    for (int e = P.pblk_off[b]; e < P.pblk_off[b + 1]; e++) {
        const int k = P.pblk_list[e] >> 1, swapped = P.pblk_list[e] & 1;   // swapped: the factor's second camera is i
        const double* Ja = P.pair_J + 48 * (size_t)k + (swapped ? 6 : 0);
        const double* Jb = P.pair_J + 48 * (size_t)k + (swapped ? 0 : 6);
        for (int a = 0; a < 6; a++)
            for (int c = 0; c < 6; c++) { double v = 0.0; for (int r = 0; r < 4; r++) v += Ja[12 * r + a] * Jb[12 * r + c]; B[6 * a + c] += v; }
    }
    const int I = i / EX_TB, J = j / EX_TB;
    double* out = P.ex_S + (size_t)P.ex_tile_of[I * P.ex_NT + J] * EX_TD * EX_TD + (size_t)(6 * (i - I * EX_TB)) * EX_TD + 6 * (j - J * EX_TB);
    const bool add = P.pblk_assembled[b] != 0;
    for (int a = 0; a < 6; a++)
        for (int c = 0; c < 6; c++) out[(size_t)a * EX_TD + c] = (add ? out[(size_t)a * EX_TD + c] : 0.0) + B[6 * a + c];
}
void launch_pair_blocks(const DevProblem& P, cudaStream_t st) {
    if (P.Npair && P.Npblk) k_pair_blocks<<<cdiv(P.Npblk, 128), 128, 0, st>>>(P);
}

// model cost change of the factors, -sum m.(r' + m/2) with m = J1' dc1 + J2' dc2 + Ji' d(intrinsics): ADDED to the replicated
// scalar k_model_reduce has just written
__global__ void __launch_bounds__(RED_BLOCK) k_pair_model(DevProblem P, const double* __restrict__ x) {
    __shared__ double sh[33];
    double acc = 0.0;
    for (int k = threadIdx.x; k < P.Npair; k += blockDim.x) {
        const int c1 = P.pair_cam[2 * k], c2 = P.pair_cam[2 * k + 1];
        const double* J = P.pair_J + 48 * (size_t)k;
        for (int i = 0; i < 4; i++) {
            double m = 0.0;
            for (int a = 0; a < 6; a++) m += J[12 * i + a] * x[6 * (size_t)c1 + a] + J[12 * i + 6 + a] * x[6 * (size_t)c2 + a];
            if (P.Ng && P.pair_Ji) {
                const double* Ji = P.pair_Ji + 64 * (size_t)k + 16 * i;
                for (int sb = 0; sb < 2; sb++) {
                    const int g = pair_group(P, sb ? c2 : c1);
                    if (g < 0) continue;
                    const int base = 6 * P.grp_blk[g];
                    for (int c = 0; c < 8; c++) if (P.grp_live[8 * g + c]) m += Ji[8 * sb + c] * x[base + c];
                }
            }
            acc += m * (P.pair_r[4 * (size_t)k + i] + 0.5 * m);
        }
    }
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) P.scalars[SC_MODEL_PRIOR_CAM] -= acc;
}
void launch_pair_model(const DevProblem& P, const double* x, cudaStream_t st) {
    if (P.Npair) k_pair_model<<<1, RED_BLOCK, 0, st>>>(P, x);
}

// inspection hook: raw residuals (4) and raw pose Jacobians (4 x 12) of every factor, before the loss
__global__ void __launch_bounds__(128) k_pair_raw(DevProblem P, double* res, double* jac) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= P.Npair) return;
    const int c1 = P.pair_cam[2 * k], c2 = P.pair_cam[2 * k + 1];
    const double* uv = P.pair_uv + 4 * (size_t)k;
    const double* w1 = P.pair_info ? P.pair_info + 8 * (size_t)k : nullptr;
    double d1[3], d2[3];
    dir_from_uv(uv[0], uv[1], P.intr + 10 * (size_t)P.cam_model[c1], d1);
    dir_from_uv(uv[2], uv[3], P.intr + 10 * (size_t)P.cam_model[c2], d2);
    typedef DJet<12> J12;
    J12 rp1[3], tp1[3], rp2[3], tp2[3], jd1[3], jd2[3], out[4];
    for (int a = 0; a < 3; a++) {
        rp1[a] = J12(P.cam_r[3 * c1 + a], a); tp1[a] = J12(P.cam_t[3 * c1 + a], 3 + a);
        rp2[a] = J12(P.cam_r[3 * c2 + a], 6 + a); tp2[a] = J12(P.cam_t[3 * c2 + a], 9 + a);
        jd1[a] = J12(d1[a]); jd2[a] = J12(d2[a]);
    }
    uv2uv_residual<J12>(rp1, tp1, rp2, tp2, jd1, jd2, w1, w1 ? w1 + 4 : nullptr, out);
    for (int i = 0; i < 4; i++) {
        res[4 * (size_t)k + i] = out[i].a;
        for (int a = 0; a < 12; a++) jac[(4 * (size_t)k + i) * 12 + a] = out[i].v[a];
    }
}
void launch_pair_raw(const DevProblem& P, double* res, double* jac, cudaStream_t st) {
    if (P.Npair) k_pair_raw<<<cdiv(P.Npair, 128), 128, 0, st>>>(P, res, jac);
}

}  // namespace sba
