/*
 * sba_oracle.cpp -- CPU ORACLE (test infrastructure, never shipped).
 *
 * Restates, with no third-party dependency, the arithmetic of steviapp's SBA
 * hot path.  Each function cites the reference lines it follows.  Jacobians
 * are produced the way the reference produces them: forward-mode automatic
 * differentiation (a small Jet type standing in for ceres::Jet) of the SAME
 * expression tree as the reference functors.  The product's CUDA kernels use
 * closed-form derivatives instead, which makes this an independent check.
 *
 * PARITY STATUS: functors pinned by the reference's known-answer vectors
 * (tests/test_oracle_golden.py); LM/Schur/PCG = restatement of Ceres' published
 * algorithm, "parity unpinned" (see sba_oracle.h).
 *
 * External code restated here (absent from /root/reference):
 *   - LibStevi StereoVision::Geometry @ 8781d6b784726195c9363e357278ba2cd9c63c51
 *     (CMakeLists.txt:104-109): RigidBodyTransform, angle-axis rotation,
 *     projectPoints, radial/tangential/inverseSkew distortion.  Formulas are
 *     anchored on in-tree evidence: edges/edgeparametrizedxyz2uv.cpp:195-229
 *     (analytic Jacobian of the same lens model) and the call sites cited below.
 *   - Ceres Solver (find_package(Ceres), CMakeLists.txt:131): trust-region
 *     Levenberg-Marquardt, Schur elimination, SCHUR_JACOBI-preconditioned CG.
 */
#include "sba_oracle.h"
#include "sba_oracle_functors.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace oracle_detail;

namespace {

// One observation: residual (2), Jc (2x6 row-major), Jp (2x3).  Autodiff over
// the 9 free scalars (r, t, X); intrinsic blocks constant
// (FixedParameter::CameraInternal, pinholecameraprojectormodule.cpp:358-361).
inline bool eval_obs(const oracle_problem* p, int o, double* res, double* Jc, double* Jp) {
    typedef Jet<9> J9;
    const int c = p->obs_cam[o], q = p->obs_pt[o];
    const double* in = p->intrinsics + ORACLE_INTR_STRIDE * p->cam_model[c];
    const int sph = p->cam_cost_kind ? p->cam_cost_kind[c] : 0;
    const double* info = p->obs_info ? p->obs_info + 4 * o : nullptr;
    const double* uv = p->obs_uv + 2 * o;
    if (!Jc && !Jp) {
        double r[3], t[3], x[3], f[1], pp[2], ks[3], ts[2], Bs[2];
        for (int i = 0; i < 3; i++) { r[i] = p->cam_r[3 * c + i]; t[i] = p->cam_t[3 * c + i]; x[i] = p->pts[3 * q + i]; }
        f[0] = in[0]; pp[0] = in[1]; pp[1] = in[2]; ks[0] = in[3]; ks[1] = in[4]; ks[2] = in[5];
        ts[0] = in[6]; ts[1] = in[7]; Bs[0] = in[8]; Bs[1] = in[9];
        return uv_cost<double>(r, t, x, f, pp, ks, ts, Bs, uv, info, sph, res);
    }
    J9 r[3], t[3], x[3], f[1], pp[2], ks[3], ts[2], Bs[2], out[2];
    for (int i = 0; i < 3; i++) {
        r[i] = J9(p->cam_r[3 * c + i], i);
        t[i] = J9(p->cam_t[3 * c + i], 3 + i);
        x[i] = J9(p->pts[3 * q + i], 6 + i);
    }
    f[0] = J9(in[0]); pp[0] = J9(in[1]); pp[1] = J9(in[2]);
    ks[0] = J9(in[3]); ks[1] = J9(in[4]); ks[2] = J9(in[5]);
    ts[0] = J9(in[6]); ts[1] = J9(in[7]); Bs[0] = J9(in[8]); Bs[1] = J9(in[9]);
    if (!uv_cost<J9>(r, t, x, f, pp, ks, ts, Bs, uv, info, sph, out)) return false;
    for (int k = 0; k < 2; k++) {
        res[k] = out[k].a;
        if (Jc) for (int j = 0; j < 6; j++) Jc[6 * k + j] = out[k].v[j];
        if (Jp) for (int j = 0; j < 3; j++) Jp[3 * k + j] = out[k].v[6 + j];
    }
    return true;
}

// ---------------------------------------------------------------------------
// Linearised system around the current point
// ---------------------------------------------------------------------------
struct System {
    const oracle_problem* p;
    int nt;
    std::vector<int> cam_off, cam_obs, pt_off, pt_obs;       // CSR by camera / by point (stable order)
    std::vector<double> res, Jc, Jp;                          // per observation
    std::vector<double> pres;                                 // 3 per prior
    std::vector<double> Hcc, gc, Hpp, gp;                     // undamped blocks (36/6/9/3)
    std::vector<double> Hinv;                                 // damped Hpp^-1 (9)
    std::vector<double> Sdiag_inv;                            // 36 per camera
    std::vector<double> lam_c, lam_p;                         // LM damping diagonals
    std::vector<double> scale_c, scale_p;                     // jacobi scaling
    std::vector<double> ytmp;
    std::vector<std::vector<int>> cam_priors, pt_priors;

    void build_index() {
        const int Nc = p->n_cams, Np = p->n_pts, No = p->n_obs;
        cam_off.assign(Nc + 1, 0); pt_off.assign(Np + 1, 0);
        for (int o = 0; o < No; o++) { cam_off[p->obs_cam[o] + 1]++; pt_off[p->obs_pt[o] + 1]++; }
        for (int c = 0; c < Nc; c++) cam_off[c + 1] += cam_off[c];
        for (int q = 0; q < Np; q++) pt_off[q + 1] += pt_off[q];
        cam_obs.resize(No); pt_obs.resize(No);
        std::vector<int> cc(cam_off.begin(), cam_off.end() - 1), pc(pt_off.begin(), pt_off.end() - 1);
        for (int o = 0; o < No; o++) { cam_obs[cc[p->obs_cam[o]]++] = o; pt_obs[pc[p->obs_pt[o]]++] = o; }
        cam_priors.assign(Nc, {}); pt_priors.assign(Np, {});
        for (int i = 0; i < p->n_priors; i++) {
            if (p->prior_kind[i] == 0) pt_priors[p->prior_index[i]].push_back(i);
            else cam_priors[p->prior_index[i]].push_back(i);
        }
        res.resize(2 * (size_t)No); Jc.resize(12 * (size_t)No); Jp.resize(6 * (size_t)No);
        pres.resize(3 * (size_t)p->n_priors);
        Hcc.resize(36 * (size_t)Nc); gc.resize(6 * (size_t)Nc); Hpp.resize(9 * (size_t)Np); gp.resize(3 * (size_t)Np);
        Hinv.resize(9 * (size_t)Np); Sdiag_inv.resize(36 * (size_t)Nc);
        lam_c.assign(6 * (size_t)Nc, 0.0); lam_p.assign(3 * (size_t)Np, 0.0);
        scale_c.assign(6 * (size_t)Nc, 1.0); scale_p.assign(3 * (size_t)Np, 1.0);
        ytmp.resize(3 * (size_t)Np);
    }
    bool cam_is_fixed(int c) const { return p->cam_fixed && p->cam_fixed[c]; }
    bool pt_is_fixed(int q) const { return p->pt_fixed && p->pt_fixed[q]; }
    const double* prior_x(int i) const {
        const int k = p->prior_kind[i], idx = p->prior_index[i];
        return k == 0 ? p->pts + 3 * idx : (k == 1 ? p->cam_r + 3 * idx : p->cam_t + 3 * idx);
    }

    // cost only (residual evaluation); returns false on functor failure
    bool cost_only(double* cost) const {
        const int No = p->n_obs;
        double sum = 0.0; int bad = 0;
#pragma omp parallel for num_threads(nt) reduction(+ : sum, bad) schedule(static)
        for (int o = 0; o < No; o++) {
            double r[2];
            if (!eval_obs(p, o, r, nullptr, nullptr)) { bad++; continue; }
            sum += r[0] * r[0] + r[1] * r[1];
        }
        for (int i = 0; i < p->n_priors; i++) {
            const double* A = p->prior_A + 9 * i; const double* b = p->prior_b + 3 * i; const double* x = prior_x(i);
            for (int k = 0; k < 3; k++) {
                double r = 0; for (int j = 0; j < 3; j++) r += A[3 * k + j] * (x[j] - b[j]);   // fixedsizenormalprior.h:89-96
                sum += r * r;
            }
        }
        *cost = 0.5 * sum;
        return bad == 0;
    }

    // residuals + Jacobians + block-diagonal normal equations
    bool linearize(double* cost) {
        const int No = p->n_obs, Nc = p->n_cams, Np = p->n_pts;
        int bad = 0;
#pragma omp parallel for num_threads(nt) reduction(+ : bad) schedule(static)
        for (int o = 0; o < No; o++) {
            if (!eval_obs(p, o, &res[2 * (size_t)o], &Jc[12 * (size_t)o], &Jp[6 * (size_t)o])) { bad++; continue; }
            if (cam_is_fixed(p->obs_cam[o])) for (int j = 0; j < 12; j++) Jc[12 * (size_t)o + j] = 0.0;
            if (pt_is_fixed(p->obs_pt[o])) for (int j = 0; j < 6; j++) Jp[6 * (size_t)o + j] = 0.0;
        }
        if (bad) return false;
        double sum = 0.0;
        for (size_t i = 0; i < res.size(); i++) sum += res[i] * res[i];
        for (int i = 0; i < p->n_priors; i++) {
            const double* A = p->prior_A + 9 * i; const double* b = p->prior_b + 3 * i; const double* x = prior_x(i);
            for (int k = 0; k < 3; k++) {
                double r = 0; for (int j = 0; j < 3; j++) r += A[3 * k + j] * (x[j] - b[j]);
                pres[3 * i + k] = r; sum += r * r;
            }
        }
        *cost = 0.5 * sum;
#pragma omp parallel for num_threads(nt) schedule(dynamic, 16)
        for (int c = 0; c < Nc; c++) {
            double* H = &Hcc[36 * (size_t)c]; double* g = &gc[6 * (size_t)c];
            std::fill(H, H + 36, 0.0); std::fill(g, g + 6, 0.0);
            for (int k = cam_off[c]; k < cam_off[c + 1]; k++) {
                const int o = cam_obs[k]; const double* J = &Jc[12 * (size_t)o]; const double* r = &res[2 * (size_t)o];
                for (int a = 0; a < 6; a++) {
                    g[a] += J[a] * r[0] + J[6 + a] * r[1];
                    for (int b = 0; b < 6; b++) H[6 * a + b] += J[a] * J[b] + J[6 + a] * J[6 + b];
                }
            }
            if (!cam_is_fixed(c)) for (int i : cam_priors[c]) {
                const double* A = p->prior_A + 9 * i; const int off = p->prior_kind[i] == 1 ? 0 : 3;
                for (int a = 0; a < 3; a++) {
                    for (int k = 0; k < 3; k++) g[off + a] += A[3 * k + a] * pres[3 * i + k];
                    for (int b = 0; b < 3; b++) for (int k = 0; k < 3; k++) H[6 * (off + a) + off + b] += A[3 * k + a] * A[3 * k + b];
                }
            }
        }
#pragma omp parallel for num_threads(nt) schedule(static)
        for (int q = 0; q < Np; q++) {
            double* H = &Hpp[9 * (size_t)q]; double* g = &gp[3 * (size_t)q];
            std::fill(H, H + 9, 0.0); std::fill(g, g + 3, 0.0);
            for (int k = pt_off[q]; k < pt_off[q + 1]; k++) {
                const int o = pt_obs[k]; const double* J = &Jp[6 * (size_t)o]; const double* r = &res[2 * (size_t)o];
                for (int a = 0; a < 3; a++) {
                    g[a] += J[a] * r[0] + J[3 + a] * r[1];
                    for (int b = 0; b < 3; b++) H[3 * a + b] += J[a] * J[b] + J[3 + a] * J[3 + b];
                }
            }
            if (!pt_is_fixed(q)) for (int i : pt_priors[q]) {
                const double* A = p->prior_A + 9 * i;
                for (int a = 0; a < 3; a++) {
                    for (int k = 0; k < 3; k++) g[a] += A[3 * k + a] * pres[3 * i + k];
                    for (int b = 0; b < 3; b++) for (int k = 0; k < 3; k++) H[3 * a + b] += A[3 * k + a] * A[3 * k + b];
                }
            }
        }
        return true;
    }

    // Ceres jacobi_scaling [external]: computed once at iteration 0,
    // scale_j = 1 / (1 + ||J col j||).
    void compute_scaling(bool enabled) {
        const int Nc = p->n_cams, Np = p->n_pts;
        for (int c = 0; c < Nc; c++) for (int a = 0; a < 6; a++)
            scale_c[6 * (size_t)c + a] = enabled ? 1.0 / (1.0 + std::sqrt(Hcc[36 * (size_t)c + 7 * a])) : 1.0;
        for (int q = 0; q < Np; q++) for (int a = 0; a < 3; a++)
            scale_p[3 * (size_t)q + a] = enabled ? 1.0 / (1.0 + std::sqrt(Hpp[9 * (size_t)q + 4 * a])) : 1.0;
    }

    // Ceres LevenbergMarquardtStrategy [external]: D^2 = clamp(diag(J~^T J~), min, max) / radius
    // in the scaled variables; expressed in the unscaled variables the damping is D^2 / scale^2.
    void compute_damping(double radius, double dmin, double dmax) {
        const int Nc = p->n_cams, Np = p->n_pts;
        for (int c = 0; c < Nc; c++) for (int a = 0; a < 6; a++) {
            const double s = scale_c[6 * (size_t)c + a];
            const double d = std::min(std::max(Hcc[36 * (size_t)c + 7 * a] * s * s, dmin), dmax);
            lam_c[6 * (size_t)c + a] = d / radius / (s * s);
        }
        for (int q = 0; q < Np; q++) for (int a = 0; a < 3; a++) {
            const double s = scale_p[3 * (size_t)q + a];
            const double d = std::min(std::max(Hpp[9 * (size_t)q + 4 * a] * s * s, dmin), dmax);
            lam_p[3 * (size_t)q + a] = d / radius / (s * s);
        }
    }

    // per-point damped inverse, SchurEliminator e-block step [external]; in-tree
    // analogue modules/g2o/g2o/core/block_solver.hpp:371-400.
    bool invert_points() {
        const int Np = p->n_pts; int bad = 0;
#pragma omp parallel for num_threads(nt) reduction(+ : bad) schedule(static)
        for (int q = 0; q < Np; q++) {
            double H[9];
            for (int i = 0; i < 9; i++) H[i] = Hpp[9 * (size_t)q + i];
            for (int a = 0; a < 3; a++) H[4 * a] += lam_p[3 * (size_t)q + a];
            if (pt_is_fixed(q)) { for (int i = 0; i < 9; i++) Hinv[9 * (size_t)q + i] = 0.0; continue; }
            if (!inv3_sym(H, &Hinv[9 * (size_t)q])) bad++;
        }
        return bad == 0;
    }

    // reduced rhs  gt = gc - sum_p W_p Hpp^-1 gp   (block_solver.hpp:420-447)
    void reduced_rhs(double* gt) {
        const int Nc = p->n_cams, Np = p->n_pts;
#pragma omp parallel for num_threads(nt) schedule(static)
        for (int q = 0; q < Np; q++) {
            const double* Hi = &Hinv[9 * (size_t)q]; const double* g = &gp[3 * (size_t)q];
            for (int a = 0; a < 3; a++) ytmp[3 * (size_t)q + a] = Hi[3 * a] * g[0] + Hi[3 * a + 1] * g[1] + Hi[3 * a + 2] * g[2];
        }
#pragma omp parallel for num_threads(nt) schedule(dynamic, 16)
        for (int c = 0; c < Nc; c++) {
            double acc[6]; for (int a = 0; a < 6; a++) acc[a] = gc[6 * (size_t)c + a];
            for (int k = cam_off[c]; k < cam_off[c + 1]; k++) {
                const int o = cam_obs[k]; const double* J = &Jc[12 * (size_t)o]; const double* B = &Jp[6 * (size_t)o];
                const double* y = &ytmp[3 * (size_t)p->obs_pt[o]];
                const double b0 = B[0] * y[0] + B[1] * y[1] + B[2] * y[2], b1 = B[3] * y[0] + B[4] * y[1] + B[5] * y[2];
                for (int a = 0; a < 6; a++) acc[a] -= J[a] * b0 + J[6 + a] * b1;
            }
            for (int a = 0; a < 6; a++) gt[6 * (size_t)c + a] = acc[a];
        }
    }

    // y = S x  (implicit Schur operator)
    void schur_matvec(const double* x, double* y) {
        const int Nc = p->n_cams, Np = p->n_pts;
#pragma omp parallel for num_threads(nt) schedule(static)
        for (int q = 0; q < Np; q++) {
            double u[3] = {0, 0, 0};
            for (int k = pt_off[q]; k < pt_off[q + 1]; k++) {
                const int o = pt_obs[k]; const double* J = &Jc[12 * (size_t)o]; const double* B = &Jp[6 * (size_t)o];
                const double* xc = x + 6 * (size_t)p->obs_cam[o];
                double a0 = 0, a1 = 0;
                for (int a = 0; a < 6; a++) { a0 += J[a] * xc[a]; a1 += J[6 + a] * xc[a]; }
                for (int a = 0; a < 3; a++) u[a] += B[a] * a0 + B[3 + a] * a1;
            }
            const double* Hi = &Hinv[9 * (size_t)q];
            for (int a = 0; a < 3; a++) ytmp[3 * (size_t)q + a] = Hi[3 * a] * u[0] + Hi[3 * a + 1] * u[1] + Hi[3 * a + 2] * u[2];
        }
#pragma omp parallel for num_threads(nt) schedule(dynamic, 16)
        for (int c = 0; c < Nc; c++) {
            const double* H = &Hcc[36 * (size_t)c]; const double* xc = x + 6 * (size_t)c;
            double acc[6];
            for (int a = 0; a < 6; a++) {
                double s = lam_c[6 * (size_t)c + a] * xc[a];
                for (int b = 0; b < 6; b++) s += H[6 * a + b] * xc[b];
                acc[a] = s;
            }
            for (int k = cam_off[c]; k < cam_off[c + 1]; k++) {
                const int o = cam_obs[k]; const double* J = &Jc[12 * (size_t)o]; const double* B = &Jp[6 * (size_t)o];
                const double* yy = &ytmp[3 * (size_t)p->obs_pt[o]];
                const double b0 = B[0] * yy[0] + B[1] * yy[1] + B[2] * yy[2], b1 = B[3] * yy[0] + B[4] * yy[1] + B[5] * yy[2];
                for (int a = 0; a < 6; a++) acc[a] -= J[a] * b0 + J[6 + a] * b1;
            }
            for (int a = 0; a < 6; a++) y[6 * (size_t)c + a] = acc[a];
        }
    }

    // inverse of the 6x6 diagonal blocks of S: Ceres SCHUR_JACOBI [external]
    // (6x6 pose blocks); g2o analogue solvers/pcg/linear_solver_pcg.hpp:100-104.
    bool schur_diag_inverse() {
        const int Nc = p->n_cams; int bad = 0;
#pragma omp parallel for num_threads(nt) reduction(+ : bad) schedule(dynamic, 16)
        for (int c = 0; c < Nc; c++) {
            double D[36];
            for (int i = 0; i < 36; i++) D[i] = Hcc[36 * (size_t)c + i];
            for (int a = 0; a < 6; a++) D[7 * a] += lam_c[6 * (size_t)c + a];
            for (int k = cam_off[c]; k < cam_off[c + 1]; k++) {
                const int o = cam_obs[k]; const double* J = &Jc[12 * (size_t)o]; const double* B = &Jp[6 * (size_t)o];
                const double* Hi = &Hinv[9 * (size_t)p->obs_pt[o]];
                double BH[6];
                for (int r = 0; r < 2; r++) for (int a = 0; a < 3; a++)
                    BH[3 * r + a] = B[3 * r] * Hi[a] + B[3 * r + 1] * Hi[3 + a] + B[3 * r + 2] * Hi[6 + a];
                double T[4];
                for (int r = 0; r < 2; r++) for (int s = 0; s < 2; s++)
                    T[2 * r + s] = BH[3 * r] * B[3 * s] + BH[3 * r + 1] * B[3 * s + 1] + BH[3 * r + 2] * B[3 * s + 2];
                for (int a = 0; a < 6; a++) {
                    const double t0 = J[a] * T[0] + J[6 + a] * T[2], t1 = J[a] * T[1] + J[6 + a] * T[3];
                    for (int b = 0; b < 6; b++) D[6 * a + b] -= t0 * J[b] + t1 * J[6 + b];
                }
            }
            if (!inv6_spd(D, &Sdiag_inv[36 * (size_t)c])) bad++;
        }
        return bad == 0;
    }

    // dense S (small scenes only), for the Cholesky validation path
    void dense_schur(std::vector<double>& S) {
        const int Nc = p->n_cams, Np = p->n_pts; const size_t n = 6 * (size_t)Nc;
        S.assign(n * n, 0.0);
        for (int c = 0; c < Nc; c++) for (int a = 0; a < 6; a++) {
            for (int b = 0; b < 6; b++) S[(6 * c + a) * n + 6 * c + b] = Hcc[36 * (size_t)c + 6 * a + b];
            S[(6 * c + a) * n + 6 * c + a] += lam_c[6 * (size_t)c + a];
        }
        for (int q = 0; q < Np; q++) {
            const double* Hi = &Hinv[9 * (size_t)q];
            for (int k1 = pt_off[q]; k1 < pt_off[q + 1]; k1++) for (int k2 = pt_off[q]; k2 < pt_off[q + 1]; k2++) {
                const int o1 = pt_obs[k1], o2 = pt_obs[k2];
                const int c1 = p->obs_cam[o1], c2 = p->obs_cam[o2];
                double W1[18], W2[18];   // 6x3
                for (int a = 0; a < 6; a++) for (int b = 0; b < 3; b++) {
                    W1[3 * a + b] = Jc[12 * (size_t)o1 + a] * Jp[6 * (size_t)o1 + b] + Jc[12 * (size_t)o1 + 6 + a] * Jp[6 * (size_t)o1 + 3 + b];
                    W2[3 * a + b] = Jc[12 * (size_t)o2 + a] * Jp[6 * (size_t)o2 + b] + Jc[12 * (size_t)o2 + 6 + a] * Jp[6 * (size_t)o2 + 3 + b];
                }
                for (int a = 0; a < 6; a++) {
                    double WH[3];
                    for (int b = 0; b < 3; b++) WH[b] = W1[3 * a] * Hi[b] + W1[3 * a + 1] * Hi[3 + b] + W1[3 * a + 2] * Hi[6 + b];
                    for (int b = 0; b < 6; b++)
                        S[(6 * c1 + a) * n + 6 * c2 + b] -= WH[0] * W2[3 * b] + WH[1] * W2[3 * b + 1] + WH[2] * W2[3 * b + 2];
                }
            }
        }
    }
};

double dot(const std::vector<double>& a, const std::vector<double>& b) {
    double s = 0; for (size_t i = 0; i < a.size(); i++) s += a[i] * b[i]; return s;
}

double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // namespace

extern "C" {

int oracle_num_threads(void) { return nthreads(0); }

// Upper bound of the OpenMP team for every later call (overrides an OMP_NUM_THREADS inherited from a launcher: torchrun
// exports OMP_NUM_THREADS=1 to its workers).  n <= 0: number of online processors.
int oracle_set_max_threads(int n) {
#ifdef _OPENMP
    if (n <= 0) n = omp_get_num_procs();
    omp_set_num_threads(n);
    return omp_get_max_threads();
#else
    (void)n;
    return 1;
#endif
}

void oracle_default_options(oracle_options* o) {
    // set by the reference: modularsbasolver.cpp:897-924 (tolerances 1e-8 / 1e-10);
    // everything else Ceres defaults [external].
    o->max_iterations = 50;
    o->function_tolerance = 1e-8;
    o->parameter_tolerance = 1e-10;
    o->gradient_tolerance = 1e-10;
    o->initial_radius = 1e4;
    o->max_radius = 1e16;
    o->min_radius = 1e-32;
    o->min_lm_diagonal = 1e-6;
    o->max_lm_diagonal = 1e32;
    o->min_relative_decrease = 1e-3;
    o->jacobi_scaling = 1;
    o->pcg_rel_tol = 1e-10;
    o->pcg_max_iters = 500;
    o->linear_solver = 0;
    o->num_threads = 1;   // ceres::Solver::Options::num_threads is never set by the reference
}

void oracle_dir_from_uv(const double* uv, const double* in, double* dir3) {
    dir_from_uv<double>(uv, in + 0, in + 1, in + 3, in + 6, in + 8, dir3);
}

int oracle_uv_from_dir(const double* xn, const double* in, double* uv) {
    // Newton on g(uv) = dirFromUV(uv).xy / f - xn, Jacobian by Jet<2> over uv
    // is not available (uv is a constant of the functor), so use Jets on the
    // normalised coordinates instead: solve corrected(x,y) = xn, then apply skew.
    typedef Jet<2> J2;
    const double f = in[0], ppx = in[1], ppy = in[2];
    const double* ks = in + 3; const double* ts = in + 6;
    double x = xn[0], y = xn[1];
    for (int it = 0; it < 50; it++) {
        J2 X(x, 0), Y(y, 1);
        J2 r2 = X * X + Y * Y;
        J2 rad = r2 * (ks[0] + r2 * (ks[1] + r2 * ks[2]));
        J2 xy = X * Y;
        J2 cx = X + X * rad + 2.0 * xy * ts[0] + (r2 + 2.0 * X * X) * ts[1];
        J2 cy = Y + Y * rad + (r2 + 2.0 * Y * Y) * ts[0] + 2.0 * xy * ts[1];
        const double e0 = cx.a - xn[0], e1 = cy.a - xn[1];
        const double det = cx.v[0] * cy.v[1] - cx.v[1] * cy.v[0];
        if (!(std::fabs(det) > 1e-300)) return 1;
        const double dx = (cy.v[1] * e0 - cx.v[1] * e1) / det, dy = (-cy.v[0] * e0 + cx.v[0] * e1) / det;
        x -= dx; y -= dy;
        if (std::fabs(dx) + std::fabs(dy) < 1e-16 * (1.0 + std::fabs(x) + std::fabs(y))) break;
    }
    // forward skew with B := ts (the quirk): u = (f + B1) x + B2 y + ppx, v = f y + ppy
    uv[0] = (f + ts[0]) * x + ts[1] * y + ppx;
    uv[1] = f * y + ppy;
    return 0;
}

// test hooks on the LibStevi restatements, pinned by the reference's own vectors
// (tests/test_geometry/testRotations.cpp:25-78, tests/test_functional_utils/main.cpp:198-218)
// PinholePushbroomUVProjector::dirFromUV, sbamodules/pinholecameraprojectormodule.h:86-110 (params f[1], pp[1], as[6], bs[6])
void oracle_pushbroom_dir(const double* uv, double sensor_width, double f, double pp, const double* as, const double* bs, double* dir3) {
    const double s = (uv[0] - sensor_width / 2.0) / sensor_width;
    const double s2 = s * s, s3 = s2 * s, s4 = s3 * s, s5 = s4 * s;
    const double du = as[0] + as[1] * s + as[2] * s2 + as[3] * s3 + as[4] * s4 + as[5] * s5;
    const double dv = bs[0] + bs[1] * s + bs[2] * s2 + bs[3] * s3 + bs[4] * s4 + bs[5] * s5;
    dir3[0] = uv[0] - pp + du; dir3[1] = uv[1] + dv; dir3[2] = f;
}
void oracle_angle_axis_rotate(const double* r, const double* x, double* out) { angle_axis_rotate<double>(r, x, out); }
void oracle_rigid_inverse(const double* r, const double* t, double* ri, double* ti) { rigid_inverse<double>(r, t, ri, ti); }

int oracle_evaluate(const oracle_problem* p, int num_threads, double* cost, double* residuals,
                    double* jac_cam, double* jac_pt, uint8_t* failed) {
    const int nt = nthreads(num_threads);
    const int No = p->n_obs;
    double sum = 0.0; int bad = 0;
#pragma omp parallel for num_threads(nt) reduction(+ : sum, bad) schedule(static)
    for (int o = 0; o < No; o++) {
        double r[2] = {0, 0}, Jc[12], Jp[6];
        const bool ok = eval_obs(p, o, r, jac_cam ? Jc : nullptr, jac_pt ? Jp : nullptr);
        if (failed) failed[o] = ok ? 0 : 1;
        if (!ok) { bad++; r[0] = r[1] = 0; std::fill(Jc, Jc + 12, 0.0); std::fill(Jp, Jp + 6, 0.0); }
        sum += r[0] * r[0] + r[1] * r[1];
        if (residuals) { residuals[2 * (size_t)o] = r[0]; residuals[2 * (size_t)o + 1] = r[1]; }
        if (jac_cam) std::memcpy(jac_cam + 12 * (size_t)o, Jc, sizeof(Jc));
        if (jac_pt) std::memcpy(jac_pt + 6 * (size_t)o, Jp, sizeof(Jp));
    }
    for (int i = 0; i < p->n_priors; i++) {
        const double* A = p->prior_A + 9 * i; const double* b = p->prior_b + 3 * i;
        const int k = p->prior_kind[i], idx = p->prior_index[i];
        const double* x = k == 0 ? p->pts + 3 * idx : (k == 1 ? p->cam_r + 3 * idx : p->cam_t + 3 * idx);
        for (int a = 0; a < 3; a++) { double r = 0; for (int j = 0; j < 3; j++) r += A[3 * a + j] * (x[j] - b[j]); sum += r * r; }
    }
    if (cost) *cost = 0.5 * sum;
    return bad ? 1 : 0;
}

int oracle_evaluate_intrinsics_jac(const oracle_problem* p, double* jac) {
    typedef Jet<10> J10;
    for (int o = 0; o < p->n_obs; o++) {
        const int c = p->obs_cam[o], q = p->obs_pt[o];
        const double* in = p->intrinsics + ORACLE_INTR_STRIDE * p->cam_model[c];
        const int sph = p->cam_cost_kind ? p->cam_cost_kind[c] : 0;
        J10 r[3], t[3], x[3], par[10], out[2];
        for (int i = 0; i < 3; i++) { r[i] = J10(p->cam_r[3 * c + i]); t[i] = J10(p->cam_t[3 * c + i]); x[i] = J10(p->pts[3 * q + i]); }
        for (int i = 0; i < 10; i++) par[i] = J10(in[i], i);
        const bool ok = uv_cost<J10>(r, t, x, par, par + 1, par + 3, par + 6, par + 8, p->obs_uv + 2 * o,
                                     p->obs_info ? p->obs_info + 4 * o : nullptr, sph, out);
        for (int k = 0; k < 2; k++) for (int j = 0; j < 10; j++) jac[20 * (size_t)o + 10 * k + j] = ok ? out[k].v[j] : 0.0;
    }
    return 0;
}

// n independent samples of the lever-arm projection cost: outputs res[2n], jac_cam[12n] (d/dr d/dt), jac_la[12n]
// (d/d boresight, d/d leverArm), jac_pt[6n], failed[n]; duals over the 15 free scalars
int oracle_eval_leverarm(int32_t n, const double* r, const double* t, const double* b, const double* l, const double* r_off,
                         const double* t_off, const double* pts, const double* intr10, const double* uv, const double* info,
                         int spherical, double* res, double* jac_cam, double* jac_la, double* jac_pt, uint8_t* failed) {
    typedef Jet<15> J;
    int bad = 0;
    for (int32_t o = 0; o < n; o++) {
        J jr[3], jt[3], jb[3], jl[3], x[3], f[1], pp[2], ks[3], ts[2], Bs[2], out[2];
        for (int i = 0; i < 3; i++) {
            jr[i] = J(r[3 * o + i], i); jt[i] = J(t[3 * o + i], 3 + i); jb[i] = J(b[3 * o + i], 6 + i); jl[i] = J(l[3 * o + i], 9 + i);
            x[i] = J(pts[3 * o + i], 12 + i);
        }
        f[0] = J(intr10[0]); pp[0] = J(intr10[1]); pp[1] = J(intr10[2]);
        ks[0] = J(intr10[3]); ks[1] = J(intr10[4]); ks[2] = J(intr10[5]);
        ts[0] = J(intr10[6]); ts[1] = J(intr10[7]); Bs[0] = J(intr10[8]); Bs[1] = J(intr10[9]);
        const bool ok = uv_cost_leverarm<J>(jr, jt, jb, jl, r_off + 3 * o, t_off + 3 * o, x, f, pp, ks, ts, Bs, uv + 2 * o,
                                            info ? info + 4 * o : nullptr, spherical, out);
        if (failed) failed[o] = ok ? 0 : 1;
        if (!ok) bad++;
        for (int k = 0; k < 2; k++) {
            res[2 * o + k] = ok ? out[k].a : 0.0;
            for (int j = 0; j < 6; j++) jac_cam[12 * (size_t)o + 6 * k + j] = ok ? out[k].v[j] : 0.0;
            for (int j = 0; j < 6; j++) jac_la[12 * (size_t)o + 6 * k + j] = ok ? out[k].v[6 + j] : 0.0;
            for (int j = 0; j < 3; j++) jac_pt[6 * (size_t)o + 3 * k + j] = ok ? out[k].v[12 + j] : 0.0;
        }
    }
    return bad;
}

// n independent samples of the PoseTransform (constant sensor offset) projection cost, SURVEY 8(f) row 3: per sample pose
// (r, t), offset (r_off, t_off), point, pixel, optional 2x2 stiffness; one intrinsics record.  Outputs: res[2n], jac_cam[12n]
// (rows d/dr(3) d/dt(3)), jac_pt[6n], failed[n].  Duals over the 9 free scalars, as AutoDiffCostFunction<..., 2, 3, 3, 3, ...>.
int oracle_eval_offset(int32_t n, const double* r, const double* t, const double* r_off, const double* t_off, const double* pts,
                       const double* intr10, const double* uv, const double* info, int spherical,
                       double* res, double* jac_cam, double* jac_pt, uint8_t* failed) {
    typedef Jet<9> J9;
    int bad = 0;
    for (int32_t o = 0; o < n; o++) {
        J9 jr[3], jt[3], x[3], f[1], pp[2], ks[3], ts[2], Bs[2], out[2];
        for (int i = 0; i < 3; i++) { jr[i] = J9(r[3 * o + i], i); jt[i] = J9(t[3 * o + i], 3 + i); x[i] = J9(pts[3 * o + i], 6 + i); }
        f[0] = J9(intr10[0]); pp[0] = J9(intr10[1]); pp[1] = J9(intr10[2]);
        ks[0] = J9(intr10[3]); ks[1] = J9(intr10[4]); ks[2] = J9(intr10[5]);
        ts[0] = J9(intr10[6]); ts[1] = J9(intr10[7]); Bs[0] = J9(intr10[8]); Bs[1] = J9(intr10[9]);
        const bool ok = uv_cost_offset<J9>(jr, jt, r_off + 3 * o, t_off + 3 * o, x, f, pp, ks, ts, Bs, uv + 2 * o,
                                           info ? info + 4 * o : nullptr, spherical, out);
        if (failed) failed[o] = ok ? 0 : 1;
        if (!ok) bad++;
        for (int k = 0; k < 2; k++) {
            res[2 * o + k] = ok ? out[k].a : 0.0;
            for (int j = 0; j < 6; j++) jac_cam[12 * (size_t)o + 6 * k + j] = ok ? out[k].v[j] : 0.0;
            for (int j = 0; j < 3; j++) jac_pt[6 * (size_t)o + 3 * k + j] = ok ? out[k].v[6 + j] : 0.0;
        }
    }
    return bad;
}

int oracle_normal_blocks(const oracle_problem* p, int num_threads, double* gc, double* gp, double* Hcc, double* Hpp) {
    System s; s.p = p; s.nt = nthreads(num_threads); s.build_index();
    double cost;
    if (!s.linearize(&cost)) return 1;
    if (gc) std::copy(s.gc.begin(), s.gc.end(), gc);
    if (gp) std::copy(s.gp.begin(), s.gp.end(), gp);
    if (Hcc) std::copy(s.Hcc.begin(), s.Hcc.end(), Hcc);
    if (Hpp) std::copy(s.Hpp.begin(), s.Hpp.end(), Hpp);
    return 0;
}

int oracle_schur_matvec(const oracle_problem* p, int num_threads, const double* lam_c, const double* lam_p,
                        const double* x, double* y, double* rhs) {
    System s; s.p = p; s.nt = nthreads(num_threads); s.build_index();
    double cost;
    if (!s.linearize(&cost)) return 1;
    std::copy(lam_c, lam_c + 6 * (size_t)p->n_cams, s.lam_c.begin());
    std::copy(lam_p, lam_p + 3 * (size_t)p->n_pts, s.lam_p.begin());
    if (!s.invert_points()) return 2;
    if (y) s.schur_matvec(x, y);
    if (rhs) s.reduced_rhs(rhs);
    return 0;
}

// Trust-region Levenberg-Marquardt as Ceres' TrustRegionMinimizer +
// LevenbergMarquardtStrategy run it [external; restated from the published
// algorithm, options as set at modularsbasolver.cpp:897-924].  In-tree reading
// for the same family: modules/g2o/g2o/core/optimization_algorithm_levenberg.cpp:57-160.
int oracle_solve(oracle_problem* p, const oracle_options* opt, oracle_summary* sum) {
    const double t_start = now_s();
    std::memset(sum, 0, sizeof(*sum));
    System s; s.p = p; s.nt = nthreads(opt->num_threads); s.build_index();
    const int Nc = p->n_cams, Np = p->n_pts, No = p->n_obs;
    const size_t nc6 = 6 * (size_t)Nc, np3 = 3 * (size_t)Np;

    double cost;
    double t0 = now_s();
    if (!s.linearize(&cost)) { sum->status = -1; return -1; }   // initial evaluation failed
    sum->seconds_evaluate += now_s() - t0;
    s.compute_scaling(opt->jacobi_scaling != 0);
    sum->initial_cost = cost;
    sum->cost_history[0] = cost;
    double radius = opt->initial_radius, decrease_factor = 2.0;
    sum->radius_history[0] = radius;

    auto grad_max = [&]() {
        double m = 0;
        for (int c = 0; c < Nc; c++) if (!s.cam_is_fixed(c)) for (int a = 0; a < 6; a++) m = std::max(m, std::fabs(s.gc[6 * (size_t)c + a]));
        for (int q = 0; q < Np; q++) if (!s.pt_is_fixed(q)) for (int a = 0; a < 3; a++) m = std::max(m, std::fabs(s.gp[3 * (size_t)q + a]));
        return m;
    };
    if (grad_max() <= opt->gradient_tolerance) { sum->status = 1; sum->final_cost = cost; return 0; }

    std::vector<double> gt(nc6), dc(nc6), dp(np3), r(nc6), z(nc6), pv(nc6), qv(nc6);
    std::vector<double> r0(3 * (size_t)Nc), t0v(3 * (size_t)Nc), x0(np3);
    int invalid_in_a_row = 0;
    int it = 0;
    sum->status = 0;
    while (it < opt->max_iterations) {
        it++;
        double tl = now_s();
        s.compute_damping(radius, opt->min_lm_diagonal, opt->max_lm_diagonal);
        bool lin_ok = s.invert_points();
        int pcg_it = 0;
        if (lin_ok) {
            s.reduced_rhs(gt.data());
            for (size_t i = 0; i < nc6; i++) gt[i] = -gt[i];   // solve S dc = -gt
            if (opt->linear_solver == 1) {
                std::vector<double> S; s.dense_schur(S);
                lin_ok = chol(S.data(), (int)nc6);
                if (lin_ok) { dc = gt; chol_solve(S.data(), (int)nc6, dc.data()); }
            } else {
                lin_ok = s.schur_diag_inverse();
                if (lin_ok) {
                    // block-Jacobi PCG, x0 = 0 (linear_solver_pcg.hpp:84-165 analogue)
                    std::fill(dc.begin(), dc.end(), 0.0);
                    r = gt;
                    auto precond = [&](const std::vector<double>& in, std::vector<double>& out) {
                        for (int c = 0; c < Nc; c++) for (int a = 0; a < 6; a++) {
                            double v = 0; for (int b = 0; b < 6; b++) v += s.Sdiag_inv[36 * (size_t)c + 6 * a + b] * in[6 * (size_t)c + b];
                            out[6 * (size_t)c + a] = v;
                        }
                    };
                    precond(r, z);
                    pv = z;
                    double rz = dot(r, z);
                    const double rz0 = rz;
                    const double thresh = opt->pcg_rel_tol * opt->pcg_rel_tol * rz0;
                    while (pcg_it < opt->pcg_max_iters && rz > thresh && rz0 > 0.0) {
                        s.schur_matvec(pv.data(), qv.data());
                        const double alpha = rz / dot(pv, qv);
                        for (size_t i = 0; i < nc6; i++) { dc[i] += alpha * pv[i]; r[i] -= alpha * qv[i]; }
                        precond(r, z);
                        const double rz_new = dot(r, z);
                        const double beta = rz_new / rz;
                        for (size_t i = 0; i < nc6; i++) pv[i] = z[i] + beta * pv[i];
                        rz = rz_new;
                        pcg_it++;
                    }
                }
            }
        }
        sum->pcg_iterations_total += pcg_it;
        double model_change = 0.0;
        double step2 = 0.0, x2 = 0.0;
        if (lin_ok) {
            for (int c = 0; c < Nc; c++) if (s.cam_is_fixed(c)) for (int a = 0; a < 6; a++) dc[6 * (size_t)c + a] = 0.0;
            // back substitution dp = -Hpp^-1 (gp + W^T dc)   (block_solver.hpp:455-477)
#pragma omp parallel for num_threads(s.nt) schedule(static)
            for (int q = 0; q < Np; q++) {
                double u[3] = {s.gp[3 * (size_t)q], s.gp[3 * (size_t)q + 1], s.gp[3 * (size_t)q + 2]};
                for (int k = s.pt_off[q]; k < s.pt_off[q + 1]; k++) {
                    const int o = s.pt_obs[k]; const double* J = &s.Jc[12 * (size_t)o]; const double* B = &s.Jp[6 * (size_t)o];
                    const double* xc = &dc[6 * (size_t)p->obs_cam[o]];
                    double a0 = 0, a1 = 0;
                    for (int a = 0; a < 6; a++) { a0 += J[a] * xc[a]; a1 += J[6 + a] * xc[a]; }
                    for (int a = 0; a < 3; a++) u[a] += B[a] * a0 + B[3 + a] * a1;
                }
                const double* Hi = &s.Hinv[9 * (size_t)q];
                for (int a = 0; a < 3; a++) dp[3 * (size_t)q + a] = -(Hi[3 * a] * u[0] + Hi[3 * a + 1] * u[1] + Hi[3 * a + 2] * u[2]);
            }
            // model_cost_change = -m.(r + m/2), m = J delta  (Ceres TrustRegionMinimizer [external])
            double mc = 0.0;
#pragma omp parallel for num_threads(s.nt) reduction(+ : mc) schedule(static)
            for (int o = 0; o < No; o++) {
                const double* J = &s.Jc[12 * (size_t)o]; const double* B = &s.Jp[6 * (size_t)o];
                const double* xc = &dc[6 * (size_t)p->obs_cam[o]]; const double* xp = &dp[3 * (size_t)p->obs_pt[o]];
                double m0 = 0, m1 = 0;
                for (int a = 0; a < 6; a++) { m0 += J[a] * xc[a]; m1 += J[6 + a] * xc[a]; }
                for (int a = 0; a < 3; a++) { m0 += B[a] * xp[a]; m1 += B[3 + a] * xp[a]; }
                mc += m0 * (s.res[2 * (size_t)o] + 0.5 * m0) + m1 * (s.res[2 * (size_t)o + 1] + 0.5 * m1);
            }
            for (int i = 0; i < p->n_priors; i++) {
                const int k = p->prior_kind[i], idx = p->prior_index[i];
                if (k == 0 ? s.pt_is_fixed(idx) : s.cam_is_fixed(idx)) continue;
                const double* d = k == 0 ? &dp[3 * (size_t)idx] : (k == 1 ? &dc[6 * (size_t)idx] : &dc[6 * (size_t)idx + 3]);
                const double* A = p->prior_A + 9 * i;
                for (int a = 0; a < 3; a++) {
                    const double m = A[3 * a] * d[0] + A[3 * a + 1] * d[1] + A[3 * a + 2] * d[2];
                    mc += m * (s.pres[3 * i + a] + 0.5 * m);
                }
            }
            model_change = -mc;
            for (int c = 0; c < Nc; c++) if (!s.cam_is_fixed(c)) for (int a = 0; a < 3; a++) {
                step2 += dc[6 * (size_t)c + a] * dc[6 * (size_t)c + a] + dc[6 * (size_t)c + 3 + a] * dc[6 * (size_t)c + 3 + a];
                x2 += p->cam_r[3 * c + a] * p->cam_r[3 * c + a] + p->cam_t[3 * c + a] * p->cam_t[3 * c + a];
            }
            for (int q = 0; q < Np; q++) if (!s.pt_is_fixed(q)) for (int a = 0; a < 3; a++) {
                step2 += dp[3 * (size_t)q + a] * dp[3 * (size_t)q + a];
                x2 += p->pts[3 * (size_t)q + a] * p->pts[3 * (size_t)q + a];
            }
        }
        sum->seconds_linear += now_s() - tl;
        sum->pcg_iters_history[std::min(it, ORACLE_MAX_HISTORY - 1)] = pcg_it;

        bool step_valid = lin_ok && model_change > 0.0 && std::isfinite(model_change);
        bool accepted = false;
        if (!step_valid) {
            if (++invalid_in_a_row >= 5) { sum->status = -2; break; }   // max_num_consecutive_invalid_steps
        } else {
            invalid_in_a_row = 0;
            if (std::sqrt(step2) <= opt->parameter_tolerance * (std::sqrt(x2) + opt->parameter_tolerance)) {
                sum->status = 1; sum->iterations = it;
                sum->cost_history[std::min(it, ORACLE_MAX_HISTORY - 1)] = cost;
                sum->radius_history[std::min(it, ORACLE_MAX_HISTORY - 1)] = radius;
                break;
            }
            // candidate x + delta (ambient plus, no manifold: modularsbasolver.cpp:837-838)
            std::copy(p->cam_r, p->cam_r + 3 * (size_t)Nc, r0.begin());
            std::copy(p->cam_t, p->cam_t + 3 * (size_t)Nc, t0v.begin());
            std::copy(p->pts, p->pts + np3, x0.begin());
            for (int c = 0; c < Nc; c++) for (int a = 0; a < 3; a++) {
                p->cam_r[3 * c + a] += dc[6 * (size_t)c + a];
                p->cam_t[3 * c + a] += dc[6 * (size_t)c + 3 + a];
            }
            for (size_t i = 0; i < np3; i++) p->pts[i] += dp[i];
            double new_cost = 0.0;
            t0 = now_s();
            const bool eval_ok = s.cost_only(&new_cost);
            sum->seconds_evaluate += now_s() - t0;
            const double cost_change = cost - new_cost;
            const double rho = eval_ok ? cost_change / model_change : -1.0;
            if (eval_ok && std::fabs(cost_change) <= opt->function_tolerance * cost) {
                // FunctionToleranceReached: Ceres returns before applying the candidate
                std::copy(r0.begin(), r0.end(), p->cam_r); std::copy(t0v.begin(), t0v.end(), p->cam_t); std::copy(x0.begin(), x0.end(), p->pts);
                sum->status = 1; sum->iterations = it;
                sum->cost_history[std::min(it, ORACLE_MAX_HISTORY - 1)] = cost;
                sum->radius_history[std::min(it, ORACLE_MAX_HISTORY - 1)] = radius;
                break;
            }
            if (eval_ok && rho > opt->min_relative_decrease) {
                accepted = true;
                sum->successful_steps++;
                t0 = now_s();
                if (!s.linearize(&cost)) { sum->status = -3; break; }
                sum->seconds_evaluate += now_s() - t0;
                // StepAccepted: radius /= max(1/3, 1 - (2 rho - 1)^3)
                const double q3 = 2.0 * rho - 1.0;
                radius = radius / std::max(1.0 / 3.0, 1.0 - q3 * q3 * q3);
                radius = std::min(opt->max_radius, radius);
                decrease_factor = 2.0;
            } else {
                std::copy(r0.begin(), r0.end(), p->cam_r); std::copy(t0v.begin(), t0v.end(), p->cam_t); std::copy(x0.begin(), x0.end(), p->pts);
            }
        }
        if (!step_valid) {
            // LevenbergMarquardtStrategy::StepIsInvalid [external]: the radius is halved, decrease_factor is left alone
            radius *= 0.5;
        } else if (!accepted) {
            // StepRejected
            radius = radius / decrease_factor;
            decrease_factor *= 2.0;
        }
        sum->iterations = it;
        sum->cost_history[std::min(it, ORACLE_MAX_HISTORY - 1)] = cost;
        sum->radius_history[std::min(it, ORACLE_MAX_HISTORY - 1)] = radius;
        if (accepted && grad_max() <= opt->gradient_tolerance) { sum->status = 1; break; }
        if (radius < opt->min_radius) { sum->status = 1; break; }
    }
    sum->final_cost = cost;
    sum->seconds_total = now_s() - t_start;
    return 0;
}

// Batched multi-view DLT (inhomogeneous form): per view, with the world->sensor
// pose (R^T, -R^T t) of InvertPose and normalised ray (x, y) = dirFromUV.xy / dir.z,
// the two rows  (x r3 - r1) . X = x c3 - c1,  (y r3 - r2) . X = y c3 - c2
// (r_i rows of R^T, c = R^T t); the 3x3 normal equations are solved in closed form.
// The reference has only pairwise forms on this side (sbainitializer.cpp:987-1063
// [not compiled], modularuvprojection.h:741-757); this is the batched analogue.
int oracle_triangulate(const oracle_problem* p, int32_t n_points, const int32_t* view_off, const int32_t* view_cam,
                       const double* view_uv, int num_threads, double* xyz, uint8_t* ok_out) {
    const int nt = nthreads(num_threads);
#pragma omp parallel for num_threads(nt) schedule(static)
    for (int i = 0; i < n_points; i++) {
        double A[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, b[3] = {0, 0, 0};
        for (int k = view_off[i]; k < view_off[i + 1]; k++) {
            const int c = view_cam[k];
            const double* in = p->intrinsics + ORACLE_INTR_STRIDE * p->cam_model[c];
            double d[3];
            oracle_dir_from_uv(view_uv + 2 * (size_t)k, in, d);
            const double x = d[0] / d[2], y = d[1] / d[2];
            // rows of R^T = R(-r): rotate the unit vectors with r (columns of R)
            double R[9];
            for (int j = 0; j < 3; j++) {
                double e[3] = {0, 0, 0}, col[3]; e[j] = 1.0;
                angle_axis_rotate<double>(p->cam_r + 3 * c, e, col);
                for (int r = 0; r < 3; r++) R[3 * r + j] = col[r];
            }
            const double* t = p->cam_t + 3 * c;
            double cc[3];   // c = R^T t
            for (int j = 0; j < 3; j++) cc[j] = R[j] * t[0] + R[3 + j] * t[1] + R[6 + j] * t[2];
            double row0[3], row1[3];
            for (int j = 0; j < 3; j++) { row0[j] = x * R[3 * j + 2] - R[3 * j]; row1[j] = y * R[3 * j + 2] - R[3 * j + 1]; }
            const double h0 = x * cc[2] - cc[0], h1 = y * cc[2] - cc[1];
            for (int a = 0; a < 3; a++) {
                b[a] += row0[a] * h0 + row1[a] * h1;
                for (int bb = 0; bb < 3; bb++) A[3 * a + bb] += row0[a] * row0[bb] + row1[a] * row1[bb];
            }
        }
        double Ai[9];
        const bool ok = inv3_sym(A, Ai);
        for (int a = 0; a < 3; a++) xyz[3 * (size_t)i + a] = ok ? Ai[3 * a] * b[0] + Ai[3 * a + 1] * b[1] + Ai[3 * a + 2] * b[2] : 0.0;
        if (ok_out) ok_out[i] = ok ? 1 : 0;
    }
    return 0;
}

}  // extern "C"
