/*
 * sba_oracle_ext.cpp -- CPU ORACLE (test infrastructure, never shipped), extended problem class.
 *
 * Restates the rows SURVEY 8(f) lists after the camera / point path, on the same forward-mode duals as
 * sba_oracle.cpp (the expression trees are the reference functors'):
 *   f1  free intrinsic blocks f[1], pp[2], k[3], t[2], B[2] per camera model with per-block constness and
 *       FixedSizeNormalPrior<n,n> priors (sbamodules/pinholecameraprojectormodule.cpp:105-122, 302-533);
 *   f3  constant sensor offsets and free boresight / lever-arm blocks (PoseTransformCost, LeverArmCost,
 *       PoseTransformLeverArmCost: sbamodules/pinholecameraprojectormodule.cpp:18-25, 147-214;
 *       costfunctors/posedecoratorfunctors.h:148-264, 628-700);
 *   f2  UV <-> UV cross-projection factors (UV2UVCost, costfunctors/modularuvprojection.h:660-795; built at
 *       modularsbasolver.cpp:1616-1739 with both poses inverted) under ceres::HuberLoss / CauchyLoss / ArctanLoss(1)
 *       (sbamodules/correspondencessetsbamodule.cpp:2459-2471);
 *   f4  marginal covariance of the point blocks (ModularSBASolver::std_step, modularsbasolver.cpp:952-1004; pairs
 *       requested by sbamodules/landmarkssbamodule.cpp:264-297).
 *
 * The linear algebra is DENSE (the reduced system over cameras + shared blocks is formed explicitly), so this file
 * is for scenes of up to a few hundred cameras.  Layout of the reduced system ("extended cameras", 6 columns each):
 *   real cameras [rAxis | t], then per camera model with at least one free block two blocks [f ppx ppy k1 k2 k3] and
 *   [t1 t2 - - - -], then one block [boresight | leverArm] per free mounting.  Columns of constant blocks and the
 *   padding columns are dead (unit diagonal, zero rhs); the B block never enters: d res / d B == 0 (the skew is built from
 *   the tangential block, pinholecameraprojectormodule.h:46-47), so its Gauss-Newton step is exactly zero in Ceres too.
 *   The preconditioner of the PCG is the inverse of the 6x6 diagonal blocks of S in THIS layout.
 *
 * PARITY STATUS: functors pinned by the reference's zero-cost vectors (tests/test_oracle_golden.py: uv projector,
 * simplePnP, imagePair); LM / losses / covariance restate Ceres' published algorithms -- "parity unpinned".
 */
#include "sba_oracle.h"
#include "sba_oracle_functors.h"

#include <array>
#include <chrono>
#include <cstdio>
#include <limits>
#include <vector>

using namespace oracle_detail;

namespace {

constexpr int NV = 25;   // duals of one projection observation: pose 6, point 3, intrinsics 10, boresight + lever arm 6
typedef Jet<NV> J25;
constexpr int NP = 32;   // duals of one UV <-> UV factor: pose1 6, pose2 6, intrinsics1 10, intrinsics2 10
typedef Jet<NP> J32;

// UV2UVCost::operator() (costfunctors/modularuvprojection.h:694-776) behind InvertPoseDynamic on both poses
// (modularsbasolver.cpp:1519,1648): r1, t1, r2, t2 are the PARAMETER blocks (sensor -> world).
template <class T>
inline void uv2uv_cost(const T* rp1, const T* tp1, const T* rp2, const T* tp2, const T* in1, const T* in2,
                       const double* uv1, const double* uv2, const double* info1, const double* info2, T* res) {
    T r1[3], t1[3], r2[3], t2[3];
    rigid_inverse(rp1, tp1, r1, t1);               // world2sensor1 (InvertPose)
    rigid_inverse(rp2, tp2, r2, t2);               // world2sensor2
    T s1w_r[3], s1w_t[3];
    rigid_inverse(r1, t1, s1w_r, s1w_t);           // sensor12world = world2sensor1.inverse()  (":709")
    // sensor12sensor2 = world2sensor2 * sensor12world (":716"): x -> R2 (R1w x + t1w) + t2
    auto rot12 = [&](const T* x, T* out) { T a[3]; angle_axis_rotate(s1w_r, x, a); angle_axis_rotate(r2, a, out); };
    T t12[3];
    { T a[3]; angle_axis_rotate(r2, s1w_t, a); for (int i = 0; i < 3; i++) t12[i] = a[i] + t2[i]; }
    T d1[3], d2[3];
    dir_from_uv(uv1, in1, in1 + 1, in1 + 3, in1 + 6, in1 + 8, d1);
    dir_from_uv(uv2, in2, in2 + 1, in2 + 3, in2 + 6, in2 + 8, d2);
    T a1[3];
    rot12(d1, a1);                                  // direction1CamFrame (":738")
    // [a1 -d2] [x; y] = -t12 in the least-squares sense (":741-753", householderQr().solve): Householder QR of the 3x2 matrix
    T M[3][2], b[3];
    for (int i = 0; i < 3; i++) { M[i][0] = a1[i]; M[i][1] = -d2[i]; b[i] = -t12[i]; }
    for (int k = 0; k < 2; k++) {
        T nrm2 = T(0.0);
        for (int i = k; i < 3; i++) nrm2 = nrm2 + M[i][k] * M[i][k];
        T nrm = sqrt(nrm2);
        T alpha = val(M[k][k]) > 0.0 ? -nrm : nrm;
        T v[3];
        for (int i = k; i < 3; i++) v[i] = M[i][k];
        v[k] = v[k] - alpha;
        T vv = T(0.0);
        for (int i = k; i < 3; i++) vv = vv + v[i] * v[i];
        auto reflect = [&](auto get, auto set) {
            T dotp = T(0.0);
            for (int i = k; i < 3; i++) dotp = dotp + v[i] * get(i);
            T f = 2.0 * dotp / vv;
            for (int i = k; i < 3; i++) set(i, get(i) - f * v[i]);
        };
        if (val(vv) > 0.0) {
            for (int c = k; c < 2; c++) reflect([&](int i) { return M[i][c]; }, [&](int i, const T& x) { M[i][c] = x; });
            reflect([&](int i) { return b[i]; }, [&](int i, const T& x) { b[i] = x; });
        }
    }
    T y = b[1] / M[1][1];
    T x = (b[0] - M[0][1] * y) / M[0][0];
    T inter[3];
    for (int i = 0; i < 3; i++) inter[i] = (t12[i] + x * a1[i] + y * d2[i]) / 2.0;   // intersectionCam2 (":755-758")
    // intersectionCam1 = sensor12sensor2.inverse() * intersectionCam2 (":760"): R1w^T (R2^T (p - t2) - t1w)
    T c1[3];
    {
        T q[3], nr2[3] = {-r2[0], -r2[1], -r2[2]}, nr1[3] = {-s1w_r[0], -s1w_r[1], -s1w_r[2]}, u[3];
        for (int i = 0; i < 3; i++) q[i] = inter[i] - t2[i];
        angle_axis_rotate(nr2, q, u);
        for (int i = 0; i < 3; i++) u[i] = u[i] - s1w_t[i];
        angle_axis_rotate(nr1, u, c1);
    }
    // ":762-774" -- the reference scales projCam1 by direction2Img2.z and projCam2 by direction1Img1.z (kept)
    T e1x = c1[0] / c1[2] * d2[2] - d1[0], e1y = c1[1] / c1[2] * d2[2] - d1[1];
    T e2x = inter[0] / inter[2] * d1[2] - d2[0], e2y = inter[1] / inter[2] * d1[2] - d2[1];
    const double I[4] = {1, 0, 0, 1};
    const double* w1 = info1 ? info1 : I; const double* w2 = info2 ? info2 : I;
    res[0] = w1[0] * e1x + w1[1] * e1y; res[1] = w1[2] * e1x + w1[3] * e1y;
    res[2] = w2[0] * e2x + w2[1] * e2y; res[3] = w2[2] * e2x + w2[3] * e2y;
}

// ceres::HuberLoss / CauchyLoss / ArctanLoss::Evaluate [external, ceres/loss_function.cc restated]: rho(s), rho', rho''
inline void loss_eval(int kind, double a, double s, double* rho) {
    const double tiny = std::numeric_limits<double>::min();
    if (kind == 1) {
        const double b = a * a;
        if (s > b) { const double r = std::sqrt(s); rho[0] = 2.0 * a * r - b; rho[1] = std::max(tiny, a / r); rho[2] = -rho[1] / (2.0 * s); }
        else { rho[0] = s; rho[1] = 1.0; rho[2] = 0.0; }
    } else if (kind == 2) {
        const double b = a * a, c = 1.0 / b, sum = 1.0 + s * c, inv = 1.0 / sum;
        rho[0] = b * std::log(sum); rho[1] = std::max(tiny, inv); rho[2] = -c * inv * inv;
    } else if (kind == 3) {
        const double b = 1.0 / (a * a), sum = 1.0 + s * s * b, inv = 1.0 / sum;
        rho[0] = a * std::atan2(s, a); rho[1] = std::max(tiny, inv); rho[2] = -2.0 * s * b * inv * inv;
    } else { rho[0] = s; rho[1] = 1.0; rho[2] = 0.0; }
}
// ceres::internal::Corrector [external, ceres/corrector.cc restated]: residual and Jacobian scaling of a robustified block
struct Corrector {
    double sqrt_rho1, residual_scaling, alpha_sq_norm;
    Corrector(double sq_norm, const double* rho) {
        sqrt_rho1 = std::sqrt(rho[1]);
        if (sq_norm == 0.0 || rho[2] <= 0.0) { residual_scaling = sqrt_rho1; alpha_sq_norm = 0.0; return; }
        const double D = 1.0 + 2.0 * sq_norm * rho[2] / rho[1];
        const double alpha = 1.0 - std::sqrt(D);
        residual_scaling = sqrt_rho1 / (1.0 - alpha);
        alpha_sq_norm = alpha / sq_norm;
    }
};

struct Ext {
    const oracle_ext_problem* q;
    const oracle_problem* p;
    int nt = 1;
    int Nc, Np, No, Nm, Nmt, Npair, Nce;
    std::vector<int> model_blk, mount_blk;        // first extended block of a model (two blocks) / block of a mounting, or -1
    std::vector<uint8_t> live;                    // 6 * Nce: column takes part in the solve
    std::vector<int> pt_off, pt_obs;
    // linearisation
    std::vector<double> res, Jc, Jp, Ji, Jl;      // per observation: 2, 2x6, 2x3, 2x10, 2x6
    std::vector<double> pres, pJ;                 // per pair: 4 (corrected), 4 x 32 (corrected)
    std::vector<double> prior_res, ipr;           // 3 per prior, 3 per intrinsics prior
    std::vector<double> H, g;                     // reduced-side J^T J (n x n) and J^T r (n), n = 6 Nce, undamped
    std::vector<double> Hpp, gp, Hinv;            // 9 / 3 / 9 per point
    std::vector<std::vector<std::pair<int, std::array<double, 18>>>> Wp;   // per point: (extended block, 6x3 = J_blk^T Jp summed)
    std::vector<double> scale, lam, scale_p, lam_p;

    bool cam_fixed(int c) const { return p->cam_fixed && p->cam_fixed[c]; }
    bool pt_fixed(int i) const { return p->pt_fixed && p->pt_fixed[i]; }
    bool intr_block_fixed(int m, int blk) const { return !q->intr_rw || !q->intr_fixed_mask || (q->intr_fixed_mask[m] >> blk & 1); }
    static int intr_block_of(int k) { return k == 0 ? 0 : (k < 3 ? 1 : (k < 6 ? 2 : (k < 8 ? 3 : 4))); }
    bool mount_free(int k) const { return !(q->mount_fixed && q->mount_fixed[k]); }

    void setup() {
        p = &q->base;
        Nc = p->n_cams; Np = p->n_pts; No = p->n_obs; Nm = p->n_models; Nmt = q->n_mounts; Npair = q->n_pairs;
        model_blk.assign(Nm, -1); mount_blk.assign(std::max(Nmt, 1), -1);
        int e = Nc;
        for (int m = 0; m < Nm; m++) {
            bool any = false;
            for (int b = 0; b < 4; b++) any = any || !intr_block_fixed(m, b);   // B never enters
            if (any) { model_blk[m] = e; e += 2; }
        }
        for (int k = 0; k < Nmt; k++) if (mount_free(k)) mount_blk[k] = e++;
        Nce = e;
        live.assign(6 * (size_t)Nce, 0);
        for (int c = 0; c < Nc; c++) if (!cam_fixed(c)) for (int a = 0; a < 6; a++) live[6 * c + a] = 1;
        for (int m = 0; m < Nm; m++) if (model_blk[m] >= 0)
            for (int k = 0; k < 8; k++) if (!intr_block_fixed(m, intr_block_of(k))) live[6 * model_blk[m] + k] = 1;   // k = 6, 7 fall into the second block
        for (int k = 0; k < Nmt; k++) if (mount_blk[k] >= 0) for (int a = 0; a < 6; a++) live[6 * mount_blk[k] + a] = 1;
        pt_off.assign(Np + 1, 0);
        for (int o = 0; o < No; o++) pt_off[p->obs_pt[o] + 1]++;
        for (int i = 0; i < Np; i++) pt_off[i + 1] += pt_off[i];
        pt_obs.resize(No);
        { std::vector<int> cur(pt_off.begin(), pt_off.end() - 1); for (int o = 0; o < No; o++) pt_obs[cur[p->obs_pt[o]]++] = o; }
        res.resize(2 * (size_t)No); Jc.resize(12 * (size_t)No); Jp.resize(6 * (size_t)No); Ji.resize(20 * (size_t)No); Jl.resize(12 * (size_t)No);
        pres.resize(4 * (size_t)Npair); pJ.resize(4 * (size_t)NP * Npair);
        prior_res.resize(3 * (size_t)p->n_priors); ipr.resize(3 * (size_t)q->n_intr_priors);
        const size_t n = 6 * (size_t)Nce;
        H.resize(n * n); g.resize(n); Hpp.resize(9 * (size_t)Np); gp.resize(3 * (size_t)Np); Hinv.resize(9 * (size_t)Np);
        Wp.resize(Np);
        scale.assign(n, 1.0); lam.assign(n, 0.0); scale_p.assign(3 * (size_t)Np, 1.0); lam_p.assign(3 * (size_t)Np, 0.0);
    }

    // one projection observation on duals (vars: 0..5 pose, 6..8 point, 9..18 intrinsics, 19..24 boresight | lever arm)
    template <class T>
    bool obs_functor(int o, const T* r, const T* t, const T* x, const T* in, const T* bl, T* out) const {
        const int c = p->obs_cam[o];
        const int sph = p->cam_cost_kind ? p->cam_cost_kind[c] : 0;
        const double* info = p->obs_info ? p->obs_info + 4 * (size_t)o : nullptr;
        const double* uv = p->obs_uv + 2 * (size_t)o;
        const double* off = q->cam_offset ? q->cam_offset + 6 * (size_t)c : nullptr;
        const bool has_off = off && (off[0] * off[0] + off[1] * off[1] + off[2] * off[2] >= 1e-12 || off[3] * off[3] + off[4] * off[4] + off[5] * off[5] >= 1e-12);
        const int mt = q->cam_mount ? q->cam_mount[c] : -1;
        const double zero[3] = {0, 0, 0};
        // the four branches of PinholeCamProjModule::addProjectionCostFunctionImpl (pinholecameraprojectormodule.cpp:58-214)
        if (mt >= 0) return uv_cost_leverarm<T>(r, t, bl, bl + 3, has_off ? off : zero, has_off ? off + 3 : zero, x, in, in + 1, in + 3, in + 6, in + 8, uv, info, sph, out);
        if (has_off) return uv_cost_offset<T>(r, t, off, off + 3, x, in, in + 1, in + 3, in + 6, in + 8, uv, info, sph, out);
        return uv_cost<T>(r, t, x, in, in + 1, in + 3, in + 6, in + 8, uv, info, sph, out);
    }
    bool eval_obs(int o, bool jac, double* r2, double* jc, double* jp, double* ji, double* jl) const {
        const int c = p->obs_cam[o], i = p->obs_pt[o];
        const double* in = p->intrinsics + 10 * (size_t)p->cam_model[c];
        const int mt = q->cam_mount ? q->cam_mount[c] : -1;
        if (!jac) {
            double r[3], t[3], x[3], par[10], bl[6] = {0, 0, 0, 0, 0, 0};
            for (int a = 0; a < 3; a++) { r[a] = p->cam_r[3 * c + a]; t[a] = p->cam_t[3 * c + a]; x[a] = p->pts[3 * (size_t)i + a]; }
            for (int a = 0; a < 10; a++) par[a] = in[a];
            if (mt >= 0) for (int a = 0; a < 3; a++) { bl[a] = q->mount_b[3 * mt + a]; bl[3 + a] = q->mount_l[3 * mt + a]; }
            return obs_functor<double>(o, r, t, x, par, bl, r2);
        }
        J25 r[3], t[3], x[3], par[10], bl[6], out[2];
        for (int a = 0; a < 3; a++) { r[a] = J25(p->cam_r[3 * c + a], a); t[a] = J25(p->cam_t[3 * c + a], 3 + a); x[a] = J25(p->pts[3 * (size_t)i + a], 6 + a); }
        for (int a = 0; a < 10; a++) par[a] = J25(in[a], 9 + a);
        for (int a = 0; a < 6; a++) bl[a] = J25(0.0);
        if (mt >= 0) for (int a = 0; a < 3; a++) { bl[a] = J25(q->mount_b[3 * mt + a], 19 + a); bl[3 + a] = J25(q->mount_l[3 * mt + a], 22 + a); }
        if (!obs_functor<J25>(o, r, t, x, par, bl, out)) return false;
        for (int k = 0; k < 2; k++) {
            r2[k] = out[k].a;
            for (int a = 0; a < 6; a++) jc[6 * k + a] = out[k].v[a];
            for (int a = 0; a < 3; a++) jp[3 * k + a] = out[k].v[6 + a];
            for (int a = 0; a < 10; a++) ji[10 * k + a] = out[k].v[9 + a];
            for (int a = 0; a < 6; a++) jl[6 * k + a] = out[k].v[19 + a];
        }
        return true;
    }
    // one UV <-> UV factor: raw residuals (4), raw Jacobian 4 x 32 (pose1, pose2, intrinsics of camera 1, of camera 2)
    void eval_pair(int k, bool jac, double* r4, double* J) const {
        const int c1 = q->pair_cam[2 * k], c2 = q->pair_cam[2 * k + 1];
        const double* in1 = p->intrinsics + 10 * (size_t)p->cam_model[c1];
        const double* in2 = p->intrinsics + 10 * (size_t)p->cam_model[c2];
        const double* uv = q->pair_uv + 4 * (size_t)k;
        const double* info = q->pair_info ? q->pair_info + 8 * (size_t)k : nullptr;
        if (!jac) {
            double a1[10], a2[10];
            for (int a = 0; a < 10; a++) { a1[a] = in1[a]; a2[a] = in2[a]; }
            uv2uv_cost<double>(p->cam_r + 3 * c1, p->cam_t + 3 * c1, p->cam_r + 3 * c2, p->cam_t + 3 * c2, a1, a2, uv, uv + 2, info, info ? info + 4 : nullptr, r4);
            return;
        }
        J32 r1[3], t1[3], r2[3], t2[3], a1[10], a2[10], out[4];
        for (int a = 0; a < 3; a++) {
            r1[a] = J32(p->cam_r[3 * c1 + a], a); t1[a] = J32(p->cam_t[3 * c1 + a], 3 + a);
            r2[a] = J32(p->cam_r[3 * c2 + a], 6 + a); t2[a] = J32(p->cam_t[3 * c2 + a], 9 + a);
        }
        for (int a = 0; a < 10; a++) { a1[a] = J32(in1[a], 12 + a); a2[a] = J32(in2[a], 22 + a); }
        uv2uv_cost<J32>(r1, t1, r2, t2, a1, a2, uv, uv + 2, info, info ? info + 4 : nullptr, out);
        for (int i = 0; i < 4; i++) { r4[i] = out[i].a; for (int a = 0; a < NP; a++) J[NP * i + a] = out[i].v[a]; }
    }
    // r = A (x - b), A = dim x dim row-major in the leading entries of a 9-array (fixedsizenormalprior.h:89-96)
    static void prior_eval(const double* A, const double* b, const double* x, int dim, double* r) {
        for (int k = 0; k < dim; k++) { double v = 0; for (int j = 0; j < dim; j++) v += A[dim * k + j] * (x[j] - b[j]); r[k] = v; }
        for (int k = dim; k < 3; k++) r[k] = 0.0;
    }
    static void intr_block_range(int blk, int* first, int* dim) {
        static const int f[5] = {0, 1, 3, 6, 8}, d[5] = {1, 2, 3, 2, 2};
        *first = f[blk]; *dim = d[blk];
    }
    const double* prior_x(int i) const {
        const int k = p->prior_kind[i], idx = p->prior_index[i];
        return k == 0 ? p->pts + 3 * (size_t)idx : (k == 1 ? p->cam_r + 3 * idx : p->cam_t + 3 * idx);
    }
    double priors_cost(bool store) {
        double sum = 0.0;
        for (int i = 0; i < p->n_priors; i++) {
            double r[3];
            prior_eval(p->prior_A + 9 * (size_t)i, p->prior_b + 3 * (size_t)i, prior_x(i), 3, r);
            for (int k = 0; k < 3; k++) { sum += r[k] * r[k]; if (store) prior_res[3 * (size_t)i + k] = r[k]; }
        }
        for (int i = 0; i < q->n_intr_priors; i++) {
            int first, dim;
            intr_block_range(q->ip_block[i], &first, &dim);
            double r[3];
            prior_eval(q->ip_A + 9 * (size_t)i, q->ip_b + 3 * (size_t)i, p->intrinsics + 10 * (size_t)q->ip_model[i] + first, dim, r);
            for (int k = 0; k < 3; k++) { sum += r[k] * r[k]; if (store) ipr[3 * (size_t)i + k] = r[k]; }
        }
        return sum;
    }
    bool cost_only(double* cost) {
        double sum = 0.0; int bad = 0;
#pragma omp parallel for num_threads(nt) reduction(+ : sum, bad) schedule(static)
        for (int o = 0; o < No; o++) {
            double r[2];
            if (!eval_obs(o, false, r, nullptr, nullptr, nullptr, nullptr)) { bad++; continue; }
            sum += r[0] * r[0] + r[1] * r[1];
        }
        for (int k = 0; k < Npair; k++) {
            double r[4], rho[3];
            eval_pair(k, false, r, nullptr);
            loss_eval(q->loss_kind, q->loss_a, r[0] * r[0] + r[1] * r[1] + r[2] * r[2] + r[3] * r[3], rho);
            sum += rho[0];
        }
        sum += priors_cost(false);
        *cost = 0.5 * sum;
        return bad == 0;
    }
    // extended column of dual variable v of observation o (or -1: not part of the reduced side / dead)
    int col_of(int c, int v) const {
        if (v < 6) return 6 * c + v;
        if (v >= 9 && v < 19) {
            const int k = v - 9, m = p->cam_model[c];
            if (k >= 8 || model_blk[m] < 0) return -1;
            return 6 * model_blk[m] + k;
        }
        if (v >= 19) {
            const int mt = q->cam_mount ? q->cam_mount[c] : -1;
            if (mt < 0 || mount_blk[mt] < 0) return -1;
            return 6 * mount_blk[mt] + (v - 19);
        }
        return -1;
    }
    bool linearize(double* cost) {
        const size_t n = 6 * (size_t)Nce;
        int bad = 0;
#pragma omp parallel for num_threads(nt) reduction(+ : bad) schedule(static)
        for (int o = 0; o < No; o++)
            if (!eval_obs(o, true, &res[2 * (size_t)o], &Jc[12 * (size_t)o], &Jp[6 * (size_t)o], &Ji[20 * (size_t)o], &Jl[12 * (size_t)o])) bad++;
        if (bad) return false;
        double sum = 0.0;
        for (size_t i = 0; i < res.size(); i++) sum += res[i] * res[i];
        std::fill(H.begin(), H.end(), 0.0); std::fill(g.begin(), g.end(), 0.0);
        std::fill(Hpp.begin(), Hpp.end(), 0.0); std::fill(gp.begin(), gp.end(), 0.0);
        // projection observations: reduced-side columns (pose, intrinsics, lever arm) and the point block
        for (int i = 0; i < Np; i++) {
            Wp[i].clear();
            const bool pfix = pt_fixed(i);
            for (int kk = pt_off[i]; kk < pt_off[i + 1]; kk++) {
                const int o = pt_obs[kk], c = p->obs_cam[o];
                double row[2][NV];
                for (int k = 0; k < 2; k++) {
                    for (int a = 0; a < 6; a++) row[k][a] = Jc[12 * (size_t)o + 6 * k + a];
                    for (int a = 0; a < 3; a++) row[k][6 + a] = pfix ? 0.0 : Jp[6 * (size_t)o + 3 * k + a];
                    for (int a = 0; a < 10; a++) row[k][9 + a] = Ji[20 * (size_t)o + 10 * k + a];
                    for (int a = 0; a < 6; a++) row[k][19 + a] = Jl[12 * (size_t)o + 6 * k + a];
                }
                int cols[NV];
                for (int v = 0; v < NV; v++) { cols[v] = (v >= 6 && v < 9) ? -1 : col_of(c, v); if (cols[v] >= 0 && !live[cols[v]]) cols[v] = -1; }
                const double* r = &res[2 * (size_t)o];
                for (int v = 0; v < NV; v++) {
                    if (cols[v] < 0) continue;
                    g[cols[v]] += row[0][v] * r[0] + row[1][v] * r[1];
                    for (int w = 0; w < NV; w++) if (cols[w] >= 0) H[(size_t)cols[v] * n + cols[w]] += row[0][v] * row[0][w] + row[1][v] * row[1][w];
                }
                for (int a = 0; a < 3; a++) {
                    gp[3 * (size_t)i + a] += row[0][6 + a] * r[0] + row[1][6 + a] * r[1];
                    for (int b = 0; b < 3; b++) Hpp[9 * (size_t)i + 3 * a + b] += row[0][6 + a] * row[0][6 + b] + row[1][6 + a] * row[1][6 + b];
                }
                // W blocks: (extended block, 6 x 3)
                for (int v = 0; v < NV; v++) {
                    if (cols[v] < 0) continue;
                    const int blk = cols[v] / 6, a6 = cols[v] % 6;
                    size_t j = 0;
                    for (; j < Wp[i].size(); j++) if (Wp[i][j].first == blk) break;
                    if (j == Wp[i].size()) { Wp[i].push_back({blk, {}}); Wp[i][j].second.fill(0.0); }
                    for (int b = 0; b < 3; b++) Wp[i][j].second[3 * a6 + b] += row[0][v] * row[0][6 + b] + row[1][v] * row[1][6 + b];
                }
            }
        }
        // UV <-> UV factors under the loss (corrected residuals / Jacobians: Gauss-Newton sees rho' J^T J, rho' J^T r)
        for (int k = 0; k < Npair; k++) {
            double r[4], J[4 * NP], rho[3];
            eval_pair(k, true, r, J);
            const double s = r[0] * r[0] + r[1] * r[1] + r[2] * r[2] + r[3] * r[3];
            loss_eval(q->loss_kind, q->loss_a, s, rho);
            sum += rho[0];
            const Corrector cor(s, rho);
            double Jr[NP];
            for (int a = 0; a < NP; a++) Jr[a] = r[0] * J[a] + r[1] * J[NP + a] + r[2] * J[2 * NP + a] + r[3] * J[3 * NP + a];
            for (int i = 0; i < 4; i++) for (int a = 0; a < NP; a++) pJ[((size_t)k * 4 + i) * NP + a] = cor.sqrt_rho1 * (J[NP * i + a] - cor.alpha_sq_norm * r[i] * Jr[a]);
            for (int i = 0; i < 4; i++) pres[4 * (size_t)k + i] = cor.residual_scaling * r[i];
            const int c1 = q->pair_cam[2 * k], c2 = q->pair_cam[2 * k + 1];
            int cols[NP];
            for (int a = 0; a < 6; a++) { cols[a] = 6 * c1 + a; cols[6 + a] = 6 * c2 + a; }
            for (int a = 0; a < 10; a++) { cols[12 + a] = col_of(c1, 9 + a); cols[22 + a] = col_of(c2, 9 + a); }
            for (int a = 0; a < NP; a++) if (cols[a] >= 0 && !live[cols[a]]) cols[a] = -1;
            const double* Jk = &pJ[(size_t)k * 4 * NP]; const double* rk = &pres[4 * (size_t)k];
            for (int a = 0; a < NP; a++) {
                if (cols[a] < 0) continue;
                for (int i = 0; i < 4; i++) g[cols[a]] += Jk[NP * i + a] * rk[i];
                for (int b = 0; b < NP; b++) {
                    if (cols[b] < 0) continue;
                    double v = 0; for (int i = 0; i < 4; i++) v += Jk[NP * i + a] * Jk[NP * i + b];
                    H[(size_t)cols[a] * n + cols[b]] += v;
                }
            }
        }
        sum += priors_cost(true);
        for (int i = 0; i < p->n_priors; i++) {
            const int k = p->prior_kind[i], idx = p->prior_index[i];
            const double* A = p->prior_A + 9 * (size_t)i;
            if (k == 0) {
                if (pt_fixed(idx)) continue;
                for (int a = 0; a < 3; a++) {
                    for (int r = 0; r < 3; r++) gp[3 * (size_t)idx + a] += A[3 * r + a] * prior_res[3 * (size_t)i + r];
                    for (int b = 0; b < 3; b++) for (int r = 0; r < 3; r++) Hpp[9 * (size_t)idx + 3 * a + b] += A[3 * r + a] * A[3 * r + b];
                }
            } else {
                if (cam_fixed(idx)) continue;
                const int c0 = 6 * idx + (k == 1 ? 0 : 3);
                for (int a = 0; a < 3; a++) {
                    for (int r = 0; r < 3; r++) g[c0 + a] += A[3 * r + a] * prior_res[3 * (size_t)i + r];
                    for (int b = 0; b < 3; b++) for (int r = 0; r < 3; r++) H[(size_t)(c0 + a) * n + c0 + b] += A[3 * r + a] * A[3 * r + b];
                }
            }
        }
        for (int i = 0; i < q->n_intr_priors; i++) {
            const int m = q->ip_model[i], blk = q->ip_block[i];
            int first, dim;
            intr_block_range(blk, &first, &dim);
            if (blk == 4 || model_blk[m] < 0 || intr_block_fixed(m, blk)) continue;
            const double* A = q->ip_A + 9 * (size_t)i;
            const int c0 = 6 * model_blk[m] + first;
            for (int a = 0; a < dim; a++) {
                for (int r = 0; r < dim; r++) g[c0 + a] += A[dim * r + a] * ipr[3 * (size_t)i + r];
                for (int b = 0; b < dim; b++) for (int r = 0; r < dim; r++) H[(size_t)(c0 + a) * n + c0 + b] += A[dim * r + a] * A[dim * r + b];
            }
        }
        *cost = 0.5 * sum;
        return true;
    }
    void compute_scaling(bool enabled) {
        const size_t n = 6 * (size_t)Nce;
        for (size_t j = 0; j < n; j++) scale[j] = (enabled && live[j]) ? 1.0 / (1.0 + std::sqrt(H[j * n + j])) : 1.0;
        for (size_t j = 0; j < 3 * (size_t)Np; j++) scale_p[j] = enabled ? 1.0 / (1.0 + std::sqrt(Hpp[9 * (j / 3) + 4 * (j % 3)])) : 1.0;
    }
    void compute_damping(double radius, double dmin, double dmax) {
        const size_t n = 6 * (size_t)Nce;
        for (size_t j = 0; j < n; j++) {
            const double s = scale[j];
            lam[j] = live[j] ? std::min(std::max(H[j * n + j] * s * s, dmin), dmax) / radius / (s * s) : 0.0;
        }
        for (size_t j = 0; j < 3 * (size_t)Np; j++) {
            const double s = scale_p[j];
            lam_p[j] = std::min(std::max(Hpp[9 * (j / 3) + 4 * (j % 3)] * s * s, dmin), dmax) / radius / (s * s);
        }
    }
    bool invert_points(bool damped) {
        int bad = 0;
        for (int i = 0; i < Np; i++) {
            double Hd[9];
            for (int k = 0; k < 9; k++) Hd[k] = Hpp[9 * (size_t)i + k];
            if (damped) for (int a = 0; a < 3; a++) Hd[4 * a] += lam_p[3 * (size_t)i + a];
            if (pt_fixed(i)) { for (int k = 0; k < 9; k++) Hinv[9 * (size_t)i + k] = 0.0; continue; }
            if (!inv3_sym(Hd, &Hinv[9 * (size_t)i])) bad++;
        }
        return bad == 0;
    }
    // S = H + D - sum_p W Hinv W^T (dead columns: unit diagonal), rhs = -(g - sum_p W Hinv gp)
    void reduced(bool damped, std::vector<double>& S, std::vector<double>& rhs) {
        const size_t n = 6 * (size_t)Nce;
        S = H; rhs.assign(n, 0.0);
        for (size_t j = 0; j < n; j++) { rhs[j] = -g[j]; if (damped) S[j * n + j] += lam[j]; if (!live[j]) { S[j * n + j] = 1.0; rhs[j] = 0.0; } }
        for (int i = 0; i < Np; i++) {
            const double* Hi = &Hinv[9 * (size_t)i];
            const auto& W = Wp[i];
            std::vector<std::array<double, 18>> WH(W.size());
            for (size_t a = 0; a < W.size(); a++)
                for (int r = 0; r < 6; r++) for (int b = 0; b < 3; b++)
                    WH[a][3 * r + b] = W[a].second[3 * r] * Hi[b] + W[a].second[3 * r + 1] * Hi[3 + b] + W[a].second[3 * r + 2] * Hi[6 + b];
            for (size_t a = 0; a < W.size(); a++) {
                for (int r = 0; r < 6; r++) {
                    const size_t row = 6 * (size_t)W[a].first + r;
                    rhs[row] += WH[a][3 * r] * gp[3 * (size_t)i] + WH[a][3 * r + 1] * gp[3 * (size_t)i + 1] + WH[a][3 * r + 2] * gp[3 * (size_t)i + 2];
                    for (size_t b = 0; b < W.size(); b++)
                        for (int c = 0; c < 6; c++)
                            S[row * n + 6 * (size_t)W[b].first + c] -= WH[a][3 * r] * W[b].second[3 * c] + WH[a][3 * r + 1] * W[b].second[3 * c + 1] + WH[a][3 * r + 2] * W[b].second[3 * c + 2];
                }
            }
        }
    }
};

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// block-Jacobi (6x6 diagonal blocks of S) preconditioned CG on the dense S, x0 = 0; same recurrences and exit rule as oracle_solve
int pcg_dense(const std::vector<double>& S, const std::vector<double>& rhs, int nblk, double tol, int max_it, std::vector<double>& x, bool* ok) {
    const size_t n = 6 * (size_t)nblk;
    std::vector<double> Minv(36 * (size_t)nblk), r = rhs, z(n), pv(n), qv(n);
    *ok = true;
    for (int b = 0; b < nblk; b++) {
        double D[36];
        for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) D[6 * i + j] = S[(6 * (size_t)b + i) * n + 6 * (size_t)b + j];
        if (!inv6_spd(D, &Minv[36 * (size_t)b])) { *ok = false; return 0; }
    }
    auto precond = [&](const std::vector<double>& in, std::vector<double>& out) {
        for (int b = 0; b < nblk; b++) for (int i = 0; i < 6; i++) {
            double v = 0; for (int j = 0; j < 6; j++) v += Minv[36 * (size_t)b + 6 * i + j] * in[6 * (size_t)b + j];
            out[6 * (size_t)b + i] = v;
        }
    };
    auto dot = [&](const std::vector<double>& a, const std::vector<double>& b) { double s = 0; for (size_t i = 0; i < n; i++) s += a[i] * b[i]; return s; };
    x.assign(n, 0.0);
    precond(r, z); pv = z;
    double rz = dot(r, z);
    const double rz0 = rz, thresh = tol * tol * rz0;
    int it = 0;
    while (it < max_it && rz > thresh && rz0 > 0.0) {
        for (size_t i = 0; i < n; i++) { double v = 0; const double* row = &S[i * n]; for (size_t j = 0; j < n; j++) v += row[j] * pv[j]; qv[i] = v; }
        const double alpha = rz / dot(pv, qv);
        for (size_t i = 0; i < n; i++) { x[i] += alpha * pv[i]; r[i] -= alpha * qv[i]; }
        precond(r, z);
        const double rz_new = dot(r, z), beta = rz_new / rz;
        for (size_t i = 0; i < n; i++) pv[i] = z[i] + beta * pv[i];
        rz = rz_new; it++;
    }
    return it;
}

}  // namespace

extern "C" {

int oracle_ext_layout(const oracle_ext_problem* q, int32_t* n_ext, int32_t* model_block, int32_t* mount_block, uint8_t* live) {
    Ext e; e.q = q; e.setup();
    if (n_ext) *n_ext = e.Nce;
    if (model_block) for (int m = 0; m < e.Nm; m++) model_block[m] = e.model_blk[m];
    if (mount_block) for (int k = 0; k < e.Nmt; k++) mount_block[k] = e.mount_blk[k];
    if (live) std::copy(e.live.begin(), e.live.end(), live);
    return 0;
}

int oracle_ext_evaluate(const oracle_ext_problem* q, double* cost, double* residuals, double* jac_cam, double* jac_pt, double* jac_intr,
                        double* jac_lever, double* pair_res, double* pair_jac) {
    Ext e; e.q = q; e.setup();
    int bad = 0;
    double sum = 0.0;
    for (int o = 0; o < e.No; o++) {
        double r[2] = {0, 0}, jc[12] = {0}, jp[6] = {0}, ji[20] = {0}, jl[12] = {0};
        if (!e.eval_obs(o, true, r, jc, jp, ji, jl)) { bad++; r[0] = r[1] = 0; std::fill(jc, jc + 12, 0.0); std::fill(jp, jp + 6, 0.0); std::fill(ji, ji + 20, 0.0); std::fill(jl, jl + 12, 0.0); }
        sum += r[0] * r[0] + r[1] * r[1];
        if (residuals) { residuals[2 * (size_t)o] = r[0]; residuals[2 * (size_t)o + 1] = r[1]; }
        if (jac_cam) std::memcpy(jac_cam + 12 * (size_t)o, jc, sizeof(jc));
        if (jac_pt) std::memcpy(jac_pt + 6 * (size_t)o, jp, sizeof(jp));
        if (jac_intr) std::memcpy(jac_intr + 20 * (size_t)o, ji, sizeof(ji));
        if (jac_lever) std::memcpy(jac_lever + 12 * (size_t)o, jl, sizeof(jl));
    }
    for (int k = 0; k < e.Npair; k++) {
        double r[4], J[4 * NP], rho[3];
        e.eval_pair(k, true, r, J);
        loss_eval(q->loss_kind, q->loss_a, r[0] * r[0] + r[1] * r[1] + r[2] * r[2] + r[3] * r[3], rho);
        sum += rho[0];
        if (pair_res) std::memcpy(pair_res + 4 * (size_t)k, r, sizeof(r));
        if (pair_jac) std::memcpy(pair_jac + 4 * (size_t)NP * k, J, sizeof(J));
    }
    sum += e.priors_cost(false);
    if (cost) *cost = 0.5 * sum;
    return bad ? 1 : 0;
}

// dense reduced system at the current point for a given trust-region radius (radius <= 0: undamped, no scaling)
int oracle_ext_reduced(const oracle_ext_problem* q, const oracle_options* opt, double radius, double* S_out, double* rhs_out) {
    Ext e; e.q = q; e.setup();
    double cost;
    if (!e.linearize(&cost)) return 1;
    const bool damped = radius > 0.0;
    if (damped) { e.compute_scaling(opt->jacobi_scaling != 0); e.compute_damping(radius, opt->min_lm_diagonal, opt->max_lm_diagonal); }
    if (!e.invert_points(damped)) return 2;
    std::vector<double> S, rhs;
    e.reduced(damped, S, rhs);
    if (S_out) std::copy(S.begin(), S.end(), S_out);
    if (rhs_out) std::copy(rhs.begin(), rhs.end(), rhs_out);
    return 0;
}

// marginal covariance of every free point block at the current point: (J^T J)^-1 restricted to the block
// = Hpp^-1 + Hpp^-1 W^T S^-1 W Hpp^-1 with the UNDAMPED reduced system (ceres::Covariance, modularsbasolver.cpp:966-991)
int oracle_ext_point_covariance(const oracle_ext_problem* q, double* cov9) {
    Ext e; e.q = q; e.setup();
    double cost;
    if (!e.linearize(&cost)) return 1;
    if (!e.invert_points(false)) return 2;
    std::vector<double> S, rhs;
    e.reduced(false, S, rhs);
    const size_t n = 6 * (size_t)e.Nce;
    std::vector<double> L = S;
    if (!chol(L.data(), (int)n)) return 3;
    for (int i = 0; i < e.Np; i++) {
        double* C = cov9 + 9 * (size_t)i;
        const double* Hi = &e.Hinv[9 * (size_t)i];
        for (int k = 0; k < 9; k++) C[k] = Hi[k];
        if (e.pt_fixed(i)) continue;
        // V = W Hinv scattered into n x 3, X = S^-1 V, C += V^T X
        std::vector<double> V(3 * n, 0.0);
        for (const auto& w : e.Wp[i])
            for (int r = 0; r < 6; r++) for (int b = 0; b < 3; b++)
                V[b * n + 6 * (size_t)w.first + r] = w.second[3 * r] * Hi[b] + w.second[3 * r + 1] * Hi[3 + b] + w.second[3 * r + 2] * Hi[6 + b];
        std::vector<double> X = V;
        for (int b = 0; b < 3; b++) chol_solve(L.data(), (int)n, &X[b * n]);
        for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) {
            double v = 0; for (size_t j = 0; j < n; j++) v += V[a * n + j] * X[b * n + j];
            C[3 * a + b] += v;
        }
    }
    return 0;
}

// Trust-region LM as oracle_solve runs it (same radius rules, same tests), over the extended parameter set
int oracle_ext_solve(oracle_ext_problem* q, const oracle_options* opt, oracle_summary* sum) {
    const double t_start = now_s();
    std::memset(sum, 0, sizeof(*sum));
    Ext e; e.q = q; e.nt = nthreads(opt->num_threads); e.setup();
    oracle_problem* p = &q->base;
    const int Nc = e.Nc, Np = e.Np, Nm = e.Nm, Nmt = e.Nmt;
    const size_t n = 6 * (size_t)e.Nce, np3 = 3 * (size_t)Np;
    double cost;
    if (!e.linearize(&cost)) { sum->status = -1; return -1; }
    e.compute_scaling(opt->jacobi_scaling != 0);
    sum->initial_cost = cost; sum->cost_history[0] = cost;
    double radius = opt->initial_radius, decrease_factor = 2.0;
    sum->radius_history[0] = radius;
    auto grad_max = [&]() {
        double m = 0;
        for (size_t j = 0; j < n; j++) if (e.live[j]) m = std::max(m, std::fabs(e.g[j]));
        for (int i = 0; i < Np; i++) if (!e.pt_fixed(i)) for (int a = 0; a < 3; a++) m = std::max(m, std::fabs(e.gp[3 * (size_t)i + a]));
        return m;
    };
    if (grad_max() <= opt->gradient_tolerance) { sum->status = 1; sum->final_cost = cost; return 0; }
    std::vector<double> S, rhs, dx, dp(np3);
    std::vector<double> sv_r(3 * (size_t)Nc), sv_t(3 * (size_t)Nc), sv_x(np3), sv_i(10 * (size_t)Nm), sv_b(3 * (size_t)std::max(Nmt, 1)), sv_l(3 * (size_t)std::max(Nmt, 1));
    int it = 0, invalid_in_a_row = 0;
    sum->status = 0;
    auto hist = [&](int i, double c, double r) { const int k = std::min(i, ORACLE_MAX_HISTORY - 1); sum->cost_history[k] = c; sum->radius_history[k] = r; };
    while (it < opt->max_iterations) {
        it++;
        e.compute_damping(radius, opt->min_lm_diagonal, opt->max_lm_diagonal);
        bool lin_ok = e.invert_points(true);
        int pcg_it = 0;
        if (lin_ok) {
            e.reduced(true, S, rhs);
            if (opt->linear_solver == 1) {
                std::vector<double> L = S;
                lin_ok = chol(L.data(), (int)n);
                if (lin_ok) { dx = rhs; chol_solve(L.data(), (int)n, dx.data()); }
            } else {
                pcg_it = pcg_dense(S, rhs, e.Nce, opt->pcg_rel_tol, opt->pcg_max_iters, dx, &lin_ok);
            }
        }
        sum->pcg_iterations_total += pcg_it;
        sum->pcg_iters_history[std::min(it, ORACLE_MAX_HISTORY - 1)] = pcg_it;
        double model_change = 0.0, step2 = 0.0, x2 = 0.0;
        if (lin_ok) {
            for (size_t j = 0; j < n; j++) if (!e.live[j]) dx[j] = 0.0;
            // back substitution dp = -Hinv (gp + W^T dx)
            for (int i = 0; i < Np; i++) {
                double u[3] = {e.gp[3 * (size_t)i], e.gp[3 * (size_t)i + 1], e.gp[3 * (size_t)i + 2]};
                for (const auto& w : e.Wp[i]) for (int r = 0; r < 6; r++) for (int b = 0; b < 3; b++) u[b] += w.second[3 * r + b] * dx[6 * (size_t)w.first + r];
                const double* Hi = &e.Hinv[9 * (size_t)i];
                for (int a = 0; a < 3; a++) dp[3 * (size_t)i + a] = -(Hi[3 * a] * u[0] + Hi[3 * a + 1] * u[1] + Hi[3 * a + 2] * u[2]);
            }
            // model cost change -sum m.(r + m/2), m = J delta, over every residual block (corrected ones for the factors)
            double mc = 0.0;
            for (int o = 0; o < e.No; o++) {
                const int c = p->obs_cam[o], i = p->obs_pt[o];
                double m[2] = {0, 0};
                for (int k = 0; k < 2; k++) {
                    for (int v = 0; v < NV; v++) {
                        if (v >= 6 && v < 9) continue;
                        const int col = e.col_of(c, v);
                        if (col < 0 || !e.live[col]) continue;
                        const double J = v < 6 ? e.Jc[12 * (size_t)o + 6 * k + v] : (v < 19 ? e.Ji[20 * (size_t)o + 10 * k + v - 9] : e.Jl[12 * (size_t)o + 6 * k + v - 19]);
                        m[k] += J * dx[col];
                    }
                    if (!e.pt_fixed(i)) for (int a = 0; a < 3; a++) m[k] += e.Jp[6 * (size_t)o + 3 * k + a] * dp[3 * (size_t)i + a];
                }
                mc += m[0] * (e.res[2 * (size_t)o] + 0.5 * m[0]) + m[1] * (e.res[2 * (size_t)o + 1] + 0.5 * m[1]);
            }
            for (int k = 0; k < e.Npair; k++) {
                const int c1 = q->pair_cam[2 * k], c2 = q->pair_cam[2 * k + 1];
                int cols[NP];
                for (int a = 0; a < 6; a++) { cols[a] = 6 * c1 + a; cols[6 + a] = 6 * c2 + a; }
                for (int a = 0; a < 10; a++) { cols[12 + a] = e.col_of(c1, 9 + a); cols[22 + a] = e.col_of(c2, 9 + a); }
                for (int i = 0; i < 4; i++) {
                    double m = 0;
                    for (int a = 0; a < NP; a++) if (cols[a] >= 0 && e.live[cols[a]]) m += e.pJ[((size_t)k * 4 + i) * NP + a] * dx[cols[a]];
                    mc += m * (e.pres[4 * (size_t)k + i] + 0.5 * m);
                }
            }
            for (int i = 0; i < p->n_priors; i++) {
                const int k = p->prior_kind[i], idx = p->prior_index[i];
                if (k == 0 ? e.pt_fixed(idx) : e.cam_fixed(idx)) continue;
                const double* d = k == 0 ? &dp[3 * (size_t)idx] : &dx[6 * (size_t)idx + (k == 1 ? 0 : 3)];
                const double* A = p->prior_A + 9 * (size_t)i;
                for (int a = 0; a < 3; a++) {
                    const double m = A[3 * a] * d[0] + A[3 * a + 1] * d[1] + A[3 * a + 2] * d[2];
                    mc += m * (e.prior_res[3 * (size_t)i + a] + 0.5 * m);
                }
            }
            for (int i = 0; i < q->n_intr_priors; i++) {
                const int m_ = q->ip_model[i], blk = q->ip_block[i];
                int first, dim;
                Ext::intr_block_range(blk, &first, &dim);
                if (blk == 4 || e.model_blk[m_] < 0 || e.intr_block_fixed(m_, blk)) continue;
                const double* d = &dx[6 * (size_t)e.model_blk[m_] + first];
                const double* A = q->ip_A + 9 * (size_t)i;
                for (int a = 0; a < dim; a++) {
                    double m = 0; for (int j = 0; j < dim; j++) m += A[dim * a + j] * d[j];
                    mc += m * (e.ipr[3 * (size_t)i + a] + 0.5 * m);
                }
            }
            model_change = -mc;
            // step / parameter norms over the live columns (the B block never enters: its step is zero and its value is left out of |x|)
            for (size_t j = 0; j < n; j++) if (e.live[j]) step2 += dx[j] * dx[j];
            for (int c = 0; c < Nc; c++) if (!e.cam_fixed(c)) for (int a = 0; a < 3; a++) x2 += p->cam_r[3 * c + a] * p->cam_r[3 * c + a] + p->cam_t[3 * c + a] * p->cam_t[3 * c + a];
            for (int m = 0; m < Nm; m++) if (e.model_blk[m] >= 0) for (int k = 0; k < 8; k++) if (e.live[6 * (size_t)e.model_blk[m] + k]) x2 += p->intrinsics[10 * m + k] * p->intrinsics[10 * m + k];
            for (int k = 0; k < Nmt; k++) if (e.mount_blk[k] >= 0) for (int a = 0; a < 3; a++) x2 += q->mount_b[3 * k + a] * q->mount_b[3 * k + a] + q->mount_l[3 * k + a] * q->mount_l[3 * k + a];
            for (int i = 0; i < Np; i++) if (!e.pt_fixed(i)) for (int a = 0; a < 3; a++) { step2 += dp[3 * (size_t)i + a] * dp[3 * (size_t)i + a]; x2 += p->pts[3 * (size_t)i + a] * p->pts[3 * (size_t)i + a]; }
        }
        const bool step_valid = lin_ok && model_change > 0.0 && std::isfinite(model_change);
        bool accepted = false;
        auto restore = [&]() {
            std::copy(sv_r.begin(), sv_r.end(), p->cam_r); std::copy(sv_t.begin(), sv_t.end(), p->cam_t); std::copy(sv_x.begin(), sv_x.end(), p->pts);
            if (q->intr_rw) std::copy(sv_i.begin(), sv_i.end(), q->intr_rw);
            if (Nmt) { std::copy(sv_b.begin(), sv_b.begin() + 3 * (size_t)Nmt, q->mount_b); std::copy(sv_l.begin(), sv_l.begin() + 3 * (size_t)Nmt, q->mount_l); }
        };
        if (!step_valid) {
            if (++invalid_in_a_row >= 5) { sum->status = -2; sum->iterations = it; break; }
        } else {
            invalid_in_a_row = 0;
            if (std::sqrt(step2) <= opt->parameter_tolerance * (std::sqrt(x2) + opt->parameter_tolerance)) { sum->status = 1; sum->iterations = it; hist(it, cost, radius); break; }
            std::copy(p->cam_r, p->cam_r + 3 * (size_t)Nc, sv_r.begin()); std::copy(p->cam_t, p->cam_t + 3 * (size_t)Nc, sv_t.begin());
            std::copy(p->pts, p->pts + np3, sv_x.begin());
            if (q->intr_rw) std::copy(q->intr_rw, q->intr_rw + 10 * (size_t)Nm, sv_i.begin());
            if (Nmt) { std::copy(q->mount_b, q->mount_b + 3 * (size_t)Nmt, sv_b.begin()); std::copy(q->mount_l, q->mount_l + 3 * (size_t)Nmt, sv_l.begin()); }
            for (int c = 0; c < Nc; c++) for (int a = 0; a < 3; a++) { p->cam_r[3 * c + a] += dx[6 * (size_t)c + a]; p->cam_t[3 * c + a] += dx[6 * (size_t)c + 3 + a]; }
            for (size_t i = 0; i < np3; i++) p->pts[i] += dp[i];
            for (int m = 0; m < Nm; m++) if (e.model_blk[m] >= 0) for (int k = 0; k < 8; k++) q->intr_rw[10 * m + k] += dx[6 * (size_t)e.model_blk[m] + k];
            for (int k = 0; k < Nmt; k++) if (e.mount_blk[k] >= 0) for (int a = 0; a < 3; a++) { q->mount_b[3 * k + a] += dx[6 * (size_t)e.mount_blk[k] + a]; q->mount_l[3 * k + a] += dx[6 * (size_t)e.mount_blk[k] + 3 + a]; }
            double new_cost = 0.0;
            const bool eval_ok = e.cost_only(&new_cost);
            const double cost_change = cost - new_cost;
            const double rho = eval_ok ? cost_change / model_change : -1.0;
            if (eval_ok && std::fabs(cost_change) <= opt->function_tolerance * cost) { restore(); sum->status = 1; sum->iterations = it; hist(it, cost, radius); break; }
            if (eval_ok && rho > opt->min_relative_decrease) {
                accepted = true;
                sum->successful_steps++;
                if (!e.linearize(&cost)) { sum->status = -3; break; }
                const double q3 = 2.0 * rho - 1.0;
                radius = std::min(opt->max_radius, radius / std::max(1.0 / 3.0, 1.0 - q3 * q3 * q3));
                decrease_factor = 2.0;
            } else {
                restore();
            }
        }
        if (!step_valid) radius *= 0.5;
        else if (!accepted) { radius = radius / decrease_factor; decrease_factor *= 2.0; }
        sum->iterations = it;
        hist(it, cost, radius);
        if (accepted && grad_max() <= opt->gradient_tolerance) { sum->status = 1; break; }
        if (radius < opt->min_radius) { sum->status = 1; break; }
    }
    sum->final_cost = cost;
    sum->seconds_total = now_s() - t_start;
    return 0;
}

}  // extern "C"
