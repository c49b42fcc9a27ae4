// sba_jet.cuh -- forward-mode dual numbers for the few cost functors whose closed-form derivative is not worth its length
// (the UV <-> UV cross-projection cost: a 3x2 Householder least-squares solve between two rotated rays).  Value + N partials in
// registers / local memory; __host__ __device__ so tests/hostmath can run the same code against the oracle without a GPU.
#pragma once
#include <math.h>

#include "sba_math.cuh"

namespace sba {

template <int N>
struct DJet {
    double a;
    double v[N];
    SBA_HD DJet() : a(0.0) { for (int i = 0; i < N; i++) v[i] = 0.0; }
    SBA_HD DJet(double x) : a(x) { for (int i = 0; i < N; i++) v[i] = 0.0; }
    SBA_HD DJet(double x, int k) : a(x) { for (int i = 0; i < N; i++) v[i] = 0.0; v[k] = 1.0; }
};
template <int N> SBA_HD DJet<N> operator+(const DJet<N>& x, const DJet<N>& y) { DJet<N> r; r.a = x.a + y.a; for (int i = 0; i < N; i++) r.v[i] = x.v[i] + y.v[i]; return r; }
template <int N> SBA_HD DJet<N> operator-(const DJet<N>& x, const DJet<N>& y) { DJet<N> r; r.a = x.a - y.a; for (int i = 0; i < N; i++) r.v[i] = x.v[i] - y.v[i]; return r; }
template <int N> SBA_HD DJet<N> operator-(const DJet<N>& x) { DJet<N> r; r.a = -x.a; for (int i = 0; i < N; i++) r.v[i] = -x.v[i]; return r; }
template <int N> SBA_HD DJet<N> operator*(const DJet<N>& x, const DJet<N>& y) { DJet<N> r; r.a = x.a * y.a; for (int i = 0; i < N; i++) r.v[i] = x.a * y.v[i] + x.v[i] * y.a; return r; }
template <int N> SBA_HD DJet<N> operator/(const DJet<N>& x, const DJet<N>& y) { DJet<N> r; const double inv = 1.0 / y.a; r.a = x.a * inv; for (int i = 0; i < N; i++) r.v[i] = (x.v[i] - r.a * y.v[i]) * inv; return r; }
template <int N> SBA_HD DJet<N> operator+(const DJet<N>& x, double y) { DJet<N> r = x; r.a += y; return r; }
template <int N> SBA_HD DJet<N> operator-(const DJet<N>& x, double y) { DJet<N> r = x; r.a -= y; return r; }
template <int N> SBA_HD DJet<N> operator-(double y, const DJet<N>& x) { DJet<N> r = -x; r.a += y; return r; }
template <int N> SBA_HD DJet<N> operator*(const DJet<N>& x, double y) { DJet<N> r; r.a = x.a * y; for (int i = 0; i < N; i++) r.v[i] = x.v[i] * y; return r; }
template <int N> SBA_HD DJet<N> operator*(double y, const DJet<N>& x) { return x * y; }
template <int N> SBA_HD DJet<N> operator/(const DJet<N>& x, double y) { return x * (1.0 / y); }
template <int N> SBA_HD DJet<N> jsqrt(const DJet<N>& x) { DJet<N> r; r.a = sqrt(x.a); const double d = 0.5 / r.a; for (int i = 0; i < N; i++) r.v[i] = x.v[i] * d; return r; }
template <int N> SBA_HD DJet<N> jsin(const DJet<N>& x) { DJet<N> r; r.a = sin(x.a); const double c = cos(x.a); for (int i = 0; i < N; i++) r.v[i] = c * x.v[i]; return r; }
template <int N> SBA_HD DJet<N> jcos(const DJet<N>& x) { DJet<N> r; r.a = cos(x.a); const double s = -sin(x.a); for (int i = 0; i < N; i++) r.v[i] = s * x.v[i]; return r; }
SBA_HD double jsqrt(double x) { return sqrt(x); }
SBA_HD double jsin(double x) { return sin(x); }
SBA_HD double jcos(double x) { return cos(x); }
SBA_HD double jval(double x) { return x; }
template <int N> SBA_HD double jval(const DJet<N>& x) { return x.a; }

// R(r) x, Rodrigues with the small-angle series (the rotation LibStevi's RigidBodyTransform applies; same statement as pose_cache)
template <class T>
SBA_HD void jet_rotate(const T* r, const T* x, T* out) {
    const T th2 = r[0] * r[0] + r[1] * r[1] + r[2] * r[2];
    T a, b;
    if (jval(th2) > 1e-6) {
        const T th = jsqrt(th2);
        a = jsin(th) / th;
        b = (1.0 - jcos(th)) / th2;
    } else {
        a = 1.0 - th2 * (1.0 / 6.0) * (1.0 - th2 * (1.0 / 20.0) * (1.0 - th2 * (1.0 / 42.0)));
        b = 0.5 * (1.0 - th2 * (1.0 / 12.0) * (1.0 - th2 * (1.0 / 30.0) * (1.0 - th2 * (1.0 / 56.0))));
    }
    const T c0 = r[1] * x[2] - r[2] * x[1], c1 = r[2] * x[0] - r[0] * x[2], c2 = r[0] * x[1] - r[1] * x[0];
    const T e0 = r[1] * c2 - r[2] * c1, e1 = r[2] * c0 - r[0] * c2, e2 = r[0] * c1 - r[1] * c0;
    out[0] = x[0] + a * c0 + b * e0; out[1] = x[1] + a * c1 + b * e1; out[2] = x[2] + a * c2 + b * e2;
}

// UV2UVCost::operator() (costfunctors/modularuvprojection.h:694-776) behind InvertPoseDynamic on both poses
// (modularsbasolver.cpp:1519,1648).  rp, tp = the pose PARAMETER blocks (sensor -> world); d1, d2 = the view rays
// dirFromUV(uv; intrinsics) of the two pixels (variables of their own here: the chain to the intrinsic blocks is closed form);
// w1, w2 = 2x2 stiffness row-major or nullptr.  res[4].  With sensor2world poses (R1, t1), (R2, t2):
//   sensor1 -> sensor2:  x -> R2^T (R1 x + t1 - t2);  the two rays t12 + x a1 and y d2 are intersected in the least-squares
//   sense (3x2 Householder QR, ":741-753"), the midpoint is projected into both sensors and compared with the rays;
//   the reference scales the projection into camera 1 by d2.z and the one into camera 2 by d1.z (":765-771", kept).
template <class T>
SBA_HD void uv2uv_residual(const T* rp1, const T* tp1, const T* rp2, const T* tp2, const T* d1, const T* d2,
                           const double* w1, const double* w2, T* res) {
    const T nr2[3] = {-rp2[0], -rp2[1], -rp2[2]}, nr1[3] = {-rp1[0], -rp1[1], -rp1[2]};
    T a1[3], t12[3];
    {
        T u[3];
        jet_rotate(rp1, d1, u);
        jet_rotate(nr2, u, a1);                       // direction1CamFrame
        const T dt[3] = {tp1[0] - tp2[0], tp1[1] - tp2[1], tp1[2] - tp2[2]};
        jet_rotate(nr2, dt, t12);                     // sensor12sensor2.t
    }
    T M[3][2], b[3];
    for (int i = 0; i < 3; i++) { M[i][0] = a1[i]; M[i][1] = -d2[i]; b[i] = -t12[i]; }
    for (int k = 0; k < 2; k++) {
        T nrm2 = M[k][k] * M[k][k];
        for (int i = k + 1; i < 3; i++) nrm2 = nrm2 + M[i][k] * M[i][k];
        const T nrm = jsqrt(nrm2);
        const T alpha = jval(M[k][k]) > 0.0 ? -nrm : nrm;
        T v[3];
        for (int i = k; i < 3; i++) v[i] = M[i][k];
        v[k] = v[k] - alpha;
        T vv = v[k] * v[k];
        for (int i = k + 1; i < 3; i++) vv = vv + v[i] * v[i];
        if (jval(vv) > 0.0) {
            for (int c = k; c < 3; c++) {             // columns k..1 of M, then b (c == 2)
                T dotp = v[k] * (c < 2 ? M[k][c] : b[k]);
                for (int i = k + 1; i < 3; i++) dotp = dotp + v[i] * (c < 2 ? M[i][c] : b[i]);
                const T f = 2.0 * dotp / vv;
                for (int i = k; i < 3; i++) { if (c < 2) M[i][c] = M[i][c] - f * v[i]; else b[i] = b[i] - f * v[i]; }
            }
        }
    }
    const T y = b[1] / M[1][1];
    const T x = (b[0] - M[0][1] * y) / M[0][0];
    T inter[3], c1[3];
    for (int i = 0; i < 3; i++) inter[i] = (t12[i] + x * a1[i] + y * d2[i]) / 2.0;      // intersectionCam2
    {   // intersectionCam1 = sensor12sensor2.inverse() * intersectionCam2 = R1^T (R2 p + t2 - t1)
        T u[3];
        jet_rotate(rp2, inter, u);
        for (int i = 0; i < 3; i++) u[i] = u[i] + tp2[i] - tp1[i];
        jet_rotate(nr1, u, c1);
    }
    const T e1x = c1[0] / c1[2] * d2[2] - d1[0], e1y = c1[1] / c1[2] * d2[2] - d1[1];
    const T e2x = inter[0] / inter[2] * d1[2] - d2[0], e2y = inter[1] / inter[2] * d1[2] - d2[1];
    if (w1) { res[0] = w1[0] * e1x + w1[1] * e1y; res[1] = w1[2] * e1x + w1[3] * e1y; } else { res[0] = e1x; res[1] = e1y; }
    if (w2) { res[2] = w2[0] * e2x + w2[1] * e2y; res[3] = w2[2] * e2x + w2[3] * e2y; } else { res[2] = e2x; res[3] = e2y; }
}

// rho(s), rho', rho'' of ceres::HuberLoss / CauchyLoss / ArctanLoss (a = scale; the reference passes 1,
// sbamodules/correspondencessetsbamodule.cpp:2459-2471); kind 0 = no loss
SBA_HD void loss_rho(int kind, double a, double s, double* rho) {
    const double tiny = 2.2250738585072014e-308;
    if (kind == 1) {
        const double b = a * a;
        if (s > b) { const double r = sqrt(s); rho[0] = 2.0 * a * r - b; rho[1] = fmax(tiny, a / r); rho[2] = -rho[1] / (2.0 * s); }
        else { rho[0] = s; rho[1] = 1.0; rho[2] = 0.0; }
    } else if (kind == 2) {
        const double b = a * a, c = 1.0 / b, sum = 1.0 + s * c, inv = 1.0 / sum;
        rho[0] = b * log(sum); rho[1] = fmax(tiny, inv); rho[2] = -c * inv * inv;
    } else if (kind == 3) {
        const double b = 1.0 / (a * a), sum = 1.0 + s * s * b, inv = 1.0 / sum;
        rho[0] = a * atan2(s, a); rho[1] = fmax(tiny, inv); rho[2] = -2.0 * s * b * inv * inv;
    } else { rho[0] = s; rho[1] = 1.0; rho[2] = 0.0; }
}
// ceres::internal::Corrector: r' = scale_r r, J' = sqrt(rho') (J - alpha_sq r (r^T J))
SBA_HD void loss_corrector(double sq_norm, const double* rho, double* sqrt_rho1, double* residual_scaling, double* alpha_sq_norm) {
    *sqrt_rho1 = sqrt(rho[1]);
    if (sq_norm == 0.0 || rho[2] <= 0.0) { *residual_scaling = *sqrt_rho1; *alpha_sq_norm = 0.0; return; }
    const double D = 1.0 + 2.0 * sq_norm * rho[2] / rho[1];
    const double alpha = 1.0 - sqrt(D);
    *residual_scaling = *sqrt_rho1 / (1.0 - alpha);
    *alpha_sq_norm = alpha / sq_norm;
}

}  // namespace sba
