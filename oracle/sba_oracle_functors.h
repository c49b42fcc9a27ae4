/*
 * sba_oracle_functors.h -- CPU ORACLE (test infrastructure, never shipped): the forward-mode dual type and the
 * restated reference functors, shared by sba_oracle.cpp (camera / point LM at any scale) and sba_oracle_ext.cpp
 * (free intrinsics, sensor offsets / lever arms, UV <-> UV factors, robust losses, covariance: dense reduced system).
 * Every function cites the reference lines it follows; see sba_oracle.cpp for the parity status.
 */
#ifndef SBA_ORACLE_FUNCTORS_H
#define SBA_ORACLE_FUNCTORS_H

#include <algorithm>
#include <cmath>
#include <cstring>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace oracle_detail {

// ----------------------------------------------------------------------------
// Jet<N>: value + N partial derivatives (stand-in for ceres::Jet<double,N>).
// ----------------------------------------------------------------------------
template <int N>
struct Jet {
    double a;
    double v[N];
    Jet() : a(0.0) { for (int i = 0; i < N; i++) v[i] = 0.0; }
    Jet(double x) : a(x) { for (int i = 0; i < N; i++) v[i] = 0.0; }
    Jet(double x, int k) : a(x) { for (int i = 0; i < N; i++) v[i] = 0.0; v[k] = 1.0; }
};
template <int N> inline Jet<N> operator+(const Jet<N>& x, const Jet<N>& y) { Jet<N> r; r.a = x.a + y.a; for (int i = 0; i < N; i++) r.v[i] = x.v[i] + y.v[i]; return r; }
template <int N> inline Jet<N> operator-(const Jet<N>& x, const Jet<N>& y) { Jet<N> r; r.a = x.a - y.a; for (int i = 0; i < N; i++) r.v[i] = x.v[i] - y.v[i]; return r; }
template <int N> inline Jet<N> operator-(const Jet<N>& x) { Jet<N> r; r.a = -x.a; for (int i = 0; i < N; i++) r.v[i] = -x.v[i]; return r; }
template <int N> inline Jet<N> operator*(const Jet<N>& x, const Jet<N>& y) { Jet<N> r; r.a = x.a * y.a; for (int i = 0; i < N; i++) r.v[i] = x.a * y.v[i] + x.v[i] * y.a; return r; }
template <int N> inline Jet<N> operator/(const Jet<N>& x, const Jet<N>& y) { Jet<N> r; const double inv = 1.0 / y.a; r.a = x.a * inv; for (int i = 0; i < N; i++) r.v[i] = (x.v[i] - r.a * y.v[i]) * inv; return r; }
template <int N> inline Jet<N> operator+(const Jet<N>& x, double y) { Jet<N> r = x; r.a += y; return r; }
template <int N> inline Jet<N> operator+(double y, const Jet<N>& x) { Jet<N> r = x; r.a += y; return r; }
template <int N> inline Jet<N> operator-(const Jet<N>& x, double y) { Jet<N> r = x; r.a -= y; return r; }
template <int N> inline Jet<N> operator-(double y, const Jet<N>& x) { Jet<N> r = -x; r.a += y; return r; }
template <int N> inline Jet<N> operator*(const Jet<N>& x, double y) { Jet<N> r; r.a = x.a * y; for (int i = 0; i < N; i++) r.v[i] = x.v[i] * y; return r; }
template <int N> inline Jet<N> operator*(double y, const Jet<N>& x) { return x * y; }
template <int N> inline Jet<N> operator/(const Jet<N>& x, double y) { return x * (1.0 / y); }
template <int N> inline Jet<N> operator/(double y, const Jet<N>& x) { return Jet<N>(y) / x; }
template <int N> inline Jet<N> sqrt(const Jet<N>& x) { Jet<N> r; r.a = std::sqrt(x.a); const double d = 0.5 / r.a; for (int i = 0; i < N; i++) r.v[i] = x.v[i] * d; return r; }
template <int N> inline Jet<N> sin(const Jet<N>& x) { Jet<N> r; r.a = std::sin(x.a); const double c = std::cos(x.a); for (int i = 0; i < N; i++) r.v[i] = c * x.v[i]; return r; }
template <int N> inline Jet<N> cos(const Jet<N>& x) { Jet<N> r; r.a = std::cos(x.a); const double s = -std::sin(x.a); for (int i = 0; i < N; i++) r.v[i] = s * x.v[i]; return r; }
template <int N> inline Jet<N> atan2(const Jet<N>& y, const Jet<N>& x) {
    Jet<N> r; r.a = std::atan2(y.a, x.a);
    const double inv = 1.0 / (x.a * x.a + y.a * y.a);
    for (int i = 0; i < N; i++) r.v[i] = (x.a * y.v[i] - y.a * x.v[i]) * inv;
    return r;
}
inline double val(double x) { return x; }
template <int N> inline double val(const Jet<N>& x) { return x.a; }
using std::atan2; using std::cos; using std::sin; using std::sqrt;

// ----------------------------------------------------------------------------
// LibStevi restatement [external]: angle-axis rotation R(r) x.
// R(r)x = x + a (r x x) + b r x (r x x), a = sin(th)/th, b = (1-cos th)/th^2,
// with the Taylor series of a, b below th^2 = 1e-6 (Rodrigues with small-angle
// fallback; same family as the in-tree BAL functor
// modules/g2o/g2o/examples/bal/bal_example.cpp:174-190).
// ----------------------------------------------------------------------------
template <class T>
inline void cross3(const T* a, const T* b, T* c) {
    c[0] = a[1] * b[2] - a[2] * b[1];
    c[1] = a[2] * b[0] - a[0] * b[2];
    c[2] = a[0] * b[1] - a[1] * b[0];
}

template <class T>
inline void angle_axis_rotate(const T* r, const T* x, T* out) {
    T th2 = r[0] * r[0] + r[1] * r[1] + r[2] * r[2];
    T a, b;
    if (val(th2) > 1e-6) {
        T th = sqrt(th2);
        a = sin(th) / th;
        b = (1.0 - cos(th)) / th2;
    } else {
        a = 1.0 - th2 * (1.0 / 6.0) * (1.0 - th2 * (1.0 / 20.0) * (1.0 - th2 * (1.0 / 42.0)));
        b = 0.5 * (1.0 - th2 * (1.0 / 12.0) * (1.0 - th2 * (1.0 / 30.0) * (1.0 - th2 * (1.0 / 56.0))));
    }
    T rx[3], rrx[3];
    cross3(r, x, rx);
    cross3(r, rx, rrx);
    for (int i = 0; i < 3; i++) out[i] = x[i] + a * rx[i] + b * rrx[i];
}

// RigidBodyTransform::inverse [external]: (r,t) -> (-r, -R(-r) t);
// call site costfunctors/posedecoratorfunctors.h:1125-1127.
template <class T>
inline void rigid_inverse(const T* r, const T* t, T* ri, T* ti) {
    for (int i = 0; i < 3; i++) ri[i] = -r[i];
    T rt[3];
    angle_axis_rotate(ri, t, rt);
    for (int i = 0; i < 3; i++) ti[i] = -rt[i];
}

// PinholeUVProjector::dirFromUV, sbamodules/pinholecameraprojectormodule.h:28-71.
// params: f[1], pp[2], ks[3], ts[2], Bs[2].  QUIRK kept: B_dist is built from
// ts (":46-47"), Bs is read but never used.
// inverseSkewDistortion / radialDistortion / tangentialDistortion [external,
// LibStevi]; forms from edges/edgeparametrizedxyz2uv.cpp:195-229:
//   u = f x + B1 x + B2 y + ppx ; v = f y + ppy
//   radial = p (k1 r^2 + k2 r^4 + k3 r^6)
//   tangential = (2xy t1 + (r^2+2x^2) t2, (r^2+2y^2) t1 + 2xy t2)
template <class T>
inline void dir_from_uv(const double* uv, const T* f, const T* pp, const T* ks,
                        const T* ts, const T* /*Bs*/, T* dir) {
    const T& cam_f = f[0];
    const T B1 = ts[0], B2 = ts[1];
    T y = (uv[1] - pp[1]) / cam_f;
    T x = (uv[0] - pp[0] - B2 * y) / (cam_f + B1);
    T r2 = x * x + y * y;
    T rad = r2 * (ks[0] + r2 * (ks[1] + r2 * ks[2]));
    T xy = x * y;
    T dtx = 2.0 * xy * ts[0] + (r2 + 2.0 * x * x) * ts[1];
    T dty = (r2 + 2.0 * y * y) * ts[0] + 2.0 * xy * ts[1];
    T cx = x + x * rad + dtx;
    T cy = y + y * rad + dty;
    dir[0] = cx * cam_f;
    dir[1] = cy * cam_f;
    dir[2] = cam_f;  // T(1) * f
}

// UV2ParametrizedXYZCost::operator(), costfunctors/modularuvprojection.h:368-448,
// wrapped by InvertPose<.,0>, costfunctors/posedecoratorfunctors.h:1106-1155.
// r, t are the PARAMETER blocks (sensor->world).  Returns false when the plane
// cost sees Pbar.z < 0 (":395-399").
// the part of UV2ParametrizedXYZCost::operator() behind world2sensor * lm_pos (costfunctors/modularuvprojection.h:395-448)
template <class T>
inline bool uv_cost_sensor(const T* Pbar, const T* f, const T* pp, const T* ks, const T* ts, const T* Bs,
                           const double* uv, const double* info, int spherical, T* residual) {
    if (!spherical) {
        if (val(Pbar[2]) < 0.0) return false;
    }
    T dir[3];
    dir_from_uv(uv, f, pp, ks, ts, Bs, dir);
    T e0, e1;
    if (!spherical) {
        T p0 = Pbar[0] / Pbar[2], p1 = Pbar[1] / Pbar[2];   // projectPoints (":401")
        e0 = p0 * dir[2] - dir[0];                           // ":414-416"
        e1 = p1 * dir[2] - dir[1];
    } else {
        // ":417-432"
        T alpha1 = atan2(dir[0], dir[2]);
        T c1 = cos(alpha1), s1 = sin(alpha1);
        T d1x = c1 * Pbar[2] + s1 * Pbar[0];
        T d1y = c1 * Pbar[0] - s1 * Pbar[2];
        T alpha2 = atan2(d1y, d1x);
        e0 = alpha2 * sqrt(dir[0] * dir[0] + dir[2] * dir[2]) / c1;
        T beta1 = atan2(dir[1], dir[2]);
        T c2 = cos(beta1), s2 = sin(beta1);
        T d2x = c2 * Pbar[2] + s2 * Pbar[1];
        T d2y = c2 * Pbar[1] - s2 * Pbar[2];
        T beta2 = atan2(d2y, d2x);
        e1 = beta2 * sqrt(dir[1] * dir[1] + dir[2] * dir[2]) / c2;
    }
    if (info) {                                              // ":437-438"
        residual[0] = info[0] * e0 + info[1] * e1;
        residual[1] = info[2] * e0 + info[3] * e1;
    } else {
        residual[0] = e0;
        residual[1] = e1;
    }
    return true;
}

template <class T>
inline bool uv_cost(const T* r, const T* t, const T* lm, const T* f, const T* pp,
                    const T* ks, const T* ts, const T* Bs, const double* uv,
                    const double* info, int spherical, T* residual) {
    T ri[3], ti[3];
    rigid_inverse(r, t, ri, ti);                 // InvertPose
    T Pbar[3];
    angle_axis_rotate(ri, lm, Pbar);             // world2sensor * lm_pos (":394")
    for (int i = 0; i < 3; i++) Pbar[i] = Pbar[i] + ti[i];
    return uv_cost_sensor(Pbar, f, pp, ks, ts, Bs, uv, info, spherical, residual);
}

// PoseTransform<UVCost, SourceToInitial> (costfunctors/posedecoratorfunctors.h:628-700, built at
// sbamodules/pinholecameraprojectormodule.cpp:23,197-214): the UV cost evaluated at transformed = pose * offset, a constant
// sensor -> body transform (r_off, t_off).  RigidBodyTransform::operator* [external, LibStevi] composes
// x -> R(r)(R(r_off) x + t_off) + t; the cost only ever uses the composed pose through world2sensor * lm, so the
// restatement applies the two inverses in turn instead of forming the composed axis-angle (exp(log(R)) = R):
// Pbar = R(-r_off) ((R(-r) lm - R(-r) t) - t_off).
template <class T>
inline bool uv_cost_offset(const T* r, const T* t, const double* r_off, const double* t_off, const T* lm, const T* f, const T* pp,
                           const T* ks, const T* ts, const T* Bs, const double* uv, const double* info, int spherical, T* residual) {
    T ri[3], ti[3];
    rigid_inverse(r, t, ri, ti);
    T Pb[3];
    angle_axis_rotate(ri, lm, Pb);
    for (int i = 0; i < 3; i++) Pb[i] = Pb[i] + ti[i] - t_off[i];
    T ro[3] = {T(-r_off[0]), T(-r_off[1]), T(-r_off[2])};
    T Pbar[3];
    angle_axis_rotate(ro, Pb, Pbar);
    return uv_cost_sensor(Pbar, f, pp, ks, ts, Bs, uv, info, spherical, residual);
}

// PoseTransform<LeverArm<UVCost, Body2World | Sensor2Body, 0>, SourceToInitial, 0> (costfunctors/posedecoratorfunctors.h:148-264 for
// ApplyLeverArm -- sensor2world = body2world * sensor2body at ":208", alias at ":500" --, :628-700 for PoseTransform; built at
// sbamodules/pinholecameraprojectormodule.cpp:21,25,147-166 with the blocks pose r, pose t, boresight, lever arm, point, intrinsics):
// the UV cost at (pose * offset) * (boresight, leverArm).  As in uv_cost_offset the inverses are applied in turn:
// Pbar = R(-b) ( R(-r_off) (R(-r)(lm - t) - t_off) - l ).
template <class T>
inline bool uv_cost_leverarm(const T* r, const T* t, const T* b, const T* l, const double* r_off, const double* t_off, const T* lm,
                             const T* f, const T* pp, const T* ks, const T* ts, const T* Bs, const double* uv, const double* info,
                             int spherical, T* residual) {
    T ri[3], ti[3];
    rigid_inverse(r, t, ri, ti);
    T Pb[3];
    angle_axis_rotate(ri, lm, Pb);
    for (int i = 0; i < 3; i++) Pb[i] = Pb[i] + ti[i] - t_off[i];
    T ro[3] = {T(-r_off[0]), T(-r_off[1]), T(-r_off[2])};
    T Pm[3];
    angle_axis_rotate(ro, Pb, Pm);
    for (int i = 0; i < 3; i++) Pm[i] = Pm[i] - l[i];
    T bi[3] = {-b[0], -b[1], -b[2]};
    T Pbar[3];
    angle_axis_rotate(bi, Pm, Pbar);
    return uv_cost_sensor(Pbar, f, pp, ks, ts, Bs, uv, info, spherical, residual);
}

inline int nthreads(int req) {
#ifdef _OPENMP
    int mx = omp_get_max_threads();
    return (req <= 0 || req > mx) ? mx : req;
#else
    (void)req;
    return 1;
#endif
}

// ---------------------------------------------------------------------------
// small dense helpers
// ---------------------------------------------------------------------------
// inverse of a symmetric 3x3 (full row-major storage) by adjugate
inline bool inv3_sym(const double* A, double* Ai) {
    const double a = A[0], b = A[1], c = A[2], d = A[4], e = A[5], f = A[8];
    const double C00 = d * f - e * e, C01 = c * e - b * f, C02 = b * e - c * d;
    const double det = a * C00 + b * C01 + c * C02;
    if (!(std::fabs(det) > 0.0) || !std::isfinite(det)) return false;
    const double id = 1.0 / det;
    Ai[0] = C00 * id; Ai[1] = C01 * id; Ai[2] = C02 * id;
    Ai[3] = Ai[1]; Ai[4] = (a * f - c * c) * id; Ai[5] = (b * c - a * e) * id;
    Ai[6] = Ai[2]; Ai[7] = Ai[5]; Ai[8] = (a * d - b * b) * id;
    return true;
}

// Cholesky of an n x n SPD matrix (row-major, in place lower); returns false if not PD
inline bool chol(double* A, int n) {
    for (int j = 0; j < n; j++) {
        double d = A[j * n + j];
        for (int k = 0; k < j; k++) d -= A[j * n + k] * A[j * n + k];
        if (!(d > 0.0)) return false;
        d = std::sqrt(d);
        A[j * n + j] = d;
        for (int i = j + 1; i < n; i++) {
            double s = A[i * n + j];
            for (int k = 0; k < j; k++) s -= A[i * n + k] * A[j * n + k];
            A[i * n + j] = s / d;
        }
    }
    return true;
}
inline void chol_solve(const double* L, int n, double* b) {
    for (int i = 0; i < n; i++) { double s = b[i]; for (int k = 0; k < i; k++) s -= L[i * n + k] * b[k]; b[i] = s / L[i * n + i]; }
    for (int i = n - 1; i >= 0; i--) { double s = b[i]; for (int k = i + 1; k < n; k++) s -= L[k * n + i] * b[k]; b[i] = s / L[i * n + i]; }
}
inline bool inv6_spd(const double* A, double* Ai) {
    double L[36];
    std::memcpy(L, A, sizeof(L));
    if (!chol(L, 6)) return false;
    for (int j = 0; j < 6; j++) {
        double e[6] = {0, 0, 0, 0, 0, 0};
        e[j] = 1.0;
        chol_solve(L, 6, e);
        for (int i = 0; i < 6; i++) Ai[i * 6 + j] = e[i];
    }
    return true;
}


}  // namespace oracle_detail
#endif
