// sba_math.cuh -- per-observation / per-block FP64 math of the SBA hot path.
//
// Closed-form (not autodiff) derivatives of the reference functor
//   InvertPose<UV2ParametrizedXYZCost<PinholeUVProjector,Flags,1,2,3,2,2>,0>
// (sbamodules/pinholecameraprojectormodule.cpp:18-19).  Everything here is
// __host__ __device__ so that tests/hostmath can compile the SAME expressions
// with g++ and check them against the oracle without a GPU; the product only
// ever runs them inside the kernels of sba_kernels.cu.
#pragma once

#include <math.h>

#if defined(__CUDACC__)
#define SBA_HD __host__ __device__ __forceinline__
#else
#define SBA_HD inline
#endif

namespace sba {

// camera cache record (doubles): 0..8 Rt = R(r)^T row-major (world->sensor
// rotation), 9..11 t (sensor->world translation), 12..20 Jr = right Jacobian of
// SO(3) at r, 21..23 padding (record = 192 B, sector aligned)
constexpr int CAM_CACHE_STRIDE = 24;

// a = sin(th)/th, b = (1-cos th)/th^2, c = (th - sin th)/th^3 with series near 0.
SBA_HD void so3_coeffs(double th2, double* a, double* b, double* c) {
    if (th2 > 1e-4) {
        const double th = sqrt(th2);
        double s, co;
#if defined(__CUDA_ARCH__)
        sincos(th, &s, &co);
#else
        s = sin(th); co = cos(th);
#endif
        const double sh = sin(0.5 * th);
        *a = s / th;
        *b = 2.0 * sh * sh / th2;            // (1 - cos th)/th^2 without cancellation
        if (th2 > 0.25) {
            *c = (th - s) / (th2 * th);
        } else {
            // (th - sin th)/th^3 series: cancellation-free below th = 0.5
            *c = (1.0 / 6.0) * (1.0 - th2 * (1.0 / 20.0) * (1.0 - th2 * (1.0 / 42.0) * (1.0 - th2 * (1.0 / 72.0) *
                 (1.0 - th2 * (1.0 / 110.0) * (1.0 - th2 * (1.0 / 156.0) * (1.0 - th2 * (1.0 / 210.0)))))));
        }
        (void)co;
    } else {
        *a = 1.0 - th2 * (1.0 / 6.0) * (1.0 - th2 * (1.0 / 20.0) * (1.0 - th2 * (1.0 / 42.0)));
        *b = 0.5 * (1.0 - th2 * (1.0 / 12.0) * (1.0 - th2 * (1.0 / 30.0) * (1.0 - th2 * (1.0 / 56.0))));
        *c = (1.0 / 6.0) * (1.0 - th2 * (1.0 / 20.0) * (1.0 - th2 * (1.0 / 42.0) * (1.0 - th2 * (1.0 / 72.0))));
    }
}

// K0: per-camera cache.  Parameters (r,t) are sensor->world; InvertPose
// (costfunctors/posedecoratorfunctors.h:1106-1155) hands the functor
// (r',t') = (-r, -R(-r)t), so Pbar = R(r)^T (X - t)  (modularuvprojection.h:389-394).
//   R(r)   = I + a [r]x + b [r]x^2
//   dPbar/dr = [Pbar]x Jr(r),  Jr(r) = I - b [r]x + c [r]x^2   (right Jacobian of SO(3);
//   equals the Gallego-Yezzi closed form (r r^T - (R - I)[r]x)/|r|^2 premultiplied by R^T)
SBA_HD void pose_cache(const double* r, const double* t, double* out) {
    const double x = r[0], y = r[1], z = r[2];
    const double th2 = x * x + y * y + z * z;
    double a, b, c;
    so3_coeffs(th2, &a, &b, &c);
    // [r]x^2 = r r^T - th2 I
    const double xx = x * x - th2, yy = y * y - th2, zz = z * z - th2;
    const double xy = x * y, xz = x * z, yz = y * z;
    // R = I + a K + b K2, K = [r]x = [[0,-z,y],[z,0,-x],[-y,x,0]];  store R^T
    const double R00 = 1.0 + b * xx, R01 = -a * z + b * xy, R02 = a * y + b * xz;
    const double R10 = a * z + b * xy, R11 = 1.0 + b * yy, R12 = -a * x + b * yz;
    const double R20 = -a * y + b * xz, R21 = a * x + b * yz, R22 = 1.0 + b * zz;
    out[0] = R00; out[1] = R10; out[2] = R20;
    out[3] = R01; out[4] = R11; out[5] = R21;
    out[6] = R02; out[7] = R12; out[8] = R22;
    out[9] = t[0]; out[10] = t[1]; out[11] = t[2];
    // Jr = I - b K + c K2
    out[12] = 1.0 + c * xx; out[13] = b * z + c * xy;  out[14] = -b * y + c * xz;
    out[15] = -b * z + c * xy; out[16] = 1.0 + c * yy; out[17] = b * x + c * yz;
    out[18] = b * y + c * xz;  out[19] = -b * x + c * yz; out[20] = 1.0 + c * zz;
    out[21] = 0.0; out[22] = 0.0; out[23] = 0.0;
}

// PinholeUVProjector::dirFromUV (sbamodules/pinholecameraprojectormodule.h:28-71)
// intr = f, pp[2], k[3], t[2], B[2]; skew built from t (":46-47"), B unused.
SBA_HD void dir_from_uv(double u, double v, const double* in, double* d) {
    const double f = in[0];
    const double y = (v - in[2]) / f;
    const double x = (u - in[1] - in[7] * y) / (f + in[6]);
    const double r2 = x * x + y * y;
    const double rad = r2 * (in[3] + r2 * (in[4] + r2 * in[5]));
    const double xy = x * y;
    const double dtx = 2.0 * xy * in[6] + (r2 + 2.0 * x * x) * in[7];
    const double dty = (r2 + 2.0 * y * y) * in[6] + 2.0 * xy * in[7];
    d[0] = (x + x * rad + dtx) * f;
    d[1] = (y + y * rad + dty) * f;
    d[2] = f;
}

// The same ray with the two reciprocals of the lens model, 1 / f and 1 / (f + t1), supplied by the caller (table intr_r, once per
// camera model and linearisation: k_intr_recip): no division per observation in K1 / K1c / K1p.  K1 and K1p both use THIS form, so the
// camera-order and point-order copies of A and res stay bitwise identical; against dir_from_uv it differs by rounding only.
SBA_HD void dir_from_uv_r(double u, double v, const double* in, double inv_f, double inv_fx, double* d) {
    const double f = in[0];
    const double y = (v - in[2]) * inv_f;
    const double x = (u - in[1] - in[7] * y) * inv_fx;
    const double r2 = x * x + y * y;
    const double rad = r2 * (in[3] + r2 * (in[4] + r2 * in[5]));
    const double xy = x * y;
    const double dtx = 2.0 * xy * in[6] + (r2 + 2.0 * x * x) * in[7];
    const double dty = (r2 + 2.0 * y * y) * in[6] + 2.0 * xy * in[7];
    d[0] = (x + x * rad + dtx) * f;
    d[1] = (y + y * rad + dty) * f;
    d[2] = f;
}

// dirFromUV and its Jacobian with respect to (f, ppx, ppy, k1, k2, k3, t1, t2): dd[8 r + c] = d d_r / d intr_c, r = x, y, z
// (d_z = f).  Same expressions as in obs_intr_jac_pbar below.
SBA_HD void dir_from_uv_jac(double u, double v, const double* in, double* d, double* dd) {
    const double f = in[0], t1 = in[6], t2 = in[7];
    const double y = (v - in[2]) / f;
    const double D = f + t1;
    const double x = (u - in[1] - t2 * y) / D;
    const double r2 = x * x + y * y;
    const double rad = r2 * (in[3] + r2 * (in[4] + r2 * in[5]));
    const double Rp = in[3] + r2 * (2.0 * in[4] + 3.0 * r2 * in[5]);
    const double xy = x * y;
    const double g0 = x + x * rad + 2.0 * xy * t1 + (r2 + 2.0 * x * x) * t2;
    const double g1 = y + y * rad + (r2 + 2.0 * y * y) * t1 + 2.0 * xy * t2;
    const double g0x = 1.0 + rad + 2.0 * x * x * Rp + 2.0 * y * t1 + 6.0 * x * t2, g0y = 2.0 * xy * Rp + 2.0 * x * t1 + 2.0 * y * t2;
    const double g1x = g0y, g1y = 1.0 + rad + 2.0 * y * y * Rp + 6.0 * y * t1 + 2.0 * x * t2;
    const double y_f = -y / f, y_py = -1.0 / f;
    const double x_f = (t2 * y / f - x) / D, x_px = -1.0 / D, x_py = t2 / (f * D), x_t1 = -x / D, x_t2 = -y / D;
    d[0] = g0 * f; d[1] = g1 * f; d[2] = f;
    dd[0] = g0 + f * (g0x * x_f + g0y * y_f); dd[8] = g1 + f * (g1x * x_f + g1y * y_f);
    dd[1] = f * g0x * x_px;                   dd[9] = f * g1x * x_px;
    dd[2] = f * (g0x * x_py + g0y * y_py);    dd[10] = f * (g1x * x_py + g1y * y_py);
    dd[3] = f * x * r2;                       dd[11] = f * y * r2;
    dd[4] = f * x * r2 * r2;                  dd[12] = f * y * r2 * r2;
    dd[5] = f * x * r2 * r2 * r2;             dd[13] = f * y * r2 * r2 * r2;
    dd[6] = f * (2.0 * xy + g0x * x_t1);            dd[14] = f * (r2 + 2.0 * y * y + g1x * x_t1);
    dd[7] = f * (r2 + 2.0 * x * x + g0x * x_t2);    dd[15] = f * (2.0 * xy + g1x * x_t2);
    dd[16] = 1.0;
    for (int c = 1; c < 8; c++) dd[16 + c] = 0.0;
}

// One observation.  cc = camera cache record, X = point, d = view ray,
// W = 2x2 stiffness (row-major) or nullptr (identity).
// Outputs: res[2]; if JAC: Jc[12] = rows [d/dr(3) d/dt(3)], Jp[6] = rows d/dX(3).
// aux (optional, 5 doubles): Pbar[3] and the two per-axis scale constants of Pi (plane: d_z, d_z;
// spherical: (d_x^2+d_z^2)/d_z, (d_y^2+d_z^2)/d_z) -- everything the factored operator pass needs
// to rebuild Pi, hence A = Pi R^T and C = Pi [Pbar]x Jr, without reading them.
// Returns false when the plane cost sees Pbar.z < 0 (modularuvprojection.h:395-399).
template <bool JAC>
SBA_HD bool obs_eval(const double* cc, const double* X, const double* d, const double* W,
                     int spherical, double* res, double* Jc, double* Jp, double* aux = nullptr) {
    const double vx = X[0] - cc[9], vy = X[1] - cc[10], vz = X[2] - cc[11];
    const double Px = cc[0] * vx + cc[1] * vy + cc[2] * vz;
    const double Py = cc[3] * vx + cc[4] * vy + cc[5] * vz;
    const double Pz = cc[6] * vx + cc[7] * vy + cc[8] * vz;
    double e0, e1;
    // Pi = d e / d Pbar, rows (p00, 0, p02), (0, p11, p12)
    double p00, p02, p11, p12;
    if (!spherical) {
        if (Pz < 0.0) return false;
        const double iz = 1.0 / Pz;
        const double px = Px * iz, py = Py * iz;
        e0 = px * d[2] - d[0];                       // ":414-416"
        e1 = py * d[2] - d[1];
        if (JAC) {
            const double s = d[2] * iz;
            p00 = s; p11 = s; p02 = -s * px; p12 = -s * py;
            if (aux) { aux[3] = d[2]; aux[4] = d[2]; }
        }
    } else {
        // ":417-432": alpha1 = atan2(dx,dz), rotate (Pz,Px) by alpha1, angle of the result,
        // scaled by |(dx,dz)| / cos(alpha1) = (dx^2+dz^2)/dz.
        const double nx2 = d[0] * d[0] + d[2] * d[2], ny2 = d[1] * d[1] + d[2] * d[2];
        const double inx = 1.0 / sqrt(nx2), iny = 1.0 / sqrt(ny2);
        const double cx = d[2] * inx, sx = d[0] * inx;
        const double cy = d[2] * iny, sy = d[1] * iny;
        const double kx = nx2 / d[2], ky = ny2 / d[2];
        const double a0 = cx * Pz + sx * Px, a1 = cx * Px - sx * Pz;
        const double b0 = cy * Pz + sy * Py, b1 = cy * Py - sy * Pz;
        e0 = atan2(a1, a0) * kx;
        e1 = atan2(b1, b0) * ky;
        if (JAC) {
            const double qx = kx / (Px * Px + Pz * Pz), qy = ky / (Py * Py + Pz * Pz);
            p00 = qx * Pz; p02 = -qx * Px;
            p11 = qy * Pz; p12 = -qy * Py;
            if (aux) { aux[3] = kx; aux[4] = ky; }
        }
    }
    if (JAC && aux) { aux[0] = Px; aux[1] = Py; aux[2] = Pz; }
    if (W) {
        res[0] = W[0] * e0 + W[1] * e1;
        res[1] = W[2] * e0 + W[3] * e1;
    } else {
        res[0] = e0; res[1] = e1;
    }
    if (JAC) {
        // A = Pi * Rt (2x3) = d e / d X ; d e / d t = -A
        double A[6];
        A[0] = p00 * cc[0] + p02 * cc[6]; A[1] = p00 * cc[1] + p02 * cc[7]; A[2] = p00 * cc[2] + p02 * cc[8];
        A[3] = p11 * cc[3] + p12 * cc[6]; A[4] = p11 * cc[4] + p12 * cc[7]; A[5] = p11 * cc[5] + p12 * cc[8];
        // B = Pi * [Pbar]x (2x3)
        const double B0 = -p02 * Py, B1 = -p00 * Pz + p02 * Px, B2 = p00 * Py;
        const double B3 = p11 * Pz - p12 * Py, B4 = p12 * Px, B5 = -p11 * Px;
        // C = B * Jr (2x3) = d e / d r
        double C[6];
        C[0] = B0 * cc[12] + B1 * cc[15] + B2 * cc[18];
        C[1] = B0 * cc[13] + B1 * cc[16] + B2 * cc[19];
        C[2] = B0 * cc[14] + B1 * cc[17] + B2 * cc[20];
        C[3] = B3 * cc[12] + B4 * cc[15] + B5 * cc[18];
        C[4] = B3 * cc[13] + B4 * cc[16] + B5 * cc[19];
        C[5] = B3 * cc[14] + B4 * cc[17] + B5 * cc[20];
        if (W) {
            for (int j = 0; j < 3; j++) {
                const double c0 = C[j], c1 = C[3 + j], a0 = A[j], a1 = A[3 + j];
                Jc[j] = W[0] * c0 + W[1] * c1;       Jc[6 + j] = W[2] * c0 + W[3] * c1;
                const double x0 = W[0] * a0 + W[1] * a1, x1 = W[2] * a0 + W[3] * a1;
                Jp[j] = x0; Jp[3 + j] = x1;
                Jc[3 + j] = -x0; Jc[9 + j] = -x1;
            }
        } else {
            for (int j = 0; j < 3; j++) {
                Jc[j] = C[j]; Jc[6 + j] = C[3 + j];
                Jp[j] = A[j]; Jp[3 + j] = A[3 + j];
                Jc[3 + j] = -A[j]; Jc[9 + j] = -A[3 + j];
            }
        }
    }
    return true;
}

// One observation through a CONSTANT sensor offset (PoseTransform<UVCost, SourceToInitial>,
// costfunctors/posedecoratorfunctors.h:628-700, built at sbamodules/pinholecameraprojectormodule.cpp:197-214): the UV cost
// at pose * offset, i.e. Pbar = Roff^T (R^T (X - t) - toff).  cc = cache record of the BODY pose, off = Roff^T row-major [9]
// and toff [3].  Same outputs as obs_eval.  Closed form:  A = Pi Roff^T R^T = d e / d X,  d e / d t = -A,
// d e / d r = (Pi Roff^T) [R^T (X - t)]x Jr  -- everything downstream of the linearisation stays written in A and the body pose.
// Math only so far: checked against the oracle's duals on the host (tests/test_hostmath.py); no kernel calls it yet.
template <bool JAC>
SBA_HD bool obs_eval_offset(const double* cc, const double* off, const double* X, const double* d, const double* W,
                            int spherical, double* res, double* Jc, double* Jp) {
    const double vx = X[0] - cc[9], vy = X[1] - cc[10], vz = X[2] - cc[11];
    const double Bx = cc[0] * vx + cc[1] * vy + cc[2] * vz;      // point in the body frame
    const double By = cc[3] * vx + cc[4] * vy + cc[5] * vz;
    const double Bz = cc[6] * vx + cc[7] * vy + cc[8] * vz;
    const double ux = Bx - off[9], uy = By - off[10], uz = Bz - off[11];
    const double Px = off[0] * ux + off[1] * uy + off[2] * uz;   // point in the sensor frame
    const double Py = off[3] * ux + off[4] * uy + off[5] * uz;
    const double Pz = off[6] * ux + off[7] * uy + off[8] * uz;
    double e0, e1, p00 = 0.0, p02 = 0.0, p11 = 0.0, p12 = 0.0;
    if (!spherical) {
        if (Pz < 0.0) return false;
        const double iz = 1.0 / Pz, px = Px * iz, py = Py * iz;
        e0 = px * d[2] - d[0];
        e1 = py * d[2] - d[1];
        if (JAC) { const double s = d[2] * iz; p00 = s; p11 = s; p02 = -s * px; p12 = -s * py; }
    } else {
        const double nx2 = d[0] * d[0] + d[2] * d[2], ny2 = d[1] * d[1] + d[2] * d[2];
        const double inx = 1.0 / sqrt(nx2), iny = 1.0 / sqrt(ny2);
        const double cx = d[2] * inx, sx = d[0] * inx, cy = d[2] * iny, sy = d[1] * iny;
        const double kx = nx2 / d[2], ky = ny2 / d[2];
        e0 = atan2(cx * Px - sx * Pz, cx * Pz + sx * Px) * kx;
        e1 = atan2(cy * Py - sy * Pz, cy * Pz + sy * Py) * ky;
        if (JAC) {
            const double qx = kx / (Px * Px + Pz * Pz), qy = ky / (Py * Py + Pz * Pz);
            p00 = qx * Pz; p02 = -qx * Px; p11 = qy * Pz; p12 = -qy * Py;
        }
    }
    if (W) { res[0] = W[0] * e0 + W[1] * e1; res[1] = W[2] * e0 + W[3] * e1; }
    else { res[0] = e0; res[1] = e1; }
    if (JAC) {
        // Q = Pi Roff^T (2x3), rows (p00, 0, p02) and (0, p11, p12) times Roff^T
        double Q[6], A[6], C[6];
        for (int j = 0; j < 3; j++) { Q[j] = p00 * off[j] + p02 * off[6 + j]; Q[3 + j] = p11 * off[3 + j] + p12 * off[6 + j]; }
        for (int k = 0; k < 2; k++) {
            const double q0 = Q[3 * k], q1 = Q[3 * k + 1], q2 = Q[3 * k + 2];
            for (int j = 0; j < 3; j++) A[3 * k + j] = q0 * cc[j] + q1 * cc[3 + j] + q2 * cc[6 + j];     // Q R^T
            const double b0 = q1 * Bz - q2 * By, b1 = -q0 * Bz + q2 * Bx, b2 = q0 * By - q1 * Bx;        // q [B]x
            for (int j = 0; j < 3; j++) C[3 * k + j] = b0 * cc[12 + j] + b1 * cc[15 + j] + b2 * cc[18 + j];   // ... Jr
        }
        for (int j = 0; j < 3; j++) {
            const double c0 = C[j], c1 = C[3 + j], a0 = A[j], a1 = A[3 + j];
            const double y0 = W ? W[0] * c0 + W[1] * c1 : c0, y1 = W ? W[2] * c0 + W[3] * c1 : c1;
            const double x0 = W ? W[0] * a0 + W[1] * a1 : a0, x1 = W ? W[2] * a0 + W[3] * a1 : a1;
            Jc[j] = y0; Jc[6 + j] = y1;
            Jp[j] = x0; Jp[3 + j] = x1;
            Jc[3 + j] = -x0; Jc[9 + j] = -x1;
        }
    }
    return true;
}

// One observation through a constant offset AND a free boresight / lever arm (SURVEY 8(f) row 3;
// PoseTransform<LeverArm<UVCost, Body2World | Sensor2Body, 0>, SourceToInitial, 0>, costfunctors/posedecoratorfunctors.h:148-264,500,
// built at sbamodules/pinholecameraprojectormodule.cpp:147-166): sensor2world = (pose * offset) * (boresight b, lever arm l), i.e.
// Pbar = Rb^T (Roff^T (R^T (X - t) - toff) - l).  cc = cache record of the body pose, off = Roff^T [9] + toff [3],
// cl = pose_cache(b, l) (Rb^T, l, Jr(b)).  Outputs as obs_eval plus Jl[12] = rows [d/db(3) d/dl(3)].  Closed form, with
// G = Pi Rb^T:  d e / d l = -G,  d e / d b = Pi [Pbar]x Jr(b),  A = G Roff^T R^T = d e / d X = -d e / d t,
// d e / d r = (G Roff^T) [R^T (X - t)]x Jr(r).  Math only so far (host parity test); no kernel calls it yet.
template <bool JAC>
SBA_HD bool obs_eval_leverarm(const double* cc, const double* off, const double* cl, const double* X, const double* d, const double* W,
                              int spherical, double* res, double* Jc, double* Jl, double* Jp) {
    const double vx = X[0] - cc[9], vy = X[1] - cc[10], vz = X[2] - cc[11];
    const double Bx = cc[0] * vx + cc[1] * vy + cc[2] * vz;      // body frame
    const double By = cc[3] * vx + cc[4] * vy + cc[5] * vz;
    const double Bz = cc[6] * vx + cc[7] * vy + cc[8] * vz;
    const double ux = Bx - off[9], uy = By - off[10], uz = Bz - off[11];
    const double mx = off[0] * ux + off[1] * uy + off[2] * uz - cl[9];    // mounting frame, lever arm removed
    const double my = off[3] * ux + off[4] * uy + off[5] * uz - cl[10];
    const double mz = off[6] * ux + off[7] * uy + off[8] * uz - cl[11];
    const double Px = cl[0] * mx + cl[1] * my + cl[2] * mz;      // sensor frame
    const double Py = cl[3] * mx + cl[4] * my + cl[5] * mz;
    const double Pz = cl[6] * mx + cl[7] * my + cl[8] * mz;
    double e0, e1, p00 = 0.0, p02 = 0.0, p11 = 0.0, p12 = 0.0;
    if (!spherical) {
        if (Pz < 0.0) return false;
        const double iz = 1.0 / Pz, px = Px * iz, py = Py * iz;
        e0 = px * d[2] - d[0];
        e1 = py * d[2] - d[1];
        if (JAC) { const double s = d[2] * iz; p00 = s; p11 = s; p02 = -s * px; p12 = -s * py; }
    } else {
        const double nx2 = d[0] * d[0] + d[2] * d[2], ny2 = d[1] * d[1] + d[2] * d[2];
        const double inx = 1.0 / sqrt(nx2), iny = 1.0 / sqrt(ny2);
        const double cx = d[2] * inx, sx = d[0] * inx, cy = d[2] * iny, sy = d[1] * iny;
        const double kx = nx2 / d[2], ky = ny2 / d[2];
        e0 = atan2(cx * Px - sx * Pz, cx * Pz + sx * Px) * kx;
        e1 = atan2(cy * Py - sy * Pz, cy * Pz + sy * Py) * ky;
        if (JAC) {
            const double qx = kx / (Px * Px + Pz * Pz), qy = ky / (Py * Py + Pz * Pz);
            p00 = qx * Pz; p02 = -qx * Px; p11 = qy * Pz; p12 = -qy * Py;
        }
    }
    if (W) { res[0] = W[0] * e0 + W[1] * e1; res[1] = W[2] * e0 + W[3] * e1; }
    else { res[0] = e0; res[1] = e1; }
    if (JAC) {
        double G[6], Q[6], A[6], C[6], Cb[6];
        const double pi0[3] = {p00, 0.0, p02}, pi1[3] = {0.0, p11, p12};
        for (int k = 0; k < 2; k++) {
            const double* pi = k ? pi1 : pi0;
            for (int j = 0; j < 3; j++) G[3 * k + j] = pi[0] * cl[j] + pi[1] * cl[3 + j] + pi[2] * cl[6 + j];                 // Pi Rb^T
            for (int j = 0; j < 3; j++) Q[3 * k + j] = G[3 * k] * off[j] + G[3 * k + 1] * off[3 + j] + G[3 * k + 2] * off[6 + j];   // G Roff^T
            for (int j = 0; j < 3; j++) A[3 * k + j] = Q[3 * k] * cc[j] + Q[3 * k + 1] * cc[3 + j] + Q[3 * k + 2] * cc[6 + j];      // Q R^T
            const double q0 = Q[3 * k], q1 = Q[3 * k + 1], q2 = Q[3 * k + 2];
            const double b0 = q1 * Bz - q2 * By, b1 = -q0 * Bz + q2 * Bx, b2 = q0 * By - q1 * Bx;                              // q [B]x
            for (int j = 0; j < 3; j++) C[3 * k + j] = b0 * cc[12 + j] + b1 * cc[15 + j] + b2 * cc[18 + j];                    // ... Jr(r)
            const double s0 = pi[1] * Pz - pi[2] * Py, s1 = -pi[0] * Pz + pi[2] * Px, s2 = pi[0] * Py - pi[1] * Px;            // pi [Pbar]x
            for (int j = 0; j < 3; j++) Cb[3 * k + j] = s0 * cl[12 + j] + s1 * cl[15 + j] + s2 * cl[18 + j];                   // ... Jr(b)
        }
        for (int j = 0; j < 3; j++) {
            const double w00 = W ? W[0] : 1.0, w01 = W ? W[1] : 0.0, w10 = W ? W[2] : 0.0, w11 = W ? W[3] : 1.0;
            const double a0 = w00 * A[j] + w01 * A[3 + j], a1 = w10 * A[j] + w11 * A[3 + j];
            Jc[j] = w00 * C[j] + w01 * C[3 + j];      Jc[6 + j] = w10 * C[j] + w11 * C[3 + j];
            Jc[3 + j] = -a0;                          Jc[9 + j] = -a1;
            Jp[j] = a0;                               Jp[3 + j] = a1;
            Jl[j] = w00 * Cb[j] + w01 * Cb[3 + j];    Jl[6 + j] = w10 * Cb[j] + w11 * Cb[3 + j];
            Jl[3 + j] = -(w00 * G[j] + w01 * G[3 + j]); Jl[9 + j] = -(w10 * G[j] + w11 * G[3 + j]);
        }
    }
    return true;
}

// d res / d (f, pp[2], k[3], t[2], B[2]) of one observation, 2 x 10 row-major -- the Jacobian blocks the reference's
// AutoDiffCostFunction<..., 2, 3,3,3, 1,2,3,2,2> produces for the five intrinsic blocks when they are free
// (sbamodules/pinholecameraprojectormodule.cpp:105-122,302-533).  Closed form:
//   res = W e(Pbar, d),  d = dirFromUV(uv; intr)  ->  d res / d intr = W (d e / d d) (d d / d intr);
// the B columns are identically zero (the skew is built from the tangential block, pinholecameraprojectormodule.h:46-47).
// Returns false (and zeros) when the plane cost fails (Pbar.z < 0).
SBA_HD bool obs_intr_jac_pbar(double Px, double Py, double Pz, double u, double v, const double* in, const double* W,
                              int spherical, double* J);
SBA_HD bool obs_intr_jac(const double* cc, const double* X, double u, double v, const double* in, const double* W,
                         int spherical, double* J) {
    const double vx = X[0] - cc[9], vy = X[1] - cc[10], vz = X[2] - cc[11];
    return obs_intr_jac_pbar(cc[0] * vx + cc[1] * vy + cc[2] * vz, cc[3] * vx + cc[4] * vy + cc[5] * vz,
                             cc[6] * vx + cc[7] * vy + cc[8] * vz, u, v, in, W, spherical, J);
}
// same from the sensor-frame point Pbar (what the camera-order planes hold; with a sensor offset / lever arm Pbar is the point in
// the SENSOR frame, so this form serves every projector variant)
SBA_HD bool obs_intr_jac_pbar(double Px, double Py, double Pz, double u, double v, const double* in, const double* W,
                              int spherical, double* J) {
    for (int k = 0; k < 20; k++) J[k] = 0.0;
    // ---- d and d d / d intr (3 x 8; d_z = f)
    const double f = in[0], t1 = in[6], t2 = in[7];
    const double y = (v - in[2]) / f;
    const double D = f + t1;
    const double x = (u - in[1] - t2 * y) / D;
    const double r2 = x * x + y * y;
    const double rad = r2 * (in[3] + r2 * (in[4] + r2 * in[5]));
    const double Rp = in[3] + r2 * (2.0 * in[4] + 3.0 * r2 * in[5]);   // d rad / d r2
    const double xy = x * y;
    const double g0 = x + x * rad + 2.0 * xy * t1 + (r2 + 2.0 * x * x) * t2;
    const double g1 = y + y * rad + (r2 + 2.0 * y * y) * t1 + 2.0 * xy * t2;
    const double g0x = 1.0 + rad + 2.0 * x * x * Rp + 2.0 * y * t1 + 6.0 * x * t2, g0y = 2.0 * xy * Rp + 2.0 * x * t1 + 2.0 * y * t2;
    const double g1x = g0y, g1y = 1.0 + rad + 2.0 * y * y * Rp + 6.0 * y * t1 + 2.0 * x * t2;
    // x, y as functions of (f, ppx, ppy, t1, t2): columns 0, 1, 2, 6, 7
    const double y_f = -y / f, y_py = -1.0 / f;
    const double x_f = (t2 * y / f - x) / D, x_px = -1.0 / D, x_py = t2 / (f * D), x_t1 = -x / D, x_t2 = -y / D;
    double dd0[8], dd1[8];   // d d_x / d intr[c], d d_y / d intr[c], c = f, ppx, ppy, k1, k2, k3, t1, t2
    dd0[0] = g0 + f * (g0x * x_f + g0y * y_f); dd1[0] = g1 + f * (g1x * x_f + g1y * y_f);
    dd0[1] = f * g0x * x_px;                   dd1[1] = f * g1x * x_px;
    dd0[2] = f * (g0x * x_py + g0y * y_py);    dd1[2] = f * (g1x * x_py + g1y * y_py);
    dd0[3] = f * x * r2;                       dd1[3] = f * y * r2;
    dd0[4] = f * x * r2 * r2;                  dd1[4] = f * y * r2 * r2;
    dd0[5] = f * x * r2 * r2 * r2;             dd1[5] = f * y * r2 * r2 * r2;
    dd0[6] = f * (2.0 * xy + g0x * x_t1);              dd1[6] = f * (r2 + 2.0 * y * y + g1x * x_t1);
    dd0[7] = f * (r2 + 2.0 * x * x + g0x * x_t2);      dd1[7] = f * (2.0 * xy + g1x * x_t2);
    const double d0 = g0 * f, d1 = g1 * f, d2 = f;
    // ---- d e / d d (2 x 3): rows (ex0, 0, ex2), (0, ey1, ey2)
    double ex0, ex2, ey1, ey2;
    if (!spherical) {
        if (Pz < 0.0) return false;
        ex0 = -1.0; ey1 = -1.0; ex2 = Px / Pz; ey2 = Py / Pz;      // e = d_z Pbar_xy / Pbar_z - d_xy
    } else {
        // e_a = theta_a s_a, theta_a = atan2(Pbar_a, Pbar_z) - atan2(d_a, d_z) (as the rotated pair of obs_eval), s_a = (d_a^2 + d_z^2) / d_z:
        // d e_a / d d_a = -1 + 2 theta_a d_a / d_z,  d e_a / d d_z = d_a / d_z + theta_a (1 - d_a^2 / d_z^2)
        const double nx2 = d0 * d0 + d2 * d2, ny2 = d1 * d1 + d2 * d2;
        const double inx = 1.0 / sqrt(nx2), iny = 1.0 / sqrt(ny2);
        const double cx = d2 * inx, sx = d0 * inx, cy = d2 * iny, sy = d1 * iny;
        const double thx = atan2(cx * Px - sx * Pz, cx * Pz + sx * Px), thy = atan2(cy * Py - sy * Pz, cy * Pz + sy * Py);
        const double ax = d0 / d2, ay = d1 / d2;
        ex0 = -1.0 + 2.0 * thx * ax; ex2 = ax + thx * (1.0 - ax * ax);
        ey1 = -1.0 + 2.0 * thy * ay; ey2 = ay + thy * (1.0 - ay * ay);
    }
    double E0[8], E1[8];
    for (int c = 0; c < 8; c++) { E0[c] = ex0 * dd0[c]; E1[c] = ey1 * dd1[c]; }
    E0[0] += ex2; E1[0] += ey2;   // d d_z / d f = 1
    for (int c = 0; c < 8; c++) {
        J[c] = W ? W[0] * E0[c] + W[1] * E1[c] : E0[c];
        J[10 + c] = W ? W[2] * E0[c] + W[3] * E1[c] : E1[c];
    }
    return true;
}


// d res / d (boresight b[3], leverArm l[3]) of one observation behind a free mounting (LeverArm<UVCost,
// Body2World | Sensor2Body>, costfunctors/posedecoratorfunctors.h:148-264), 2 x 6 row-major, from the SENSOR-frame point
// Pbar = Rb^T (m - l): with G = Pi Rb^T,  d e / d l = -G,  d e / d b = Pi [Pbar]x Jr(b).  cl = pose_cache(b, l).
// Returns false (and zeros) when the plane cost fails.
SBA_HD bool obs_lever_jac_pbar(double Px, double Py, double Pz, const double* d, const double* cl, const double* W,
                               int spherical, double* J) {
    for (int k = 0; k < 12; k++) J[k] = 0.0;
    double p00, p02, p11, p12;
    if (!spherical) {
        if (Pz < 0.0) return false;
        const double iz = 1.0 / Pz, s = d[2] * iz;
        p00 = s; p11 = s; p02 = -s * (Px * iz); p12 = -s * (Py * iz);
    } else {
        const double kx = (d[0] * d[0] + d[2] * d[2]) / d[2], ky = (d[1] * d[1] + d[2] * d[2]) / d[2];
        const double qx = kx / (Px * Px + Pz * Pz), qy = ky / (Py * Py + Pz * Pz);
        p00 = qx * Pz; p02 = -qx * Px; p11 = qy * Pz; p12 = -qy * Py;
    }
    double E[12];
    const double pi0[3] = {p00, 0.0, p02}, pi1[3] = {0.0, p11, p12};
    for (int k = 0; k < 2; k++) {
        const double* pi = k ? pi1 : pi0;
        const double s0 = pi[1] * Pz - pi[2] * Py, s1 = -pi[0] * Pz + pi[2] * Px, s2 = pi[0] * Py - pi[1] * Px;   // pi [Pbar]x
        for (int j = 0; j < 3; j++) {
            E[6 * k + j] = s0 * cl[12 + j] + s1 * cl[15 + j] + s2 * cl[18 + j];                     // ... Jr(b)
            E[6 * k + 3 + j] = -(pi[0] * cl[j] + pi[1] * cl[3 + j] + pi[2] * cl[6 + j]);          // -Pi Rb^T
        }
    }
    for (int c = 0; c < 6; c++) {
        J[c] = W ? W[0] * E[c] + W[1] * E[6 + c] : E[c];
        J[6 + c] = W ? W[2] * E[c] + W[3] * E[6 + c] : E[6 + c];
    }
    return true;
}

// Pi = d e / d Pbar rebuilt from Pbar and the aux scale constants (identity stiffness):
// rows (p00, 0, p02), (0, p11, p12).  Same expressions as in obs_eval.
SBA_HD void pi_from_pbar(double Px, double Py, double Pz, double kx, double ky, int spherical,
                         double* p00, double* p02, double* p11, double* p12) {
    if (!spherical) {
        const double iz = 1.0 / Pz;
        const double s = kx * iz;
        *p00 = s; *p11 = s; *p02 = -s * (Px * iz); *p12 = -s * (Py * iz);
    } else {
        const double qx = kx / (Px * Px + Pz * Pz), qy = ky / (Py * Py + Pz * Pz);
        *p00 = qx * Pz; *p02 = -qx * Px; *p11 = qy * Pz; *p12 = -qy * Py;
    }
}

// symmetric 3x3 inverse by adjugate; H = (h00,h01,h02,h11,h12,h22) -> same packing.
// Returns false if the determinant is 0 or not finite.
SBA_HD bool inv3_sym(const double* H, double* Hi) {
    const double a = H[0], b = H[1], c = H[2], d = H[3], e = H[4], f = H[5];
    const double C00 = d * f - e * e, C01 = c * e - b * f, C02 = b * e - c * d;
    const double det = a * C00 + b * C01 + c * C02;
    if (!(fabs(det) > 0.0) || !isfinite(det)) return false;
    const double id = 1.0 / det;
    Hi[0] = C00 * id; Hi[1] = C01 * id; Hi[2] = C02 * id;
    Hi[3] = (a * f - c * c) * id; Hi[4] = (b * c - a * e) * id; Hi[5] = (a * d - b * b) * id;
    return true;
}

// packed upper-triangular index of a symmetric 6x6 (21 entries, row-major upper)
SBA_HD int sym6(int i, int j) {
    if (i > j) { const int t = i; i = j; j = t; }
    return i * 6 - (i * (i - 1)) / 2 + (j - i);
}

// inverse of an SPD 6x6 (full row-major in/out) by Cholesky; false if not PD
SBA_HD bool inv6_spd(const double* A, double* Ai) {
    double L[36];
    for (int i = 0; i < 36; i++) L[i] = A[i];
    for (int j = 0; j < 6; j++) {
        double dd = L[j * 6 + j];
        for (int k = 0; k < j; k++) dd -= L[j * 6 + k] * L[j * 6 + k];
        if (!(dd > 0.0)) return false;
        dd = sqrt(dd);
        L[j * 6 + j] = dd;
        for (int i = j + 1; i < 6; i++) {
            double s = L[i * 6 + j];
            for (int k = 0; k < j; k++) s -= L[i * 6 + k] * L[j * 6 + k];
            L[i * 6 + j] = s / dd;
        }
    }
    for (int j = 0; j < 6; j++) {
        double e[6];
        for (int i = 0; i < 6; i++) e[i] = (i == j) ? 1.0 : 0.0;
        for (int i = 0; i < 6; i++) { double s = e[i]; for (int k = 0; k < i; k++) s -= L[i * 6 + k] * e[k]; e[i] = s / L[i * 6 + i]; }
        for (int i = 5; i >= 0; i--) { double s = e[i]; for (int k = i + 1; k < 6; k++) s -= L[k * 6 + i] * e[k]; e[i] = s / L[i * 6 + i]; }
        for (int i = 0; i < 6; i++) Ai[i * 6 + j] = e[i];
    }
    return true;
}

// dirFromUV divided by its z component (= f): the normalised view ray (x, y, 1) the DLT needs, with the two reciprocals
// 1 / f and 1 / (f + t1) supplied by the caller (per camera MODEL, not per view): no division per view.
SBA_HD void dir_from_uv_xy(double u, double v, const double* in, double inv_f, double inv_fx, double* xo, double* yo) {
    const double y = (v - in[2]) * inv_f;
    const double x = (u - in[1] - in[7] * y) * inv_fx;
    const double r2 = x * x + y * y;
    const double rad = r2 * (in[3] + r2 * (in[4] + r2 * in[5]));
    const double xy = x * y;
    *xo = x + x * rad + 2.0 * xy * in[6] + (r2 + 2.0 * x * x) * in[7];
    *yo = y + y * rad + (r2 + 2.0 * y * y) * in[6] + 2.0 * xy * in[7];
}
// DLT camera record: 8 doubles = the unit quaternion (w, x, y, z) of Rt (the world -> sensor rotation of the cache record) and
// c = Rt t -- 64 bytes, two sectors per view instead of the three of {Rt, t}; the kernel rebuilds Rt from the quaternion
// (18 flops) and no longer computes c per view.  Shepperd's branch on the largest of (trace, R00, R11, R22).
SBA_HD void dlt_record(const double* cc, double* rec) {
    const double m00 = cc[0], m01 = cc[1], m02 = cc[2], m10 = cc[3], m11 = cc[4], m12 = cc[5], m20 = cc[6], m21 = cc[7], m22 = cc[8];
    const double tr = m00 + m11 + m22;
    double w, x, y, z;
    if (tr >= m00 && tr >= m11 && tr >= m22) {
        const double s = 2.0 * sqrt(1.0 + tr);
        w = 0.25 * s; x = (m21 - m12) / s; y = (m02 - m20) / s; z = (m10 - m01) / s;
    } else if (m00 >= m11 && m00 >= m22) {
        const double s = 2.0 * sqrt(1.0 + m00 - m11 - m22);
        w = (m21 - m12) / s; x = 0.25 * s; y = (m01 + m10) / s; z = (m02 + m20) / s;
    } else if (m11 >= m22) {
        const double s = 2.0 * sqrt(1.0 + m11 - m00 - m22);
        w = (m02 - m20) / s; x = (m01 + m10) / s; y = 0.25 * s; z = (m12 + m21) / s;
    } else {
        const double s = 2.0 * sqrt(1.0 + m22 - m00 - m11);
        w = (m10 - m01) / s; x = (m02 + m20) / s; y = (m12 + m21) / s; z = 0.25 * s;
    }
    const double n = 1.0 / sqrt(w * w + x * x + y * y + z * z);
    rec[0] = w * n; rec[1] = x * n; rec[2] = y * n; rec[3] = z * n;
    rec[4] = m00 * cc[9] + m01 * cc[10] + m02 * cc[11];
    rec[5] = m10 * cc[9] + m11 * cc[10] + m12 * cc[11];
    rec[6] = m20 * cc[9] + m21 * cc[10] + m22 * cc[11];
    rec[7] = 0.0;
}
// one DLT view from the record: rows of Rt rebuilt from the quaternion, ray normalised to (x, y, 1)
SBA_HD void dlt_accumulate_rec(double qw, double qx, double qy, double qz, double c0, double c1, double c2, double x, double y, double* A,
                               double* b) {
    const double xx = qx * qx, yy = qy * qy, zz = qz * qz, xy = qx * qy, xz = qx * qz, yz = qy * qz, wx = qw * qx, wy = qw * qy, wz = qw * qz;
    const double m00 = 1.0 - 2.0 * (yy + zz), m01 = 2.0 * (xy - wz), m02 = 2.0 * (xz + wy);
    const double m10 = 2.0 * (xy + wz), m11 = 1.0 - 2.0 * (xx + zz), m12 = 2.0 * (yz - wx);
    const double m20 = 2.0 * (xz - wy), m21 = 2.0 * (yz + wx), m22 = 1.0 - 2.0 * (xx + yy);
    const double r0x = x * m20 - m00, r0y = x * m21 - m01, r0z = x * m22 - m02;
    const double r1x = y * m20 - m10, r1y = y * m21 - m11, r1z = y * m22 - m12;
    const double h0 = x * c2 - c0, h1 = y * c2 - c1;
    b[0] += r0x * h0 + r1x * h1; b[1] += r0y * h0 + r1y * h1; b[2] += r0z * h0 + r1z * h1;
    A[0] += r0x * r0x + r1x * r1x; A[1] += r0x * r0y + r1x * r1y; A[2] += r0x * r0z + r1x * r1z;
    A[3] += r0y * r0y + r1y * r1y; A[4] += r0y * r0z + r1y * r1z; A[5] += r0z * r0z + r1z * r1z;
}

// dlt_accumulate for a ray already normalised to (x, y, 1)
SBA_HD void dlt_accumulate_xy(const double* cc, double x, double y, double* A, double* b) {
    const double c0 = cc[0] * cc[9] + cc[1] * cc[10] + cc[2] * cc[11];
    const double c1 = cc[3] * cc[9] + cc[4] * cc[10] + cc[5] * cc[11];
    const double c2 = cc[6] * cc[9] + cc[7] * cc[10] + cc[8] * cc[11];
    const double r0x = x * cc[6] - cc[0], r0y = x * cc[7] - cc[1], r0z = x * cc[8] - cc[2];
    const double r1x = y * cc[6] - cc[3], r1y = y * cc[7] - cc[4], r1z = y * cc[8] - cc[5];
    const double h0 = x * c2 - c0, h1 = y * c2 - c1;
    b[0] += r0x * h0 + r1x * h1; b[1] += r0y * h0 + r1y * h1; b[2] += r0z * h0 + r1z * h1;
    A[0] += r0x * r0x + r1x * r1x; A[1] += r0x * r0y + r1x * r1y; A[2] += r0x * r0z + r1x * r1z;
    A[3] += r0y * r0y + r1y * r1y; A[4] += r0y * r0z + r1y * r1z; A[5] += r0z * r0z + r1z * r1z;
}

// One DLT view: accumulates the inhomogeneous normal equations of the two rows
//   (x r3 - r1) . X = x c3 - c1,  (y r3 - r2) . X = y c3 - c2
// (r_i rows of R^T, c = R^T t, (x,y) = dirFromUV.xy / dir.z).  A packed sym 3x3.
SBA_HD void dlt_accumulate(const double* cc, const double* d, double* A, double* b) {
    const double x = d[0] / d[2], y = d[1] / d[2];
    const double c0 = cc[0] * cc[9] + cc[1] * cc[10] + cc[2] * cc[11];
    const double c1 = cc[3] * cc[9] + cc[4] * cc[10] + cc[5] * cc[11];
    const double c2 = cc[6] * cc[9] + cc[7] * cc[10] + cc[8] * cc[11];
    const double r0x = x * cc[6] - cc[0], r0y = x * cc[7] - cc[1], r0z = x * cc[8] - cc[2];
    const double r1x = y * cc[6] - cc[3], r1y = y * cc[7] - cc[4], r1z = y * cc[8] - cc[5];
    const double h0 = x * c2 - c0, h1 = y * c2 - c1;
    b[0] += r0x * h0 + r1x * h1; b[1] += r0y * h0 + r1y * h1; b[2] += r0z * h0 + r1z * h1;
    A[0] += r0x * r0x + r1x * r1x; A[1] += r0x * r0y + r1x * r1y; A[2] += r0x * r0z + r1x * r1z;
    A[3] += r0y * r0y + r1y * r1y; A[4] += r0y * r0z + r1y * r1z; A[5] += r0z * r0z + r1z * r1z;
}

}  // namespace sba
