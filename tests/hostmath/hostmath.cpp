// hostmath.cpp -- TEST INFRASTRUCTURE: compiles the product's __host__ __device__ math header
// (steviapp_b200/csrc/sba_math.cuh) with g++ so that the closed-form derivatives can be checked
// against the oracle's autodiff on a machine without a GPU.  Never part of the product path.
#include <stdint.h>

#include "../../steviapp_b200/csrc/sba_math.cuh"
#include "../../steviapp_b200/csrc/sba_jet.cuh"

extern "C" {

// per observation i: camera (r,t) index cam[i], point pt[i]
int hm_evaluate(int32_t n_obs, const int32_t* cam, const int32_t* pt, const double* cam_r, const double* cam_t,
                const uint8_t* kind, const int32_t* cam_model, const double* intr, const double* pts,
                const double* uv, const double* info, double* res, double* jc, double* jp, uint8_t* failed) {
    int bad = 0;
    for (int32_t i = 0; i < n_obs; i++) {
        double cc[sba::CAM_CACHE_STRIDE];
        sba::pose_cache(cam_r + 3 * cam[i], cam_t + 3 * cam[i], cc);
        double d[3];
        sba::dir_from_uv(uv[2 * i], uv[2 * i + 1], intr + 10 * cam_model[cam[i]], d);
        double r[2] = {0, 0}, Jc[12] = {0}, Jp[6] = {0};
        const bool ok = sba::obs_eval<true>(cc, pts + 3 * pt[i], d, info ? info + 4 * i : nullptr, kind ? kind[cam[i]] : 0, r, Jc, Jp);
        if (!ok) { bad++; for (int k = 0; k < 12; k++) Jc[k] = 0; for (int k = 0; k < 6; k++) Jp[k] = 0; r[0] = r[1] = 0; }
        if (failed) failed[i] = ok ? 0 : 1;
        res[2 * i] = r[0]; res[2 * i + 1] = r[1];
        for (int k = 0; k < 12; k++) jc[12 * i + k] = Jc[k];
        for (int k = 0; k < 6; k++) jp[6 * i + k] = Jp[k];
    }
    return bad;
}

// intrinsics Jacobian (2 x 10 per observation), SURVEY 8(f) row 1
int hm_intrinsics_jac(int32_t n_obs, const int32_t* cam, const int32_t* pt, const double* cam_r, const double* cam_t,
                      const uint8_t* kind, const int32_t* cam_model, const double* intr, const double* pts,
                      const double* uv, const double* info, double* jac) {
    int bad = 0;
    for (int32_t i = 0; i < n_obs; i++) {
        double cc[sba::CAM_CACHE_STRIDE];
        sba::pose_cache(cam_r + 3 * cam[i], cam_t + 3 * cam[i], cc);
        if (!sba::obs_intr_jac(cc, pts + 3 * pt[i], uv[2 * i], uv[2 * i + 1], intr + 10 * cam_model[cam[i]],
                               info ? info + 4 * i : nullptr, kind ? kind[cam[i]] : 0, jac + 20 * (size_t)i)) bad++;
    }
    return bad;
}

// constant sensor offset: n independent samples; off_rt = Roff^T row-major [9] + toff [3] per sample
int hm_eval_offset(int32_t n, const double* r, const double* t, const double* off_rt, const double* pts, const double* intr10,
                   const double* uv, const double* info, int spherical, double* res, double* jc, double* jp, uint8_t* failed) {
    int bad = 0;
    for (int32_t i = 0; i < n; i++) {
        double cc[sba::CAM_CACHE_STRIDE], d[3];
        sba::pose_cache(r + 3 * i, t + 3 * i, cc);
        sba::dir_from_uv(uv[2 * i], uv[2 * i + 1], intr10, d);
        double rr[2] = {0, 0}, Jc[12] = {0}, Jp[6] = {0};
        const bool ok = sba::obs_eval_offset<true>(cc, off_rt + 12 * i, pts + 3 * i, d, info ? info + 4 * i : nullptr, spherical, rr, Jc, Jp);
        if (!ok) { bad++; for (int k = 0; k < 12; k++) Jc[k] = 0; for (int k = 0; k < 6; k++) Jp[k] = 0; rr[0] = rr[1] = 0; }
        if (failed) failed[i] = ok ? 0 : 1;
        res[2 * i] = rr[0]; res[2 * i + 1] = rr[1];
        for (int k = 0; k < 12; k++) jc[12 * i + k] = Jc[k];
        for (int k = 0; k < 6; k++) jp[6 * i + k] = Jp[k];
    }
    return bad;
}

// free boresight b / lever arm l behind a constant offset: n independent samples
int hm_eval_leverarm(int32_t n, const double* r, const double* t, const double* b, const double* l, const double* off_rt, const double* pts,
                     const double* intr10, const double* uv, const double* info, int spherical, double* res, double* jc, double* jl,
                     double* jp, uint8_t* failed) {
    int bad = 0;
    for (int32_t i = 0; i < n; i++) {
        double cc[sba::CAM_CACHE_STRIDE], cl[sba::CAM_CACHE_STRIDE], d[3];
        sba::pose_cache(r + 3 * i, t + 3 * i, cc);
        sba::pose_cache(b + 3 * i, l + 3 * i, cl);
        sba::dir_from_uv(uv[2 * i], uv[2 * i + 1], intr10, d);
        double rr[2] = {0, 0}, Jc[12] = {0}, Jl[12] = {0}, Jp[6] = {0};
        const bool ok = sba::obs_eval_leverarm<true>(cc, off_rt + 12 * i, cl, pts + 3 * i, d, info ? info + 4 * i : nullptr, spherical, rr, Jc, Jl, Jp);
        if (!ok) { bad++; for (int k = 0; k < 12; k++) { Jc[k] = 0; Jl[k] = 0; } for (int k = 0; k < 6; k++) Jp[k] = 0; rr[0] = rr[1] = 0; }
        if (failed) failed[i] = ok ? 0 : 1;
        res[2 * i] = rr[0]; res[2 * i + 1] = rr[1];
        for (int k = 0; k < 12; k++) { jc[12 * i + k] = Jc[k]; jl[12 * i + k] = Jl[k]; }
        for (int k = 0; k < 6; k++) jp[6 * i + k] = Jp[k];
    }
    return bad;
}

int hm_inv6(const double* A, double* Ai) { return sba::inv6_spd(A, Ai) ? 0 : 1; }
int hm_inv3(const double* H6, double* Hi6) { return sba::inv3_sym(H6, Hi6) ? 0 : 1; }
void hm_pose_cache(const double* r, const double* t, double* out24) { sba::pose_cache(r, t, out24); }


// UV <-> UV factor (sba_jet.cuh: the functor the pair kernels run): raw residuals (4) and the Jacobian 4 x 28 with respect to
// pose1 (6), pose2 (6) and the 8 intrinsic columns of camera 1's and camera 2's model (chain rule through the view rays)
void hm_uv2uv(int32_t n, const double* r1, const double* t1, const double* r2, const double* t2, const double* intr1, const double* intr2,
              const double* uv, const double* info, double* res, double* jac) {
    typedef sba::DJet<18> J;
    for (int32_t k = 0; k < n; k++) {
        double d1[3], d2[3], dd1[24], dd2[24];
        sba::dir_from_uv_jac(uv[4 * k], uv[4 * k + 1], intr1, d1, dd1);
        sba::dir_from_uv_jac(uv[4 * k + 2], uv[4 * k + 3], intr2, d2, dd2);
        J a[3], b[3], c[3], d[3], e[3], f[3], out[4];
        for (int i = 0; i < 3; i++) {
            a[i] = J(r1[3 * k + i], i); b[i] = J(t1[3 * k + i], 3 + i); c[i] = J(r2[3 * k + i], 6 + i); d[i] = J(t2[3 * k + i], 9 + i);
            e[i] = J(d1[i], 12 + i); f[i] = J(d2[i], 15 + i);
        }
        sba::uv2uv_residual<J>(a, b, c, d, e, f, info ? info + 8 * k : nullptr, info ? info + 8 * k + 4 : nullptr, out);
        for (int i = 0; i < 4; i++) {
            res[4 * k + i] = out[i].a;
            for (int j = 0; j < 12; j++) jac[(4 * k + i) * 28 + j] = out[i].v[j];
            for (int j = 0; j < 8; j++) {
                jac[(4 * k + i) * 28 + 12 + j] = out[i].v[12] * dd1[j] + out[i].v[13] * dd1[8 + j] + out[i].v[14] * dd1[16 + j];
                jac[(4 * k + i) * 28 + 20 + j] = out[i].v[15] * dd2[j] + out[i].v[16] * dd2[8 + j] + out[i].v[17] * dd2[16 + j];
            }
        }
    }
}
void hm_loss(int kind, double a, double s, double* rho3, double* corr3) {
    sba::loss_rho(kind, a, s, rho3);
    sba::loss_corrector(s, rho3, corr3, corr3 + 1, corr3 + 2);
}

}
