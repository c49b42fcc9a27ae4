// hostmath.cpp -- TEST INFRASTRUCTURE: compiles the product's __host__ __device__ math header
// (steviapp_b200/csrc/sba_math.cuh) with g++ so that the closed-form derivatives can be checked
// against the oracle's autodiff on a machine without a GPU.  Never part of the product path.
#include <stdint.h>

#include "../../steviapp_b200/csrc/sba_math.cuh"

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

}
