// sba_kernels_misc.cu -- parameter update / restore, output gathers for the
// evaluate hook, layout conversion and K5 batched DLT triangulation.
#include <algorithm>

#include <cstdlib>

#include "sba_kernel_util.cuh"

namespace sba {

// candidate x + delta: ambient plus on every block (no manifold is attached to rAxis in the
// reference: AddParameterBlock(ptr,3), modularsbasolver.cpp:837-838).  Saves the old values.
__global__ void k_apply_step(DevProblem P, const double* __restrict__ dc) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P.Nc) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const double r = P.cam_r[3 * (size_t)i + a], t = P.cam_t[3 * (size_t)i + a];
            P.cam_saved[6 * (size_t)i + a] = r;
            P.cam_saved[6 * (size_t)i + 3 + a] = t;
            P.cam_r[3 * (size_t)i + a] = r + dc[6 * (size_t)i + a];
            P.cam_t[3 * (size_t)i + a] = t + dc[6 * (size_t)i + 3 + a];
        }
    }
    if (i < P.Np) {
        double* X = P.pts + 4 * (size_t)i;
        const double* d = P.dp + 4 * (size_t)i;
        double* S = P.pts_saved + 4 * (size_t)i;
        const double2 x01 = *reinterpret_cast<const double2*>(X), x23 = *reinterpret_cast<const double2*>(X + 2);
        const double2 d01 = ldg2(d), d23 = ldg2(d + 2);
        st2(S, x01.x, x01.y); st2(S + 2, x23.x, x23.y);
        st2(X, x01.x + d01.x, x01.y + d01.y); st2(X + 2, x23.x + d23.x, x23.y);
    }
}
void launch_apply_step(const DevProblem& P, const double* dc, cudaStream_t st) {
    const int n = max(P.Nc, P.Np);
    k_apply_step<<<cdiv(n, 256), 256, 0, st>>>(P, dc);
}

__global__ void k_restore(DevProblem P) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P.Nc) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            P.cam_r[3 * (size_t)i + a] = P.cam_saved[6 * (size_t)i + a];
            P.cam_t[3 * (size_t)i + a] = P.cam_saved[6 * (size_t)i + 3 + a];
        }
    }
    if (i < P.Np) {
        const double* S = P.pts_saved + 4 * (size_t)i;
        double* X = P.pts + 4 * (size_t)i;
        const double2 s01 = ldg2(S), s23 = ldg2(S + 2);
        st2(X, s01.x, s01.y); st2(X + 2, s23.x, s23.y);
    }
}
void launch_restore(const DevProblem& P, cudaStream_t st) {
    const int n = max(P.Nc, P.Np);
    k_restore<<<cdiv(n, 256), 256, 0, st>>>(P);
}

// evaluate hook: caller-order residuals and full Jacobians (Jc = [C | -A], Jp = A)
__global__ void k_gather_outputs(DevProblem P, const int32_t* __restrict__ slot_of_obs, double* res_out,
                                 double* jc_out, double* jp_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.No) return;
    const int si = slot_of_obs[i];
    const size_t s = si, Ns = (size_t)P.Ns;
    if (res_out) { res_out[2 * (size_t)i] = P.res[s]; res_out[2 * (size_t)i + 1] = P.res[Ns + s]; }
    double A[6], C[6];
#pragma unroll
    for (int a = 0; a < 6; a++) A[a] = P.Ac[(size_t)a * Ns + s];
    if (jc_out) {
        int lo = 0, hi = P.Nt - 1;
        while (lo < hi) {   // camera of the slot: last tile with tile_start <= s
            const int mid = (lo + hi + 1) >> 1;
            if (__ldg(P.tile_start + mid) <= si) lo = mid; else hi = mid - 1;
        }
        const CamRQ crq = load_cam_rq(P, __ldg(P.tile_cam + lo));
        rebuild_C(A, P.Pc[s], P.Pc[Ns + s], P.Pc[2 * Ns + s], crq, C);
    }
#pragma unroll
    for (int row = 0; row < 2; row++)
#pragma unroll
        for (int a = 0; a < 3; a++) {
            if (jc_out) {
                jc_out[12 * (size_t)i + 6 * row + a] = C[3 * row + a];
                jc_out[12 * (size_t)i + 6 * row + 3 + a] = -A[3 * row + a];
            }
            if (jp_out) jp_out[6 * (size_t)i + 3 * row + a] = A[3 * row + a];
        }
}
void launch_gather_outputs(const DevProblem& P, const int32_t* slot_of_obs, double* res_out, double* jc_out,
                           double* jp_out, cudaStream_t st) {
    if (P.No == 0) return;
    k_gather_outputs<<<cdiv(P.No, 256), 256, 0, st>>>(P, slot_of_obs, res_out, jc_out, jp_out);
}

// intrinsics Jacobian hook: d res / d (f, pp, k, t, B), 2 x 10 per observation, caller order.
// Thread per observation; the camera of a slot is found by bisection of the tile starts (slots are camera-sorted).
__global__ void __launch_bounds__(128) k_intrinsics_jacobian(DevProblem P, const int32_t* __restrict__ slot_of_obs, double* jac_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.No) return;
    const int s = slot_of_obs[i];
    int lo = 0, hi = P.Nt - 1;
    while (lo < hi) {   // last tile with tile_start <= s
        const int mid = (lo + hi + 1) >> 1;
        if (__ldg(P.tile_start + mid) <= s) lo = mid; else hi = mid - 1;
    }
    const int cam = __ldg(P.tile_cam + lo), pt = P.obs_pt[s];
    const size_t Ns = (size_t)P.Ns;
    double W[4];
    if (P.obs_info) for (int k = 0; k < 4; k++) W[k] = P.obs_info[(size_t)k * Ns + s];
    double J[20];
    obs_intr_jac(P.cam_cache + (size_t)CAM_CACHE_STRIDE * cam, P.pts + 4 * (size_t)pt, P.obs_uv[2 * (size_t)s], P.obs_uv[2 * (size_t)s + 1],
                 P.intr + 10 * P.cam_model[cam], P.obs_info ? W : nullptr, P.cam_kind[cam], J);
    for (int k = 0; k < 20; k++) jac_out[20 * (size_t)i + k] = J[k];
}
void launch_intrinsics_jacobian(const DevProblem& P, const int32_t* slot_of_obs, double* jac_out, cudaStream_t st) {
    if (P.No) k_intrinsics_jacobian<<<cdiv(P.No, 128), 128, 0, st>>>(P, slot_of_obs, jac_out);
}

// constant-sensor-offset hook: residuals and Jacobians of every observation with the per-camera offset
// off[12 cam] = Roff^T row-major + toff, caller order; failed evaluations (plane cost behind the composed sensor pose) write
// zeros and are counted.  Thread per observation, same camera lookup as the intrinsics hook.
__global__ void __launch_bounds__(128) k_evaluate_offset(DevProblem P, const int32_t* __restrict__ slot_of_obs, const double* __restrict__ off,
                                                         double* res_out, double* jc_out, double* jp_out, int* n_failed) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.No) return;
    const int s = slot_of_obs[i];
    int lo = 0, hi = P.Nt - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (__ldg(P.tile_start + mid) <= s) lo = mid; else hi = mid - 1;
    }
    const int cam = __ldg(P.tile_cam + lo), pt = P.obs_pt[s];
    const size_t Ns = (size_t)P.Ns;
    double W[4], d[3], r[2] = {0.0, 0.0}, Jc[12], Jp[6];
    if (P.obs_info) for (int k = 0; k < 4; k++) W[k] = P.obs_info[(size_t)k * Ns + s];
    dir_from_uv(P.obs_uv[2 * (size_t)s], P.obs_uv[2 * (size_t)s + 1], P.intr + 10 * P.cam_model[cam], d);
    const bool ok = obs_eval_offset<true>(P.cam_cache + (size_t)CAM_CACHE_STRIDE * cam, off + 12 * (size_t)cam, P.pts + 4 * (size_t)pt, d,
                                          P.obs_info ? W : nullptr, P.cam_kind[cam], r, Jc, Jp);
    if (!ok) {
        atomicAdd(n_failed, 1);
        r[0] = r[1] = 0.0;
        for (int k = 0; k < 12; k++) Jc[k] = 0.0;
        for (int k = 0; k < 6; k++) Jp[k] = 0.0;
    }
    res_out[2 * (size_t)i] = r[0]; res_out[2 * (size_t)i + 1] = r[1];
    for (int k = 0; k < 12; k++) jc_out[12 * (size_t)i + k] = Jc[k];
    for (int k = 0; k < 6; k++) jp_out[6 * (size_t)i + k] = Jp[k];
}
void launch_evaluate_offset(const DevProblem& P, const int32_t* slot_of_obs, const double* off, double* res_out, double* jc_out,
                            double* jp_out, int* n_failed, cudaStream_t st) {
    if (P.No) k_evaluate_offset<<<cdiv(P.No, 128), 128, 0, st>>>(P, slot_of_obs, off, res_out, jc_out, jp_out, n_failed);
}

__global__ void k_pack_points(const double* __restrict__ xyz3, double* pts4, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    st2(pts4 + 4 * (size_t)i, xyz3[3 * (size_t)i], xyz3[3 * (size_t)i + 1]);
    st2(pts4 + 4 * (size_t)i + 2, xyz3[3 * (size_t)i + 2], 0.0);
}
void launch_pack_points(const double* xyz3, double* pts4, int n, cudaStream_t st) {
    if (n) k_pack_points<<<cdiv(n, 256), 256, 0, st>>>(xyz3, pts4, n);
}
__global__ void k_unpack_points(const double* __restrict__ pts4, double* xyz3, int n, int stride_in) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
#pragma unroll
    for (int a = 0; a < 3; a++) xyz3[3 * (size_t)i + a] = pts4[(size_t)stride_in * i + a];
}
void launch_unpack_points(const double* pts4, double* xyz3, int n, cudaStream_t st) {
    if (n) k_unpack_points<<<cdiv(n, 256), 256, 0, st>>>(pts4, xyz3, n, 4);
}

// expand packed symmetric blocks to the oracle's full row-major shapes (test hook)
__global__ void k_unpack_blocks(DevProblem P, double* Hcc36, double* Hpp9, double* gp3) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (Hcc36 && i < P.Nc) {
        for (int a = 0; a < 6; a++)
            for (int b = 0; b < 6; b++) Hcc36[36 * (size_t)i + 6 * a + b] = P.Hcc[21 * (size_t)i + sym6(a, b)];
    }
    if (i < P.Np) {
        if (Hpp9) {
            const double* H = P.Hpp + 8 * (size_t)i;
            double* o = Hpp9 + 9 * (size_t)i;
            o[0] = H[0]; o[1] = H[1]; o[2] = H[2]; o[3] = H[1]; o[4] = H[3]; o[5] = H[4]; o[6] = H[2]; o[7] = H[4]; o[8] = H[5];
        }
        if (gp3) for (int a = 0; a < 3; a++) gp3[3 * (size_t)i + a] = P.gp[4 * (size_t)i + a];
    }
}
void launch_unpack_blocks(const DevProblem& P, double* Hcc36, double* Hpp9, double* gp3, cudaStream_t st) {
    const int n = max(P.Nc, P.Np);
    if (n) k_unpack_blocks<<<cdiv(n, 256), 256, 0, st>>>(P, Hcc36, Hpp9, gp3);
}

// test hook: explicit damping diagonals (lam_p given as 3*Np AoS)
__global__ void k_set_lambda(DevProblem P, const double* lam_c, const double* lam_p3) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 6 * P.Nc) P.lam_c[i] = lam_c[i];
    if (i < P.Np) {
        for (int a = 0; a < 3; a++) P.lam_p[4 * (size_t)i + a] = lam_p3[3 * (size_t)i + a];
        P.lam_p[4 * (size_t)i + 3] = 0.0;
    }
}
void launch_set_lambda(const DevProblem& P, const double* lam_c, const double* lam_p3, cudaStream_t st) {
    const int n = max(6 * P.Nc, P.Np);
    if (n) k_set_lambda<<<cdiv(n, 256), 256, 0, st>>>(P, lam_c, lam_p3);
}

// ---------------------------------------------------------------------------
// K5 batched multi-view DLT: thread per point, inhomogeneous normal equations accumulated over its views in registers,
// closed-form symmetric 3x3 solve.  Per view: 16 B pixel + 4 B camera index streamed (both prefetched one view ahead), the
// camera's DLT record gathered from L1 / L2: 64 B = two sectors {unit quaternion of Rt, Rt t} (k_dlt_records, once per call)
// instead of the 96 B of {Rt, t}; Rt is rebuilt from the quaternion in registers.  NO division per view: the ray is
// normalised by construction (dirFromUV.xy / dirFromUV.z, where z = f cancels) and the two reciprocals of the lens model
// are taken once per camera MODEL (re-read only when the model changes between two views of the point: one model per
// Camera datablock in the reference, almost never).  24 B + 1 B out per point.  SURVEY 8(d) bytes: 72 B per view (pose + ray)
// + 24 B per point.  Round-2 steps on 2 M points / 10 M views (scripts/prof_dlt.py): 0.287 ms -> 0.243 (divisions) -> 0.203
// (record) -> 0.194 (pixel prefetch).  (Eight lanes per point with a shuffle tree was 2x slower; the camera table in shared
// memory and a 64-register cap were slower too.)
// ---------------------------------------------------------------------------
__global__ void k_dlt_records(DevProblem P) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= P.Nc) return;
    double cc[12], rec[8];
#pragma unroll
    for (int i = 0; i < 12; i++) cc[i] = P.cam_cache[(size_t)CAM_CACHE_STRIDE * c + i];
    dlt_record(cc, rec);
#pragma unroll
    for (int i = 0; i < 8; i++) P.dlt_rec[8 * (size_t)c + i] = rec[i];
}
__global__ void __launch_bounds__(256) k_triangulate(DevProblem P, int n_points, const int32_t* __restrict__ view_off,
                                                     const int32_t* __restrict__ view_cam,
                                                     const double* __restrict__ view_uv, double* xyz_out,
                                                     uint8_t* ok_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_points) return;
    double A[6] = {0, 0, 0, 0, 0, 0}, b[3] = {0, 0, 0};
    const int k0 = __ldg(view_off + i), k1 = __ldg(view_off + i + 1);
    int model = -1;
    double intr[10], inv_f = 0.0, inv_fx = 0.0;   // the two reciprocals of the lens model: once per camera model, not per view
    int cam_next = k0 < k1 ? __ldg(view_cam + k0) : 0;
    double2 uv_next = k0 < k1 ? ldg2(view_uv + 2 * (size_t)k0) : make_double2(0.0, 0.0);
    for (int k = k0; k < k1; k++) {
        const int cam = cam_next;
        const double2 uv = uv_next;
        if (k + 1 < k1) {   // next index and pixel in flight during this view's gather and arithmetic
            cam_next = __ldg(view_cam + k + 1);
            uv_next = ldg2(view_uv + 2 * (size_t)(k + 1));
        }
        const double* rc = P.dlt_rec + 8 * (size_t)cam;
        const d4 q = ld256(rc), cv = ld256(rc + 4);   // 64 bytes: two sectors
        const int m = __ldg(P.cam_model + cam);
        if (m != model) {
            model = m;
            const double* in = P.intr + 10 * m;
#pragma unroll
            for (int j = 0; j < 5; j++) { const double2 v = ldg2(in + 2 * j); intr[2 * j] = v.x; intr[2 * j + 1] = v.y; }
            inv_f = 1.0 / intr[0]; inv_fx = 1.0 / (intr[0] + intr[6]);
        }
        double x, y;   // dirFromUV.xy / dirFromUV.z: the factor f of the reference's ray cancels, no division per view
        dir_from_uv_xy(uv.x, uv.y, intr, inv_f, inv_fx, &x, &y);
        dlt_accumulate_rec(q.x, q.y, q.z, q.w, cv.x, cv.y, cv.z, x, y, A, b);
    }
    double Ai[6];
    const bool ok = inv3_sym(A, Ai);
    double x0 = 0.0, x1 = 0.0, x2 = 0.0;
    if (ok) {
        x0 = Ai[0] * b[0] + Ai[1] * b[1] + Ai[2] * b[2];
        x1 = Ai[1] * b[0] + Ai[3] * b[1] + Ai[4] * b[2];
        x2 = Ai[2] * b[0] + Ai[4] * b[1] + Ai[5] * b[2];
    }
    xyz_out[3 * (size_t)i] = x0; xyz_out[3 * (size_t)i + 1] = x1; xyz_out[3 * (size_t)i + 2] = x2;
    if (ok_out) ok_out[i] = ok ? 1 : 0;
}
void launch_triangulate(const DevProblem& P, int n_points, const int32_t* view_off, const int32_t* view_cam,
                        const double* view_uv, double* xyz_out, uint8_t* ok_out, cudaStream_t st) {
    // (a variant with the camera table in shared memory, persistent 1024-thread blocks, was measured slower: 1.84 vs 1.49 ms)
    if (!n_points) return;
    if (P.Nc) k_dlt_records<<<cdiv(P.Nc, 128), 128, 0, st>>>(P);   // after the pose cache (the callers launch it first)
    k_triangulate<<<cdiv(n_points, 256), 256, 0, st>>>(P, n_points, view_off, view_cam, view_uv, xyz_out, ok_out);
}

}  // namespace sba
