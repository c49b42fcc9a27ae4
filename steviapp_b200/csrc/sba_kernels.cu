// sba_kernels.cu -- hand-written sm_100a FP64 kernels of the SBA hot path.
//
// All reductions are deterministic: fixed-shape warp-shuffle trees inside a
// warp tile, fixed-order sums over the tiles of a camera, single-block trees
// for scalars.  No floating-point atomics anywhere (integer flag stores only).
//
// Reference arithmetic restated (closed form instead of autodiff):
//   K1/K1p : InvertPose<UV2ParametrizedXYZCost<PinholeUVProjector,..>>  (sbamodules/
//            pinholecameraprojectormodule.cpp:18-19, costfunctors/modularuvprojection.h:368-448)
//   K2..K4 : the Schur elimination / SCHUR_JACOBI PCG / back-substitution that ceres::Solve
//            performs for the reference (modularsbasolver.cpp:938); in-tree analogue
//            modules/g2o/g2o/core/block_solver.hpp:329-479, solvers/pcg/linear_solver_pcg.hpp:84-165
#include <cstdlib>

#include "sba_kernel_util.cuh"

namespace sba {

// ---------------------------------------------------------------------------
// K0 pose cache: thread per camera
// ---------------------------------------------------------------------------
__global__ void k_mount_cache(DevProblem P) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= P.Nmt) return;
    double out[CAM_CACHE_STRIDE];
    pose_cache(P.mount + 6 * (size_t)k, P.mount + 6 * (size_t)k + 3, out);
#pragma unroll
    for (int i = 0; i < CAM_CACHE_STRIDE; i++) P.mount_cache[(size_t)CAM_CACHE_STRIDE * k + i] = out[i];
}
// With a constant sensor offset (Roff, toff) and / or a mounting (boresight b, lever arm l) the record holds the SENSOR pose
//   Rs = R Roff Rb,  ts = t + R (toff + Roff l)        (sensor2world = (pose * offset) * (b, l),
//   costfunctors/posedecoratorfunctors.h:208,628-700; sbamodules/pinholecameraprojectormodule.cpp:147-214)
// so that Pbar = Rs^T (X - ts) and A = Pi Rs^T = d res / d X = -d res / d t hold as without them; Jr stays the BODY one and
// cam_Q = (R Jr)^T | ts - t, because d res / d rAxis = A [X - t]x (R Jr) with the body pose: the readers add ts - t to Rs Pbar.
__global__ void k_pose_cache(DevProblem P) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= P.Nc) return;
    double out[CAM_CACHE_STRIDE];
    pose_cache(P.cam_r + 3 * c, P.cam_t + 3 * c, out);
    out[21] = P.intr[10 * P.cam_model[c]];   // f of the camera's model: d_z of the plane cost (factored camera pass)
    // Q = (R Jr)^T: C = A [X - t]x Q^T (explicit reduced system, sba_kernels_explicit.cu).  R[i][k] = Rt[3k+i]
    double* Q = P.cam_Q + 12 * (size_t)c;
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int a = 0; a < 3; a++) Q[3 * r + a] = out[a] * out[12 + r] + out[3 + a] * out[15 + r] + out[6 + a] * out[18 + r];
    double dt[3] = {0.0, 0.0, 0.0};
    const int mt = P.cam_mount ? P.cam_mount[c] : -1;
    if (P.cam_off || mt >= 0) {
        double M[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, cb[3] = {0, 0, 0};   // M = sensor <- body rotation (Rb^T Roff^T), cb = sensor origin in the body frame
        if (P.cam_off) {
            const double* off = P.cam_off + 12 * (size_t)c;
#pragma unroll
            for (int i = 0; i < 9; i++) M[i] = off[i];
            cb[0] = off[9]; cb[1] = off[10]; cb[2] = off[11];
        }
        if (mt >= 0) {
            const double* cl = P.mount_cache + (size_t)CAM_CACHE_STRIDE * mt;   // Rb^T, l
            // cb += Roff l = M_off^T l ; M = Rb^T M_off
            double c2[3], M2[9];
#pragma unroll
            for (int i = 0; i < 3; i++) c2[i] = cb[i] + M[i] * cl[9] + M[3 + i] * cl[10] + M[6 + i] * cl[11];
#pragma unroll
            for (int i = 0; i < 3; i++)
#pragma unroll
                for (int j = 0; j < 3; j++) M2[3 * i + j] = cl[3 * i] * M[j] + cl[3 * i + 1] * M[3 + j] + cl[3 * i + 2] * M[6 + j];
#pragma unroll
            for (int i = 0; i < 9; i++) M[i] = M2[i];
            cb[0] = c2[0]; cb[1] = c2[1]; cb[2] = c2[2];
        }
        // dt = R cb (R = Rt^T), Rs^T = M Rt
#pragma unroll
        for (int i = 0; i < 3; i++) dt[i] = out[i] * cb[0] + out[3 + i] * cb[1] + out[6 + i] * cb[2];
        double Rs[9];
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
            for (int j = 0; j < 3; j++) Rs[3 * i + j] = M[3 * i] * out[j] + M[3 * i + 1] * out[3 + j] + M[3 * i + 2] * out[6 + j];
#pragma unroll
        for (int i = 0; i < 9; i++) out[i] = Rs[i];
        out[9] += dt[0]; out[10] += dt[1]; out[11] += dt[2];
    }
    Q[9] = dt[0]; Q[10] = dt[1]; Q[11] = dt[2];
#pragma unroll
    for (int i = 0; i < CAM_CACHE_STRIDE; i++) P.cam_cache[(size_t)CAM_CACHE_STRIDE * c + i] = out[i];
}
__global__ void k_intr_recip(DevProblem P) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= P.Nm) return;
    const double f = P.intr[10 * (size_t)m], t1 = P.intr[10 * (size_t)m + 6];
    P.intr_r[2 * (size_t)m] = 1.0 / f;
    P.intr_r[2 * (size_t)m + 1] = 1.0 / (f + t1);
}
void launch_pose_cache(const DevProblem& P, cudaStream_t st) {
    if (P.Nm && P.intr_r) k_intr_recip<<<cdiv(P.Nm, 128), 128, 0, st>>>(P);
    if (P.Nmt) k_mount_cache<<<cdiv(P.Nmt, 128), 128, 0, st>>>(P);
    k_pose_cache<<<cdiv(P.Nc, 128), 128, 0, st>>>(P);
}

// ---------------------------------------------------------------------------
// K1 residual + Jacobian, camera order.  One warp per tile (<= 256 consecutive
// slots of ONE camera): camera cache record and intrinsics staged in shared
// memory once per tile, (u,v) fetched with one 128-bit load per observation,
// the point with two 128-bit loads (one 32 B sector), outputs written as
// coalesced SoA planes.  JAC=false is the cost-only variant K1c (trial steps).
// Algorithmic bytes per observation: in 16 (uv) + 4 (pt); out 16 (res) + 48 (A)
// + 24 (Pbar) = 108 (+4 B/obs index) = 112 B/obs.  SURVEY 8(d) counts 184 B/obs for the fully
// materialised Jacobian; here J_t = -J_X is not stored and J_r = A [R Pbar]x Q^T is rebuilt by its
// readers from A and Pbar (rebuild_C), so 72 B/obs of it never exist.
// ---------------------------------------------------------------------------
// one observation of the tile: residual, A, Pbar (and the d_z scale) into the camera-order planes
template <bool JAC, bool INFO>
__device__ __forceinline__ void k1_slot(const DevProblem& P, const double* cc, const double* intr, int spherical, size_t Ns, int s,
                                        double2 uv, const double (&X)[3], double& acc, int& bad) {
    double d[3];
    dir_from_uv_r(uv.x, uv.y, intr, intr[10], intr[11], d);
    double W[4] = {1.0, 0.0, 0.0, 1.0};
    if (INFO) {
#pragma unroll
        for (int i = 0; i < 4; i++) W[i] = __ldcs(P.obs_info + i * Ns + s);
    }
    double r[2] = {0.0, 0.0}, Jc[12], Jp[6], aux[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
    const bool ok = obs_eval<JAC>(cc, X, d, INFO ? W : nullptr, spherical, r, Jc, Jp, JAC ? aux : nullptr);
    if (!ok) {
        bad = 1;
        aux[0] = aux[1] = aux[2] = 0.0;   // Pbar = 0 marks a failed row for the factored pass
        r[0] = r[1] = 0.0;
        if (JAC) {
#pragma unroll
            for (int i = 0; i < 12; i++) Jc[i] = 0.0;
#pragma unroll
            for (int i = 0; i < 6; i++) Jp[i] = 0.0;
        }
    }
    acc += r[0] * r[0] + r[1] * r[1];
    if (JAC) {
        stcs(P.res + (s), r[0]);
        stcs(P.res + (Ns + s), r[1]);
#pragma unroll
        for (int j = 0; j < 3; j++) {
            stcs(P.Ac + ((size_t)j * Ns + s), Jp[j]);
            stcs(P.Ac + ((size_t)(3 + j) * Ns + s), Jp[3 + j]);
        }
        stcs(P.Pc + (s), aux[0]); stcs(P.Pc + (Ns + s), aux[1]); stcs(P.Pc + (2 * Ns + s), aux[2]);
        if (P.Kc) { stcs(P.Kc + (s), aux[3]); stcs(P.Kc + (Ns + s), aux[4]); }
    }
}
// PIPE: the loads of the lane's next observation (index two steps ahead, pixel and point one step ahead) are in flight
// during the arithmetic of the current one -- the kernel is bound by the latency of the index -> point-gather chain
// (ncu r2f: FP64 pipe 20 %, DRAM 46 %, 42 % warps active), not by bytes or flops.
template <bool JAC, bool PIPE, int MINB, bool INFO>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32, MINB) k_residual_jacobian(DevProblem P) {
    __shared__ double s_cam[WARPS_PER_BLOCK][CAM_CACHE_STRIDE + SBA_INTR_PAD + 2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * WARPS_PER_BLOCK + warp;
    if (tile >= P.Nt) return;
    const int cam = __ldg(P.tile_cam + tile);
    const int start = __ldg(P.tile_start + tile), len = __ldg(P.tile_len + tile);
    double* cc = s_cam[warp];
    if (lane < CAM_CACHE_STRIDE) cc[lane] = __ldg(P.cam_cache + (size_t)CAM_CACHE_STRIDE * cam + lane);
    {
        const int m = __ldg(P.cam_model + cam);
        if (lane < 10) cc[CAM_CACHE_STRIDE + lane] = __ldg(P.intr + 10 * m + lane);
        else if (lane < 12) cc[CAM_CACHE_STRIDE + lane] = __ldg(P.intr_r + 2 * m + (lane - 10));   // 1 / f, 1 / (f + t1)
    }
    __syncwarp();
    const double* intr = cc + CAM_CACHE_STRIDE;
    const int spherical = P.cam_kind[cam];
    const size_t Ns = (size_t)P.Ns;
    const int end = start + len;
    double acc = 0.0;
    int bad = 0;
    if (!PIPE) {
        for (int s = start + lane; s < end; s += 32) {
            const int pt = __ldcs(P.obs_pt + s);
            if (pt < 0) continue;   // padding slot: planes stay zero
            const double2 uv = ldcs2(P.obs_uv + 2 * (size_t)s);
            const d4 Xp = ld256(P.pts + 4 * (size_t)pt);   // one 32-byte sector, one request
            const double X[3] = {Xp.x, Xp.y, Xp.z};
            k1_slot<JAC, INFO>(P, cc, intr, spherical, Ns, s, uv, X, acc, bad);
        }
    } else {
        int s = start + lane;
        int pt1 = s < end ? __ldcs(P.obs_pt + s) : -1;
        int pt2 = s + 32 < end ? __ldcs(P.obs_pt + s + 32) : -1;
        double2 uv1 = s < end ? ldcs2(P.obs_uv + 2 * (size_t)s) : make_double2(0.0, 0.0);
        d4 X1 = ld256(P.pts + 4 * (size_t)max(pt1, 0));
        for (; s < end; s += 32) {
            const int pt = pt1;
            const double2 uv = uv1;
            const double X[3] = {X1.x, X1.y, X1.z};
            pt1 = pt2;
            pt2 = s + 64 < end ? __ldcs(P.obs_pt + s + 64) : -1;
            if (s + 32 < end) {
                uv1 = ldcs2(P.obs_uv + 2 * (size_t)(s + 32));
                X1 = ld256(P.pts + 4 * (size_t)max(pt1, 0));
            }
            if (pt < 0) continue;
            k1_slot<JAC, INFO>(P, cc, intr, spherical, Ns, s, uv, X, acc, bad);
        }
    }
    acc = warp_sum(acc);
    bad = __any_sync(FULL, bad);
    if (lane == 0) {
        P.tile_cost[tile] = acc;
        if (bad) P.flags[FL_CHEIRALITY] = 1;
    }
}
void launch_residual_jacobian(const DevProblem& P, bool jac, cudaStream_t st) {
    if (P.Nt == 0) return;
    const int nb = cdiv(P.Nt, WARPS_PER_BLOCK);
    const char* e = std::getenv("SBA_K1_PIPE");   // measurement knobs
    const bool pipe = e ? std::atoi(e) != 0 : true;
    const char* e2 = std::getenv("SBA_K1_MINB");
    // measured on the Venice shape (scripts/ab_k1.py): K1 0.1268 ms with the prefetch at 3 blocks / SM (80 registers, no spills)
    // vs 0.1367 without it (64 registers, 4 blocks) and 0.162 with it at 64 registers (spills); K1c 0.049 vs 0.066
    const int minb = e2 ? std::atoi(e2) : (pipe ? (jac ? 3 : 4) : 4);
    const int nt = WARPS_PER_BLOCK * 32;
#define K1_LAUNCH(J, PP, MB)                                                        \
    do {                                                                            \
        if (P.obs_info) k_residual_jacobian<J, PP, MB, true><<<nb, nt, 0, st>>>(P); \
        else k_residual_jacobian<J, PP, MB, false><<<nb, nt, 0, st>>>(P);           \
    } while (0)
    if (jac) {
        if (pipe) { if (minb == 4) K1_LAUNCH(true, true, 4); else if (minb == 2) K1_LAUNCH(true, true, 2); else K1_LAUNCH(true, true, 3); }
        else { if (minb == 3) K1_LAUNCH(true, false, 3); else K1_LAUNCH(true, false, 4); }
    } else {
        if (pipe) { if (minb == 4) K1_LAUNCH(false, true, 4); else if (minb == 2) K1_LAUNCH(false, true, 2); else K1_LAUNCH(false, true, 3); }
        else { if (minb == 3) K1_LAUNCH(false, false, 3); else K1_LAUNCH(false, false, 4); }
    }
#undef K1_LAUNCH
}

}  // namespace sba
