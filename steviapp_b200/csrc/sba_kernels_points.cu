// sba_kernels_points.cu -- the point-order ("q slot") kernels: K1p linearisation fused with the K2a point
// blocks, and pass 1 of the implicit Schur operator / back-substitution.
//
// One thread per q slot; the qA planes are written / read as coalesced 256-byte runs per warp.  The sum
// over the observations of a point is a segmented warp reduction: points with k <= 32 observations never
// straddle a 32-slot window (layout builder), so a fixed-shape shuffle-down tree keyed on the point index
// gives every point's sum in its first lane -- deterministic, no atomics.  Points with k > 32 ("heavy")
// own whole windows; their per-window partials go to heavy_part and a second tiny kernel folds them in
// window order.
#include <algorithm>

#include "sba_kernel_util.cuh"

namespace sba {

// Segmented warp reduction keyed on the point index: after the call the first lane of every run of equal keys holds
// the run's sum (fixed shape -> deterministic).  One shuffle + one ballot find the run boundaries; `rem` = slots from
// this lane to the end of its run, so "lane + o belongs to my run" is just o < rem (no key shuffle per step).
__device__ __forceinline__ int seg_setup(int key, int lane, bool* head) {
    const int prev = __shfl_up_sync(FULL, key, 1);
    *head = lane == 0 || prev != key;
    const unsigned heads = __ballot_sync(FULL, *head);
    const unsigned above = lane == 31 ? 0u : heads >> (lane + 1);
    return above ? __ffs(above) : 32 - lane;
}
template <int N>
__device__ __forceinline__ void seg_reduce(double (&v)[N], int rem) {
    const int longest = __reduce_max_sync(FULL, rem);   // warp-uniform: steps that no run of this warp needs are skipped
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        if (o >= longest) break;
        const bool take = o < rem;
#pragma unroll
        for (int i = 0; i < N; i++) {
            const double t = __shfl_down_sync(FULL, v[i], o);
            if (take) v[i] += t;
        }
    }
}

// ---------------------------------------------------------------------------
// K1p + K2a: linearisation in point order fused with the point blocks.  Same expressions as K1
// (sba_math.cuh), so both orders hold bitwise identical A and res.  The thread keeps its A and res in
// registers, writes the six qA planes (the only point-order copy the operator needs) and feeds
//   Hpp = sum A^T A (packed sym),  gp = sum A^T res
// straight into the segmented warp reduction -- res and C are never written in point order.
// In: 16 (uv) + 8 (pt, cam idx) per slot, camera record gathered from L2 (96 B + intrinsics), point from
// L1/L2; out: 48 B per slot as 6 coalesced planes + 96 B per point.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void store_point_blocks(const DevProblem& P, int pt, const double* v) {
    const bool fixed = P.pt_fixed[pt];
    double* Hp = P.Hpp + 8 * (size_t)pt;
    double* gp = P.gp + 4 * (size_t)pt;
    if (fixed) {
        st2(Hp, 0.0, 0.0); st2(Hp + 2, 0.0, 0.0); st2(Hp + 4, 0.0, 0.0); st2(Hp + 6, 0.0, 0.0);
        st2(gp, 0.0, 0.0); st2(gp + 2, 0.0, 0.0);
    } else {
        st2(Hp, v[0], v[1]); st2(Hp + 2, v[2], v[3]); st2(Hp + 4, v[4], v[5]); st2(Hp + 6, 0.0, 0.0);
        st2(gp, v[6], v[7]); st2(gp + 2, v[8], 0.0);
    }
}
// one q slot; `crec(cam, ccr)` fetches Rt (9) and t (3) of the camera (global gather or shared-memory table)
template <class Fetch>
__device__ __forceinline__ void linearize_slot(const DevProblem& P, int q, int lane, Fetch crec) {
    const int pt = __ldcs(P.q_pt + q);
    double v[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    if (pt >= 0) {   // padding: planes stay zero
        const int cam = __ldcs(P.q_cam + q);
        const d4 Xp = ld256(P.pts + 4 * (size_t)pt);
        const double X[3] = {Xp.x, Xp.y, Xp.z};
        // only Rt and t are needed (C is not kept in point order: its expressions, hence Jr, are dead code here)
        double ccr[22];
        crec(cam, ccr);
#pragma unroll
        for (int i = 12; i < 22; i++) ccr[i] = 0.0;
        const int mdl = P.Nm == 1 ? 0 : __ldg(P.cam_model + cam);
        const double* in = P.intr + 10 * mdl;
        double intr[10];
#pragma unroll
        for (int i = 0; i < 5; i++) { const double2 t = ldg2(in + 2 * i); intr[2 * i] = t.x; intr[2 * i + 1] = t.y; }
        const double2 rcp = ldg2(P.intr_r + 2 * mdl);   // 1 / f, 1 / (f + t1): the same table K1 reads (bitwise equal A and res)
        const double2 uv = ldcs2(P.q_uv + 2 * (size_t)q);
        double d[3];
        dir_from_uv_r(uv.x, uv.y, intr, rcp.x, rcp.y, d);
        double W[4];
        const double* Wp = nullptr;
        if (P.q_info) {
            const double2 w01 = ldcs2(P.q_info + 4 * (size_t)q), w23 = ldcs2(P.q_info + 4 * (size_t)q + 2);
            W[0] = w01.x; W[1] = w01.y; W[2] = w23.x; W[3] = w23.y;
            Wp = W;
        }
        double r[2] = {0.0, 0.0}, Jc[12], A[6];
        const bool ok = obs_eval<true>(ccr, X, d, Wp, P.cam_kind[cam], r, Jc, A);
        if (!ok) {
            r[0] = r[1] = 0.0;
#pragma unroll
            for (int i = 0; i < 6; i++) A[i] = 0.0;
        }
        const size_t Nq = (size_t)P.Nq;
#pragma unroll
        for (int j = 0; j < 6; j++) stcs(P.qA + ((size_t)j * Nq + q), A[j]);
        v[0] = A[0] * A[0] + A[3] * A[3]; v[1] = A[0] * A[1] + A[3] * A[4]; v[2] = A[0] * A[2] + A[3] * A[5];
        v[3] = A[1] * A[1] + A[4] * A[4]; v[4] = A[1] * A[2] + A[4] * A[5]; v[5] = A[2] * A[2] + A[5] * A[5];
        v[6] = A[0] * r[0] + A[3] * r[1]; v[7] = A[1] * r[0] + A[4] * r[1]; v[8] = A[2] * r[0] + A[5] * r[1];
    }
    bool head;
    const int rem = seg_setup(pt, lane, &head);
    seg_reduce<9>(v, rem);
    if (pt >= 0 && head) {
        const int h = __ldg(P.pt_heavy + pt);
        if (h < 0) {
            store_point_blocks(P, pt, v);
        } else {
            double* o = P.heavy_part + (size_t)HP * (__ldg(P.heavy_w0 + h) + ((q - __ldg(P.pt_q0 + pt)) >> 5));
#pragma unroll
            for (int i = 0; i < 9; i++) o[i] = v[i];
        }
    }
}
// variant G: three 256-bit gathers of the camera record per slot (any camera count)
__global__ void __launch_bounds__(256) k_linearize_q(DevProblem P) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;   // Nq is a multiple of 32: whole warps in range
    if (q >= P.Nq) return;
    linearize_slot(P, q, threadIdx.x & 31, [&](int cam, double* ccr) {
        const double* cc = P.cam_cache + (size_t)CAM_CACHE_STRIDE * cam;
        const d4 c0 = ld256(cc), c1 = ld256(cc + 4), c2 = ld256(cc + 8);
        ccr[0] = c0.x; ccr[1] = c0.y; ccr[2] = c0.z; ccr[3] = c0.w; ccr[4] = c1.x; ccr[5] = c1.y; ccr[6] = c1.z; ccr[7] = c1.w;
        ccr[8] = c2.x; ccr[9] = c2.y; ccr[10] = c2.z; ccr[11] = c2.w;
    });
}
// variant S: the whole table {Rt, t} (96 B per camera) staged in shared memory by persistent blocks (one per SM): the
// per-slot gather becomes six LDS.128 instead of 96 L1 wavefronts per warp -- the kernel is bound by the L1 / MIO wavefront
// rate of its gathers, not by HBM bytes.  Used when 96 * Nc fits the SM's shared memory (Venice: 171 KB).
__global__ void __launch_bounds__(1024, 1) k_linearize_q_s(DevProblem P) {
    extern __shared__ double2 s_cr[];   // 6 double2 per camera
    for (int i = threadIdx.x; i < 6 * P.Nc; i += blockDim.x) {
        const int c = i / 6, j = i - 6 * c;
        s_cr[i] = *reinterpret_cast<const double2*>(P.cam_cache + (size_t)CAM_CACHE_STRIDE * c + 2 * j);
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < P.Nq; q += gridDim.x * blockDim.x)   // Nq % 32 == 0: whole warps
        linearize_slot(P, q, lane, [&](int cam, double* ccr) {
#pragma unroll
            for (int j = 0; j < 6; j++) { const double2 t = s_cr[6 * cam + j]; ccr[2 * j] = t.x; ccr[2 * j + 1] = t.y; }
        });
}
__global__ void k_point_blocks_heavy(DevProblem P) {
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= P.Nh) return;
    double v[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int w = P.heavy_w0[h]; w < P.heavy_w0[h + 1]; w++)
#pragma unroll
        for (int i = 0; i < 9; i++) v[i] += P.heavy_part[(size_t)HP * w + i];
    store_point_blocks(P, P.heavy_pt[h], v);
}
void launch_linearize_points(const DevProblem& P, cudaStream_t st) {
    if (P.Nq == 0) return;
    static int sms = 0, smem_max = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        cudaFuncSetAttribute(k_linearize_q_s, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
    }
    const size_t table = 96 * (size_t)P.Nc;
    if (P.pp_variant != 1 && table <= (size_t)smem_max && P.Nq >= 256 * 1024)
        k_linearize_q_s<<<std::min(sms, cdiv(P.Nq, 1024)), 1024, table, st>>>(P);
    else
        k_linearize_q<<<cdiv(P.Nq, 256), 256, 0, st>>>(P);
    if (P.Nh) k_point_blocks_heavy<<<cdiv(P.Nh, 128), 128, 0, st>>>(P);
}

// ---------------------------------------------------------------------------
// operator pass 1 / back-substitution.  x must be zero at fixed cameras.
//   MODE 0: y_p = Hinv * sum_o A^T (C x_r - A x_t)
//   MODE 2: dp  = -Hinv * (gp + sum_o A^T (C x_r - A x_t))   (block_solver.hpp:455-477 analogue)
// C is not read: C = A [X - t]x (R Jr) exactly (Pbar = R^T (X - t), [R^T d]x = R^T [d]x R), hence
//   C x_r - A x_t = A ( (X - t) x a_c - x_t ),   a_c = R Jr x_r  (per camera, k_cam_vec),
// which holds with any stiffness (A = W Pi R^T carries it).  56 B/slot in (A planes + pt, cam idx); per-camera
// operand (a_c, g_c: 64 B record = two 256-bit loads, 114 KB for 1778 cameras: L1/L2) and the point (one 256-bit
// load, shared by the point's lanes) gathered; 64 B/point in (Hinv [+ gp]) and 32 B/point out, by the head lanes.
// ---------------------------------------------------------------------------
__global__ void k_cam_vec(DevProblem P, const double* __restrict__ x) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= P.Nc) return;
    double x6[6];
#pragma unroll
    for (int a = 0; a < 6; a++) x6[a] = x[6 * (size_t)c + a];
    cam_vec_store(P, c, x6);
}
void launch_cam_vec(const DevProblem& P, const double* x, cudaStream_t st) {
    k_cam_vec<<<cdiv(P.Nc, 128), 128, 0, st>>>(P, x);
}
template <int MODE>
__device__ __forceinline__ void finish_point_pass(const DevProblem& P, int pt, double u0, double u1, double u2) {
    if (MODE == 2) {
        const d4 g = ld256(P.gp + 4 * (size_t)pt);
        u0 += g.x; u1 += g.y; u2 += g.z;
    }
    const double* Hp = P.Hinv + 8 * (size_t)pt;
    const d4 h0 = ld256(Hp), h1 = ld256(Hp + 4);   // (h00 h01 h02 h11) (h12 h22 - -)
    double y0 = h0.x * u0 + h0.y * u1 + h0.z * u2;
    double y1 = h0.y * u0 + h0.w * u1 + h1.x * u2;
    double y2 = h0.z * u0 + h1.x * u1 + h1.y * u2;
    double* o = (MODE == 2 ? P.dp : P.y_p) + 4 * (size_t)pt;
    if (MODE == 2) { y0 = -y0; y1 = -y1; y2 = -y2; }
    st256(o, y0, y1, y2, 0.0);
}
// one q slot: `cvec(cam, a, g)` fetches the camera operand (global gather or shared-memory table)
template <int MODE, class Fetch>
__device__ __forceinline__ void point_pass_slot(const DevProblem& P, int q, int lane, Fetch cvec) {
    const int pt = __ldcs(P.q_pt + q);
    const size_t Nq = (size_t)P.Nq;
    double u[3] = {0.0, 0.0, 0.0};
    if (pt >= 0) {
        const int cam = __ldcs(P.q_cam + q);
        const d4 X = ld256(P.pts + 4 * (size_t)pt);
        double a[3], g[3];
        cvec(cam, a, g);
        double A[6];
#pragma unroll
        for (int i = 0; i < 6; i++) A[i] = __ldcs(P.qA + (size_t)i * Nq + q);
        const double d0 = X.x - P.center[0], d1 = X.y - P.center[1], d2 = X.z - P.center[2];
        // w = (X - c0) x a_c - g_c
        const double w0 = d1 * a[2] - d2 * a[1] - g[0];
        const double w1 = d2 * a[0] - d0 * a[2] - g[1];
        const double w2 = d0 * a[1] - d1 * a[0] - g[2];
        const double m0 = A[0] * w0 + A[1] * w1 + A[2] * w2;
        const double m1 = A[3] * w0 + A[4] * w1 + A[5] * w2;
        u[0] = A[0] * m0 + A[3] * m1;
        u[1] = A[1] * m0 + A[4] * m1;
        u[2] = A[2] * m0 + A[5] * m1;
    }
    bool head;
    const int rem = seg_setup(pt, lane, &head);
    seg_reduce<3>(u, rem);
    if (pt >= 0 && head) {
        const int h = __ldg(P.pt_heavy + pt);
        if (h < 0) {
            finish_point_pass<MODE>(P, pt, u[0], u[1], u[2]);
        } else {
            double* o = P.heavy_part + (size_t)HP * (__ldg(P.heavy_w0 + h) + ((q - __ldg(P.pt_q0 + pt)) >> 5));
            o[0] = u[0]; o[1] = u[1]; o[2] = u[2];
        }
    }
}
// variant G: operand gathered from global memory (two 256-bit loads per slot; any camera count)
template <int MODE>
__global__ void __launch_bounds__(256) k_point_pass_q(DevProblem P) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= P.Nq) return;
    point_pass_slot<MODE>(P, q, threadIdx.x & 31, [&](int cam, double* a, double* g) {
        const double* cv = P.cam_vec + (size_t)CAM_VEC_STRIDE * cam;
        const d4 c0 = ld256(cv), c1 = ld256(cv + 4);   // (a0 a1 a2 g0) (g1 g2 - -)
        a[0] = c0.x; a[1] = c0.y; a[2] = c0.z; g[0] = c0.w; g[1] = c1.x; g[2] = c1.y;
    });
}
// variant S: the whole operand table (48 B per camera) staged in shared memory by persistent blocks -- the gather
// then costs 3 LDS.128 per slot instead of 64 L1 wavefronts per warp, which is what bounds variant G (the pass is
// limited by the L1/MIO wavefront rate, not by HBM bytes).  Used when 48 * Nc fits the SM's shared memory.
template <int MODE>
__global__ void __launch_bounds__(1024, 2) k_point_pass_s(DevProblem P) {
    extern __shared__ double2 s_cv[];   // 3 double2 per camera: (a0 a1) (a2 g0) (g1 g2)
    for (int i = threadIdx.x; i < 3 * P.Nc; i += blockDim.x) {
        const int c = i / 3, j = i - 3 * c;
        s_cv[i] = *reinterpret_cast<const double2*>(P.cam_vec + (size_t)CAM_VEC_STRIDE * c + 2 * j);
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < P.Nq; q += gridDim.x * blockDim.x)   // Nq % 32 == 0: whole warps
        point_pass_slot<MODE>(P, q, lane, [&](int cam, double* a, double* g) {
            const double2 c0 = s_cv[3 * cam], c1 = s_cv[3 * cam + 1], c2 = s_cv[3 * cam + 2];
            a[0] = c0.x; a[1] = c0.y; a[2] = c1.x; g[0] = c1.y; g[1] = c2.x; g[2] = c2.y;
        });
}
template <int MODE>
__global__ void k_point_pass_heavy(DevProblem P) {
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= P.Nh) return;
    double u0 = 0.0, u1 = 0.0, u2 = 0.0;
    for (int w = P.heavy_w0[h]; w < P.heavy_w0[h + 1]; w++) {
        u0 += P.heavy_part[(size_t)HP * w]; u1 += P.heavy_part[(size_t)HP * w + 1]; u2 += P.heavy_part[(size_t)HP * w + 2];
    }
    finish_point_pass<MODE>(P, P.heavy_pt[h], u0, u1, u2);
}
// MODE 1: y_p = Hinv * gp, thread per point (no observation data)
__global__ void __launch_bounds__(256) k_point_hinv_g(DevProblem P) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P.Np) return;
    const double* g = P.gp + 4 * (size_t)p;
    const double2 g01 = ldg2(g), g23 = ldg2(g + 2);
    const double* Hp = P.Hinv + 8 * (size_t)p;
    const double2 h01 = ldg2(Hp), h23 = ldg2(Hp + 2), h45 = ldg2(Hp + 4);
    double* o = P.y_p + 4 * (size_t)p;
    st2(o, h01.x * g01.x + h01.y * g01.y + h23.x * g23.x, h01.y * g01.x + h23.y * g01.y + h45.x * g23.x);
    st2(o + 2, h23.x * g01.x + h45.x * g01.y + h45.y * g23.x, 0.0);
}
// back-substitution of the free points that carry priors but no observation (no q slot): dp = -Hinv gp
__global__ void k_point_lonely(DevProblem P) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= P.Nlonely) return;
    finish_point_pass<2>(P, P.lonely_pt[k], 0.0, 0.0, 0.0);
}
void launch_point_pass(const DevProblem& P, const double* x, int mode, cudaStream_t st, bool table_ready) {
    if (P.Np == 0) return;
    if (mode == 1) { k_point_hinv_g<<<cdiv(P.Np, 256), 256, 0, st>>>(P); return; }
    if (mode == 2 && P.Nlonely) k_point_lonely<<<cdiv(P.Nlonely, 128), 128, 0, st>>>(P);
    if (!table_ready) launch_cam_vec(P, x, st);
    // shared-memory table variant when the table fits (persistent blocks, 1 or 2 per SM)
    static int sms = 0, smem_max = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        cudaFuncSetAttribute(k_point_pass_s<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
        cudaFuncSetAttribute(k_point_pass_s<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
    }
    const size_t table = 48 * (size_t)P.Nc;
    const bool use_s = P.pp_variant != 1 && table <= (size_t)smem_max && P.Nq >= 64 * 1024;
    if (use_s) {
        const int per_sm = 2 * (table + 1024) <= (size_t)(smem_max + 1024) ? 2 : 1;   // the SM has smem_max + 1 KB in total
        const int nb = std::min(sms * per_sm, cdiv(P.Nq, 1024));
        if (mode == 0) k_point_pass_s<0><<<nb, 1024, table, st>>>(P);
        else k_point_pass_s<2><<<nb, 1024, table, st>>>(P);
    } else {
        const int nb = cdiv(P.Nq, 256);
        if (mode == 0) k_point_pass_q<0><<<nb, 256, 0, st>>>(P);
        else k_point_pass_q<2><<<nb, 256, 0, st>>>(P);
    }
    if (P.Nh) {
        if (mode == 0) k_point_pass_heavy<0><<<cdiv(P.Nh, 128), 128, 0, st>>>(P);
        else k_point_pass_heavy<2><<<cdiv(P.Nh, 128), 128, 0, st>>>(P);
    }
}

}  // namespace sba
