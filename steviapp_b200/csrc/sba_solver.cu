// sba_solver.cu -- ingest + layout building + LM driver (host side, C++).
//
// Mirrors ModularSBASolver's life cycle (libs/sparsesolver/modularsbasolver.cpp):
//   init()      :679-802  -> set_* / add_* / finalize()
//   opt_step()  :884-951  -> solve()   (the whole LM, as the single ceres::Solve call does)
//   writeResults:1005-1090 -> download_parameters() (caller arrays updated in place)
#include "sba_solver.h"
#include "sba_math.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace sba {

#define CU(call)                                                     \
    do {                                                             \
        cudaError_t e__ = (call);                                    \
        if (e__ != cudaSuccess) return cuda_fail(e__, #call);        \
    } while (0)

Solver::Solver(const sba_options& o) : opt(o) {}

Solver::~Solver() {
    if (st) cudaStreamSynchronize(st);
    release_peer_exchange();
    release_peer_tiles();
    unpin_all();
    comm.destroy();
    timer.destroy();
    arena.release();
    if (h_scalars) cudaFreeHost(h_scalars);
    if (h_flags) cudaFreeHost(h_flags);
    if (h_stage) cudaFreeHost(h_stage);
    if (st) cudaStreamDestroy(st);
}

int Solver::fail(int code, const std::string& msg) {
    err = msg;
    if (opt.verbose) fprintf(stderr, "[sba_b200] error %d: %s\n", code, msg.c_str());
    return code;
}
int Solver::cuda_fail(cudaError_t e, const char* what) {
    return fail(SBA_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

int Solver::open_device() {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(SBA_ERR_NO_DEVICE, std::string("no CUDA device (this engine has no CPU fallback): ") +
                                           (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0"));
    if (opt.device >= 0) CU(cudaSetDevice(opt.device));
    CU(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    CU(cudaMallocHost((void**)&h_scalars, SC_N * sizeof(double)));
    CU(cudaMallocHost((void**)&h_flags, FL_N * sizeof(int32_t)));
    return SBA_OK;
}

// ---- timers ---------------------------------------------------------------
cudaEvent_t PhaseTimer::get() {
    if (used == pool.size()) {
        cudaEvent_t e;
        cudaEventCreate(&e);
        pool.push_back(e);
    }
    return pool[used++];
}
void PhaseTimer::begin(int phase, cudaStream_t st) {
    Span s{get(), nullptr, phase};
    cudaEventRecord(s.a, st);
    spans.push_back(s);
}
void PhaseTimer::end(cudaStream_t st) {
    Span& s = spans.back();
    s.b = get();
    cudaEventRecord(s.b, st);
}
void PhaseTimer::collect(double* ms) {
    for (int i = 0; i < PH_N; i++) ms[i] = 0.0;
    for (Span& s : spans) {
        if (!s.b) continue;
        cudaEventSynchronize(s.b);
        float t = 0.f;
        cudaEventElapsedTime(&t, s.a, s.b);
        ms[s.phase] += t;
    }
    reset();
}
void PhaseTimer::destroy() {
    for (cudaEvent_t e : pool) cudaEventDestroy(e);
    pool.clear();
    spans.clear();
    used = 0;
}

// ---- ingest ---------------------------------------------------------------
int Solver::set_intrinsics(int32_t n, const double* intr10) {
    if (n <= 0 || !intr10) return fail(SBA_ERR_INVALID, "set_intrinsics: need at least one camera model");
    if (finalized && n != Nm) return fail(SBA_ERR_INVALID, "set_intrinsics: model count changed after finalize");
    Nm = n;
    intr.assign(intr10, intr10 + (size_t)SBA_INTR_STRIDE * n);
    if (finalized) CU(cudaMemcpyAsync(P.intr, intr.data(), intr.size() * sizeof(double), cudaMemcpyHostToDevice, st));
    return SBA_OK;
}

int Solver::set_poses(double* raxis, double* t, int32_t n, const int32_t* model, const uint8_t* fixed,
                      const uint8_t* kind) {
    if (n <= 0 || !raxis || !t) return fail(SBA_ERR_INVALID, "set_poses: null arrays or n <= 0");
    if (finalized && n != Nc) return fail(SBA_ERR_INVALID, "set_poses: pose count changed after finalize");
    if (raxis != h_raxis || t != h_t) unpin_all();
    h_raxis = raxis; h_t = t; Nc = n;
    if (!finalized) {
        cam_model.assign(n, 0); cam_fixed.assign(n, 0); cam_kind.assign(n, 0);
        if (model) cam_model.assign(model, model + n);
        if (fixed) cam_fixed.assign(fixed, fixed + n);
        if (kind) cam_kind.assign(kind, kind + n);
    }
    return SBA_OK;
}

int Solver::set_points(double* xyz, int32_t n, const uint8_t* fixed) {
    if (n < 0 || (n > 0 && !xyz)) return fail(SBA_ERR_INVALID, "set_points: null array");
    if (finalized && n != Np) return fail(SBA_ERR_INVALID, "set_points: point count changed after finalize");
    if (xyz != h_xyz) unpin_all();
    h_xyz = xyz; Np = n;
    if (!finalized) {
        pt_fixed.assign(n, 0);
        if (fixed) pt_fixed.assign(fixed, fixed + n);
    }
    return SBA_OK;
}

int Solver::add_observations(int32_t n, const int32_t* cam, const int32_t* pt, const double* uv, const double* info) {
    if (finalized) return fail(SBA_ERR_INVALID, "add_observations after finalize");
    if (n < 0 || (n > 0 && (!cam || !pt || !uv))) return fail(SBA_ERR_INVALID, "add_observations: null arrays");
    if (n == 0) return SBA_OK;
    const size_t old = obs_cam.size();
    if ((size_t)old + (size_t)n > 0x7fffffe0u) return fail(SBA_ERR_INVALID, "add_observations: more than 2^31 observations on one rank");
    obs_cam.insert(obs_cam.end(), cam, cam + n);
    obs_pt.insert(obs_pt.end(), pt, pt + n);
    obs_uv.insert(obs_uv.end(), uv, uv + 2 * (size_t)n);
    if (info) {
        if (!has_info) {   // earlier batches had identity stiffness
            obs_info.assign(4 * old, 0.0);
            for (size_t i = 0; i < old; i++) { obs_info[4 * i] = 1.0; obs_info[4 * i + 3] = 1.0; }
            has_info = true;
        }
        obs_info.insert(obs_info.end(), info, info + 4 * (size_t)n);
    } else if (has_info) {
        for (int32_t i = 0; i < n; i++) { const double I[4] = {1, 0, 0, 1}; obs_info.insert(obs_info.end(), I, I + 4); }
    }
    return SBA_OK;
}

int Solver::add_priors(int32_t n, const int32_t* kind, const int32_t* index, const double* A9, const double* b3) {
    if (finalized) return fail(SBA_ERR_INVALID, "add_priors after finalize");
    if (n < 0 || (n > 0 && (!kind || !index || !A9 || !b3))) return fail(SBA_ERR_INVALID, "add_priors: null arrays");
    for (int32_t i = 0; i < n; i++)
        if (kind[i] < 0 || kind[i] > 2) return fail(SBA_ERR_INVALID, "add_priors: kind must be 0, 1 or 2");
    prior_kind.insert(prior_kind.end(), kind, kind + n);
    prior_index.insert(prior_index.end(), index, index + n);
    prior_A.insert(prior_A.end(), A9, A9 + 9 * (size_t)n);
    prior_b.insert(prior_b.end(), b3, b3 + 3 * (size_t)n);
    return SBA_OK;
}

}  // namespace sba

#include "sba_solver_shared.inc"
#include "sba_solver_layout.inc"
#include "sba_solver_explicit.inc"
#include "sba_solver_lm.inc"
#include "sba_solver_hooks.inc"
#include "sba_solver_pairs.inc"
