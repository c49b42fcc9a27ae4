// sba_solver.h -- host side of libsba_b200.so: problem ingest, layout building,
// the Levenberg-Marquardt driver and the NCCL plumbing.  C++ only; the C ABI in
// include/sba_b200.h is a thin shell over this class.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "sba_b200.h"
#include "sba_device.h"

namespace sba {

// NCCL loaded lazily with dlopen (the single-GPU path has no NCCL dependency)
struct Comm {
    bool active = false;
    int rank = 0, nranks = 1;
    void* comm = nullptr;
    int init(int rank, int nranks, const uint8_t* id128, std::string& err);
    int allreduce_sum(double* buf, size_t n, cudaStream_t st, std::string& err);
    int allreduce_max(double* buf, size_t n, cudaStream_t st, std::string& err);
    int reduce_scatter_sum(const double* send, double* recv, size_t count, cudaStream_t st, std::string& err);
    int allgather_bytes(const void* send, void* recv, size_t bytes_per_rank, cudaStream_t st, std::string& err);
    void destroy();
    static int unique_id(uint8_t* id128, std::string& err);
};

struct DeviceArena {
    std::vector<void*> ptrs;
    size_t bytes = 0;
    template <class T>
    cudaError_t alloc(T** p, size_t n, bool zero = true) {
        *p = nullptr;
        if (n == 0) n = 1;
        cudaError_t e = cudaMalloc((void**)p, n * sizeof(T));
        if (e != cudaSuccess) return e;
        ptrs.push_back(*p);
        bytes += n * sizeof(T);
        if (zero) e = cudaMemset(*p, 0, n * sizeof(T));
        return e;
    }
    void release() {
        for (void* p : ptrs) cudaFree(p);
        ptrs.clear();
        bytes = 0;
    }
};

enum Phase { PH_LINEARIZE = 0, PH_SCHUR, PH_PCG, PH_UPDATE, PH_H2D, PH_D2H, PH_N };

struct PhaseTimer {
    struct Span { cudaEvent_t a, b; int phase; };
    std::vector<Span> spans;
    std::vector<cudaEvent_t> pool;
    size_t used = 0;
    cudaEvent_t get();
    void begin(int phase, cudaStream_t st);
    void end(cudaStream_t st);
    void collect(double* ms_by_phase);   // synchronises
    void reset() { spans.clear(); used = 0; }
    void destroy();
};

class Solver {
public:
    explicit Solver(const sba_options& o);
    ~Solver();
    int open_device();

    int set_intrinsics(int32_t n, const double* intr10);
    // shared blocks (sba_solver_shared.inc)
    int set_intrinsics_blocks(int32_t n, double* intr10, const uint8_t* fixed_mask);
    int add_intrinsics_priors(int32_t n, const int32_t* model, const int32_t* block, const double* A9, const double* b3);
    int set_sensor_offsets(const double* offset6);
    int set_lever_arms(int32_t n_mounts, double* boresight, double* lever, const uint8_t* fixed, const int32_t* cam_mount);
    int shared_info(int32_t* n_ext, int32_t* n_groups, int32_t* model_block, int32_t* mount_block) const;
    // UV <-> UV factors (sba_solver_pairs.inc)
    int add_uv_pairs(int32_t n, const int32_t* cam1, const int32_t* cam2, const double* uv1, const double* uv2, const double* info1, const double* info2);
    int set_pair_loss(int32_t kind, double a);
    int evaluate_pairs(double* res4, double* jac48);
    int point_covariances(double* cov9);        // ModularSBASolver::std_step (sba_solver_pairs.inc)
    int direct_solve();                         // Cholesky of the explicit S -> pcg_x (opt.linear_solver == 1)
    int set_poses(double* raxis, double* t, int32_t n, const int32_t* model, const uint8_t* fixed, const uint8_t* kind);
    int set_points(double* xyz, int32_t n, const uint8_t* fixed);
    int add_observations(int32_t n, const int32_t* cam, const int32_t* pt, const double* uv, const double* info);
    int add_priors(int32_t n, const int32_t* kind, const int32_t* index, const double* A9, const double* b3);
    int finalize();
    int upload_parameters();
    int download_parameters();
    int evaluate(double* cost, double* residuals, double* jac_cam, double* jac_pt);
    int evaluate_intrinsics_jacobian(double* jac20);
    int evaluate_offset(const double* off12, double* res, double* jac_cam, double* jac_pt, int32_t* n_failed);
    int normal_blocks(double* gc, double* gp, double* Hcc, double* Hpp);
    int schur_matvec(const double* lam_c, const double* lam_p, const double* x, double* y, double* rhs);
    int solve(int copy_params, sba_summary* sum);
    int time_stage(int stage, int reps, double* ms);
    int ensure_dense_tiles();      // sba_solver_lm.inc: panels -> dense tiles when stale (hooks, dense factorisation, tile mat-vec)
    int triangulate(int32_t n_points, const int32_t* view_off, const int32_t* view_cam, const double* view_uv,
                    double* xyz_out, uint8_t* ok_out, double* ms_kernel);
    int comm_init(int rank, int nranks, const uint8_t* id128);

    sba_options opt;
    std::string err;
    sba_iteration_callback cb = nullptr;
    void* cb_user = nullptr;
    DevProblem P{};
    bool finalized = false;
    // the tile / PCG / per-camera vector kernels see the pseudo-cameras of the shared blocks as cameras
    DevProblem Pe() const { DevProblem e = P; e.Nc = P.Nce; return e; }

private:
    int fail(int code, const std::string& msg);
    int cuda_fail(cudaError_t e, const char* what);
    int read_scalars();                      // device scalars/flags -> h_scalars/h_flags (synchronises)
    int linearize(double* cost, bool* eval_ok);            // K0,K1,K1p,K2a,K2b (+ all-reduce) at the current point
    int cost_only(double* cost, bool* eval_ok);            // K0,K1c
    int build_rhs_and_preconditioner(bool* ok);            // damped inverse, reduced rhs, Minv
    int schur_apply(const double* x, double* y);           // y = S x  (x zero at fixed cameras)
    int run_pcg(int* iters);
    int setup_peer_tiles();                  // sba_solver_hooks.inc: IPC mapping of every rank's partial tiles (pull-reduce of the build)
    int setup_peer_exchange();               // sba_solver_hooks.inc: IPC mailboxes of the fused PCG all-reduce (multi-GPU)
    int build_explicit_structure(const std::vector<int32_t>& pt_cnt, const std::vector<int32_t>& pt_q0,
                                 const std::vector<int32_t>& q_cam);   // sba_solver_explicit.inc

    // host copies (caller order)
    double* h_raxis = nullptr; double* h_t = nullptr; double* h_xyz = nullptr;   // caller-owned
    int32_t Nc = 0, Np = 0, Nm = 0;
    std::vector<int32_t> cam_model; std::vector<uint8_t> cam_fixed, cam_kind, pt_fixed;
    std::vector<double> intr;
    std::vector<int32_t> obs_cam, obs_pt; std::vector<double> obs_uv, obs_info; bool has_info = false, info_missing = false;
    std::vector<int32_t> prior_kind, prior_index; std::vector<double> prior_A, prior_b;
    std::vector<int32_t> slot_of_obs;
    int32_t* d_slot_of_obs = nullptr;
    // shared blocks: caller-owned arrays updated in place like the pose / point blocks
    double* h_intr_rw = nullptr;                       // n_models * 10 (set_intrinsics_blocks), else nullptr: intrinsics constant
    std::vector<uint8_t> intr_fixed_mask;              // per model, bit k: block k (f, pp, k, t, B) constant
    std::vector<int32_t> ipr_model, ipr_block; std::vector<double> ipr_A, ipr_b;
    std::vector<double> cam_offset6;                   // per camera (r_off, t_off), empty: identity
    double* h_mount_b = nullptr; double* h_mount_l = nullptr; int32_t Nmt = 0;
    std::vector<uint8_t> mount_fixed; std::vector<int32_t> cam_mount;
    int32_t Nce = 0;
    double* dn_D = nullptr; double* dn_Linv = nullptr; double* dn_work = nullptr; double* dn_x = nullptr; int dn_n = 0;   // dense factor of S (lazy)
    int dense_alloc();
    std::vector<int32_t> pair_c; std::vector<double> pair_px, pair_w; bool pair_has_info = false;
    int32_t pair_loss_kind = 0; double pair_loss_scale = 1.0;
    std::vector<int32_t> pblk_i_h, pblk_j_h;          // distinct camera pairs of the factors
    int upload_pair_tables();
    std::vector<int32_t> model_block, mount_block;     // first pseudo-camera of a free model / mounting, -1: constant
    int plan_shared_blocks();                          // groups + pseudo-cameras (before the allocations of finalize)
    int upload_shared_tables();                        // device side of the shared blocks (after the allocations)
    int shared_linearize();                            // W, sh_H, sh_g (+ all-reduce), Hcc / gc of the pseudo-cameras
    int shared_build();                                // Y, border, Schur terms (+ all-reduce), Sd / rhs / tile blocks of the pseudo-cameras

    DeviceArena arena;
    cudaStream_t st = nullptr;
    double* h_scalars = nullptr;   // pinned, SC_N
    int32_t* h_flags = nullptr;    // pinned, FL_N
    double* d_red = nullptr;       // small device buffer for scalar all-reduces
    double* h_stage = nullptr;     // pinned staging for parameters
    size_t h_stage_n = 0;
    double* d_stage = nullptr; size_t d_stage_n = 0;
    double* d_HG = nullptr;        // Hcc | gc | tail(2) contiguous
    Comm comm;
    PhaseTimer timer;
    int launches = 0;
    bool scaling_done = false;
    void* xchg_mailbox = nullptr;  // this rank's mailbox of the peer-memory all-reduce (own cudaMalloc: exported through CUDA IPC)
    std::vector<void*> ipc_opened; // peer mailboxes opened with cudaIpcOpenMemHandle
    void release_peer_exchange();
    void* ex_S_ipc = nullptr;             // multi-GPU: the partial tiles in an allocation of their own (exported through CUDA IPC)
    std::vector<void*> ipc_tiles_opened;  // peer partial tiles opened with cudaIpcOpenMemHandle
    void release_peer_tiles();
    struct Pinned { const void* p; size_t bytes; bool ok; };
    std::vector<Pinned> pinned;    // caller parameter arrays registered with cudaHostRegister (sba_solver_layout.inc)
    bool pin_caller(const void* p, size_t bytes);
    void unpin_all();
    bool tiles_stale = false;      // direct mode: the panels were rebuilt since the dense tiles were last refreshed
};

// sba_noalloc.cu: stateless DLT on device pointers, all scratch in one static __device__ array (sba_triangulate_noalloc)
int triangulate_noalloc(int32_t n_cams, const double* raxis, const double* t, const int32_t* cam_model, int32_t n_models,
                        const double* intr10, int32_t n_points, int32_t n_views, const int32_t* view_off, const int32_t* view_cam,
                        const double* view_uv, double* xyz_out, uint8_t* ok_out, void* stream, std::string& err);

}  // namespace sba
