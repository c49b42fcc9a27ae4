// sba_capi.cpp -- the extern "C" shell of include/sba_b200.h over sba::Solver.
#include <cstring>
#include <new>
#include <string>

#include "sba_solver.h"

struct sba_solver {
    sba::Solver impl;
    explicit sba_solver(const sba_options& o) : impl(o) {}
};

static thread_local std::string g_create_error;

#define GUARD(s) do { if (!(s)) return SBA_ERR_INVALID; } while (0)

extern "C" {

const char* sba_version(void) { return "sba-b200 0.1 (sm_100a, fp64)"; }

void sba_default_options(sba_options* o) {
    if (!o) return;
    std::memset(o, 0, sizeof(*o));
    o->device = -1;
    o->max_iterations = 50;
    o->function_tolerance = 1e-8;     // setFuncTolerance, modularsbasolver.cpp:921
    o->parameter_tolerance = 1e-10;   // setParamsTolerance, :922
    o->gradient_tolerance = 1e-10;
    o->initial_radius = 1e4;
    o->max_radius = 1e16;
    o->min_radius = 1e-32;
    o->min_lm_diagonal = 1e-6;
    o->max_lm_diagonal = 1e32;
    o->min_relative_decrease = 1e-3;
    o->jacobi_scaling = 1;
    o->pcg_rel_tol = 1e-10;
    o->pcg_max_iters = 500;
    o->pcg_check_every = 1;
    o->operator_mode = 0;
    o->verbose = 0;
    o->linear_solver = 0;
    o->pcg_eta = 0.0;
}

int sba_create(const sba_options* o, sba_solver** out) {
    if (!out) return SBA_ERR_INVALID;
    *out = nullptr;
    sba_options opt;
    if (o) opt = *o; else sba_default_options(&opt);
    sba_solver* s = new (std::nothrow) sba_solver(opt);
    if (!s) { g_create_error = "out of host memory"; return SBA_ERR_INVALID; }
    const int rc = s->impl.open_device();
    if (rc) { g_create_error = s->impl.err; delete s; return rc; }
    *out = s;
    return SBA_OK;
}
void sba_destroy(sba_solver* s) { delete s; }
const char* sba_last_error(const sba_solver* s) { return s ? s->impl.err.c_str() : "null handle"; }
const char* sba_create_error(void) { return g_create_error.c_str(); }

int sba_set_options(sba_solver* s, const sba_options* o) { GUARD(s && o); const int dev = s->impl.opt.device; s->impl.opt = *o; s->impl.opt.device = dev; return SBA_OK; }
int sba_set_intrinsics(sba_solver* s, int32_t n, const double* intr10) { GUARD(s); return s->impl.set_intrinsics(n, intr10); }
int sba_set_intrinsics_blocks(sba_solver* s, int32_t n, double* intr10, const uint8_t* fixed_mask) { GUARD(s); return s->impl.set_intrinsics_blocks(n, intr10, fixed_mask); }
int sba_add_intrinsics_priors(sba_solver* s, int32_t n, const int32_t* model, const int32_t* block, const double* A9, const double* b3) {
    GUARD(s); return s->impl.add_intrinsics_priors(n, model, block, A9, b3);
}
int sba_set_sensor_offsets(sba_solver* s, const double* offset6) { GUARD(s); return s->impl.set_sensor_offsets(offset6); }
int sba_set_lever_arms(sba_solver* s, int32_t n_mounts, double* boresight, double* lever, const uint8_t* fixed, const int32_t* cam_mount) {
    GUARD(s); return s->impl.set_lever_arms(n_mounts, boresight, lever, fixed, cam_mount);
}
int sba_add_uv_pairs(sba_solver* s, int32_t n, const int32_t* cam1, const int32_t* cam2, const double* uv1, const double* uv2,
                     const double* info1, const double* info2) { GUARD(s); return s->impl.add_uv_pairs(n, cam1, cam2, uv1, uv2, info1, info2); }
int sba_set_pair_loss(sba_solver* s, int32_t kind, double a) { GUARD(s); return s->impl.set_pair_loss(kind, a); }
int sba_evaluate_pairs(sba_solver* s, double* res4, double* jac48) { GUARD(s); return s->impl.evaluate_pairs(res4, jac48); }
int sba_point_covariances(sba_solver* s, double* cov9) { GUARD(s); return s->impl.point_covariances(cov9); }
int sba_shared_info(const sba_solver* s, int32_t* n_ext, int32_t* n_groups, int32_t* model_block, int32_t* mount_block) {
    GUARD(s); return s->impl.shared_info(n_ext, n_groups, model_block, mount_block);
}
int sba_set_poses(sba_solver* s, double* raxis, double* t, int32_t n, const int32_t* cam_model, const uint8_t* fixed,
                  const uint8_t* cost_kind) { GUARD(s); return s->impl.set_poses(raxis, t, n, cam_model, fixed, cost_kind); }
int sba_set_points(sba_solver* s, double* xyz, int32_t n, const uint8_t* fixed) { GUARD(s); return s->impl.set_points(xyz, n, fixed); }
int sba_add_observations(sba_solver* s, int32_t n, const int32_t* cam_idx, const int32_t* pt_idx, const double* uv,
                         const double* info) { GUARD(s); return s->impl.add_observations(n, cam_idx, pt_idx, uv, info); }
int sba_add_priors(sba_solver* s, int32_t n, const int32_t* kind, const int32_t* index, const double* A9,
                   const double* b3) { GUARD(s); return s->impl.add_priors(n, kind, index, A9, b3); }
int sba_finalize(sba_solver* s) { GUARD(s); return s->impl.finalize(); }
int sba_upload_parameters(sba_solver* s) { GUARD(s); return s->impl.upload_parameters(); }
int sba_download_parameters(sba_solver* s) { GUARD(s); return s->impl.download_parameters(); }
int sba_evaluate(sba_solver* s, double* cost, double* residuals, double* jac_cam, double* jac_pt) {
    GUARD(s); return s->impl.evaluate(cost, residuals, jac_cam, jac_pt);
}
int sba_normal_blocks(sba_solver* s, double* gc, double* gp, double* Hcc, double* Hpp) { GUARD(s); return s->impl.normal_blocks(gc, gp, Hcc, Hpp); }
int sba_schur_matvec(sba_solver* s, const double* lam_c, const double* lam_p, const double* x, double* y, double* rhs) {
    GUARD(s); return s->impl.schur_matvec(lam_c, lam_p, x, y, rhs);
}
int sba_solve(sba_solver* s, int32_t copy_params, sba_summary* summary) { GUARD(s && summary); return s->impl.solve(copy_params, summary); }
int sba_set_iteration_callback(sba_solver* s, sba_iteration_callback cb, void* user) { GUARD(s); s->impl.cb = cb; s->impl.cb_user = user; return SBA_OK; }
int sba_time_stage(sba_solver* s, int32_t stage, int32_t reps, double* ms) { GUARD(s && ms); return s->impl.time_stage(stage, reps, ms); }
int sba_triangulate(sba_solver* s, int32_t n_points, const int32_t* view_off, const int32_t* view_cam, const double* view_uv,
                    double* xyz_out, uint8_t* ok_out, double* ms_kernel) {
    GUARD(s); return s->impl.triangulate(n_points, view_off, view_cam, view_uv, xyz_out, ok_out, ms_kernel);
}
int sba_triangulate_noalloc(int32_t n_cams, const double* raxis, const double* t, const int32_t* cam_model, int32_t n_models,
                            const double* intr10, int32_t n_points, int32_t n_views, const int32_t* view_off,
                            const int32_t* view_cam, const double* view_uv, double* xyz_out, uint8_t* ok_out, void* stream) {
    std::string err;
    const int rc = sba::triangulate_noalloc(n_cams, raxis, t, cam_model, n_models, intr10, n_points, n_views, view_off, view_cam,
                                            view_uv, xyz_out, ok_out, stream, err);
    if (rc) g_create_error = err;
    return rc;
}
int sba_evaluate_intrinsics_jacobian(sba_solver* s, double* jac_intr) { GUARD(s); return s->impl.evaluate_intrinsics_jacobian(jac_intr); }
int sba_evaluate_offset(sba_solver* s, const double* off12, double* residuals, double* jac_cam, double* jac_pt, int32_t* n_failed) {
    GUARD(s); return s->impl.evaluate_offset(off12, residuals, jac_cam, jac_pt, n_failed);
}

int sba_comm_unique_id(uint8_t id128[128]) {
    std::string err;
    if (sba::Comm::unique_id(id128, err)) { g_create_error = err; return SBA_ERR_COMM; }
    return SBA_OK;
}
int sba_comm_init(sba_solver* s, int32_t rank, int32_t nranks, const uint8_t id128[128]) { GUARD(s && id128); return s->impl.comm_init(rank, nranks, id128); }

int sba_partition_points(int32_t n_pts, const int32_t* obs_per_point, int32_t nranks, int32_t* begin) {
    if (n_pts < 0 || nranks <= 0 || !begin || (n_pts > 0 && !obs_per_point)) return SBA_ERR_INVALID;
    int64_t total = 0;
    for (int32_t p = 0; p < n_pts; p++) total += obs_per_point[p];
    // contiguous ranges; rank r ends at the first point where the running count reaches (r+1)/nranks of the total
    begin[0] = 0;
    int32_t p = 0;
    int64_t run = 0;
    for (int32_t r = 1; r < nranks; r++) {
        const int64_t target = (total * r + nranks - 1) / nranks;
        while (p < n_pts && run < target) run += obs_per_point[p++];
        begin[r] = p;
    }
    begin[nranks] = n_pts;
    return SBA_OK;
}

int sba_problem_sizes(const sba_solver* s, int32_t* n_cams, int32_t* n_pts, int32_t* n_obs, int32_t* n_slots, int32_t* n_tiles) {
    GUARD(s);
    const sba::DevProblem& P = s->impl.P;
    if (n_cams) *n_cams = P.Nc; if (n_pts) *n_pts = P.Np; if (n_obs) *n_obs = P.No;
    if (n_slots) *n_slots = P.Ns; if (n_tiles) *n_tiles = P.Nt;
    return SBA_OK;
}

int sba_explicit_info(const sba_solver* s, int32_t* enabled, int32_t* n_tiles, int32_t* n_blocks, int32_t* n_contributions) {
    GUARD(s);
    const sba::DevProblem& P = s->impl.P;
    if (enabled) *enabled = P.ex_on; if (n_tiles) *n_tiles = P.ex_ntiles;
    if (n_blocks) *n_blocks = P.ex_nblocks; if (n_contributions) *n_contributions = P.ex_ncontrib;
    return SBA_OK;
}

int sba_explicit_shard(const sba_solver* s, int32_t* first_tile, int32_t* n_own, int32_t* peer_exchange, int32_t* persistent_pcg) {
    GUARD(s);
    const sba::DevProblem& P = s->impl.P;
    if (first_tile) *first_tile = P.ex_on ? P.ex_own0 : 0; if (n_own) *n_own = P.ex_on ? P.ex_nown : 0;
    if (peer_exchange) *peer_exchange = P.xchg_n > 0 ? 1 : 0;
    if (persistent_pcg) *persistent_pcg = (P.ex_on && P.ps_on) ? 1 : 0;
    return SBA_OK;
}

int sba_explicit_sparse_info(const sba_solver* s, int32_t* enabled, int32_t* n_panels, int32_t* n_chunks, int64_t* n_blocks, int64_t* stream_bytes) {
    GUARD(s);
    const sba::DevProblem& P = s->impl.P;
    const bool on = P.ex_on && P.ps_on && P.sp_on;
    if (enabled) *enabled = on ? 1 : 0; if (n_panels) *n_panels = on ? P.sp_npanel : 0; if (n_chunks) *n_chunks = on ? P.sp_nchunk : 0;
    if (n_blocks) *n_blocks = on ? P.sp_nblocks : 0; if (stream_bytes) *stream_bytes = on ? P.sp_bytes : 0;
    return SBA_OK;
}

int sba_explicit_tiles(sba_solver* s, double* tiles, int32_t* tile_i, int32_t* tile_j, double* rhs) {
    GUARD(s);
    const sba::DevProblem& P = s->impl.P;
    if (!P.ex_on || P.ex_Sown != P.ex_S) return SBA_ERR_INVALID;   // single-GPU inspection hook (a rank holds only its slice)
    s->impl.ensure_dense_tiles();   // direct mode: the build keeps only the panels current
    cudaDeviceSynchronize();
    const size_t n = (size_t)P.ex_ntiles;
    if (tiles && cudaMemcpy(tiles, P.ex_S, n * 384 * 384 * sizeof(double), cudaMemcpyDeviceToHost) != cudaSuccess) return SBA_ERR_CUDA;
    if (tile_i && cudaMemcpy(tile_i, P.ex_tile_I, n * sizeof(int32_t), cudaMemcpyDeviceToHost) != cudaSuccess) return SBA_ERR_CUDA;
    if (tile_j && cudaMemcpy(tile_j, P.ex_tile_J, n * sizeof(int32_t), cudaMemcpyDeviceToHost) != cudaSuccess) return SBA_ERR_CUDA;
    if (rhs && cudaMemcpy(rhs, P.rhs, 6 * (size_t)P.Nce * sizeof(double), cudaMemcpyDeviceToHost) != cudaSuccess) return SBA_ERR_CUDA;
    return SBA_OK;
}

}  // extern "C"
