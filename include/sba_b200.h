/*
 * sba_b200.h -- C ABI of libsba_b200.so, the B200-native (sm_100a, FP64 CUDA)
 * bundle-adjustment engine that drops in for the hot path of steviapp's
 * libs/sparsesolver SBA solver.
 *
 * Every entry point names the reference interface it replaces (paths relative
 * to the reference tree, french-paragon/steviapp @ 08eee50c).  Plain pointers
 * and sizes only; no exception crosses this boundary; every function that can
 * fail returns an int status (0 = ok, <0 = error, text via sba_last_error).
 *
 * There is NO CPU fallback: if no CUDA device is usable sba_create fails with
 * SBA_ERR_NO_DEVICE.
 *
 * Conventions kept from the reference:
 *   - pose parameter blocks are sensor->world: rAxis[3] (axis-angle, ambient
 *     R^3, no manifold) and t[3] (PoseNode, libs/sparsesolver/modularsbasolver.h:593-599;
 *     AddParameterBlock(ptr,3) at modularsbasolver.cpp:837-838);
 *   - intrinsics per camera model are the five blocks f[1], pp[2], k[3], t[2],
 *     B[2] in that order (sbamodules/pinholecameraprojectormodule.cpp:112-122),
 *     10 doubles; B is read but unused and the skew uses t (reference quirk,
 *     sbamodules/pinholecameraprojectormodule.h:46-47);
 *   - cost = 1/2 sum r^2 (Ceres convention);
 *   - parameter arrays stay owned by the caller and are UPDATED IN PLACE by
 *     sba_solve, as ceres::Solve does with the registered blocks.
 */
#ifndef SBA_B200_H
#define SBA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SBA_INTR_STRIDE 10

/* status codes */
#define SBA_OK 0
#define SBA_ERR_NO_DEVICE (-1)
#define SBA_ERR_CUDA (-2)
#define SBA_ERR_INVALID (-3)      /* bad argument / call order */
#define SBA_ERR_EVAL_FAILED (-4)  /* initial point: a plane-cost functor "returned false" (Pbar.z < 0) */
#define SBA_ERR_NUMERIC (-5)      /* singular point block / non-PD preconditioner block */
#define SBA_ERR_COMM (-6)         /* NCCL unavailable or failed */

/* cost kinds: UV2ParametrizedXYZCost Flags (costfunctors/modularuvprojection.h:340-457),
 * selected per solver by the Qt property useSphericalCostPinholeCamProjCost
 * (sbamodules/pinholecameraprojectormodule.cpp:28,257-271); per camera here so
 * mixed scenes (BASELINE config 4) run in one problem. */
#define SBA_COST_PLANE 0
#define SBA_COST_SPHERICAL 1

/* prior targets: FixedSizeNormalPrior<3,3> on a point (landmarkssbamodule.cpp:171-202)
 * or on a pose block (imagealignementsbamodule.cpp:256-286) */
#define SBA_PRIOR_POINT 0
#define SBA_PRIOR_CAM_R 1
#define SBA_PRIOR_CAM_T 2

/* Replaces the ModularSBASolver constructor arguments and setters
 * (modularsbasolver.h:672,876-884) plus the ceres::Solver::Options the
 * reference fills at modularsbasolver.cpp:897-924; everything it leaves at the
 * Ceres default is exposed with that default. */
typedef struct sba_options {
    int32_t device;               /* CUDA device ordinal; -1 = current device */
    int32_t max_iterations;       /* setOptimizationSteps */
    double function_tolerance;    /* setFuncTolerance (1e-8)  */
    double parameter_tolerance;   /* setParamsTolerance (1e-10) */
    double gradient_tolerance;    /* Ceres default 1e-10 */
    double initial_radius;        /* 1e4 */
    double max_radius;            /* 1e16 */
    double min_radius;            /* 1e-32 */
    double min_lm_diagonal;       /* 1e-6 */
    double max_lm_diagonal;       /* 1e32 */
    double min_relative_decrease; /* 1e-3 */
    int32_t jacobi_scaling;       /* 1 */
    double pcg_rel_tol;           /* relative tolerance of the reduced solve (sqrt(r.z / r0.z0)) */
    int32_t pcg_max_iters;        /* Ceres max_linear_solver_iterations = 500 */
    int32_t pcg_check_every;      /* host reads the convergence scalar every n PCG iterations (1 = exact oracle iteration count) */
    int32_t operator_mode;        /* reduced camera system S inside the PCG.
                                     0 = auto: EXPLICIT S (dense 384x384 tiles over the upper block triangle, built per LM
                                         iteration from static contribution slots) when its tiles are a smaller stream than the
                                         implicit passes, it fits in HBM and the run is single-GPU; else implicit as in 2;
                                     1 = implicit operator reading the materialised A, C (132 B/obs camera pass);
                                     2 = implicit operator, factored camera pass when every stiffness is identity
                                         (rebuilds A, C from Pbar: 60 B/obs), else as 1;
                                     3 = explicit S forced (error if it cannot be built) */
    int32_t verbose;
    int32_t linear_solver;        /* reduced solve: 0 = SCHUR_JACOBI-preconditioned CG (ceres::ITERATIVE_SCHUR), 1 = Cholesky of the explicit
                                     reduced camera system on the device -- the reference's default solver type, ceres::SPARSE_SCHUR
                                     (modularsbasolver.h:672, modularsbasolver.cpp:898), is a factorisation of the same S; needs
                                     operator_mode 0 / 3, one GPU */
    double pcg_eta;               /* 0: CG to pcg_rel_tol.  > 0: inexact Newton as Ceres runs ITERATIVE_SCHUR (eta = 0.1 by default there): CG stops at
                                     sqrt(r.z / r0.z0) <= max(pcg_eta, pcg_rel_tol) -- a throughput setting, not a parity one */
} sba_options;

#define SBA_MAX_HISTORY 512
typedef struct sba_summary {
    int32_t status;               /* 0 = iteration cap, 1 = converged, <0 failure */
    int32_t iterations;           /* LM iterations executed */
    int32_t successful_steps;
    int32_t pcg_iterations_total;
    double initial_cost, final_cost;
    double cost_history[SBA_MAX_HISTORY];
    double radius_history[SBA_MAX_HISTORY];
    int32_t pcg_iters_history[SBA_MAX_HISTORY];
    /* device time, CUDA events on the solver's stream, milliseconds */
    double ms_total;              /* whole LM loop, inputs already in HBM */
    double ms_linearize;          /* K1 residual+Jacobian launches */
    double ms_schur;              /* K2 point blocks + camera blocks + preconditioner */
    double ms_pcg;                /* K3 */
    double ms_update;             /* K4 back-substitution, update, trial cost */
    double ms_h2d, ms_d2h;        /* host<->device copies done by sba_solve */
    int32_t kernel_launches;      /* kernels launched by the last sba_solve */
} sba_summary;

/* one LM iteration report; replaces ModularSBASolverIterationCallback
 * (modularsbasolver.cpp:30-56): return 0 to continue, non-zero to abort */
typedef struct sba_iteration_info {
    int32_t iteration;
    double cost, cost_change, radius;
    int32_t step_valid, step_accepted, pcg_iterations;
} sba_iteration_info;
typedef int (*sba_iteration_callback)(const sba_iteration_info*, void* user);

typedef struct sba_solver sba_solver;

void sba_default_options(sba_options* o);

/* ModularSBASolver::ModularSBASolver + init() prologue (modularsbasolver.cpp:679-706) */
int sba_create(const sba_options* o, sba_solver** out);
/* ModularSBASolver::cleanup (modularsbasolver.cpp:1091-1143) */
void sba_destroy(sba_solver* s);
const char* sba_last_error(const sba_solver* s);
/* error text for a failed sba_create (no handle yet) */
const char* sba_create_error(void);

/* Camera models: PinholeCamProjModule::init intrinsic blocks
 * (sbamodules/pinholecameraprojectormodule.cpp:302-533), held constant
 * (FixedParameter::CameraInternal, :358-361,406-408). */
int sba_set_intrinsics(sba_solver* s, int32_t n_models, const double* intr10);

/* Camera models with FREE intrinsic blocks -- the reference's default: PinholeCamProjModule::init adds the
 * five blocks f[1], pp[2], k[3], t[2], B[2] per Camera and holds a block constant only when the Camera is fixed,
 * FixedParameter::CameraInternal is set or the distortion model is switched off
 * (sbamodules/pinholecameraprojectormodule.cpp:302-533; constant branches :358-361,404-408,456-460,503-507).
 * intr10 stays owned by the caller and is UPDATED IN PLACE by sba_solve (as Ceres updates the module's blocks);
 * fixed_mask[m] bit k set -> block k of model m constant (k = 0 f, 1 pp, 2 k, 3 t, 4 B); NULL = everything constant.
 * The B block has an identically zero Jacobian (the skew is built from the tangential block,
 * pinholecameraprojectormodule.h:46-47): its step is zero whether it is free or not.  Free blocks join the reduced camera
 * system as pseudo-cameras (sba_shared_info); they need the explicit reduced system (operator_mode 0 or 3). */
int sba_set_intrinsics_blocks(sba_solver* s, int32_t n_models, double* intr10, const uint8_t* fixed_mask);
/* FixedSizeNormalPrior<n,n> on an intrinsic block (pinholecameraprojectormodule.cpp:327-345,365-378,421-438,470-485,517-531):
 * r = A (x - b), A = dim x dim row-major in the leading entries of its 9 doubles, b in the leading entries of its 3. */
int sba_add_intrinsics_priors(sba_solver* s, int32_t n, const int32_t* model, const int32_t* block, const double* A9, const double* b3);
/* Constant sensor -> body offsets: offset6 = n_cams x (r_off[3] axis-angle, t_off[3]); the projection
 * cost is evaluated at pose * offset (PoseTransformCost, pinholecameraprojectormodule.cpp:23,197-214;
 * costfunctors/posedecoratorfunctors.h:628-700).  NULL = identity for every camera.  Call after sba_set_poses. */
int sba_set_sensor_offsets(sba_solver* s, const double* offset6);
/* Free boresight / lever-arm blocks (LeverArmCost / PoseTransformLeverArmCost, pinholecameraprojectormodule.cpp:21,25,60-99,
 * 147-166; costfunctors/posedecoratorfunctors.h:148-264): sensor2world = (pose * offset) * (boresight, leverArm).
 * boresight / lever are n_mounts x 3, caller-owned, UPDATED IN PLACE; cam_mount[c] = mounting of camera c or -1;
 * fixed[k] != 0 holds mounting k constant.  Call after sba_set_poses. */
int sba_set_lever_arms(sba_solver* s, int32_t n_mounts, double* boresight, double* lever, const uint8_t* fixed, const int32_t* cam_mount);
/* layout of the reduced camera system after finalize: n_ext extended cameras (6 columns each) = the poses followed by two
 * pseudo-cameras [f ppx ppy k1 k2 k3] [t1 t2 - - - -] per free model and one [boresight | leverArm] per free mounting;
 * model_block[m] / mount_block[k] = first pseudo-camera or -1.  Any pointer may be NULL.  Measurement / test hook. */
int sba_shared_info(const sba_solver* s, int32_t* n_ext, int32_t* n_groups, int32_t* model_block, int32_t* mount_block);

/* UV <-> UV cross-projection factors: ProjectorModule::addCrossProjectionCostFunction
 * (modularsbasolver.h / modularsbasolver.cpp:1616-1739) with the plain decorators (identity offsets, no lever arms), i.e. the
 * functor InvertPoseDynamic<InvertPoseDynamic<UV2UVCost>> (costfunctors/modularuvprojection.h:660-795): factor k links poses
 * cam1[k], cam2[k] through the pixels uv1, uv2 (2 n each) with stiffness info1, info2 (4 n each, row-major, both NULL =
 * identity); 4 residuals, no landmark.  The intrinsic blocks of the two cameras' models take part when they are free. */
int sba_add_uv_pairs(sba_solver* s, int32_t n, const int32_t* cam1, const int32_t* cam2, const double* uv1, const double* uv2,
                     const double* info1, const double* info2);
/* loss on those factors: 0 none, 1 ceres::HuberLoss(a), 2 CauchyLoss(a), 3 ArctanLoss(a) -- the reference passes a = 1
 * (sbamodules/correspondencessetsbamodule.cpp:2459-2471); applied as Ceres' Corrector does (residual / Jacobian rescaling) */
#define SBA_LOSS_NONE 0
#define SBA_LOSS_HUBER 1
#define SBA_LOSS_CAUCHY 2
#define SBA_LOSS_ARCTAN 3
int sba_set_pair_loss(sba_solver* s, int32_t kind, double a);
/* raw residuals (4 per factor) and raw pose Jacobians (4 x 12 row-major: d/d rAxis1, t1, rAxis2, t2) before the loss;
 * ceres::Problem::Evaluate on those residual blocks.  Parity hook. */
int sba_evaluate_pairs(sba_solver* s, double* res4, double* jac48);

/* ModularSBASolver::std_step (modularsbasolver.cpp:952-1004): marginal covariance of the point blocks at the current
 * parameters -- ceres::Covariance over the (pos, pos) pairs LandmarksSBAModule::requestUncertainty asks for
 * (sbamodules/landmarkssbamodule.cpp:264-297).  cov9 = n_pts x 3 x 3 row-major (zeros for constant points).  From the
 * undamped normal equations by a dense Cholesky + selected inverse of the reduced camera system on the device; returns
 * SBA_ERR_NUMERIC ("Failed to compute covariance matrix", the reference's message) when the system is rank deficient. */
int sba_point_covariances(sba_solver* s, double* cov9);

/* Pose blocks: getNodeForFrame + initManagedParameters
 * (modularsbasolver.cpp:233-306,804-870).  raxis/t are 3*n AoS, caller-owned,
 * updated in place by sba_solve.  fixed -> SetParameterBlockConstant. */
int sba_set_poses(sba_solver* s, double* raxis, double* t, int32_t n,
                  const int32_t* cam_model, const uint8_t* fixed,
                  const uint8_t* cost_kind);

/* Point blocks: getNodeForLandmark (modularsbasolver.cpp:196-231). */
int sba_set_points(sba_solver* s, double* xyz, int32_t n, const uint8_t* fixed);

/* Bulk form of ProjectorModule::addProjectionCostFunction
 * (modularsbasolver.h:169-177; PinholeCamProjModule impl
 * sbamodules/pinholecameraprojectormodule.cpp:46-272): observation i links
 * pose cam_idx[i] and point pt_idx[i] with pixel uv[2i..2i+1] and 2x2
 * stiffness info[4i..4i+3] (row-major; NULL = identity).  May be called
 * repeatedly; observations accumulate in call order. */
int sba_add_observations(sba_solver* s, int32_t n, const int32_t* cam_idx,
                         const int32_t* pt_idx, const double* uv,
                         const double* info);

/* problem.AddResidualBlock(new FixedSizeNormalPrior<3,3>(A,b), nullptr, ptr)
 * (costfunctors/fixedsizenormalprior.h:75-112): r = A (x - b), A row-major 3x3. */
int sba_add_priors(sba_solver* s, int32_t n, const int32_t* kind,
                   const int32_t* index, const double* A9, const double* b3);

/* End of init(): sorts observations by camera (32-slot aligned segments),
 * builds the by-point index and uploads everything to HBM.  Called implicitly
 * by the first evaluate/solve. */
int sba_finalize(sba_solver* s);

/* Re-upload the caller's current parameter values (warm start semantics of
 * modularsbasolver.cpp:265-302) / read the device values back. */
int sba_upload_parameters(sba_solver* s);
int sba_download_parameters(sba_solver* s);

/* ceres::Problem::Evaluate as the reference tests call it through
 * ceresProblem() (modularsbasolver.h:856, tests/test_sba_solver/main.cpp:219-234).
 * Outputs in the caller's observation order, any may be NULL:
 * residuals 2*n_obs, jac_cam n_obs*2*6 row-major [d/dr | d/dt], jac_pt n_obs*2*3.
 * Returns SBA_OK, or 1 if some plane observation has Pbar.z < 0 (those rows are 0). */
int sba_evaluate(sba_solver* s, double* cost, double* residuals,
                 double* jac_cam, double* jac_pt);

/* Test hooks on the linear algebra (same shapes as the oracle's):
 * gc 6*n_cams, gp 3*n_pts, Hcc 36*n_cams, Hpp 9*n_pts (undamped). */
int sba_normal_blocks(sba_solver* s, double* gc, double* gp, double* Hcc, double* Hpp);
/* Jacobian of the residuals with respect to the five intrinsic blocks f[1], pp[2], k[3], t[2], B[2] of each observation's
 * camera model, n_obs x 2 x 10 row-major, caller order, at the current parameters: the blocks the reference's
 * AutoDiffCostFunction<..., 2, 3,3,3, 1,2,3,2,2> (sbamodules/pinholecameraprojectormodule.cpp:105-122) differentiates when
 * the intrinsics are free (PinholeCamProjModule::init, :302-533).  The B columns are identically zero (the skew is built
 * from the tangential block, pinholecameraprojectormodule.h:46-47).  First half of SURVEY 8(f) row 1: the LM itself still
 * holds the intrinsics constant (FixedParameter::CameraInternal). */
int sba_evaluate_intrinsics_jacobian(sba_solver* s, double* jac_intr);

/* Residuals and Jacobians with a CONSTANT sensor offset per camera: the cost the reference builds as
 * PoseTransform<UVCost, SourceToInitial> (costfunctors/posedecoratorfunctors.h:628-700,
 * sbamodules/pinholecameraprojectormodule.cpp:197-214), i.e. the UV cost at pose * offset.  off12 = per camera Roff^T
 * (row-major 3x3) followed by toff; outputs in caller order: residuals [2 n_obs], jac_cam [n_obs x 2 x 6] (d/drAxis, d/dt of
 * the BODY pose), jac_pt [n_obs x 2 x 3]; n_failed (may be NULL) counts plane-cost evaluations behind the composed sensor
 * pose (zeros written).  Inspection / parity hook: the LM itself does not take offsets yet. */
int sba_evaluate_offset(sba_solver* s, const double* off12, double* residuals, double* jac_cam, double* jac_pt, int32_t* n_failed);

/* y = S x and reduced rhs for explicit damping diagonals lam_c (6*n_cams), lam_p (3*n_pts) */
int sba_schur_matvec(sba_solver* s, const double* lam_c, const double* lam_p,
                     const double* x, double* y, double* rhs);

/* ModularSBASolver::opt_step -> ceres::Solve (modularsbasolver.cpp:884-951).
 * copy_params: 1 = upload the caller's arrays first and write results back
 * (the reference's observable behaviour); 0 = run on the values already
 * resident in HBM and leave the result there (sba_download_parameters reads it). */
int sba_solve(sba_solver* s, int32_t copy_params, sba_summary* summary);
int sba_set_iteration_callback(sba_solver* s, sba_iteration_callback cb, void* user);
int sba_set_options(sba_solver* s, const sba_options* o);

/* Kernel-level timing hook for bench.py: runs `reps` back-to-back launches of
 * one pipeline stage on the resident problem and returns the mean device time
 * per launch in ms (CUDA events on the solver's stream).
 * stage: 0 = K1 residual+Jacobian, 1 = K1c cost only, 2 = K1p/K2a (fused) + K2b, 3 = reduced-system mat-vec (the
 * operator the PCG applies: explicit tiles or implicit passes), 4 = preconditioner build, 5 = K1p/K2a, 6 = implicit
 * point pass, 7 = implicit camera pass, 9 = K2b camera blocks, 10 = explicit tile mat-vec kernel, 11 = explicit build
 * (contributions + assembly), 12 = fused PCG vector kernel alone, 13 = one whole PCG iteration (tile mat-vec + fused
 * kernel). */
int sba_time_stage(sba_solver* s, int32_t stage, int32_t reps, double* ms_per_launch);

/* Batched multi-view DLT triangulation (adjacent to the LM loop; reference has
 * pairwise forms only: sbainitializer.cpp:987-1063, modularuvprojection.h:741-757).
 * Uses the poses/intrinsics registered on the handle.  views of point i are
 * view_off[i]..view_off[i+1]. */
int sba_triangulate(sba_solver* s, int32_t n_points, const int32_t* view_off,
                    const int32_t* view_cam, const double* view_uv,
                    double* xyz_out, uint8_t* ok_out, double* ms_kernel);

/* Multi-GPU (one process per GPU): points and their observations are sharded
 * by the caller (sba_partition_points), cameras replicated; the library
 * all-reduces camera blocks, mat-vec partials and scalars with NCCL.
 * unique_id is a 128-byte ncclUniqueId made by rank 0 (sba_comm_unique_id)
 * and distributed by the launcher. */
int sba_comm_unique_id(uint8_t id128[128]);
int sba_comm_init(sba_solver* s, int32_t rank, int32_t nranks, const uint8_t id128[128]);

/* Host-only helper: contiguous point ranges balanced by observation count.
 * begin has nranks+1 entries. */
int sba_partition_points(int32_t n_pts, const int32_t* obs_per_point,
                         int32_t nranks, int32_t* begin);

/* sizes of the resident problem (after finalize): n_slots is the padded
 * camera-sorted slot count, n_tiles the warp tiles */
int sba_problem_sizes(const sba_solver* s, int32_t* n_cams, int32_t* n_pts,
                      int32_t* n_obs, int32_t* n_slots, int32_t* n_tiles);

/* explicit reduced camera system (operator_mode 0 auto / 3): whether the PCG applies S from the dense 384x384 tiles,
 * and the static sizes of that structure (tiles over the upper block triangle, off-diagonal 6x6 blocks with a common
 * point, point-pair contributions).  Replaces nothing in the reference: Ceres keeps the same numbers private in its
 * SchurEliminator / BlockRandomAccessSparseMatrix (behind modularsbasolver.cpp:938); exposed for measurement. */
int sba_explicit_info(const sba_solver* s, int32_t* enabled, int32_t* n_tiles,
                      int32_t* n_blocks, int32_t* n_contributions);

/* multi-GPU view of the explicit system: the summed tiles are sharded in contiguous ranges (one reduce-scatter per LM
 * iteration); this rank applies the tiles first_tile .. first_tile + n_own - 1.  peer_exchange = 1 when the per-iteration
 * all-reduce of the 6 n_cams operator partial runs inside the fused PCG kernel over CUDA-IPC peer mailboxes (NVLink),
 * 0 when it goes through NCCL (or on one GPU).  persistent_pcg = 1 when the whole PCG solve of an LM iteration runs as
 * one persistent cooperative kernel (k_pcg_solve).  Measurement only; no reference counterpart. */
int sba_explicit_shard(const sba_solver* s, int32_t* first_tile, int32_t* n_own, int32_t* peer_exchange,
                       int32_t* persistent_pcg);

/* block-sparse copy of the tiles (persistent PCG kernel): the CG iterations stream, per tile and block row, only the
 * 6x6 blocks that can be non-zero (two cameras that share a point) as PANELS of 6 rows with a 16-bit column map; consecutive
 * panels of a tile travel in CHUNKS of <= 40 KB (a panel directory + the panels, one bulk copy each).  enabled = 1 when that copy is what k_pcg_solve reads
 * (default: a rank's tiles >= 64 MB and panels <= 85 % of them; SBA_PS_SPARSE=0/1 forces); n_blocks = blocks kept; stream_bytes =
 * bytes one CG iteration reads.  Same role as the block structure of Ceres' BlockRandomAccessSparseMatrix behind
 * modularsbasolver.cpp:938.  Measurement only. */
int sba_explicit_sparse_info(const sba_solver* s, int32_t* enabled, int32_t* n_panels, int32_t* n_chunks,
                             int64_t* n_blocks, int64_t* stream_bytes);

/* inspection hook: copies the tiles of the explicit reduced camera system as last built (n_tiles * 384 * 384 doubles,
 * row-major tiles; tile t covers camera tiles (tile_i[t], tile_j[t]), i <= j; diagonal tiles hold S_ii / 2 on their
 * block diagonal and zeros below) and the reduced right-hand side (6 * n_cams).  Any pointer may be NULL.  The analogue
 * in the reference is reading Ceres' reduced system, which Ceres does not expose; used by tests and scripts only. */
int sba_explicit_tiles(sba_solver* s, double* tiles, int32_t* tile_i, int32_t* tile_j, double* rhs);

/* ---- BAL text problems.  Format and camera model of the reference's in-tree BAL example
 * (modules/g2o/g2o/examples/bal/bal_example.cpp:271-340 reads the file, :168-204 is the projection): header
 * "cameras points observations", one "camera point x y" line per observation, 9 doubles per camera (Rodrigues r, t, f, k1,
 * k2; world -> camera, looking down -z), 3 per point.  Host code only: these calls need no CUDA device.
 * Errors: SBA_ERR_INVALID with the text in sba_bal_error() (unreadable file, bad header, index out of bounds -- the
 * reference asserts the same bounds at :312-316 --, truncated file). */
typedef struct sba_bal sba_bal;
int sba_bal_open(const char* path, sba_bal** out);
void sba_bal_close(sba_bal* b);
const char* sba_bal_error(void);
int sba_bal_sizes(const sba_bal* b, int32_t* n_cams, int32_t* n_pts, int32_t* n_obs);
/* the file's own arrays, unconverted (any pointer may be NULL): cam_idx / pt_idx [n_obs], xy [2 n_obs], cam9 [9 n_cams], xyz [3 n_pts] */
int sba_bal_raw(const sba_bal* b, int32_t* cam_idx, int32_t* pt_idx, double* xy, double* cam9, double* xyz);
/* the problem in the conventions of this header, ready for sba_set_intrinsics (n_cams models {f, 0...}) / sba_set_poses /
 * sba_set_points / sba_add_observations: sensor -> world poses looking down +z (R_e = R^T diag(1,-1,-1), t_e = -R^T t),
 * pixels with the file's radial distortion removed and v negated (f, k1, k2 constant: FixedParameter::CameraInternal).
 * n_undistort_failed (may be NULL) counts observations whose radius equation did not converge (kept distorted). */
int sba_bal_convert(const sba_bal* b, double* raxis, double* t, double* intr10, int32_t* cam_model, double* xyz,
                    int32_t* cam_idx, int32_t* pt_idx, double* uv, int32_t* n_undistort_failed);
/* the file's OWN objective, 1/2 sum |pixel(cam9, xyz) - xy|^2 with the BAL projection (bal_example.cpp:168-204) and its RMS in
 * pixels, for cameras cam9 / points xyz (NULL = the file's): costs in BAL's metric before / after a solve.  The engine
 * minimises the residual with the distortion moved to the measurement (first-order equivalent, see sba_bal_convert). */
int sba_bal_cost(const sba_bal* b, const double* cam9, const double* xyz, double* cost, double* rms_px);
/* engine-convention poses (after a solve) back to BAL cameras, f / k1 / k2 from the file */
int sba_bal_cameras_from_engine(const sba_bal* b, const double* raxis, const double* t, double* cam9_out);
/* writes a BAL file (%.17g: reading it back reproduces every double) */
int sba_bal_write(const char* path, int32_t n_cams, int32_t n_pts, int32_t n_obs, const int32_t* cam_idx, const int32_t* pt_idx,
                  const double* xy, const double* cam9, const double* xyz);

/* Push-broom sensors on the pinhole kernels, intrinsics held constant.  Replaces
 * PinholePushbroomUVProjector::dirFromUV (sbamodules/pinholecameraprojectormodule.h:79-116): d = (u - pp + du(s), v + dv(s), f),
 * s = (u - W/2)/W, du / dv = degree-5 polynomials with coefficients as6 / bs6.  Writes the pixels (d_x, d_y) that, with the
 * intrinsics record {f, 0, ...} written to intr10_out (may be NULL), give the same view ray through the pinhole model, so
 * the observations can go to sba_add_observations unchanged.  Host code only. */
int sba_pushbroom_equivalent_uv(int32_t n, const double* uv, double sensor_width, double f, double pp, const double* as6,
                                const double* bs6, double* uv_out, double* intr10_out);

/* ---- The reference's project files (".steviapproj").  Replaces Project::load / save
 * (libs/datablocks/project.cpp:93-221) and the part of the SBA modules that turns data blocks into the problem:
 * GenericSBAGraphReductor (sparsesolver/sbagraphreductor.cpp:172-194: items with fewer observations than degrees of
 * freedom drop out, to the fixed point), LandmarksSBAModule::init (sbamodules/landmarkssbamodule.cpp:111-205),
 * ModularSBASolver::getNodeForFrame (modularsbasolver.cpp:233-305), ImageAlignementSBAModule::init
 * (sbamodules/imagealignementsbamodule.cpp:206-440), PinholeCamProjModule::init (pinholecameraprojectormodule.cpp:302-533),
 * and the three writeResults (landmarkssbamodule.cpp:207-262, imagealignementsbamodule.cpp:446-490,
 * pinholecameraprojectormodule.cpp:535-574).  Host code only: no CUDA device needed.
 * Errors: SBA_ERR_INVALID with the text in sba_project_error() (unreadable file, malformed JSON, a landmark with a
 * geographic CRS -- its conversion needs PROJ --, an image attached to a trajectory and a mounting).
 * Engine indices: models = the Cameras some kept image uses (order of first use), images / landmarks in file order. */
typedef struct sba_project sba_project;
#define SBA_PROJECT_FIX_CAMERA_INTERNAL 1   /* FixedParameter::CameraInternal (fixedpreoptimizedparameters.h) */
int sba_project_open(const char* path, sba_project** out);
void sba_project_close(sba_project* p);
const char* sba_project_error(void);
/* any pointer may be NULL.  n_intrinsics_priors is the count with every intrinsic block free (an upper bound);
 * n_images_without_camera counts enabled, constrained images that were left out because no Camera is assigned. */
int sba_project_sizes(const sba_project* p, int32_t* n_models, int32_t* n_images, int32_t* n_landmarks, int32_t* n_obs,
                      int32_t* n_priors, int32_t* n_intrinsics_priors, int32_t* n_images_without_camera);
/* data-block ids ("id" in the file) of the engine's models / images / landmarks */
int sba_project_ids(const sba_project* p, int64_t* model_id, int64_t* image_id, int64_t* landmark_id);
/* The problem in the conventions of this header (any pointer may be NULL; the four prior arrays and the four
 * intrinsics-prior arrays go together): intr10 [10 n_models] + model_fixed_mask [n_models] for sba_set_intrinsics_blocks,
 * raxis / t [3 n_images] + cam_model + cam_fixed for sba_set_poses, pts [3 n_landmarks] + pt_fixed for sba_set_points,
 * obs_* for sba_add_observations (obs_info [4 n_obs], diagonal), prior_* for sba_add_priors, ipr_* for
 * sba_add_intrinsics_priors (*n_intrinsics_priors = how many were written).  fixed_flags: SBA_PROJECT_FIX_CAMERA_INTERNAL. */
int sba_project_convert(const sba_project* p, int32_t fixed_flags, double* intr10, uint8_t* model_fixed_mask, double* raxis, double* t,
                        int32_t* cam_model, uint8_t* cam_fixed, double* pts, uint8_t* pt_fixed, int32_t* obs_cam, int32_t* obs_pt,
                        double* obs_uv, double* obs_info, int32_t* prior_kind, int32_t* prior_index, double* prior_A9, double* prior_b3,
                        int32_t* ipr_model, int32_t* ipr_block, double* ipr_A9, double* ipr_b3, int32_t* n_intrinsics_priors);
/* writeResults of the three modules into the loaded document: "op" / "or" of the free images, "op" of the landmarks, the
 * optimised "o*" parameters of the free Cameras; every value passes through float, as the modules store it.  Any array NULL
 * = that part is left alone (raxis and t go together). */
int sba_project_write_results(sba_project* p, const double* raxis, const double* t, const double* pts, const double* intr10);
/* Project::save: the document as indented JSON with sorted keys (what QJsonDocument::toJson writes) */
int sba_project_save(const sba_project* p, const char* path);

/* ---- Triangulation without any device allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  The same kernels as sba_triangulate (pose cache + batched multi-view DLT), stateless: no handle, every
 * argument except the sizes is a DEVICE pointer, the work is enqueued on `stream` (a cudaStream_t; NULL = the default stream)
 * and the call returns without synchronising.  It calls no cudaMalloc / cudaFree / allocating library routine: the camera
 * cache and the sanitised index arrays live in ONE file-scope `static __device__` array of SBA_NOALLOC_SCRATCH_BYTES that
 * every GPU loading the library holds until exit (csrc/sba_noalloc.cu).  Limits: n_cams <= SBA_NOALLOC_MAX_CAMS, n_points <=
 * SBA_NOALLOC_MAX_POINTS, n_views <= SBA_NOALLOC_MAX_VIEWS; beyond them SBA_ERR_INVALID with the limits in sba_create_error()
 * (the error text of handle-less calls) -- split the batch, or use sba_triangulate.
 *   raxis, t [3 n_cams] sensor -> world poses; cam_model [n_cams] or NULL (all cameras use model 0); intr10 [10 n_models];
 *   view_off [n_points + 1] (views of point i are view_off[i] .. view_off[i+1], n_views = view_off[n_points]), view_cam [n_views],
 *   view_uv [2 n_views]; xyz_out [3 n_points]; ok_out [n_points] or NULL.
 * A point with an offset, camera or camera-model index out of range gets xyz = 0, ok = 0 (nothing is read out of bounds).
 * The scratch is shared by all calls on a device and keeps the last call's contents: every region a call reads is rewritten by
 * that call first (nothing has to start at zero), calls on ONE stream are safe back to back, calls that overlap on the GPU
 * (different streams, no dependency) RACE.  The first call on a device looks the array's address up (cudaGetSymbolAddress):
 * make it outside a capture -- n_points = 0 does only that. */
#define SBA_NOALLOC_MAX_CAMS 4096
#define SBA_NOALLOC_MAX_POINTS 262144
#define SBA_NOALLOC_MAX_VIEWS 1048576
#define SBA_NOALLOC_SCRATCH_BYTES (6912 * 1024)
int sba_triangulate_noalloc(int32_t n_cams, const double* raxis, const double* t, const int32_t* cam_model, int32_t n_models,
                            const double* intr10, int32_t n_points, int32_t n_views, const int32_t* view_off,
                            const int32_t* view_cam, const double* view_uv, double* xyz_out, uint8_t* ok_out, void* stream);

const char* sba_version(void);

#ifdef __cplusplus
}
#endif
#endif
