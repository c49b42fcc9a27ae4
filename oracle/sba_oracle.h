/*
 * sba_oracle.h -- C interface of the CPU ORACLE for the SBA hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This is a dependency-free CPU restatement of the
 * arithmetic that french-paragon/steviapp performs on the path
 *   ModularSBASolver::opt_step -> ceres::Solve
 * (libs/sparsesolver/modularsbasolver.cpp:884-951) with the cost functors
 *   InvertPose<UV2ParametrizedXYZCost<PinholeUVProjector,Flags,1,2,3,2,2>,0>
 * (libs/sparsesolver/sbamodules/pinholecameraprojectormodule.cpp:18-19).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
 * arm may load it.  The product (libsba_b200.so) never links or calls it.
 *
 * PARITY STATUS: residual functors are pinned by the reference's own
 * known-answer vectors (tests/test_uv_projectors/main.cpp:37-186,
 * tests/test_sba_solver/main.cpp:61-236).  The LM / Schur / PCG part restates
 * Ceres Solver (un-vendored system dependency, version unpinned,
 * CMakeLists.txt:131) from its published algorithm: "parity unpinned" for the
 * LM iterates -- no reference test asserts any BA iterate or final cost.
 */
#ifndef SBA_ORACLE_H
#define SBA_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* intrinsics record per camera model: f, pp[2], k[3], t[2], B[2]  (10 doubles)
 * block order of pinholecameraprojectormodule.cpp:112-122 */
#define ORACLE_INTR_STRIDE 10

typedef struct {
    int32_t n_cams, n_pts, n_obs, n_models, n_priors;
    double* cam_r;                /* 3*n_cams, sensor->world axis-angle (PoseNode::rAxis, modularsbasolver.h:593-599) */
    double* cam_t;                /* 3*n_cams, sensor->world translation (PoseNode::t) */
    const int32_t* cam_model;     /* n_cams, index into intrinsics */
    const uint8_t* cam_fixed;     /* n_cams or NULL (SetParameterBlockConstant, modularsbasolver.cpp:822-846) */
    const uint8_t* cam_cost_kind; /* n_cams or NULL: 0 plane, 1 spherical (useSphericalCostPinholeCamProjCost) */
    double* pts;                  /* 3*n_pts (PositionNode::pos) */
    const uint8_t* pt_fixed;      /* n_pts or NULL */
    const double* intrinsics;     /* n_models*10 */
    const int32_t* obs_cam;       /* n_obs */
    const int32_t* obs_pt;        /* n_obs */
    const double* obs_uv;         /* 2*n_obs */
    const double* obs_info;       /* 4*n_obs row-major 2x2, or NULL = identity */
    const int32_t* prior_kind;    /* n_priors: 0 point, 1 camera rAxis, 2 camera t */
    const int32_t* prior_index;   /* n_priors */
    const double* prior_A;        /* 9*n_priors row-major (FixedSizeNormalPrior<3,3>) */
    const double* prior_b;        /* 3*n_priors */
} oracle_problem;

typedef struct {
    int32_t max_iterations;
    double function_tolerance, parameter_tolerance, gradient_tolerance;
    double initial_radius, max_radius, min_radius;
    double min_lm_diagonal, max_lm_diagonal;
    double min_relative_decrease;
    int32_t jacobi_scaling;
    double pcg_rel_tol;
    int32_t pcg_max_iters;
    int32_t linear_solver;     /* 0 = block-Jacobi PCG on implicit S, 1 = dense Cholesky of S (small scenes) */
    int32_t num_threads;       /* <=0: all */
} oracle_options;

#define ORACLE_MAX_HISTORY 512
typedef struct {
    int32_t status;            /* 0 no-convergence(max iters), 1 converged, <0 failure */
    int32_t iterations;        /* LM iterations executed (successful + unsuccessful) */
    int32_t successful_steps;
    int32_t pcg_iterations_total;
    double initial_cost, final_cost;
    double cost_history[ORACLE_MAX_HISTORY];   /* cost after iteration i (i=0: initial) */
    double radius_history[ORACLE_MAX_HISTORY];
    int32_t pcg_iters_history[ORACLE_MAX_HISTORY];
    double seconds_total, seconds_evaluate, seconds_linear;
} oracle_summary;

void oracle_default_options(oracle_options* o);

/* PinholeUVProjector::dirFromUV, pinholecameraprojectormodule.h:28-71 */
void oracle_dir_from_uv(const double* uv, const double* intr10, double* dir3);

/* forward model: numerically inverts dirFromUV (Newton to 1e-15) -- used by the
 * synthetic generators only. xn = normalised undistorted (P.x/P.z, P.y/P.z). */
int oracle_uv_from_dir(const double* xn2, const double* intr10, double* uv2);

/* R(r) x and RigidBodyTransform::inverse (LibStevi restatements) */
void oracle_angle_axis_rotate(const double* r, const double* x, double* out);
void oracle_rigid_inverse(const double* r, const double* t, double* ri, double* ti);

/* residuals (2*n_obs), jac_cam (n_obs*2*6 row-major: [d/dr(3) d/dt(3)]),
 * jac_pt (n_obs*2*3), all in the caller's observation order; any output may be
 * NULL.  cost = 1/2 sum r^2 including priors.  Returns 0, or 1 if a plane
 * observation has Pbar.z < 0 (functor returned false,
 * modularuvprojection.h:395-399); failed[] (n_obs, may be NULL) flags them. */
int oracle_evaluate(const oracle_problem* p, int num_threads, double* cost,
                    double* residuals, double* jac_cam, double* jac_pt,
                    uint8_t* failed);

/* Jacobian w.r.t. the five intrinsic blocks (n_obs*2*10) -- "next" row f1. */
void oracle_pushbroom_dir(const double* uv, double sensor_width, double f, double pp, const double* as6, const double* bs6, double* dir3);
int oracle_evaluate_intrinsics_jac(const oracle_problem* p, double* jac_intr);
/* PoseTransform<LeverArm<UVCost>> (free boresight + lever arm behind a constant offset; costfunctors/posedecoratorfunctors.h:148-264,500,
 * sbamodules/pinholecameraprojectormodule.cpp:147-166), n independent samples; returns the number of failed evaluations */
int oracle_eval_leverarm(int32_t n, const double* r, const double* t, const double* b, const double* l, const double* r_off,
                         const double* t_off, const double* pts, const double* intr10, const double* uv, const double* info,
                         int spherical, double* res, double* jac_cam, double* jac_la, double* jac_pt, uint8_t* failed);
/* PoseTransform<UVCost, SourceToInitial> (constant sensor offset; costfunctors/posedecoratorfunctors.h:628-700,
 * sbamodules/pinholecameraprojectormodule.cpp:197-214), n independent samples; returns the number of failed evaluations */
int oracle_eval_offset(int32_t n, const double* r, const double* t, const double* r_off, const double* t_off, const double* pts,
                       const double* intr10, const double* uv, const double* info, int spherical,
                       double* res, double* jac_cam, double* jac_pt, uint8_t* failed);

/* Ceres-style Levenberg-Marquardt; updates cam_r, cam_t, pts in place. */
int oracle_solve(oracle_problem* p, const oracle_options* o, oracle_summary* s);

/* gradient g = J^T r (camera part 6*n_cams, point part 3*n_pts) and the
 * block-diagonal of J^T J (Hcc 36*n_cams row-major, Hpp 9*n_pts). */
int oracle_normal_blocks(const oracle_problem* p, int num_threads, double* gc,
                         double* gp, double* Hcc, double* Hpp);

/* Implicit Schur operator y = S x for damping given directly as diagonal
 * lam_c (6*n_cams), lam_p (3*n_pts) (test hook for the PCG mat-vec). */
int oracle_schur_matvec(const oracle_problem* p, int num_threads,
                        const double* lam_c, const double* lam_p,
                        const double* x, double* y, double* rhs);

/* Multi-view DLT triangulation: for point i, views view_off[i]..view_off[i+1]
 * give camera index + uv; cost kind ignored (rays from dirFromUV).
 * Inhomogeneous normal equations sum (I - d d^T/|d|^2) over world-frame rays. */
int oracle_triangulate(const oracle_problem* cams, int32_t n_points,
                       const int32_t* view_off, const int32_t* view_cam,
                       const double* view_uv, int num_threads, double* xyz_out,
                       uint8_t* ok_out);

/* ---- extended problem class (sba_oracle_ext.cpp): free intrinsic blocks, sensor offsets / lever arms, UV <-> UV factors
 * under a robust loss, point covariances.  Dense reduced system: scenes of up to a few hundred cameras. ---- */
typedef struct {
    oracle_problem base;              /* cameras, points, projection observations, 3-D priors as above */
    double* intr_rw;                  /* NULL: intrinsics constant; else the same n_models*10 array as base.intrinsics, updated in place */
    const uint8_t* intr_fixed_mask;   /* n_models or NULL (= all constant): bit k set -> block k constant, k = 0 f, 1 pp, 2 k, 3 t, 4 B
                                         (SetParameterBlockConstant rules of pinholecameraprojectormodule.cpp:356-507) */
    int32_t n_intr_priors;            /* FixedSizeNormalPrior<n,n> on an intrinsic block (pinholecameraprojectormodule.cpp:327-345,365-378,...) */
    const int32_t* ip_model;          /* n_intr_priors */
    const int32_t* ip_block;          /* n_intr_priors: block index as above */
    const double* ip_A;               /* 9 per prior: dim x dim row-major in the leading entries */
    const double* ip_b;               /* 3 per prior */
    const double* cam_offset;         /* 6*n_cams (r_off, t_off): constant sensor -> body transform, or NULL (identity) */
    const int32_t* cam_mount;         /* n_cams or NULL: lever-arm block of the camera, -1 = none */
    int32_t n_mounts;
    double* mount_b;                  /* 3*n_mounts boresight (axis-angle), updated in place */
    double* mount_l;                  /* 3*n_mounts lever arm, updated in place */
    const uint8_t* mount_fixed;       /* n_mounts or NULL */
    int32_t n_pairs;                  /* UV <-> UV factors (UV2UVCost) */
    const int32_t* pair_cam;          /* 2*n_pairs */
    const double* pair_uv;            /* 4*n_pairs: uv1, uv2 */
    const double* pair_info;          /* 8*n_pairs: info1, info2 row-major 2x2, or NULL */
    int32_t loss_kind;                /* on the factors: 0 none, 1 Huber, 2 Cauchy, 3 Arctan (correspondencessetsbamodule.cpp:2459-2471) */
    double loss_a;                    /* scale (the reference passes 1) */
} oracle_ext_problem;

/* layout of the reduced system: n_ext extended cameras; first block of every model (-1: all its blocks constant), block of
 * every mounting (-1: constant), live flag of every column (6*n_ext) */
int oracle_ext_layout(const oracle_ext_problem* q, int32_t* n_ext, int32_t* model_block, int32_t* mount_block, uint8_t* live);
/* residuals 2*n_obs, jac_cam 12, jac_pt 6, jac_intr 20, jac_lever 12 per observation; pair_res 4, pair_jac 4 x 32 per factor
 * (raw, before the loss; columns pose1 6, pose2 6, intrinsics of camera 1 10, of camera 2 10); cost = 1/2 (sum r^2 + sum rho) */
int oracle_ext_evaluate(const oracle_ext_problem* q, double* cost, double* residuals, double* jac_cam, double* jac_pt,
                        double* jac_intr, double* jac_lever, double* pair_res, double* pair_jac);
/* dense reduced system (6 n_ext)^2 and rhs at the current point for a trust-region radius (<= 0: undamped) */
int oracle_ext_reduced(const oracle_ext_problem* q, const oracle_options* opt, double radius, double* S, double* rhs);
/* 3x3 marginal covariance of every point block (ModularSBASolver::std_step) */
int oracle_ext_point_covariance(const oracle_ext_problem* q, double* cov9);
int oracle_ext_solve(oracle_ext_problem* q, const oracle_options* o, oracle_summary* s);

int oracle_num_threads(void);
/* caps the OpenMP team of every later call at n threads (n <= 0: all online processors), whatever OMP_NUM_THREADS the
 * process inherited; returns the new maximum */
int oracle_set_max_threads(int n);

#ifdef __cplusplus
}
#endif
#endif
