// sba_device.h -- resident (HBM) data layout of one rank's problem shard and the
// launch wrappers of sba_kernels.cu.  See DESIGN.md "Data layout in HBM".
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace sba {

constexpr int TILE_SLOTS = 256;      // max observation slots per warp tile (multiple of 32)
constexpr int WARPS_PER_BLOCK = 8;   // tile kernels: 256 threads
constexpr int RED_BLOCK = 1024;      // single-block deterministic reductions
constexpr int HP = 12;               // doubles per heavy-window partial
constexpr int TP = 28;               // doubles per tile partial
constexpr int SBA_MAX_GROUPS = 12;   // free camera models + free mountings of one problem (shared blocks, sba_kernels_shared.cu)

// All arrays are device pointers.  "slot" = position in the camera-sorted,
// 32-aligned observation order (padding slots have obs_pt = -1 and all-zero J).
// "q slot" = position in the point-sorted, window-packed order.
//
// The Jacobian is stored compressed: for every observation
//   A = d res / d X (2x3)  and  C = d res / d rAxis (2x3);   d res / d t = -A exactly,
// so Jc = [C | -A], Jp = A.
//   camera order : SoA planes res[2], C[6], A[6] over slots   (coalesced warp-per-tile / thread-per-slot passes)
//   point order  : SoA planes qA[6] over q slots (thread-per-slot passes, segmented warp reductions)
struct DevProblem {
    int32_t Nc, Np, No, Ns, Nt, Nm, Nprior;
    int32_t cp_variant;   // camera pass (factored): gathers in flight per lane 0 -> 2 (default), 1 -> 4, 2 -> 1 (measurement knob)
    int32_t pp_variant;   // point pass: 0 = shared-memory operand table when it fits, 1 = always gather from global
    double center[3];     // fixed reference position (mean camera position at finalize) of the point-pass operand
    // cameras (replicated on every rank)
    double* cam_r;        // 3*Nc AoS (PoseNode::rAxis)
    double* cam_t;        // 3*Nc AoS
    double* cam_cache;    // 24*Nc, see sba_math.cuh
    int32_t* cam_model;   // Nc
    uint8_t* cam_fixed;   // Nc
    uint8_t* cam_kind;    // Nc (0 plane, 1 spherical)
    double* intr;         // 10*Nm
    // points (this rank's shard), 4 doubles per point (x,y,z,pad): one 32 B sector per gather
    double* pts;          // 4*Np
    uint8_t* pt_fixed;    // Np
    // observations, camera-sorted slots
    double* obs_uv;       // 2*Ns interleaved (u,v): one 128-bit load per observation
    int32_t* obs_pt;      // Ns, -1 = padding
    double* obs_info;     // 4 planes of Ns, or nullptr (identity stiffness)
    // warp tiles: consecutive slots of one camera, length multiple of 32, <= TILE_SLOTS
    int32_t* tile_cam;    // Nt
    int32_t* tile_start;  // Nt
    int32_t* tile_len;    // Nt
    int32_t* cam_tile_off;// Nc+1 (tiles of camera c are cam_tile_off[c]..cam_tile_off[c+1])
    // point order ("q slots"): observations grouped by point; a point with k <= 32 observations never
    // straddles a 32-slot window (padding slots have q_pt = -1), a point with k > 32 ("heavy") owns whole
    // windows.  One thread per q slot, warp-shuffle segmented reductions per point.
    int32_t Nq;           // padded q-slot count (multiple of 32)
    int32_t Nh;           // heavy points
    int32_t* pt_off;      // Np+1: CSR offsets in unpadded point order (k = pt_off[p+1]-pt_off[p])
    int32_t* pt_q0;       // Np: first q slot of the point
    int32_t* pt_heavy;    // Np: -1, or index into the heavy arrays
    int32_t* heavy_pt;    // Nh
    int32_t* heavy_w0;    // Nh+1: first entry of the point's windows in heavy_part
    double* heavy_part;   // 12 doubles per heavy window
    int32_t* q_pt;        // Nq, -1 = padding
    int32_t* q_cam;       // Nq
    double* q_uv;         // 2*Nq interleaved (static copy)
    double* q_info;       // 4*Nq AoS or nullptr
    // materialised linearisation
    double* res;          // 2 planes of Ns   } camera order
    double* Ac;           // 6 planes of Ns   }   (d res / d X)
    double* Pc;           // 3 planes of Ns: Pbar.  d res / d rAxis = C = A [R Pbar]x Q^T is rebuilt from A, Pbar and the camera
                          //   (rebuild_C, sba_kernel_util.cuh) instead of being stored
    double* Kc;           // 2 planes of Ns: Pi scale constants of spherical cameras, or nullptr (factored camera pass)
    int32_t cp_factored;  // 1: the camera pass of the implicit operator rebuilds Pi, A, C from Pbar alone (identity stiffness)
    double* qA;           // 6 planes of Nq: point order.  C is NOT kept in point order: C x_r = A ((X - t) x (R Jr x_r)),
                          // so the point pass needs only A, the point and a 9-double per-camera operand (cam_vec)
    // ---- explicit reduced camera system (operator_mode 0 auto / 3 forced; see sba_kernels_explicit.cu) ----
    // S is kept as dense 384x384 tiles (64 cameras x 64 cameras) over the UPPER block triangle: tile (I,J), I <= J, holds
    // the 6x6 blocks S_ij, i in I, j in J, i < j; diagonal tiles hold S_ii / 2 on their diagonal and zeros below, so
    // y_I += T x_J, y_J += T^T x_I reproduces the symmetric product with ONE read of every stored byte.
    int32_t ex_on;        // 1: PCG applies S from the tiles
    int32_t ex_NT;        // camera tiles per dimension = ceil(Nc / 64)
    int32_t ex_ntiles;    // allocated (non-empty) tiles
    int32_t ex_nblocks;   // off-diagonal blocks (i < j) with at least one common point
    int32_t ex_ncontrib;  // point-pair contributions: sum_p k_p (k_p - 1) / 2
    int32_t* ex_tile_of;  // ex_NT * ex_NT: tile index of (I,J), I <= J, or -1
    int32_t* ex_tile_I;   // ex_ntiles
    int32_t* ex_tile_J;   // ex_ntiles
    double* ex_S;         // row-major 384x384 tiles: all ex_ntiles of them (multi-GPU: this rank's PARTIAL sums, padded to a
                          // multiple of the rank count for the reduce-scatter)
    // multi-GPU: the summed tiles are sharded in contiguous ranges; rank r owns the global tiles ex_own0 .. ex_own0+ex_nown-1
    // in ex_Sown (local tile t = global tile - ex_own0) and applies only those.  Single GPU: ex_Sown = ex_S, all tiles.
    int32_t ex_own0, ex_nown;
    double* ex_Sown;
    int4* ex_rec;         // ex_ncontrib: {qa, qb, dest, point} in one 16-byte record (what the producer kernel reads)
    int32_t* ex_blk_i;    // ex_nblocks
    int32_t* ex_blk_j;    // ex_nblocks
    int32_t* ex_blk_off;  // ex_nblocks: first contribution of the block in the block-sorted buffer
    int32_t* ex_blk_len;  // ex_nblocks: run length; the block list is ordered by run length, longest first
    // direct mode (single GPU, block-sparse panels, no shared blocks / factors): the assembly and the diagonal kernel write the
    // panels of the persistent PCG kernel themselves; the dense tiles are brought up to date on demand (decompaction)
    int32_t ex_direct;
    int64_t* ex_blk_dst;  // ex_nblocks: offset (doubles) in sp_data of element (0, 0) of the block inside its panel
    int32_t* ex_blk_m6;   // ex_nblocks: row stride of that panel
    int64_t* ex_dg_dst;   // Nc(e): same for the diagonal block of every camera
    int32_t* ex_dg_m6;
    double* ex_contrib;   // 12 * ex_ncontrib: compact contributions (K 3x3, X), block-sorted; expanded to D_i K D_j^T by the assembly
    double* ex_row;       // ex_nown * 384: row partials of the tile mat-vec (y_I += T x_J), local tile order
    double* ex_col;       // ex_nown * 6 * 384: column partials per 64-row chunk (y_J += T^T x_I); = ex_row + ex_nown * 384
    int32_t* ex_fold_ptr; // ex_NT + 1
    int32_t* ex_fold_off; // offsets (doubles, relative to ex_row) of the partial vectors of every camera tile, fixed order
    double* cam_Q;        // 12*Nc: Q = (R Jr)^T row-major (9) + pad, written by K0: C = A [X - t]x Q^T
    double* dlt_rec;      // 8*Nc: quaternion of Rt + Rt t per camera, written by launch_triangulate for its kernel (sba_math.cuh dlt_record)
    double* intr_r;       // 2*Nm: 1 / f and 1 / (f + t1) of every camera model, refreshed with the pose cache (k_intr_recip)
    // ---- peer-memory all-reduce of the PCG operator partial (multi-GPU; sba_kernels_pcg.cu, PM_XCHG) ----
    // every rank owns a mailbox, mapped into every process (CUDA IPC, stores travel over NVLink / NVSwitch):
    //   double   data[2][xchg_n][xchg_stride]     (parity of the exchange number, source rank, vector element)
    //   uint32_t flag[2][xchg_n][xchg_nflag]      (= exchange number once the source rank's block has pushed its slice)
    int32_t xchg_n;       // ranks (0 = exchange through NCCL instead)
    int32_t xchg_rank;
    int32_t xchg_stride;  // doubles per (parity, source) buffer = pcg_blocks * 192
    int64_t xchg_ll_off;  // doubles from the mailbox base to the low-latency region of the persistent PCG kernel:
                          //   uint4 ll[2][xchg_n][ps_nitems][128] = {lo, seq, hi, seq} per value (flag travels with the data)
    int32_t xchg_nflag;   // flags per (parity, source) = pcg_blocks
    double* xchg_peer[8]; // mailbox base of every rank (own entry = local pointer)
    // multi-GPU build without ncclReduceScatter: the partial tiles of every rank are mapped into every process (CUDA IPC) and
    // the owner of a slice PULLS it from all ranks over NVLink, summing in rank order (k_explicit_pull_reduce)
    int32_t ex_pull;
    double* ex_peer_S[8]; // ex_S of every rank (own entry = local pointer)
    uint32_t* xchg_ctr;   // local: number of exchanges completed so far (identical on every rank)
    // ---- persistent PCG kernel (sba_kernels_pcgsolve.cu): static schedule of the tile mat-vec and of the fold ----
    int32_t ps_on;        // 1: run_pcg launches k_pcg_solve once per LM iteration
    int32_t ps_grid;      // resident CTAs (1 per SM)
    int32_t ps_nseg;      // segments = runs of 8-row groups inside one tile
    int32_t ps_nitems;    // items of 96 vector elements (16 cameras), item b owned by CTA b
    int32_t* ps_seg_ptr;  // ps_grid + 1: segments of CTA b
    int32_t* ps_seg;      // 4 per segment: local tile, first 8-row group, end group, pad
    double* ps_part;      // ex_nown * 384 row partials (local tile, row), then ps_nseg * 384 column partials (segment, column)
    int32_t* ps_fold_ptr; // ex_NT + 1
    int32_t* ps_fold_off; // per camera tile K: offsets into ps_part of the partial vectors that contribute to K, fixed order
    double* ps_pq;        // ps_grid: per-CTA part of p^T S p / 2
    double* ps_rz;        // ps_nitems: per-item part of r.z
    double* ps_clock;     // 4: ns CTA 0 spent in (mat-vec, barrier 1, fold + vector part, barrier 2) during the last solve
    double* pcg_p2;       // 6*Nc: second copy of p (directions alternate between pcg_p and pcg_p2)
    // ---- block-sparse copy of the tiles for the persistent PCG kernel (single GPU; sba_kernels_pcgsolve.cu, SPARSE) ----
    // 31 % of the 6x6 blocks of the Venice-shape S are structurally zero.  After every build the dense tiles are compacted into
    // PANELS: per (tile, block row) the non-zero block columns only, 6 rows x m6 doubles row-major, plus a 16-bit map
    // panel column -> tile column; the CG iterations stream the panels instead of the dense rows.
    int32_t sp_on;
    int32_t sp_npanel;    // non-empty block rows
    int32_t sp_nseg;      // segments = runs of panels of one CTA inside one tile
    int32_t sp_nchunk;
    int64_t sp_bytes;     // bytes of the panel stream (what one CG iteration reads)
    int64_t sp_nblocks;   // 6x6 blocks kept (structurally non-zero, upper triangle + diagonal)
    int4* sp_chunk;       // per chunk = consecutive panels of a tile that travel as one bulk copy: x stream offset / 16, y bytes, z directory offset / 16 in sp_idx, w directory bytes
    int4* sp_panel;       // per panel: x stream offset and y map offset in 16-byte units, z m6 = 6 * non-zero blocks, w tile * 64 + block row
    double* sp_data;      // the stream the CG iterations read, chunk after chunk: directory {panels, first panel mod 8, offset / 128 of every panel}, then per
                          // panel a 16-byte header {first row in the tile, m6}, 6 rows x m6 doubles, the column map; everything on 128-byte boundaries
    uint16_t* sp_idx;     // static part in the same order: directories, headers and column maps
    int32_t* sp_seg_ptr;  // ps_grid + 1
    int32_t* sp_seg;      // 4 per segment: local tile, first chunk, end chunk, pad
    double* sp_part;      // ex_nown * 384 row partials, then sp_nseg * 384 column partials
    int32_t* sp_fold_ptr; // ex_NT + 1
    int32_t* sp_fold_off;
    // priors, grouped by target: group g covers priors pg_off[g]..pg_off[g+1]
    int32_t Ngroup;
    int32_t* pg_off;      // Ngroup+1
    int32_t* pg_kind;     // Ngroup (0 point, 1 cam r, 2 cam t)
    int32_t* pg_index;    // Ngroup
    double* prior_A;      // 9*Nprior (sorted by group)
    double* prior_b;      // 3*Nprior
    double* prior_res;    // 3*Nprior
    int32_t Nlonely;      // free points that carry priors but no observation (a ground-control landmark no image sees): they head
    int32_t* lonely_pt;   //   no q slot, so the back-substitution handles them in a list kernel (dp = -Hinv gp)
    // ---- shared parameter blocks: free intrinsic blocks per camera model and free boresight / lever-arm
    // blocks per mounting.  They enter the reduced camera system as PSEUDO-CAMERAS (6 columns each) behind the real ones:
    // extended cameras 0 .. Nce-1, every per-camera array below is allocated for Nce entries, and the tile / PCG kernels are
    // launched on a copy of this struct with Nc = Nce.  A GROUP = one free model (8 columns f ppx ppy k1 k2 k3 | t1 t2, two
    // pseudo-cameras, the second one with four dead columns) or one free mounting (6 columns b | l, one pseudo-camera);
    // internally every group carries 8 columns.  See sba_kernels_shared.cu.
    int32_t Nce;          // extended cameras (= Nc without shared blocks)
    int32_t Ng;           // groups
    int32_t Nmt;          // mountings
    int32_t grp_kind[SBA_MAX_GROUPS];    // 0 camera model, 1 mounting          } small tables carried in the struct itself
    int32_t grp_index[SBA_MAX_GROUPS];   // model / mounting                    } (kernel parameter space), also read by
    int32_t grp_blk[SBA_MAX_GROUPS];     // first pseudo-camera of the group    } the launch wrappers on the host
    uint8_t grp_live[8 * SBA_MAX_GROUPS];   // column is free (constant blocks and padding are dead: unit diagonal, zero rhs)
    int32_t* cam_grp;     // 2*Nc: group of the camera's model, of its mounting (-1: none / constant)
    int32_t* cam_mount;   // Nc: mounting of the camera or -1 (nullptr: no mountings)
    double* cam_off;      // 12*Nc: constant sensor -> body transform per camera, Roff^T row-major + toff (nullptr: identity)
    double* mount;        // 6*Nmt: boresight b (axis-angle) | lever arm l
    double* mount_cache;  // 24*Nmt: pose_cache(b, l)
    double* mount_saved;  // 6*Nmt
    double* intr_saved;   // 10*Nm
    double* sh_W;         // [Ng][Np][24]: W = sum_o A_o^T G_o (3 x 8 row-major), G = d res / d group columns
    double* sh_Y;         // [Ng][Np][24]: (Hpp + D)^-1 W
    double* sh_heavy;     // 24 per heavy window (points with more than 32 observations)
    double* sh_H;         // [Ng][Ng][64]: sum_o G_g^T G_g' (undamped; + priors on the diagonal groups)  } all-reduced
    double* sh_g;         // [Ng][8]: sum_o G_g^T res (+ priors)                                           } together
    double* sh_acc;       // per LM iteration, all-reduced: border [Ng][Nc][48] | SS [Ng][Ng][64] = sum_p W_g^T Y_g' | rS [Ng][8] = sum_p W_g^T Hinv gp
    double* sh_gd;        // 2 planes of Ns: G_o delta of the current step (model cost change)
    double* sh_part;      // per-block partials of the group reductions (72 doubles per block)
    int32_t sh_nblk;      // blocks of the persistent point-reduction kernels
    int32_t Nipr;         // priors on intrinsic blocks (FixedSizeNormalPrior<n,n>, pinholecameraprojectormodule.cpp:327-345 ...)
    int32_t* ipr_model;   // Nipr
    int32_t* ipr_block;   // Nipr: 0 f, 1 pp, 2 k, 3 t, 4 B
    double* ipr_A;        // 9*Nipr: dim x dim row-major in the leading entries
    double* ipr_b;        // 3*Nipr
    double* ipr_res;      // 3*Nipr
    // ---- UV <-> UV cross-projection factors between two poses (sba_kernels_pairs.cu), replicated on every rank ----
    int32_t Npair;
    int32_t pair_loss;    // 0 none, 1 Huber, 2 Cauchy, 3 Arctan
    double pair_loss_a;
    int32_t* pair_cam;    // 2*Npair
    double* pair_uv;      // 4*Npair: uv1, uv2
    double* pair_info;    // 8*Npair: info1, info2 (2x2 row-major) or nullptr
    double* pair_rho;     // Npair: rho(|r|^2)
    double* pair_r;       // 4*Npair: corrected residuals
    double* pair_J;       // 48*Npair: corrected Jacobians 4 x 12 (pose 1 | pose 2), zero at constant cameras
    double* pair_Ji;      // 64*Npair: 4 x 16, the 8 intrinsic columns of camera 1's model | of camera 2's (nullptr: no free model)
    double* pair_border;  // [Ng][Nc][48]: sum over a camera's factors of Jc^T Ji (added to the border of the shared blocks)
    int32_t* pair_inc_off;// Nc+1: incidence list of every camera ...
    int32_t* pair_inc;    // ... entries 2 * factor + side
    int32_t Npblk;        // distinct camera pairs (i < j)
    int32_t* pblk_i; int32_t* pblk_j; int32_t* pblk_off;   // Npblk (+1): factors of the pair ...
    int32_t* pblk_list;   // ... entries 2 * factor + (1: the factor's SECOND camera is i)
    uint8_t* pblk_assembled;   // Npblk: the block also receives point contributions (add) or not (assign)
    // normal-equation blocks
    double* Hpp;          // 8*Np AoS: packed symmetric 3x3 (6) + pad, undamped
    double* gp;           // 4*Np AoS
    double* Hinv;         // 8*Np AoS: damped inverse, packed symmetric
    double* Hcc;          // 21*Nc AoS packed symmetric upper, undamped   } contiguous: one
    double* gc;           // 6*Nc AoS                                     } all-reduce payload
    double* Sd;           // 21*Nc AoS: sum_o Jc^T (Jp Hinv Jp^T) Jc  (partial on each rank)
    double* Minv;         // 36*Nc AoS: inverse of the diagonal blocks of S (SCHUR_JACOBI)
    double* lam_c;        // 6*Nc
    double* lam_p;        // 4*Np AoS
    double* scale_c;      // 6*Nc
    double* scale_p;      // 4*Np AoS
    // work vectors
    double* tile_part;    // TP*Nt tile partials
    double* mv_part;      // 6 per tile: camera-pass partial of the tile, per-camera factors already applied
    double* cam_vec;      // 8*Nc: operand of the point pass, (a_c = R Jr x_r, g_c = x_t + (t - center) x a_c, pad) for the current x
    double* tile_cost;    // Nt
    double* y_p;          // 4*Np AoS point-side temporary of the implicit operator
    double* dp;           // 4*Np AoS step of the points
    double* pts_saved;    // 4*Np
    double* cam_saved;    // 6*Nc
    double* xz;           // 6*Nc: input vector with fixed cameras zeroed
    double* pcg_x, *pcg_r, *pcg_z, *pcg_p, *pcg_q, *rhs;   // 6*Nc each
    double* pcg_part;     // 2 * pcg_blocks per-block partials of p.q and r.z
    double* block_part;   // 4*NB per-block partial sums for scalar reductions over points
    int32_t NB;           // number of point blocks (ceil(Np/256))
    double* scalars;      // SC_N doubles
    int32_t* flags;       // FL_N ints
};

enum { SC_COST = 0, SC_RZ = 1, SC_RZ0 = 2, SC_PQ = 3, SC_MODEL = 4, SC_STEP2 = 5, SC_X2 = 6, SC_GMAX = 7,
       SC_COST_OBS = 8, SC_COST_PRIOR = 9, SC_MODEL_PRIOR = 10, SC_COST_PRIOR_CAM = 11, SC_MODEL_PRIOR_CAM = 12,
       SC_STEP2_CAM = 13, SC_X2_CAM = 14, SC_GMAX_CAM = 15, SC_RZ_NEXT = 16, SC_N = 32 };
enum { FL_CHEIRALITY = 0, FL_SINGULAR_PT = 1, FL_NOT_PD = 2, FL_PCG_ITERS = 3, FL_PCG_ITERS_NEXT = 4, FL_XCHG_TIMEOUT = 5, FL_XCHG_DEAD = 6, FL_N = 8 };

// ---- launch wrappers (sba_kernels.cu); all asynchronous on `st` ----
void launch_pose_cache(const DevProblem& P, cudaStream_t st);
void launch_residual_jacobian(const DevProblem& P, bool jac, cudaStream_t st);   // K1 / K1c, camera order
void launch_linearize_points(const DevProblem& P, cudaStream_t st);              // K1p+K2a: point-order A planes, Hpp, gp
void launch_prior_eval(const DevProblem& P, int store, cudaStream_t st);         // prior residuals (stored if store) + their cost
void launch_pack_tail(const DevProblem& P, double* tail, cudaStream_t st);
void launch_cost_reduce(const DevProblem& P, cudaStream_t st);                   // tile costs -> SC_COST_OBS
void launch_camera_blocks(const DevProblem& P, cudaStream_t st);                 // K2b: tile partials + per-camera reduce -> Hcc, gc
void launch_preq_finish(const DevProblem& P, double* H21, double* g6, cudaStream_t st);   // per camera: T (.) T^T, T (.) with T = diag(Q, I)
void launch_prior_blocks(const DevProblem& P, int cameras, cudaStream_t st);     // add A^T A, A^T r of the priors
void launch_scaling(const DevProblem& P, int enabled, cudaStream_t st);
void launch_damping(const DevProblem& P, double radius, double dmin, double dmax, cudaStream_t st);
void launch_invert_points(const DevProblem& P, cudaStream_t st);
void launch_schur_diag(const DevProblem& P, cudaStream_t st);                    // -> Sd (partial sum of this rank)
void launch_schur_diag_invert(const DevProblem& P, cudaStream_t st);             // Minv = (Hcc + lam - Sd)^-1
// implicit operator pieces.  mode 0: y_p = Hinv * sum Jp^T Jc x ; mode 1: y_p = Hinv gp ;
// mode 2: dp = -(Hinv (gp + sum Jp^T Jc x))
// table_ready: cam_vec already holds the operand of x (the PCG kernels refresh it whenever they update p)
void launch_point_pass(const DevProblem& P, const double* x, int mode, cudaStream_t st, bool table_ready = false);
void launch_cam_vec(const DevProblem& P, const double* x, cudaStream_t st);
// out = -sum_o Jc^T Jp y_p (fixed cameras: 0); if finish: out += hsign * (Hcc x) + damp * lam_c .* x + base.
// Multi-GPU runs all-reduce the partial first and call launch_camera_finish afterwards.
void launch_camera_pass(const DevProblem& P, const double* x, const double* base, double hsign, int damp, int finish,
                        double* out, cudaStream_t st);
void launch_camera_finish(const DevProblem& P, const double* x, const double* base, double hsign, int damp, double* out,
                          cudaStream_t st);
void launch_zero_fixed(const DevProblem& P, const double* x, double* xz, cudaStream_t st);   // xz = x, 0 at fixed cameras
void launch_negate(double* x, int n, cudaStream_t st);
void launch_pcg_init(const DevProblem& P, cudaStream_t st);       // x=0, r=rhs, z=Minv r, p=z, rz, rz0
int pcg_blocks(const DevProblem& P);
// PCG step "mode": where the local part of q = S p comes from, whether it is all-reduced through the peer mailboxes
// inside the kernel, and whether (Hcc + lam) p still has to be added (implicit operator; the explicit tiles contain it)
enum { PM_SRC_TILES = 0,   // - sum of the camera-pass tile partials (mv_part)
       PM_SRC_Q = 1,       // pcg_q as it stands (all-reduced by NCCL)
       PM_SRC_FOLD = 2,    // fold of the explicit mat-vec partials of this rank's tiles
       PM_SRC_MASK = 3,
       PM_XCHG = 4,        // sum the local part over the ranks through the peer mailboxes (fused kernel only)
       PM_HCC = 8 };
bool launch_pcg_fused(const DevProblem& P, int mode, double tol2, int max_iters, cudaStream_t st);   // A+B+C in one cooperative launch
bool pcg_fused_available(const DevProblem& P);
int pcg_solve_grid();                                                   // resident CTAs of the persistent PCG kernel (0 = unavailable)
int pcg_solve_items(int n_cams);
int pcg_solve_slot();                                                   // mailbox doubles per item
bool launch_pcg_solve(const DevProblem& P, double tol2, int max_iters, cudaStream_t st);   // whole PCG solve, one cooperative launch
int pcg_xchg_nflag(const DevProblem& P);
int pcg_xchg_stride(const DevProblem& P);                               // doubles per (parity, source) mailbox buffer
size_t pcg_xchg_bytes(const DevProblem& P, int nranks);                 // mailbox size of the fused-kernel exchange (data + flags)
size_t pcg_xchg_ll_bytes(const DevProblem& P, int nranks);              // + low-latency region of the persistent kernel
void launch_pcg_A(const DevProblem& P, int mode, cudaStream_t st);    // q (from tile partials or all-reduced) + p.q partials
void launch_pcg_BC(const DevProblem& P, double tol2, int max_iters, cudaStream_t st);   // alpha, x, r, z | beta, p (frozen once converged)
void launch_camera_pass_tiles(const DevProblem& P, cudaStream_t st);   // operator pass 2, tile partials only
void launch_model_change(const DevProblem& P, const double* dc, cudaStream_t st);   // SC_MODEL, SC_STEP2, SC_X2 (local part)
void launch_apply_step(const DevProblem& P, const double* dc, cudaStream_t st);     // save + x += delta
void launch_restore(const DevProblem& P, cudaStream_t st);
void launch_grad_max(const DevProblem& P, cudaStream_t st);
void launch_gather_outputs(const DevProblem& P, const int32_t* slot_of_obs, double* res_out, double* jc_out,
                           double* jp_out, cudaStream_t st);
void launch_intrinsics_jacobian(const DevProblem& P, const int32_t* slot_of_obs, double* jac_out, cudaStream_t st);   // 2 x 10 per observation
void launch_evaluate_offset(const DevProblem& P, const int32_t* slot_of_obs, const double* off, double* res_out, double* jc_out,
                            double* jp_out, int* n_failed, cudaStream_t st);   // constant sensor offset per camera (12 doubles)
void launch_unpack_blocks(const DevProblem& P, double* Hcc36, double* Hpp9, double* gp3, cudaStream_t st);
void launch_set_lambda(const DevProblem& P, const double* lam_c, const double* lam_p3, cudaStream_t st);
void launch_pack_points(const double* xyz3, double* pts4, int n, cudaStream_t st);
void launch_unpack_points(const double* pts4, double* xyz3, int n, cudaStream_t st);
// explicit reduced camera system
void launch_explicit_build(const DevProblem& P, cudaStream_t st);     // contributions -> off-diagonal blocks of ex_S (needs Hinv of this LM iteration)
void launch_explicit_sparse_dirs(const DevProblem& P, cudaStream_t st);   // chunk directories -> panel stream (once, at finalize)
void launch_explicit_pull_reduce(const DevProblem& P, int nranks, int rank, size_t slice_doubles, cudaStream_t st);   // ex_Sown = sum over the ranks of their slice of ex_S
void launch_explicit_decompact(const DevProblem& P, cudaStream_t st);   // panels -> dense tiles (direct mode: for the hooks and the dense factorisation)
void launch_explicit_compact(const DevProblem& P, cudaStream_t st);   // dense tiles -> block-sparse panels (sp_on)
void launch_explicit_diag(const DevProblem& P, cudaStream_t st);      // diagonal blocks of the owned tiles (needs Hcc, Sd, lam_c; multi-GPU: after the reduce-scatter)
void launch_explicit_matvec(const DevProblem& P, const double* x, cudaStream_t st, int pcg = 0, double tol2 = 0.0,
                            int max_iters = 0);   // tile partials ex_row / ex_col of S x (pcg: skip once the PCG is frozen)
void launch_explicit_gather(const DevProblem& P, double* y, cudaStream_t st);         // y = sum of the partials (test hook; PCG folds them in k_pcg_A)
void launch_triangulate(const DevProblem& P, int n_points, const int32_t* view_off, const int32_t* view_cam,
                        const double* view_uv, double* xyz_out, uint8_t* ok_out, cudaStream_t st);
// dense Cholesky of the explicit reduced system: direct solve and point covariances (sba_kernels_dense.cu)
void launch_dense_from_tiles(const DevProblem& P, double* D, int n, cudaStream_t st);   // P with Nc = extended cameras; n = padded side (multiple of 384)
int dense_cholesky(double* D, int n, double* Linv_all, double* work, int* flag, cudaStream_t st);   // in place, lower factor; returns launches
int dense_solve(const double* L, int n, const double* Linv_all, double* x, cudaStream_t st);
int dense_inverse(const double* L, int n, const double* Linv_all, double* Z, double* Sinv, double* work, cudaStream_t st);
void launch_point_covariance(const DevProblem& P, const double* Sinv, int n, double* cov9, cudaStream_t st);
// UV <-> UV factors (sba_kernels_pairs.cu)
void launch_pair_eval(const DevProblem& P, int store, cudaStream_t st);          // residuals / corrected Jacobians (store) and the cost, added to SC_COST_PRIOR_CAM
void launch_pair_camera_blocks(const DevProblem& P, cudaStream_t st);           // Hcc, gc (+ sh_H, sh_g, pair_border with free intrinsics)
void launch_pair_blocks(const DevProblem& P, cudaStream_t st);                  // J_i^T J_j into the tiles (after the assembly)
void launch_pair_model(const DevProblem& P, const double* x, cudaStream_t st);  // model cost change, added to SC_MODEL_PRIOR_CAM
void launch_pair_raw(const DevProblem& P, double* res, double* jac, cudaStream_t st);   // inspection hook
// shared parameter blocks (sba_kernels_shared.cu)
int shared_reduce_blocks();                                               // blocks of the persistent point-reduction kernels
void launch_shared_linearize(const DevProblem& P, cudaStream_t st);       // per linearisation: W per group, sh_H / sh_g partial sums (this rank)
void launch_shared_lin_finish(const DevProblem& P, cudaStream_t st);      // priors into sh_H / sh_g, Hcc / gc of the pseudo-cameras
void launch_shared_build(const DevProblem& P, cudaStream_t st);           // per LM iteration: Y = Hinv W, border / SS / rS partial sums into sh_acc
void launch_shared_build_finish(const DevProblem& P, int write_tiles, cudaStream_t st);   // Sd, rhs of the pseudo-cameras; border / cross blocks into the tiles
void launch_shared_backsub(const DevProblem& P, const double* x, cudaStream_t st);        // dp -= sum_g Y_g delta_g; G_o delta planes (sh_gd)
void launch_shared_stats(const DevProblem& P, const double* x, cudaStream_t st);          // step / parameter norms and prior model change of the shared blocks (adds to the replicated scalars)
void launch_shared_apply(const DevProblem& P, const double* x, cudaStream_t st);          // save + update intrinsics / mountings
void launch_shared_restore(const DevProblem& P, cudaStream_t st);
void launch_shared_prior_eval(const DevProblem& P, int store, cudaStream_t st);           // intrinsics priors: residuals (stored if store), cost added to SC_COST_PRIOR_CAM

}  // namespace sba
