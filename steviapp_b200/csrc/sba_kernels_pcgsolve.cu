// sba_kernels_pcgsolve.cu -- the WHOLE reduced-system solve in one persistent cooperative kernel.
//
// Same SCHUR_JACOBI preconditioned CG as sba_kernels_pcg.cu (the loop ceres::Solve runs for ITERATIVE_SCHUR behind
// modularsbasolver.cpp:938; in-tree analogue modules/g2o/g2o/solvers/pcg/linear_solver_pcg.hpp:84-165) on the explicit
// tiles of sba_kernels_explicit.cu, but the ~135 iterations of a Venice-scale LM step no longer cost two kernel
// launches, five grid-wide dependencies and a host poll each:
//
//   * one launch per LM iteration; one CTA per SM stays resident and loops over the CG iterations;
//   * the tile mat-vec is split STATICALLY into equal runs of 8-row groups (one run per CTA, cut into segments at
//     tile boundaries), so all CTAs finish together -- no wave quantisation, and a rank that owns only a few tiles
//     (8 GPUs: 51) still fills every SM;
//   * a CTA's run is ONE contiguous byte range of the tile array: a producer warp streams it with 12 KB bulk
//     copies (cp.async.bulk -> shared memory, mbarrier complete_tx) through a 16-stage / 192 KB ring, so the bytes in
//     flight per SM are set by shared memory, not by registers; the 8 consumer warps each own two stages (a private
//     double buffer of 4-row half groups) and do the two FMAs per element from shared memory;
//   * p.q = p^T S p is accumulated by the mat-vec itself (2 sum_rows x_I[r] (T x_J)[r]), so alpha needs no second
//     grid-wide reduction: TWO grid barriers per CG iteration (after the mat-vec, after r.z);
//   * p is never re-broadcast: the mat-vec rebuilds p = z + beta p_old on the fly (same expression, same bits as the
//     owner's copy), so there is no barrier between the direction update and the next mat-vec;
//   * x, r, z, p and the preconditioner rows of an element live in the registers of its owner thread for the whole
//     solve (one "item" of 96 elements = 16 cameras per CTA);
//   * multi-GPU: the item's 96 partial q values and the rank's p.q partial go to the peers' mailboxes with plain
//     NVLink stores and are summed in rank order (see peer exchange in sba_kernels_pcg.cu) -- inside the same kernel.
//
// Determinism: static schedule, fixed-shape trees, fixed list orders; no floating-point atomics.  All ranks compute
// bitwise identical scalars, so the loop exits are uniform across CTAs and ranks.
#include <cooperative_groups.h>

#include <algorithm>
#include <cstdlib>

#include "sba_kernel_util.cuh"

namespace sba {

namespace cg = cooperative_groups;

constexpr int PS_CWARPS = 8;                    // consumer warps
constexpr int PS_THREADS = 32 * (PS_CWARPS + 1);   // + one producer warp
constexpr int PS_ITEM = 96;        // vector elements per item (16 cameras; divides the 384 of a camera tile)
constexpr int PS_CHAINS = PS_THREADS / PS_ITEM;   // 3 fold chains per element: all 288 threads fold
constexpr int PS_SLOT = 128;       // mailbox doubles per item: 96 q values + the p.q partial
constexpr int PS_STAGES = 2 * PS_CWARPS;        // ring stages, two per consumer warp
constexpr int PS_HALF_ROWS = 4;                 // rows per stage
constexpr uint32_t PS_STAGE_BYTES = PS_HALF_ROWS * EX_TD * sizeof(double);   // 12288
static_assert(PS_THREADS == PS_CHAINS * PS_ITEM, "fold roles cover the CTA");
// block-sparse variant (SP): the unit of work is a PANEL = the non-zero 6x6 blocks of one block row of a tile: 16-byte
// header {first row, m6}, 6 rows x m6 doubles (m6 <= 384), the 16-bit tile column of every panel column.  The unit of
// transfer is a CHUNK = directory + consecutive panels of one tile, <= 40 KB, ONE bulk copy (the producer lane needs
// ~0.3 us per copy whatever its size: with one copy per panel it was the bound).  Chunks have very different sizes, so the
// ring is a BYTE ring: the producer places chunk after chunk contiguously (wrapping early when one does not fit at the
// end), one mbarrier pair per in-flight chunk (SP_SLOTS of them).  EVERY consumer warp visits every chunk in order (the
// panels of a chunk are dealt round-robin to the warps, by panel number mod 8) and arrives on its empty barrier (count 8):
// no warp can be more than one phase ahead of a slot's barrier, so the parity waits are unambiguous.
constexpr int SP_SLOTS = PS_STAGES;
constexpr uint32_t SP_RING_BYTES = PS_STAGES * PS_STAGE_BYTES;
constexpr uint32_t SP_CHUNK_MAX = 65536;   // upper bound of what sba_solver_explicit.inc builds (default 40 KB)
static_assert(SP_RING_BYTES >= 3 * SP_CHUNK_MAX, "three maximal chunks fit the ring");

struct PsSmem {
    double ring[PS_STAGES][PS_HALF_ROWS * EX_TD];   // 192 KB
    double col[PS_CWARPS][EX_TD];                   // column partials of the consumer warps of one segment
    double xI[EX_TD];                               // p of the segment's row tile
    double xJ[EX_TD];                               // p of the segment's column tile (SP only: gathered through the column map)
    double v[PS_CHAINS][PS_ITEM];                   // chain sums of the fold / r of the item
    double sh[33];
    double pq;
    double pqw[PS_CWARPS];
    unsigned long long full[PS_STAGES], empty[PS_STAGES];
    unsigned int poff[SP_SLOTS];                    // SP: byte offset of the slot's chunk in the ring
};

__device__ __forceinline__ uint32_t ps_saddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void ps_mbar_init(unsigned long long* b, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(ps_saddr(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void ps_mbar_expect_tx(unsigned long long* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(ps_saddr(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void ps_mbar_arrive(unsigned long long* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(ps_saddr(b)) : "memory");
}
__device__ __forceinline__ void ps_mbar_wait(unsigned long long* b, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" :: "r"(ps_saddr(b)), "r"(parity) : "memory");
}
// 1-D bulk copy global -> shared (TMA engine, no tensor map: the source run is contiguous), completion on the mbarrier
__device__ __forceinline__ void ps_bulk_load(void* dst, const void* src, uint32_t bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(ps_saddr(dst)), "l"(src), "r"(bytes), "r"(ps_saddr(bar)) : "memory");
}
// same with an L2 eviction-priority hint (createpolicy): evict_last for the part of the run that is to stay L2-resident
// across the CG iterations, evict_first for the part that only streams through
__device__ __forceinline__ void ps_bulk_load_hint(void* dst, const void* src, uint32_t bytes, unsigned long long* bar, uint64_t policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 :: "r"(ps_saddr(dst)), "l"(src), "r"(bytes), "r"(ps_saddr(bar)), "l"(policy) : "memory");
}
__device__ __forceinline__ long long ps_clock_now() {   // ordered against memory operations (measurement only)
    long long t;
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) :: "memory");
    return t;
}
__device__ __forceinline__ void ps_consumer_sync() { asm volatile("bar.sync 1, %0;" :: "n"(32 * PS_CWARPS) : "memory"); }

__device__ __forceinline__ uint32_t ps_ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void ps_st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
// p = z + beta p_old, the ONE expression used by the owners and by every mat-vec CTA (no contraction: same bits)
__device__ __forceinline__ double ps_dir(double z, double beta, double p_old) { return __dadd_rn(z, __dmul_rn(beta, p_old)); }

// fold of a per-CTA partial array (written by other CTAs of this launch: L1 bypassed); same value in every thread
__device__ __forceinline__ double ps_fold(const double* part, int n, double* sh) {
    double v = 0.0;
    for (int k = threadIdx.x; k < n; k += PS_THREADS) v += __ldcg(part + k);
    return block_sum(v, sh);
}

__device__ __forceinline__ double ps_warp_sum8(const double (&v)[8], int lane) {
    const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
    double r[4];
#pragma unroll
    for (int i = 0; i < 4; i++) r[i] = (b4 ? v[4 + i] : v[i]) + __shfl_xor_sync(FULL, b4 ? v[i] : v[4 + i], 16);
    double p[2];
#pragma unroll
    for (int i = 0; i < 2; i++) p[i] = (b3 ? r[2 + i] : r[i]) + __shfl_xor_sync(FULL, b3 ? r[i] : r[2 + i], 8);
    double q = (b2 ? p[1] : p[0]) + __shfl_xor_sync(FULL, b2 ? p[0] : p[1], 4);
    q += __shfl_xor_sync(FULL, q, 2);
    q += __shfl_xor_sync(FULL, q, 1);
    return q;   // total of value (lane >> 2) in every lane
}

template <bool SP>
__global__ void __launch_bounds__(PS_THREADS, 1) k_pcg_solve(DevProblem P, double tol2, int max_iters, int pin_hg, int pin_f16, int ps_mode) {
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(128) unsigned char ps_smem_raw[];
    PsSmem& sm = *reinterpret_cast<PsSmem*>(ps_smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, b = blockIdx.x;
    const bool producer = warp == PS_CWARPS;
    const int n = 6 * P.Nc;
    const int G = P.xchg_n, me = P.xchg_rank;
    if (tid == 0) {
        for (int s = 0; s < PS_STAGES; s++) { ps_mbar_init(&sm.full[s], 1); ps_mbar_init(&sm.empty[s], SP ? PS_CWARPS : 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // ---- owner role: item b = elements 96 b .. 96 b + 95, thread tid < 96 owns element i
    const bool has_item = b < P.ps_nitems;
    const int e = tid % PS_ITEM, ch = tid / PS_ITEM;   // fold role: element e, chain ch
    const int i = b * PS_ITEM + e;
    const bool owner = has_item && tid < PS_ITEM && i < n;
    const int a6 = i % 6;
    double x_i = 0.0, r_i = 0.0, z_i = 0.0, p_i = 0.0, M[6] = {0, 0, 0, 0, 0, 0};
    if (owner) {
        const double* Mr = P.Minv + 36 * (size_t)(i / 6) + 6 * a6;
#pragma unroll
        for (int k = 0; k < 6; k++) M[k] = Mr[k];
        r_i = P.rhs[i];
    }
    // z = Minv r of the item (r of the other elements of the camera through shared memory), and the item's r.z
    auto precondition = [&]() -> double {
        __syncthreads();
        if (tid < PS_ITEM) sm.v[0][tid] = r_i;
        __syncthreads();
        double rz = 0.0;
        if (owner) {
            const double* rc = &sm.v[0][e - a6];
            double z = 0.0;
#pragma unroll
            for (int k = 0; k < 6; k++) z += M[k] * rc[k];
            z_i = z;
            __stcg(P.pcg_z + i, z);
            rz = r_i * z;
        }
        return block_sum(rz, sm.sh);
    };
    // ---- init: x = 0, r = rhs, z = Minv r, p = z (beta = 0 with p_old = z), rz0 = r.z
    {
        const double rzp = precondition();
        if (owner) { p_i = z_i; __stcg(P.pcg_p + i, z_i); }
        if (has_item && tid == 0) __stcg(P.ps_rz + b, rzp);
    }
    __threadfence();
    grid.sync();
    double rz = ps_fold(P.ps_rz, P.ps_nitems, sm.sh);
    const double rz0 = rz;
    double beta = 0.0;
    int it = 0, cur = 0;
    uint32_t seq = G > 1 ? *P.xchg_ctr : 0u;
    const int32_t* seg_tab = SP ? P.sp_seg : P.ps_seg;
    double* part = SP ? P.sp_part : P.ps_part;
    const int32_t* fold_ptr = SP ? P.sp_fold_ptr : P.ps_fold_ptr;
    const int32_t* fold_off = SP ? P.sp_fold_off : P.ps_fold_off;
    const int seg0 = (SP ? P.sp_seg_ptr : P.ps_seg_ptr)[b], seg1 = (SP ? P.sp_seg_ptr : P.ps_seg_ptr)[b + 1];
    const size_t colbase = (size_t)P.ex_nown * EX_TD;
    // this CTA's run: groups gb0 .. gb0 + ng - 1 of the rank's 8-row groups = one contiguous byte range of the tiles
    int ng = 0;
    const double* run = nullptr;
    int q_first = 0;   // SP: first panel of the run (ng = number of panels)
    if (seg1 > seg0) {
        const int4 S0 = __ldg(reinterpret_cast<const int4*>(seg_tab) + seg0);
        if (SP) q_first = S0.y; else run = P.ex_Sown + ((size_t)S0.x * 48 + S0.y) * 8 * EX_TD;
        for (int sg = seg0; sg < seg1; sg++) { const int4 S = __ldg(reinterpret_cast<const int4*>(seg_tab) + sg); ng += S.z - S.y; }
    }
    double* const ring_flat = &sm.ring[0][0];
    unsigned int jdone = 0;   // groups streamed in the previous iterations: stage = (2 * group number + half) % 16
    // ---- ring prefetch (ps_mode bit 0): S does not change during the solve and the ring is empty at the end of every mat-vec
    // (block barrier), so the producer warp refills it with the FIRST chunks / stages of the run right away -- one copy per
    // lane, positions static -- and the copies land while the CTA sits in the two grid barriers and the fold, when DRAM is
    // otherwise idle.  The producer loop of the next iteration starts behind them.  ps_mode bit 1: the L2-resident share
    // (evict_last) is spread evenly over the rest of the run (pin_f16 / 65536 of its bytes) instead of leading it, so that
    // L2 hits and DRAM streams overlap in time in every CTA.
    uint64_t pol_keep, pol_stream;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_keep));
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_stream));
    // Block-sparse panels only: on the dense tiles neither changed the time (measured), so that variant keeps its plain loop.
    const bool do_prefetch = SP && (ps_mode & 1), do_interleave = SP && (ps_mode & 2);
    int npre = 0;
    uint32_t pre_at = 0, pre_B = 0, pre_total = 0;
    int pre_dx = 0;
    if constexpr (SP) {
        if (producer && do_prefetch) {
            const bool cand = lane < ng && lane < SP_SLOTS;
            const int4 dsc = cand ? __ldg(P.sp_chunk + q_first + lane) : make_int4(0, 0, 0, 0);
            uint32_t incl = (uint32_t)dsc.y;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += t; }
            npre = __popc(__ballot_sync(FULL, cand && incl <= SP_RING_BYTES));   // a prefix of the lanes: the sums grow
            pre_B = (uint32_t)dsc.y; pre_at = incl - pre_B; pre_dx = dsc.x;
            pre_total = npre > 0 ? __shfl_sync(FULL, incl, npre - 1) : 0u;
        }
    }
    auto prefetch = [&](unsigned int jbase) {   // producer warp, ring empty
        if constexpr (SP) {
            if (lane < npre) {
                const unsigned int H = jbase + lane, slot = H % SP_SLOTS;
                *reinterpret_cast<volatile unsigned int*>(&sm.poff[slot]) = pre_at;
                ps_mbar_expect_tx(&sm.full[slot], pre_B);
                ps_bulk_load_hint(reinterpret_cast<unsigned char*>(ring_flat) + pre_at,
                                  reinterpret_cast<const unsigned char*>(P.sp_data) + 16 * (size_t)(unsigned int)pre_dx, pre_B, &sm.full[slot], pol_stream);
            }
            __syncwarp();
        }
    };
    if constexpr (SP) { if (producer && npre > 0) prefetch(0u); }
    // phase clock of CTA 0 (ns, summed over the iterations): mat-vec, barrier 1, fold + vector part, barrier 2 + scalars
    auto now = [&]() -> long long { long long t = 0; if (tid == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; };
    long long t_m = 0, t_s1 = 0, t_f = 0, t_s2 = 0, t_a = now();
    long long w_cons = 0, w_comp = 0;   // SP: cycles warp 0 waited for chunks / spent on their panels

    while (rz > tol2 * rz0 && rz0 > 0.0 && it < max_iters) {
        const double* p_old = cur ? P.pcg_p2 : P.pcg_p;
        double* p_new = cur ? P.pcg_p : P.pcg_p2;
        // ---- direction of this iteration (owner copy; the mat-vec below rebuilds the same values from z, beta, p_old)
        if (owner) { p_i = ps_dir(z_i, beta, p_i); __stcg(p_new + i, p_i); }
        // ---- M: tile mat-vec over this CTA's run
        if (producer && SP) {
            // The descriptors are fetched 32 at a time by the whole warp (next batch in flight while lane 0 issues the current
            // one); pin_hg = bytes of the run to keep in L2.  All panels of the earlier iterations are consumed (block barrier
            // at the end of M), so every iteration starts with an empty ring.
            // With the ring prefetch the first npre chunks are already in flight (oldest at offset 0, head behind the last).
            unsigned int bytes_done = 0, tailH = jdone;
            uint32_t head = pre_total, tail = 0;
            unsigned long long credit = 0;   // interleaved L2 share: a chunk is kept when the credit covers it
            volatile unsigned int* poff = sm.poff;
            // lane 0 alone (no warp-level reconvergence between two copies: a shuffle after the bulk copy was measured to hold
            // the next copy back until the previous one had landed); the descriptors of the next 8 panels are in registers
            if (lane == 0) {
                constexpr int PF = 8;
                int4 d[PF];
#pragma unroll
                for (int j = 0; j < PF; j++) d[j] = npre + j < ng ? __ldg(P.sp_chunk + q_first + npre + j) : make_int4(0, 0, 0, 0);
                for (int base = npre; base < ng; base += PF) {
#pragma unroll
                    for (int j = 0; j < PF; j++) {
                        if (base + j >= ng) break;
                        const int dx = d[j].x;
                        const uint32_t B = (uint32_t)d[j].y;   // a multiple of 128: panels start on 128-byte boundaries (stream and ring)
                        if (base + j + PF < ng) d[j] = __ldg(P.sp_chunk + q_first + base + j + PF);
                        const unsigned int H = jdone + base + j, slot = H % SP_SLOTS;
                        auto free_oldest = [&]() {   // panels are released in any order by the warps, reclaimed in order here
                            ps_mbar_wait(&sm.empty[tailH % SP_SLOTS], (tailH / SP_SLOTS) & 1u);
                            tailH++;
                            tail = tailH == H ? head : poff[tailH % SP_SLOTS];
                        };
                        while ((int)(H - tailH) >= SP_SLOTS) free_oldest();   // the slot's previous panel is consumed
                        uint32_t at;
                        while (true) {
                            if (tailH == H) { head = tail = 0; at = 0; break; }
                            if (head >= tail) {
                                if (head + B <= SP_RING_BYTES) { at = head; break; }
                                if (B < tail) { at = 0; break; }             // strict: head == tail means empty
                            } else if (head + B < tail) { at = head; break; }
                            free_oldest();
                        }
                        head = at + B;
                        poff[slot] = at;
                        ps_mbar_expect_tx(&sm.full[slot], B);
                        bool keep = bytes_done < (unsigned int)pin_hg;
                        if (do_interleave) {
                            credit += (unsigned long long)B * (unsigned int)pin_f16;
                            keep = credit >= ((unsigned long long)B << 16);
                            if (keep) credit -= (unsigned long long)B << 16;
                        }
                        ps_bulk_load_hint(reinterpret_cast<unsigned char*>(ring_flat) + at, reinterpret_cast<const unsigned char*>(P.sp_data) + 16 * (size_t)(unsigned int)dx,
                                          B, &sm.full[slot], keep ? pol_keep : pol_stream);
                        bytes_done += B;
                    }
                }
            }
        } else if (producer) {
            if (lane == 0) {
                // the first pin_hg half groups of every run are loaded evict_last: S does not change during the solve, so that
                // share of the tiles (sized to the L2 by the host) is served from L2 from the second CG iteration on; the rest
                // streams through evict_first and cannot push it out
                if constexpr (SP) { } else
                for (unsigned int h = 0; h < 2u * ng; h++) {
                    const unsigned int H = 2u * jdone + h, s = H % PS_STAGES, use = H / PS_STAGES;
                    if (use > 0) ps_mbar_wait(&sm.empty[s], (use - 1) & 1u);   // the owner warp has released the previous contents
                    ps_mbar_expect_tx(&sm.full[s], PS_STAGE_BYTES);
                    ps_bulk_load_hint(sm.ring[s], run + (size_t)h * PS_HALF_ROWS * EX_TD, PS_STAGE_BYTES, &sm.full[s],
                                      (int)h < pin_hg ? pol_keep : pol_stream);
                }
            }
        } else {
            double pq_lane = 0.0;
            unsigned int jseg = jdone;   // number of the first group of the segment
            for (int sg = seg0; sg < seg1; sg++) {
                const int4 S = __ldg(reinterpret_cast<const int4*>(seg_tab) + sg);   // local tile, first group, end group (SP: panels)
                const int tile = S.x, g0 = S.y, g1 = S.z;
                const int I = __ldg(P.ex_tile_I + P.ex_own0 + tile), J = __ldg(P.ex_tile_J + P.ex_own0 + tile);
                ps_consumer_sync();   // xI / col of the previous segment are free
                if constexpr (SP) {
                    for (int r = tid; r < EX_TD; r += 32 * PS_CWARPS) {
                        const int gi = EX_TD * I + r, gj = EX_TD * J + r;
                        sm.xI[r] = gi < n ? ps_dir(__ldcg(P.pcg_z + gi), beta, __ldcg(p_old + gi)) : 0.0;
                        sm.xJ[r] = gj < n ? ps_dir(__ldcg(P.pcg_z + gj), beta, __ldcg(p_old + gj)) : 0.0;
                    }
#pragma unroll
                    for (int k = 0; k < 12; k++) sm.col[warp][lane + 32 * k] = 0.0;
                    ps_consumer_sync();
                    // every warp visits every chunk of the segment; of its panels it takes those whose number is = warp (mod 8);
                    // a warp's column sums live in its row of sm.col
                    double* colw = sm.col[warp];
                    for (unsigned int o = 0; o < (unsigned int)(g1 - g0); o++) {
                        const unsigned int H = jseg + o, s = H % SP_SLOTS;
                        const long long w0 = ps_clock_now();
                        ps_mbar_wait(&sm.full[s], (H / SP_SLOTS) & 1u);
                        const uint32_t at = *reinterpret_cast<volatile unsigned int*>(&sm.poff[s]);   // written before the producer's arrive (release) on full[s]
                        const long long w1 = ps_clock_now();
                        w_cons += w1 - w0;
                        const unsigned char* cbase = reinterpret_cast<const unsigned char*>(ring_flat) + at;
                        const uint16_t* dir = reinterpret_cast<const uint16_t*>(cbase);
                        const int np = dir[0];
                        for (int pk = (warp + PS_CWARPS - dir[1]) % PS_CWARPS; pk < np; pk += PS_CWARPS) {
                        const unsigned char* pbase = cbase + 128u * dir[2 + pk];
                        const int row0 = reinterpret_cast<const uint16_t*>(pbase)[0], m6 = reinterpret_cast<const uint16_t*>(pbase)[1];
                        const double* buf = reinterpret_cast<const double*>(pbase + 16);
                        const uint16_t* cmap = reinterpret_cast<const uint16_t*>(buf + 6 * m6);
                        double xi[6], racc[6];
#pragma unroll
                        for (int r = 0; r < 6; r++) { xi[r] = sm.xI[row0 + r]; racc[r] = 0.0; }
                        int c = lane;
                        for (; c + 32 < m6; c += 64) {   // two columns per lane in flight (distinct columns: no aliasing in colw; four in flight measured no faster)
                            const int ca = cmap[c], cb = cmap[c + 32];
                            const double xa = sm.xJ[ca], xb = sm.xJ[cb];
                            const double wa = colw[ca], wb = colw[cb];
                            double va[6], vb[6];
#pragma unroll
                            for (int r = 0; r < 6; r++) { va[r] = buf[r * m6 + c]; vb[r] = buf[r * m6 + c + 32]; }
#pragma unroll
                            for (int r = 0; r < 6; r++) { racc[r] += va[r] * xa; racc[r] += vb[r] * xb; }
                            colw[ca] = wa + ((va[0] * xi[0] + va[1] * xi[1] + va[2] * xi[2]) + (va[3] * xi[3] + va[4] * xi[4] + va[5] * xi[5]));
                            colw[cb] = wb + ((vb[0] * xi[0] + vb[1] * xi[1] + vb[2] * xi[2]) + (vb[3] * xi[3] + vb[4] * xi[4] + vb[5] * xi[5]));
                        }
                        if (c < m6) {
                            const int ca = cmap[c];
                            const double xa = sm.xJ[ca], wa = colw[ca];
                            double va[6];
#pragma unroll
                            for (int r = 0; r < 6; r++) va[r] = buf[r * m6 + c];
#pragma unroll
                            for (int r = 0; r < 6; r++) racc[r] += va[r] * xa;
                            colw[ca] = wa + ((va[0] * xi[0] + va[1] * xi[1] + va[2] * xi[2]) + (va[3] * xi[3] + va[4] * xi[4] + va[5] * xi[5]));
                        }
                        int j6;
                        const double s6 = warp_sum6(racc, lane, &j6);
                        if ((lane & 3) == 0 && j6 < 6) {
                            __stcg(part + (size_t)tile * EX_TD + row0 + j6, s6);
                            pq_lane += sm.xI[row0 + j6] * s6;
                        }
                        }
                        __syncwarp();
                        if (lane == 0) ps_mbar_arrive(&sm.empty[s]);
                        w_comp += ps_clock_now() - w1;
                    }
                } else {
                for (int r = 8 * g0 + tid; r < 8 * g1; r += 32 * PS_CWARPS) {
                    const int gi = EX_TD * I + r;
                    sm.xI[r] = gi < n ? ps_dir(__ldcg(P.pcg_z + gi), beta, __ldcg(p_old + gi)) : 0.0;
                }
                double xJ[12], col[12];
#pragma unroll
                for (int k = 0; k < 12; k++) {
                    const int gj = EX_TD * J + lane + 32 * k;
                    xJ[k] = gj < n ? ps_dir(__ldcg(P.pcg_z + gj), beta, __ldcg(p_old + gj)) : 0.0;
                    col[k] = 0.0;
                }
                ps_consumer_sync();
                // groups of the segment whose number is = warp (mod 8)
                const unsigned int first = (warp + PS_CWARPS - (jseg % PS_CWARPS)) % PS_CWARPS;
                for (unsigned int o = first; o < (unsigned int)(g1 - g0); o += PS_CWARPS) {
                    const unsigned int j = jseg + o;
                    const uint32_t parity = (j / PS_CWARPS) & 1u;   // use number of the warp's two stages
                    const int row0 = 8 * (g0 + (int)o);
                    double racc[8];
#pragma unroll
                    for (int hf = 0; hf < 2; hf++) {
                        const int s = 2 * warp + hf;
                        ps_mbar_wait(&sm.full[s], parity);
                        const double* buf = sm.ring[s];
#pragma unroll
                        for (int rr = 0; rr < PS_HALF_ROWS; rr++) {
                            const double xi = sm.xI[row0 + 4 * hf + rr];   // warp-uniform
                            double acc0 = 0.0, acc1 = 0.0;
#pragma unroll
                            for (int k = 0; k < 12; k += 2) {
                                const double v0 = buf[rr * EX_TD + lane + 32 * k], v1 = buf[rr * EX_TD + lane + 32 * (k + 1)];
                                acc0 += v0 * xJ[k]; acc1 += v1 * xJ[k + 1];
                                col[k] += v0 * xi; col[k + 1] += v1 * xi;
                            }
                            racc[4 * hf + rr] = acc0 + acc1;
                        }
                        __syncwarp();
                        if (lane == 0) ps_mbar_arrive(&sm.empty[s]);
                    }
                    const double s8 = ps_warp_sum8(racc, lane);
                    if ((lane & 3) == 0) {
                        const int r = row0 + (lane >> 2);
                        __stcg(part + (size_t)tile * EX_TD + r, s8);
                        pq_lane += sm.xI[r] * s8;
                    }
                }
#pragma unroll
                for (int k = 0; k < 12; k++) sm.col[warp][lane + 32 * k] = col[k];
                }
                ps_consumer_sync();
                for (int c = tid; c < EX_TD; c += 32 * PS_CWARPS) {
                    double acc = sm.col[0][c];
#pragma unroll
                    for (int w = 1; w < PS_CWARPS; w++) acc += sm.col[w][c];   // warp order: fixed
                    __stcg(part + colbase + (size_t)sg * EX_TD + c, acc);
                }
                jseg += (unsigned int)(g1 - g0);
            }
            // this CTA's part of p^T S p / 2: warp trees, then the warps in order
            pq_lane = warp_sum(pq_lane);
            if (lane == 0) sm.pqw[warp] = pq_lane;
        }
        jdone += (unsigned int)ng;
        __syncthreads();
        if constexpr (SP) { if (producer && npre > 0) prefetch(jdone); }   // every chunk of this iteration is consumed: refill the ring for the next one
        if (tid == 0) {
            double v = sm.pqw[0];
#pragma unroll
            for (int w = 1; w < PS_CWARPS; w++) v += sm.pqw[w];
            __stcg(P.ps_pq + b, v);
        }
        long long t_b = now(); t_m += t_b - t_a;
        __threadfence();
        grid.sync();
        t_a = now(); t_s1 += t_a - t_b;
        // ---- F: q of the item (fold of the partials that touch its camera tile), alpha, x, r, z, item part of r.z
        if (has_item) {
            const int K = (b * PS_ITEM) / EX_TD;
            const int m0 = __ldg(fold_ptr + K), m1 = __ldg(fold_ptr + K + 1);
            {   // chain ch sums the entries m0 + ch, m0 + ch + 3, ... in list order
                const double* base = part + (b * PS_ITEM - K * EX_TD + e);
                double acc = 0.0;
                int m = m0 + ch;
                for (; m + 15 * PS_CHAINS < m1; m += 16 * PS_CHAINS) {   // 16 loads in flight
                    double v[16];
#pragma unroll
                    for (int k = 0; k < 16; k++) v[k] = __ldcg(base + __ldg(fold_off + m + PS_CHAINS * k));
#pragma unroll
                    for (int k = 0; k < 16; k++) acc += v[k];
                }
                for (; m < m1; m += PS_CHAINS) acc += __ldcg(base + __ldg(fold_off + m));
                sm.v[ch][e] = acc;
            }
            double pq = ps_fold(P.ps_pq, gridDim.x, sm.sh);   // this rank's p^T S p / 2 (barriers inside: sm.v is visible below)
            double q_i = 0.0;
            if (tid < PS_ITEM) q_i = (sm.v[0][tid] + sm.v[1][tid]) + sm.v[2][tid];
            if (G > 1) {
                // sum q and p.q over the ranks, in rank order, through the peer mailboxes -- low-latency protocol: every 8-byte
                // value travels as ONE 16-byte store {lo, seq, hi, seq}, so the flag arrives with the data (each 8-byte half
                // validates itself) and no fence / second round trip is needed: one NVLink write latency per exchange.
                // Slots alternate with the parity of seq (a rank is at most one exchange ahead of a peer).
                seq++;
                if (tid <= PS_ITEM) {
                    const double mine = tid < PS_ITEM ? q_i : pq;
                    const unsigned long long bits = (unsigned long long)__double_as_longlong(mine);
                    const uint32_t lo = (uint32_t)bits, hi = (uint32_t)(bits >> 32);
                    const size_t slot = ((size_t)((seq & 1u) * G) * P.ps_nitems + b) * PS_SLOT + tid;   // + source * nitems * PS_SLOT
                    const size_t src_stride = (size_t)P.ps_nitems * PS_SLOT;
                    for (int g = 0; g < G; g++)
                        if (g != me) {
                            uint4* dst = reinterpret_cast<uint4*>(P.xchg_peer[g] + P.xchg_ll_off) + slot + (size_t)me * src_stride;
                            asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" :: "l"(dst), "r"(lo), "r"(seq), "r"(hi), "r"(seq) : "memory");
                        }
                    double acc = 0.0;
                    const long long t0 = clock64();
                    for (int g = 0; g < G; g++) {
                        if (g == me) { acc += mine; continue; }
                        const uint4* srcp = reinterpret_cast<const uint4*>(P.xchg_peer[me] + P.xchg_ll_off) + slot + (size_t)g * src_stride;
                        uint32_t a0, f0, a1, f1;
                        while (true) {
                            asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(a0), "=r"(f0), "=r"(a1), "=r"(f1) : "l"(srcp) : "memory");
                            if (f0 == seq && f1 == seq) break;
                            if (clock64() - t0 > 6000000000LL) { P.flags[FL_XCHG_TIMEOUT] = 1; break; }   // ~3 s: a peer died
                        }
                        acc += __longlong_as_double((long long)(((unsigned long long)a1 << 32) | a0));
                    }
                    if (tid < PS_ITEM) q_i = acc; else sm.pq = acc;
                }
                __syncthreads();
                pq = sm.pq;
            }
            const double alpha = rz / (2.0 * pq);
            if (owner) { x_i += alpha * p_i; r_i -= alpha * q_i; }
            const double rzp = precondition();
            if (tid == 0) __stcg(P.ps_rz + b, rzp);
        }
        t_b = now(); t_f += t_b - t_a;
        __threadfence();
        grid.sync();
        // ---- beta and the convergence scalars: every CTA folds the same partials in the same order
        const double rz_new = ps_fold(P.ps_rz, P.ps_nitems, sm.sh);
        beta = rz_new / rz;
        rz = rz_new;
        it++;
        cur ^= 1;
        t_a = now(); t_s2 += t_a - t_b;
        if (G > 1 && *reinterpret_cast<volatile int32_t*>(P.flags + FL_XCHG_TIMEOUT)) break;   // set before the barrier: uniform
    }
    if constexpr (SP) {
        if (producer && lane < npre) {   // the prefetched copies of the iteration that did not happen must land before the CTA exits
            const unsigned int H = jdone + lane;
            ps_mbar_wait(&sm.full[H % SP_SLOTS], (H / SP_SLOTS) & 1u);
        }
    }
    if (owner) P.pcg_x[i] = x_i;
    if (b == 0 && tid == 0) {
        P.scalars[SC_RZ] = rz; P.scalars[SC_RZ_NEXT] = rz; P.scalars[SC_RZ0] = rz0;
        P.flags[FL_PCG_ITERS] = it; P.flags[FL_PCG_ITERS_NEXT] = it;
        if (G > 1) *P.xchg_ctr = seq;
        P.ps_clock[0] = (double)t_m; P.ps_clock[1] = (double)t_s1; P.ps_clock[2] = (double)t_f; P.ps_clock[3] = (double)t_s2;
    }
    if (tid == 0) {   // per-CTA mat-vec time and SM (measurement only)
        unsigned int smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        P.ps_clock[4 + 2 * b] = (double)t_m; P.ps_clock[5 + 2 * b] = (double)smid;
        if (SP) { P.ps_clock[4 + 2 * gridDim.x + 2 * b] = (double)w_cons; P.ps_clock[5 + 2 * gridDim.x + 2 * b] = (double)w_comp; }
    }
}

static int ps_grid_blocks() {
    static int blocks_of[64];   // per device ordinal (0 = not probed yet, -1 = unavailable)
    int dev = 0;
    cudaGetDevice(&dev);
    int& cached = blocks_of[dev & 63];
    if (cached != 0) return cached < 0 ? 0 : cached;
    int blocks = -1;
    {
        int coop = 0, sms = 0, per_sm = 0, smem_max = 0;
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        blocks = 0;
        if (coop && smem_max >= (int)sizeof(PsSmem) &&
            cudaFuncSetAttribute(k_pcg_solve<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(PsSmem)) == cudaSuccess &&
            cudaFuncSetAttribute(k_pcg_solve<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(PsSmem)) == cudaSuccess &&
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_pcg_solve<false>, PS_THREADS, sizeof(PsSmem)) == cudaSuccess && per_sm >= 1 &&
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_pcg_solve<true>, PS_THREADS, sizeof(PsSmem)) == cudaSuccess && per_sm >= 1)
            blocks = sms;
        cudaGetLastError();
    }
    cached = blocks > 0 ? blocks : -1;
    return blocks;
}
int pcg_solve_grid() { return ps_grid_blocks(); }
int pcg_solve_items(int n_cams) { return (6 * n_cams + PS_ITEM - 1) / PS_ITEM; }
int pcg_solve_slot() { return PS_SLOT; }
bool launch_pcg_solve(const DevProblem& P, double tol2, int max_iters, cudaStream_t st) {
    if (!P.ps_on || P.ps_grid != ps_grid_blocks()) return false;
    DevProblem Pc = P;
    // share of the rank's tiles kept L2-resident across the CG iterations (SBA_PS_L2_MB, default 80 MB of the 126 MB L2),
    // as the same number of leading 12 KB half groups of every CTA's run
    const double l2_mb = std::getenv("SBA_PS_L2_MB") ? std::atof(std::getenv("SBA_PS_L2_MB")) : 80.0;
    int pin_hg = (int)(l2_mb * 1e6 / ((double)P.ps_grid * PS_STAGE_BYTES));
    if (P.sp_on) pin_hg = (int)(l2_mb * 1e6 / (double)P.ps_grid);   // bytes of every run
    // SBA_PS_MODE: bit 0 ring prefetch during the barriers / fold, bit 1 L2-resident share interleaved over the run
    // (block-sparse panels only, default both on: measured 57.2 -> 56.2 us per CG iteration on the Venice shape; the dense-tile
    // variant ignores it -- neither changed its time).  pin_f16 / 65536 = share of the bytes behind the prefetched ones to keep.
    const int ps_mode_env = std::getenv("SBA_PS_MODE") ? std::atoi(std::getenv("SBA_PS_MODE")) : -1;
    const int ps_mode = ps_mode_env >= 0 ? ps_mode_env : 3;
    const double frac = P.sp_on ? l2_mb * 1e6 / std::max((double)P.sp_bytes - ((ps_mode & 1) ? (double)P.ps_grid * SP_RING_BYTES : 0.0), 1.0) : 0.0;
    int pin_f16 = (int)(std::min(std::max(frac, 0.0), 1.0) * 65536.0);
    int mode = ps_mode;
    void* args[] = {&Pc, &tol2, &max_iters, &pin_hg, &pin_f16, &mode};
    return cudaLaunchCooperativeKernel(P.sp_on ? (void*)k_pcg_solve<true> : (void*)k_pcg_solve<false>, dim3(P.ps_grid), dim3(PS_THREADS), args, sizeof(PsSmem), st) == cudaSuccess;
}

}  // namespace sba
