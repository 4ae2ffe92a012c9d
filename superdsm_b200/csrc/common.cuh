// Shared definitions of the superdsm_b200 CUDA library (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/superdsm_b200.h"

namespace sdsm {

constexpr int NSLOT      = 25;    // (2*NC+1)^2 neighbour cells with NC = 2 >= ceil(R / subsample)
constexpr int NC         = 2;
constexpr int GROW       = 28;    // floats per pixel row of the G~ table: NSLOT rounded up to whole 16-byte groups
constexpr int LOCAL_MAX  = 32;    // 6 polynomial + 25 slots, padded
constexpr int SCRH       = 496;   // packed lower triangle of a 31 x 31 local Hessian
constexpr int SCRG       = 32;    // local gradient (31) + value
constexpr int TILE_LD    = 36;    // row stride (doubles) of the per-warp tile T[local parameter][pixel of the chunk]: = 4 mod 16, so that
                                  // both the column writes (lane = pixel) and the DMMA fragment reads are bank-conflict free
constexpr int SOLVE_BLOCK = 256;  // k_solve: persistent CTAs, every task is executed by one CTA
constexpr int EVAL_BLOCK = SOLVE_BLOCK;   // evaluation task: one warp per pixel cell
constexpr int EVAL_WARPS = EVAL_BLOCK / 32;
constexpr int EVAL_CELLS = 32;    // most non-empty pixel cells per evaluation task: the CTA's warps take them one by one
// per-warp shared memory of the evaluation items: STAGES staging buffers, each one chunk of 32 pixels brought in by three 1-D
// TMA bulk copies (cp.async.bulk, mbarrier-tracked): the pixels' rows of the G~ panel [32][stride], y [32 (+2)], row|col
// [32 (+4)]; the Hessian pass reuses the space for its tile T[32][TILE_LD] + dd[32].  Then xi[32] (the xi values the cell
// sees) and u[32] (-y theta of the chunk's pixels, handed from the pixel lanes to the slot lanes).
constexpr int SLOTJ_INTS = 64;    // per-warp shared memory of the evaluation items (ints): grid index of the cell's active slots [32], slot of each active entry [32]
constexpr int STAGES = 3;
constexpr int STAGE_G_BYTES = GROW * 32 * 4;                  // 3584
constexpr int STAGE_Y_BYTES = 34 * 8;                         // 272: the copy starts at an even pixel index (16-byte alignment)
constexpr int STAGE_RC_BYTES = 36 * 4;                        // 144: ... at a pixel index that is a multiple of four
constexpr int STAGE_BYTES = STAGE_G_BYTES + STAGE_Y_BYTES + STAGE_RC_BYTES;    // 4000
static_assert(STAGE_BYTES % 16 == 0 && STAGE_G_BYTES % 16 == 0 && (STAGE_G_BYTES + STAGE_Y_BYTES) % 16 == 0, "TMA destinations are 16-byte aligned");
constexpr int WARP_AREA_DOUBLES = STAGES * STAGE_BYTES / 8;      // 1500 >= 32 * TILE_LD + 32
static_assert(WARP_AREA_DOUBLES >= 32 * TILE_LD + 32, "the Hessian tile must fit the staging area");
constexpr int XI_OFF = WARP_AREA_DOUBLES;
constexpr int U_OFF = WARP_AREA_DOUBLES + 32;
constexpr int TILE_DOUBLES = WARP_AREA_DOUBLES + 64;
// row stride (floats) of a cell's block of the G~ panel with `na` active slots: a multiple of 4 (16-byte groups) that is
// = 4 mod 8, so that the 16-byte reads of 8 consecutive rows fall into 8 different bank groups
__host__ __device__ __forceinline__ int grow_stride(int na) { return na <= 4 ? 4 : (na <= 12 ? 12 : (na <= 20 ? 20 : GROW)); }
constexpr int CTRL_BLOCK = SOLVE_BLOCK;   // control task: one CTA advances one candidate's state machine
constexpr int CTRL_WARPS = CTRL_BLOCK / 32;
#ifndef SDSM_SMALL_N_MAX
#define SDSM_SMALL_N_MAX 104      // (148 would fill the 100 KB of dynamic shared memory a CTA of k_solve has anyway; measured: no gain,
#endif                            //  172.6 vs 173.3 ms per step -- the factorisations of mid-size systems are not what bounds the step)
constexpr int SMALL_N_MAX = SDSM_SMALL_N_MAX;
constexpr int SMEM_L_CAP = SMALL_N_MAX * (SMALL_N_MAX + 1) / 2;  // doubles of shared memory for the packed Cholesky factor  // systems up to this size are factorised in shared memory, larger ones blocked in global memory / L2
constexpr int BIG_BLOCK  = SOLVE_BLOCK;
constexpr int N_COUNTERS = 32;
constexpr int SPEC_MAX   = 8;     // line-search trial points one request can evaluate (speculative backtracking, ctrl.cuh)
constexpr int NVEC       = 12 + 2 * SPEC_MAX;    // N-vectors of the interior-point solver: 12 + the trial points and their gradients

// one task queue: two rings (0: normal, 1: urgent) of `mask + 1` slots each, one ring of 2 (mask + 1) arrival events
struct Queue {
    uint2*    task;        // [2 cap]  x = type << 28 | candidate, y = first item of the task
    uint32_t* seq;         // [2 cap]  slot sequence numbers
    uint32_t* ev;          // [2 cap]  arrival events (sequence numbers only)
    unsigned long long* hot;        // [N_HOT * HOT_STRIDE] head / tail / availability counters, one cache line each
    uint32_t  mask, ev_mask;
};

struct Img { const double* y; const int32_t* atoms; int32_t H, W; };
struct ImgSet { Img v[SDSM_MAX_IMAGES]; };      // passed to the kernels by value (constant bank): no device-side table to maintain

struct Par {
    double alpha, epsilon, scale, sparsity_tol, abstol, reltol, feastol, hess_tol;
    int32_t sub, R, has_smooth, init_elliptical, maxiters, skip_single_fg;
    const float* psf;   // device, (R+1)^2
};

struct CInfo {
    int32_t hp, wp;        // compacted bounding box (rows / columns of the box without any region pixel removed, dsm.py:185-186)
    int32_t ch, cw;        // cells per dimension
    int32_t ncells, m;     // m = number of deformation parameters (0: null matrix)
    int32_t n_fg, flags;   // pixels with y > 0; bit0: region pixel count mismatch, bit1: zero row in G~
    int32_t n_ne, pad;     // number of non-empty cells (entries of cell_list)
};

// State of one candidate's fit between two control tasks (while its evaluation tasks run).  The interior-point
// fields are the local variables of cvxopt's cpl (oracle/cpsolver.py) that live across an evaluation of F.
struct CState {
    double t, s, z, d, lm;
    double th1, th2, th3, pres0, dres0;
    double phi0s, dphi0s, gap0s, step0s, dsdz0s, sigma0s, eta0s;
    double sv_t, sv_s, sv_z, sv_d, sv_lm, sv_dt, sv_dz, sv_ds, sv_resx, sv_rznl, sv_rxt;
    double gap, rxt, resx, rznl, resznl, phi, dphi, lmsq;
    double bt, uz0, vt, gux, gux_res, gex;
    double sigma, eta, step, dt, dz, ds, dsdz, dz2, ds2;
    double val1, psi_last, energy;
    double z_spec;                         // the z a speculative Hessian was scaled with (spec_h)
    double f_trial, psi_trial;             // scaled value / psi of the last line-search trial (= the next iterate when it is accepted)
    unsigned long long ns_ctrl;            // diagnostics: time spent in this candidate's control tasks so far (ns)
    int32_t pc, stage, iters, relaxed, pass, use_xi;
    int32_t req_vec, req_gvec, req_h;      // the pending evaluation: point, where the gradient goes, Hessian wanted
    int32_t status, bits, it1, it2, it3, nf, nh, have1, r_ipm;
    // speculative backtracking: the pending / last trial request holds spec_n points (x + step beta^k dx, k < spec_n), spec_k is
    // the one being examined; ls_cnt = trials examined in the current line search, ls_prev = in the previous one (the predictor)
    int32_t spec_n, spec_k, ls_cnt, ls_prev;
    int32_t spec_h, spec_pad;              // spec_h: the last point of the pending / last trial request was evaluated with its Hessian
    // statistics (sdsm_eval_counts): trials cp evaluates F at that were decided without evaluation / trial points evaluated but
    // never examined / accepted trials whose Hessian came with them (one pass instead of two) / Hessians evaluated in vain
    int32_t n_skipped, n_unused, n_merged, n_hwasted;
};

// what the evaluation tasks need to know about a candidate (written by the control task that files the request)
struct EvalReq {
    const double* x;       // point to evaluate (6 + m values, or 6)
    double hfac;           // factor of the Hessian rows the T_HG tasks write: z * scale / n (objects.py:380), or 1 (diagnostics)
    int32_t want_h, use_xi;  // want_h: 0 value + gradient, 1 value + gradient + Hessian, 2 Hessian only (value and gradient at this
                             // point are already known: the accepted trial of the line search was evaluated at it)
    int32_t cpt, cpw;      // non-empty cells per CTA-wide evaluation task (chains that stay in one CTA, diagnostics) / per warp item
    int32_t npts, ld;      // points to evaluate: x + k * ld for k < npts, results in the k-th copy of scrG / xig; with npts > 1,
                           // want_h is about the last point
};

// Workspace carved out of the caller's buffer.  Per-pixel / per-cell / per-table arrays are indexed with the
// candidate's offsets (sdsm_cand.pix_off / cell_off / tab_off / vec_off / mat_off).
struct WS {
    double*   pix_y;       // [P]      y of the region's pixels, cell-major
    uint32_t* pix_rc;      // [P]      compacted (row | col << 16) inside the bounding box
    float*    pix_s;       // [P]      float32 row sum of the un-normalised G~ row (dsm.py:192)
    int32_t*  pix_dist;    // [P]      chessboard distance to the sub-sample grid (dsm.py:169-179)
    float*    gval;        // [28 P]   G~ values: cell (block at GROW * first pixel) -> pixel -> active slot (rows of na4 floats, zero-padded)
    int32_t*  cell_start;  // [C]      first pixel of each cell (cell_cap + 1 entries per candidate)
    uint32_t* cell_amask;  // [C]      bit s: neighbour cell s holds a grid point with a non-zero weight in this cell
    int32_t*  cellmap;     // [C]      grid index of the cell's grid point or -1
    uint32_t* grid_rc;     // [C]      grid points, row-major order = column order of G~
    uint32_t* grid_tmp;    // [C]
    int32_t*  cell_list;   // [C]      the non-empty cells of each candidate (CInfo.n_ne entries)
    double*   xtab;        // [T]      normalised coordinates x1 of compacted rows, then x2 of compacted columns (image.py:6-21)
    uint16_t* inv;         // [T]      compacted row -> box row, then compacted col -> box col
    uint16_t* fwd;         // [T]      box row -> compacted row (0xFFFF: empty), then columns
    CInfo*    cinfo;       // [n_cand]
    CState*   cstate;      // [n_cand]
    EvalReq*  req;         // [n_cand]
    size_t    sg_stride, xig_stride;   // doubles between the SPEC_MAX copies of scrG / xig (one per trial point of a request)
    double*   scrG;        // [C * 32]   per-cell value ([31]) + theta gradient ([0..5]) of the last evaluation
    double*   xig;         // [C * 25]   xi gradient of the last evaluation: entry [j * 25 + s] = contribution to grid point j of the
                           //            cell at neighbour position s of j's cell (zero where there is none; zeroed per batch)
    double*   scrH;        // [C * 496]  per-cell packed local Hessian of the last evaluation
    double*   vecs;        // [V]      NVEC vectors of length 6 + cell_cap per candidate
    double*   mats;        // [M]      H, H2 (N x N) and the packed Cholesky factor per candidate
    // device-side scheduler (sched.cuh): qc = control tasks (consumed by whole CTAs), qe = evaluation and Hessian-row items
    // (consumed by single warps)
    Queue     qc, qe;
    int32_t*  pending;     // [n_cand] items of the candidate's current stage that have not finished yet
    int32_t*  prio;        // [n_cand] ring the candidate's tasks go to (1: its chain is on the critical path of the batch)
    unsigned long long* glob;       // [N_GLOB * HOT_STRIDE] candidates done / abort flag, one cache line each
    unsigned long long* counters;   // [N_COUNTERS]  busy-cycle / task statistics (sched.cuh)
    int32_t   n_cand;
};

struct Layout {
    size_t total;
    size_t o_pix_y, o_pix_rc, o_pix_s, o_pix_dist, o_gval, o_cell_start, o_cell_amask, o_cellmap, o_grid_rc, o_grid_tmp, o_cell_list,
           o_xtab, o_inv, o_fwd, o_cinfo, o_cstate, o_req, o_scrG, o_xig, o_scrH, o_vecs, o_mats, o_qc_task, o_qc_seq, o_qc_ev, o_qc_hot, o_qe_task, o_qe_seq, o_qe_ev, o_qe_hot, o_pending, o_prio, o_glob, o_counters;
    int64_t C, qc_cap, qe_cap;
    int32_t n_cand;
};

inline size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }
constexpr int N_GLOB_LINES = 4;   // >= N_GLOB (common.cuh, queue section): cache lines of the solver's global flags

// P, C, T, V, M: totals of the per-pixel, per-cell, per-table, per-vector and per-matrix arrays of the batch
// K = upper bound of the number of tasks that can be in flight (sdsm_plan_batch: sum over candidates of the tasks of one stage)
inline Layout make_layout(int64_t P, int64_t C, int64_t T, int64_t V, int64_t M, int64_t K, int32_t n_cand)
{
    Layout L{};
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes); return at; };
    L.o_pix_y = take(sizeof(double) * P);
    L.o_pix_rc = take(sizeof(uint32_t) * P);
    L.o_pix_s = take(sizeof(float) * P);
    L.o_pix_dist = take(sizeof(int32_t) * P);
    L.o_gval = take(sizeof(float) * GROW * P);
    L.o_cell_start = take(sizeof(int32_t) * C);
    L.o_cell_amask = take(sizeof(uint32_t) * C);
    L.o_cellmap = take(sizeof(int32_t) * C);
    L.o_grid_rc = take(sizeof(uint32_t) * C);
    L.o_grid_tmp = take(sizeof(uint32_t) * C);
    L.o_cell_list = take(sizeof(int32_t) * C);
    L.o_xtab = take(sizeof(double) * T);
    L.o_inv = take(sizeof(uint16_t) * T);
    L.o_fwd = take(sizeof(uint16_t) * T);
    L.o_cinfo = take(sizeof(CInfo) * (size_t)n_cand);
    L.o_cstate = take(sizeof(CState) * (size_t)n_cand);
    L.o_req = take(sizeof(EvalReq) * (size_t)n_cand);
    L.o_scrG = take(sizeof(double) * SCRG * C * SPEC_MAX);
    L.o_xig = take(sizeof(double) * NSLOT * C * SPEC_MAX);
    L.o_scrH = take(sizeof(double) * SCRH * C);
    L.o_vecs = take(sizeof(double) * V);
    L.o_mats = take(sizeof(double) * M);
    int64_t qc_cap = 1024, qe_cap = 1024;
    while (qc_cap < n_cand + 1024) qc_cap <<= 1;
    while (qe_cap < K + 1024) qe_cap <<= 1;
    L.o_qc_task = take(sizeof(uint2) * 2 * qc_cap);
    L.o_qc_seq = take(sizeof(uint32_t) * 2 * qc_cap);
    L.o_qc_ev = take(sizeof(uint32_t) * 2 * qc_cap);
    L.o_qc_hot = take(sizeof(unsigned long long) * 8 * 16);
    L.o_qe_task = take(sizeof(uint2) * 2 * qe_cap);
    L.o_qe_seq = take(sizeof(uint32_t) * 2 * qe_cap);
    L.o_qe_ev = take(sizeof(uint32_t) * 2 * qe_cap);
    L.o_qe_hot = take(sizeof(unsigned long long) * 8 * 16);
    L.o_pending = take(sizeof(int32_t) * (size_t)n_cand);
    L.o_prio = take(sizeof(int32_t) * (size_t)n_cand);
    L.o_glob = take(sizeof(unsigned long long) * N_GLOB_LINES * 16);
    L.o_counters = take(sizeof(unsigned long long) * N_COUNTERS);
    L.C = C;
    L.qc_cap = qc_cap;
    L.qe_cap = qe_cap;
    L.n_cand = n_cand;
    L.total = o;
    return L;
}

inline WS make_ws(void* base, const Layout& L)
{
    char* b = (char*)base;
    WS w;
    w.pix_y = (double*)(b + L.o_pix_y);
    w.pix_rc = (uint32_t*)(b + L.o_pix_rc);
    w.pix_s = (float*)(b + L.o_pix_s);
    w.pix_dist = (int32_t*)(b + L.o_pix_dist);
    w.gval = (float*)(b + L.o_gval);
    w.cell_start = (int32_t*)(b + L.o_cell_start);
    w.cell_amask = (uint32_t*)(b + L.o_cell_amask);
    w.cellmap = (int32_t*)(b + L.o_cellmap);
    w.grid_rc = (uint32_t*)(b + L.o_grid_rc);
    w.grid_tmp = (uint32_t*)(b + L.o_grid_tmp);
    w.cell_list = (int32_t*)(b + L.o_cell_list);
    w.xtab = (double*)(b + L.o_xtab);
    w.inv = (uint16_t*)(b + L.o_inv);
    w.fwd = (uint16_t*)(b + L.o_fwd);
    w.cinfo = (CInfo*)(b + L.o_cinfo);
    w.cstate = (CState*)(b + L.o_cstate);
    w.req = (EvalReq*)(b + L.o_req);
    w.sg_stride = (size_t)SCRG * (size_t)L.C; w.xig_stride = (size_t)NSLOT * (size_t)L.C;
    w.scrG = (double*)(b + L.o_scrG);
    w.xig = (double*)(b + L.o_xig);
    w.scrH = (double*)(b + L.o_scrH);
    w.vecs = (double*)(b + L.o_vecs);
    w.mats = (double*)(b + L.o_mats);
    w.qc.task = (uint2*)(b + L.o_qc_task); w.qc.seq = (uint32_t*)(b + L.o_qc_seq); w.qc.ev = (uint32_t*)(b + L.o_qc_ev);
    w.qc.hot = (unsigned long long*)(b + L.o_qc_hot); w.qc.mask = (uint32_t)(L.qc_cap - 1); w.qc.ev_mask = (uint32_t)(2 * L.qc_cap - 1);
    w.qe.task = (uint2*)(b + L.o_qe_task); w.qe.seq = (uint32_t*)(b + L.o_qe_seq); w.qe.ev = (uint32_t*)(b + L.o_qe_ev);
    w.qe.hot = (unsigned long long*)(b + L.o_qe_hot); w.qe.mask = (uint32_t)(L.qe_cap - 1); w.qe.ev_mask = (uint32_t)(2 * L.qe_cap - 1);
    w.pending = (int32_t*)(b + L.o_pending);
    w.prio = (int32_t*)(b + L.o_prio);
    w.glob = (unsigned long long*)(b + L.o_glob);
    w.counters = (unsigned long long*)(b + L.o_counters);
    w.n_cand = L.n_cand;
    return w;
}

// ---------------------------------------------------------------------------------------------------------------
// block-level helpers (fixed reduction trees: results do not depend on scheduling)

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Sum of K values per thread over the block; result in out[0..K) (shared), visible to all threads on return.
template <int BLOCK, int K>
__device__ __forceinline__ void block_sum(double (&v)[K], double* s_part /* [BLOCK/32][K] */, double* out /* [K] */)
{
    constexpr int NW = BLOCK / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        double r = warp_sum(v[k]);
        if (lane == 0) s_part[warp * K + k] = r;
    }
    __syncthreads();
    for (int k = threadIdx.x; k < K; k += BLOCK) {
        double r = 0;
#pragma unroll
        for (int w = 0; w < NW; ++w) r += s_part[w * K + k];
        out[k] = r;
    }
    __syncthreads();
}

template <int BLOCK>
__device__ __forceinline__ double block_sum1(double v, double* s_part, double* out)
{
    double a[1] = {v};
    block_sum<BLOCK, 1>(a, s_part, out);
    return out[0];
}

template <int BLOCK>
__device__ __forceinline__ int block_sum_int(int v, int* s_part /* [BLOCK/32 + 1] */)
{
    constexpr int NW = BLOCK / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if (lane == 0) s_part[warp] = v;
    __syncthreads();
    if (threadIdx.x == 0) { int r = 0; for (int w = 0; w < NW; ++w) r += s_part[w]; s_part[NW] = r; }
    __syncthreads();
    return s_part[NW];
}

template <int BLOCK>
__device__ __forceinline__ unsigned long long block_min_u64(unsigned long long v, unsigned long long* s_part /* [BLOCK/32 + 1] */)
{
    constexpr int NW = BLOCK / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { unsigned long long w = __shfl_xor_sync(0xffffffffu, v, o); v = w < v ? w : v; }
    __syncthreads();
    if (lane == 0) s_part[warp] = v;
    __syncthreads();
    if (threadIdx.x == 0) { unsigned long long r = s_part[0]; for (int w = 1; w < NW; ++w) r = s_part[w] < r ? s_part[w] : r; s_part[NW] = r; }
    __syncthreads();
    return s_part[NW];
}

__device__ __forceinline__ bool fp_member(int lab, const int32_t* fp, int n)
{
    if (lab == 0) return false;
    for (int k = 0; k < n; ++k) if (fp[k] == lab) return true;
    return false;
}

// ---------------------------------------------------------------------------------------------------------------
// Device-side task queues (struct Queue): two bounded MPMC rings of tasks (ring 1 = urgent, served first) behind one ring
// of arrival events.  A candidate's tasks all go to the ring ws.prio[c] names: the candidates whose chains bound the
// batch's makespan (the largest regions, and chains that turn out long) are not held up behind the bulk of the batch.
//
// Every ring uses 64-bit tickets and per-slot sequence numbers: slot t & mask is free for the producer of ticket t when
// its sequence number equals (uint32) t, and full for the consumer of ticket t when it equals (uint32) t + 1; the consumer
// hands it on with t + capacity.  The rings are sized above the number of tasks that can be in flight (sdsm_plan_batch),
// so producers only ever wait for a consumer that has claimed its ticket and is about to read the slot.
//
// A producer publishes the task in its ring first, raises the ring's availability count, and then publishes one arrival
// event.  An idle consumer takes an event ticket and sleeps on ITS OWN event slot (no shared line is polled); once the
// event is there, at least one task is waiting for it in one of the rings -- not necessarily the one whose arrival woke
// it up -- and it takes the oldest urgent task if there is one, else the oldest normal task.  Everything is wait-free
// (fetch-and-add only): compare-and-swap loops on the heads collapse when a few hundred consumers wake up together.
// The hot counters sit on cache lines of their own (stride HOT_STRIDE).

enum { Q_HEAD = 0, Q_TAIL = 1, Q_DONE = 2, Q_ABORT = 3, Q_CYC_EVAL = 4, Q_CYC_HG, Q_CYC_CTRL, Q_CYC_IDLE, Q_N_EVAL, Q_N_HG, Q_N_CTRL, Q_CYC_TOTAL, Q_T0,
       Q_CYC_WIDLE = 13 /* idle cycles of the evaluation warps */, Q_CYC_WTOTAL = 14 /* lifetime cycles of the evaluation warps */, Q_CYC_WFIN = 15 /* cycles of the evaluation warps in finish_item */,
       Q_CYC_KIND = 16 /* [4]: evaluation-item cycles by kind = use_xi * 2 + want_h */, Q_N_KIND = 20 /* [4]: items by kind */,
       Q_HEAD_HI = 24, Q_TAIL_HI = 25, Q_N_PROMOTED = 26, Q_CYC_CTRL_BIG = 27 /* control tasks of systems > SMALL_N_MAX */, Q_N_CTRL_BIG = 28, Q_CYC_CHOL_BIG = 29, Q_N_CHOL_BIG = 30, Q_CYC_SOLVE_BIG = 31 /* first of the two solves per iteration */ };
constexpr int HOT_STRIDE = 16;   // unsigned long longs = 128 bytes
enum { H_EV_HEAD = 0, H_EV_TAIL = 1, H_HEAD_LO = 2, H_TAIL_LO = 3, H_HEAD_HI = 4, H_TAIL_HI = 5, H_AVAIL_LO = 6, H_AVAIL_HI = 7, N_HOT = 8 };
enum { G_DONE = 0, G_ABORT = 1, G_PROGRESS = 2 /* globaltimer of the last finished control task */, N_GLOB = 3 };
enum { T_CTRL = 0, T_EVAL = 1, T_HG = 2 };
constexpr uint32_t T_SHIFT = 28, T_CMASK = (1u << T_SHIFT) - 1u;

__device__ __forceinline__ unsigned long long* hot(const Queue& q, int which) { return q.hot + which * HOT_STRIDE; }

__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p)
{
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u32(uint32_t* p, uint32_t v)
{
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ bool q_aborted(const WS& ws) { return ld_relaxed_u64(ws.glob + G_ABORT * HOT_STRIDE) != 0ull; }
__device__ __forceinline__ void q_abort(const WS& ws) { atomicExch(ws.glob + G_ABORT * HOT_STRIDE, 1ull); }
__device__ __forceinline__ unsigned long long q_done(const WS& ws) { return ld_relaxed_u64(ws.glob + G_DONE * HOT_STRIDE); }

// Watchdog: the solver is aborted when NO control task has finished anywhere for `timeout_ns` (a lost task, a chain that
// never ends) -- not when a long but healthy batch simply takes longer than that.
__device__ __forceinline__ void q_progress(const WS& ws)
{
    unsigned long long now;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(ws.glob + G_PROGRESS * HOT_STRIDE), "l"(now) : "memory");
}
__device__ __forceinline__ bool q_stalled(const WS& ws, unsigned long long t_start, unsigned long long timeout_ns)
{
    unsigned long long now;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
    const unsigned long long last = ld_relaxed_u64(ws.glob + G_PROGRESS * HOT_STRIDE);
    const unsigned long long since = last > t_start ? last : t_start;
    return now > since && now - since > timeout_ns;
}

// one thread: publishes one task under ticket `t` of `ring` and its arrival event under event ticket `e`
__device__ __forceinline__ void q_put(const WS& ws, const Queue& q, int ring, unsigned long long t, unsigned long long e, uint32_t type, uint32_t c, uint32_t first)
{
    const uint32_t slot = ((uint32_t)t & q.mask) + (ring ? q.mask + 1u : 0u);
    while (ld_acquire_u32(q.seq + slot) != (uint32_t)t) {
        if (q_aborted(ws)) return;
        __nanosleep(32);
    }
    q.task[slot] = make_uint2((type << T_SHIFT) | c, first);
    st_release_u32(q.seq + slot, (uint32_t)t + 1u);
    __threadfence();
    atomicAdd(hot(q, ring ? H_AVAIL_HI : H_AVAIL_LO), 1ull);
    const uint32_t es = (uint32_t)e & q.ev_mask;
    while (ld_acquire_u32(q.ev + es) != (uint32_t)e) {
        if (q_aborted(ws)) return;
        __nanosleep(32);
    }
    st_release_u32(q.ev + es, (uint32_t)e + 1u);
}

constexpr uint32_t PT_SHIFT = 24;          // evaluation items: task.y = trial point << PT_SHIFT | first cell
// whole CTA: publishes `count` tasks (type, c, first = k * stride for k < count; with per_pt: point k / per_pt, first (k % per_pt) * stride).  Everything the CTA wrote before the call
// is visible to the consumers of the tasks.  s_base: two unsigned long longs of shared memory.
template <int BLOCK>
__device__ __forceinline__ void q_push_range(const WS& ws, const Queue& q, int ring, uint32_t type, uint32_t c, int count, int stride, unsigned long long* s_base,
                                             int per_pt = 0x7fffffff)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        s_base[0] = atomicAdd(hot(q, ring ? H_TAIL_HI : H_TAIL_LO), (unsigned long long)count);
        s_base[1] = atomicAdd(hot(q, H_EV_TAIL), (unsigned long long)count);
    }
    __syncthreads();
    const unsigned long long base = s_base[0], ebase = s_base[1];
    for (int k = threadIdx.x; k < count; k += BLOCK) q_put(ws, q, ring, base + k, ebase + k, type, c, ((uint32_t)(k / per_pt) << PT_SHIFT) | (uint32_t)((k % per_pt) * stride));
}

// whole warp: the same for a warp (everything the WARP wrote before the call is visible to the consumers)
__device__ __forceinline__ void q_push_range_warp(const WS& ws, const Queue& q, int ring, uint32_t type, uint32_t c, int count, int stride)
{
    const int lane = threadIdx.x & 31;
    unsigned long long base = 0, ebase = 0;
    __threadfence();
    __syncwarp();
    if (lane == 0) {
        base = atomicAdd(hot(q, ring ? H_TAIL_HI : H_TAIL_LO), (unsigned long long)count);
        ebase = atomicAdd(hot(q, H_EV_TAIL), (unsigned long long)count);
    }
    base = __shfl_sync(0xffffffffu, base, 0);
    ebase = __shfl_sync(0xffffffffu, ebase, 0);
    for (int k = lane; k < count; k += 32) q_put(ws, q, ring, base + k, ebase + k, type, c, (uint32_t)(k * stride));
    __syncwarp();
}

// one thread: publishes a single task
__device__ __forceinline__ void q_push_one(const WS& ws, const Queue& q, int ring, uint32_t type, uint32_t c, uint32_t first)
{
    const unsigned long long t = atomicAdd(hot(q, ring ? H_TAIL_HI : H_TAIL_LO), 1ull);
    const unsigned long long e = atomicAdd(hot(q, H_EV_TAIL), 1ull);
    q_put(ws, q, ring, t, e, type, c, first);
}

// one thread: takes the next task of the queue, urgent ring first.  Returns false when every candidate is finished, or when
// the watchdog fired: an idle consumer that sees no control task finish anywhere for `timeout_ns` sets the abort flag and everybody leaves.
__device__ __forceinline__ bool q_pop(const WS& ws, const Queue& q, unsigned long long n_target, unsigned long long timeout_ns,
                                      unsigned long long t_start, uint2* task)
{
    // (1) wait for an arrival event of our own
    const unsigned long long e = atomicAdd(hot(q, H_EV_HEAD), 1ull);
    const uint32_t es = (uint32_t)e & q.ev_mask;
    unsigned ns = 32, spins = 0;
    while (ld_acquire_u32(q.ev + es) != (uint32_t)e + 1u) {
        if (q_done(ws) >= n_target) return false;
        if (q_aborted(ws)) return false;
        if ((++spins & 1023u) == 0 && q_stalled(ws, t_start, timeout_ns)) { q_abort(ws); return false; }
        __nanosleep(ns);
        if (ns < 512) ns <<= 1;
    }
    st_release_u32(q.ev + es, (uint32_t)e + q.ev_mask + 1u);
    // (2) a task is waiting for us: the oldest urgent one if there is any, else the oldest normal one.  H_AVAIL_* count the
    //     published tasks nobody has a right to yet: whoever decrements a positive count owns the ring's next ticket.
    for (;;) {
#pragma unroll
        for (int ring = 1; ring >= 0; --ring) {
            unsigned long long* avail = hot(q, ring ? H_AVAIL_HI : H_AVAIL_LO);
            if ((long long)ld_relaxed_u64(avail) <= 0) continue;
            if ((long long)atomicAdd(avail, ~0ull) <= 0) { atomicAdd(avail, 1ull); continue; }     // lost the race: give it back
            const unsigned long long h = atomicAdd(hot(q, ring ? H_HEAD_HI : H_HEAD_LO), 1ull);
            const uint32_t slot = ((uint32_t)h & q.mask) + (ring ? q.mask + 1u : 0u);
            while (ld_acquire_u32(q.seq + slot) != (uint32_t)h + 1u) {             // an older ticket still being published
                if (q_aborted(ws)) return false;
                __nanosleep(20);
            }
            *task = q.task[slot];
            st_release_u32(q.seq + slot, (uint32_t)h + q.mask + 1u);
            return true;
        }
        if (q_aborted(ws)) return false;
    }
}

}  // namespace sdsm
