// K3: k_solve -- the persistent solver kernel.  Every candidate's fit is a chain of dependent steps
//     control task -> evaluation items (one per 1-4 pixel cells) [-> Hessian-row items] -> control task -> ...
// (a few dozen to a few thousand links, objects.py:361-412 around cvxopt's cp), and the chains of a batch differ in length
// and width by orders of magnitude.  Instead of lock-step launches, one persistent grid pulls work from device-side
// queues (common.cuh): whoever finishes the last item of a candidate's stage publishes its next stage, so every candidate
// advances at its own pace.
//
// The CTAs have one of two roles.  Control CTAs (the first n_ctrl_ctas) take control tasks from ws.qc, one CTA per task:
// the interior-point bookkeeping, Cholesky factorisations and triangular solves need a CTA's worth of threads and
// barriers.  The WARPS of all other CTAs are independent consumers of ws.qe: an evaluation item is a few pixel cells, a
// Hessian-row item a few grid points, each handled by one warp from start to end -- no CTA-wide barrier is ever waited
// on in the evaluation path, so a warp with a short item never idles behind a long one.
// Candidates with few cells run their whole chain inside one control CTA (no queue hop between the links).
//
// Nothing ever blocks on another task: a task only waits for a queue slot whose previous occupant has already been
// claimed by a consumer.  The kernel ends when every candidate has reported completion (G_DONE == n_target), or when the
// watchdog fires (G_ABORT; sdsm_solver_stats reports it).
#pragma once
#include "common.cuh"
#include "eval.cuh"
#include "ctrl.cuh"

namespace sdsm {

constexpr size_t SOLVE_SMEM_MIN = EVAL_SMEM > BIG_SMEM ? EVAL_SMEM : BIG_SMEM;
static_assert(sizeof(double) * (size_t)(SMEM_L_CAP + CTRL_BPAD + CTRL_WARPS * 32 + 32) + sizeof(int) * 8 + CTRL_WARPS * sizeof(CState) <= EVAL_SMEM,
              "the shared-memory Cholesky must not enlarge the CTA's shared memory beyond what the evaluation warps need (2 CTAs per SM)");
__host__ __device__ inline size_t solve_smem_bytes(int lcap)
{
    const size_t a = ctrl_smem_bytes(lcap);
    return a > SOLVE_SMEM_MIN ? a : SOLVE_SMEM_MIN;
}

struct SolveArgs {
    int n_target;            // candidates that must finish
    int lcap;                // shared-memory capacity (doubles) for a packed Cholesky factor
    int solo_cells;          // candidates with at most this many non-empty cells run their chain inside one CTA
    int task_small_cells;    // candidates with at most this many non-empty cells get one cell per evaluation item
    int hi_evals;            // a candidate that has needed more evaluations of F than this becomes urgent (ring 1)
    int spec_max;            // trial points per line-search request (1: none speculative)
    int spec_retry;          // points of the second request of a line search whose first request had one point
    int spec_solo;           // chains inside one control CTA speculate too
    int spec_skip;           // first-pass line searches that do not depend on F run without evaluations
    int spec_hess;           // a trial that will probably be accepted is evaluated with its Hessian
    int spec_prio;           // of the chains that go through the queue, only urgent ones (ring 1) make requests on an unconfident prediction
    int n_ctrl_ctas;         // CTAs [0, n_ctrl_ctas) run control tasks, the warps of the others evaluation / Hessian-row items
    unsigned long long timeout_ns;   // watchdog: the kernel is aborted when no control task has finished for this long
    double* trace;           // diagnostics (STG_TRACE) or nullptr
};

__device__ __forceinline__ unsigned long long globaltimer_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

__device__ __forceinline__ void seed_queue(const Queue& q, uint32_t i, uint32_t stride, int n_lo, int n_hi)
{
    const uint32_t cap = q.mask + 1u;
    for (; i < 2u * cap; i += stride) {
        q.ev[i] = (int)i < n_lo + n_hi ? i + 1u : i;                     // one arrival event per seeded task
        if (i >= cap) continue;
        q.seq[cap + i] = (int)i < n_hi ? i + 1u : i;
        q.seq[i] = (int)i < n_lo ? i + 1u : i;
    }
}

// Seeds the queues: the control tasks that start the candidates order[0 .. n_first) (most expensive first); the first n_hi
// of them are urgent from the start (ring 1), the others go to ring 0.  The evaluation queue starts empty.
__global__ void k_seed(WS ws, const int32_t* __restrict__ order, int n_first, int single, int n_hi)
{
    const uint32_t i0 = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
    seed_queue(ws.qc, i0, stride, n_first - n_hi, n_hi);
    seed_queue(ws.qe, i0, stride, 0, 0);
    const uint32_t cap = ws.qc.mask + 1u;
    for (int i = i0; i < n_first; i += stride) {
        const int c = order ? order[i] : single;
        ws.prio[c] = i < n_hi ? 1 : 0;
        ws.qc.task[i < n_hi ? cap + i : i - n_hi] = make_uint2(((uint32_t)T_CTRL << T_SHIFT) | (uint32_t)c, 0u);
    }
    if (i0 == 0) {
        for (int k = 0; k < N_COUNTERS; ++k) ws.counters[k] = 0ull;
        for (int k = 0; k < N_HOT * HOT_STRIDE; ++k) { ws.qc.hot[k] = 0ull; ws.qe.hot[k] = 0ull; }
        for (int k = 0; k < N_GLOB * HOT_STRIDE; ++k) ws.glob[k] = 0ull;
        *hot(ws.qc, H_TAIL_LO) = (unsigned long long)(n_first - n_hi);
        *hot(ws.qc, H_TAIL_HI) = (unsigned long long)n_hi;
        *hot(ws.qc, H_EV_TAIL) = (unsigned long long)n_first;
        *hot(ws.qc, H_AVAIL_LO) = (unsigned long long)(n_first - n_hi);
        *hot(ws.qc, H_AVAIL_HI) = (unsigned long long)n_hi;
        ws.counters[Q_T0] = globaltimer_ns();
    }
}

// one warp: the evaluation item (c, first) = cells cell_list[first .. first + cpw) of candidate c's pending request
__device__ __forceinline__ void eval_item(const Par& par, const sdsm_cand* __restrict__ cands, const WS& ws, int c, int first, int pt, double* Tt, int* slotj)
{
    const int count = min(ws.req[c].cpw, ws.cinfo[c].n_ne - first);
    for (int i = 0; i < count; ++i) eval_cell(par, cands, ws, c, first + i, pt, Tt, slotj);
}

// the warp that finished an item of candidate c's current stage: the last one to finish starts the next stage
__device__ __forceinline__ void finish_item(const WS& ws, uint32_t type, int c)
{
    const int lane = threadIdx.x & 31;
    __threadfence();
    __syncwarp();
    int next = 0;                                            // 0: not the last, 1: control task, > 1: that many T_HG items
    if (lane == 0) {
        const int old = atomicSub(ws.pending + c, 1);
        if (old == 1) {
            __threadfence();                                 // the other items' results, before the next stage is published
            const EvalReq rq = ws.req[c];
            next = 1;
            if (type == T_EVAL && rq.want_h && rq.use_xi) {
                const int nt = (ws.cinfo[c].m + HG_ITEM_ROWS - 1) / HG_ITEM_ROWS;
                ws.pending[c] = nt;
                next = 1 + nt;
            }
        }
    }
    next = __shfl_sync(0xffffffffu, next, 0);
    if (next == 0) return;
    const int ring = ws.prio[c];
    if (next == 1) { if (lane == 0) q_push_one(ws, ws.qc, ring, T_CTRL, (uint32_t)c, 0u); }
    else q_push_range_warp(ws, ws.qe, ring, T_HG, (uint32_t)c, next - 1, HG_ITEM_ROWS);
}

__global__ void __launch_bounds__(SOLVE_BLOCK, 2)
k_solve(const __grid_constant__ ImgSet imgs, Par par, const sdsm_cand* __restrict__ cands, WS ws, sdsm_out out, SolveArgs args)
{
    constexpr int BLOCK = SOLVE_BLOCK;
    extern __shared__ __align__(16) char smem_solve[];
    __shared__ __align__(16) int s_slotj[EVAL_WARPS * WARP_INTS];
    __shared__ uint2 s_task;
    __shared__ int s_flag, s_next;
    __shared__ unsigned long long s_base[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned long long n_target = (unsigned long long)args.n_target;
    const unsigned long long t_start = globaltimer_ns();
    const long long c_begin = clock64();
    eval_sync_init(s_slotj);

    if ((int)blockIdx.x >= args.n_ctrl_ctas) {
        // ---- evaluation warps: independent consumers of ws.qe ----
        double* Tt = (double*)smem_solve + warp * TILE_DOUBLES;      // TILE_DOUBLES >= HG_ACC
        int* slotj = s_slotj + warp * WARP_INTS;
        // statistics: four counters in registers (anything indexed dynamically would live in local memory)
        unsigned long long cyc_idle = 0, cyc_eval = 0, cyc_hg = 0;
        unsigned n_popped = 0, n_eval = 0, n_hg = 0;
        for (;;) {
            uint2 task = make_uint2(0u, 0u);
            int ok = 0;
            const long long c0 = clock64();
            if (lane == 0) {
                if ((++n_popped & 255u) == 0 && q_stalled(ws, t_start, args.timeout_ns)) q_abort(ws);
                ok = q_pop(ws, ws.qe, n_target, args.timeout_ns, t_start, &task) ? 1 : 0;
            }
            ok = __shfl_sync(0xffffffffu, ok, 0);
            if (!ok) break;
            task.x = __shfl_sync(0xffffffffu, task.x, 0);
            task.y = __shfl_sync(0xffffffffu, task.y, 0);
            const long long c1 = clock64();
            cyc_idle += (unsigned long long)(c1 - c0);
            const uint32_t type = task.x >> T_SHIFT;
            const int c = (int)(task.x & T_CMASK);
            if (type == T_EVAL) {
                eval_item(par, cands, ws, c, (int)(task.y & ((1u << PT_SHIFT) - 1u)), (int)(task.y >> PT_SHIFT), Tt, slotj);
                finish_item(ws, type, c);
                cyc_eval += (unsigned long long)(clock64() - c1); n_eval += 1;
            } else {
                const int m = ws.cinfo[c].m;
                for (int j = (int)task.y; j < (int)task.y + HG_ITEM_ROWS && j < m; ++j) { hg_row(par, cands, ws, c, j, Tt); __syncwarp(); }
                finish_item(ws, type, c);
                cyc_hg += (unsigned long long)(clock64() - c1); n_hg += 1;
            }
        }
        if (lane == 0) {
            atomicAdd(ws.counters + Q_CYC_EVAL, cyc_eval); atomicAdd(ws.counters + Q_N_EVAL, (unsigned long long)n_eval);
            atomicAdd(ws.counters + Q_CYC_HG, cyc_hg); atomicAdd(ws.counters + Q_N_HG, (unsigned long long)n_hg);
            atomicAdd(ws.counters + Q_CYC_WIDLE, cyc_idle);
            atomicAdd(ws.counters + Q_CYC_WTOTAL, (unsigned long long)(clock64() - c_begin));
        }
        return;
    }

    // ---- control CTAs: one control task at a time, from ws.qc ----
    unsigned long long cyc_ctrl = 0, cyc_idle = 0, n_ctrl = 0, n_promoted = 0, cyc_big = 0, n_big = 0;
    unsigned n_popped = 0;
    for (;;) {
        if (tid == 0) {
            const long long c0 = clock64();
            uint2 task;
            if ((++n_popped & 255u) == 0 && q_stalled(ws, t_start, args.timeout_ns)) q_abort(ws);
            const bool ok = q_pop(ws, ws.qc, n_target, args.timeout_ns, t_start, &task);
            cyc_idle += (unsigned long long)(clock64() - c0);
            s_task = task;
            s_flag = ok ? 1 : 0;
            s_next = 0;
        }
        __syncthreads();
        const uint2 task = s_task;
        const bool ok = s_flag != 0;
        __syncthreads();
        if (!ok) break;
        const int c = (int)(task.x & T_CMASK);
        const long long c0 = clock64();
        {
            const sdsm_cand cd = cands[c];
            const Img img = imgs.v[cd.img];
            const CInfo ci = ws.cinfo[c];
            for (unsigned link = 1;; ++link) {
                if ((link & 63u) == 0) {                    // watchdog of a chain that never leaves this CTA
                    if (tid == 0) s_flag = q_aborted(ws) ? 1 : 0;       // (a chain inside one CTA makes progress by definition: every link is a finished control task)
                    __syncthreads();
                    const bool stop = s_flag != 0;
                    __syncthreads();
                    if (stop) { if (tid == 0) q_abort(ws); break; }
                }
                const unsigned long long t_link = tid == 0 ? globaltimer_ns() : 0ull;
                const bool big = (ws.cstate[c].use_xi ? 6 + ci.m : 6) > SMALL_N_MAX;
                Sm sm;
                bool requested;
                if (big) carve_smem_big(sm, smem_solve); else carve_smem(sm, smem_solve, args.lcap);
                // (the same criterion as `solo` below, before the step: only a heuristic, the results do not depend on it)
                const bool solo_before = ci.n_ne <= EVAL_WARPS || (ci.n_ne <= args.solo_cells && ws.cstate[c].nf + ws.cstate[c].nh <= args.hi_evals);
                // chains inside this CTA: the extra points run on warps of this CTA that a single point leaves idle; chains that go through
                // the queue: extra points compete with everybody's evaluations, so (spec_prio) only the urgent ones get them
                const int spec_max = (solo_before && !args.spec_solo) ? 1 : args.spec_max;
                const int spec_eager = solo_before || !args.spec_prio || ws.prio[c];
                requested = ctrl_step<BLOCK>(par, img, cd, ci, ws, c, sm, out, args.trace, args.task_small_cells, spec_max, spec_eager, args.spec_retry, args.spec_hess, args.spec_skip, big);
                if (tid == 0) { ws.cstate[c].ns_ctrl += globaltimer_ns() - t_link; q_progress(ws); }
                if (!requested) {
                    if (tid == 0) {
                        ws.pending[c] = (int32_t)((globaltimer_ns() - ws.counters[Q_T0]) / 1000ull);   // diagnostics: finish time, us
                        ws.prio[c] = (int32_t)(ws.cstate[c].ns_ctrl / 1000ull);                        // diagnostics: time in control tasks, us
                        __threadfence();
                        atomicAdd(ws.glob + G_DONE * HOT_STRIDE, 1ull);
                    }
                    break;
                }
                // Short chains of small candidates stay in this CTA (no queue hops).  A chain that turns out long leaves it:
                // it then finishes after the bulk of the batch, when the evaluation warps are idle and spreading its cells
                // over them makes every link several times shorter.
                const bool solo = ci.n_ne <= EVAL_WARPS || (ci.n_ne <= args.solo_cells && ws.cstate[c].nf + ws.cstate[c].nh <= args.hi_evals);
                if (!solo) {
                    // publishing is the LAST thing a control task does: the candidate's next control task may start on another
                    // SM as soon as the evaluations have finished
                    const int cpw = ws.req[c].cpw;
                    const int per_pt = (ci.n_ne + cpw - 1) / cpw;
                    const int n_items = per_pt * ws.req[c].npts;            // the cells of every trial point of the request
                    if (tid == 0) {
                        ws.pending[c] = n_items;
                        // a chain that turns out long becomes urgent: it would otherwise finish long after the bulk of the batch
                        if (!ws.prio[c]) { const CState* S = ws.cstate + c; if (S->nf + S->nh > args.hi_evals) { ws.prio[c] = 1; ++n_promoted; } }
                    }
                    __syncthreads();
                    q_push_range<BLOCK>(ws, ws.qe, ws.prio[c], T_EVAL, (uint32_t)c, n_items, cpw, s_base, per_pt);
                    break;
                }
                // the whole chain in this CTA: evaluate here and go on
                const EvalReq rq = ws.req[c];
                const int cpt = rq.cpt;
                const int n_tasks = (ci.n_ne + cpt - 1) / cpt;
                for (int k = 0; k < n_tasks; ++k) {
                    if (tid == 0) s_next = 0;
                    __syncthreads();
                    eval_task(par, cands, ws, c, k * cpt, (double*)smem_solve, s_slotj, &s_next, rq.npts);
                    __syncthreads();
                }
                if (rq.want_h && rq.use_xi)
                    for (int j = 0; j < ci.m; j += HG_WARPS) { hg_task(par, cands, ws, c, j, (double*)smem_solve); __syncthreads(); }
            }
        }
        if (tid == 0) {
            const unsigned long long dc = (unsigned long long)(clock64() - c0);
            cyc_ctrl += dc; n_ctrl += 1;
            if (6 + ws.cinfo[c].m > SMALL_N_MAX) { cyc_big += dc; n_big += 1; }
        }
    }
    if (tid == 0) {
        atomicAdd(ws.counters + Q_CYC_CTRL, cyc_ctrl); atomicAdd(ws.counters + Q_CYC_IDLE, cyc_idle); atomicAdd(ws.counters + Q_N_CTRL, n_ctrl);
        atomicAdd(ws.counters + Q_CYC_TOTAL, (unsigned long long)(clock64() - c_begin));
        if (n_promoted) atomicAdd(ws.counters + Q_N_PROMOTED, n_promoted);
        atomicAdd(ws.counters + Q_CYC_CTRL_BIG, cyc_big); atomicAdd(ws.counters + Q_N_CTRL_BIG, n_big);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// diagnostics for the parity tests (single candidate of the last solved batch; plain grid-stride kernels, no queue)

// prepares the evaluation of F (with Hessian, unscaled) of candidate c at xin
__global__ void __launch_bounds__(SOLVE_BLOCK)
k_diag_eval_begin(Par par, const sdsm_cand* __restrict__ cands, int c, WS ws, const double* __restrict__ xin, int want_h)
{
    const int tid = threadIdx.x;
    const sdsm_cand cd = cands[c];
    const CInfo ci = ws.cinfo[c];
    Prob P;
    setup_problem(P, par, cd, ci, ws);
    const int N = P.N;
    double* nx = P.V + (size_t)V_NX * P.ld;
    int bad = 0;
    for (int i = tid; i < N; i += SOLVE_BLOCK) { nx[i] = xin[i]; bad |= !isfinite(xin[i]); }
    for (size_t i = tid; i < (size_t)N * N; i += SOLVE_BLOCK) P.H[i] = 0.0;
    bad = __syncthreads_or(bad);
    if (tid == 0) {
        EvalReq rq; rq.x = nx; rq.hfac = 1.0; rq.want_h = want_h; rq.use_xi = P.m > 0 ? 1 : 0; rq.cpt = EVAL_CELLS; rq.cpw = 1; rq.npts = 1; rq.ld = 0;
        ws.req[c] = rq;
        ws.cstate[c].pc = bad ? -1 : PC_DONE;
    }
}

__global__ void __launch_bounds__(SOLVE_BLOCK)
k_diag_eval_cells(Par par, const sdsm_cand* __restrict__ cands, int c, WS ws)
{
    extern __shared__ __align__(16) char smem_solve[];
    __shared__ __align__(16) int s_slotj[EVAL_WARPS * WARP_INTS];
    __shared__ int s_next;
    eval_sync_init(s_slotj);
    const int n_ne = ws.cinfo[c].n_ne;
    for (int first = blockIdx.x * EVAL_CELLS; first < n_ne; first += gridDim.x * EVAL_CELLS) {
        if (threadIdx.x == 0) s_next = 0;
        __syncthreads();
        eval_task(par, cands, ws, c, first, (double*)smem_solve, s_slotj, &s_next);
        __syncthreads();
    }
}

__global__ void __launch_bounds__(SOLVE_BLOCK)
k_diag_eval_rows(Par par, const sdsm_cand* __restrict__ cands, int c, WS ws)
{
    extern __shared__ __align__(16) char smem_solve[];
    const int m = ws.cinfo[c].m;
    for (int first = blockIdx.x * HG_WARPS; first < m; first += gridDim.x * HG_WARPS) {
        hg_task(par, cands, ws, c, first, (double*)smem_solve);
        __syncthreads();
    }
}

__global__ void __launch_bounds__(SOLVE_BLOCK)
k_diag_eval_end(Par par, const sdsm_cand* __restrict__ cands, int c, WS ws, double* value, double* grad, double* Hout, int lcap)
{
    extern __shared__ __align__(16) char smem_solve[];
    Sm sm;
    carve_smem(sm, smem_solve, lcap);
    const int tid = threadIdx.x;
    const sdsm_cand cd = cands[c];
    const CInfo ci = ws.cinfo[c];
    Prob P;
    setup_problem(P, par, cd, ci, ws);
    P.scale_cp = 1.0;
    const int N = P.N;
    double* g = P.V + (size_t)V_NG * P.ld;
    const double* xe = P.V + (size_t)V_NX * P.ld;
    double psi = 0;
    double f;
    if (ws.cstate[c].pc == -1) f = NAN;
    else if (P.m > 0) f = Hout ? gather_cells<SOLVE_BLOCK, true>(P, sm, xe, g, 1.0, &psi) : gather_cells<SOLVE_BLOCK, false>(P, sm, xe, g, 1.0, &psi);
    else f = Hout ? gather_poly<SOLVE_BLOCK, true>(P, sm, g, 1.0, &psi) : gather_poly<SOLVE_BLOCK, false>(P, sm, g, 1.0, &psi);
    __syncthreads();
    if (tid == 0) *value = f;
    for (int i = tid; i < N; i += SOLVE_BLOCK) grad[i] = g[i];
    if (Hout) for (size_t i = tid; i < (size_t)N * N; i += SOLVE_BLOCK) Hout[i] = P.H[i];
}

// prepares one interior-point solve of candidate c from x0 (STG_TRACE): k_solve continues it at PC_IPMSTART and records
// the iterates
__global__ void __launch_bounds__(SOLVE_BLOCK)
k_diag_trace_begin(Par par, const sdsm_cand* __restrict__ cands, int c, int use_xi_, WS ws, const double* __restrict__ x0)
{
    const int tid = threadIdx.x;
    const sdsm_cand cd = cands[c];
    const CInfo ci = ws.cinfo[c];
    Prob P;
    setup_problem(P, par, cd, ci, ws);
    const int use_xi = (use_xi_ && P.m > 0) ? 1 : 0;
    const int N = use_xi ? P.N : 6;
    double* x = P.V + (size_t)V_X * P.ld;
    for (int i = tid; i < N; i += SOLVE_BLOCK) x[i] = x0[i];
    if (use_xi) for (size_t i = tid; i < (size_t)N * N; i += SOLVE_BLOCK) P.H[i] = 0.0;
    __syncthreads();
    if (tid == 0) {
        CState S = ws.cstate[c];
        S.use_xi = use_xi; S.stage = STG_TRACE; S.r_ipm = 0; S.pc = PC_TRACESTART;
        ws.cstate[c] = S;
    }
}

__global__ void k_diag_trace_end(Par par, const sdsm_cand* __restrict__ cands, int c, WS ws, double* xout, int32_t* status_iters)
{
    const sdsm_cand cd = cands[c];
    const CState* S = ws.cstate + c;
    const int ld = 6 + (par.has_smooth ? cd.cell_cap : 0);
    const int N = S->use_xi ? 6 + ws.cinfo[c].m : 6;
    const double* x = ws.vecs + cd.vec_off + (size_t)V_X * ld;
    for (int i = threadIdx.x; i < N; i += blockDim.x) xout[i] = x[i];
    if (threadIdx.x == 0) { status_iters[0] = S->r_ipm; status_iters[1] = S->iters; }
}

// ---------------------------------------------------------------------------------------------------------------
// measurement aid (bench.py, roofline of the evaluation tasks): one sweep of the evaluation tasks over EVERY candidate of
// the last solved batch at its solution, outside the solver.  k_sweep_prep lists the tasks, k_sweep runs them.

__global__ void __launch_bounds__(SOLVE_BLOCK)
k_sweep_prep(Par par, const sdsm_cand* __restrict__ cands, int n_cand, WS ws, int want_h, int rows)
{
    __shared__ unsigned long long s_base;
    for (int c = blockIdx.x; c < n_cand; c += gridDim.x) {
        const sdsm_cand cd = cands[c];
        const CInfo ci = ws.cinfo[c];
        const bool live = !ci.flags && !(ci.n_fg == 1 && par.skip_single_fg) && ws.cstate[c].pc == PC_DONE;
        const int count = !live ? 0 : (rows ? (ci.m + HG_WARPS - 1) / HG_WARPS : (ci.n_ne + EVAL_CELLS - 1) / EVAL_CELLS);
        __syncthreads();
        if (threadIdx.x == 0) {
            const int ld = 6 + (par.has_smooth ? cd.cell_cap : 0);
            EvalReq rq; rq.x = ws.vecs + cd.vec_off + (size_t)V_X * ld; rq.hfac = 1.0; rq.want_h = want_h; rq.use_xi = ci.m > 0 ? 1 : 0; rq.cpt = EVAL_CELLS; rq.cpw = 1; rq.npts = 1; rq.ld = 0;
            ws.req[c] = rq;
            s_base = atomicAdd(ws.counters + Q_TAIL, (unsigned long long)count);
        }
        __syncthreads();
        const unsigned long long base = s_base;
        for (int k = threadIdx.x; k < count; k += SOLVE_BLOCK)
            ws.qe.task[base + k] = make_uint2((uint32_t)c, (uint32_t)(k * (rows ? HG_WARPS : EVAL_CELLS)));
    }
}

__global__ void __launch_bounds__(SOLVE_BLOCK, 2)
k_sweep(Par par, const sdsm_cand* __restrict__ cands, WS ws, int rows)
{
    extern __shared__ __align__(16) char smem_solve[];
    __shared__ __align__(16) int s_slotj[EVAL_WARPS * WARP_INTS];
    __shared__ int s_next;
    eval_sync_init(s_slotj);
    const unsigned long long n_tasks = ws.counters[Q_TAIL];
    for (unsigned long long i = blockIdx.x; i < n_tasks; i += gridDim.x) {
        const uint2 task = ws.qe.task[i];
        if (threadIdx.x == 0) s_next = 0;
        __syncthreads();
        if (rows) hg_task(par, cands, ws, (int)task.x, (int)task.y, (double*)smem_solve);
        else eval_task(par, cands, ws, (int)task.x, (int)task.y, (double*)smem_solve, s_slotj, &s_next);
        __syncthreads();
    }
}

}  // namespace sdsm
