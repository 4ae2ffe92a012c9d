// K3b: control tasks -- one CTA advances one candidate between two evaluations of F.  It gathers the per-cell results of the
// evaluation tasks into value / gradient / Hessian, advances the candidate's state machine
//   elliptical initialisation (objects.py:321-358) -> DSM solve (objects.py:394-410, dsm.py:431-490) -> result
// up to the next point where F must be evaluated, files that request (evaluation tasks, sched.cuh) and returns.
// The interior-point iterations restate cvxopt.solvers.cp (oracle/cpsolver.py, which this mirrors line by line); the
// local variables of that loop live in CState across tasks.
#pragma once
#include "common.cuh"
#include "eval.cuh"

namespace sdsm {

constexpr int MAX_RELAXED = 8;
constexpr int EVAL_BUDGET = 50000;   // evaluations of F per candidate (a regular solve needs < 4000: 100 iterations, few trials each)
constexpr double CP_STEP = 0.99, CP_BETA = 0.5, CP_ALPHA = 0.01;

enum { V_X = 0, V_G, V_DX, V_T, V_NX, V_NG, V_SX, V_SDX, V_RD, V_X0, V_E, V_W,
       V_SPX /* SPEC_MAX line-search trial points */, V_SPG = V_SPX + SPEC_MAX /* and their gradients */ };
static_assert(V_SPG + SPEC_MAX == NVEC, "vector count");
enum { PC_START = 0, PC_TOP, PC_RESTORED, PC_TRIAL, PC_VAL1, PC_VALINIT, PC_JX, PC_J0, PC_FINAL, PC_DONE, PC_TRACESTART };
enum { STG_ELL1 = 0, STG_ELL2, STG_DSM, STG_TRACE };

struct Prob {
    int n, m, N, ncells, cw, ch, ld;
    const double* py; const uint32_t* prc;
    const int32_t* cstart; const uint32_t* amask; const int32_t* cellmap; const uint32_t* grid; const int32_t* clist;
    int n_ne;
    const double* xr; const double* xc;
    double scale_cp, alpha, eps, tol, hess_tol;
    double *V, *H, *H2, *Lg;
    const double *scrG, *scrH, *xig;
    int sub;
};

struct Sm {
    double* Lf;        // [lcap + CTRL_BPAD] packed Cholesky factor when it fits (else it lives in global memory), followed by /
                       // replaced by the right-hand side of the triangular solves
    double* red_part;  // [CTRL_WARPS * 32]
    double* red_out;   // [32]
    int*    flag;      // [8]
    int     lcap;      // doubles available for the factor
    CState* cst;       // [CTRL_WARPS] the candidate's state while a control task runs, one copy per warp (see ctrl_step)
    double *D, *Xj, *yv;   // large systems only: diagonal block, panel block, solved block (see cholesky_big)
    double *Wt;            // large systems only: one CB x CB_LD tile per warp (coalesced staging of global rows)
};
constexpr int CTRL_BPAD = 160;     // >= SMALL_N_MAX: the right-hand side of the triangular solves sits behind the factor
__host__ __device__ inline size_t ctrl_smem_bytes(int lcap) { return sizeof(double) * (size_t)(lcap + CTRL_BPAD + CTRL_WARPS * 32 + 32) + sizeof(int) * 8 + CTRL_WARPS * sizeof(CState); }
// factor capacity (doubles) a batch whose largest system has n_max unknowns asks for: the whole packed factor if n_max <= 104
static_assert(CTRL_BPAD >= SMALL_N_MAX, "the solves' right-hand side must fit behind the factor");
__host__ __device__ inline int ctrl_lcap(int n_max) { const int need = n_max * (n_max + 1) / 2; return need <= SMEM_L_CAP ? (need < 64 ? 64 : need) : SMEM_L_CAP; }

__device__ __forceinline__ void carve_smem(Sm& sm, char* base, int lcap)
{
    sm.Lf = (double*)base;
    sm.red_part = sm.Lf + lcap + CTRL_BPAD;
    sm.red_out = sm.red_part + CTRL_WARPS * 32;
    sm.flag = (int*)(sm.red_out + 32);
    sm.cst = (CState*)(sm.flag + 8);
    sm.lcap = lcap;
    sm.D = sm.Xj = sm.yv = sm.Wt = nullptr;
}

__device__ __forceinline__ void setup_problem(Prob& P, const Par& par, const sdsm_cand& cd, const CInfo& ci, const WS& ws)
{
    P.n = cd.n; P.m = ci.m; P.N = 6 + ci.m; P.ncells = ci.ncells; P.cw = ci.cw; P.ch = ci.ch;
    P.ld = 6 + (par.has_smooth ? cd.cell_cap : 0);
    P.py = ws.pix_y + cd.pix_off; P.prc = ws.pix_rc + cd.pix_off;
    P.cstart = ws.cell_start + cd.cell_off; P.amask = ws.cell_amask + cd.cell_off; P.cellmap = ws.cellmap + cd.cell_off;
    P.grid = ws.grid_rc + cd.cell_off; P.clist = ws.cell_list + cd.cell_off; P.n_ne = ci.n_ne;
    P.xr = ws.xtab + cd.tab_off; P.xc = P.xr + cd.h;
    P.scale_cp = par.scale / (double)cd.n;                                   // objects.py:380
    P.alpha = par.alpha; P.eps = par.epsilon; P.tol = par.sparsity_tol; P.hess_tol = par.hess_tol; P.sub = par.sub;
    P.V = ws.vecs + cd.vec_off;
    const size_t nn = (size_t)P.ld * P.ld;
    P.H = ws.mats + cd.mat_off; P.H2 = P.H + nn; P.Lg = P.H2 + nn;
    P.scrG = ws.scrG + (size_t)cd.cell_off * SCRG; P.scrH = ws.scrH + (size_t)cd.cell_off * SCRH;
    P.xig = ws.xig + (size_t)cd.cell_off * NSLOT;
}

// slot index (0..24) of the cell at offset (da, db) from a cell, and the local index of that slot in the cell's vector
__device__ __forceinline__ int local_index(uint32_t am, int da, int db)
{
    const int s = (da + NC) * 5 + (db + NC);
    if (!((am >> s) & 1u)) return -1;
    return 6 + __popc(am & ((1u << s) - 1u));
}

// ---------------------------------------------------------------------------------------------------------------
// Gathers the per-cell results of k_eval.  g: scaled gradient out; H (if HESS) = hz * scale_cp * Hess psi, full symmetric,
// row stride N.  Returns the scaled value; *psi_out = unscaled psi.  Fixed summation order.

template <int BLOCK, bool HESS>
__device__ double gather_poly(const Prob& P, Sm& sm, double* g, double hz, double* psi_out)
{
    const int tid = threadIdx.x;
    constexpr int K = HESS ? 28 : 7;
    double acc[K];
#pragma unroll
    for (int k = 0; k < K; ++k) acc[k] = 0;
    for (int i = tid; i < P.n_ne; i += BLOCK) {
        const int cell = P.clist[i];
        const double* sg = P.scrG + (size_t)cell * SCRG;
        acc[0] += sg[31];
#pragma unroll
        for (int k = 0; k < 6; ++k) acc[1 + k] += sg[k];
        if (HESS) {
            const double* sh = P.scrH + (size_t)cell * SCRH;
#pragma unroll
            for (int k = 0; k < 21; ++k) acc[7 + k] += sh[k];
        }
    }
    block_sum<BLOCK, K>(acc, sm.red_part, sm.red_out);
    const double psi = sm.red_out[0];
    const double sc = P.scale_cp;
    if (tid < 6) g[tid] = sc * sm.red_out[1 + tid];
    if (HESS && tid < 21) {
        int a = 0, t = tid;
        while (t >= a + 1) { t -= a + 1; ++a; }
        const double v = hz * sc * sm.red_out[7 + tid];
        P.H[a * 6 + t] = v;
        P.H[t * 6 + a] = v;
    }
    __syncthreads();
    *psi_out = psi;
    return sc * psi;
}

template <int BLOCK, bool HESS>
__device__ double gather_cells(const Prob& P, Sm& sm, const double* x, double* g, double hz, double* psi_out)
{
    const int tid = threadIdx.x;
    const int m = P.m, N = P.N, cw = P.cw, ch = P.ch;
    const double sc = P.scale_cp;
    __syncthreads();
    // (1) value, theta gradient, 6x6 block, regulariser: sums over all cells / grid points
    constexpr int K = HESS ? 29 : 8;
    double acc[K];
#pragma unroll
    for (int k = 0; k < K; ++k) acc[k] = 0;
    for (int i = tid; i < P.n_ne; i += BLOCK) {
        const int cell = P.clist[i];
        const double* sg = P.scrG + (size_t)cell * SCRG;
        acc[0] += sg[31];
#pragma unroll
        for (int k = 0; k < 6; ++k) acc[1 + k] += sg[k];
        if (HESS) {
            const double* sh = P.scrH + (size_t)cell * SCRH;
#pragma unroll
            for (int k = 0; k < 21; ++k) acc[8 + k] += sh[k];
        }
    }
    for (int j = tid; j < m; j += BLOCK) { const double v = x[6 + j]; acc[7] += sqrt(v * v + P.eps); }
    block_sum<BLOCK, K>(acc, sm.red_part, sm.red_out);
    double obj2 = P.alpha * sm.red_out[7] - P.alpha * sqrt(P.eps) * (double)m;   // dsm.py:325-331
    if (obj2 < 0) obj2 = 0;
    const double psi = sm.red_out[0] + obj2;
    if (tid < 6) g[tid] = sc * sm.red_out[1 + tid];
    if (HESS && tid < 21) {
        int a = 0, t = tid;
        while (t >= a + 1) { t -= a + 1; ++a; }
        double u = sm.red_out[8 + tid];
        if (a != t && fabs(u) < P.hess_tol) u = 0.0;                           // dsm.py:376-382 (the diagonal is kept)
        const double v = hz * sc * u;
        P.H[a * N + t] = v;
        P.H[t * N + a] = v;
    }
    // (2) xi gradient: the evaluation tasks left the contributions of the (up to 25) cells around grid point j in 25
    //     contiguous slots (zero where there is no cell); fixed summation order
    for (int j = tid; j < m; j += BLOCK) {
        const double* r = P.xig + (size_t)j * NSLOT;
        double v[NSLOT];
#pragma unroll
        for (int k = 0; k < NSLOT; ++k) v[k] = r[k];
        double s = 0;
#pragma unroll
        for (int k = 0; k < NSLOT; ++k) s += v[k];
        const double xv = x[6 + j];
        s += P.alpha * (xv / sqrt(xv * xv + P.eps));                             // dsm.py:350
        g[6 + j] = sc * s;
    }
    // the xi-theta and xi-xi blocks of the Hessian are assembled by the T_HG tasks (one warp per grid point), before this task runs
    __syncthreads();
    *psi_out = psi;
    return sc * psi;
}

// Hessian-only evaluations (EvalReq.want_h == 2): only the theta-theta block is left to gather (the other rows are assembled by
// the T_HG tasks); value and gradient at this point are the accepted line-search trial's.
template <int BLOCK>
__device__ void gather_hessian_block(const Prob& P, Sm& sm, double hz, int N, double prune)
{
    const int tid = threadIdx.x;
    double acc[21];
#pragma unroll
    for (int k = 0; k < 21; ++k) acc[k] = 0;
    for (int i = tid; i < P.n_ne; i += BLOCK) {
        const double* sh = P.scrH + (size_t)P.clist[i] * SCRH;
#pragma unroll
        for (int k = 0; k < 21; ++k) acc[k] += sh[k];
    }
    __syncthreads();
    block_sum<BLOCK, 21>(acc, sm.red_part, sm.red_out);
    if (tid < 21) {
        int a = 0, t = tid;
        while (t >= a + 1) { t -= a + 1; ++a; }
        double u = sm.red_out[tid];
        if (a != t && fabs(u) < prune) u = 0.0;
        const double v = hz * P.scale_cp * u;
        P.H[a * N + t] = v;
        P.H[t * N + a] = v;
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------------------------
// T_HG tasks: rows 6.. of the Hessian (the xi-theta and xi-xi blocks, dsm.py:369-374) of a candidate that asked for one.
// One warp per grid point j: it visits the <= 25 cells that see j in a fixed order and adds their local
// Hessian rows into 81 + 6 accumulators (the grid points of the 9 x 9 cells around j's cell, and theta), then writes row
// 6 + j of H = hz * scale_cp * Hess psi (both triangles; entries outside the 9 x 9 neighbourhood stay zero).

constexpr int HG_BLOCK = SOLVE_BLOCK;
constexpr int HG_WARPS = HG_BLOCK / 32;
constexpr int HG_ACC = 96;
constexpr int HG_ITEM_ROWS = 2;      // grid points per queued Hessian-row item (one warp)
constexpr int WARP_ITEM_CELLS = 4;   // cells per queued evaluation item of a large candidate (one warp)

// one warp: row 6 + j of candidate c's Hessian.  acc: HG_ACC doubles of shared memory of the warp's own.
__device__ __forceinline__ void hg_row(const Par& par, const sdsm_cand* __restrict__ cands, const WS& ws, int c, int j, double* acc)
{
    const int lane = threadIdx.x & 31;
    constexpr int SP = 4 * NC + 1;   // 9
    const sdsm_cand* cd = cands + c;
    const int64_t cell_off = cd->cell_off;
    const CInfo ci = ws.cinfo[c];
    if (j >= ci.m) return;
    const int cw = ci.cw, ch = ci.ch, N = 6 + ci.m;
    const EvalReq rq = ws.req[c];
    const double fac = rq.hfac;
    const int32_t* cstart = ws.cell_start + cell_off;
    const uint32_t* amask = ws.cell_amask + cell_off;
    const int32_t* cellmap = ws.cellmap + cell_off;
    const double* scrH = ws.scrH + (size_t)cell_off * SCRH;
    double* H = ws.mats + cd->mat_off;
    const uint32_t gp = ws.grid_rc[cell_off + j];
    const int ja = (int)(gp & 0xffffu) / par.sub, jb = (int)(gp >> 16) / par.sub;
    __syncwarp();
    for (int t = lane; t < HG_ACC; t += 32) acc[t] = 0.0;
    __syncwarp();
    for (int da = -NC; da <= NC; ++da) {
        const int a = ja - da;                       // cell = j's cell - (da, db): it sees j in slot (da, db)
        if (a < 0 || a >= ch) continue;
        for (int db = -NC; db <= NC; ++db) {
            const int b = jb - db;
            if (b < 0 || b >= cw) continue;
            const int cell = a * cw + b;
            if (cstart[cell + 1] == cstart[cell]) continue;
            const uint32_t am = amask[cell];
            const int lj = local_index(am, da, db);
            if (lj < 0) continue;
            const int L = 6 + __popc(am);
            if (lane < L) {
                const double v = scrH[(size_t)cell * SCRH + (lane >= lj ? tri(lane, lj) : tri(lj, lane))];
                int t;
                if (lane < 6) t = SP * SP + lane;
                else {
                    const int s = __fns(am, 0, lane - 6 + 1);           // slot of local entry `lane`
                    t = (s / 5 - NC - da + 2 * NC) * SP + (s % 5 - NC - db + 2 * NC);
                }
                acc[t] += v;
            }
            __syncwarp();
        }
    }
    const double* x = rq.x + (size_t)(rq.npts - 1) * rq.ld;          // the point the Hessian is evaluated at: the last of the request
    for (int t = lane; t < SP * SP; t += 32) {
        const int ka = ja + t / SP - 2 * NC, kb = jb + t % SP - 2 * NC;
        if (ka < 0 || ka >= ch || kb < 0 || kb >= cw) continue;
        const int k = cellmap[ka * cw + kb];
        if (k < 0) continue;
        double s = acc[t];
        if (k == j) {
            const double v = x[6 + j];
            const double t3 = v * v, t2 = sqrt(t3 + par.epsilon);
            double gd = par.alpha * (1.0 / t2 - t3 / (t2 * t2 * t2));      // dsm.py:372-374
            if (gd < 0) gd = 0;
            s += gd;
        }
        else if (fabs(s) < par.hess_tol) s = 0.0;                            // dsm.py:376-382: small off-diagonal entries dropped
        H[(size_t)(6 + j) * N + 6 + k] = s * fac;
    }
    if (lane < 6) {
        double s = acc[SP * SP + lane];
        if (fabs(s) < par.hess_tol) s = 0.0;
        s *= fac;
        H[(size_t)(6 + j) * N + lane] = s;
        H[(size_t)lane * N + 6 + j] = s;
    }
}

// CTA-wide task (c, first): the CTA's warps take the grid points first .. first + HG_WARPS of candidate c (chains that stay in
// one CTA, diagnostics).  s_acc: HG_WARPS * HG_ACC doubles.
__device__ __forceinline__ void hg_task(const Par& par, const sdsm_cand* __restrict__ cands, const WS& ws, int c, int first, double* s_acc)
{
    const int warp = threadIdx.x >> 5;
    hg_row(par, cands, ws, c, first + warp, s_acc + warp * HG_ACC);
}

// ---------------------------------------------------------------------------------------------------------------
// dense Cholesky of the lower triangle of H (N x N, row-major, ld = N) into packed L; the diagonal of L is stored as its
// reciprocal (the triangular solves only multiply by it).
// Fails (returns false) on a non-positive or non-finite pivot, as lapack.potrf / cholmod.numeric do.

template <int BLOCK>
__device__ bool cholesky(const double* H, int N, double* L, Sm& sm)
{
    constexpr int NW = BLOCK / 32;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    __syncthreads();
    for (int i = warp; i < N; i += NW)
        for (int j = lane; j <= i; j += 32) L[tri(i, j)] = H[(size_t)i * N + j];
    if (tid == 0) sm.flag[0] = 0;
    __syncthreads();
    for (int k = 0; k < N; ++k) {
        if (tid == 0) {
            const double dk = L[tri(k, k)];
            if (!(dk > 0.0) || !isfinite(dk)) sm.flag[0] = 1;
            else L[tri(k, k)] = 1.0 / sqrt(dk);
        }
        __syncthreads();
        if (sm.flag[0]) return false;
        const double rk = L[tri(k, k)];
        for (int i = k + 1 + tid; i < N; i += BLOCK) L[tri(i, k)] *= rk;
        __syncthreads();
        for (int i = k + 1 + warp; i < N; i += NW) {
            const double lik = L[tri(i, k)];
            double* row = L + tri(i, 0);
            for (int j = k + 1 + lane; j <= i; j += 32) row[j] = fma(-lik, L[tri(j, k)], row[j]);
        }
        __syncthreads();
    }
    return true;
}

// b := (L L^T)^-1 b, by warp 0.  Callers synchronise the block before and after.
__device__ void chol_solve_warp(const double* L, int N, double* b)
{
    const int lane = threadIdx.x & 31;
    for (int k = 0; k < N; ++k) {
        const double yk = b[k] * L[tri(k, k)];
        __syncwarp();
        for (int i = k + 1 + lane; i < N; i += 32) b[i] = fma(-L[tri(i, k)], yk, b[i]);
        if (lane == 0) b[k] = yk;
        __syncwarp();
    }
    for (int k = N - 1; k >= 0; --k) {
        const double xk = b[k] * L[tri(k, k)];
        __syncwarp();
        const double* row = L + tri(k, 0);
        for (int i = lane; i < k; i += 32) b[i] = fma(-row[i], xk, b[i]);
        if (lane == 0) b[k] = xk;
        __syncwarp();
    }
}

// v := (L L^T)^-1 v for a vector in global memory: staged through shared memory (bv, capacity bvcap) when it fits
template <int BLOCK>
__device__ void chol_solve(const double* L, int N, double* v, double* bv, int bvcap)
{
    const int tid = threadIdx.x;
    __syncthreads();
    if (N <= bvcap) {
        for (int i = tid; i < N; i += BLOCK) bv[i] = v[i];
        __syncthreads();
        if (tid < 32) chol_solve_warp(L, N, bv);
        __syncthreads();
        for (int i = tid; i < N; i += BLOCK) v[i] = bv[i];
    } else {
        if (tid < 32) chol_solve_warp(L, N, v);
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------------------------
// Large systems (N > SMALL_N_MAX): blocked right-looking Cholesky of the lower triangle of H into A (N x N row-major,
// global memory / L2), 32-wide panels staged in shared memory; same conventions (reciprocal diagonal, failure on a
// non-positive pivot).  One CTA of BLOCK threads.

constexpr int CB = 32;               // panel width
constexpr int CB_LD = CB + 1;      // row stride of tiles accessed by rows and by columns
constexpr int TL = CB + 4;         // row stride of tiles read as DMMA fragments (= 4 mod 16: conflict-free)

constexpr size_t BIG_SMEM = sizeof(double) * ((2 + BIG_BLOCK / 32) * CB * TL + CB + (BIG_BLOCK / 32) * 32 + 32) + sizeof(int) * 8 + (BIG_BLOCK / 32) * sizeof(CState);

__device__ __forceinline__ void carve_smem_big(Sm& sm, char* base)
{
    sm.D = (double*)base;
    sm.Xj = sm.D + CB * TL;
    sm.Wt = sm.Xj + CB * TL;
    sm.yv = sm.Wt + (BIG_BLOCK / 32) * CB * TL;
    sm.red_part = sm.yv + CB;
    sm.red_out = sm.red_part + (BIG_BLOCK / 32) * 32;
    sm.flag = (int*)(sm.red_out + 32);
    sm.cst = (CState*)(sm.flag + 8);
    sm.Lf = nullptr;
    sm.lcap = 0;
}

// Left-looking by 32-column panels: the rows of panel k first receive the contributions of all earlier panels
//   A[i, k-panel] -= sum_j A[i, j-panel] A[k-rows, j-panel]^T
// as 32 x 32 x 32 tile products on the FP64 tensor cores (DMMA.8x8x4), then the diagonal block is factorised by warp 0 and
// the rows below are solved against it.  Global memory (the factor lives in L2) is only touched with coalesced rows, staged
// through shared-memory tiles: two tiles of A[k-rows, .] shared by the CTA per step, one tile per warp for its own rows.
// (The right-looking version this replaces spent 1-3 ms per factorisation at N = 300-450 waiting for L2 round trips.)
#ifdef SDSM_CHOL_TRACE
__device__ int chol_trace_count = 0;
#define CHOL_T(k) { const long long now_ = clock64(); ct_[k] += now_ - cl_; cl_ = now_; }
#else
#define CHOL_T(k)
#endif

template <int BLOCK>
__device__ bool cholesky_big(const double* H, int N, double* A, Sm& sb)
{
#ifdef SDSM_CHOL_TRACE
    long long ct_[5] = {0, 0, 0, 0, 0}, cl_ = clock64();
#endif
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = BLOCK / 32;
    const int fr = lane >> 2, fk = lane & 3;                 // DMMA fragment coordinates (eval.cuh::dmma884)
    double* W = sb.Wt + warp * CB * TL;
    __syncthreads();
    // No copy H -> A up front: whatever touches a panel first (its first update step, or for panel 0 the diagonal block and
    // the solve below it) reads H and writes A.
    if (tid == 0) sb.flag[0] = 0;
    __syncthreads();
    CHOL_T(0);
    for (int k0 = 0; k0 < N; k0 += CB) {
        const int kw = min(CB, N - k0);
        const int nrb = (N - k0 + CB - 1) / CB;              // row blocks of the panel; block 0 holds the diagonal block
        // (1) contributions of the earlier panels, two at a time
        for (int j0 = 0; j0 < k0; j0 += 2 * CB) {
            const int nt = (j0 + CB < k0) ? 2 : 1;
            __syncthreads();                                 // the tiles of the previous step are no longer in use
            {
                // all loads first, then the stores (the compiler may not reorder them itself: it cannot rule out aliasing)
                constexpr int PER = 2 * CB * CB / BLOCK;
                double v[PER];
#pragma unroll
                for (int u = 0; u < PER; ++u) {
                    const int e = tid + u * BLOCK;
                    const int t = e / (CB * CB), r = (e / CB) % CB, q = e % CB;
                    v[u] = (e < nt * CB * CB && r < kw) ? A[(size_t)(k0 + r) * N + j0 + t * CB + q] : 0.0;
                }
#pragma unroll
                for (int u = 0; u < PER; ++u) {
                    const int e = tid + u * BLOCK;
                    const int t = e / (CB * CB), r = (e / CB) % CB, q = e % CB;
                    if (e < nt * CB * CB) (t ? sb.Xj : sb.D)[r * TL + q] = v[u];
                }
            }
            __syncthreads();
            for (int ib = warp; ib < nrb; ib += NW) {
                const int i0 = k0 + ib * CB, nr = min(CB, N - i0);
                double c[4][4][2];
#pragma unroll
                for (int I = 0; I < 4; ++I)
#pragma unroll
                    for (int J = 0; J < 4; ++J) { c[I][J][0] = 0.0; c[I][J][1] = 0.0; }
                for (int t = 0; t < nt; ++t) {
                    const double* T = t ? sb.Xj : sb.D;
                    __syncwarp();
#pragma unroll
                    for (int rb = 0; rb < CB; rb += 16) {
                        double v[16];
#pragma unroll
                        for (int u = 0; u < 16; ++u) v[u] = rb + u < nr ? A[(size_t)(i0 + rb + u) * N + j0 + t * CB + lane] : 0.0;
#pragma unroll
                        for (int u = 0; u < 16; ++u) W[(rb + u) * TL + lane] = v[u];
                    }
                    __syncwarp();
#pragma unroll 2
                    for (int kk = 0; kk < 8; ++kk) {
                        double af[4], bf[4];
#pragma unroll
                        for (int X = 0; X < 4; ++X) { af[X] = W[(8 * X + fr) * TL + 4 * kk + fk]; bf[X] = T[(8 * X + fr) * TL + 4 * kk + fk]; }
#pragma unroll
                        for (int I = 0; I < 4; ++I)
#pragma unroll
                            for (int J = 0; J < 4; ++J) dmma884(c[I][J][0], c[I][J][1], af[I], bf[J]);
                    }
                }
                // subtract: staged through the warp's tile so that the rows are updated with coalesced accesses
                __syncwarp();
#pragma unroll
                for (int I = 0; I < 4; ++I)
#pragma unroll
                    for (int J = 0; J < 4; ++J) { W[(8 * I + fr) * TL + 8 * J + 2 * fk] = c[I][J][0]; W[(8 * I + fr) * TL + 8 * J + 2 * fk + 1] = c[I][J][1]; }
                __syncwarp();
                if (lane < kw) {
#pragma unroll
                    for (int rb = 0; rb < CB; rb += 16) {
                        double v[16];
#pragma unroll
                        for (int u = 0; u < 16; ++u) { const int r = rb + u; v[u] = (r < nr && k0 + lane <= i0 + r) ? (j0 == 0 ? H : A)[(size_t)(i0 + r) * N + k0 + lane] : 0.0; }
#pragma unroll
                        for (int u = 0; u < 16; ++u) { const int r = rb + u; if (r < nr && k0 + lane <= i0 + r) A[(size_t)(i0 + r) * N + k0 + lane] = v[u] - W[r * TL + lane]; }   // lower triangle only
                    }
                }
            }
        }
        __syncthreads();
        CHOL_T(1);
        // (2) diagonal block -> shared memory, factorised by warp 0 (lane = row)
        {
            const double* S0 = k0 == 0 ? H : A;
            double v[CB * CB / BLOCK];
#pragma unroll
            for (int u = 0; u < CB * CB / BLOCK; ++u) { const int e = tid + u * BLOCK, r = e / CB, q = e % CB; v[u] = (r < kw && q <= r) ? S0[(size_t)(k0 + r) * N + k0 + q] : 0.0; }
#pragma unroll
            for (int u = 0; u < CB * CB / BLOCK; ++u) { const int e = tid + u * BLOCK, r = e / CB, q = e % CB; if (r < kw && q <= r) sb.D[r * CB_LD + q] = v[u]; }
        }
        __syncthreads();
        {
            // right-looking on the shared-memory block, the whole CTA per column: thread (i, s) = (tid / 8, tid % 8) updates
            // the entries j = cc + 1 + s, + 8, ... <= i of row cc + 1 + i from the UNSCALED column cc; the scaled column
            // is written after the barrier (nobody reads column cc again), so there is one barrier per column.
            const int ri = tid >> 3, rs = tid & 7;
            for (int cc = 0; cc < kw; ++cc) {
                const double dk = sb.D[cc * CB_LD + cc];
                if (!(dk > 0.0) || !isfinite(dk)) { if (tid == 0) sb.flag[0] = 1; break; }      // uniform: every thread reads the same value
                const double rinv = rsqrt(dk);
                const int i = cc + 1 + ri;
                double li = 0.0;
                if (i < kw) {
                    li = sb.D[i * CB_LD + cc] * rinv;
                    double* row = sb.D + i * CB_LD;
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int j = cc + 1 + rs + 8 * u;
                        if (j <= i) row[j] = fma(-li, sb.D[j * CB_LD + cc] * rinv, row[j]);
                    }
                }
                __syncthreads();
                if (i < kw && rs == 0) sb.D[i * CB_LD + cc] = li;
                if (tid == 0) sb.D[cc * CB_LD + cc] = rinv;
            }
        }
        __syncthreads();
        CHOL_T(2);
        if (sb.flag[0]) return false;
        for (int e = tid; e < kw * kw; e += BLOCK) { const int r = e / kw, q = e - r * kw; if (q <= r) A[(size_t)(k0 + r) * N + k0 + q] = sb.D[r * CB_LD + q]; }
        const int r0 = k0 + kw;
        if (r0 >= N) break;
        // (3) rows below the block: X = A[i, k0:k0+kw] D^-T.  A warp takes 32 rows at a time: the 32 x kw tile is read and
        //     written back row by row (lane = column: coalesced) through the warp's tile and solved with lane = row (row
        //     stride CB_LD = 33: both accesses are free of bank conflicts).
        for (int i0 = r0 + warp * CB; i0 < N; i0 += NW * CB) {
            const int nr = min(CB, N - i0);
            __syncwarp();
            {
                const double* S0 = k0 == 0 ? H : A;
#pragma unroll
                for (int rb = 0; rb < CB; rb += 16) {
                    double v[16];
#pragma unroll
                    for (int u = 0; u < 16; ++u) v[u] = (rb + u < nr && lane < kw) ? S0[(size_t)(i0 + rb + u) * N + k0 + lane] : 0.0;
#pragma unroll
                    for (int u = 0; u < 16; ++u) if (rb + u < nr) W[(rb + u) * CB_LD + lane] = v[u];
                }
            }
            __syncwarp();
            if (lane < nr) {
                double r[CB];
#pragma unroll
                for (int q = 0; q < CB; ++q) r[q] = W[lane * CB_LD + q];
#pragma unroll
                for (int q = 0; q < CB; ++q) {
                    if (q < kw) {
                        double v = r[q];
#pragma unroll
                        for (int p = 0; p < q; ++p) v = fma(-r[p], sb.D[q * CB_LD + p], v);
                        r[q] = v * sb.D[q * CB_LD + q];
                    }
                }
#pragma unroll
                for (int q = 0; q < CB; ++q) W[lane * CB_LD + q] = r[q];
            }
            __syncwarp();
            if (lane < kw) for (int r = 0; r < nr; ++r) A[(size_t)(i0 + r) * N + k0 + lane] = W[r * CB_LD + lane];
        }
        // the next panel's step (1) starts with a CTA barrier
#ifdef SDSM_CHOL_TRACE
        __syncthreads();
#endif
        CHOL_T(3);
    }
#ifdef SDSM_CHOL_TRACE
    if (tid == 0 && atomicAdd(&chol_trace_count, 1) % 500 == 0)
        printf("cholesky_big N %d: copy %lld update %lld diag %lld solve %lld cycles\n", N, ct_[0], ct_[1], ct_[2], ct_[3]);
#endif
    return true;
}

// v := (A A^T)^-1 v with the factor of cholesky_big; blocked forward / backward substitution, v in global memory.
// Per 32-block: the contribution of the part already solved is gathered by all warps (coalesced rows of A), the diagonal
// block is staged in shared memory and solved by warp 0.
template <int BLOCK>
__device__ void chol_solve_big(const double* A, int N, double* v, Sm& sb)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = BLOCK / 32;
    __syncthreads();
    // forward: L y = v
    for (int k0 = 0; k0 < N; k0 += CB) {
        const int kw = min(CB, N - k0);
        {   // diagonal block -> shared memory (loads batched in registers: one L2 round trip instead of four)
            double dv[CB * CB / BLOCK];
#pragma unroll
            for (int u = 0; u < CB * CB / BLOCK; ++u) { const int e = tid + u * BLOCK, r = e / CB, q = e % CB; dv[u] = (r < kw && q <= r) ? A[(size_t)(k0 + r) * N + k0 + q] : 0.0; }
#pragma unroll
            for (int u = 0; u < CB * CB / BLOCK; ++u) { const int e = tid + u * BLOCK, r = e / CB, q = e % CB; if (r < kw && q <= r) sb.D[r * CB_LD + q] = dv[u]; }
        }
        for (int r = warp; r < kw; r += NW) {                // y_k = v_k - L[k-rows, 0:k0] y[0:k0], a warp per row
            const double* ar = A + (size_t)(k0 + r) * N;
            double a = 0;
            double a1 = 0, a2 = 0, a3 = 0;
            int q = lane;
            for (; q + 96 < k0; q += 128) {
                const double l0 = ar[q], l1 = ar[q + 32], l2 = ar[q + 64], l3 = ar[q + 96];
                const double v0 = v[q], v1 = v[q + 32], v2 = v[q + 64], v3 = v[q + 96];
                a = fma(l0, v0, a); a1 = fma(l1, v1, a1); a2 = fma(l2, v2, a2); a3 = fma(l3, v3, a3);
            }
            for (; q < k0; q += 32) a = fma(ar[q], v[q], a);
            a = warp_sum((a + a1) + (a2 + a3));
            if (lane == 0) sb.yv[r] = v[k0 + r] - a;
        }
        __syncthreads();
        if (warp == 0) {
            double b = lane < kw ? sb.yv[lane] : 0.0;
            for (int cc = 0; cc < kw; ++cc) {
                const double yk = __shfl_sync(0xffffffffu, b, cc) * sb.D[cc * CB_LD + cc];
                if (lane == cc) b = yk;
                else if (lane > cc && lane < kw) b = fma(-sb.D[lane * CB_LD + cc], yk, b);
            }
            if (lane < kw) v[k0 + lane] = b;
        }
        __syncthreads();
    }
    // backward: L^T x = y
    for (int k0 = ((N - 1) / CB) * CB; k0 >= 0; k0 -= CB) {
        const int kw = min(CB, N - k0);
        {   // diagonal block -> shared memory (loads batched in registers: one L2 round trip instead of four)
            double dv[CB * CB / BLOCK];
#pragma unroll
            for (int u = 0; u < CB * CB / BLOCK; ++u) { const int e = tid + u * BLOCK, r = e / CB, q = e % CB; dv[u] = (r < kw && q <= r) ? A[(size_t)(k0 + r) * N + k0 + q] : 0.0; }
#pragma unroll
            for (int u = 0; u < CB * CB / BLOCK; ++u) { const int e = tid + u * BLOCK, r = e / CB, q = e % CB; if (r < kw && q <= r) sb.D[r * CB_LD + q] = dv[u]; }
        }
        {   // x_k = y_k - L[rows below, k-columns]^T x[rows below]: lane = column (coalesced), the rows dealt to the warps
            double a = 0;
            if (lane < kw) {
                double a1 = 0, a2 = 0, a3 = 0;
                int i = k0 + kw + warp;
                for (; i + 3 * NW < N; i += 4 * NW) {
                    const double l0 = A[(size_t)i * N + k0 + lane], l1 = A[(size_t)(i + NW) * N + k0 + lane];
                    const double l2 = A[(size_t)(i + 2 * NW) * N + k0 + lane], l3 = A[(size_t)(i + 3 * NW) * N + k0 + lane];
                    a = fma(l0, v[i], a); a1 = fma(l1, v[i + NW], a1); a2 = fma(l2, v[i + 2 * NW], a2); a3 = fma(l3, v[i + 3 * NW], a3);
                }
                for (; i < N; i += NW) a = fma(A[(size_t)i * N + k0 + lane], v[i], a);
                a = (a + a1) + (a2 + a3);
            }
            sb.red_part[warp * 32 + lane] = a;
        }
        __syncthreads();
        if (warp == 0) {
            double b = 0;
            if (lane < kw) {
                double a = 0;
#pragma unroll
                for (int w = 0; w < NW; ++w) a += sb.red_part[w * 32 + lane];
                b = v[k0 + lane] - a;
            }
            for (int cc = kw - 1; cc >= 0; --cc) {
                const double xk = __shfl_sync(0xffffffffu, b, cc) * sb.D[cc * CB_LD + cc];
                if (lane == cc) b = xk;
                else if (lane < cc) b = fma(-sb.D[cc * CB_LD + lane], xk, b);
            }
            if (lane < kw) v[k0 + lane] = b;
        }
        __syncthreads();
    }
}

template <int BLOCK>
__device__ double bdot(const double* a, const double* b, int N, Sm& sm)
{
    double s = 0;
    for (int i = threadIdx.x; i < N; i += BLOCK) s = fma(a[i], b[i], s);
    return block_sum1<BLOCK>(s, sm.red_part, sm.red_out);
}

// ---------------------------------------------------------------------------------------------------------------
// data-driven ellipse of objects.py:287-296 / dsm.py:96-111 (second elliptical pass).  Returns false if it raises.

template <int BLOCK>
__device__ bool estimate_initialization(const Prob& P, Sm& sm, const Img& img, const sdsm_cand& cd, const uint16_t* inv_r,
                                        const uint16_t* inv_c, double* out6)
{
    double a[3] = {0, 0, 0};
    for (int i = threadIdx.x; i < P.n; i += BLOCK) {
        if (P.py[i] > 0) {
            const uint32_t rc = P.prc[i];
            a[0] += (double)(cd.r0 + inv_r[rc & 0xffffu]);
            a[1] += (double)(cd.c0 + inv_c[rc >> 16]);
            a[2] += 1.0;
        }
    }
    block_sum<BLOCK, 3>(a, sm.red_part, sm.red_out);
    const double cnt = sm.red_out[2];
    if (!(cnt > 0)) return false;                       // center_of_mass of an empty mask: NaN -> IndexError
    const double cr = rint(sm.red_out[0] / cnt), cc = rint(sm.red_out[1] / cnt);   // np.round: half to even
    __syncthreads();
    const double zr = img.H > 1 ? (double)(img.H - 1) : 1.0, zc = img.W > 1 ? (double)(img.W - 1) : 1.0;
    const double c1 = cr / zr, c2 = cc / zc;
    double b[2] = {0, 0};
    for (int i = threadIdx.x; i < P.n; i += BLOCK) {
        if (P.py[i] > 0) {
            const uint32_t rc = P.prc[i];
            b[0] += P.xr[rc & 0xffffu] - c1;
            b[1] += P.xc[rc >> 16] - c2;
        }
    }
    block_sum<BLOCK, 2>(b, sm.red_part, sm.red_out);
    const double m1 = sm.red_out[0] / cnt, m2 = sm.red_out[1] / cnt;
    __syncthreads();
    double v[2] = {0, 0};
    for (int i = threadIdx.x; i < P.n; i += BLOCK) {
        if (P.py[i] > 0) {
            const uint32_t rc = P.prc[i];
            const double e1 = (P.xr[rc & 0xffffu] - c1) - m1, e2 = (P.xc[rc >> 16] - c2) - m2;
            v[0] += e1 * e1;
            v[1] += e2 * e2;
        }
    }
    block_sum<BLOCK, 2>(v, sm.red_part, sm.red_out);
    const double h1 = fmax(sqrt(sm.red_out[0] / cnt), 1e-8), h2 = fmax(sqrt(sm.red_out[1] / cnt), 1e-8);
    __syncthreads();
    const double A1 = 1.0 / (h1 * h1), A2 = 1.0 / (h2 * h2);
    const double bb1 = A1 * c1, bb2 = A2 * c2;
    const double cc0 = (c1 * bb1 + c2 * bb2) - 1.0;
    if (threadIdx.x == 0) { out6[0] = -A1; out6[1] = -A2; out6[2] = -0.0; out6[3] = bb1; out6[4] = bb2; out6[5] = -cc0; }
    __syncthreads();
    return true;
}

// ---------------------------------------------------------------------------------------------------------------
// the candidate's state machine

// internal (never stored across rounds) continuation points
enum { PC_IPMDONE = 32, PC_BEGIN_DSM, PC_EST_INIT, PC_FINISH, PC_KKT, PC_PASS, PC_TRIALREQ, PC_ACCEPT, PC_IPMSTART };

__device__ __forceinline__ void ipm_begin(CState& S)
{
    S.t = 0.0; S.s = 1.0; S.z = 1.0; S.d = 1.0; S.lm = 1.0;
    S.relaxed = 0; S.iters = 0;
    S.spec_n = 1; S.spec_k = 0; S.ls_cnt = 0; S.ls_prev = 1; S.spec_h = 0; S.spec_pad = 0; S.z_spec = 0.0;
    S.th1 = S.th2 = S.th3 = 0.0; S.pres0 = S.dres0 = 1.0;
    S.phi0s = S.dphi0s = S.gap0s = S.step0s = S.dsdz0s = S.sigma0s = S.eta0s = 0.0;
    S.sv_t = S.sv_s = S.sv_z = S.sv_d = S.sv_lm = S.sv_dt = S.sv_dz = S.sv_ds = S.sv_resx = S.sv_rznl = S.sv_rxt = 0.0;
}

// Prepares the evaluation of F at vector `vec`: records the request in S (it is published by publish_request when the control
// task ends).  Returns false, recording nothing, if the point holds a NaN (the reference's Energy raises on NaN).
template <int BLOCK>
__device__ bool prepare_request(const Prob& P, CState& S, int vec, int gvec, int want_h, bool budgeted)
{
    const int tid = threadIdx.x;
    const int N = S.use_xi ? P.N : 6;
    const double* x = P.V + (size_t)vec * P.ld;
    // NaN only: an infinite parameter (a relaxed excursion that overflowed) passes the reference's check (dsm.py:72 asserts on
    // NaN), F evaluates to NaN / inf there, the factorisation of the non-finite KKT matrix fails and cp restores the saved state
    int bad = 0;
    for (int i = tid; i < N; i += BLOCK) bad |= isnan(x[i]);
    if (__syncthreads_or(bad)) return false;
    // A solve whose line search never terminates (NaN merit values) runs into the reference's SIGALRM time-out
    // (dsm.py:483-489, cp_timeout = 300 s) and raises; here the evaluation budget plays that role.
    if (budgeted && S.nf + S.nh >= EVAL_BUDGET) return false;
    S.req_vec = vec; S.req_gvec = gvec; S.req_h = want_h;
    S.spec_n = 1; S.spec_k = 0; S.spec_h = 0;                   // one point, as asked (PC_TRIALREQ may add speculative work)
    if (want_h) S.nh += 1; else S.nf += 1;                      // evaluation counters (reported in out.info)
    return true;
}

// Advances candidate c by one control task: consumes the evaluation it asked for, runs up to the next evaluation request.
// Returns true if an evaluation was requested (ws.req[c] describes it; the caller schedules it), false if the candidate is
// finished.  first = false: there is an evaluation to consume (always true except for the task that starts the candidate).
// BIG (runtime, uniform over the CTA): the system is larger than the shared-memory factor (blocked Cholesky in L2).  One
// instantiation serves both: the state machine is inlined once (the kernel's code size is an instruction-cache matter).
template <int BLOCK>
__device__ __forceinline__ bool ctrl_step(const Par& par, const Img& img, const sdsm_cand& cd, const CInfo& ci, const WS& ws, int c,
                          Sm& sm, const sdsm_out& out, double* trace, int task_small_cells, int spec_max, int spec_eager, int spec_retry, int spec_hess, int spec_skip, const bool BIG)
{
    const int tid = threadIdx.x;
    // The candidate's state lives in shared memory while the task runs, one copy per warp: all threads compute the same
    // scalars (warp-uniform code: the lanes of a warp read and write their copy in lock-step), and the ~60 doubles neither
    // fit the registers next to everything else nor belong in per-thread local memory (256 copies = 130 KB, i.e. L2 latency
    // on every access: ncu showed control tasks of small systems at 39 us).
    static_assert(sizeof(CState) % 8 == 0, "CState is copied in 8-byte words");
    CState& S = sm.cst[tid >> 5];
    {
        const unsigned long long* src = reinterpret_cast<const unsigned long long*>(ws.cstate + c);
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(&S);
        for (int i = tid & 31; i < (int)(sizeof(CState) / 8); i += 32) dst[i] = src[i];
        __syncwarp();
    }
    Prob P;
    setup_problem(P, par, cd, ci, ws);
    const int ld = P.ld;
    double* x = P.V + V_X * ld;  double* g = P.V + V_G * ld;   double* dx = P.V + V_DX * ld; double* tv = P.V + V_T * ld;
    double* sx = P.V + V_SX * ld; double* sdx = P.V + V_SDX * ld;
    double* x0 = P.V + V_X0 * ld; double* ev = P.V + V_E * ld; double* w6 = P.V + V_W * ld;
    int32_t* info = out.d_info + (size_t)c * 8;

    double f0 = 0.0;
    // there is an evaluation to consume (always, except for the task that starts the candidate); PC_TRIAL sets it again to look
    // at the next point of a speculative request
    bool need_gather = S.pc != PC_START && S.pc != PC_TRACESTART;
    bool hres_alt = false;
    bool yield = false;
    bool requested = false;

#define REQUEST(vec_, gvec_, h_, next_, onbad_)                                                       \
    do {                                                                                              \
        if (prepare_request<BLOCK>(P, S, vec_, gvec_, h_, (next_) == PC_TOP || (next_) == PC_TRIAL || (next_) == PC_RESTORED)) { S.pc = next_; yield = true; requested = true; } \
        else { onbad_; }                                                                              \
    } while (0)

    while (!yield) {
        if (need_gather) {
            need_gather = false;
            double psi = 0.0;
            const int pt = S.pc == PC_TRIAL ? S.spec_k : 0;          // the point of the request whose results are wanted
            const int req_h = S.pc == PC_TRIAL ? 0 : S.req_h;        // (a trial evaluated with its Hessian: that is gathered when it is accepted)
            double* gout = P.V + (size_t)(S.req_gvec + pt) * ld;
            P.scrG = ws.scrG + (size_t)pt * ws.sg_stride + (size_t)cd.cell_off * SCRG;
            P.xig = ws.xig + (size_t)pt * ws.xig_stride + (size_t)cd.cell_off * NSLOT;
            if (req_h == 2) {
                // Hessian only: value and gradient are those of the accepted trial (same point, same bits: PC_ACCEPT)
                gather_hessian_block<BLOCK>(P, sm, S.z, S.use_xi ? P.N : 6, S.use_xi ? P.hess_tol : 0.0);
                f0 = S.f_trial;
                psi = S.psi_trial;
            }
            else if (!S.use_xi) f0 = req_h ? gather_poly<BLOCK, true>(P, sm, gout, S.z, &psi) : gather_poly<BLOCK, false>(P, sm, gout, S.z, &psi);
            else {
                const double* xe = P.V + (size_t)(S.req_vec + pt) * ld;
                f0 = req_h ? gather_cells<BLOCK, true>(P, sm, xe, gout, S.z, &psi) : gather_cells<BLOCK, false>(P, sm, xe, gout, S.z, &psi);
            }
            S.psi_last = psi;
        }
        const int Nc = S.use_xi ? P.N : 6;
        const int tri_n = Nc * (Nc + 1) / 2;
        const bool l_sm = !BIG && tri_n <= sm.lcap && Nc <= CTRL_BPAD;
        double* Lf = l_sm ? sm.Lf : P.Lg;
        double* bv = l_sm ? sm.Lf + tri_n : sm.Lf;
        const int bvcap = l_sm ? CTRL_BPAD : sm.lcap + CTRL_BPAD;
        switch (S.pc) {
        case PC_START: {
            if (tid == 0) {
                info[0] = ci.m; info[1] = -1; info[2] = -1; info[3] = -1; info[4] = ci.n_fg; info[5] = 0; info[6] = 0; info[7] = 0;
                out.d_fgbox[c * 4 + 0] = -1; out.d_fgbox[c * 4 + 1] = -1; out.d_fgbox[c * 4 + 2] = -1; out.d_fgbox[c * 4 + 3] = -1;
                out.d_onboundary[c] = 0;
                out.d_energy[c] = 0.0;
            }
            S.status = SDSM_ST_OPTIMAL; S.bits = 0; S.it1 = S.it2 = S.it3 = -1; S.nf = S.nh = 0; S.have1 = 0; S.r_ipm = 0;
            S.n_skipped = S.n_unused = S.n_merged = S.n_hwasted = 0;
            S.val1 = 0; S.psi_last = 0; S.energy = 0;
            if (ci.flags) { S.status = SDSM_ST_BADREGION; if (tid == 0) out.d_status[c] = S.status; S.pc = PC_DONE; yield = true; break; }
            if (ci.n_fg == 1 && par.skip_single_fg) { S.status = SDSM_ST_SKIPPED; if (tid == 0) out.d_status[c] = S.status; S.pc = PC_DONE; yield = true; break; }   // objects.py:184-191
            for (int i = tid; i < P.N; i += BLOCK) x[i] = 0.0;
            __syncthreads();
            if (par.init_elliptical) { S.stage = STG_ELL1; S.use_xi = 0; S.pc = PC_IPMSTART; }
            else S.pc = PC_BEGIN_DSM;
            break;
        }
        case PC_TRACESTART: { S.pc = PC_IPMSTART; break; }      // diagnostics: one solve from the x the host provided
        case PC_IPMSTART: {
            ipm_begin(S);
            REQUEST(V_X, V_G, 1, PC_TOP, { S.r_ipm = 2; S.pc = PC_IPMDONE; });
            break;
        }
        // ---- cvxopt.solvers.cp, scalar-epigraph specialisation (oracle/cpsolver.py::cp) ----
        case PC_TOP: {                       // F(x), Df(x) and H = z * Hess f(x) are in (f0, g, P.H)
            const double f = f0 - S.t;
            S.gap = S.s * S.z;
            S.rxt = 1.0 - S.z;
            double gg;
            { double a = 0; for (int i = tid; i < Nc; i += BLOCK) { const double r = S.z * g[i]; a += r * r; } gg = block_sum1<BLOCK>(a, sm.red_part, sm.red_out); }
            S.resx = sqrt(gg + S.rxt * S.rxt);
            S.rznl = S.s + f; S.resznl = fabs(S.rznl);
            const double pcost = S.t, dcost = pcost + S.z * S.rznl - S.gap;
            double relgap = -1.0;   // None
            if (pcost < 0.0) relgap = S.gap / -pcost; else if (dcost > 0.0) relgap = S.gap / dcost;
            double pres = S.resznl, dres = S.resx;
            if (S.iters == 0) {
                const double resx0 = fmax(1.0, S.resx), resznl0 = fmax(1.0, S.resznl);
                S.pres0 = fmax(1.0, pres); S.dres0 = fmax(1.0, dres);
                S.th1 = 1.0 / S.gap; S.th2 = 1.0 / resx0; S.th3 = 1.0 / resznl0;
            }
            S.phi = S.th1 * S.gap + S.th2 * S.resx + S.th3 * S.resznl;
            pres /= S.pres0; dres /= S.dres0;
            if (trace && tid == 0) {
                double* tr = trace + (size_t)S.iters * 8;
                tr[0] = f0; tr[1] = S.gap; tr[2] = pres; tr[3] = dres; tr[4] = S.t; tr[5] = S.s; tr[6] = S.z; tr[7] = (double)S.relaxed;
            }
            if ((pres <= par.feastol && dres <= par.feastol && (S.gap <= par.abstol || (relgap >= 0.0 && relgap <= par.reltol))) || S.iters == par.maxiters) {
                S.r_ipm = (S.iters == par.maxiters) ? 1 : 0;
                S.pc = PC_IPMDONE;
                break;
            }
            if (S.iters == 0) { S.d = sqrt(S.s / S.z); S.lm = sqrt(S.s * S.z); }
            S.lmsq = S.lm * S.lm;
            const long long tc0 = clock64();
            const bool ok = BIG ? cholesky_big<BLOCK>(P.H, Nc, P.Lg, sm) : cholesky<BLOCK>(P.H, Nc, Lf, sm);
            if (BIG && tid == 0) { atomicAdd(ws.counters + Q_CYC_CHOL_BIG, (unsigned long long)(clock64() - tc0)); atomicAdd(ws.counters + Q_N_CHOL_BIG, 1ull); }
            hres_alt = false;
            if (ok) { S.pc = PC_KKT; break; }
            if (S.iters == 0) { S.r_ipm = 2; S.pc = PC_IPMDONE; break; }                 // ValueError("Rank(A) < p or ...")
            // cvxopt restores the saved state for 0 < relaxed < MAX_RELAXED only and gives up ('unknown') when the factorisation
            // fails right after the last relaxed step.  Whether a numerically singular z * Hess f "fails" there is decided by
            // the last bits of the factorisation (LAPACK / CHOLMOD / numpy / this one all differ), and when it happens to
            // succeed cp restores the saved state one step later anyway; so the saved state is restored in both cases
            // (oracle/cpsolver.py RESTORE_POLICY = 'saved').
            if (0 < S.relaxed && S.relaxed <= MAX_RELAXED) {
                // Restart a standard line search from the last saved state: F is re-evaluated at the restored point.
                S.phi = S.phi0s; S.gap = S.gap0s;
                S.d = S.sv_d; S.lm = S.sv_lm; S.t = S.sv_t; S.s = S.sv_s; S.z = S.sv_z; S.lmsq = S.lm * S.lm;
                __syncthreads();
                for (int i = tid; i < Nc; i += BLOCK) x[i] = sx[i];
                S.rxt = S.sv_rxt; S.resx = S.sv_resx; S.rznl = S.sv_rznl; S.resznl = fabs(S.rznl);
                S.relaxed = -1;
                __syncthreads();
                REQUEST(V_X, V_G, 1, PC_RESTORED, { S.r_ipm = 2; S.pc = PC_IPMDONE; });
                break;
            }
            S.r_ipm = 1; S.pc = PC_IPMDONE;                                               // 'unknown' (singular KKT matrix)
            break;
        }
        case PC_RESTORED: {
            const bool ok = BIG ? cholesky_big<BLOCK>(P.H, Nc, P.Lg, sm) : cholesky<BLOCK>(P.H, Nc, Lf, sm);
            if (!ok) { S.r_ipm = 1; S.pc = PC_IPMDONE; break; }
            // cvxopt would refine the Newton system below against the gradient and Hessian of the abandoned point (cpl's residual
            // closures keep them): garbage, possibly non-finite.  The 'saved' policy refines against the restored point's.
            hres_alt = false;
            S.pc = PC_KKT;
            break;
        }
        case PC_KKT: {
            // KKT solves (cp's epigraph wrapper around kkt_chol2 + one step of iterative refinement).  With eta = 0 the
            // x-part of both passes is the same vector: it is computed once (bit-identical to recomputing it).
            const double* Hres = hres_alt ? P.H2 : P.H;
            const double* gres = hres_alt ? sdx : g;
            const double z = S.z, d = S.d;
            S.bt = -S.rxt;
            const double bt = S.bt;
            __syncthreads();
            for (int i = tid; i < Nc; i += BLOCK) { const double bx = -(z * g[i]); tv[i] = bx + g[i] * bt; }
            const long long ts0 = clock64();
            if (BIG) chol_solve_big<BLOCK>(P.Lg, Nc, tv, sm); else chol_solve<BLOCK>(Lf, Nc, tv, bv, bvcap);
            if (BIG && tid == 0) atomicAdd(ws.counters + Q_CYC_SOLVE_BIG, (unsigned long long)(clock64() - ts0));
            S.uz0 = -d * bt;
            const double wz3 = S.uz0 / d;
            S.vt = bt + wz3;
            const double vt = S.vt;
            {   // vx = bx - H ux - g wz3, then ex = K^-1 (vx + g vt)
                constexpr int NW = BLOCK / 32;
                const int lane = tid & 31, warp = tid >> 5;
                for (int i = warp; i < Nc; i += NW) {
                    double a = 0;
                    const double* hr = Hres + (size_t)i * Nc;
                    for (int j = lane; j < Nc; j += 32) a = fma(hr[j], tv[j], a);
                    a = warp_sum(a);
                    if (lane == 0) { const double vx = -(z * g[i]) - a - gres[i] * wz3; ev[i] = vx + g[i] * vt; }
                }
            }
            S.gux = bdot<BLOCK>(g, tv, Nc, sm);
            S.gux_res = hres_alt ? bdot<BLOCK>(gres, tv, Nc, sm) : S.gux;
            if (BIG) chol_solve_big<BLOCK>(P.Lg, Nc, ev, sm); else chol_solve<BLOCK>(Lf, Nc, ev, bv, bvcap);
            S.gex = bdot<BLOCK>(g, ev, Nc, sm);
            for (int i = tid; i < Nc; i += BLOCK) dx[i] = tv[i] + ev[i];
            __syncthreads();
            S.sigma = 0.0; S.eta = 0.0; S.step = 1.0; S.pass = 0;
            S.pc = PC_PASS;
            break;
        }
        case PC_PASS: {
            const double d = S.d, lm = S.lm;
            const double mu = S.gap;
            const double bs = -S.lmsq + S.sigma * mu;
            const double bz = -(1.0 - S.eta) * S.rznl;
            // f4_no_ir
            const double bs1 = bs / lm, bz1 = bz - d * bs1;
            double uz = S.uz0, ut = S.gux + d * d * S.bt - bz1, us = bs1 - uz;
            // residual + second f4_no_ir
            const double vz = bz - (S.gux_res - ut) - d * us;
            const double vs = bs - lm * (uz + us);
            const double vs1 = vs / lm, vz1 = vz - d * vs1;
            const double ez = -d * S.vt, et = S.gex + d * d * S.vt - vz1, es = vs1 - ez;
            ut += et; uz += ez; us += es;
            S.dt = ut; S.dz = uz; S.ds = us;
            S.dsdz = S.ds * S.dz;
            S.dz2 = S.dz / d; S.ds2 = S.ds * d;
            S.ds /= lm; S.dz /= lm;
            const double tmax = fmax(0.0, fmax(-S.ds, -S.dz));
            S.step = (tmax == 0.0) ? 1.0 : fmin(1.0, CP_STEP / tmax);
            S.phi = S.th1 * S.gap + S.th2 * S.resx + S.th3 * S.resznl;
            S.dphi = (S.pass == 0) ? -S.phi : (-S.th1 * (1 - S.sigma) * S.gap - S.th2 * (1 - S.eta) * S.resx - S.th3 * (1 - S.eta) * S.resznl);
            if (S.ls_cnt > 0) S.ls_prev = S.ls_cnt;          // a new line search starts: the last one examined that many trials
            S.ls_cnt = 0;
            S.pc = PC_TRIALREQ;
            break;
        }
        case PC_TRIALREQ: {
            if (spec_skip && S.pass == 0 && 0 <= S.relaxed && S.relaxed < MAX_RELAXED) {
                // First pass while relaxed steps are allowed: cp accepts or rejects the trial on newgap alone, which is a function
                // of the step (PC_TRIAL: the merit function is only consulted when relaxed is -1 or MAX_RELAXED).  cp evaluates F
                // at every trial point all the same, but nothing of what F returns there is ever used (the second pass overwrites
                // it), so the whole first-pass search runs here, without evaluations.  What an evaluation can do besides returning
                // values is kept: a NaN in the point raises (prepare_request), the budget stands in for the time-out, and the
                // trials are counted as the evaluations the reference makes.
                for (;;) {
                    const double step = S.step;
                    int bad = 0;
                    for (int i = tid; i < Nc; i += BLOCK) bad |= isnan(x[i] + step * dx[i]);
                    if (__syncthreads_or(bad) || S.nf + S.nh >= EVAL_BUDGET) { S.r_ipm = 2; S.pc = PC_IPMDONE; break; }
                    S.nf += 1; S.ls_cnt += 1; S.n_skipped += 1;
                    const double newgap = (1.0 - (1.0 - S.sigma) * step) * S.gap + step * step * S.dsdz;
                    if (newgap <= (1.0 - CP_ALPHA * step) * S.gap) {
                        const double r = newgap / S.gap;
                        S.sigma = fmin(r, r * r * r);
                        S.eta = 0.0;
                        S.pass = 1; S.pc = PC_PASS;
                        break;
                    }
                    S.step *= CP_BETA;
                }
                break;
            }
            // Speculative backtracking.  The line search examines x + step dx, x + beta step dx, ... one after the other, and every
            // trial costs a round trip through the evaluation warps.  One request therefore evaluates the next K points of that
            // sequence (the same expressions in the same arithmetic as K sequential requests would use); PC_TRIAL then examines
            // them in order and stops at the first one the line search accepts, so decisions and results are those of the
            // sequential search, bit for bit.  K is predicted: line searches that follow each other (the two passes of an
            // iteration in particular) tend to need the same number of trials.
            // The first request of the second pass is a confident one (its predictor is the first pass of the same iteration);
            // the others are made only for candidates allowed to waste evaluations (spec_eager).
            // A second-pass trial made while relaxed steps are allowed is accepted whatever F returns (PC_TRIAL): one point.
            const bool sure = S.pass == 1 && 0 <= S.relaxed && S.relaxed < MAX_RELAXED;
            int K = 1;
            if (sure) K = 1;
            else if (S.ls_cnt == 0) { if (spec_eager || S.pass == 1) K = S.ls_prev; }
            else if (spec_eager) K = S.ls_cnt < spec_retry ? spec_retry : S.ls_cnt;
            K = K < 1 ? 1 : (K > spec_max ? spec_max : K);
            __syncthreads();
            {
                double st = S.step;
                for (int k = 0; k < K; ++k) {
                    double* px = P.V + (size_t)(V_SPX + k) * ld;
                    for (int i = tid; i < Nc; i += BLOCK) px[i] = x[i] + st * dx[i];
                    st *= CP_BETA;
                }
            }
            __syncthreads();
            REQUEST(V_SPX, V_SPG, 0, PC_TRIAL, { S.r_ipm = 2; S.pc = PC_IPMDONE; });
            const bool want_spec_h = requested && spec_hess && (sure || (S.pass == 1 && S.ls_cnt == 0 && K == S.ls_prev));
            if (requested && K > 1) {
                // a later point that holds a NaN ends the request before it (the sequential search would have stopped there)
                int Ke = 1;
                for (int k = 1; k < K; ++k) {
                    const double* px = P.V + (size_t)(V_SPX + k) * ld;
                    int bad = 0;
                    for (int i = tid; i < Nc; i += BLOCK) bad |= isnan(px[i]);
                    if (__syncthreads_or(bad)) break;
                    Ke = k + 1;
                }
                S.spec_n = Ke;
            }
            if (want_spec_h && S.spec_n == K) {
                // The Hessian with the trial.  The trial is either sure to be accepted, or the first pass of this iteration examined
                // K trials, so that this pass almost certainly accepts its K-th too (the two passes of an iteration need the same
                // number of trials nearly always); the accepted point is the next iterate, whose Hessian is the next thing asked
                // for: the last point of the request is evaluated with it (one pass over the pixels instead of two, one round
                // trip less).  The rows the T_HG tasks write are
                // scaled with the z of the next iterate, which only depends on the step: PC_ACCEPT's expressions.  H of the
                // current iterate is not needed any more (it is factored, and the Newton system is solved).
                double st = S.step;
                for (int k = 1; k < K; ++k) st *= CP_BETA;
                const double ds = (1.0 + st * S.ds) * S.lm;
                const double dz = (1.0 + st * S.dz) * S.lm;
                const double sq_s = sqrt(ds), sq_z = sqrt(dz);
                const double d_new = S.d * sq_s / sq_z;
                const double lm_new = sq_s * sq_z;
                S.z_spec = lm_new / d_new;
                S.req_h = 1; S.spec_h = 1;
            }
            if (requested) S.n_unused += S.spec_n;       // (every examined point takes itself off again: PC_TRIAL)
            break;
        }
        case PC_TRIAL: {                     // F and Df at x + step dx are in (f0, ng)
            const double step = S.step;
            const double* ng = P.V + (size_t)(V_SPG + S.spec_k) * ld;
            bool resumed = false;
            S.ls_cnt += 1; S.n_unused -= 1;
            S.f_trial = f0; S.psi_trial = S.psi_last;
            const double newt = S.t + step * S.dt, newz = S.z + step * S.dz2, news = S.s + step * S.ds2;
            const double newf = f0 - newt;
            double ngg;
            { double a = 0; for (int i = tid; i < Nc; i += BLOCK) { const double r = newz * ng[i]; a += r * r; } ngg = block_sum1<BLOCK>(a, sm.red_part, sm.red_out); }
            const double nrxt = 1.0 - newz;
            const double newresx = sqrt(ngg + nrxt * nrxt);
            const double newresznl = fabs(news + newf);
            const double newgap = (1.0 - (1.0 - S.sigma) * step) * S.gap + step * step * S.dsdz;
            const double newphi = S.th1 * newgap + S.th2 * newresx + S.th3 * newresznl;
            bool backtrack = true;
            if (S.pass == 0) {
                if (newgap <= (1.0 - CP_ALPHA * step) * S.gap && ((0 <= S.relaxed && S.relaxed < MAX_RELAXED) || newphi <= S.phi + CP_ALPHA * step * S.dphi)) {
                    backtrack = false;
                    const double r = newgap / S.gap;
                    S.sigma = fmin(r, r * r * r);
                    S.eta = 0.0;
                } else S.step *= CP_BETA;
            } else {
                if (S.relaxed == -1) {
                    if (newphi <= S.phi + CP_ALPHA * step * S.dphi) backtrack = false; else S.step *= CP_BETA;
                } else if (S.relaxed == 0) {
                    if (newphi <= S.phi + CP_ALPHA * step * S.dphi) S.relaxed = 0;
                    else {
                        S.phi0s = S.phi; S.dphi0s = S.dphi; S.gap0s = S.gap; S.step0s = step;
                        S.sv_d = S.d; S.sv_lm = S.lm; S.sv_t = S.t; S.sv_s = S.s; S.sv_z = S.z;
                        S.sv_rxt = S.rxt; S.sv_resx = S.resx; S.sv_rznl = S.rznl;
                        __syncthreads();
                        for (int i = tid; i < Nc; i += BLOCK) { sx[i] = x[i]; sdx[i] = dx[i]; }
                        __syncthreads();
                        S.sv_dt = S.dt; S.sv_dz = S.dz; S.sv_ds = S.ds;
                        S.dsdz0s = S.dsdz; S.sigma0s = S.sigma; S.eta0s = S.eta;
                        S.relaxed = 1;
                    }
                    backtrack = false;
                } else if (S.relaxed < MAX_RELAXED) {
                    if (newphi <= S.phi0s + CP_ALPHA * S.step0s * S.dphi0s) S.relaxed = 0; else S.relaxed += 1;
                    backtrack = false;
                } else {   // relaxed == MAX_RELAXED
                    if (newphi <= S.phi0s + CP_ALPHA * S.step0s * S.dphi0s) { backtrack = false; S.relaxed = 0; }
                    else if (newphi >= S.phi0s) {
                        // resume the last saved line search
                        S.phi = S.phi0s; S.dphi = S.dphi0s; S.gap = S.gap0s; S.step = S.step0s;
                        S.d = S.sv_d; S.lm = S.sv_lm; S.lmsq = S.lm * S.lm; S.t = S.sv_t; S.s = S.sv_s; S.z = S.sv_z;
                        S.rxt = S.sv_rxt; S.resx = S.sv_resx; S.rznl = S.sv_rznl; S.resznl = fabs(S.rznl);
                        __syncthreads();
                        for (int i = tid; i < Nc; i += BLOCK) { x[i] = sx[i]; dx[i] = sdx[i]; }
                        __syncthreads();
                        S.dt = S.sv_dt; S.dz = S.sv_dz; S.ds = S.sv_ds;
                        S.dz2 = S.dz * S.lm / S.d; S.ds2 = S.ds * S.lm * S.d;
                        S.dsdz = S.dsdz0s; S.sigma = S.sigma0s; S.eta = S.eta0s;
                        S.relaxed = -1;
                        resumed = true;
                    } else if (newphi <= S.phi + CP_ALPHA * step * S.dphi) { backtrack = false; S.relaxed = -1; }
                    else { S.r_ipm = 2; S.pc = PC_IPMDONE; break; }   // cvxopt would repeat this trial until the reference's SIGALRM time-out fires
                }
            }
            // a Hessian evaluated with the request is the last point's: of no use if the search ends before or goes on after it
            if (S.spec_h && (backtrack ? S.spec_k + 1 >= S.spec_n || resumed : S.spec_k + 1 != S.spec_n)) { S.spec_h = 0; S.n_hwasted += 1; }
            if (backtrack) {
                if (resumed) { S.ls_cnt = 0; S.pc = PC_TRIALREQ; }                     // another point, another direction
                else if (S.spec_k + 1 < S.spec_n) {
                    // the next trial is already evaluated: what a request for it would have checked, then look at it
                    if (S.nf + S.nh >= EVAL_BUDGET) { S.r_ipm = 2; S.pc = PC_IPMDONE; }
                    else { S.spec_k += 1; S.nf += 1; need_gather = true; }
                }
                else S.pc = PC_TRIALREQ;
            }
            else if (S.pass == 0) { S.pass = 1; S.pc = PC_PASS; }
            else S.pc = PC_ACCEPT;
            break;
        }
        case PC_ACCEPT: {
            __syncthreads();
            for (int i = tid; i < Nc; i += BLOCK) x[i] = x[i] + S.step * dx[i];
            S.t = S.t + S.step * S.dt;
            const double ds = (1.0 + S.step * S.ds) * S.lm;
            const double dz = (1.0 + S.step * S.dz) * S.lm;
            const double sq_s = sqrt(ds), sq_z = sqrt(dz);
            S.d = S.d * sq_s / sq_z;
            S.lm = sq_s * sq_z;
            S.s = S.lm * S.d;
            S.z = S.lm / S.d;
            S.iters += 1;
            // x is the accepted trial point (the same expression in the same arithmetic as PC_TRIALREQ), where value and gradient
            // are known (S.f_trial, ng): the reference evaluates F there again -- with identical results, dsm.py:388-418 caches them
            // -- so only the Hessian is asked for
            { const double* ng = P.V + (size_t)(V_SPG + S.spec_k) * ld; for (int i = tid; i < Nc; i += BLOCK) g[i] = ng[i]; }
            __syncthreads();
            if (S.spec_h && S.z_spec == S.z) {
                // the Hessian at x came with the trial: what the request for it would have checked and counted, then on to PC_TOP
                S.spec_h = 0;
                if (S.nf + S.nh >= EVAL_BUDGET) { S.r_ipm = 2; S.pc = PC_IPMDONE; break; }
                S.nh += 1; S.n_merged += 1;
                gather_hessian_block<BLOCK>(P, sm, S.z, Nc, S.use_xi ? P.hess_tol : 0.0);
                f0 = S.f_trial; S.psi_last = S.psi_trial;
                S.pc = PC_TOP;
                break;
            }
            S.spec_h = 0;
            REQUEST(V_X, V_G, 2, PC_TOP, { S.r_ipm = 2; S.pc = PC_IPMDONE; });
            break;
        }
        // ---- the driver around the solves: objects.py:321-358 (elliptical initialisation), 361-412 (cvxprog) ----
        case PC_IPMDONE: {
            const int r = S.r_ipm;
            if (S.stage == STG_TRACE) { S.pc = PC_DONE; yield = true; break; }
            if (S.stage == STG_ELL1) {
                S.it1 = S.iters; S.bits |= r; S.have1 = (r != 2);
                if (r == 0) { S.pc = PC_BEGIN_DSM; break; }
                if (S.have1) {
                    __syncthreads();
                    for (int i = tid; i < 6; i += BLOCK) w6[i] = x[i];        // pass-1 solution
                    __syncthreads();
                    REQUEST(V_X, V_G, 0, PC_VAL1, { S.val1 = NAN; S.pc = PC_EST_INIT; });
                } else { S.val1 = INFINITY; S.pc = PC_EST_INIT; }
                break;
            }
            if (S.stage == STG_ELL2) {
                S.it2 = S.iters; S.bits |= r << 2;
                if (r == 2) {
                    if (!S.have1) { S.status = SDSM_ST_ERROR; S.pc = PC_FINISH; break; }   // CvxprogError (objects.py:351-353)
                    __syncthreads();
                    for (int i = tid; i < 6; i += BLOCK) x[i] = w6[i];
                    __syncthreads();
                }
                S.pc = PC_BEGIN_DSM;
                break;
            }
            // STG_DSM (objects.py:393-410)
            S.it3 = S.iters; S.bits |= r << 4;
            if (r == 0) { S.energy = S.psi_last; S.pc = PC_FINISH; break; }
            if (r == 1) {
                // status 'unknown': accepted unless the energy got worse than at the initialisation (objects.py:399-403);
                // J(x) is the value of the evaluation that ended the solve
                S.energy = S.psi_last;
                REQUEST(V_X0, V_NG, 0, PC_J0, { S.pc = PC_FINISH; });
                break;
            }
            // the solver raised: fall back to the initialisation
            S.status = SDSM_ST_FALLBACK;
            __syncthreads();
            for (int i = tid; i < Nc; i += BLOCK) x[i] = x0[i];
            __syncthreads();
            REQUEST(V_X, V_NG, 0, PC_FINAL, { S.energy = NAN; S.pc = PC_FINISH; });
            break;
        }
        case PC_VAL1: { S.val1 = S.psi_last; S.pc = PC_EST_INIT; break; }
        case PC_EST_INIT: {
            const uint16_t* inv_r = ws.inv + cd.tab_off; const uint16_t* inv_c = inv_r + cd.h;
            const bool ok = estimate_initialization<BLOCK>(P, sm, img, cd, inv_r, inv_c, x);
            if (!ok) { S.status = SDSM_ST_ERROR; S.pc = PC_FINISH; break; }                // the reference raises here
            REQUEST(V_X, V_G, 0, PC_VALINIT, { S.stage = STG_ELL2; S.pc = PC_IPMSTART; });
            break;
        }
        case PC_VALINIT: {
            const double init_val = S.psi_last;
            if (init_val > S.val1) {                                                        // "initialization worse than previous solution"
                __syncthreads();
                for (int i = tid; i < 6; i += BLOCK) x[i] = w6[i];
                __syncthreads();
                S.pc = PC_BEGIN_DSM;
            } else { S.stage = STG_ELL2; S.pc = PC_IPMSTART; }
            break;
        }
        case PC_BEGIN_DSM: {
            S.stage = STG_DSM;
            S.use_xi = P.m > 0 ? 1 : 0;
            const int Nd = S.use_xi ? P.N : 6;
            __syncthreads();
            for (int i = tid + 6; i < Nd; i += BLOCK) x[i] = 0.0;
            __syncthreads();
            for (int i = tid; i < Nd; i += BLOCK) x0[i] = x[i];
            if (S.use_xi) for (size_t i = tid; i < (size_t)Nd * Nd; i += BLOCK) P.H[i] = 0.0;
            __syncthreads();
            S.pc = PC_IPMSTART;
            break;
        }
        case PC_J0: {
            const double j0 = S.psi_last;
            if (S.energy > j0) {
                S.status = SDSM_ST_FALLBACK;
                S.energy = j0;
                __syncthreads();
                for (int i = tid; i < Nc; i += BLOCK) x[i] = x0[i];
                __syncthreads();
            }
            S.pc = PC_FINISH;
            break;
        }
        case PC_FINAL: { S.energy = S.psi_last; S.pc = PC_FINISH; break; }
        case PC_FINISH: {
            if (S.status == SDSM_ST_OPTIMAL || S.status == SDSM_ST_FALLBACK) {
                double* pout = out.d_params + cd.par_off;
                const int Nf = 6 + P.m;
                for (int i = tid; i < Nf; i += BLOCK) pout[i] = (i < Nc) ? x[i] : 0.0;
            }
            if (tid == 0) {
                out.d_status[c] = S.status;
                out.d_energy[c] = (S.status == SDSM_ST_OPTIMAL || S.status == SDSM_ST_FALLBACK) ? S.energy : 0.0;
                info[1] = S.it1; info[2] = S.it2; info[3] = S.it3; info[5] = S.bits; info[6] = S.nf; info[7] = S.nh;
            }
            S.pc = PC_DONE;
            yield = true;
            break;
        }
        default: { yield = true; break; }    // PC_DONE
        }
    }
#undef REQUEST
    __syncthreads();
    if (tid < 32) {
        const unsigned long long* src = reinterpret_cast<const unsigned long long*>(&S);
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(ws.cstate + c);
        for (int i = tid; i < (int)(sizeof(CState) / 8); i += 32) dst[i] = src[i];
    }
    if (tid == 0) {
        if (requested) {
            EvalReq rq;
            rq.x = P.V + (size_t)S.req_vec * P.ld; rq.hfac = (S.spec_h ? S.z_spec : S.z) * P.scale_cp; rq.want_h = S.req_h; rq.use_xi = S.use_xi;
            rq.cpt = P.n_ne <= task_small_cells ? EVAL_WARPS : EVAL_CELLS;
            rq.cpw = P.n_ne <= task_small_cells ? 1 : (P.n_ne <= 4 * task_small_cells ? 2 : WARP_ITEM_CELLS);
            rq.npts = S.spec_n; rq.ld = P.ld;
            ws.req[c] = rq;
        }
    }
    __syncthreads();
    return requested;
}

// ---------------------------------------------------------------------------------------------------------------
// kernels (the persistent solver kernel is in sched.cuh)

// Result rasterisation (objects.py:198-211): foreground inside omega, its bounding box, on_boundary.  One CTA per candidate.
constexpr int FIN_BLOCK = 256;
__global__ void __launch_bounds__(FIN_BLOCK)
k_finish(const __grid_constant__ ImgSet imgs, Par par, const sdsm_cand* __restrict__ cands, int n_cand, WS ws, sdsm_out out)
{
    constexpr int BLOCK = FIN_BLOCK;
    __shared__ int s_box[4];
    __shared__ int s_any;
    __shared__ int s_slotj[(BLOCK / 32) * 32];
    const int tid = threadIdx.x;
    if (blockIdx.x == 0 && tid == 0 && out.d_solver) {
        out.d_solver[0] = ws.glob[G_DONE * HOT_STRIDE];
        out.d_solver[1] = ws.glob[G_ABORT * HOT_STRIDE];
    }
    for (int c = blockIdx.x; c < n_cand; c += gridDim.x) {
        const int status = out.d_status[c];
        if (status != SDSM_ST_OPTIMAL && status != SDSM_ST_FALLBACK) continue;
        const sdsm_cand cd = cands[c];
        const Img img = imgs.v[cd.img];
        const CInfo ci = ws.cinfo[c];
        const double* x = out.d_params + cd.par_off;
        const bool use_xi = ci.m > 0;
        const int32_t* cstart = ws.cell_start + cd.cell_off;
        const uint32_t* amask = ws.cell_amask + cd.cell_off;
        const int32_t* cellmap = ws.cellmap + cd.cell_off;
        const uint32_t* prc_c = ws.pix_rc + cd.pix_off;
        const float* gv = ws.gval + (size_t)GROW * cd.pix_off;
        const uint16_t* inv_r = ws.inv + cd.tab_off; const uint16_t* inv_c = inv_r + cd.h;
        // model in the pixel coordinates of the image padded by 1 (dsm.py:113-128 with roi = whole image, pad = 1)
        const double gr = 1.0 / ((double)img.H - 1.0), gc = 1.0 / ((double)img.W - 1.0);
        const double A11 = x[0], A22 = x[1], A12 = x[2], pb1 = x[3], pb2 = x[4], pc = x[5];
        const double v1 = -gr, v2 = -gc;
        const double M11 = gr * A11 * gr, M12 = gr * A12 * gc, M22 = gc * A22 * gc;
        const double mb1 = gr * (A11 * v1 + A12 * v2 + pb1), mb2 = gc * (A12 * v1 + A22 * v2 + pb2);
        const double mc = (v1 * (A11 * v1 + A12 * v2) + v2 * (A12 * v1 + A22 * v2)) + 2 * (pb1 * v1 + pb2 * v2) + pc;
        __syncthreads();
        if (tid == 0) { s_box[0] = 0x7fffffff; s_box[1] = -1; s_box[2] = 0x7fffffff; s_box[3] = -1; s_any = 0; }
        __syncthreads();
        {
            // foreground inside omega: S > 0 (objects.py:200-204); G~ xi through the cell structure
            constexpr int NW = BLOCK / 32;
            const int lane = tid & 31, warp = tid >> 5;
            int* slotj = s_slotj + warp * 32;
            uint8_t* fg = out.d_fgmask + cd.fg_off;
            for (int cell = warp; cell < ci.ncells; cell += NW) {
                const int st = cstart[cell], npix = cstart[cell + 1] - st;
                if (npix == 0) continue;
                int na = 0;
                if (use_xi) {
                    const uint32_t am = amask[cell];
                    na = __popc(am);
                    const int ca = cell / ci.cw, cb = cell - ca * ci.cw;
                    __syncwarp();
                    if (lane < na) { const int sl = nth_set_bit(am, lane); slotj[lane] = cellmap[(ca + sl / 5 - NC) * ci.cw + (cb + sl % 5 - NC)]; }
                    __syncwarp();
                }
                const float* gcell = gv + (size_t)GROW * st;
                for (int p = lane; p < npix; p += 32) {
                    const uint32_t rc = prc_c[st + p];
                    const int br = inv_r[rc & 0xffffu], bc = inv_c[rc >> 16];
                    const double X1 = (double)(cd.r0 + br + 1), X2 = (double)(cd.c0 + bc + 1);
                    double S = (X1 * M11 + X2 * M12) * X1 + (X1 * M12 + X2 * M22) * X2 + 2 * (X1 * mb1 + X2 * mb2) + mc;
                    double sxi = 0;
                    for (int a = 0; a < na; ++a) sxi += (double)gcell[(size_t)p * grow_stride(na) + a] * x[6 + slotj[a]];
                    S += sxi;
                    if (S > 0) {
                        fg[(size_t)br * cd.w + bc] = 1;
                        atomicMin(&s_box[0], br); atomicMax(&s_box[1], br); atomicMin(&s_box[2], bc); atomicMax(&s_box[3], bc);
                    }
                }
            }
            // on_boundary: S > 0 anywhere on the 1-pixel pad ring (objects.py:209); only the polynomial lives there
            const int HP = img.H + 2, WP = img.W + 2;
            int any = 0;
            for (int k = tid; k < 2 * WP + 2 * HP; k += BLOCK) {
                double X1, X2;
                if (k < WP) { X1 = 0; X2 = k; }
                else if (k < 2 * WP) { X1 = HP - 1; X2 = k - WP; }
                else if (k < 2 * WP + HP) { X1 = k - 2 * WP; X2 = 0; }
                else { X1 = k - 2 * WP - HP; X2 = WP - 1; }
                const double S = (X1 * M11 + X2 * M12) * X1 + (X1 * M12 + X2 * M22) * X2 + 2 * (X1 * mb1 + X2 * mb2) + mc;
                any |= (S > 0);
            }
            if (any) s_any = 1;
        }
        __syncthreads();
        if (tid == 0) {
            out.d_onboundary[c] = (uint8_t)s_any;
            if (s_box[1] >= 0) { out.d_fgbox[c * 4 + 0] = s_box[0]; out.d_fgbox[c * 4 + 1] = s_box[1]; out.d_fgbox[c * 4 + 2] = s_box[2]; out.d_fgbox[c * 4 + 3] = s_box[3]; }
        }
    }
}

// register-resident FP64 FMA peak (the FP64 roofline denominator; MEASURED_PEAKS.json has none)
__global__ void __launch_bounds__(256) k_dfma_peak(double* out, int iters)
{
    double a[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = 1.0 + 1e-9 * (threadIdx.x + k);
    const double b = 1.0000001, c = 1e-9;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = fma(a[k], b, c);
    }
    double r = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) r += a[k];
    if (r == 12345.678) out[0] = r;
}

__global__ void k_export(Par par, const sdsm_cand* __restrict__ cands, int c, WS ws, uint32_t* rc, int32_t* cols, float* vals, uint32_t* grid)
{
    const sdsm_cand cd = cands[c];
    const CInfo ci = ws.cinfo[c];
    const int32_t* cstart = ws.cell_start + cd.cell_off;
    const uint32_t* prc = ws.pix_rc + cd.pix_off;
    const float* gv = ws.gval + (size_t)GROW * cd.pix_off;
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < ci.m; j += gridDim.x * blockDim.x) grid[j] = ws.grid_rc[cd.cell_off + j];
    for (int cell = blockIdx.x; cell < ci.ncells; cell += gridDim.x) {
        const int st = cstart[cell], npix = cstart[cell + 1] - st;
        const uint32_t am = ci.m > 0 ? ws.cell_amask[cd.cell_off + cell] : 0u;
        const int na = __popc(am);
        const int ca = cell / ci.cw, cb = cell - ca * ci.cw;
        for (int p = threadIdx.x; p < npix; p += blockDim.x) {
            rc[st + p] = prc[st + p];
            for (int a = 0; a < NSLOT; ++a) {
                int col = -1; float v = 0.f;
                if (a < na) {
                    const int s = nth_set_bit(am, a);
                    col = ws.cellmap[cd.cell_off + (ca + s / 5 - NC) * ci.cw + (cb + s % 5 - NC)];
                    v = gv[(size_t)GROW * st + (size_t)p * grow_stride(na) + a];
                }
                cols[(size_t)(st + p) * NSLOT + a] = col;
                vals[(size_t)(st + p) * NSLOT + a] = v;
            }
        }
    }
}

}  // namespace sdsm
