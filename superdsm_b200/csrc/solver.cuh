// K3: persistent per-candidate solver.  One CTA fits one candidate from start to finish:
//   elliptical initialisation (objects.py:321-358) -> DSM solve (objects.py:394-410, dsm.py:431-490) -> result
//   rasterisation (objects.py:198-211).
// The interior-point iterations restate cvxopt.solvers.cp (see oracle/cpsolver.py, which this mirrors line by line).
//
// Energy psi, gradient and Hessian (dsm.py:291-385) are evaluated cell by cell: all pixels of a subsample x subsample
// cell see the same <= 25 grid points, so a cell has a dense local parameter vector of length L = 6 + #active slots.
// A warp owns a cell; per chunk of 32 pixels each lane builds its pixel's local vector into a shared-memory tile, then
// the lanes switch roles and own outputs: lane e accumulates local gradient entry e, and 4x4 register blocks of the
// local Hessian.  Local results go to a per-cell scratch and are gathered into the candidate's gradient / Hessian in a
// fixed order -- no atomics, bit-reproducible.
#pragma once
#include "common.cuh"

namespace sdsm {

constexpr double NEG_LOG_DBL_MAX = -709.782712893384;   // -log(finfo(float64).max), dsm.py:298
constexpr int MAX_RELAXED = 8;
constexpr double CP_STEP = 0.99, CP_BETA = 0.5, CP_ALPHA = 0.01;

enum { V_X = 0, V_G, V_DX, V_T, V_NX, V_NG, V_SX, V_SDX, V_RD, V_X0, V_E, V_W };

struct Prob {
    int n, m, N, ncells, cw, ch;
    const double* py; const uint32_t* prc; const float* gv;
    const int32_t* cstart; const uint32_t* amask; const int32_t* cellmap; const uint32_t* grid;
    const double* xr; const double* xc;
    double scale_cp, alpha, eps, tol;
    double *H, *H2, *Lg, *scrH, *scrG;
    int sub;
};

struct Sm {
    double* tile;      // [SOLVER_WARPS * TILE_DOUBLES], aliased by the packed Cholesky factor
    double* vec;       // [NVEC * NV_SMEM]
    double* red_part;  // [SOLVER_WARPS * 32]
    double* red_out;   // [32]
    int*    slotj;     // [SOLVER_WARPS * 32]
    int*    flag;      // [4]
    double* scal;      // [8]  scal[0] = unscaled psi of the last evaluation
};

__device__ __forceinline__ int tri(int i, int j) { return i * (i + 1) / 2 + j; }   // i >= j

__device__ __forceinline__ void logistic(double y, double S, double& phi, double& theta)
{
    const double t = y * S;
    if (t >= NEG_LOG_DBL_MAX) {
        const double h = exp(-t);
        phi = log(1.0 + h);          // dsm.py:321 (log, not log1p)
        theta = h / (1.0 + h);       // dsm.py:310
    } else {                          // exp would overflow: dsm.py:299,309,322
        phi = -t;
        theta = 1.0;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// polynomial-only model (m = 0): thread-strided pixels, register accumulators, one block reduction

template <int BLOCK, bool HESS>
__device__ void eval_poly(const Prob& P, Sm& sm, const double* x, double* out /* shared [28] */)
{
    const double a1 = x[0], a2 = x[1], a3 = x[2], b1 = x[3], b2 = x[4], c0 = x[5];
    constexpr int K = HESS ? 28 : 7;
    double acc[K];
#pragma unroll
    for (int k = 0; k < K; ++k) acc[k] = 0;
    for (int i = threadIdx.x; i < P.n; i += BLOCK) {
        const uint32_t rc = P.prc[i];
        const double x1 = P.xr[rc & 0xffffu], x2 = P.xc[rc >> 16], y = P.py[i];
        const double S = (x1 * a1 + x2 * a3) * x1 + (x1 * a3 + x2 * a2) * x2 + 2 * (x1 * b1 + x2 * b2) + c0;
        double phi, theta;
        logistic(y, S, phi, theta);
        const double u = -y * theta;
        const double q[6] = {x1 * x1, x2 * x2, 2 * (x1 * x2), 2 * x1, 2 * x2, 1.0};
        acc[0] += phi;
#pragma unroll
        for (int k = 0; k < 6; ++k) acc[1 + k] += u * q[k];
        if (HESS) {
            double kappa = theta - theta * theta;
            if (kappa < P.tol) kappa = 0;
            const double dd = kappa * y * y;
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                const double da = dd * q[a];
#pragma unroll
                for (int b = 0; b <= a; ++b) acc[7 + a * (a + 1) / 2 + b] += da * q[b];
            }
        }
    }
    block_sum<BLOCK, K>(acc, sm.red_part, out);
}

// ---------------------------------------------------------------------------------------------------------------
// deformable model (m > 0): warp per cell

__device__ __forceinline__ int nth_set_bit(uint32_t mask, int n)
{
    for (int k = 0; k < n; ++k) mask &= mask - 1;
    return __ffs(mask) - 1;
}

template <int BLOCK, bool HESS>
__device__ void eval_cells(const Prob& P, Sm& sm, const double* x)
{
    constexpr int NW = BLOCK / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double* Tt = sm.tile + warp * TILE_DOUBLES;
    double* U = Tt + 32 * TILE_LD;
    double* DD = U + 32;
    int* slotj = sm.slotj + warp * 32;
    const double a1 = x[0], a2 = x[1], a3 = x[2], b1 = x[3], b2 = x[4], c0 = x[5];
    const double* xi = x + 6;

    for (int cell = warp; cell < P.ncells; cell += NW) {
        const int st = P.cstart[cell], npix = P.cstart[cell + 1] - st;
        if (npix == 0) continue;
        const uint32_t am = P.amask[cell];
        const int na = __popc(am), L = 6 + na;
        const int ca = cell / P.cw, cb = cell - ca * P.cw;
        __syncwarp();
        if (lane < na) {
            const int s = nth_set_bit(am, lane);
            slotj[lane] = P.cellmap[(ca + s / 5 - NC) * P.cw + (cb + s % 5 - NC)];
        }
        __syncwarp();
        const float* gcell = P.gv + (size_t)NSLOT * st;
        const int Lpad = (L + 3) & ~3;
        // Hessian block ownership: 4x4 blocks (A, B), B <= A, of the Lpad x Lpad local matrix
        const int nb = Lpad >> 2, nblk = nb * (nb + 1) / 2;
        int A0 = 0, B0 = 0, A1 = 0, B1 = 0;
        if (HESS) {
            int t = lane;
            while (t >= A0 + 1) { t -= A0 + 1; ++A0; }
            B0 = t;
            t = lane + 32;
            while (t >= A1 + 1) { t -= A1 + 1; ++A1; }
            B1 = t;
        }
        double h0[16], h1[16];
        if (HESS) {
#pragma unroll
            for (int k = 0; k < 16; ++k) { h0[k] = 0; h1[k] = 0; }
        }
        double acc_g = 0, acc_v = 0;

        for (int p0 = 0; p0 < npix; p0 += 32) {
            const int p = p0 + lane;
            const bool act = p < npix;
            double* row = Tt + lane * TILE_LD;
            double y = 0, S = 0, x1 = 0, x2 = 0;
            if (act) {
                const uint32_t rc = P.prc[st + p];
                x1 = P.xr[rc & 0xffffu];
                x2 = P.xc[rc >> 16];
                y = P.py[st + p];
                S = (x1 * a1 + x2 * a3) * x1 + (x1 * a3 + x2 * a2) * x2 + 2 * (x1 * b1 + x2 * b2) + c0;
            }
            row[0] = x1 * x1; row[1] = x2 * x2; row[2] = 2 * (x1 * x2); row[3] = 2 * x1; row[4] = 2 * x2; row[5] = act ? 1.0 : 0.0;
            double sx = 0;
            for (int a = 0; a < na; ++a) {
                const double g = act ? (double)gcell[(size_t)a * npix + p] : 0.0;
                sx += g * xi[slotj[a]];
                row[6 + a] = g;
            }
            for (int e = L; e < Lpad; ++e) row[e] = 0;
            S += sx;
            double phi = 0, theta = 0;
            if (act) logistic(y, S, phi, theta);
            acc_v += phi;
            const double u = -y * theta;
            U[lane] = u;
            if (HESS) {
                double kappa = theta - theta * theta;
                if (kappa < P.tol) kappa = 0;
                DD[lane] = act ? kappa * y * y : 0.0;
            }
            __syncwarp();
            if (lane < L) {
                double s = 0;
                if (lane < 6 || P.tol == 0) {
#pragma unroll 8
                    for (int pp = 0; pp < 32; ++pp) s += Tt[pp * TILE_LD + lane] * U[pp];
                } else {
                    // the xi part of the gradient drops |u| < sparsity_tol (dsm.py:346); the theta part does not
                    for (int pp = 0; pp < 32; ++pp) { const double uu = U[pp]; s += Tt[pp * TILE_LD + lane] * (fabs(uu) < P.tol ? 0.0 : uu); }
                }
                acc_g += s;
            }
            if (HESS) {
                if (lane < nblk) {
                    const double* ra = Tt + 4 * A0;
                    const double* rb = Tt + 4 * B0;
#pragma unroll 4
                    for (int pp = 0; pp < 32; ++pp) {
                        const double dd = DD[pp];
                        const double2 a01 = *reinterpret_cast<const double2*>(ra + pp * TILE_LD);
                        const double2 a23 = *reinterpret_cast<const double2*>(ra + pp * TILE_LD + 2);
                        const double2 b01 = *reinterpret_cast<const double2*>(rb + pp * TILE_LD);
                        const double2 b23 = *reinterpret_cast<const double2*>(rb + pp * TILE_LD + 2);
                        const double av[4] = {a01.x * dd, a01.y * dd, a23.x * dd, a23.y * dd};
                        const double bv[4] = {b01.x, b01.y, b23.x, b23.y};
#pragma unroll
                        for (int i = 0; i < 4; ++i)
#pragma unroll
                            for (int j = 0; j < 4; ++j) h0[i * 4 + j] += av[i] * bv[j];
                    }
                }
                if (lane + 32 < nblk) {
                    const double* ra = Tt + 4 * A1;
                    const double* rb = Tt + 4 * B1;
#pragma unroll 4
                    for (int pp = 0; pp < 32; ++pp) {
                        const double dd = DD[pp];
                        const double2 a01 = *reinterpret_cast<const double2*>(ra + pp * TILE_LD);
                        const double2 a23 = *reinterpret_cast<const double2*>(ra + pp * TILE_LD + 2);
                        const double2 b01 = *reinterpret_cast<const double2*>(rb + pp * TILE_LD);
                        const double2 b23 = *reinterpret_cast<const double2*>(rb + pp * TILE_LD + 2);
                        const double av[4] = {a01.x * dd, a01.y * dd, a23.x * dd, a23.y * dd};
                        const double bv[4] = {b01.x, b01.y, b23.x, b23.y};
#pragma unroll
                        for (int i = 0; i < 4; ++i)
#pragma unroll
                            for (int j = 0; j < 4; ++j) h1[i * 4 + j] += av[i] * bv[j];
                    }
                }
            }
            __syncwarp();
        }
        // local results -> per-cell scratch
        acc_v = warp_sum(acc_v);
        double* sg = P.scrG + (size_t)cell * SCRG;
        sg[lane] = (lane == 31) ? acc_v : (lane < L ? acc_g : 0.0);
        if (HESS) {
            double* sh = P.scrH + (size_t)cell * SCRH;
            if (lane < nblk) {
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int r = 4 * A0 + i, q = 4 * B0 + j;
                        if (q <= r && r < L) sh[tri(r, q)] = h0[i * 4 + j];
                    }
            }
            if (lane + 32 < nblk) {
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int r = 4 * A1 + i, q = 4 * B1 + j;
                        if (q <= r && r < L) sh[tri(r, q)] = h1[i * 4 + j];
                    }
            }
        }
    }
}

// slot index (0..24) of the cell at offset (da, db) from a cell, and the local index of that slot in the cell's vector
__device__ __forceinline__ int local_index(uint32_t am, int da, int db)
{
    const int s = (da + NC) * 5 + (db + NC);
    if (!((am >> s) & 1u)) return -1;
    return 6 + __popc(am & ((1u << s) - 1u));
}

// Gathers the per-cell results into value / gradient / Hessian of the candidate.  g: scaled gradient out.
// Returns the scaled value; sm.scal[0] = unscaled psi.
template <int BLOCK, bool HESS>
__device__ double gather_cells(const Prob& P, Sm& sm, const double* x, double* g, double hz)
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
    for (int cell = tid; cell < P.ncells; cell += BLOCK) {
        if (P.cstart[cell + 1] == P.cstart[cell]) continue;
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
        const double v = hz * sc * sm.red_out[8 + tid];
        P.H[a * N + t] = v;
        P.H[t * N + a] = v;
    }
    // (2) xi gradient: grid point j collects from the cells around its own
    for (int j = tid; j < m; j += BLOCK) {
        const uint32_t gp = P.grid[j];
        const int ja = (int)(gp & 0xffffu) / P.sub, jb = (int)(gp >> 16) / P.sub;
        double s = 0;
        for (int da = -NC; da <= NC; ++da) {
            const int a = ja - da;
            if (a < 0 || a >= ch) continue;
            for (int db = -NC; db <= NC; ++db) {
                const int b = jb - db;
                if (b < 0 || b >= cw) continue;
                const int cell = a * cw + b;
                if (P.cstart[cell + 1] == P.cstart[cell]) continue;
                const int li = local_index(P.amask[cell], da, db);
                if (li >= 0) s += P.scrG[(size_t)cell * SCRG + li];
            }
        }
        const double v = x[6 + j];
        s += P.alpha * (v / sqrt(v * v + P.eps));                             // dsm.py:350
        g[6 + j] = sc * s;
    }
    if (HESS) {
        // (3) xi-theta block
        for (int it = tid; it < m * 6; it += BLOCK) {
            const int j = it / 6, q = it - j * 6;
            const uint32_t gp = P.grid[j];
            const int ja = (int)(gp & 0xffffu) / P.sub, jb = (int)(gp >> 16) / P.sub;
            double s = 0;
            for (int da = -NC; da <= NC; ++da) {
                const int a = ja - da;
                if (a < 0 || a >= ch) continue;
                for (int db = -NC; db <= NC; ++db) {
                    const int b = jb - db;
                    if (b < 0 || b >= cw) continue;
                    const int cell = a * cw + b;
                    if (P.cstart[cell + 1] == P.cstart[cell]) continue;
                    const int li = local_index(P.amask[cell], da, db);
                    if (li >= 0) s += P.scrH[(size_t)cell * SCRH + tri(li, q)];
                }
            }
            s *= hz * sc;
            P.H[(6 + j) * N + q] = s;
            P.H[q * N + 6 + j] = s;
        }
        // (4) xi-xi block: pairs of grid points whose cells are at most 2 NC apart share pixels
        constexpr int SP = 4 * NC + 1;
        for (int it = tid; it < m * SP * SP; it += BLOCK) {
            const int j = it / (SP * SP), e = it - j * (SP * SP);
            const int ea = e / SP - 2 * NC, eb = e % SP - 2 * NC;
            const uint32_t gp = P.grid[j];
            const int ja = (int)(gp & 0xffffu) / P.sub, jb = (int)(gp >> 16) / P.sub;
            const int ka = ja + ea, kb = jb + eb;
            if (ka < 0 || ka >= ch || kb < 0 || kb >= cw) continue;
            const int k = P.cellmap[ka * cw + kb];
            if (k < 0 || k > j) continue;
            double s = 0;
            for (int da = -NC; da <= NC; ++da) {
                const int a = ja - da;                       // cell c = j's cell - (da, db)
                if (a < 0 || a >= ch || abs(ka - a) > NC) continue;
                for (int db = -NC; db <= NC; ++db) {
                    const int b = jb - db;
                    if (b < 0 || b >= cw || abs(kb - b) > NC) continue;
                    const int cell = a * cw + b;
                    if (P.cstart[cell + 1] == P.cstart[cell]) continue;
                    const uint32_t am = P.amask[cell];
                    const int lj = local_index(am, da, db);
                    const int lk = local_index(am, ka - a, kb - b);
                    if (lj >= 0 && lk >= 0) s += P.scrH[(size_t)cell * SCRH + (lj >= lk ? tri(lj, lk) : tri(lk, lj))];
                }
            }
            if (k == j) {
                const double v = x[6 + j];
                const double t3 = v * v, t2 = sqrt(t3 + P.eps);
                double gd = P.alpha * (1.0 / t2 - t3 / (t2 * t2 * t2));      // dsm.py:372-374
                if (gd < 0) gd = 0;
                s += gd;
            }
            s *= hz * sc;
            P.H[(6 + j) * N + 6 + k] = s;
            P.H[(6 + k) * N + 6 + j] = s;
        }
    }
    __syncthreads();
    if (tid == 0) sm.scal[0] = psi;
    return sc * psi;
}

// value + gradient (+ Hessian = hz * scale * Hess psi into P.H, full symmetric) at x.  Sets sm.flag[1] if x is not finite.
template <int BLOCK>
__device__ __noinline__ double eval(const Prob& P, Sm& sm, const double* x, double* g, bool want_h, double hz, bool use_xi)
{
    const int tid = threadIdx.x;
    const int N = use_xi ? P.N : 6;
    __syncthreads();
    int bad = 0;
    for (int i = tid; i < N; i += BLOCK) bad |= !isfinite(x[i]);
    if (bad) sm.flag[1] = 1;
    __syncthreads();
    if (sm.flag[1]) return 0.0;
    if (tid == 0) sm.flag[want_h ? 3 : 2] += 1;          // evaluation counters (reported in out.info)
    if (!use_xi) {
        double* out = sm.red_out;
        if (want_h) eval_poly<BLOCK, true>(P, sm, x, out); else eval_poly<BLOCK, false>(P, sm, x, out);
        const double psi = out[0];
        const double sc = P.scale_cp;
        if (tid < 6) g[tid] = sc * out[1 + tid];
        if (want_h && tid < 21) {
            int a = 0, t = tid;
            while (t >= a + 1) { t -= a + 1; ++a; }
            const double v = hz * sc * out[7 + tid];
            P.H[a * 6 + t] = v;
            P.H[t * 6 + a] = v;
        }
        __syncthreads();
        if (tid == 0) sm.scal[0] = psi;
        __syncthreads();
        return sc * psi;
    }
    if (want_h) { eval_cells<BLOCK, true>(P, sm, x); const double f = gather_cells<BLOCK, true>(P, sm, x, g, hz); __syncthreads(); return f; }
    eval_cells<BLOCK, false>(P, sm, x);
    const double f = gather_cells<BLOCK, false>(P, sm, x, g, hz);
    __syncthreads();
    return f;
}

// ---------------------------------------------------------------------------------------------------------------
// dense Cholesky of the lower triangle of H (N x N, row-major, ld = N) into packed L; rdiag = 1 / diag(L).
// Fails (returns false) on a non-positive or non-finite pivot, as lapack.potrf / cholmod.numeric do.

template <int BLOCK>
__device__ bool cholesky(const double* H, int N, double* L, double* rdiag, Sm& sm)
{
    constexpr int NW = BLOCK / 32;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    __syncthreads();
    for (int i = warp; i < N; i += NW)
        for (int j = lane; j <= i; j += 32) L[tri(i, j)] = H[i * N + j];
    if (tid == 0) sm.flag[0] = 0;
    __syncthreads();
    for (int k = 0; k < N; ++k) {
        if (tid == 0) {
            const double dk = L[tri(k, k)];
            if (!(dk > 0.0) || !isfinite(dk)) sm.flag[0] = 1;
            else { const double r = sqrt(dk); L[tri(k, k)] = r; rdiag[k] = 1.0 / r; }
        }
        __syncthreads();
        if (sm.flag[0]) return false;
        const double rk = rdiag[k];
        for (int i = k + 1 + tid; i < N; i += BLOCK) L[tri(i, k)] *= rk;
        __syncthreads();
        for (int i = k + 1 + warp; i < N; i += NW) {
            const double lik = L[tri(i, k)];
            double* row = L + tri(i, 0);
            for (int j = k + 1 + lane; j <= i; j += 32) row[j] -= lik * L[tri(j, k)];
        }
        __syncthreads();
    }
    return true;
}

// b := (L L^T)^-1 b, by warp 0.  Callers synchronise the block before and after.
__device__ void chol_solve_warp(const double* L, const double* rdiag, int N, double* b)
{
    const int lane = threadIdx.x & 31;
    for (int k = 0; k < N; ++k) {
        const double yk = b[k] * rdiag[k];
        __syncwarp();
        for (int i = k + 1 + lane; i < N; i += 32) b[i] -= L[tri(i, k)] * yk;
        if (lane == 0) b[k] = yk;
        __syncwarp();
    }
    for (int k = N - 1; k >= 0; --k) {
        const double xk = b[k] * rdiag[k];
        __syncwarp();
        const double* row = L + tri(k, 0);
        for (int i = lane; i < k; i += 32) b[i] -= row[i] * xk;
        if (lane == 0) b[k] = xk;
        __syncwarp();
    }
}

template <int BLOCK>
__device__ double bdot(const double* a, const double* b, int N, Sm& sm)
{
    double s = 0;
    for (int i = threadIdx.x; i < N; i += BLOCK) s += a[i] * b[i];
    return block_sum1<BLOCK>(s, sm.red_part, sm.red_out);
}

// ---------------------------------------------------------------------------------------------------------------
// cvxopt.solvers.cp, scalar-epigraph specialisation (oracle/cpsolver.py::cp).
// x (V_X) holds the start point on entry and the returned point on exit.
// Returns 0 'optimal', 1 'unknown', 2 exception (rank error at the first factorisation, non-finite point, stalled search).

template <int BLOCK>
__device__ __noinline__ int ipm(const Prob& P, Sm& sm, double* V, int ld, bool use_xi, const Par& par, int* iters_out, double* trace = nullptr)
{
    const int tid = threadIdx.x;
    const int N = use_xi ? P.N : 6;
    double* x = V + V_X * ld;  double* g = V + V_G * ld;   double* dx = V + V_DX * ld; double* tv = V + V_T * ld;
    double* nx = V + V_NX * ld; double* ng = V + V_NG * ld; double* sx = V + V_SX * ld; double* sdx = V + V_SDX * ld;
    double* rd = V + V_RD * ld; double* ev = V + V_E * ld;
    double* Lf = (N * (N + 1) / 2 <= SMEM_L_CAP) ? sm.tile : P.Lg;

    double t = 0.0, s = 1.0, z = 1.0, d = 1.0, lm = 1.0;
    int relaxed = 0;
    double th1 = 0, th2 = 0, th3 = 0, pres0 = 1, dres0 = 1;
    double phi0s = 0, dphi0s = 0, gap0s = 0, step0s = 0, dsdz0s = 0, sigma0s = 0, eta0s = 0;
    double sv_t = 0, sv_s = 0, sv_z = 0, sv_d = 0, sv_lm = 0, sv_dt = 0, sv_dz = 0, sv_ds = 0, sv_resx = 0, sv_rznl = 0, sv_rxt = 0;
    if (tid == 0) sm.flag[1] = 0;
    __syncthreads();

    for (int iters = 0;; ++iters) {
        *iters_out = iters;
        double f0 = eval<BLOCK>(P, sm, x, g, true, z, use_xi);
        if (sm.flag[1]) return 2;
        double f = f0 - t;
        double gap = s * z;
        double rxt = 1.0 - z;
        double gg;
        { double a = 0; for (int i = tid; i < N; i += BLOCK) { const double r = z * g[i]; a += r * r; } gg = block_sum1<BLOCK>(a, sm.red_part, sm.red_out); }
        double resx = sqrt(gg + rxt * rxt);
        double rznl = s + f, resznl = fabs(rznl);
        const double pcost = t, dcost = pcost + z * rznl - gap;
        double relgap = -1.0;   // None
        if (pcost < 0.0) relgap = gap / -pcost; else if (dcost > 0.0) relgap = gap / dcost;
        double pres = resznl, dres = resx;
        if (iters == 0) {
            const double resx0 = fmax(1.0, resx), resznl0 = fmax(1.0, resznl);
            pres0 = fmax(1.0, pres); dres0 = fmax(1.0, dres);
            th1 = 1.0 / gap; th2 = 1.0 / resx0; th3 = 1.0 / resznl0;
        }
        double phi = th1 * gap + th2 * resx + th3 * resznl;
        pres /= pres0; dres /= dres0;
        if (trace && tid == 0) {
            double* tr = trace + (size_t)iters * 8;
            tr[0] = f0; tr[1] = gap; tr[2] = pres; tr[3] = dres; tr[4] = t; tr[5] = s; tr[6] = z; tr[7] = (double)relaxed;
        }
        if ((pres <= par.feastol && dres <= par.feastol && (gap <= par.abstol || (relgap >= 0.0 && relgap <= par.reltol))) || iters == par.maxiters)
            return iters == par.maxiters ? 1 : 0;

        if (iters == 0) { d = sqrt(s / z); lm = sqrt(s * z); }
        double lmsq = lm * lm;

        bool ok = cholesky<BLOCK>(P.H, N, Lf, rd, sm);
        // cpl's own Df and H (used by the residual of the iterative refinement) vs. the ones cp's KKT wrapper sees
        const double* Hres = P.H;
        const double* gres = g;
        if (!ok) {
            if (iters == 0) return 2;                                       // ValueError("Rank(A) < p or ...")
            if (0 < relaxed && relaxed < MAX_RELAXED) {
                // Restart a standard line search from the last saved state.  cvxopt re-evaluates F at the restored
                // point inside the KKT solver only: cpl's residual closures keep the Df and H of the abandoned point.
                phi = phi0s; gap = gap0s;
                d = sv_d; lm = sv_lm; t = sv_t; s = sv_s; z = sv_z; lmsq = lm * lm;
                __syncthreads();
                for (int i = tid; i < N; i += BLOCK) { x[i] = sx[i]; sdx[i] = g[i]; }
                for (size_t i = tid; i < (size_t)N * N; i += BLOCK) P.H2[i] = P.H[i];
                rxt = sv_rxt; resx = sv_resx; rznl = sv_rznl; resznl = fabs(rznl);
                relaxed = -1;
                __syncthreads();
                eval<BLOCK>(P, sm, x, g, true, z, use_xi);
                if (sm.flag[1]) return 2;
                ok = cholesky<BLOCK>(P.H, N, Lf, rd, sm);
                Hres = P.H2;
                gres = sdx;
            }
            if (!ok) return 1;                                              // 'unknown' (singular KKT matrix)
        }

        // KKT solves (cp's epigraph wrapper around kkt_chol2 + one step of iterative refinement).  With eta = 0 the
        // x-part of both passes is the same vector: it is computed once (bit-identical to recomputing it).
        const double bt = -rxt;
        for (int i = tid; i < N; i += BLOCK) { const double bx = -(z * g[i]); tv[i] = bx + g[i] * bt; }
        __syncthreads();
        if (tid < 32) chol_solve_warp(Lf, rd, N, tv);
        __syncthreads();
        const double uz0 = -d * bt;
        const double wz3 = uz0 / d;
        const double vt = bt + wz3;
        {   // vx = bx - H ux - g wz3, then ex = K^-1 (vx + g vt)
            constexpr int NW = BLOCK / 32;
            const int lane = tid & 31, warp = tid >> 5;
            for (int i = warp; i < N; i += NW) {
                double a = 0;
                const double* hr = Hres + (size_t)i * N;
                for (int j = lane; j < N; j += 32) a += hr[j] * tv[j];
                a = warp_sum(a);
                if (lane == 0) { const double vx = -(z * g[i]) - a - gres[i] * wz3; ev[i] = vx + g[i] * vt; }
            }
        }
        const double gux = bdot<BLOCK>(g, tv, N, sm);
        const double gux_res = (gres == g) ? gux : bdot<BLOCK>(gres, tv, N, sm);
        if (tid < 32) chol_solve_warp(Lf, rd, N, ev);
        __syncthreads();
        const double gex = bdot<BLOCK>(g, ev, N, sm);
        for (int i = tid; i < N; i += BLOCK) dx[i] = tv[i] + ev[i];
        __syncthreads();

        double sigma = 0.0, eta = 0.0, step = 1.0, dt = 0, dz = 0, ds = 0;
        for (int pass = 0; pass < 2; ++pass) {
            const double mu = gap;
            const double bs = -lmsq + sigma * mu;
            const double bz = -(1.0 - eta) * rznl;
            // f4_no_ir
            const double bs1 = bs / lm, bz1 = bz - d * bs1;
            double uz = uz0, ut = gux + d * d * bt - bz1, us = bs1 - uz;
            // residual + second f4_no_ir
            const double vz = bz - (gux_res - ut) - d * us;
            const double vs = bs - lm * (uz + us);
            const double vs1 = vs / lm, vz1 = vz - d * vs1;
            const double ez = -d * vt, et = gex + d * d * vt - vz1, es = vs1 - ez;
            ut += et; uz += ez; us += es;
            dt = ut; dz = uz; ds = us;

            double dsdz = ds * dz;
            double dz2 = dz / d, ds2 = ds * d;
            ds /= lm; dz /= lm;
            const double tmax = fmax(0.0, fmax(-ds, -dz));
            step = (tmax == 0.0) ? 1.0 : fmin(1.0, CP_STEP / tmax);
            phi = th1 * gap + th2 * resx + th3 * resznl;
            double dphi = (pass == 0) ? -phi : (-th1 * (1 - sigma) * gap - th2 * (1 - eta) * resx - th3 * (1 - eta) * resznl);

            bool backtrack = true;
            while (backtrack) {
                for (int i = tid; i < N; i += BLOCK) nx[i] = x[i] + step * dx[i];
                const double newt = t + step * dt, newz = z + step * dz2, news = s + step * ds2;
                const double nf0 = eval<BLOCK>(P, sm, nx, ng, false, 0.0, use_xi);
                if (sm.flag[1]) return 2;
                const double newf = nf0 - newt;
                double ngg;
                { double a = 0; for (int i = tid; i < N; i += BLOCK) { const double r = newz * ng[i]; a += r * r; } ngg = block_sum1<BLOCK>(a, sm.red_part, sm.red_out); }
                const double nrxt = 1.0 - newz;
                const double newresx = sqrt(ngg + nrxt * nrxt);
                const double newresznl = fabs(news + newf);
                const double newgap = (1.0 - (1.0 - sigma) * step) * gap + step * step * dsdz;
                const double newphi = th1 * newgap + th2 * newresx + th3 * newresznl;
                if (pass == 0) {
                    if (newgap <= (1.0 - CP_ALPHA * step) * gap && ((0 <= relaxed && relaxed < MAX_RELAXED) || newphi <= phi + CP_ALPHA * step * dphi)) {
                        backtrack = false;
                        const double r = newgap / gap;
                        sigma = fmin(r, r * r * r);
                        eta = 0.0;
                    } else step *= CP_BETA;
                } else {
                    if (relaxed == -1) {
                        if (newphi <= phi + CP_ALPHA * step * dphi) backtrack = false; else step *= CP_BETA;
                    } else if (relaxed == 0) {
                        if (newphi <= phi + CP_ALPHA * step * dphi) relaxed = 0;
                        else {
                            phi0s = phi; dphi0s = dphi; gap0s = gap; step0s = step;
                            sv_d = d; sv_lm = lm; sv_t = t; sv_s = s; sv_z = z;
                            sv_rxt = rxt; sv_resx = resx; sv_rznl = rznl;
                            for (int i = tid; i < N; i += BLOCK) { sx[i] = x[i]; sdx[i] = dx[i]; }
                            sv_dt = dt; sv_dz = dz; sv_ds = ds;
                            dsdz0s = dsdz; sigma0s = sigma; eta0s = eta;
                            relaxed = 1;
                        }
                        backtrack = false;
                    } else if (relaxed < MAX_RELAXED) {
                        if (newphi <= phi0s + CP_ALPHA * step0s * dphi0s) relaxed = 0; else relaxed += 1;
                        backtrack = false;
                    } else {   // relaxed == MAX_RELAXED
                        if (newphi <= phi0s + CP_ALPHA * step0s * dphi0s) { backtrack = false; relaxed = 0; }
                        else if (newphi >= phi0s) {
                            // resume the last saved line search
                            phi = phi0s; dphi = dphi0s; gap = gap0s; step = step0s;
                            d = sv_d; lm = sv_lm; lmsq = lm * lm; t = sv_t; s = sv_s; z = sv_z;
                            rxt = sv_rxt; resx = sv_resx; rznl = sv_rznl; resznl = fabs(rznl);
                            __syncthreads();
                            for (int i = tid; i < N; i += BLOCK) { x[i] = sx[i]; dx[i] = sdx[i]; }
                            __syncthreads();
                            dt = sv_dt; dz = sv_dz; ds = sv_ds;
                            dz2 = dz * lm / d; ds2 = ds * lm * d;
                            dsdz = dsdz0s; sigma = sigma0s; eta = eta0s;
                            relaxed = -1;
                        } else if (newphi <= phi + CP_ALPHA * step * dphi) { backtrack = false; relaxed = -1; }
                        else return 2;   // cvxopt would repeat this trial until the reference's SIGALRM time-out fires
                    }
                }
            }
        }
        // accept the step
        __syncthreads();
        for (int i = tid; i < N; i += BLOCK) x[i] = x[i] + step * dx[i];
        t = t + step * dt;
        ds = (1.0 + step * ds) * lm;
        dz = (1.0 + step * dz) * lm;
        const double sq_s = sqrt(ds), sq_z = sqrt(dz);
        d = d * sq_s / sq_z;
        lm = sq_s * sq_z;
        s = lm * d;
        z = lm / d;
        __syncthreads();
    }
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

template <int BLOCK>
__device__ void setup_problem(Prob& P, const Par& par, const sdsm_cand& cd, const CInfo& ci, const WS& ws, char* cta)
{
    P.n = cd.n; P.m = ci.m; P.N = 6 + ci.m; P.ncells = ci.ncells; P.cw = ci.cw; P.ch = ci.ch;
    P.py = ws.pix_y + cd.pix_off; P.prc = ws.pix_rc + cd.pix_off; P.gv = ws.gval + (size_t)NSLOT * cd.pix_off;
    P.cstart = ws.cell_start + cd.cell_off; P.amask = ws.cell_amask + cd.cell_off; P.cellmap = ws.cellmap + cd.cell_off;
    P.grid = ws.grid_rc + cd.cell_off;
    P.xr = ws.xtab + cd.tab_off; P.xc = P.xr + cd.h;
    P.scale_cp = par.scale / (double)cd.n;                                   // objects.py:380
    P.alpha = par.alpha; P.eps = par.epsilon; P.tol = par.sparsity_tol; P.sub = par.sub;
    P.H = (double*)(cta + ws.off_H); P.H2 = (double*)(cta + ws.off_H2); P.Lg = (double*)(cta + ws.off_L);
    P.scrH = (double*)(cta + ws.off_scrH); P.scrG = (double*)(cta + ws.off_scrG);
}

__device__ __forceinline__ void carve_smem(Sm& sm, char* base)
{
    sm.tile = (double*)base;
    sm.vec = sm.tile + SOLVER_WARPS * TILE_DOUBLES;
    sm.red_part = sm.vec + NVEC * NV_SMEM;
    sm.red_out = sm.red_part + SOLVER_WARPS * 32;
    sm.scal = sm.red_out + 32;
    sm.slotj = (int*)(sm.scal + 8);
    sm.flag = sm.slotj + SOLVER_WARPS * 32;
}
constexpr size_t SOLVER_SMEM = sizeof(double) * (SOLVER_WARPS * TILE_DOUBLES + NVEC * NV_SMEM + SOLVER_WARPS * 32 + 32 + 8)
                             + sizeof(int) * (SOLVER_WARPS * 32 + 8);

__global__ void __launch_bounds__(SOLVER_BLOCK, 2)
k_solve(Img img, Par par, const sdsm_cand* __restrict__ cands, int n_cand, const int32_t* __restrict__ order, WS ws, sdsm_out out)
{
    constexpr int BLOCK = SOLVER_BLOCK;
    extern __shared__ __align__(16) char smem_raw[];
    __shared__ int s_next;
    __shared__ int s_box[4];
    __shared__ int s_any;
    Sm sm;
    carve_smem(sm, smem_raw);
    const int tid = threadIdx.x;
    char* cta = ws.cta + (size_t)blockIdx.x * ws.cta_bytes;

    while (true) {
        __syncthreads();
        if (tid == 0) s_next = atomicAdd(&ws.counters[0], 1);
        __syncthreads();
        const int it = s_next;
        if (it >= n_cand) break;
        const int c = order[it];
        const sdsm_cand cd = cands[c];
        const CInfo ci = ws.cinfo[c];
        int32_t* info = out.d_info + (size_t)c * 8;
        if (tid == 0) {
            sm.flag[2] = 0; sm.flag[3] = 0;
            info[0] = ci.m; info[1] = -1; info[2] = -1; info[3] = -1; info[4] = ci.n_fg; info[5] = 0; info[6] = 0; info[7] = 0;
            out.d_fgbox[c * 4 + 0] = -1; out.d_fgbox[c * 4 + 1] = -1; out.d_fgbox[c * 4 + 2] = -1; out.d_fgbox[c * 4 + 3] = -1;
            out.d_onboundary[c] = 0;
            out.d_energy[c] = 0.0;
        }
        if (ci.flags) { if (tid == 0) out.d_status[c] = SDSM_ST_BADREGION; continue; }
        if (ci.n_fg == 1) { if (tid == 0) out.d_status[c] = SDSM_ST_SKIPPED; continue; }     // objects.py:184-191

        Prob P;
        setup_problem<BLOCK>(P, par, cd, ci, ws, cta);
        const int N = P.N;
        const bool use_xi = P.m > 0;
        double* V; int ld;
        if (N <= NV_SMEM) { V = sm.vec; ld = NV_SMEM; } else { V = (double*)(cta + ws.off_vec); ld = N; }
        double* x = V + V_X * ld; double* x0 = V + V_X0 * ld; double* w6 = V + V_W * ld;
        const uint16_t* inv_r = ws.inv + cd.tab_off; const uint16_t* inv_c = inv_r + cd.h;
        int status = SDSM_ST_OPTIMAL;
        int it1 = -1, it2 = -1, it3 = -1, bits = 0;

        // ---- elliptical initialisation (objects.py:321-358) ----
        for (int i = tid; i < N; i += BLOCK) x[i] = 0.0;
        __syncthreads();
        if (par.init_elliptical) {
            const int r1 = ipm<BLOCK>(P, sm, V, ld, false, par, &it1);
            bits |= r1;
            bool have = (r1 != 2);
            double val = INFINITY;
            if (have) {
                eval<BLOCK>(P, sm, x, V + V_G * ld, false, 0.0, false);
                val = sm.flag[1] ? NAN : sm.scal[0];
                for (int i = tid; i < 6; i += BLOCK) w6[i] = x[i];            // pass-1 solution
            }
            __syncthreads();
            if (!have || r1 != 0) {
                const bool ok = estimate_initialization<BLOCK>(P, sm, img, cd, inv_r, inv_c, x);
                if (!ok) status = SDSM_ST_ERROR;                             // the reference raises here
                else {
                    if (tid == 0) sm.flag[1] = 0;
                    __syncthreads();
                    eval<BLOCK>(P, sm, x, V + V_G * ld, false, 0.0, false);
                    const double init_val = sm.flag[1] ? NAN : sm.scal[0];
                    __syncthreads();
                    if (init_val > val) {                                    // "initialization worse than previous solution"
                        for (int i = tid; i < 6; i += BLOCK) x[i] = w6[i];
                    } else {
                        const int r2 = ipm<BLOCK>(P, sm, V, ld, false, par, &it2);
                        bits |= r2 << 2;
                        if (r2 == 2) {
                            if (!have) status = SDSM_ST_ERROR;               // CvxprogError (objects.py:351-353)
                            else for (int i = tid; i < 6; i += BLOCK) x[i] = w6[i];
                        }
                    }
                }
                __syncthreads();
            }
        }
        if (status == SDSM_ST_ERROR) {
            if (tid == 0) { out.d_status[c] = status; info[1] = it1; info[2] = it2; info[5] = bits; }
            continue;
        }

        // ---- DSM solve (objects.py:393-410) ----
        __syncthreads();
        for (int i = tid + 6; i < N; i += BLOCK) x[i] = 0.0;
        __syncthreads();
        for (int i = tid; i < N; i += BLOCK) x0[i] = x[i];
        if (use_xi) for (size_t i = tid; i < (size_t)N * N; i += BLOCK) P.H[i] = 0.0;
        if (tid == 0) sm.flag[1] = 0;
        __syncthreads();
        const int r3 = ipm<BLOCK>(P, sm, V, ld, use_xi, par, &it3);
        bits |= r3 << 4;
        bool fallback = (r3 == 2);
        if (r3 == 1) {
            if (tid == 0) sm.flag[1] = 0;
            __syncthreads();
            eval<BLOCK>(P, sm, x, V + V_G * ld, false, 0.0, use_xi);
            const double jx = sm.scal[0];
            __syncthreads();
            eval<BLOCK>(P, sm, x0, V + V_G * ld, false, 0.0, use_xi);
            const double j0 = sm.scal[0];
            __syncthreads();
            if (jx > j0) fallback = true;
        }
        if (fallback) {
            status = SDSM_ST_FALLBACK;
            __syncthreads();
            for (int i = tid; i < N; i += BLOCK) x[i] = x0[i];
        }
        if (tid == 0) sm.flag[1] = 0;
        __syncthreads();

        // ---- result (objects.py:198-211) ----
        eval<BLOCK>(P, sm, x, V + V_G * ld, false, 0.0, use_xi);
        const double energy = sm.scal[0];
        double* pout = out.d_params + cd.par_off;
        for (int i = tid; i < N; i += BLOCK) pout[i] = x[i];
        // model in the pixel coordinates of the image padded by 1 (dsm.py:113-128 with roi = whole image, pad = 1)
        const double gr = 1.0 / ((double)img.H - 1.0), gc = 1.0 / ((double)img.W - 1.0);
        const double A11 = x[0], A22 = x[1], A12 = x[2], pb1 = x[3], pb2 = x[4], pc = x[5];
        const double v1 = -gr, v2 = -gc;
        const double M11 = gr * A11 * gr, M12 = gr * A12 * gc, M22 = gc * A22 * gc;
        const double mb1 = gr * (A11 * v1 + A12 * v2 + pb1), mb2 = gc * (A12 * v1 + A22 * v2 + pb2);
        const double mc = (v1 * (A11 * v1 + A12 * v2) + v2 * (A12 * v1 + A22 * v2)) + 2 * (pb1 * v1 + pb2 * v2) + pc;
        if (tid == 0) { s_box[0] = 0x7fffffff; s_box[1] = -1; s_box[2] = 0x7fffffff; s_box[3] = -1; s_any = 0; }
        __syncthreads();
        {
            // foreground inside omega: S > 0 (objects.py:200-204); G~ xi through the cell structure
            constexpr int NW = BLOCK / 32;
            const int lane = tid & 31, warp = tid >> 5;
            int* slotj = sm.slotj + warp * 32;
            uint8_t* fg = out.d_fgmask + cd.fg_off;
            for (int cell = warp; cell < P.ncells; cell += NW) {
                const int st = P.cstart[cell], npix = P.cstart[cell + 1] - st;
                if (npix == 0) continue;
                int na = 0;
                if (use_xi) {
                    const uint32_t am = P.amask[cell];
                    na = __popc(am);
                    const int ca = cell / P.cw, cb = cell - ca * P.cw;
                    __syncwarp();
                    if (lane < na) { const int sl = nth_set_bit(am, lane); slotj[lane] = P.cellmap[(ca + sl / 5 - NC) * P.cw + (cb + sl % 5 - NC)]; }
                    __syncwarp();
                }
                const float* gcell = P.gv + (size_t)NSLOT * st;
                for (int p = lane; p < npix; p += 32) {
                    const uint32_t rc = P.prc[st + p];
                    const int br = inv_r[rc & 0xffffu], bc = inv_c[rc >> 16];
                    const double X1 = (double)(cd.r0 + br + 1), X2 = (double)(cd.c0 + bc + 1);
                    double S = (X1 * M11 + X2 * M12) * X1 + (X1 * M12 + X2 * M22) * X2 + 2 * (X1 * mb1 + X2 * mb2) + mc;
                    double sxi = 0;
                    for (int a = 0; a < na; ++a) sxi += (double)gcell[(size_t)a * npix + p] * x[6 + slotj[a]];
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
            out.d_status[c] = status;
            out.d_energy[c] = energy;
            out.d_onboundary[c] = (uint8_t)s_any;
            if (s_box[1] >= 0) { out.d_fgbox[c * 4 + 0] = s_box[0]; out.d_fgbox[c * 4 + 1] = s_box[1]; out.d_fgbox[c * 4 + 2] = s_box[2]; out.d_fgbox[c * 4 + 3] = s_box[3]; }
            info[1] = it1; info[2] = it2; info[3] = it3; info[5] = bits; info[6] = sm.flag[2]; info[7] = sm.flag[3];
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// diagnostics for the parity tests

__global__ void __launch_bounds__(SOLVER_BLOCK, 2)
k_eval(Par par, const sdsm_cand* __restrict__ cands, int c, WS ws, const double* __restrict__ xin, double* value, double* grad, double* Hout)
{
    constexpr int BLOCK = SOLVER_BLOCK;
    extern __shared__ __align__(16) char smem_raw[];
    Sm sm;
    carve_smem(sm, smem_raw);
    const int tid = threadIdx.x;
    const sdsm_cand cd = cands[c];
    const CInfo ci = ws.cinfo[c];
    Prob P;
    setup_problem<BLOCK>(P, par, cd, ci, ws, ws.cta);
    P.scale_cp = 1.0;
    const int N = P.N;
    double* V = (double*)(ws.cta + ws.off_vec);
    double* x = V; double* g = V + N;
    for (int i = tid; i < N; i += BLOCK) x[i] = xin[i];
    for (size_t i = tid; i < (size_t)N * N; i += BLOCK) P.H[i] = 0.0;
    if (tid == 0) sm.flag[1] = 0;
    __syncthreads();
    const double f = eval<BLOCK>(P, sm, x, g, true, 1.0, P.m > 0);
    __syncthreads();
    if (tid == 0) *value = f;
    for (int i = tid; i < N; i += BLOCK) grad[i] = g[i];
    for (size_t i = tid; i < (size_t)N * N; i += BLOCK) Hout[i] = P.H[i];
}

// runs one interior-point solve of candidate c (packed by the last batch) from x0 and records the iterates
__global__ void __launch_bounds__(SOLVER_BLOCK, 2)
k_trace(Par par, const sdsm_cand* __restrict__ cands, int c, int use_xi_, WS ws, const double* __restrict__ x0, double* trace,
        double* xout, int32_t* status_iters)
{
    constexpr int BLOCK = SOLVER_BLOCK;
    extern __shared__ __align__(16) char smem_raw[];
    Sm sm;
    carve_smem(sm, smem_raw);
    const int tid = threadIdx.x;
    const sdsm_cand cd = cands[c];
    const CInfo ci = ws.cinfo[c];
    Prob P;
    setup_problem<BLOCK>(P, par, cd, ci, ws, ws.cta);
    const bool use_xi = use_xi_ && P.m > 0;
    const int N = use_xi ? P.N : 6;
    double* V; int ld;
    if (P.N <= NV_SMEM) { V = sm.vec; ld = NV_SMEM; } else { V = (double*)(ws.cta + ws.off_vec); ld = P.N; }
    for (int i = tid; i < N; i += BLOCK) V[V_X * ld + i] = x0[i];
    if (use_xi) for (size_t i = tid; i < (size_t)N * N; i += BLOCK) P.H[i] = 0.0;
    __syncthreads();
    int iters = -1;
    const int r = ipm<BLOCK>(P, sm, V, ld, use_xi, par, &iters, trace);
    __syncthreads();
    for (int i = tid; i < N; i += BLOCK) xout[i] = V[V_X * ld + i];
    if (tid == 0) { status_iters[0] = r; status_iters[1] = iters; }
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
    const float* gv = ws.gval + (size_t)NSLOT * cd.pix_off;
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
                    v = gv[(size_t)NSLOT * st + (size_t)a * npix + p];
                }
                cols[(size_t)(st + p) * NSLOT + a] = col;
                vals[(size_t)(st + p) * NSLOT + a] = v;
            }
        }
    }
}

}  // namespace sdsm
