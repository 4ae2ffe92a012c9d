// K3a: evaluation tasks -- the energy psi, its gradient and (on request) its Hessian (dsm.py:291-385).  A work item is one
// non-empty pixel cell of one candidate; a warp owns an item, a task (sched.cuh) is EVAL_WARPS consecutive items.
//
// All pixels of a subsample x subsample cell see the same <= 25 grid points, so a cell has a dense local parameter vector
// of length L = 6 + #active slots.  Per chunk of 32 pixels each lane builds its pixel's local row into a shared-memory
// tile, then the lanes switch roles and own outputs: lane e accumulates local gradient entry e, and 4x4 register blocks
// of the local Hessian.  Local results go to the cell's scratch (scrG / scrH); k_ctrl gathers them into the candidate's
// value / gradient / Hessian in a fixed order -- no floating-point atomics, bit-reproducible whatever the schedule.
#pragma once
#include "common.cuh"

namespace sdsm {

constexpr double NEG_LOG_DBL_MAX = -709.782712893384;   // -log(finfo(float64).max), dsm.py:298

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

__device__ __forceinline__ int nth_set_bit(uint32_t mask, int n)
{
    for (int k = 0; k < n; ++k) mask &= mask - 1;
    return __ffs(mask) - 1;
}

// ---------------------------------------------------------------------------------------------------------------
// polynomial-only model (6 parameters): each lane keeps its pixels' sums in registers, one warp reduction per cell.
// scrG[0..5] = gradient, scrG[31] = value, scrH[0..20] = packed lower triangle of the 6 x 6 block.

template <bool HESS>
__device__ __forceinline__ void eval_cell_poly(const double* __restrict__ py, const uint32_t* __restrict__ prc, int npix,
                                               const double* __restrict__ xr, const double* __restrict__ xc, const double* x,
                                               double tol, double* sg, double* sh)
{
    const int lane = threadIdx.x & 31;
    const double a1 = x[0], a2 = x[1], a3 = x[2], b1 = x[3], b2 = x[4], c0 = x[5];
    constexpr int K = HESS ? 28 : 7;
    double acc[K];
#pragma unroll
    for (int k = 0; k < K; ++k) acc[k] = 0;
    for (int i = lane; i < npix; i += 32) {
        const uint32_t rc = prc[i];
        const double x1 = xr[rc & 0xffffu], x2 = xc[rc >> 16], y = py[i];
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
            if (kappa < tol) kappa = 0;
            const double dd = kappa * y * y;
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                const double da = dd * q[a];
#pragma unroll
                for (int b = 0; b <= a; ++b) acc[7 + a * (a + 1) / 2 + b] += da * q[b];
            }
        }
    }
#pragma unroll
    for (int k = 0; k < K; ++k) acc[k] = warp_sum(acc[k]);
    if (lane == 0) {
        sg[31] = acc[0];
#pragma unroll
        for (int k = 0; k < 6; ++k) sg[k] = acc[1 + k];
        if (HESS) {
#pragma unroll
            for (int k = 0; k < 21; ++k) sh[k] = acc[7 + k];
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// deformable model: warp per cell, shared-memory tile of local rows

template <bool HESS>
__device__ __forceinline__ void eval_cell_xi(const double* __restrict__ py, const uint32_t* __restrict__ prc, const float* __restrict__ gcell,
                                             int npix, uint32_t am, const double* __restrict__ xr, const double* __restrict__ xc,
                                             const double* x, double tol, double* Tt, const int* slotj,
                                             double* sg, double* sh)
{
    const int lane = threadIdx.x & 31;
    double* U = Tt + 32 * TILE_LD;
    double* DD = U + 32;
    const double a1 = x[0], a2 = x[1], a3 = x[2], b1 = x[3], b2 = x[4], c0 = x[5];
    const double* xi = x + 6;
    const int na = __popc(am), L = 6 + na;
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
    // the xi values this cell sees, one per lane
    const double xi_l = lane < na ? xi[slotj[lane]] : 0.0;

    for (int p0 = 0; p0 < npix; p0 += 32) {
        const int p = p0 + lane;
        const bool act = p < npix;
        double* row = Tt + lane * TILE_LD;
        double y = 0, S = 0, x1 = 0, x2 = 0;
        if (act) {
            const uint32_t rc = prc[p];
            x1 = xr[rc & 0xffffu];
            x2 = xc[rc >> 16];
            y = py[p];
            S = (x1 * a1 + x2 * a3) * x1 + (x1 * a3 + x2 * a2) * x2 + 2 * (x1 * b1 + x2 * b2) + c0;
        }
        row[0] = x1 * x1; row[1] = x2 * x2; row[2] = 2 * (x1 * x2); row[3] = 2 * x1; row[4] = 2 * x2; row[5] = act ? 1.0 : 0.0;
        double sx = 0;
        for (int a = 0; a < na; ++a) {
            const double g = act ? (double)gcell[(size_t)a * npix + p] : 0.0;
            sx += g * __shfl_sync(0xffffffffu, xi_l, a);
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
            if (kappa < tol) kappa = 0;
            DD[lane] = act ? kappa * y * y : 0.0;
        }
        __syncwarp();
        if (lane < L) {
            double s = 0;
            if (lane < 6 || tol == 0) {
#pragma unroll 8
                for (int pp = 0; pp < 32; ++pp) s += Tt[pp * TILE_LD + lane] * U[pp];
            } else {
                // the xi part of the gradient drops |u| < sparsity_tol (dsm.py:346); the theta part does not
                for (int pp = 0; pp < 32; ++pp) { const double uu = U[pp]; s += Tt[pp * TILE_LD + lane] * (fabs(uu) < tol ? 0.0 : uu); }
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
    sg[lane] = (lane == 31) ? acc_v : (lane < L ? acc_g : 0.0);
    if (HESS) {
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

// ---------------------------------------------------------------------------------------------------------------

constexpr size_t EVAL_SMEM = sizeof(double) * EVAL_WARPS * TILE_DOUBLES;

// Evaluation task (c, first): the CTA's warps take the non-empty cells cell_list[first .. first + EVAL_WARPS) of candidate c,
// one warp per cell, and leave the per-cell value / gradient / Hessian in scrG / scrH.  smem_eval: EVAL_SMEM bytes.
__device__ __forceinline__ void eval_task(const Par& par, const sdsm_cand* __restrict__ cands, const WS& ws, int c, int first,
                                          double* smem_eval, int* s_slotj)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double* Tt = smem_eval + warp * TILE_DOUBLES;
    int* slotj = s_slotj + warp * 32;
    const sdsm_cand* cd = cands + c;
    const int64_t pix_off = cd->pix_off, cell_off = cd->cell_off, tab_off = cd->tab_off;
    const CInfo ci = ws.cinfo[c];
    const int idx = first + warp;
    if (idx >= ci.n_ne) return;
    const int cell = ws.cell_list[cell_off + idx];
    const int h = cd->h;
    const EvalReq rq = ws.req[c];
    const int32_t* cstart = ws.cell_start + cell_off;
    const int st = cstart[cell], npix = cstart[cell + 1] - st;
    const double* py = ws.pix_y + pix_off + st;
    const uint32_t* prc = ws.pix_rc + pix_off + st;
    const double* xr = ws.xtab + tab_off;
    const double* xc = xr + h;
    double* sg = ws.scrG + (size_t)(cell_off + cell) * SCRG;
    double* sh = ws.scrH + (size_t)(cell_off + cell) * SCRH;
    if (!rq.use_xi) {
        if (rq.want_h) eval_cell_poly<true>(py, prc, npix, xr, xc, rq.x, par.sparsity_tol, sg, sh);
        else eval_cell_poly<false>(py, prc, npix, xr, xc, rq.x, par.sparsity_tol, sg, sh);
        return;
    }
    const int cw = ci.cw;
    const uint32_t am = ws.cell_amask[cell_off + cell];
    const int na = __popc(am);
    const int ca = cell / cw, cb = cell - ca * cw;
    __syncwarp();
    if (lane < na) {
        const int s = nth_set_bit(am, lane);
        slotj[lane] = ws.cellmap[cell_off + (ca + s / 5 - NC) * cw + (cb + s % 5 - NC)];
    }
    __syncwarp();
    const float* gcell = ws.gval + (size_t)NSLOT * (pix_off + st);
    if (rq.want_h) eval_cell_xi<true>(py, prc, gcell, npix, am, xr, xc, rq.x, par.sparsity_tol, Tt, slotj, sg, sh);
    else eval_cell_xi<false>(py, prc, gcell, npix, am, xr, xc, rq.x, par.sparsity_tol, Tt, slotj, sg, sh);
}

}  // namespace sdsm
