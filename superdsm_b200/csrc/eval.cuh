// K3a: evaluation tasks -- the energy psi, its gradient and (on request) its Hessian (dsm.py:291-385).  A work item is one
// non-empty pixel cell of one candidate; a warp owns an item, a task (sched.cuh) is EVAL_WARPS consecutive items.
//
// All pixels of a subsample x subsample cell see the same <= 25 grid points, so a cell has a dense local parameter vector
// of length L = 6 + #active slots (<= 31).  A warp owns a cell, a lane a pixel: value and local gradient accumulate in
// registers; the local Hessian (a 32-pixel rank update per chunk) runs on the FP64 tensor cores (DMMA.8x8x4) from a
// shared-memory tile.  Local results go to the cell's scratch (scrG / scrH); the control task gathers them into the
// candidate's value / gradient / Hessian in a fixed order -- no floating-point atomics, bit-reproducible whatever the schedule.
#pragma once
#include "common.cuh"
#include "logistic.cuh"

namespace sdsm {

constexpr double NEG_LOG_DBL_MAX = -709.782712893384;   // -log(finfo(float64).max), dsm.py:298

__device__ __forceinline__ int tri(int i, int j) { return i * (i + 1) / 2 + j; }   // i >= j

// One out-of-line copy for all evaluation paths (instruction cache, and bit-identical values whoever calls): logistic.cuh.
// The pair comes back in registers (reference parameters of an out-of-line function go through the stack).
__device__ __noinline__ double2 logistic_pair(double y, double S)
{
    double2 r;
    logistic_eval(y * S, r.x, r.y);
    return r;
}

__device__ __forceinline__ void logistic(double y, double S, double& phi, double& theta)
{
    const double2 r = logistic_pair(y, S);
    phi = r.x;
    theta = r.y;
}

__device__ __forceinline__ int nth_set_bit(uint32_t mask, int n)
{
    for (int k = 0; k < n; ++k) mask &= mask - 1;
    return __ffs(mask) - 1;
}

// Sum over the warp of 32 per-lane values at once: on return lane l holds (in the return value) the warp total of v[l].
// 31 shuffles instead of 160; fixed order, so results do not depend on scheduling.  (Template recursion: every index is a
// compile-time constant, so v stays in registers.)
template <int O>
__device__ __forceinline__ void warp_sum_transposed_step(double (&v)[32], int lane)
{
    const bool up = (lane & O) != 0;
#pragma unroll
    for (int k = 0; k < O; ++k) {
        const double send = up ? v[k] : v[k + O];
        const double recv = __shfl_xor_sync(0xffffffffu, send, O);
        v[k] = (up ? v[k + O] : v[k]) + recv;
    }
    if constexpr (O > 1) warp_sum_transposed_step<O / 2>(v, lane);
}

__device__ __forceinline__ double warp_sum_transposed(double (&v)[32])
{
    warp_sum_transposed_step<16>(v, threadIdx.x & 31);
    return v[0];
}

// D (8 x 8) += A (8 x 4) B (4 x 8) in fp64 on the tensor cores: DMMA.8x8x4.  Fragments: a = A[lane / 4][lane % 4],
// b = B[lane % 4][lane / 4], c0 / c1 = C[lane / 4][2 (lane % 4) + 0 / 1].
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// a load the compiler may not merge with an earlier load of the same address
__device__ __forceinline__ float ldg_again(const float* p)
{
    float v;
    asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}

__device__ __forceinline__ double cell_S(const double (&th)[6], double x1, double x2)
{
    // explicit fma: the value must not depend on how the compiler contracts the expression in a given context
    const double r1 = fma(x1, th[0], x2 * th[2]), r2 = fma(x1, th[2], x2 * th[1]), r3 = fma(x1, th[3], x2 * th[4]);
    return fma(r1, x1, fma(r2, x2, fma(2.0, r3, th[5])));
}

// ---------------------------------------------------------------------------------------------------------------
// polynomial-only model (6 parameters): each lane keeps its pixels' sums in registers, one warp reduction per cell.
// scrG[0..5] = gradient, scrG[31] = value, scrH[0..20] = packed lower triangle of the 6 x 6 block.

template <bool HESS>
__device__ __noinline__ void eval_cell_poly(const double* __restrict__ py, const uint32_t* __restrict__ prc, int npix,
                                               const double* __restrict__ xr, const double* __restrict__ xc, const double* x,
                                               double tol, double* sg, double* sh)
{
    const int lane = threadIdx.x & 31;
    const double th[6] = {x[0], x[1], x[2], x[3], x[4], x[5]};
    double acc[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) acc[k] = 0;
    for (int i = lane; i < npix; i += 32) {
        const uint32_t rc = prc[i];
        const double x1 = xr[rc & 0xffffu], x2 = xc[rc >> 16], y = py[i];
        const double S = cell_S(th, x1, x2);
        double phi, theta;
        logistic(y, S, phi, theta);
        const double u = -y * theta;
        const double q[6] = {x1 * x1, x2 * x2, 2 * (x1 * x2), 2 * x1, 2 * x2, 1.0};
        acc[31] += phi;
#pragma unroll
        for (int k = 0; k < 6; ++k) acc[k] = fma(u, q[k], acc[k]);
        if (HESS) {
            double kappa = theta - theta * theta;
            if (kappa < tol) kappa = 0;
            const double dd = kappa * y * y;
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                const double da = dd * q[a];
#pragma unroll
                for (int b = 0; b <= a; ++b) acc[6 + a * (a + 1) / 2 + b] = fma(da, q[b], acc[6 + a * (a + 1) / 2 + b]);
            }
        }
    }
    const double tot = warp_sum_transposed(acc);     // lane: 0..5 gradient, 6..26 Hessian, 31 value
    if (lane < 6 || lane == 31) sg[lane] = tot;
    if (HESS && lane >= 6 && lane < 27) sh[lane - 6] = tot;
}

// ---------------------------------------------------------------------------------------------------------------
// deformable model: warp per cell, lane per pixel.
//
// cell_value_grad: value and local gradient accumulate in registers, one transposed warp reduction per cell.  The line
// searches of cp compare merit values of evaluations WITH and WITHOUT Hessian at (eventually) the same point and only
// terminate when they are equal, as they are in the reference where both come from the same numpy code; so both kinds of
// evaluation call this one function, and every multiply-add in it is an explicit fma (no context-dependent contraction).
//
// cell_hessian: the local Hessian  sum_p dd_p row_p row_p^T: the lanes stage their rows in a shared-memory tile
// T[parameter][pixel] and the warp multiplies the 8 x 8 output tiles of the lower triangle on the FP64 tensor cores.

// cp.async (LDGSTS): 4 / 8 bytes global -> shared without passing through registers
__device__ __forceinline__ void cp_async4(void* smem, const void* gmem)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async8(void* smem, const void* gmem)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}

// a shared-memory load the compiler may not merge with an earlier load of the same address
__device__ __forceinline__ float4 lds128_again(const float4* p)
{
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"((uint32_t)__cvta_generic_to_shared(p)));
    return v;
}

// Stages chunk k (32 pixels from p0 = 32 k) of the cell into shared memory: every lane copies its own pixel's y, rc and the
// first na4 entries of its row of G~ in 16-byte pieces (a lane only ever reads back what it copied itself, so no warp
// synchronisation is needed; rows are GROW = 28 floats apart in shared memory too, which makes both the 16-byte copies
// and the 16-byte reads conflict-free).  Lanes beyond the cell's last pixel copy the chunk's first pixel: finite data
// that the caller masks out.
__device__ __noinline__ void stage_chunk(char* stage, const double* __restrict__ py, const uint32_t* __restrict__ prc,
                                            const float* __restrict__ gcell, int npix, int na, int p0, int lane)
{
    const int p = (p0 + lane < npix) ? p0 + lane : p0;
    float* G = reinterpret_cast<float*>(stage);
    double* Y = reinterpret_cast<double*>(stage + GROW * 32 * 4);
    uint32_t* RC = reinterpret_cast<uint32_t*>(stage + GROW * 32 * 4 + 32 * 8);
    cp_async8(Y + lane, py + p);
    cp_async4(RC + lane, prc + p);
    const float* gp = gcell + (size_t)p * ((na + 3) & ~3);      // rows of the cell's block: na4 floats each
    float* gs = G + lane * GROW;
#pragma unroll
    for (int a = 0; a < GROW; a += 4)
        if (a < na) cp_async16(gs + a, gp + a);          // warp-uniform
}

// Returns, in lane e, the cell's local gradient entry e (e < 6 + na), and in lane 31 its value.
__device__ __noinline__ double cell_value_grad(const double* __restrict__ py, const uint32_t* __restrict__ prc, const float* __restrict__ gcell,
                                               int npix, int na, const double* __restrict__ xr, const double* __restrict__ xc,
                                               const double* x, double tol, char* stages, const double* XI)
{
    const int lane = threadIdx.x & 31;
    const double th[6] = {x[0], x[1], x[2], x[3], x[4], x[5]};
    double acc[32];                                   // acc[e] = local gradient entry e, acc[31] = value
#pragma unroll
    for (int k = 0; k < 32; ++k) acc[k] = 0;
    const int nchunks = (npix + 31) >> 5;
    // software pipeline: chunk k + STAGES - 1 is in flight while chunk k is evaluated
#pragma unroll
    for (int k = 0; k < STAGES - 1; ++k) {
        if (k < nchunks) stage_chunk(stages + k * STAGE_BYTES, py, prc, gcell, npix, na, 32 * k, lane);
        cp_async_commit();
    }
    int slot = 0, slot_in = STAGES - 1;
    for (int k = 0; k < nchunks; ++k) {
        if (k + STAGES - 1 < nchunks) stage_chunk(stages + slot_in * STAGE_BYTES, py, prc, gcell, npix, na, 32 * (k + STAGES - 1), lane);
        cp_async_commit();
        cp_async_wait<STAGES - 1>();                  // this lane's copies of chunk k have landed
        const char* stage = stages + slot * STAGE_BYTES;
        const float4* G4 = reinterpret_cast<const float4*>(stage) + lane * (GROW / 4);
        const bool act = 32 * k + lane < npix;
        const uint32_t rc = reinterpret_cast<const uint32_t*>(stage + GROW * 32 * 4 + 32 * 8)[lane];
        const double y = act ? reinterpret_cast<const double*>(stage + GROW * 32 * 4)[lane] : 0.0;
        const double x1 = xr[rc & 0xffffu], x2 = xc[rc >> 16];
        double S = cell_S(th, x1, x2);
        double sx0 = 0, sx1 = 0, sx2 = 0, sx3 = 0;        // XI is zero beyond the cell's last active slot, and so is the row
#pragma unroll
        for (int j = 0; j < GROW / 4; ++j) {
            if (4 * j < na) {                             // (warp-uniform)
                const float4 g = G4[j];
                const double2 xa = reinterpret_cast<const double2*>(XI)[2 * j], xb = reinterpret_cast<const double2*>(XI)[2 * j + 1];
                sx0 = fma((double)g.x, xa.x, sx0); sx1 = fma((double)g.y, xa.y, sx1);
                sx2 = fma((double)g.z, xb.x, sx2); sx3 = fma((double)g.w, xb.y, sx3);
            }
        }
        S += (sx0 + sx1) + (sx2 + sx3);
        double phi = 0, theta = 0;
        if (act) logistic(y, S, phi, theta);
        const double u = -y * theta;                  // 0 for the lanes beyond the cell's last pixel (y = 0)
        const double q[6] = {x1 * x1, x2 * x2, 2 * (x1 * x2), 2 * x1, 2 * x2, 1.0};
        acc[31] += phi;
#pragma unroll
        for (int e = 0; e < 6; ++e) acc[e] = fma(u, q[e], acc[e]);
        // the xi part of the gradient drops |u| < sparsity_tol (dsm.py:346); the theta part does not
        const double ux = (tol > 0 && fabs(u) < tol) ? 0.0 : u;
        // the row of G~ is read from shared memory a second time instead of living in 25 registers across the logistic
        // function: with the 32 accumulators that would not fit the 128 registers two resident CTAs per SM allow
#pragma unroll
        for (int j = 0; j < GROW / 4; ++j) {
            if (4 * j < na) {                             // (warp-uniform) nothing is issued for the slots beyond the cell's last one
                const float4 g = lds128_again(G4 + j);
                acc[6 + 4 * j] = fma(ux, (double)g.x, acc[6 + 4 * j]);
                if (4 * j + 1 < NSLOT) acc[7 + 4 * j] = fma(ux, (double)g.y, acc[7 + 4 * j]);
                if (4 * j + 2 < NSLOT) acc[8 + 4 * j] = fma(ux, (double)g.z, acc[8 + 4 * j]);
                if (4 * j + 3 < NSLOT) acc[9 + 4 * j] = fma(ux, (double)g.w, acc[9 + 4 * j]);
            }
        }
        slot = (slot + 1 == STAGES) ? 0 : slot + 1;
        slot_in = (slot_in + 1 == STAGES) ? 0 : slot_in + 1;
    }
    cp_async_wait<0>();
    return warp_sum_transposed(acc);                 // lane: local gradient entry, 31: value
}

__device__ __noinline__ void cell_hessian(const double* __restrict__ py, const uint32_t* __restrict__ prc, const float* __restrict__ gcell,
                                             int npix, int na, const double* __restrict__ xr, const double* __restrict__ xc,
                                             const double* x, double tol, double* Tt, const double* XI, double* sh)
{
    const int lane = threadIdx.x & 31;
    double* DD = Tt + 32 * TILE_LD;
    const double th[6] = {x[0], x[1], x[2], x[3], x[4], x[5]};
    const int L = 6 + na;
    const int nt = (L + 7) >> 3;                      // 8 x 8 tiles per side of the local Hessian
    double hc[10][2];
#pragma unroll
    for (int k = 0; k < 10; ++k) { hc[k][0] = 0; hc[k][1] = 0; }
    const int fr = lane >> 2, fk = lane & 3;          // DMMA fragment coordinates

    for (int p0 = 0; p0 < npix; p0 += 32) {
        const int p = p0 + lane;
        const bool act = p < npix;
        double y = 0, x1 = 0, x2 = 0;
        if (act) {
            const uint32_t rc = prc[p];
            x1 = xr[rc & 0xffffu];
            x2 = xc[rc >> 16];
            y = py[p];
        }
        double S = cell_S(th, x1, x2);
        double* col = Tt + lane;
        double sx0 = 0, sx1 = 0;
        {
            const float4* gp = reinterpret_cast<const float4*>(gcell + (size_t)(act ? p : 0) * ((na + 3) & ~3));
#pragma unroll
            for (int j = 0; j < GROW / 4; ++j) {
                if (4 * j < na) {                         // (warp-uniform)
                    float4 g = __ldg(gp + j);
                    if (!act) g = make_float4(0.f, 0.f, 0.f, 0.f);
                    const double g0 = g.x, g1 = g.y, g2 = g.z, g3 = g.w;
                    col[(6 + 4 * j) * TILE_LD] = g0;
                    if (4 * j + 1 < na) col[(7 + 4 * j) * TILE_LD] = g1;
                    if (4 * j + 2 < na) col[(8 + 4 * j) * TILE_LD] = g2;
                    if (4 * j + 3 < na) col[(9 + 4 * j) * TILE_LD] = g3;
                    sx0 = fma(g0, XI[4 * j], sx0); sx1 = fma(g1, XI[4 * j + 1], sx1);
                    sx0 = fma(g2, XI[4 * j + 2], sx0); sx1 = fma(g3, XI[4 * j + 3], sx1);
                }
            }
        }
        S += sx0 + sx1;
        double phi = 0, theta = 0;
        if (act) logistic(y, S, phi, theta);
        double kappa = theta - theta * theta;
        if (kappa < tol) kappa = 0;
        col[0] = x1 * x1; col[TILE_LD] = x2 * x2; col[2 * TILE_LD] = 2 * (x1 * x2); col[3 * TILE_LD] = 2 * x1; col[4 * TILE_LD] = 2 * x2;
        col[5 * TILE_LD] = act ? 1.0 : 0.0;
        for (int e = L; e < 8 * nt; ++e) col[e * TILE_LD] = 0.0;
        DD[lane] = act ? kappa * y * y : 0.0;
        __syncwarp();
        const double* frag = Tt + fr * TILE_LD + fk;
#pragma unroll 2
        for (int kk = 0; kk < 8; ++kk) {
            const double ddk = DD[4 * kk + fk];
            double f[4], fs[4];
#pragma unroll
            for (int X = 0; X < 4; ++X) {
                f[X] = (X < nt) ? frag[X * 8 * TILE_LD + 4 * kk] : 0.0;
                fs[X] = f[X] * ddk;
            }
#pragma unroll
            for (int I = 0; I < 4; ++I)
#pragma unroll
                for (int J = 0; J <= I; ++J)
                    if (I < nt) dmma884(hc[I * (I + 1) / 2 + J][0], hc[I * (I + 1) / 2 + J][1], fs[I], f[J]);
        }
        __syncwarp();
    }
#pragma unroll
    for (int I = 0; I < 4; ++I)
#pragma unroll
        for (int J = 0; J <= I; ++J) {
            if (I < nt) {
                const int r = 8 * I + fr, q0 = 8 * J + 2 * fk;
#pragma unroll
                for (int i = 0; i < 2; ++i)
                    if (q0 + i <= r && r < L) sh[tri(r, q0 + i)] = hc[I * (I + 1) / 2 + J][i];
            }
        }
}

template <bool HESS>
__device__ __forceinline__ void eval_cell_xi(const double* __restrict__ py, const uint32_t* __restrict__ prc, const float* __restrict__ gcell,
                                             int npix, uint32_t am, const double* __restrict__ xr, const double* __restrict__ xc,
                                             const double* x, double tol, double* Tt, const int* slotj,
                                             double* sg, double* xig, double* sh)
{
    const int lane = threadIdx.x & 31;
    double* XI = Tt + XI_OFF;
    const int na = __popc(am);
    __syncwarp();
    XI[lane] = lane < na ? x[6 + slotj[lane]] : 0.0;  // the xi values this cell sees (zero-padded: rows are read in groups of four)
    __syncwarp();
    const double tot = cell_value_grad(py, prc, gcell, npix, na, xr, xc, x, tol, reinterpret_cast<char*>(Tt), XI);
    if (lane < 6 || lane == 31) sg[lane] = tot;
    else if (lane < 6 + na) {
        // xi gradient: straight into the slot that grid point slotj[a] keeps for this cell (seen from the grid point's cell,
        // this cell sits at the mirrored neighbour position), so that the control task sums 25 contiguous numbers per
        // grid point instead of chasing cell -> mask -> scratch entry
        const int a = lane - 6;
        xig[(size_t)slotj[a] * NSLOT + (NSLOT - 1 - slotj[32 + a])] = tot;
    }
    __syncwarp();
    if (HESS) cell_hessian(py, prc, gcell, npix, na, xr, xc, x, tol, Tt, XI, sh);
    __syncwarp();
}

// ---------------------------------------------------------------------------------------------------------------

constexpr size_t EVAL_SMEM = sizeof(double) * EVAL_WARPS * TILE_DOUBLES;

// Evaluation task (c, first): the CTA's warps share the non-empty cells cell_list[first .. first + EVAL_CELLS) of candidate c
// -- a warp takes the next cell when it is done with its last one (cells differ in their pixel counts) -- and leave the
// per-cell value / gradient / Hessian in scrG / scrH.  smem_eval: EVAL_SMEM bytes; *s_next == 0 on entry (all threads
// synchronised after it was set), and the caller synchronises the CTA after the call.
__device__ __forceinline__ void eval_cell(const Par& par, const sdsm_cand* __restrict__ cands, const WS& ws, int c, const CInfo& ci, int idx,
                                          double* Tt, int* slotj);

__device__ __forceinline__ void eval_task(const Par& par, const sdsm_cand* __restrict__ cands, const WS& ws, int c, int first,
                                          double* smem_eval, int* s_slotj, int* s_next)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const CInfo ci = ws.cinfo[c];
    const int count = min(ws.req[c].cpt, ci.n_ne - first);
    for (;;) {
        int i = 0;
        if (lane == 0) i = atomicAdd(s_next, 1);
        i = __shfl_sync(0xffffffffu, i, 0);
        if (i >= count) break;
        eval_cell(par, cands, ws, c, ci, first + i, smem_eval + warp * TILE_DOUBLES, s_slotj + warp * SLOTJ_INTS);
    }
}

__device__ __forceinline__ void eval_cell(const Par& par, const sdsm_cand* __restrict__ cands, const WS& ws, int c, const CInfo& ci, int idx,
                                          double* Tt, int* slotj)
{
    const int lane = threadIdx.x & 31;
    const sdsm_cand* cd = cands + c;
    const int64_t pix_off = cd->pix_off, cell_off = cd->cell_off, tab_off = cd->tab_off;
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
        else {
            // value + gradient of the polynomial model = the deformable pass without slots: same arithmetic in the same
            // order as eval_cell_poly (the values of the two must agree bit for bit), but y and row|col arrive through
            // the cp.async pipeline instead of dependent loads
            const double tot = cell_value_grad(py, prc, nullptr, npix, 0, xr, xc, rq.x, par.sparsity_tol, reinterpret_cast<char*>(Tt), Tt + XI_OFF);
            if (lane < 6 || lane == 31) sg[lane] = tot;
        }
        return;
    }
    const int cw = ci.cw;
    const uint32_t am = ws.cell_amask[cell_off + cell];
    const int na = __popc(am);
    const int ca = cell / cw, cb = cell - ca * cw;
    __syncwarp();
    if (lane < NSLOT && ((am >> lane) & 1u)) slotj[32 + __popc(am & ((1u << lane) - 1u))] = lane;     // slot of the cell's e-th active entry
    __syncwarp();
    if (lane < na) {
        const int s = slotj[32 + lane];
        slotj[lane] = ws.cellmap[cell_off + (ca + s / 5 - NC) * cw + (cb + s % 5 - NC)];
    }
    __syncwarp();
    const float* gcell = ws.gval + (size_t)GROW * (pix_off + st);
    double* xig = ws.xig + (size_t)cell_off * NSLOT;
    if (rq.want_h) eval_cell_xi<true>(py, prc, gcell, npix, am, xr, xc, rq.x, par.sparsity_tol, Tt, slotj, sg, xig, sh);
    else eval_cell_xi<false>(py, prc, gcell, npix, am, xr, xc, rq.x, par.sparsity_tol, Tt, slotj, sg, xig, sh);
}

}  // namespace sdsm
