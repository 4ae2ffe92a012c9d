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

template <int O>
__device__ __forceinline__ void warp_sum_transposed_step8(double (&v)[8], int lane)
{
    const bool up = (lane & O) != 0;
#pragma unroll
    for (int k = 0; k < O; ++k) {
        const double send = up ? v[k] : v[k + O];
        const double recv = __shfl_xor_sync(0xffffffffu, send, O);
        v[k] = (up ? v[k + O] : v[k]) + recv;
    }
    if constexpr (O > 1) warp_sum_transposed_step8<O / 2>(v, lane);
}

// Sum over the warp of 8 per-lane values at once (the 6 theta-gradient entries and the value): on return every lane l holds
// the warp total of v[l & 7].  9 shuffles instead of 40; fixed order.
__device__ __forceinline__ double warp_sum8(double (&v)[8])
{
    const int lane = threadIdx.x & 31;
    warp_sum_transposed_step8<4>(v, lane);
    double r = v[0];
    r += __shfl_xor_sync(0xffffffffu, r, 8);
    r += __shfl_xor_sync(0xffffffffu, r, 16);
    return r;
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

// Value, gradient and Hessian of the polynomial model over one cell (evaluations WITH Hessian; those without go through
// cell_value_grad with no slots).  Value and gradient are accumulated and reduced exactly as cell_value_grad does it --
// same per-lane order (pixels lane, lane + 32, ...), same warp_sum8 -- so that the two agree bit for bit.
__device__ __noinline__ void eval_cell_poly_h(const double* __restrict__ py, const uint32_t* __restrict__ prc, int npix,
                                                 const double* __restrict__ xr, const double* __restrict__ xc, const double* x,
                                                 double tol, double* sg, double* sh)
{
    const int lane = threadIdx.x & 31;
    const double th[6] = {x[0], x[1], x[2], x[3], x[4], x[5]};
    double acc[8], hs[24];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0;
#pragma unroll
    for (int k = 0; k < 24; ++k) hs[k] = 0;
    for (int i = lane; i < npix; i += 32) {
        const uint32_t rc = prc[i];
        const double x1 = xr[rc & 0xffffu], x2 = xc[rc >> 16], y = py[i];
        const double S = cell_S(th, x1, x2);
        double phi, theta;
        logistic(y, S, phi, theta);
        const double u = -y * theta;
        const double q[6] = {x1 * x1, x2 * x2, 2 * (x1 * x2), 2 * x1, 2 * x2, 1.0};
        acc[6] += phi;
#pragma unroll
        for (int k = 0; k < 5; ++k) acc[k] = fma(u, q[k], acc[k]);
        acc[5] += u;
        double kappa = theta - theta * theta;
        if (kappa < tol) kappa = 0;
        const double dd = kappa * y * y;
#pragma unroll
        for (int a = 0; a < 6; ++a) {
            const double da = dd * q[a];
#pragma unroll
            for (int b = 0; b <= a; ++b) hs[a * (a + 1) / 2 + b] = fma(da, q[b], hs[a * (a + 1) / 2 + b]);
        }
    }
    const double tot = warp_sum8(acc);                // lane: 0..5 gradient, 6 value
    if (lane < 6) sg[lane] = tot;
    if (lane == 6) sg[31] = tot;
#pragma unroll
    for (int g = 0; g < 3; ++g) {
        double v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = hs[8 * g + k];
        const double t = warp_sum8(v);
        if (lane < 8 && 8 * g + lane < 21) sh[8 * g + lane] = t;
    }
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

// ---- 1-D TMA bulk copies (cp.async.bulk, SASS: UBLKCP) tracked by mbarriers ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(void* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(void* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, void* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(void* bar, uint32_t parity)
{
    const uint32_t addr = smem_u32(bar);
    uint32_t done, polls = 0;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(addr), "r"(parity) : "memory");
        if (!done && ++polls > (1u << 24)) __trap();      // a copy that never arrives must not hang the device
    } while (!done);
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Every warp owns STAGES mbarriers and the word with their phase bits, behind the SLOTJ_INTS ints of its slot tables
// (static shared memory: it survives whatever the dynamic shared memory is used for between two items).
constexpr int SYNC_INTS = 8;      // 3 x 8-byte mbarriers + phase bits + pad
constexpr int WARP_INTS = SLOTJ_INTS + SYNC_INTS;
__device__ __forceinline__ void eval_sync_init(int* s_slotj /* [EVAL_WARPS * WARP_INTS], 16-byte aligned */)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int* w = s_slotj + warp * WARP_INTS + SLOTJ_INTS;
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < STAGES; ++k) mbar_init(w + 2 * k, 1);
        w[2 * STAGES] = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        fence_proxy_async();
    }
    __syncthreads();
}

// true in exactly one lane of the (converged) warp
__device__ __forceinline__ bool elect_one()
{
    uint32_t pred;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
    return pred != 0;
}

// Phase 2 of cell_value_grad for a panel whose rows are GS floats apart (a compile-time stride: every shared-memory read has
// an immediate offset).  Lane = active slot: g[p & 3] += u[p] * G~[p][slot] over the chunk's pixels, in this order.
template <int GS>
__device__ __forceinline__ void grad_phase2(const float* __restrict__ Gs, const double* __restrict__ U, int cnt,
                                            double& g0, double& g1, double& g2, double& g3)
{
    const double2* U2 = reinterpret_cast<const double2*>(U);
#ifndef SDSM_P2_UNROLL
#define SDSM_P2_UNROLL 8          // full unrolling (32) is 6 % faster in the stand-alone sweep and 2 % slower in the solver (code size)
#endif
    if (cnt == 32) {
#pragma unroll 1
        for (int q = 0; q < 32; q += SDSM_P2_UNROLL) {
            const float* Gq = Gs + q * GS;
            const double2* Uq = U2 + (q >> 1);
#pragma unroll
            for (int p = 0; p < SDSM_P2_UNROLL; p += 4) {
                const float f0 = Gq[p * GS], f1 = Gq[(p + 1) * GS], f2 = Gq[(p + 2) * GS], f3 = Gq[(p + 3) * GS];
                const double2 ua = Uq[p >> 1], ub = Uq[(p >> 1) + 1];
                g0 = fma(ua.x, (double)f0, g0); g1 = fma(ua.y, (double)f1, g1);
                g2 = fma(ub.x, (double)f2, g2); g3 = fma(ub.y, (double)f3, g3);
            }
        }
        return;
    }
    int p = 0;
    for (; p + 4 <= cnt; p += 4) {
        const float f0 = Gs[p * GS], f1 = Gs[(p + 1) * GS], f2 = Gs[(p + 2) * GS], f3 = Gs[(p + 3) * GS];
        const double2 ua = U2[p >> 1], ub = U2[(p >> 1) + 1];
        g0 = fma(ua.x, (double)f0, g0); g1 = fma(ua.y, (double)f1, g1);
        g2 = fma(ub.x, (double)f2, g2); g3 = fma(ub.y, (double)f3, g3);
    }
    if (p < cnt) g0 = fma(U[p], (double)Gs[p * GS], g0);
    if (p + 1 < cnt) g1 = fma(U[p + 1], (double)Gs[(p + 1) * GS], g1);
    if (p + 2 < cnt) g2 = fma(U[p + 2], (double)Gs[(p + 2) * GS], g2);
}

// one lane: brings chunk k (pixels 32 k .. of the cell) into `stage`: G~ rows, y, row|col.  py / prc point at the cell's first
// pixel; a TMA source must be 16-byte aligned, so the y copy starts ay = 0/1 and the row|col copy arc = 0..3 pixels early
// (and both are rounded up to 16 bytes: the arrays are padded to multiples of 32 pixels per candidate).
__device__ __forceinline__ void stage_issue(char* stage, void* bar, const double* __restrict__ py, const uint32_t* __restrict__ prc,
                                            const float* __restrict__ gcell, int npix, int gstride, int ay, int arc, int k)
{
    const int p0 = 32 * k, cnt = min(32, npix - p0);
    const uint32_t by = (uint32_t)((cnt + ay + 1) & ~1) * 8u, brc = (uint32_t)((cnt + arc + 3) & ~3) * 4u;
    const uint32_t bg = (uint32_t)(cnt * gstride) * 4u;               // gstride = 0: polynomial model, no panel
    mbar_expect_tx(bar, by + brc + bg);
    tma_bulk_g2s(stage + STAGE_G_BYTES, py + p0 - ay, by, bar);
    tma_bulk_g2s(stage + STAGE_G_BYTES + STAGE_Y_BYTES, prc + p0 - arc, brc, bar);
    if (bg) tma_bulk_g2s(stage, gcell + (size_t)p0 * gstride, bg, bar);
}

// Value and gradient of one cell.  Two phases per chunk of 32 pixels:
//   (1) lane = pixel: S = polynomial + sum_a G~[p][a] xi[a], logistic loss -> phi, theta; the value and the six theta-gradient
//       entries accumulate in registers (7 per lane); u = -y theta goes to shared memory;
//   (2) lane = active slot a: xi-gradient entry a += sum_p u[p] G~[p][a], straight from the staged panel (one float per pixel:
//       consecutive lanes read consecutive words), four independent partial sums in a fixed order.
// So a lane never carries 32 accumulators (the 128-register budget of two resident CTAs is left to the logistic function's
// two FMA chains), and no transposed 32-value reduction ends the cell.  Both kinds of evaluation (with / without Hessian)
// call this one function, and every multiply-add in it is an explicit fma (no context-dependent contraction): the line
// searches of cp compare merit values of evaluations with and without Hessian at the same point and rely on equal bits.
// Returns (.x) in lane e < 6 the theta-gradient entry e and in lane 6 the value (lanes 8.. repeat them), (.y) in lane
// a < na the xi-gradient entry of the cell's a-th active slot.
__device__ __noinline__ double2 cell_value_grad(const double* __restrict__ py, const uint32_t* __restrict__ prc, const float* __restrict__ gcell,
                                                int npix, int na, const double* __restrict__ xr, const double* __restrict__ xc,
                                                const double* x, double tol, char* stages, double* XI, int* sync)
{
    const int lane = threadIdx.x & 31;
    const double th[6] = {x[0], x[1], x[2], x[3], x[4], x[5]};
    double acc[8];                                    // 0..5 theta gradient, 6 value
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0;
    double g0 = 0, g1 = 0, g2 = 0, g3 = 0;            // xi-gradient entry of slot `lane`: partial sums over the pixels p = 0, 1, 2, 3 mod 4
    double* U = XI + 32;
    const int nchunks = (npix + 31) >> 5;
    const int gstride = na > 0 ? grow_stride(na) : 0;
    const int ay = (int)((reinterpret_cast<uintptr_t>(py) >> 3) & 1u), arc = (int)((reinterpret_cast<uintptr_t>(prc) >> 2) & 3u);
    uint32_t phase = (uint32_t)sync[2 * STAGES];
    __syncwarp();
    if (elect_one()) {
        fence_proxy_async();                          // the staging area may have been written with ordinary stores (Hessian tile)
#pragma unroll
        for (int k = 0; k < STAGES - 1; ++k)
            if (k < nchunks) stage_issue(stages + k * STAGE_BYTES, sync + 2 * k, py, prc, gcell, npix, gstride, ay, arc, k);
    }
    int slot = 0, slot_in = STAGES - 1;
    const int la = lane < gstride ? lane : 0;         // lanes beyond the panel's row read slot 0 (and drop the result)
    for (int k = 0; k < nchunks; ++k) {
        __syncwarp();                                 // everybody is done with chunk k - 1: its buffer and U are free
        if (k + STAGES - 1 < nchunks && elect_one())
            stage_issue(stages + slot_in * STAGE_BYTES, sync + 2 * slot_in, py, prc, gcell, npix, gstride, ay, arc, k + STAGES - 1);
        mbar_wait(sync + 2 * slot, (phase >> slot) & 1u);
        phase ^= 1u << slot;
        const char* stage = stages + slot * STAGE_BYTES;
        const int cnt = min(32, npix - 32 * k);
        const bool act = lane < cnt;
        // ---- phase 1: lane = pixel ----
        const uint32_t rc = act ? reinterpret_cast<const uint32_t*>(stage + STAGE_G_BYTES + STAGE_Y_BYTES)[arc + lane] : 0u;
        const double y = act ? reinterpret_cast<const double*>(stage + STAGE_G_BYTES)[ay + lane] : 0.0;
        // (tried: the cell's rows / columns of the coordinate tables in shared memory -- the sweep gained 5 %, the solver lost
        //  4 %: one more dependent global load in front of every cell)
        const double x1 = xr[rc & 0xffffu], x2 = xc[rc >> 16];
        double S = cell_S(th, x1, x2);
        if (na > 0) {                                 // (warp-uniform)
            const float4* G4 = reinterpret_cast<const float4*>(stage) + lane * (gstride >> 2);
            double sx0 = 0, sx1 = 0, sx2 = 0, sx3 = 0;    // XI is zero beyond the cell's last active slot, and so is the row
#pragma unroll
            for (int j = 0; j < GROW / 4; ++j) {
                if (4 * j < na) {                         // (warp-uniform)
                    const float4 g = G4[j];
                    const double2 xa = reinterpret_cast<const double2*>(XI)[2 * j], xb = reinterpret_cast<const double2*>(XI)[2 * j + 1];
                    sx0 = fma((double)g.x, xa.x, sx0); sx1 = fma((double)g.y, xa.y, sx1);
                    sx2 = fma((double)g.z, xb.x, sx2); sx3 = fma((double)g.w, xb.y, sx3);
                }
            }
            S += (sx0 + sx1) + (sx2 + sx3);
        }
        double phi = 0, theta = 0;
        if (act) logistic(y, S, phi, theta);          // lanes beyond the cell's last pixel: S may be garbage, never used
        const double u = -y * theta;                  // 0 for those lanes (y = 0)
        const double q[5] = {x1 * x1, x2 * x2, 2 * (x1 * x2), 2 * x1, 2 * x2};
        acc[6] += phi;
#pragma unroll
        for (int e = 0; e < 5; ++e) acc[e] = fma(u, q[e], acc[e]);
        acc[5] += u;
        if (na > 0) {
            // the xi part of the gradient drops |u| < sparsity_tol (dsm.py:346); the theta part does not
            U[lane] = (tol > 0 && fabs(u) < tol) ? 0.0 : u;
            __syncwarp();
            // ---- phase 2: lane = active slot ----
            const float* Gs = reinterpret_cast<const float*>(stage) + la;
            if (gstride == GROW) grad_phase2<GROW>(Gs, U, cnt, g0, g1, g2, g3);
            else if (gstride == 20) grad_phase2<20>(Gs, U, cnt, g0, g1, g2, g3);
            else if (gstride == 12) grad_phase2<12>(Gs, U, cnt, g0, g1, g2, g3);
            else grad_phase2<4>(Gs, U, cnt, g0, g1, g2, g3);
        }
        slot = (slot + 1 == STAGES) ? 0 : slot + 1;
        slot_in = (slot_in + 1 == STAGES) ? 0 : slot_in + 1;
    }
    __syncwarp();
    if (lane == 0) sync[2 * STAGES] = (int)phase;
    double2 r;
    r.x = warp_sum8(acc);
    r.y = (g0 + g1) + (g2 + g3);
    return r;
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
            const float4* gp = reinterpret_cast<const float4*>(gcell + (size_t)(act ? p : 0) * grow_stride(na));
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

// MODE 0: value + gradient; 1: value + gradient + local Hessian; 2: local Hessian only
template <int MODE>
__device__ __forceinline__ void eval_cell_xi(const double* __restrict__ py, const uint32_t* __restrict__ prc, const float* __restrict__ gcell,
                                             int npix, uint32_t am, const double* __restrict__ xr, const double* __restrict__ xc,
                                             const double* x, double tol, double* Tt, int* slotj,
                                             double* sg, double* xig, double* sh)
{
    const int lane = threadIdx.x & 31;
    double* XI = Tt + XI_OFF;
    const int na = __popc(am);
    __syncwarp();
    XI[lane] = lane < na ? x[6 + slotj[lane]] : 0.0;  // the xi values this cell sees (zero-padded: rows are read in groups of four)
    __syncwarp();
    if (MODE != 2) {
        const double2 tot = cell_value_grad(py, prc, gcell, npix, na, xr, xc, x, tol, reinterpret_cast<char*>(Tt), XI, slotj + SLOTJ_INTS);
        if (lane < 6) sg[lane] = tot.x;
        if (lane == 6) sg[31] = tot.x;
        // xi gradient: straight into the slot that grid point slotj[a] keeps for this cell (seen from the grid point's cell, this
        // cell sits at the mirrored neighbour position), so that the control task sums 25 contiguous numbers per grid point
        // instead of chasing cell -> mask -> scratch entry
        if (lane < na) xig[(size_t)slotj[lane] * NSLOT + (NSLOT - 1 - slotj[32 + lane])] = tot.y;
        __syncwarp();
    }
    if (MODE != 0) cell_hessian(py, prc, gcell, npix, na, xr, xc, x, tol, Tt, XI, sh);
    __syncwarp();
}

// ---------------------------------------------------------------------------------------------------------------

constexpr size_t EVAL_SMEM = sizeof(double) * EVAL_WARPS * TILE_DOUBLES;

// Evaluation task (c, first): the CTA's warps share the non-empty cells cell_list[first .. first + EVAL_CELLS) of candidate c
// -- a warp takes the next cell when it is done with its last one (cells differ in their pixel counts) -- and leave the
// per-cell value / gradient / Hessian in scrG / scrH.  smem_eval: EVAL_SMEM bytes; *s_next == 0 on entry (all threads
// synchronised after it was set), and the caller synchronises the CTA after the call.
__device__ __forceinline__ void eval_cell(const Par& par, const sdsm_cand* __restrict__ cands, const WS& ws, int c, int idx, int pt,
                                          double* Tt, int* slotj);

__device__ __forceinline__ void eval_task(const Par& par, const sdsm_cand* __restrict__ cands, const WS& ws, int c, int first,
                                          double* smem_eval, int* s_slotj, int* s_next, int npts = 1)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int count = min(ws.req[c].cpt, ws.cinfo[c].n_ne - first);
    for (;;) {
        int i = 0;
        if (lane == 0) i = atomicAdd(s_next, 1);
        i = __shfl_sync(0xffffffffu, i, 0);
        if (i >= count * npts) break;                // the cells of all npts trial points of the request share the warps
        eval_cell(par, cands, ws, c, first + i % count, i / count, smem_eval + warp * TILE_DOUBLES, s_slotj + warp * WARP_INTS);
    }
}

__device__ __forceinline__ void eval_cell(const Par& par, const sdsm_cand* __restrict__ cands, const WS& ws, int c, int idx, int pt,
                                          double* Tt, int* slotj)
{
    const int lane = threadIdx.x & 31;
    const sdsm_cand* cd = cands + c;
    const int64_t pix_off = cd->pix_off, cell_off = cd->cell_off, tab_off = cd->tab_off;
    const int cell = ws.cell_list[cell_off + idx];
    const int h = cd->h;
    EvalReq rq = ws.req[c];
    rq.x += (size_t)pt * rq.ld;                      // the pt-th trial point of the request (pt = 0 unless it is a speculative one)
    if (pt != rq.npts - 1) rq.want_h = 0;            // a Hessian asked for with several trial points is the last point's
    const int32_t* cstart = ws.cell_start + cell_off;
    const int st = cstart[cell], npix = cstart[cell + 1] - st;
    const double* py = ws.pix_y + pix_off + st;
    const uint32_t* prc = ws.pix_rc + pix_off + st;
    const double* xr = ws.xtab + tab_off;
    const double* xc = xr + h;
    double* sg = ws.scrG + (size_t)pt * ws.sg_stride + (size_t)(cell_off + cell) * SCRG;
    double* sh = ws.scrH + (size_t)(cell_off + cell) * SCRH;
    const int cw = ws.cinfo[c].cw;
    const int ca = cell / cw, cb = cell - ca * cw;
    if (!rq.use_xi) {
        if (rq.want_h) eval_cell_poly_h(py, prc, npix, xr, xc, rq.x, par.sparsity_tol, sg, sh);
        else {
            // value + gradient of the polynomial model = the deformable pass without slots: same arithmetic in the same
            // order as eval_cell_poly_h (the values of the two must agree bit for bit), with y and row|col arriving through
            // the TMA pipeline instead of dependent loads
            const double2 tot = cell_value_grad(py, prc, nullptr, npix, 0, xr, xc, rq.x, par.sparsity_tol, reinterpret_cast<char*>(Tt), Tt + XI_OFF, slotj + SLOTJ_INTS);
            if (lane < 6) sg[lane] = tot.x;
            if (lane == 6) sg[31] = tot.x;
        }
        return;
    }
    const uint32_t am = ws.cell_amask[cell_off + cell];
    const int na = __popc(am);
    __syncwarp();
    if (lane < NSLOT && ((am >> lane) & 1u)) slotj[32 + __popc(am & ((1u << lane) - 1u))] = lane;     // slot of the cell's e-th active entry
    __syncwarp();
    if (lane < na) {
        const int s = slotj[32 + lane];
        slotj[lane] = ws.cellmap[cell_off + (ca + s / 5 - NC) * cw + (cb + s % 5 - NC)];
    }
    __syncwarp();
    const float* gcell = ws.gval + (size_t)GROW * (pix_off + st);
    double* xig = ws.xig + (size_t)pt * ws.xig_stride + (size_t)cell_off * NSLOT;
    if (rq.want_h == 2) eval_cell_xi<2>(py, prc, gcell, npix, am, xr, xc, rq.x, par.sparsity_tol, Tt, slotj, sg, xig, sh);
    else if (rq.want_h) eval_cell_xi<1>(py, prc, gcell, npix, am, xr, xc, rq.x, par.sparsity_tol, Tt, slotj, sg, xig, sh);
    else eval_cell_xi<0>(py, prc, gcell, npix, am, xr, xc, rq.x, par.sparsity_tol, Tt, slotj, sg, xig, sh);
}

}  // namespace sdsm
