// K2: the sub-sampled smooth matrix G~ of every candidate, in closed form (dsm.py:137-235), never materialised as a matrix.
//
//  * grid points = omega AND lattice, then the greedy completion of dsm.py:169-180 (chessboard distances),
//  * grid points are pairwise >= subsample apart (chessboard), so every subsample x subsample cell holds at most one;
//    a pixel only sees the grid points of the 5 x 5 cells around its own (R <= 2 subsample),
//  * M[i, j] = psf32[|dr|, |dc|] for |dr|, |dc| <= R, row sums in float32 in numpy's pairwise order over ALL m columns
//    (dsm.py:192; zeros are exact no-ops, so only the non-zeros and their column indices matter),
//  * G~[i, j] = float32 division M[i, j] / rowsum (dsm.py:193), stored as float32 per (cell, active slot, pixel).
#pragma once
#include "common.cuh"

namespace sdsm {

constexpr int GRID_BLOCK = 256;

// numpy's float32 pairwise sum over columns [lo, lo + n) of a row whose non-zeros (col[k], val[k]), k in [0, cnt), are
// sorted by column; kp is the running cursor into the non-zeros.
__device__ float pairwise_f32(const int* col, const float* val, int cnt, int& kp, int lo, int n)
{
    if (n < 8) {
        float res = 0.f;
        while (kp < cnt && col[kp] < lo + n) { res = __fadd_rn(res, val[kp]); ++kp; }
        return res;
    }
    if (n <= 128) {
        float r[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) r[k] = 0.f;
        const int body = n - (n % 8);
        while (kp < cnt && col[kp] < lo + body) {
            const int k = (col[kp] - lo) & 7;
            const float v = val[kp];
            // static indexing keeps r[] in registers
#pragma unroll
            for (int q = 0; q < 8; ++q) if (q == k) r[q] = __fadd_rn(r[q], v);
            ++kp;
        }
        float res = __fadd_rn(__fadd_rn(__fadd_rn(r[0], r[1]), __fadd_rn(r[2], r[3])),
                              __fadd_rn(__fadd_rn(r[4], r[5]), __fadd_rn(r[6], r[7])));
        while (kp < cnt && col[kp] < lo + n) { res = __fadd_rn(res, val[kp]); ++kp; }
        return res;
    }
    int n2 = n / 2;
    n2 -= n2 % 8;
    const float a = pairwise_f32(col, val, cnt, kp, lo, n2);
    const float b = pairwise_f32(col, val, cnt, kp, lo + n2, n - n2);
    return __fadd_rn(a, b);
}

__device__ __forceinline__ int cheb(uint32_t a, uint32_t b)
{
    const int dr = (int)(a & 0xffffu) - (int)(b & 0xffffu);
    const int dc = (int)(a >> 16) - (int)(b >> 16);
    return max(abs(dr), abs(dc));
}

#ifdef SDSM_GRID_TRACE
#define GRID_T0 long long gt_[5]; gt_[4] = clock64()
#define GRID_T(k) gt_[k] = clock64()
#define GRID_TPRINT if (tid == 0 && it < 3) printf("k_grid it %d n %d m %d far %d: a %lld b %lld c %lld d %lld cycles\n", it, n, m, nfar, gt_[0] - gt_[4], gt_[1] - gt_[0], gt_[2] - gt_[1], gt_[3] - gt_[2])
#else
#define GRID_T0
#define GRID_T(k)
#define GRID_TPRINT
#endif

__global__ void __launch_bounds__(GRID_BLOCK)
k_grid(Par par, const sdsm_cand* __restrict__ cands, int n_cand, const int32_t* __restrict__ order, WS ws)
{
    constexpr int BLOCK = GRID_BLOCK;
    __shared__ unsigned long long s_min[BLOCK / 32 + 1];
    __shared__ int s_cnt;
    __shared__ int s_far;
    const int tid = threadIdx.x;
    const int sub = par.sub, R = par.R;

    for (int it = blockIdx.x; it < n_cand; it += gridDim.x) {
        const int c = order[it];
        const sdsm_cand cd = cands[c];
        CInfo ci = ws.cinfo[c];
        __syncthreads();
        if (ci.flags & 1) continue;
        // null matrix: bounding box not larger than the kernel radius in some dimension (dsm.py:187)
        if (ci.hp <= R || ci.wp <= R) continue;
        const int n = cd.n, ncells = ci.ncells, cw = ci.cw, ch = ci.ch, wp = ci.wp;
        const int32_t* cstart = ws.cell_start + cd.cell_off;
        const uint32_t* prc = ws.pix_rc + cd.pix_off;
        int32_t* cellmap = ws.cellmap + cd.cell_off;
        uint32_t* gl = ws.grid_tmp + cd.cell_off;        // unordered grid points
        uint32_t* gs = ws.grid_rc + cd.cell_off;         // sorted grid points
        int32_t* far_d = ws.pix_dist + cd.pix_off;       // pixels farther than sub from every lattice point: distance,
        uint32_t* far_p = reinterpret_cast<uint32_t*>(ws.pix_s + cd.pix_off);   // and point (pix_s is written by k_grid_rows)
        if (tid == 0) { s_cnt = 0; s_far = 0; }
        GRID_T0;
        __syncthreads();

        // (a) lattice points inside omega: the cell's first pixel sits on the cell origin
        for (int cell = tid; cell < ncells; cell += BLOCK) {
            const int st = cstart[cell];
            int v = -1;
            if (cstart[cell + 1] > st) {
                const int ca = cell / cw, cb = cell - ca * cw;
                const uint32_t org = (uint32_t)(ca * sub) | ((uint32_t)(cb * sub) << 16);
                if (prc[st] == org) { v = 0; gl[atomicAdd(&s_cnt, 1)] = org; }
            }
            cellmap[cell] = v;
        }
        __syncthreads();
        int m = s_cnt;
        GRID_T(0);

        // (b) chessboard distance of every pixel to the lattice points (distance_transform_bf, dsm.py:169)
        //     Only the four lattice corners around a pixel can be closer than `sub` (every other lattice point is at least
        //     `sub` rows or columns away), and the completion below only orders the pixels whose distance is >= sub: the
        //     scan over all m lattice points is needed for the few pixels without a lattice corner inside omega.
        //     Pixels at distance >= sub are the only ones the completion looks at: they go to a compact list (point, distance).
        for (int i = tid; i < n; i += BLOCK) {
            const uint32_t p = prc[i];
            const int pr = (int)(p & 0xffffu), pc = (int)(p >> 16);
            const int a = pr / sub, b = pc / sub;
            int d = 0x7fffffff;
#pragma unroll
            for (int da = 0; da < 2; ++da)
#pragma unroll
                for (int db = 0; db < 2; ++db) {
                    const int ca = a + da, cb = b + db;
                    if (ca < ch && cb < cw && cellmap[ca * cw + cb] >= 0) d = min(d, max(abs(ca * sub - pr), abs(cb * sub - pc)));
                }
            if (d >= sub) {
                d = 0x7fffffff;
                for (int k = 0; k < m; ++k) d = min(d, cheb(p, gl[k]));
                if (d >= sub) { const int pos = atomicAdd(&s_far, 1); far_p[pos] = p; far_d[pos] = d; }
            }
        }
        __syncthreads();
        GRID_T(1);

        // (c) greedy completion: first (row-major) pixel of minimal distance among those >= subsample
        const int nfar = s_far;
        uint32_t newp = 0;
        bool have_new = false;
        while (nfar > 0) {
            unsigned long long best = ~0ull;
            for (int i = tid; i < nfar; i += BLOCK) {
                const uint32_t p = far_p[i];
                int d = far_d[i];
                if (have_new) { d = min(d, cheb(p, newp)); far_d[i] = d; }
                if (d >= sub) {
                    const unsigned long long key = ((unsigned long long)(uint32_t)d << 32) | (uint32_t)((p & 0xffffu) * (uint32_t)wp + (p >> 16));
                    best = key < best ? key : best;
                }
            }
            best = block_min_u64<BLOCK>(best, s_min);
            if (best == ~0ull) break;
            const uint32_t lin = (uint32_t)(best & 0xffffffffu);
            newp = (lin / (uint32_t)wp) | ((lin % (uint32_t)wp) << 16);
            have_new = true;
            if (tid == 0) {
                gl[m] = newp;
                const int cell = ((newp & 0xffffu) / sub) * cw + ((newp >> 16) / sub);
                cellmap[cell] = 0;
            }
            ++m;
            __syncthreads();
        }
        GRID_T(2);

        // (d) row-major order of the grid points = column order of G~
        for (int k = tid; k < m; k += BLOCK) {
            const uint32_t p = gl[k];
            const uint32_t key = (p & 0xffffu) * (uint32_t)wp + (p >> 16);
            int rank = 0;
            for (int q = 0; q < m; ++q) { const uint32_t o = gl[q]; rank += ((o & 0xffffu) * (uint32_t)wp + (o >> 16)) < key; }
            gs[rank] = p;
            cellmap[((p & 0xffffu) / sub) * cw + ((p >> 16) / sub)] = rank;
        }
        __syncthreads();

        GRID_T(3);
        if (tid == 0) {
            ci.m = m;
            ws.cinfo[c] = ci;
        }
        GRID_TPRINT;
    }
}

// K2b: the per-pixel part of the smooth matrix, split over ROWS_SPLIT CTAs per candidate (the largest region of a batch
// would otherwise keep one CTA busy for ~12 ms): CTA (it, part) takes a contiguous range of cells of candidate order[it]
// -- and with it a contiguous range of the cell-major pixel arrays.
constexpr int ROWS_SPLIT = 16;

__global__ void __launch_bounds__(GRID_BLOCK)
k_grid_rows(Par par, const sdsm_cand* __restrict__ cands, int n_cand, const int32_t* __restrict__ order, WS ws)
{
    constexpr int BLOCK = GRID_BLOCK;
    extern __shared__ float s_psf[];                    // (R+1)^2
    __shared__ int s_bad;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int sub = par.sub, R = par.R, R1 = par.R + 1;
    for (int i = tid; i < R1 * R1; i += BLOCK) s_psf[i] = par.psf[i];
    __syncthreads();

    for (int it = blockIdx.x; it < n_cand; it += gridDim.x) {
        const int c = order[it];
        const sdsm_cand cd = cands[c];
        const CInfo ci = ws.cinfo[c];
        __syncthreads();
        if (ci.flags & 1) continue;
        if (ci.hp <= R || ci.wp <= R) continue;          // null matrix
        const int ncells = ci.ncells, cw = ci.cw, ch = ci.ch, m = ci.m;
        const int per = (ncells + ROWS_SPLIT - 1) / ROWS_SPLIT;
        const int cell0 = min(ncells, (int)blockIdx.y * per), cell1 = min(ncells, cell0 + per);
        if (cell0 >= cell1) continue;
        const int32_t* cstart = ws.cell_start + cd.cell_off;
        const uint32_t* prc = ws.pix_rc + cd.pix_off;
        const int32_t* cellmap = ws.cellmap + cd.cell_off;
        const uint32_t* gs = ws.grid_rc + cd.cell_off;
        const int i0 = cstart[cell0], i1 = cstart[cell1];
        if (tid == 0) s_bad = 0;
        __syncthreads();

        // (e) float32 row sums in numpy's pairwise order
        float* ps = ws.pix_s + cd.pix_off;
        for (int i = i0 + tid; i < i1; i += BLOCK) {
            const uint32_t p = prc[i];
            const int pr = p & 0xffffu, pc = p >> 16;
            const int ca = pr / sub, cb = pc / sub;
            int col[NSLOT];
            float val[NSLOT];
            int cnt = 0;
            for (int da = -NC; da <= NC; ++da) {
                const int a = ca + da;
                if (a < 0 || a >= ch) continue;
                for (int db = -NC; db <= NC; ++db) {
                    const int b = cb + db;
                    if (b < 0 || b >= cw) continue;
                    const int j = cellmap[a * cw + b];
                    if (j < 0) continue;
                    const uint32_t g = gs[j];
                    const int dr = abs((int)(g & 0xffffu) - pr), dc = abs((int)(g >> 16) - pc);
                    if (dr > R || dc > R) continue;
                    // insertion keeps (col, val) sorted by column
                    int k = cnt++;
                    const float v = s_psf[dr * R1 + dc];
                    while (k > 0 && col[k - 1] > j) { col[k] = col[k - 1]; val[k] = val[k - 1]; --k; }
                    col[k] = j; val[k] = v;
                }
            }
            int kp = 0;
            const float s = pairwise_f32(col, val, cnt, kp, 0, m);
            ps[i] = s;
            if (!(s > 0.f)) s_bad = 1;
        }
        __syncthreads();

        // (f) G~ values per (cell, pixel, active slot) and the active-slot mask of every cell: the cell's block starts at
        //     GROW * (first pixel) and holds one row per pixel of grow_stride(#active slots) floats (16-byte groups, a stride
        //     that keeps the evaluation's 16-byte shared-memory reads free of bank conflicts), the active slots in slot
        //     order, zeros behind them
        float* gv = ws.gval + (size_t)GROW * cd.pix_off;
        uint32_t* amask = ws.cell_amask + cd.cell_off;
        for (int cell = cell0 + warp; cell < cell1; cell += BLOCK / 32) {
            const int st = cstart[cell], npix = cstart[cell + 1] - st;
            uint32_t am = 0;
            if (npix > 0) {
                const int ca = cell / cw, cb = cell - ca * cw;
                float* out = gv + (size_t)GROW * st;
                // active slots: grid points of the 5 x 5 neighbour cells within the kernel radius of some pixel of the cell
                for (int s = 0; s < NSLOT; ++s) {
                    const int a = ca + s / 5 - NC, b = cb + s % 5 - NC;
                    if (a < 0 || a >= ch || b < 0 || b >= cw) continue;
                    const int j = cellmap[a * cw + b];
                    if (j < 0) continue;
                    const uint32_t g = gs[j];
                    const int gr = g & 0xffffu, gc = g >> 16;
                    bool any = false;
                    for (int p = lane; p < npix; p += 32) {
                        const uint32_t q = prc[st + p];
                        any |= abs(gr - (int)(q & 0xffffu)) <= R && abs(gc - (int)(q >> 16)) <= R;
                    }
                    if (__any_sync(0xffffffffu, any)) am |= 1u << s;
                }
                const int na4 = grow_stride(__popc(am));     // row stride of this cell's block (floats)
                int rank = 0;
                for (int s = 0; s < NSLOT; ++s) {
                    if (!((am >> s) & 1u)) continue;
                    const uint32_t g = gs[cellmap[(ca + s / 5 - NC) * cw + (cb + s % 5 - NC)]];
                    const int gr = g & 0xffffu, gc = g >> 16;
                    for (int p = lane; p < npix; p += 32) {
                        const uint32_t q = prc[st + p];
                        const int dr = abs(gr - (int)(q & 0xffffu)), dc = abs(gc - (int)(q >> 16));
                        out[(size_t)p * na4 + rank] = (dr <= R && dc <= R) ? __fdiv_rn(s_psf[dr * R1 + dc], ps[st + p]) : 0.f;
                    }
                    ++rank;
                }
                for (int a = rank; a < na4; ++a)
                    for (int p = lane; p < npix; p += 32) out[(size_t)p * na4 + a] = 0.f;
            }
            if (lane == 0) amask[cell] = am;
        }
        __syncthreads();
        if (tid == 0 && s_bad) atomicOr(&ws.cinfo[c].flags, 2);
    }
}

}  // namespace sdsm
