// Shared definitions of the superdsm_b200 CUDA library (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/superdsm_b200.h"

namespace sdsm {

constexpr int NSLOT      = 25;    // (2*NC+1)^2 neighbour cells with NC = 2 >= ceil(R / subsample)
constexpr int NC         = 2;
constexpr int LOCAL_MAX  = 32;    // 6 polynomial + 25 slots, padded
constexpr int SCRH       = 496;   // packed lower triangle of a 31 x 31 local Hessian
constexpr int SCRG       = 32;    // local gradient (31) + value
constexpr int TILE_LD    = 34;    // row stride (doubles) of the per-warp pixel tile: conflict-free 128-bit rows
constexpr int EVAL_BLOCK = 256;   // k_eval: one warp per pixel cell
constexpr int EVAL_WARPS = EVAL_BLOCK / 32;
constexpr int TILE_DOUBLES = 32 * TILE_LD + 64;                 // tile + u[32] + dd[32]
constexpr int CTRL_BLOCK = 128;   // k_ctrl: one CTA per live candidate
constexpr int CTRL_WARPS = CTRL_BLOCK / 32;
constexpr int SMEM_L_CAP = 5504;  // doubles of shared memory for the packed Cholesky factor (N <= 104)
constexpr int SMALL_N_MAX = 104;  // systems up to this size are factorised in shared memory by k_ctrl, larger ones by k_ctrl_big
constexpr int BIG_BLOCK  = 256;   // k_ctrl_big
constexpr int NVEC       = 12;    // N-vectors of the interior-point solver

struct Img { const double* y; const int32_t* atoms; int32_t H, W; };

struct Par {
    double alpha, epsilon, scale, sparsity_tol, abstol, reltol, feastol;
    int32_t sub, R, has_smooth, init_elliptical, maxiters;
    const float* psf;   // device, (R+1)^2
};

struct CInfo {
    int32_t hp, wp;        // compacted bounding box (rows / columns of the box without any region pixel removed, dsm.py:185-186)
    int32_t ch, cw;        // cells per dimension
    int32_t ncells, m;     // m = number of deformation parameters (0: null matrix)
    int32_t n_fg, flags;   // pixels with y > 0; bit0: region pixel count mismatch, bit1: zero row in G~
    int32_t n_ne, pad;     // number of non-empty cells (entries of cell_list)
};

// State of one candidate's fit between two rounds of the lock-step solver (k_eval <-> k_ctrl).  The interior-point
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
    int32_t pc, stage, iters, relaxed, pass, use_xi;
    int32_t req_vec, req_gvec, req_h;      // the pending evaluation: point, where the gradient goes, Hessian wanted
    int32_t status, bits, it1, it2, it3, nf, nh, have1, r_ipm;
};

// what k_eval needs to know about a live candidate (written by k_ctrl when it files the request)
struct EvalReq {
    const double* x;       // point to evaluate (6 + m values, or 6)
    int32_t want_h, use_xi;
};

// Workspace carved out of the caller's buffer.  Per-pixel / per-cell / per-table arrays are indexed with the
// candidate's offsets (sdsm_cand.pix_off / cell_off / tab_off / vec_off / mat_off).
struct WS {
    double*   pix_y;       // [P]      y of the region's pixels, cell-major
    uint32_t* pix_rc;      // [P]      compacted (row | col << 16) inside the bounding box
    float*    pix_s;       // [P]      float32 row sum of the un-normalised G~ row (dsm.py:192)
    int32_t*  pix_dist;    // [P]      chessboard distance to the sub-sample grid (dsm.py:169-179)
    float*    gval;        // [25 P]   G~ values: cell -> active slot -> pixel
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
    double*   scrG;        // [C * 32]   per-cell value + local gradient of the last evaluation
    double*   scrH;        // [C * 496]  per-cell packed local Hessian of the last evaluation
    double*   vecs;        // [V]      NVEC vectors of length 6 + cell_cap per candidate
    double*   mats;        // [M]      H, H2 (N x N) and the packed Cholesky factor per candidate
    uint2*    items;       // [2 * C]  work lists of k_eval (candidate, cell), double-buffered by round parity
    uint2*    hitems;      // [2 * C]  work lists of k_hgather (candidate, grid point), double-buffered
    int32_t*  live;        // [2 * n_cand]  live candidates of a round (small systems from the front, big ones from the back)
    unsigned long long* counters;   // [16]  [p]: (#live small << 32 | #items), [2 + p]: (#live big << 32 | #hitems), p = parity
    int64_t   items_cap;
    int32_t   n_cand;
};

struct Layout {
    size_t total;
    size_t o_pix_y, o_pix_rc, o_pix_s, o_pix_dist, o_gval, o_cell_start, o_cell_amask, o_cellmap, o_grid_rc, o_grid_tmp, o_cell_list,
           o_xtab, o_inv, o_fwd, o_cinfo, o_cstate, o_req, o_scrG, o_scrH, o_vecs, o_mats, o_items, o_hitems, o_live, o_counters;
    int64_t C;
    int32_t n_cand;
};

inline size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }

// P, C, T, V, M: totals of the per-pixel, per-cell, per-table, per-vector and per-matrix arrays of the batch
inline Layout make_layout(int64_t P, int64_t C, int64_t T, int64_t V, int64_t M, int32_t n_cand)
{
    Layout L{};
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes); return at; };
    L.o_pix_y = take(sizeof(double) * P);
    L.o_pix_rc = take(sizeof(uint32_t) * P);
    L.o_pix_s = take(sizeof(float) * P);
    L.o_pix_dist = take(sizeof(int32_t) * P);
    L.o_gval = take(sizeof(float) * NSLOT * P);
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
    L.o_scrG = take(sizeof(double) * SCRG * C);
    L.o_scrH = take(sizeof(double) * SCRH * C);
    L.o_vecs = take(sizeof(double) * V);
    L.o_mats = take(sizeof(double) * M);
    L.o_items = take(sizeof(uint2) * 2 * C);
    L.o_hitems = take(sizeof(uint2) * 2 * C);
    L.o_live = take(sizeof(int32_t) * 2 * (size_t)n_cand);
    L.o_counters = take(sizeof(unsigned long long) * 16);
    L.C = C;
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
    w.scrG = (double*)(b + L.o_scrG);
    w.scrH = (double*)(b + L.o_scrH);
    w.vecs = (double*)(b + L.o_vecs);
    w.mats = (double*)(b + L.o_mats);
    w.items = (uint2*)(b + L.o_items);
    w.hitems = (uint2*)(b + L.o_hitems);
    w.live = (int32_t*)(b + L.o_live);
    w.counters = (unsigned long long*)(b + L.o_counters);
    w.items_cap = L.C;
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

}  // namespace sdsm
