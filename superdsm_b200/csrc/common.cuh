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
constexpr int SOLVER_BLOCK = 256;
constexpr int SOLVER_WARPS = SOLVER_BLOCK / 32;
constexpr int TILE_DOUBLES = 32 * TILE_LD + 64;                 // tile + u[32] + dd[32]
constexpr int SMEM_L_CAP   = SOLVER_WARPS * TILE_DOUBLES;       // doubles available for the packed Cholesky factor
constexpr int NVEC       = 12;    // N-vectors of the interior-point solver
constexpr int NV_SMEM    = 192;   // vectors live in shared memory up to this N

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
};

// Workspace carved out of the caller's buffer.  Per-pixel / per-cell / per-table arrays are indexed with the
// candidate's offsets (sdsm_cand.pix_off / cell_off / tab_off).
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
    double*   xtab;        // [T]      normalised coordinates x1 of compacted rows, then x2 of compacted columns (image.py:6-21)
    uint16_t* inv;         // [T]      compacted row -> box row, then compacted col -> box col
    uint16_t* fwd;         // [T]      box row -> compacted row (0xFFFF: empty), then columns
    CInfo*    cinfo;       // [n_cand]
    int32_t*  counters;    // [16]
    char*     cta;         // per-CTA solver workspace
    size_t    cta_bytes, off_H, off_H2, off_L, off_scrH, off_scrG, off_vec;
};

struct Layout {
    size_t total;
    size_t o_pix_y, o_pix_rc, o_pix_s, o_pix_dist, o_gval, o_cell_start, o_cell_amask, o_cellmap, o_grid_rc, o_grid_tmp,
           o_xtab, o_inv, o_fwd, o_cinfo, o_counters, o_cta;
    size_t cta_bytes, off_H, off_H2, off_L, off_scrH, off_scrG, off_vec;
};

inline size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }

inline Layout make_layout(int64_t P, int64_t C, int64_t T, int32_t n_cand, int32_t n_max, int32_t cells_max, int32_t ctas)
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
    L.o_xtab = take(sizeof(double) * T);
    L.o_inv = take(sizeof(uint16_t) * T);
    L.o_fwd = take(sizeof(uint16_t) * T);
    L.o_cinfo = take(sizeof(CInfo) * (size_t)n_cand);
    L.o_counters = take(sizeof(int32_t) * 16);
    size_t c = 0;
    auto ctake = [&](size_t bytes) { size_t at = c; c = align_up(c + bytes); return at; };
    L.off_H = ctake(sizeof(double) * (size_t)n_max * n_max);
    L.off_H2 = ctake(sizeof(double) * (size_t)n_max * n_max);
    L.off_L = ctake(sizeof(double) * ((size_t)n_max * (n_max + 1) / 2));
    L.off_scrH = ctake(sizeof(double) * (size_t)cells_max * SCRH);
    L.off_scrG = ctake(sizeof(double) * (size_t)cells_max * SCRG);
    L.off_vec = ctake(sizeof(double) * (size_t)NVEC * n_max);
    L.cta_bytes = c;
    L.o_cta = take(c * (size_t)ctas);
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
    w.xtab = (double*)(b + L.o_xtab);
    w.inv = (uint16_t*)(b + L.o_inv);
    w.fwd = (uint16_t*)(b + L.o_fwd);
    w.cinfo = (CInfo*)(b + L.o_cinfo);
    w.counters = (int32_t*)(b + L.o_counters);
    w.cta = b + L.o_cta;
    w.cta_bytes = L.cta_bytes;
    w.off_H = L.off_H; w.off_H2 = L.off_H2; w.off_L = L.off_L; w.off_scrH = L.off_scrH; w.off_scrG = L.off_scrG; w.off_vec = L.off_vec;
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
