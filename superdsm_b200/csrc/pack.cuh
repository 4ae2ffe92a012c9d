// K1: region extraction and ragged packing.
//
// For every candidate: omega = {pixels of the bounding box whose (masked) atom label is in the footprint}
// (objects.py:76-128; the candidate-independent part of the mask is folded into the atoms map by the host),
// removal of empty rows / columns (dsm.py:185-186), normalised coordinates (image.py:6-21), and packing of
// (y, row, col) cell-major: cells are subsample x subsample tiles of the compacted box, pixels row-major within a cell.
#pragma once
#include "common.cuh"

namespace sdsm {

constexpr int PACK_BLOCK = 256;

template <int BLOCK>
__device__ int compact_axis(uint16_t* fwd, uint16_t* inv, int len, int* s_scan)
{
    // fwd[i] holds 0/1 occupancy on entry; on exit fwd[i] = compacted index or 0xFFFF, inv[k] = i. Returns the count.
    const int tid = threadIdx.x;
    const int per = (len + BLOCK - 1) / BLOCK;
    const int lo = min(tid * per, len), hi = min(lo + per, len);
    int cnt = 0;
    for (int i = lo; i < hi; ++i) cnt += fwd[i] ? 1 : 0;
    __syncthreads();
    s_scan[tid] = cnt;
    __syncthreads();
    if (tid == 0) {
        int run = 0;
        for (int k = 0; k < BLOCK; ++k) { int v = s_scan[k]; s_scan[k] = run; run += v; }
        s_scan[BLOCK] = run;
    }
    __syncthreads();
    int run = s_scan[tid];
    for (int i = lo; i < hi; ++i) {
        if (fwd[i]) { inv[run] = (uint16_t)i; fwd[i] = (uint16_t)run; ++run; }
        else fwd[i] = 0xFFFF;
    }
    __syncthreads();
    return s_scan[BLOCK];
}

template <int BLOCK>
__device__ int excl_scan_i32(int32_t* a, int len, int* s_scan)
{
    const int tid = threadIdx.x;
    const int per = (len + BLOCK - 1) / BLOCK;
    const int lo = min(tid * per, len), hi = min(lo + per, len);
    int cnt = 0;
    for (int i = lo; i < hi; ++i) cnt += a[i];
    __syncthreads();
    s_scan[tid] = cnt;
    __syncthreads();
    if (tid == 0) {
        int run = 0;
        for (int k = 0; k < BLOCK; ++k) { int v = s_scan[k]; s_scan[k] = run; run += v; }
        s_scan[BLOCK] = run;
    }
    __syncthreads();
    int run = s_scan[tid];
    for (int i = lo; i < hi; ++i) { int v = a[i]; a[i] = run; run += v; }
    __syncthreads();
    return s_scan[BLOCK];
}

__global__ void __launch_bounds__(PACK_BLOCK)
k_pack(const __grid_constant__ ImgSet imgs, Par par, const sdsm_cand* __restrict__ cands, int n_cand, const int32_t* __restrict__ footprints, WS ws)
{
    constexpr int BLOCK = PACK_BLOCK;
    __shared__ int32_t s_fp[256];
    __shared__ int s_scan[BLOCK + 1];
    __shared__ int s_red[BLOCK / 32 + 1];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int sub = par.sub;

    for (int c = blockIdx.x; c < n_cand; c += gridDim.x) {
        const sdsm_cand cd = cands[c];
        const Img img = imgs.v[cd.img];
        const int h = cd.h, w = cd.w;
        const int32_t* fp = footprints + cd.fp_off;
        const int nfp = cd.fp_len;
        __syncthreads();
        if (nfp <= 256) {
            for (int k = tid; k < nfp; k += BLOCK) s_fp[k] = fp[k];
            fp = s_fp;
        }
        uint16_t* fwd_r = ws.fwd + cd.tab_off;
        uint16_t* fwd_c = fwd_r + h;
        uint16_t* inv_r = ws.inv + cd.tab_off;
        uint16_t* inv_c = inv_r + h;
        for (int i = tid; i < h + w; i += BLOCK) fwd_r[i] = 0;
        __syncthreads();

        // pass 1: occupancy of rows / columns, pixel and foreground counts
        int cnt = 0, cnt_fg = 0;
        for (int idx = tid; idx < h * w; idx += BLOCK) {
            const int r = idx / w, q = idx - r * w;
            const size_t g = (size_t)(cd.r0 + r) * img.W + (cd.c0 + q);
            if (fp_member(img.atoms[g], fp, nfp)) {
                fwd_r[r] = 1;
                fwd_c[q] = 1;
                ++cnt;
                if (img.y[g] > 0) ++cnt_fg;
            }
        }
        const int n_found = block_sum_int<BLOCK>(cnt, s_red);
        const int n_fg = block_sum_int<BLOCK>(cnt_fg, s_red);
        const int hp = compact_axis<BLOCK>(fwd_r, inv_r, h, s_scan);
        const int wp = compact_axis<BLOCK>(fwd_c, inv_c, w, s_scan);

        // normalised coordinates (image.py:19-21: division by shape - 1, zero guarded)
        double* xr = ws.xtab + cd.tab_off;
        double* xc = xr + h;
        const double zr = img.H > 1 ? (double)(img.H - 1) : 1.0, zc = img.W > 1 ? (double)(img.W - 1) : 1.0;
        for (int i = tid; i < hp; i += BLOCK) xr[i] = (double)(cd.r0 + inv_r[i]) / zr;
        for (int i = tid; i < wp; i += BLOCK) xc[i] = (double)(cd.c0 + inv_c[i]) / zc;

        const int ch = (hp + sub - 1) / sub, cw = (wp + sub - 1) / sub;
        const int ncells = ch * cw;
        int32_t* cstart = ws.cell_start + cd.cell_off;
        for (int i = tid; i <= ncells; i += BLOCK) cstart[i] = 0;
        if (tid == 0) {
            CInfo ci;
            ci.hp = hp; ci.wp = wp; ci.ch = ch; ci.cw = cw; ci.ncells = ncells; ci.m = 0; ci.n_fg = n_fg; ci.n_ne = 0; ci.pad = 0;
            ci.flags = (n_found != cd.n || n_found == 0) ? 1 : 0;
            ws.cinfo[c] = ci;
        }
        __syncthreads();
        if (n_found != cd.n) continue;   // host / device disagree on the region: flagged, never solved

        // pass 2: pixels per cell
        for (int idx = tid; idx < h * w; idx += BLOCK) {
            const int r = idx / w, q = idx - r * w;
            const size_t g = (size_t)(cd.r0 + r) * img.W + (cd.c0 + q);
            if (fp_member(img.atoms[g], fp, nfp)) {
                const int cell = (fwd_r[r] / sub) * cw + (fwd_c[q] / sub);
                atomicAdd(&cstart[cell], 1);
            }
        }
        __syncthreads();
        // the non-empty cells: the work items of the evaluation tasks
        {
            int32_t* pos = (int32_t*)(ws.grid_tmp + cd.cell_off);          // free until k_grid
            int32_t* clist = ws.cell_list + cd.cell_off;
            for (int i = tid; i <= ncells; i += BLOCK) pos[i] = (i < ncells && cstart[i] > 0) ? 1 : 0;
            __syncthreads();
            const int n_ne = excl_scan_i32<BLOCK>(pos, ncells + 1, s_scan);
            for (int i = tid; i < ncells; i += BLOCK) if (cstart[i] > 0) clist[pos[i]] = i;
            if (tid == 0) ws.cinfo[c].n_ne = n_ne;
            __syncthreads();
            // ... ordered by their number of 32-pixel chunks, largest first (stable): the cells one evaluation task hands
            // to the warps of a CTA then take about the same time, and the partly filled cells of the region's rim end up
            // together in the last tasks
            for (int i = tid; i < n_ne; i += BLOCK) {
                const int key = (cstart[clist[i]] + 31) >> 5;
                int rank = 0;
                for (int j = 0; j < n_ne; ++j) { const int kj = (cstart[clist[j]] + 31) >> 5; rank += (kj > key || (kj == key && j < i)) ? 1 : 0; }
                pos[rank] = clist[i];
            }
            __syncthreads();
            for (int i = tid; i < n_ne; i += BLOCK) clist[i] = pos[i];
        }
        __syncthreads();
        excl_scan_i32<BLOCK>(cstart, ncells + 1, s_scan);

        // pass 3: fill, one warp per cell, row-major inside the cell
        double* py = ws.pix_y + cd.pix_off;
        uint32_t* prc = ws.pix_rc + cd.pix_off;
        for (int cell = warp; cell < ncells; cell += BLOCK / 32) {
            const int ca = cell / cw, cb = cell - ca * cw;
            int base = cstart[cell];
            if (cstart[cell + 1] == base) continue;
            for (int q0 = 0; q0 < sub * sub; q0 += 32) {
                const int q = q0 + lane;
                const int u = q / sub, v = q - u * sub;
                const int rp = ca * sub + u, cp = cb * sub + v;
                bool in = (q < sub * sub) && rp < hp && cp < wp;
                double yv = 0;
                if (in) {
                    const size_t g = (size_t)(cd.r0 + inv_r[rp]) * img.W + (cd.c0 + inv_c[cp]);
                    in = fp_member(img.atoms[g], fp, nfp);
                    if (in) yv = img.y[g];
                }
                const unsigned bal = __ballot_sync(0xffffffffu, in);
                if (in) {
                    const int pos = base + __popc(bal & ((1u << lane) - 1u));
                    py[pos] = yv;
                    prc[pos] = (uint32_t)rp | ((uint32_t)cp << 16);
                }
                base += __popc(bal);
            }
        }
    }
}

}  // namespace sdsm
