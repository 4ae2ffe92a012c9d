// C ABI of the superdsm_b200 library (see include/superdsm_b200.h).  Host side: argument checks, workspace layout,
// kernel launches on the caller's stream.  No per-batch allocation, no synchronisation.
#include <cstdio>
#include <cstring>
#include <new>
#include <string>

#include "common.cuh"
#include "pack.cuh"
#include "grid.cuh"
#include "eval.cuh"
#include "ctrl.cuh"

using namespace sdsm;

struct sdsm_ctx {
    int device = 0;
    int sm_count = 148;
    Img img{nullptr, nullptr, 0, 0};
    Par par{};
    bool have_par = false;
    float* d_psf = nullptr;
    size_t psf_cap = 0;
    int64_t launches = 0;
    // totals of the last plan (the layout is a pure function of these)
    int64_t P = 0, C = 0, T = 0, V = 0, M = 0;
    int32_t n_cand = 0, n_big = 0;
    unsigned long long* h_flag = nullptr;   // pinned: size of the next work list, published by the device every few rounds
    int64_t rounds = 0;                      // rounds of the last solve (diagnostics)
    std::string err;
};

#define CUDA_TRY(ctx, expr)                                                                      \
    do {                                                                                         \
        cudaError_t e_ = (expr);                                                                 \
        if (e_ != cudaSuccess) {                                                                 \
            (ctx)->err = std::string(#expr) + ": " + cudaGetErrorString(e_);                     \
            return SDSM_E_CUDA;                                                                  \
        }                                                                                        \
    } while (0)

extern "C" int sdsm_abi_version(void) { return SDSM_ABI_VERSION; }

extern "C" int sdsm_create(sdsm_ctx** out, int device)
{
    if (!out) return SDSM_E_ARG;
    sdsm_ctx* ctx = new (std::nothrow) sdsm_ctx();
    if (!ctx) return SDSM_E_ARG;
    ctx->device = device;
    *out = ctx;
    CUDA_TRY(ctx, cudaSetDevice(device));
    cudaDeviceProp prop;
    CUDA_TRY(ctx, cudaGetDeviceProperties(&prop, device));
    ctx->sm_count = prop.multiProcessorCount;
    if (prop.major < 10) { ctx->err = "superdsm_b200 is built for sm_100a (B200) only"; return SDSM_E_UNSUPPORTED; }
    const int ctrl_max = (int)ctrl_smem_bytes(SMEM_L_CAP);
    CUDA_TRY(ctx, cudaFuncSetAttribute(k_eval, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)EVAL_SMEM));
    CUDA_TRY(ctx, cudaFuncSetAttribute(k_ctrl, cudaFuncAttributeMaxDynamicSharedMemorySize, ctrl_max));
    CUDA_TRY(ctx, cudaFuncSetAttribute(k_ctrl_big, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BIG_SMEM));
    CUDA_TRY(ctx, cudaFuncSetAttribute(k_diag_eval_begin, cudaFuncAttributeMaxDynamicSharedMemorySize, ctrl_max));
    CUDA_TRY(ctx, cudaFuncSetAttribute(k_diag_eval_end, cudaFuncAttributeMaxDynamicSharedMemorySize, ctrl_max));
    CUDA_TRY(ctx, cudaFuncSetAttribute(k_diag_trace_begin, cudaFuncAttributeMaxDynamicSharedMemorySize, ctrl_max));
    CUDA_TRY(ctx, cudaHostAlloc((void**)&ctx->h_flag, 64, cudaHostAllocDefault));
    return SDSM_OK;
}

extern "C" void sdsm_destroy(sdsm_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->d_psf) cudaFree(ctx->d_psf);
    if (ctx->h_flag) cudaFreeHost(ctx->h_flag);
    delete ctx;
}

extern "C" const char* sdsm_last_error(const sdsm_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

extern "C" int64_t sdsm_launch_count(const sdsm_ctx* ctx) { return ctx ? ctx->launches : 0; }
extern "C" int64_t sdsm_round_count(const sdsm_ctx* ctx) { return ctx ? ctx->rounds : 0; }

extern "C" int sdsm_set_image(sdsm_ctx* ctx, const double* d_y, const int32_t* d_atoms, int32_t H, int32_t W)
{
    if (!ctx || !d_y || !d_atoms || H < 2 || W < 2) return SDSM_E_ARG;
    if (H > 65535 || W > 65535) { ctx->err = "image larger than 65535 pixels in a dimension"; return SDSM_E_UNSUPPORTED; }
    ctx->img = Img{d_y, d_atoms, H, W};
    return SDSM_OK;
}

extern "C" int sdsm_set_params(sdsm_ctx* ctx, const sdsm_params* p)
{
    if (!ctx || !p) return SDSM_E_ARG;
    if (!(p->epsilon > 0) || !(p->alpha >= 0) || !(p->scale > 0) || p->subsample < 1 || !(p->sparsity_tol >= 0)) {
        ctx->err = "bad hyper-parameters (epsilon > 0, alpha >= 0, scale > 0, smooth_subsample >= 1, sparsity_tol >= 0)";
        return SDSM_E_ARG;
    }
    Par par{};
    par.alpha = p->alpha; par.epsilon = p->epsilon; par.scale = p->scale; par.sparsity_tol = p->sparsity_tol;
    par.abstol = p->abstol; par.reltol = p->reltol; par.feastol = p->feastol;
    par.sub = p->subsample; par.R = p->radius; par.has_smooth = p->has_smooth; par.init_elliptical = p->init_elliptical;
    par.maxiters = p->maxiters;
    par.psf = nullptr;
    if (p->has_smooth) {
        if (!p->h_psf || p->radius < 1) return SDSM_E_ARG;
        // a pixel only searches the 5 x 5 cells around its own for grid points: needs ceil(R / subsample) <= 2;
        // subsample <= R + 1 is what keeps every row of G~ non-empty in the reference (dsm.py:194)
        if ((p->radius + p->subsample - 1) / p->subsample > NC) {
            ctx->err = "kernel radius larger than 2 * smooth_subsample is not implemented";
            return SDSM_E_UNSUPPORTED;
        }
        if (p->radius > 255) { ctx->err = "kernel radius > 255 is not implemented"; return SDSM_E_UNSUPPORTED; }
        const size_t bytes = sizeof(float) * (size_t)(p->radius + 1) * (p->radius + 1);
        CUDA_TRY(ctx, cudaSetDevice(ctx->device));
        if (bytes > ctx->psf_cap) {
            if (ctx->d_psf) CUDA_TRY(ctx, cudaFree(ctx->d_psf));
            ctx->d_psf = nullptr; ctx->psf_cap = 0;
            CUDA_TRY(ctx, cudaMalloc(&ctx->d_psf, bytes));
            ctx->psf_cap = bytes;
        }
        CUDA_TRY(ctx, cudaMemcpy(ctx->d_psf, p->h_psf, bytes, cudaMemcpyHostToDevice));
        par.psf = ctx->d_psf;
    } else {
        // the cells only organise the pixels when there is no smooth matrix
        if (par.sub < 4) par.sub = 16;
    }
    ctx->par = par;
    ctx->have_par = true;
    return SDSM_OK;
}


extern "C" int sdsm_plan_batch(sdsm_ctx* ctx, sdsm_cand* cands, int32_t n_cand, sdsm_plan* plan)
{
    if (!ctx || !cands || !plan || n_cand < 1) return SDSM_E_ARG;
    if (!ctx->have_par || !ctx->img.y) { ctx->err = "sdsm_set_image / sdsm_set_params must precede sdsm_plan_batch"; return SDSM_E_ARG; }
    const int sub = ctx->par.sub;
    int64_t P = 0, C = 0, T = 0, V = 0, M = 0, PAR = 0, FG = 0;
    int32_t n_max = 6, cells_max = 1, n_big = 0;
    for (int32_t i = 0; i < n_cand; ++i) {
        sdsm_cand& c = cands[i];
        if (c.h < 1 || c.w < 1 || c.n < 1 || c.r0 < 0 || c.c0 < 0 || c.r0 + c.h > ctx->img.H || c.c0 + c.w > ctx->img.W || c.fp_len < 1) {
            ctx->err = "candidate " + std::to_string(i) + ": empty region or bounding box outside the image";
            return SDSM_E_ARG;
        }
        const int32_t cap = ((c.h + sub - 1) / sub) * ((c.w + sub - 1) / sub);
        const int64_t ld = 6 + (ctx->par.has_smooth ? cap : 0);
        c.cell_cap = cap;
        c.pix_off = P;  P += (c.n + 31) / 32 * 32;
        c.cell_off = C; C += cap + 1;
        c.tab_off = T;  T += (c.h + c.w + 3) / 4 * 4;
        c.par_off = PAR; PAR += ld;
        c.fg_off = FG;  FG += (int64_t)c.h * c.w;
        c.vec_off = V;  V += (int64_t)NVEC * ld;
        c.mat_off = M;  M += 3 * ld * ld;
        if (ld > SMALL_N_MAX) ++n_big;
        if (ld > n_max) n_max = (int32_t)ld;
        if (cap > cells_max) cells_max = cap;
    }
    const Layout L = make_layout(P, C, T, V, M, n_cand);
    plan->workspace_bytes = L.total;
    plan->params_len = PAR;
    plan->fgmask_len = FG;
    plan->n_max = n_max;
    plan->cells_max = cells_max;
    plan->reserved0 = n_big;
    plan->reserved1 = 0;
    ctx->n_big = n_big;
    ctx->P = P; ctx->C = C; ctx->T = T; ctx->V = V; ctx->M = M; ctx->n_cand = n_cand;
    return SDSM_OK;
}

static int check_plan(sdsm_ctx* ctx, int32_t n_cand, size_t workspace_bytes, Layout* L)
{
    if (n_cand != ctx->n_cand) { ctx->err = "batch does not match the last sdsm_plan_batch"; return SDSM_E_ARG; }
    *L = make_layout(ctx->P, ctx->C, ctx->T, ctx->V, ctx->M, n_cand);
    if (L->total > workspace_bytes) { ctx->err = "workspace too small"; return SDSM_E_WORKSPACE; }
    return SDSM_OK;
}

// Drives the lock-step rounds k_eval -> k_hgather -> k_ctrl / k_ctrl_big until no candidate is live.  `parity` names the
// work list that is ready.  The sizes of the next work lists are published to pinned memory every few rounds; the host
// only waits on that.
struct Bounds { int64_t live_small, live_big, items, hitems; };

static int run_rounds(sdsm_ctx* ctx, cudaStream_t stream, const sdsm_cand* d_cands, const WS& ws, const sdsm_out& out, int lcap,
                      Bounds bd, int parity, double* trace, int unit_scale)
{
    const size_t csmem = ctrl_smem_bytes(lcap);
    const int max_rounds = 64 * (ctx->par.maxiters + 2) * 3;
    const int64_t g_max = (int64_t)ctx->sm_count * 16;
    int round = 0;
    int burst = 4;
    while (true) {
        for (int b = 0; b < burst; ++b, ++round) {
            int64_t ge = (bd.items + EVAL_WARPS - 1) / EVAL_WARPS;
            if (ge > g_max) ge = g_max;
            if (ge < 1) ge = 1;
            k_eval<<<(unsigned)ge, EVAL_BLOCK, EVAL_SMEM, stream>>>(ctx->par, d_cands, ws, parity);
            ctx->launches++;
            if (bd.hitems > 0) {
                int64_t gh = (bd.hitems + HG_BLOCK / 32 - 1) / (HG_BLOCK / 32);
                if (gh > g_max) gh = g_max;
                k_hgather<<<(unsigned)gh, HG_BLOCK, 0, stream>>>(ctx->par, d_cands, ws, parity, unit_scale);
                ctx->launches++;
            }
            if (bd.live_big > 0) {
                k_ctrl_big<<<(unsigned)bd.live_big, BIG_BLOCK, BIG_SMEM, stream>>>(ctx->img, ctx->par, d_cands, ws, out, parity, trace);
                ctx->launches++;
            }
            if (bd.live_small > 0) {
                k_ctrl<<<(unsigned)bd.live_small, CTRL_BLOCK, csmem, stream>>>(ctx->img, ctx->par, d_cands, nullptr, 0, ws, out, parity, lcap, trace);
                ctx->launches++;
            }
            parity ^= 1;
        }
        k_publish<<<1, 1, 0, stream>>>(ws, parity, ctx->h_flag);
        ctx->launches++;
        CUDA_TRY(ctx, cudaStreamSynchronize(stream));
        const unsigned long long f0 = ((volatile unsigned long long*)ctx->h_flag)[0], f2 = ((volatile unsigned long long*)ctx->h_flag)[1];
        Bounds nb{(int64_t)(f0 >> 32), (int64_t)(f2 >> 32), (int64_t)(f0 & 0xffffffffull), (int64_t)(f2 & 0xffffffffull)};
        if (nb.live_small + nb.live_big == 0) break;
        // the live set only shrinks; a candidate can change class (small -> big) once, when its DSM stage begins
        if (nb.live_small > bd.live_small || nb.items > bd.items) { ctx->err = "work list grew: internal error"; return SDSM_E_CUDA; }
        bd.live_big = nb.live_small + nb.live_big < bd.live_big ? nb.live_small + nb.live_big : bd.live_big;
        bd.live_small = nb.live_small;
        bd.items = nb.items;
        if (round > max_rounds) { ctx->err = "solver did not terminate within the round limit"; return SDSM_E_CUDA; }
        if (burst < 8) burst *= 2;
    }
    ctx->rounds = round;
    return SDSM_OK;
}

extern "C" int sdsm_solve_batch(sdsm_ctx* ctx, const sdsm_plan* plan, const sdsm_cand* d_cands, int32_t n_cand,
                                const int32_t* d_footprints, const int32_t* d_order, void* d_workspace, size_t workspace_bytes,
                                const sdsm_out* out, void* stream_)
{
    if (!ctx || !plan || !d_cands || !d_footprints || !d_order || !d_workspace || !out) return SDSM_E_ARG;
    if (!out->d_energy || !out->d_status || !out->d_info || !out->d_params || !out->d_fgbox || !out->d_onboundary || !out->d_fgmask) return SDSM_E_ARG;
    Layout L;
    int rc = check_plan(ctx, n_cand, workspace_bytes, &L);
    if (rc) return rc;
    cudaStream_t stream = (cudaStream_t)stream_;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    WS ws = make_ws(d_workspace, L);
    CUDA_TRY(ctx, cudaMemsetAsync(ws.counters, 0, sizeof(unsigned long long) * 16, stream));
    CUDA_TRY(ctx, cudaMemsetAsync(ws.cstate, 0, sizeof(CState) * (size_t)n_cand, stream));
    CUDA_TRY(ctx, cudaMemsetAsync(out->d_fgmask, 0, (size_t)plan->fgmask_len, stream));

    const int pack_grid = n_cand < 8 * ctx->sm_count ? n_cand : 8 * ctx->sm_count;
    k_pack<<<pack_grid, PACK_BLOCK, 0, stream>>>(ctx->img, ctx->par, d_cands, n_cand, d_footprints, ws);
    ctx->launches++;
    if (ctx->par.has_smooth) {
        const size_t psf_bytes = sizeof(float) * (size_t)(ctx->par.R + 1) * (ctx->par.R + 1);
        if (psf_bytes > 48 * 1024)
            CUDA_TRY(ctx, cudaFuncSetAttribute(k_grid, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)psf_bytes));
        k_grid<<<pack_grid, GRID_BLOCK, psf_bytes, stream>>>(ctx->par, d_cands, n_cand, d_order, ws);
        ctx->launches++;
    }
    const int lcap = ctrl_lcap(plan->n_max);
    // round 0: every candidate starts (work list 0 receives the first requests)
    k_ctrl<<<n_cand, CTRL_BLOCK, ctrl_smem_bytes(lcap), stream>>>(ctx->img, ctx->par, d_cands, d_order, n_cand, ws, *out, 1, lcap, nullptr);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    {
        const Bounds bd{n_cand, ctx->n_big, ctx->C, ctx->par.has_smooth ? ctx->C : 0};
        rc = run_rounds(ctx, stream, d_cands, ws, *out, lcap, bd, 0, nullptr, 0);
    }
    if (rc) return rc;
    const int fin_grid = n_cand < 16 * ctx->sm_count ? n_cand : 16 * ctx->sm_count;
    k_finish<<<fin_grid, FIN_BLOCK, 0, stream>>>(ctx->img, ctx->par, d_cands, n_cand, ws, *out);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    return SDSM_OK;
}

extern "C" int sdsm_eval_candidate(sdsm_ctx* ctx, const sdsm_plan* plan, const sdsm_cand* d_cands, int32_t c, const double* d_x,
                                   double* d_value, double* d_grad, double* d_H, void* d_workspace, void* stream_)
{
    if (!ctx || !plan || !d_cands || !d_x || !d_value || !d_grad || !d_H || !d_workspace || c < 0 || c >= ctx->n_cand) return SDSM_E_ARG;
    const Layout L = make_layout(ctx->P, ctx->C, ctx->T, ctx->V, ctx->M, ctx->n_cand);
    WS ws = make_ws(d_workspace, L);
    cudaStream_t stream = (cudaStream_t)stream_;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const int lcap = ctrl_lcap(plan->n_max);
    const size_t csmem = ctrl_smem_bytes(lcap);
    CUDA_TRY(ctx, cudaMemsetAsync(ws.counters, 0, sizeof(unsigned long long) * 16, stream));
    k_diag_eval_begin<<<1, CTRL_BLOCK, csmem, stream>>>(ctx->par, d_cands, c, ws, d_x, 1, lcap);
    k_eval<<<ctx->sm_count, EVAL_BLOCK, EVAL_SMEM, stream>>>(ctx->par, d_cands, ws, 0);
    if (ctx->par.has_smooth) { k_hgather<<<ctx->sm_count, HG_BLOCK, 0, stream>>>(ctx->par, d_cands, ws, 0, 1); ctx->launches++; }
    k_diag_eval_end<<<1, CTRL_BLOCK, csmem, stream>>>(ctx->par, d_cands, c, ws, d_value, d_grad, d_H, lcap);
    ctx->launches += 3;
    CUDA_TRY(ctx, cudaGetLastError());
    return SDSM_OK;
}

extern "C" int sdsm_export_smooth(sdsm_ctx* ctx, const sdsm_plan* plan, const sdsm_cand* d_cands, int32_t c, uint32_t* d_rc,
                                  int32_t* d_cols, float* d_vals, uint32_t* d_grid, void* d_workspace, void* stream_)
{
    if (!ctx || !plan || !d_cands || !d_rc || !d_cols || !d_vals || !d_grid || !d_workspace || c < 0 || c >= ctx->n_cand) return SDSM_E_ARG;
    const Layout L = make_layout(ctx->P, ctx->C, ctx->T, ctx->V, ctx->M, ctx->n_cand);
    WS ws = make_ws(d_workspace, L);
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    k_export<<<64, 128, 0, (cudaStream_t)stream_>>>(ctx->par, d_cands, c, ws, d_rc, d_cols, d_vals, d_grid);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    return SDSM_OK;
}

extern "C" int sdsm_trace_solve(sdsm_ctx* ctx, const sdsm_plan* plan, const sdsm_cand* d_cands, int32_t c, int32_t use_xi,
                                const double* d_x0, double* d_trace, double* d_xout, int32_t* d_status_iters, void* d_workspace, void* stream_)
{
    if (!ctx || !plan || !d_cands || !d_x0 || !d_trace || !d_xout || !d_status_iters || !d_workspace || c < 0 || c >= ctx->n_cand) return SDSM_E_ARG;
    const Layout L = make_layout(ctx->P, ctx->C, ctx->T, ctx->V, ctx->M, ctx->n_cand);
    WS ws = make_ws(d_workspace, L);
    cudaStream_t stream = (cudaStream_t)stream_;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const int lcap = ctrl_lcap(plan->n_max);
    CUDA_TRY(ctx, cudaMemsetAsync(ws.counters, 0, sizeof(unsigned long long) * 16, stream));
    k_diag_trace_begin<<<1, CTRL_BLOCK, ctrl_smem_bytes(lcap), stream>>>(ctx->par, d_cands, c, use_xi, ws, d_x0, 1, lcap);
    ctx->launches++;
    sdsm_out none{};   // STG_TRACE never writes results
    const Bounds bd{1, 1, ctx->C, ctx->par.has_smooth ? ctx->C : 0};
    int rc = run_rounds(ctx, stream, d_cands, ws, none, lcap, bd, 0, d_trace, 0);
    if (rc) return rc;
    k_diag_trace_end<<<1, 128, 0, stream>>>(ctx->par, d_cands, c, ws, d_xout, d_status_iters);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    return SDSM_OK;
}

extern "C" int sdsm_dfma_peak(sdsm_ctx* ctx, double* d_scratch, int32_t iters, int64_t* flops, void* stream_)
{
    if (!ctx || !d_scratch || !flops || iters < 1) return SDSM_E_ARG;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const int blocks = ctx->sm_count * 8;
    k_dfma_peak<<<blocks, 256, 0, (cudaStream_t)stream_>>>(d_scratch, iters);
    ctx->launches++;
    *flops = (int64_t)blocks * 256 * 8 * 2 * (int64_t)iters;
    CUDA_TRY(ctx, cudaGetLastError());
    return SDSM_OK;
}
