// C ABI of the superdsm_b200 library (see include/superdsm_b200.h).  Host side: argument checks, workspace layout,
// kernel launches on the caller's stream.  No per-batch allocation, no synchronisation.
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>

#include "common.cuh"
#include "image.cuh"
#include "preprocess.cuh"
#include "pack.cuh"
#include "grid.cuh"
#include "eval.cuh"
#include "ctrl.cuh"
#include "sched.cuh"

using namespace sdsm;

struct sdsm_ctx {
    int device = 0;
    int sm_count = 148;
    ImgSet imgs{};
    int n_images = 0;
    Par par{};
    bool have_par = false;
    float* d_psf = nullptr;
    size_t psf_cap = 0;
    double* d_w = nullptr;                   // Gaussian weights of sdsm_preprocess
    size_t w_cap = 0;
    int64_t launches = 0;
    int solve_ctas_per_sm = 1;               // resident CTAs of k_solve per SM (occupancy query)
    int solo_cells = 0;                      // SDSM_SOLO_CELLS: chains of candidates up to this many cells stay in one CTA
    int task_small_cells = 512;              // SDSM_TASK_SMALL_CELLS: candidates up to this many cells get one cell per warp
    double timeout_s = 60.0;                 // SDSM_TIMEOUT_S: watchdog of k_solve
    double hi_frac = 0.02;                   // SDSM_HI_FRAC: this share of the batch (the most expensive candidates) is urgent from the start
    bool ctrl_ctas_fixed = false;            // SDSM_CTRL_CTAS given: no adaptation to the batch
    int ctrl_ctas = 64;                      // SDSM_CTRL_CTAS: CTAs of k_solve that run control tasks (the warps of the others evaluate)
    int hi_evals = 300;                      // SDSM_HI_EVALS: candidates that needed more evaluations than this become urgent
    int spec_max = SPEC_MAX;                 // SDSM_SPEC_MAX: trial points per line-search request (1: no speculative backtracking)
    int spec_retry = 4;                      // SDSM_SPEC_RETRY: points of the request after a first request of one point was rejected
    int spec_skip = 1;                       // SDSM_SPEC_SKIP: 1 = first-pass line searches whose decisions do not depend on F make no evaluations
    int spec_hess = 1;                       // SDSM_SPEC_HESS: 1 = a trial that will probably be accepted is evaluated with its Hessian
    int spec_prio = 1;                       // SDSM_SPEC_PRIO: 1 = of the chains that go through the queue only the urgent ones speculate
    int spec_solo = 1;                       // SDSM_SPEC_SOLO: 1 = chains inside one control CTA speculate too (their points are evaluated in that CTA)
    double ctrl_share = 0.95;                // SDSM_CTRL_SHARE: share of the remaining grid the control CTAs get per unit of solo work
    std::string err;
};

// candidates with at most this many pixel cells run their chain inside one control CTA: about 5000 pixels' worth of cells
static int solo_cells_of(const sdsm_ctx* ctx)
{
    if (ctx->solo_cells != 0) return ctx->solo_cells < 0 ? 0 : ctx->solo_cells;
    const int v = 512 / (ctx->par.sub > 0 ? ctx->par.sub : 16);
    return v < 8 ? 8 : (v > 128 ? 128 : v);
}

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
    const int smem_max = (int)solve_smem_bytes(SMEM_L_CAP);
    CUDA_TRY(ctx, cudaFuncSetAttribute(k_solve, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
    CUDA_TRY(ctx, cudaFuncSetAttribute(k_sweep, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
    CUDA_TRY(ctx, cudaFuncSetAttribute(k_diag_eval_cells, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
    CUDA_TRY(ctx, cudaFuncSetAttribute(k_diag_eval_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
    CUDA_TRY(ctx, cudaFuncSetAttribute(k_diag_eval_end, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
    CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx->solve_ctas_per_sm, k_solve, SOLVE_BLOCK, (size_t)smem_max));
    if (ctx->solve_ctas_per_sm < 1) { ctx->err = "k_solve does not fit on an SM"; return SDSM_E_UNSUPPORTED; }
    if (const char* e = getenv("SDSM_SOLO_CELLS")) ctx->solo_cells = atoi(e) > 0 ? atoi(e) : -1;     // 0 (default): 512 / smooth_subsample
    if (const char* e = getenv("SDSM_TASK_SMALL_CELLS")) ctx->task_small_cells = atoi(e);
    if (const char* e = getenv("SDSM_TIMEOUT_S")) ctx->timeout_s = atof(e);
    if (const char* e = getenv("SDSM_HI_FRAC")) ctx->hi_frac = atof(e);
    if (const char* e = getenv("SDSM_HI_EVALS")) ctx->hi_evals = atoi(e);
    if (const char* e = getenv("SDSM_SPEC_MAX")) { ctx->spec_max = atoi(e); if (ctx->spec_max < 1) ctx->spec_max = 1; if (ctx->spec_max > SPEC_MAX) ctx->spec_max = SPEC_MAX; }
    if (const char* e = getenv("SDSM_SPEC_RETRY")) { ctx->spec_retry = atoi(e); if (ctx->spec_retry < 1) ctx->spec_retry = 1; }
    if (const char* e = getenv("SDSM_SPEC_SOLO")) ctx->spec_solo = atoi(e) != 0;
    if (const char* e = getenv("SDSM_SPEC_SKIP")) ctx->spec_skip = atoi(e) != 0;
    if (const char* e = getenv("SDSM_SPEC_HESS")) ctx->spec_hess = atoi(e) != 0;
    if (const char* e = getenv("SDSM_SPEC_PRIO")) ctx->spec_prio = atoi(e) != 0;
    if (const char* e = getenv("SDSM_CTRL_SHARE")) ctx->ctrl_share = atof(e);
    if (const char* e = getenv("SDSM_CTRL_CTAS")) { ctx->ctrl_ctas = atoi(e); ctx->ctrl_ctas_fixed = true; }
    return SDSM_OK;
}

extern "C" void sdsm_destroy(sdsm_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->d_psf) cudaFree(ctx->d_psf);
    if (ctx->d_w) cudaFree(ctx->d_w);
    delete ctx;
}

extern "C" const char* sdsm_last_error(const sdsm_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

extern "C" int64_t sdsm_launch_count(const sdsm_ctx* ctx) { return ctx ? ctx->launches : 0; }

extern "C" int sdsm_set_images(sdsm_ctx* ctx, const sdsm_image* h_images, int32_t n_images)
{
    if (!ctx || !h_images || n_images < 1) return SDSM_E_ARG;
    if (n_images > SDSM_MAX_IMAGES) { ctx->err = "more than SDSM_MAX_IMAGES images in one set"; return SDSM_E_UNSUPPORTED; }
    for (int32_t i = 0; i < n_images; ++i) {
        const sdsm_image& im = h_images[i];
        if (!im.d_y || !im.d_atoms || im.H < 2 || im.W < 2) return SDSM_E_ARG;
        if (im.H > 65535 || im.W > 65535) { ctx->err = "image larger than 65535 pixels in a dimension"; return SDSM_E_UNSUPPORTED; }
    }
    for (int32_t i = 0; i < n_images; ++i) ctx->imgs.v[i] = Img{h_images[i].d_y, h_images[i].d_atoms, h_images[i].H, h_images[i].W};
    ctx->n_images = n_images;
    return SDSM_OK;
}

extern "C" int sdsm_set_image(sdsm_ctx* ctx, const double* d_y, const int32_t* d_atoms, int32_t H, int32_t W)
{
    const sdsm_image im{d_y, d_atoms, H, W};
    return sdsm_set_images(ctx, &im, 1);
}

extern "C" int sdsm_prepare_image(sdsm_ctx* ctx, const double* d_y, const uint8_t* d_ymask, const int32_t* d_atoms, int32_t H, int32_t W,
                                  double background_margin, int32_t n_labels, int32_t* d_atoms_out, uint64_t* d_count, int32_t* d_box,
                                  uint8_t* d_scratch, void* stream_)
{
    if (!ctx || !d_y || !d_ymask || !d_atoms || !d_atoms_out || !d_count || !d_box || !d_scratch || H < 2 || W < 2 || n_labels < 0) return SDSM_E_ARG;
    if (background_margin != background_margin || background_margin < 0) { ctx->err = "background_margin must be >= 0"; return SDSM_E_ARG; }
    const bool unbounded = background_margin > 1e300;
    // a margin beyond the image's extent reaches every pixel that has any foreground pixel at all: the window is the image
    const int Mi = unbounded ? 0 : (int)(background_margin < (double)(H + W) ? background_margin : (double)(H + W));            // floor
    if (Mi > EDT_MAX_MARGIN) { ctx->err = "background_margin > 65534 on an image that large is not implemented"; return SDSM_E_UNSUPPORTED; }
    cudaStream_t stream = (cudaStream_t)stream_;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const dim3 block(128), grid((W + 127) / 128, H);
    k_edt_init<<<(n_labels + 256) / 256, 256, 0, stream>>>(n_labels, H, W, (unsigned long long*)d_count, d_box);
    if (!unbounded) { k_edt_cols<<<grid, block, 0, stream>>>(d_y, H, W, Mi, (uint16_t*)d_scratch, 0.0); ctx->launches++; }
    k_edt_rows<<<grid, block, 0, stream>>>((const uint16_t*)d_scratch, d_ymask, d_atoms, H, W, Mi, unbounded ? -1.0 : background_margin, n_labels, d_atoms_out,
                                           (unsigned long long*)d_count, d_box);
    ctx->launches += 2;
    CUDA_TRY(ctx, cudaGetLastError());
    return SDSM_OK;
}

// ---- sdsm_prepare_image without caller scratch and without any allocation -------------------------------------------------
// The 16-bit vertical distances of the distance transform live in a file-scope device array: every GPU that loads the library
// holds SDSM_NOALLOC_MAX_PIXELS * 2 bytes until the process exits.  `g_prepare_scratch` below is, in HOST code, only a stand-in
// for the symbol: its device address is asked for once per device (with that device current) and handed to the kernels.
static __device__ __align__(256) unsigned char g_prepare_scratch[(size_t)SDSM_NOALLOC_MAX_PIXELS * 2];

static cudaError_t prepare_scratch_address(int device, uint8_t** out)
{
    constexpr int MAX_DEVICES = 64;
    static uint8_t* addr[MAX_DEVICES] = {};              // resolved addresses, one per device
    static std::mutex lock;
    if (device < 0 || device >= MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> hold(lock);
    if (!addr[device]) {
        int current = -1;
        cudaError_t e = cudaGetDevice(&current);
        if (e != cudaSuccess) return e;
        if (current != device) return cudaErrorInvalidDevice;       // the caller made `device` current; never resolve on another one
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_prepare_scratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        addr[device] = static_cast<uint8_t*>(p);
    }
    *out = addr[device];
    return cudaSuccess;
}

extern "C" int sdsm_prepare_image_noalloc(sdsm_ctx* ctx, const double* d_y, const uint8_t* d_ymask, const int32_t* d_atoms, int32_t H, int32_t W,
                                          double background_margin, int32_t n_labels, int32_t* d_atoms_out, uint64_t* d_count, int32_t* d_box,
                                          void* stream_)
{
    if (!ctx) return SDSM_E_ARG;
    if (H < 2 || W < 2) return SDSM_E_ARG;
    if ((long long)H * (long long)W > (long long)SDSM_NOALLOC_MAX_PIXELS) {
        ctx->err = "sdsm_prepare_image_noalloc: image of " + std::to_string(H) + " x " + std::to_string(W) + " pixels exceeds the built-in scratch of " +
                   std::to_string((long long)SDSM_NOALLOC_MAX_PIXELS) + " pixels (SDSM_NOALLOC_MAX_PIXELS); use sdsm_prepare_image with caller scratch";
        return SDSM_E_UNSUPPORTED;
    }
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    uint8_t* scratch = nullptr;
    CUDA_TRY(ctx, prepare_scratch_address(ctx->device, &scratch));
    // Nothing in the scratch has to start at zero: k_edt_cols writes every one of the H * W distances before k_edt_rows reads
    // them (same stream), and a call with an infinite margin does not touch it.  The contents of the previous call are overwritten.
    return sdsm_prepare_image(ctx, d_y, d_ymask, d_atoms, H, W, background_margin, n_labels, d_atoms_out, d_count, d_box, scratch, stream_);
}

extern "C" int sdsm_preprocess(sdsm_ctx* ctx, const double* d_graw, int32_t H, int32_t W, const double* h_w1, int32_t r1, const double* h_w2,
                               int32_t r2, double sigma2, double offset_clip_abs, double lower_clip, double* d_y, double* d_scratch, void* stream_)
{
    if (!ctx || !d_graw || !h_w1 || !h_w2 || !d_y || !d_scratch || H < 1 || W < 1 || r1 < 0 || r2 < 0 || !(sigma2 > 0)) return SDSM_E_ARG;
    const bool have_clip = offset_clip_abs < 1e300;
    const int Mi = (int)sigma2 + 1;                                      // only distances < sigma2 contribute (preprocess.py:58)
    if (have_clip && Mi > EDT_MAX_MARGIN) { ctx->err = "preprocess/sigma2 > 65533 is not implemented"; return SDSM_E_UNSUPPORTED; }
    cudaStream_t stream = (cudaStream_t)stream_;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const size_t n = (size_t)H * W, wbytes = sizeof(double) * (size_t)(2 * r1 + 1 + 2 * r2 + 1);
    if (wbytes > ctx->w_cap) {
        if (ctx->d_w) CUDA_TRY(ctx, cudaFree(ctx->d_w));
        ctx->d_w = nullptr; ctx->w_cap = 0;
        CUDA_TRY(ctx, cudaMalloc(&ctx->d_w, wbytes + 64));
        ctx->w_cap = wbytes;
    }
    double* w1 = ctx->d_w; double* w2 = w1 + 2 * r1 + 1;
    CUDA_TRY(ctx, cudaMemcpyAsync(w1, h_w1, sizeof(double) * (2 * r1 + 1), cudaMemcpyHostToDevice, stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(w2, h_w2, sizeof(double) * (2 * r2 + 1), cudaMemcpyHostToDevice, stream));
    // scratch: [0] pass buffer, [1] smooth1, [2] original offset, [3] clipped offset, [4] t0 (+ 16-bit distances and the max behind it)
    double* tmp = d_scratch; double* s1 = tmp + n; double* orig = s1 + n; double* clipped = orig + n; double* t0 = clipped + n;
    uint16_t* gv = reinterpret_cast<uint16_t*>(t0 + n);
    unsigned long long* tmax = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(gv) + ((2 * n + 15) / 16) * 16);
    const dim3 block(128), grid((W + 127) / 128, H);
    const double inf = 1e308 * 10;
    k_gauss_pass<<<grid, block, 0, stream>>>(d_graw, tmp, H, W, w1, r1, 0, inf);
    k_gauss_pass<<<grid, block, 0, stream>>>(tmp, s1, H, W, w1, r1, 1, inf);
    k_gauss_pass<<<grid, block, 0, stream>>>(d_graw, tmp, H, W, w2, r2, 0, inf);
    k_gauss_pass<<<grid, block, 0, stream>>>(tmp, orig, H, W, w2, r2, 1, inf);
    ctx->launches += 4;
    if (have_clip) {
        k_gauss_pass<<<grid, block, 0, stream>>>(d_graw, tmp, H, W, w2, r2, 0, offset_clip_abs);
        k_gauss_pass<<<grid, block, 0, stream>>>(tmp, clipped, H, W, w2, r2, 1, inf);
        CUDA_TRY(ctx, cudaMemsetAsync(tmax, 0, sizeof(unsigned long long), stream));
        k_edt_cols<<<grid, block, 0, stream>>>(d_graw, H, W, Mi, gv, offset_clip_abs);
        k_blend_weight<<<grid, block, 0, stream>>>(gv, H, W, Mi, sigma2, t0, tmax);
        ctx->launches += 4;
    }
    k_offset_image<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(s1, orig, clipped, t0, tmax, have_clip ? 1 : 0, lower_clip, n, d_y);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
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
    par.hess_tol = p->hessian_sparsity_tol > 0 ? p->hessian_sparsity_tol : 0.0;
    par.sub = p->subsample; par.R = p->radius; par.has_smooth = p->has_smooth; par.init_elliptical = p->init_elliptical;
    par.maxiters = p->maxiters;
    par.skip_single_fg = p->skip_single_fg ? 1 : 0;
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
    if (!ctx->have_par || ctx->n_images < 1) { ctx->err = "sdsm_set_image(s) / sdsm_set_params must precede sdsm_plan_batch"; return SDSM_E_ARG; }
    const int sub = ctx->par.sub;
    int64_t P = 0, C = 0, T = 0, V = 0, M = 0, PAR = 0, FG = 0, K = 0;
    int32_t n_max = 6, cells_max = 1;
    double work_all = 0, work_solo = 0;
    for (int32_t i = 0; i < n_cand; ++i) {
        sdsm_cand& c = cands[i];
        if (c.img < 0 || c.img >= ctx->n_images) { ctx->err = "candidate " + std::to_string(i) + ": image index outside the image set"; return SDSM_E_ARG; }
        const Img& im = ctx->imgs.v[c.img];
        if (c.h < 1 || c.w < 1 || c.n < 1 || c.r0 < 0 || c.c0 < 0 || c.r0 + c.h > im.H || c.c0 + c.w > im.W || c.fp_len < 1) {
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
        work_all += c.n;
        // its chain will (most likely) stay in one control CTA: non-empty cells ~ 1.6 n / sub^2 (the cells of the rim are partly filled)
        if (1.6 * c.n <= (double)solo_cells_of(ctx) * sub * sub) work_solo += c.n;
        K += 1 + cap;                                     // items of the candidate that can be queued at the same time
        if (ld > n_max) n_max = (int32_t)ld;
        if (cap > cells_max) cells_max = cap;
    }
    const Layout L = make_layout(P, C, T, V, M, K, n_cand);
    plan->workspace_bytes = L.total;
    plan->params_len = PAR;
    plan->fgmask_len = FG;
    plan->n_max = n_max;
    plan->cells_max = cells_max;
    plan->n_cand = n_cand;
    plan->reserved = 0;
    plan->totals[0] = P; plan->totals[1] = C; plan->totals[2] = T; plan->totals[3] = V; plan->totals[4] = M; plan->totals[5] = K;
    plan->solo_share = work_all > 0 ? work_solo / work_all : 0.0;
    return SDSM_OK;
}

// the workspace layout of a planned batch: a pure function of the plan (no hidden context state)
static Layout plan_layout(const sdsm_plan* plan)
{
    return make_layout(plan->totals[0], plan->totals[1], plan->totals[2], plan->totals[3], plan->totals[4], plan->totals[5], plan->n_cand);
}

static int check_plan(sdsm_ctx* ctx, const sdsm_plan* plan, int32_t n_cand, size_t workspace_bytes, Layout* L)
{
    if (n_cand != plan->n_cand || plan->n_cand < 1) { ctx->err = "batch does not match its plan"; return SDSM_E_ARG; }
    for (int k = 0; k < 6; ++k) if (plan->totals[k] < 0) { ctx->err = "corrupt plan"; return SDSM_E_ARG; }
    *L = plan_layout(plan);
    if (L->total != plan->workspace_bytes) { ctx->err = "corrupt plan"; return SDSM_E_ARG; }
    if (L->total > workspace_bytes) { ctx->err = "workspace too small"; return SDSM_E_WORKSPACE; }
    return SDSM_OK;
}

// Seeds the task queue with the control tasks that start `n_first` candidates (order, or the single candidate `single`)
// and launches the persistent solver: one grid of resident CTAs that runs until every seeded candidate has finished.
static int run_solver(sdsm_ctx* ctx, cudaStream_t stream, const sdsm_cand* d_cands, const WS& ws, const sdsm_out& out, int lcap,
                      const int32_t* d_order, int n_first, int single, double* trace, double solo_share)
{
    // urgent from the start: the candidates at the head of the processing order (most expensive first); 0 for single solves
    int n_hi = d_order ? (int)(ctx->hi_frac * n_first) : 0;
    if (n_hi > n_first) n_hi = n_first;
    k_seed<<<64, 256, 0, stream>>>(ws, d_order, n_first, single, n_hi);
    SolveArgs args{};
    args.n_target = n_first;
    args.lcap = lcap;
    args.solo_cells = solo_cells_of(ctx);
    args.task_small_cells = ctx->task_small_cells;
    args.hi_evals = ctx->hi_evals;
    args.spec_max = ctx->spec_max;
    args.spec_retry = ctx->spec_retry;
    args.spec_solo = ctx->spec_solo;
    args.spec_hess = ctx->spec_hess;
    args.spec_skip = ctx->spec_skip;
    args.spec_prio = ctx->spec_prio;
    args.n_ctrl_ctas = ctx->ctrl_ctas;
    args.timeout_ns = (unsigned long long)(ctx->timeout_s * 1e9);
    args.trace = trace;
    const int grid = ctx->sm_count * ctx->solve_ctas_per_sm;
    // batches of small candidates keep their chains inside control CTAs: give those the share of the grid their work has
    if (!ctx->ctrl_ctas_fixed && d_order) args.n_ctrl_ctas += (int)(ctx->ctrl_share * (grid - 8 - args.n_ctrl_ctas) * solo_share);
    if (args.n_ctrl_ctas < 1) args.n_ctrl_ctas = 1;
    if (args.n_ctrl_ctas > grid - 1) args.n_ctrl_ctas = grid - 1;
    k_solve<<<grid, SOLVE_BLOCK, solve_smem_bytes(lcap), stream>>>(ctx->imgs, ctx->par, d_cands, ws, out, args);
    ctx->launches += 2;
    CUDA_TRY(ctx, cudaGetLastError());
    return SDSM_OK;
}

extern "C" int sdsm_solve_batch(sdsm_ctx* ctx, const sdsm_plan* plan, const sdsm_cand* d_cands, int32_t n_cand,
                                const int32_t* d_footprints, const int32_t* d_order, void* d_workspace, size_t workspace_bytes,
                                const sdsm_out* out, void* stream_)
{
    if (!ctx || !plan || !d_cands || !d_footprints || !d_order || !d_workspace || !out) return SDSM_E_ARG;
    if (!out->d_energy || !out->d_status || !out->d_info || !out->d_params || !out->d_fgbox || !out->d_onboundary || !out->d_fgmask) return SDSM_E_ARG;
    Layout L;
    int rc = check_plan(ctx, plan, n_cand, workspace_bytes, &L);
    if (rc) return rc;
    cudaStream_t stream = (cudaStream_t)stream_;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    WS ws = make_ws(d_workspace, L);
    CUDA_TRY(ctx, cudaMemsetAsync(ws.cstate, 0, sizeof(CState) * (size_t)n_cand, stream));
    CUDA_TRY(ctx, cudaMemsetAsync(ws.xig, 0, sizeof(double) * NSLOT * (size_t)L.C * SPEC_MAX, stream));
    CUDA_TRY(ctx, cudaMemsetAsync(out->d_fgmask, 0, (size_t)plan->fgmask_len, stream));

    const int pack_grid = n_cand < 8 * ctx->sm_count ? n_cand : 8 * ctx->sm_count;
    k_pack<<<pack_grid, PACK_BLOCK, 0, stream>>>(ctx->imgs, ctx->par, d_cands, n_cand, d_footprints, ws);
    ctx->launches++;
    if (ctx->par.has_smooth) {
        const size_t psf_bytes = sizeof(float) * (size_t)(ctx->par.R + 1) * (ctx->par.R + 1);
        k_grid<<<pack_grid, GRID_BLOCK, 0, stream>>>(ctx->par, d_cands, n_cand, d_order, ws);
        if (psf_bytes > 48 * 1024)
            CUDA_TRY(ctx, cudaFuncSetAttribute(k_grid_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)psf_bytes));
        k_grid_rows<<<dim3(pack_grid, ROWS_SPLIT), GRID_BLOCK, psf_bytes, stream>>>(ctx->par, d_cands, n_cand, d_order, ws);
        ctx->launches += 2;
    }
    const int lcap = ctrl_lcap(plan->n_max);
    rc = run_solver(ctx, stream, d_cands, ws, *out, lcap, d_order, n_cand, 0, nullptr, plan->solo_share);
    if (rc) return rc;
    const int fin_grid = n_cand < 16 * ctx->sm_count ? n_cand : 16 * ctx->sm_count;
    k_finish<<<fin_grid, FIN_BLOCK, 0, stream>>>(ctx->imgs, ctx->par, d_cands, n_cand, ws, *out);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    return SDSM_OK;
}

extern "C" int sdsm_eval_candidate(sdsm_ctx* ctx, const sdsm_plan* plan, const sdsm_cand* d_cands, int32_t c, const double* d_x,
                                   double* d_value, double* d_grad, double* d_H, void* d_workspace, void* stream_)
{
    if (!ctx || !plan || !d_cands || !d_x || !d_value || !d_grad || !d_workspace || c < 0 || c >= plan->n_cand) return SDSM_E_ARG;
    const Layout L = plan_layout(plan);
    WS ws = make_ws(d_workspace, L);
    cudaStream_t stream = (cudaStream_t)stream_;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const int lcap = ctrl_lcap(plan->n_max);
    const size_t smem = solve_smem_bytes(lcap);
    k_diag_eval_begin<<<1, SOLVE_BLOCK, 0, stream>>>(ctx->par, d_cands, c, ws, d_x, d_H ? 1 : 0);
    k_diag_eval_cells<<<ctx->sm_count, SOLVE_BLOCK, smem, stream>>>(ctx->par, d_cands, c, ws);
    if (ctx->par.has_smooth && d_H) { k_diag_eval_rows<<<ctx->sm_count, SOLVE_BLOCK, smem, stream>>>(ctx->par, d_cands, c, ws); ctx->launches++; }
    k_diag_eval_end<<<1, SOLVE_BLOCK, smem, stream>>>(ctx->par, d_cands, c, ws, d_value, d_grad, d_H, lcap);
    ctx->launches += 3;
    CUDA_TRY(ctx, cudaGetLastError());
    return SDSM_OK;
}

extern "C" int sdsm_export_smooth(sdsm_ctx* ctx, const sdsm_plan* plan, const sdsm_cand* d_cands, int32_t c, uint32_t* d_rc,
                                  int32_t* d_cols, float* d_vals, uint32_t* d_grid, void* d_workspace, void* stream_)
{
    if (!ctx || !plan || !d_cands || !d_rc || !d_cols || !d_vals || !d_grid || !d_workspace || c < 0 || c >= plan->n_cand) return SDSM_E_ARG;
    const Layout L = plan_layout(plan);
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
    if (!ctx || !plan || !d_cands || !d_x0 || !d_trace || !d_xout || !d_status_iters || !d_workspace || c < 0 || c >= plan->n_cand) return SDSM_E_ARG;
    const Layout L = plan_layout(plan);
    WS ws = make_ws(d_workspace, L);
    cudaStream_t stream = (cudaStream_t)stream_;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const int lcap = ctrl_lcap(plan->n_max);
    k_diag_trace_begin<<<1, SOLVE_BLOCK, 0, stream>>>(ctx->par, d_cands, c, use_xi, ws, d_x0);
    ctx->launches++;
    sdsm_out none{};   // STG_TRACE never writes results
    int rc = run_solver(ctx, stream, d_cands, ws, none, lcap, nullptr, 1, c, d_trace, 0.0);
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

extern "C" int sdsm_solver_stats(sdsm_ctx* ctx, const sdsm_plan* plan, const void* d_workspace, uint64_t* h_stats, void* stream_)
{
    if (!ctx || !plan || !d_workspace || !h_stats) return SDSM_E_ARG;
    const Layout L = plan_layout(plan);
    WS ws = make_ws(const_cast<void*>(d_workspace), L);
    cudaStream_t stream = (cudaStream_t)stream_;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    uint64_t h_qc[N_HOT * HOT_STRIDE], h_qe[N_HOT * HOT_STRIDE], h_gl[N_GLOB * HOT_STRIDE];
    CUDA_TRY(ctx, cudaMemcpyAsync(h_stats, ws.counters, sizeof(uint64_t) * N_COUNTERS, cudaMemcpyDeviceToHost, stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(h_qc, ws.qc.hot, sizeof(h_qc), cudaMemcpyDeviceToHost, stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(h_qe, ws.qe.hot, sizeof(h_qe), cudaMemcpyDeviceToHost, stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(h_gl, ws.glob, sizeof(h_gl), cudaMemcpyDeviceToHost, stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(stream));
    h_stats[Q_HEAD] = h_qc[H_EV_HEAD * HOT_STRIDE] + h_qe[H_EV_HEAD * HOT_STRIDE];     // event tickets taken by consumers
    h_stats[Q_TAIL] = h_qc[H_EV_TAIL * HOT_STRIDE] + h_qe[H_EV_TAIL * HOT_STRIDE];     // tasks published
    h_stats[Q_HEAD_HI] = h_qc[H_HEAD_HI * HOT_STRIDE] + h_qe[H_HEAD_HI * HOT_STRIDE];  // of which urgent, taken
    h_stats[Q_TAIL_HI] = h_qc[H_TAIL_HI * HOT_STRIDE] + h_qe[H_TAIL_HI * HOT_STRIDE];  // of which urgent, published
    h_stats[Q_DONE] = h_gl[G_DONE * HOT_STRIDE]; h_stats[Q_ABORT] = h_gl[G_ABORT * HOT_STRIDE];
    return SDSM_OK;
}

extern "C" int sdsm_sweep_eval(sdsm_ctx* ctx, const sdsm_plan* plan, const sdsm_cand* d_cands, int32_t n_cand, int32_t want_h,
                               int32_t rows, int32_t repeats, void* d_workspace, void* stream_)
{
    if (!ctx || !plan || !d_cands || !d_workspace || n_cand != plan->n_cand || repeats < 1) return SDSM_E_ARG;
    const Layout L = plan_layout(plan);
    WS ws = make_ws(d_workspace, L);
    cudaStream_t stream = (cudaStream_t)stream_;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CUDA_TRY(ctx, cudaMemsetAsync(ws.counters, 0, sizeof(unsigned long long) * N_COUNTERS, stream));
    const int prep_grid = n_cand < 8 * ctx->sm_count ? n_cand : 8 * ctx->sm_count;
    k_sweep_prep<<<prep_grid, SOLVE_BLOCK, 0, stream>>>(ctx->par, d_cands, n_cand, ws, want_h, rows);
    const size_t smem = solve_smem_bytes(ctrl_lcap(plan->n_max));
    for (int r = 0; r < repeats; ++r)
        k_sweep<<<ctx->sm_count * ctx->solve_ctas_per_sm, SOLVE_BLOCK, smem, stream>>>(ctx->par, d_cands, ws, rows);
    ctx->launches += 1 + repeats;
    CUDA_TRY(ctx, cudaGetLastError());
    return SDSM_OK;
}

extern "C" int sdsm_eval_counts(sdsm_ctx* ctx, const sdsm_plan* plan, const void* d_workspace, int32_t* h_counts, void* stream_)
{
    if (!ctx || !plan || !d_workspace || !h_counts) return SDSM_E_ARG;
    const Layout L = plan_layout(plan);
    WS ws = make_ws(const_cast<void*>(d_workspace), L);
    cudaStream_t stream = (cudaStream_t)stream_;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    // the four counters are the last members of CState: one strided copy takes them out of the candidates' states
    static_assert(offsetof(CState, n_hwasted) + sizeof(int32_t) == sizeof(CState) && offsetof(CState, n_skipped) + 4 * sizeof(int32_t) == sizeof(CState), "counters last");
    CUDA_TRY(ctx, cudaMemcpy2DAsync(h_counts, 4 * sizeof(int32_t), reinterpret_cast<const char*>(ws.cstate) + offsetof(CState, n_skipped), sizeof(CState),
                                    4 * sizeof(int32_t), (size_t)plan->n_cand, cudaMemcpyDeviceToHost, stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(stream));
    return SDSM_OK;
}

extern "C" int sdsm_finish_times(sdsm_ctx* ctx, const sdsm_plan* plan, const void* d_workspace, int32_t* h_us, void* stream_)
{
    if (!ctx || !plan || !d_workspace || !h_us) return SDSM_E_ARG;
    const Layout L = plan_layout(plan);
    WS ws = make_ws(const_cast<void*>(d_workspace), L);
    cudaStream_t stream = (cudaStream_t)stream_;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    CUDA_TRY(ctx, cudaMemcpyAsync(h_us, ws.pending, sizeof(int32_t) * (size_t)plan->n_cand, cudaMemcpyDeviceToHost, stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(h_us + plan->n_cand, ws.prio, sizeof(int32_t) * (size_t)plan->n_cand, cudaMemcpyDeviceToHost, stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(stream));
    return SDSM_OK;
}
