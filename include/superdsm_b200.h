/*
 * superdsm_b200 -- C ABI of the B200-native DSM candidate solver.
 *
 * This is the drop-in boundary for SuperDSM's per-candidate convex fit.  The reference has no FFI for this path (it
 * is pure Python on top of cvxopt + MKL via ctypes); the seam this library sits behind is
 *
 *     superdsm/objects.py:270-284   _compute_objects(objects, y, atoms, x_map, lock, dsm_cfg, log_root_dir)
 *     superdsm/objects.py:361-412   cvxprog(region, scale, epsilon, alpha, smooth_amount, ...)
 *
 * and the entry points below are what a ctypes binding of that seam calls (INTEGRATION.md shows the stub).
 * One call of sdsm_solve_batch() does, for every candidate of the batch, what objects.py:177-212 (_compute_object)
 * does for one: region extraction, smooth-matrix construction, elliptical initialisation (objects.py:321-358), the
 * DSM solve (dsm.py:431-490 -> cvxopt.solvers.cp), status/fallback logic (objects.py:394-410) and the rasterisation
 * of the result (objects.py:198-211).
 *
 * Conventions: plain C types only; every pointer is caller-owned; d_* pointers are DEVICE pointers on the context's
 * device (e.g. torch tensors' data_ptr()), h_* pointers are host pointers.  The library allocates no per-batch device
 * memory: the caller provides the workspace sdsm_plan_batch() asked for.  All work is enqueued on the caller's
 * stream; nothing synchronises.  Errors are integer codes (0 = ok), solver trouble is per-candidate data
 * (status word), never a C++ exception.  One context per device; a context is not thread-safe.
 *
 * Environment (read by sdsm_create), all optional tuning knobs of the persistent solver kernel:
 *   SDSM_CTRL_CTAS (48 + a share of the grid that grows with the batch's share of small candidates)   CTAs that run control
 *                         tasks; the warps of all other CTAs run evaluation / Hessian-row items
 *   SDSM_CTRL_SHARE (0.95)  share of the rest of the grid that becomes control CTAs per unit of the batch's solo share
 *   SDSM_SOLO_CELLS (512 / smooth_subsample)  candidates with at most this many pixel cells run their whole chain inside
 *                         one control CTA
 *   SDSM_TASK_SMALL_CELLS (512)  candidates with at most this many cells get evaluation items of one cell (else 4)
 *   SDSM_HI_FRAC (0.02)   share of the batch (most expensive candidates first) whose tasks are urgent from the start
 *   SDSM_HI_EVALS (300)   a candidate that has needed more evaluations than this becomes urgent
 *   SDSM_TIMEOUT_S (60)   watchdog of the persistent solver kernel: it is aborted when no control task has finished anywhere
 *                         for this long (lack of progress, not total run time: batches may run for minutes)
 */
#ifndef SUPERDSM_B200_H
#define SUPERDSM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SDSM_ABI_VERSION 5
#define SDSM_MAX_IMAGES  32   /* images one batch can draw its candidates from (sdsm_set_images) */

/* return codes */
#define SDSM_OK              0
#define SDSM_E_ARG           1   /* bad argument */
#define SDSM_E_CUDA          2   /* CUDA runtime error, see sdsm_last_error() */
#define SDSM_E_UNSUPPORTED   3   /* configuration outside what the kernels implement, see sdsm_last_error() */
#define SDSM_E_WORKSPACE     4   /* workspace too small */

/* per-candidate status word (out.status) */
#define SDSM_ST_OPTIMAL      0   /* objects.cvxprog status 'optimal' (objects.py:401-403) */
#define SDSM_ST_FALLBACK     1   /* status 'fallback': DSM solve failed, initialisation returned (objects.py:399-410) */
#define SDSM_ST_SKIPPED      2   /* exactly one foreground pixel: not fitted (objects.py:184-191) */
#define SDSM_ST_ERROR        3   /* CvxprogError: both elliptical passes failed (objects.py:351-353), or an empty region */
#define SDSM_ST_BADREGION    4   /* pixel count differs from the host's, or a zero row in G~ (dsm.py:194 assertion) */

/* hyper-parameters: the dsm_cfg dictionary of dsmcfg.py:6-21 after host-side preparation */
typedef struct sdsm_params {
    double alpha;            /* dsm/alpha */
    double epsilon;          /* dsm/epsilon */
    double scale;            /* dsm/scale; the solver minimises scale / n * psi (objects.py:380) */
    double sparsity_tol;     /* dsm/sparsity_tol (dsm.py:346,363) */
    int32_t subsample;       /* dsm/smooth_subsample; also the edge of the pixel cells */
    int32_t radius;          /* R = kernel box // 2 (dsm.py:139,187); ignored if !has_smooth */
    int32_t has_smooth;      /* 0: smooth_amount = inf, no deformation parameters (c2freganal.py:126) */
    int32_t init_elliptical; /* 1: dsm/init == 'elliptical'; 0: zeros (objects.py:387-392) */
    const float* h_psf;      /* HOST pointer, (radius+1)^2 floats: psf32[|dr|][|dc|], the float32 Gaussian kernel of
                                dsm.py:137-142,226 (quarter of the symmetric box); may be NULL if !has_smooth */
    int32_t maxiters;        /* cvxopt default 100 */
    int32_t skip_single_fg;  /* 1: a region with exactly one foreground pixel is not fitted (SDSM_ST_SKIPPED) -- the rule of
                                _compute_object (objects.py:184-191); 0: it is fitted like any other, as cvxprog itself does
                                (objects.py:361-412, the c2f seam) */
    double abstol, reltol, feastol;   /* cvxopt defaults 1e-7, 1e-6, 1e-7 */
    double hessian_sparsity_tol;      /* dsm/hessian_sparsity_tol (dsm.py:376-382): off-diagonal entries of the Hessian of psi below this
                                         are dropped (deformable model only); 0: none */
} sdsm_params;

/* one candidate region.  The caller fills r0 .. n and img; sdsm_plan_batch() fills the offsets. */
typedef struct sdsm_cand {
    int32_t r0, c0, h, w;    /* bounding box (image coordinates) of the usable pixels of the footprint's atoms */
    int32_t fp_off, fp_len;  /* footprint = footprints[fp_off : fp_off + fp_len] (atom labels of image `img`) */
    int32_t n;               /* number of pixels of the region (sum of the atoms' usable-pixel counts) */
    int32_t cell_cap;        /* out: upper bound of the number of pixel cells */
    int32_t img;             /* index of the candidate's image in the set given to sdsm_set_images() (0 with sdsm_set_image) */
    int32_t reserved;
    int64_t pix_off;         /* out: offset into the per-pixel workspace arrays */
    int64_t cell_off;        /* out: offset into the per-cell workspace arrays */
    int64_t tab_off;         /* out: offset into the per-row/column tables */
    int64_t par_off;         /* out: offset into out.params (6 + cell_cap doubles reserved) */
    int64_t fg_off;          /* out: offset into out.fgmask (h * w bytes, row-major over the bounding box) */
    int64_t vec_off;         /* out: offset into the solver's vector workspace (12 vectors of 6 + cell_cap doubles) */
    int64_t mat_off;         /* out: offset into the solver's matrix workspace (Hessians and Cholesky factor) */
} sdsm_cand;

typedef struct sdsm_plan {
    size_t workspace_bytes;  /* device workspace sdsm_solve_batch() needs */
    int64_t params_len;      /* doubles in out.params */
    int64_t fgmask_len;      /* bytes in out.fgmask */
    int32_t n_max;           /* largest 6 + cell_cap of the batch */
    int32_t cells_max;       /* largest cell_cap of the batch */
    int32_t n_cand;          /* candidates of the batch */
    int32_t reserved;
    int64_t totals[6];       /* sizes of the workspace's per-pixel / per-cell / per-table / per-vector / per-matrix arrays and of the
                                task queues: the workspace layout is a pure function of these (every call that takes a plan and a
                                workspace derives the layout from the plan it is given, not from the context) */
    double  solo_share;      /* share of the batch's pixels in candidates whose chains stay inside one control CTA (scheduling hint) */
} sdsm_plan;

/* device output arrays of one batch, all caller-owned */
typedef struct sdsm_out {
    double*  d_energy;       /* [n_cand]   psi at the result, unscaled (objects.py:208) */
    int32_t* d_status;       /* [n_cand]   SDSM_ST_* */
    int32_t* d_info;         /* [n_cand*8] m, iterations (elliptical pass 1, pass 2, DSM), #fg pixels, cvxopt status
                                           bits, #value+gradient evaluations, #Hessian evaluations */
    double*  d_params;       /* [params_len] (theta, xi) of candidate c at cands[c].par_off, 6 + m values */
    int32_t* d_fgbox;        /* [n_cand*4] rmin, rmax, cmin, cmax of the foreground inside the bounding box, -1 if none */
    uint8_t* d_onboundary;   /* [n_cand]   objects.py:209 */
    uint8_t* d_fgmask;       /* [fgmask_len] (S > 0) AND omega over the candidate's bounding box (objects.py:200-204) */
    uint64_t* d_solver;      /* [2] or NULL: candidates the persistent solver finished, and its abort flag (non-zero: the watchdog
                                fired, results invalid) -- copied with the results, so that checking them costs no extra round trip */
} sdsm_out;

typedef struct sdsm_ctx sdsm_ctx;

int         sdsm_abi_version(void);
int         sdsm_create(sdsm_ctx** ctx, int device);
void        sdsm_destroy(sdsm_ctx* ctx);
const char* sdsm_last_error(const sdsm_ctx* ctx);

/* y: float64 HxW offset image (Image.model); atoms: int32 HxW atom labels with every pixel outside
 * y_mask AND (EDT(y <= 0) <= background_margin) set to 0 (objects.py:126-127, candidate-independent part). */
int sdsm_set_image(sdsm_ctx* ctx, const double* d_y, const int32_t* d_atoms, int32_t H, int32_t W);

/* The same for a set of images (batch mode, batch.py:288-361: the candidates of several images are solved by one launch, so
 * that the long dependency chains of one image overlap the bulk of the next).  sdsm_cand.img selects a candidate's image. */
typedef struct sdsm_image { const double* d_y; const int32_t* d_atoms; int32_t H, W; } sdsm_image;
int sdsm_set_images(sdsm_ctx* ctx, const sdsm_image* h_images, int32_t n_images);
int sdsm_set_params(sdsm_ctx* ctx, const sdsm_params* params);

/* Per-image preparation on the device (the candidate-independent part of Object.get_cvxprog_region, objects.py:126-127):
 * d_atoms_out[p] = d_atoms[p] if d_ymask[p] and distance_transform_edt(y <= 0)[p] <= background_margin, else 0 -- the map
 * sdsm_set_image() expects -- computed with an exact windowed Euclidean distance transform (bit-equal to the scipy
 * expression).  d_count[n_labels + 1] / d_box[4 (n_labels + 1)] receive, per atom label, the number of usable pixels and
 * their bounding box (r0, r1, c0, c1; end-exclusive; r0 = H, r1 = 0 if none).  background_margin = +inf keeps every masked
 * pixel.  d_scratch: 2 * H * W bytes.  d_ymask: one byte per pixel (numpy bool). */
int sdsm_prepare_image(sdsm_ctx* ctx, const double* d_y, const uint8_t* d_ymask, const int32_t* d_atoms, int32_t H, int32_t W,
                       double background_margin, int32_t n_labels, int32_t* d_atoms_out, uint64_t* d_count, int32_t* d_box,
                       uint8_t* d_scratch, void* stream);

/* sdsm_prepare_image for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same arguments, results
 * and errors as sdsm_prepare_image, minus d_scratch: the scratch is a file-scope device array inside the library, of
 * 2 * SDSM_NOALLOC_MAX_PIXELS bytes (8 MiB), held by every GPU that loads the library until the process exits.  Images with
 * H * W > SDSM_NOALLOC_MAX_PIXELS are refused with SDSM_E_UNSUPPORTED (sdsm_last_error() says why).  The call allocates and
 * frees nothing (no cudaMalloc / cudaFree, no allocating library routine); inputs and outputs are the caller's.  All calls on
 * a device share the one array: it keeps the previous call's contents (every call overwrites what it reads, nothing needs
 * clearing), and two calls that OVERLAP on the GPU (different streams, no ordering between them) race on it -- order them. */
#define SDSM_NOALLOC_MAX_PIXELS (4 * 1024 * 1024)
int sdsm_prepare_image_noalloc(sdsm_ctx* ctx, const double* d_y, const uint8_t* d_ymask, const int32_t* d_atoms, int32_t H, int32_t W,
                               double background_margin, int32_t n_labels, int32_t* d_atoms_out, uint64_t* d_count, int32_t* d_box,
                               void* stream);

/* The Preprocessing stage on the device (preprocess.py:39-68): d_y = gaussian_filter(g, sigma1) - offset, bit for bit as
 * scipy computes it.  h_w1 / h_w2: HOST arrays with the 2 r + 1 weights of scipy's 1-D Gaussian kernels for sigma1 / sigma2
 * (radius r = int(4 sigma + 0.5)).  offset_clip_abs = offset_clip * g.std() (the caller's numpy reduction; +inf: no clipping,
 * preprocess.py:50-51); lower_clip = g.mean() for preprocess/lower_clip_mean, -inf otherwise.  d_scratch: 5 H W doubles +
 * 2 H W + 32 bytes.  d_graw, d_y: float64 H x W on the device. */
int sdsm_preprocess(sdsm_ctx* ctx, const double* d_graw, int32_t H, int32_t W, const double* h_w1, int32_t r1, const double* h_w2,
                    int32_t r2, double sigma2, double offset_clip_abs, double lower_clip, double* d_y, double* d_scratch, void* stream);

/* fills the offset fields of h_cands[0..n_cand) and the plan; host-only, no CUDA call */
int sdsm_plan_batch(sdsm_ctx* ctx, sdsm_cand* h_cands, int32_t n_cand, sdsm_plan* plan);

/* d_cands / d_footprints / d_order: device copies of the planned candidates, the footprint labels and the processing
 * order (a permutation of 0..n_cand-1, most expensive first).  stream is a cudaStream_t. */
int sdsm_solve_batch(sdsm_ctx* ctx, const sdsm_plan* plan, const sdsm_cand* d_cands, int32_t n_cand,
                     const int32_t* d_footprints, const int32_t* d_order,
                     void* d_workspace, size_t workspace_bytes, const sdsm_out* out, void* stream);

/* Diagnostics for the parity tests: evaluates psi, grad psi and the lower triangle of the Hessian of candidate c of the
 * LAST solved batch (same workspace, still intact) at the given parameters; d_H is N x N row-major, N = 6 + m.
 * d_H = NULL evaluates value and gradient only (the evaluation the line searches use). */
int sdsm_eval_candidate(sdsm_ctx* ctx, const sdsm_plan* plan, const sdsm_cand* d_cands, int32_t c, const double* d_x,
                        double* d_value, double* d_grad, double* d_H, void* d_workspace, void* stream);

/* Diagnostics: exports the smooth matrix G~ of candidate c of the last batch as dense rows in the solver's pixel order:
 * d_rc[n] = compacted (row | col << 16), d_cols[n * 25] grid index or -1, d_vals[n * 25]; d_grid[m] = (row | col << 16). */
int sdsm_export_smooth(sdsm_ctx* ctx, const sdsm_plan* plan, const sdsm_cand* d_cands, int32_t c, uint32_t* d_rc,
                       int32_t* d_cols, float* d_vals, uint32_t* d_grid, void* d_workspace, void* stream);

/* Diagnostics: one interior-point solve (dsm.py:478-490) of candidate c of the last batch from d_x0 (6 + m values; only
 * the first 6 are used if !use_xi).  d_trace[(maxiters + 1) * 8] receives per iteration: f, gap, primal and dual
 * infeasibility, t, s, z, relaxed-iteration counter; d_status_iters[2] = {0 optimal | 1 unknown | 2 exception, iterations}. */
int sdsm_trace_solve(sdsm_ctx* ctx, const sdsm_plan* plan, const sdsm_cand* d_cands, int32_t c, int32_t use_xi,
                     const double* d_x0, double* d_trace, double* d_xout, int32_t* d_status_iters, void* d_workspace, void* stream);

/* Measurement aid: enqueues a register-resident FP64 FMA loop on every SM (8 CTAs of 256 threads per SM, 8 independent
 * chains per thread); *flops receives the floating-point operations of the launch.  Time it with CUDA events. */
int sdsm_dfma_peak(sdsm_ctx* ctx, double* d_scratch, int32_t iters, int64_t* flops, void* stream);

/* Diagnostics: h_us[2 * n_cand] receives, for every candidate of the last sdsm_solve_batch() on this workspace, the time (in
 * microseconds after the solver kernel started) at which its fit finished, then [n_cand ..] the time it spent in control
 * tasks (microseconds); SYNCHRONISES the stream. */
int sdsm_finish_times(sdsm_ctx* ctx, const sdsm_plan* plan, const void* d_workspace, int32_t* h_us, void* stream);

/* Diagnostics: h_counts[4 * n_cand] receives, for every candidate of the last sdsm_solve_batch() on this workspace, what the
 * solver did ahead of cvxopt's cp (oracle/cpsolver.py; results are bit-identical to the sequential search): [0] line-search
 * trials cp evaluates F at whose outcome does not depend on F (first pass, relaxed steps allowed: cpsolver.py:286-287) and that
 * were decided without an evaluation, [1] trial points evaluated ahead and never examined, [2] accepted trials whose Hessian
 * was evaluated with them (one pass over the pixels instead of two), [3] such Hessians evaluated in vain.  info[6] / info[7]
 * of sdsm_out keep counting the evaluations the REFERENCE makes; executed value+gradient passes = info[6] - [0] + [1] - [2].
 * SYNCHRONISES the stream. */
int sdsm_eval_counts(sdsm_ctx* ctx, const sdsm_plan* plan, const void* d_workspace, int32_t* h_counts, void* stream);

/* Measurement aid: `repeats` sweeps of the evaluation tasks (rows = 0; value + gradient, plus the per-cell Hessians if
 * want_h) or of the Hessian-row tasks (rows = 1) over EVERY candidate of the last solved batch at its solution, outside
 * the solver: the streaming pass the HBM / FP64 roofline of the energy kernels is quoted on.  Time it with CUDA events. */
int sdsm_sweep_eval(sdsm_ctx* ctx, const sdsm_plan* plan, const sdsm_cand* d_cands, int32_t n_cand, int32_t want_h,
                    int32_t rows, int32_t repeats, void* d_workspace, void* stream);

/* number of kernel launches this context has enqueued so far (bench.py's gpu_launches) */
int64_t sdsm_launch_count(const sdsm_ctx* ctx);

/* Diagnostics / watchdog: copies the 32 counters of the persistent solver of the last sdsm_solve_batch() on this workspace to
 * h_stats and SYNCHRONISES the stream.  [2] candidates finished, [3] != 0: the watchdog aborted the solve (results invalid),
 * [4..7] busy clock cycles summed over CTAs in evaluation / Hessian-row / control tasks and idle, [8..10] number of
 * evaluation / Hessian-row / control tasks, [11] CTA lifetime cycles summed over CTAs, [16..19] / [20..23] cycles / number of
 * the queued evaluation tasks by kind (polynomial model: value+gradient, +Hessian; deformable model: value+gradient, +Hessian). */
int sdsm_solver_stats(sdsm_ctx* ctx, const sdsm_plan* plan, const void* d_workspace, uint64_t* h_stats, void* stream);

#ifdef __cplusplus
}
#endif
#endif
