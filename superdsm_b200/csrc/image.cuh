// K0: per-image preparation on the device -- the candidate-independent part of Object.get_cvxprog_region
// (objects.py:126-127):  usable = y_mask AND (distance_transform_edt(y <= 0) <= background_margin).
// scipy's EDT gives, for a pixel with y <= 0, the Euclidean distance to the nearest pixel with y > 0 (0 for the others),
// as sqrt of the exact integer squared distance.  Only the comparison with the margin M is needed, so the exact
// two-pass separable algorithm is run on windows of +-floor(M):
//   k_edt_cols:  g(r, c)  = min |r - r'| over r' with y(r', c) > 0, |r - r'| <= floor(M)   (65535: none)
//   k_edt_rows:  d2(r, c) = min over |dc| <= floor(M) of dc^2 + g(r, c + dc)^2;  usable iff sqrt((double) d2) <= M
// -- every pixel whose true distance is <= M has its nearest y > 0 pixel inside both windows, so the thresholded result is
// exact (bit-equal to the scipy expression).  k_edt_rows also writes the masked atoms map and gathers, per atom label, the
// number of usable pixels and their bounding box (what engine.describe_candidates needs for the descriptors).
#pragma once
#include "common.cuh"

namespace sdsm {

constexpr int EDT_MAX_MARGIN = 65534;     // vertical distances are kept in 16 bits (65535: no foreground within the window)

// feature pixels: y > thr (thr = 0 for the region mask of objects.py:126-127)
__global__ void k_edt_cols(const double* __restrict__ y, int H, int W, int Mi, uint16_t* __restrict__ g, double thr)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
    if (c >= W) return;
    int best = 65535;
    if (!(y[(size_t)r * W + c] <= thr)) best = 0;            // scipy: the zero elements of (y <= thr)
    else {
        for (int d = 1; d <= Mi; ++d) {
            const bool up = r - d >= 0 && !(y[(size_t)(r - d) * W + c] <= thr);
            const bool dn = r + d < H && !(y[(size_t)(r + d) * W + c] <= thr);
            if (up || dn) { best = d; break; }
        }
    }
    g[(size_t)r * W + c] = (uint16_t)best;
}

// margin < 0 stands for +inf (every masked pixel is usable)
__global__ void k_edt_rows(const uint16_t* __restrict__ g, const uint8_t* __restrict__ ymask, const int32_t* __restrict__ atoms, int H, int W,
                           int Mi, double margin, int n_labels, int32_t* __restrict__ atoms_out, unsigned long long* __restrict__ count,
                           int32_t* __restrict__ box)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
    if (c >= W) return;
    const size_t at = (size_t)r * W + c;
    bool usable = ymask[at] != 0;
    if (usable && margin >= 0) {
        long long best = -1;
        for (int dc = -Mi; dc <= Mi; ++dc) {
            const int cc = c + dc;
            if (cc < 0 || cc >= W) continue;
            const int gv = g[(size_t)r * W + cc];
            if (gv == 65535) continue;
            const long long d2 = (long long)dc * dc + (long long)gv * gv;
            if (best < 0 || d2 < best) best = d2;
        }
        usable = best >= 0 && sqrt((double)best) <= margin;
    }
    const int lab = usable ? atoms[at] : 0;
    atoms_out[at] = lab;
    if (lab > 0 && lab <= n_labels) {
        atomicAdd(count + lab, 1ull);
        atomicMin(box + 4 * lab + 0, r); atomicMax(box + 4 * lab + 1, r + 1);
        atomicMin(box + 4 * lab + 2, c); atomicMax(box + 4 * lab + 3, c + 1);
    }
}

__global__ void k_edt_init(int n_labels, int H, int W, unsigned long long* __restrict__ count, int32_t* __restrict__ box)
{
    for (int lab = blockIdx.x * blockDim.x + threadIdx.x; lab <= n_labels; lab += gridDim.x * blockDim.x) {
        count[lab] = 0ull;
        box[4 * lab + 0] = H; box[4 * lab + 1] = 0; box[4 * lab + 2] = W; box[4 * lab + 3] = 0;
    }
}

}  // namespace sdsm
