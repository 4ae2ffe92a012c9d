// K-1: the Preprocessing stage on the device (preprocess.py:39-68): the offset image
//     y = G_sigma1 * g  -  [ (1 - t) G_sigma2 * clip(g, 0, c) + t G_sigma2 * g ],   t = (max(sigma2 - EDT(g <= c), 0) / max)^2
// bit for bit as scipy computes it.  scipy.ndimage.gaussian_filter = correlate1d along axis 0, then axis 1 (float64 in
// between), boundary mode 'reflect', and for the symmetric Gaussian kernel NI_Correlate1D sums
//     tmp = x[l] w[0];  for j = -r .. -1:  tmp += (x[l + j] + x[l - j]) w[j]
// in exactly this order (no contraction: the library is built with -fmad=false); oracle check: tests/test_gpu_round2.py.
#pragma once
#include "common.cuh"

namespace sdsm {

// scipy's 'reflect' extension (d c b a | a b c d | d c b a), any distance
__device__ __forceinline__ int reflect_index(int i, int n)
{
    const int p = 2 * n;
    i %= p;
    if (i < 0) i += p;
    return i >= n ? p - 1 - i : i;
}

// one pass along `axis` (0: down the columns, 1: along the rows).  w: the 2 r + 1 weights (device), centre w[r].
// clip_hi < +inf: the input is clip(x, 0, clip_hi) (first pass of the clipped offset, preprocess.py:54).
__global__ void k_gauss_pass(const double* __restrict__ in, double* __restrict__ out, int H, int W, const double* __restrict__ w, int r,
                             int axis, double clip_hi)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x, row = blockIdx.y;
    if (c >= W) return;
    const bool clip = clip_hi < 1e300;
    const int n = axis == 0 ? H : W, l = axis == 0 ? row : c;
    const size_t stride = axis == 0 ? (size_t)W : 1;
    const double* line = axis == 0 ? in + c : in + (size_t)row * W;
    auto at = [&](int i) { double v = line[(size_t)i * stride]; if (clip) v = fmin(fmax(v, 0.0), clip_hi); return v; };
    double tmp = at(l) * w[r];
    if (l - r >= 0 && l + r < n) {
        for (int j = -r; j < 0; ++j) tmp = tmp + (at(l + j) + at(l - j)) * w[j + r];
    } else {
        for (int j = -r; j < 0; ++j) tmp = tmp + (at(reflect_index(l + j, n)) + at(reflect_index(l - j, n))) * w[j + r];
    }
    out[(size_t)row * W + c] = tmp;
}

// t0 = max(sigma2 - EDT(g <= c), 0) from the windowed vertical distances gv (k_edt_cols with the threshold c, window Mi >= sigma2),
// and its maximum over the image (non-negative doubles order like their bit patterns).
__global__ void k_blend_weight(const uint16_t* __restrict__ gv, int H, int W, int Mi, double sigma2, double* __restrict__ t0,
                               unsigned long long* __restrict__ tmax_bits)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
    double t = 0.0;
    if (c < W) {
        long long best = -1;
        for (int dc = -Mi; dc <= Mi; ++dc) {
            const int cc = c + dc;
            if (cc < 0 || cc >= W) continue;
            const int g = gv[(size_t)r * W + cc];
            if (g == 65535) continue;
            const long long d2 = (long long)dc * dc + (long long)g * g;
            if (best < 0 || d2 < best) best = d2;
        }
        if (best >= 0) t = fmax(sigma2 - sqrt((double)best), 0.0);
        t0[(size_t)r * W + c] = t;
    }
    unsigned long long b = (unsigned long long)__double_as_longlong(t);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const unsigned long long v = __shfl_xor_sync(0xffffffffu, b, o); b = v > b ? v : b; }
    if ((threadIdx.x & 31) == 0 && b) atomicMax(tmax_bits, b);
}

// y = smooth1 - [(1 - t) clipped + t original],  t = (t0 / max)^2   (preprocess.py:58-64); lower_clip > -inf: lower_clip_mean
__global__ void k_offset_image(const double* __restrict__ smooth1, const double* __restrict__ original, const double* __restrict__ clipped,
                               const double* __restrict__ t0, const unsigned long long* __restrict__ tmax_bits, int have_clip,
                               double lower_clip, size_t n, double* __restrict__ y)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double off = original[i];
    if (have_clip) {
        const double tmax = __longlong_as_double((long long)*tmax_bits);
        double t = t0[i] / tmax;
        t = t * t;
        off = (1.0 - t) * clipped[i] + t * original[i];
    }
    if (lower_clip > -1e300) off = off > lower_clip ? off : lower_clip;      // np.max([offset, mean]): NaN-free inputs
    y[i] = smooth1[i] - off;
}

}  // namespace sdsm
