// Pipe-rate probes for sm_100a used to plan the evaluation kernels: DFMA, DMMA.8x8x4, F2F.F64.F32, MUFU-heavy exp/log.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/microbench scripts/microbench.cu ; run on the GPU box.
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) k_dfma(double* out, int iters)
{
    double a[8];
    for (int k = 0; k < 8; ++k) a[k] = 1.0 + 1e-9 * (threadIdx.x + k);
    const double b = 1.0000001, c = 1e-9;
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = fma(a[k], b, c);
    double r = 0;
    for (int k = 0; k < 8; ++k) r += a[k];
    if (r == 12345.678) out[0] = r;
}

__global__ void __launch_bounds__(256) k_dmma(double* out, int iters)
{
    double c[8][2];
    for (int k = 0; k < 8; ++k) { c[k][0] = 0; c[k][1] = 0; }
    const double a = 1.0 + 1e-9 * threadIdx.x, b = 1.0 - 1e-9 * threadIdx.x;
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int k = 0; k < 8; ++k)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[k][0]), "+d"(c[k][1]) : "d"(a), "d"(b));
    double r = 0;
    for (int k = 0; k < 8; ++k) r += c[k][0] + c[k][1];
    if (r == 12345.678) out[0] = r;
}

__global__ void __launch_bounds__(256) k_f2f(double* out, const float* in, int iters)
{
    float f[8];
    for (int k = 0; k < 8; ++k) f[k] = in[threadIdx.x + k];
    double acc = 0;
    unsigned u = 0;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            double d;
            f[k] = f[k] * 1.0000001f + 1e-7f;              // loop-variant input (FP32 pipe, full rate)
            asm volatile("cvt.f64.f32 %0, %1;" : "=d"(d) : "f"(f[k]));
            u ^= (unsigned)__double2hiint(d);
        }
    }
    acc = u;
    if (u == 0x12345678u) out[0] = acc;
}

// the same conversion with integer instructions (non-negative normal or zero inputs)
__global__ void __launch_bounds__(256) k_f2f_int(double* out, const float* in, int iters)
{
    float f[8];
    for (int k = 0; k < 8; ++k) f[k] = in[threadIdx.x + k];
    unsigned u = 0;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            unsigned b;
            f[k] = f[k] * 1.0000001f + 1e-7f;
            asm volatile("mov.b32 %0, %1;" : "=r"(b) : "f"(f[k]));
            const unsigned hi = b ? (b >> 3) + 0x38000000u : 0u, lo = b << 29;
            u ^= hi + lo;
        }
    }
    if (u == 0x12345678u) out[0] = (double)u;
}

// dependent chain: F2F feeding a DFMA accumulator (latency of the pair)
__global__ void __launch_bounds__(32) k_chain(double* out, const float* in, int iters, long long* cycles)
{
    float f = in[threadIdx.x];
    double acc = 0;
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        double d;
        asm volatile("cvt.f64.f32 %0, %1;" : "=d"(d) : "f"(f));
        acc = fma(d, 1.0000001, acc);
        f = __double2float_rn(acc) ;
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) cycles[0] = t1 - t0;
    if (acc == 12345.678) out[0] = acc;
}

__global__ void __launch_bounds__(256) k_explog(double* out, const double* in, int iters)
{
    double t = in[threadIdx.x & 31];
    double acc = 0;
    for (int i = 0; i < iters; ++i) {
        const double h = exp(-t);
        acc += log(1.0 + h) + h / (1.0 + h);
        t += 1e-3;
    }
    if (acc == 12345.678) out[0] = acc;
}

template <class F>
static float time_ms(F launch)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

int main()
{
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    const int sms = prop.multiProcessorCount, blocks = sms * 8;
    double* d_out; float* d_f; double* d_d;
    cudaMalloc(&d_out, 64); cudaMalloc(&d_f, 4096); cudaMalloc(&d_d, 4096);
    cudaMemset(d_f, 0x3f, 4096); cudaMemset(d_d, 0, 4096);
    const double clk = prop.clockRate * 1e3;
    printf("%s, %d SMs, %.0f MHz\n", prop.name, sms, clk / 1e6);
    {
        const int iters = 20000;
        const float ms = time_ms([&] { k_dfma<<<blocks, 256>>>(d_out, iters); });
        const double fma = (double)blocks * 256 * 8 * iters;
        printf("DFMA        : %8.1f GFLOP/s  %6.1f FMA/clk/SM\n", 2 * fma / ms * 1e-6, fma / (ms * 1e-3) / clk / sms);
    }
    {
        const int iters = 20000;
        const float ms = time_ms([&] { k_dmma<<<blocks, 256>>>(d_out, iters); });
        const double mma = (double)blocks * 8 * 8 * iters;           // warp-level instructions
        printf("DMMA.8x8x4  : %8.1f GFLOP/s  %6.2f clk per warp-instruction per SM\n", mma * 512 / ms * 1e-6, (ms * 1e-3) * clk * sms / mma);
    }
    {
        const int iters = 20000;
        const float ms = time_ms([&] { k_f2f<<<blocks, 256>>>(d_out, d_f, iters); });
        const double ops = (double)blocks * 256 * 8 * iters;
        printf("F2F.F64.F32 : %6.1f conversions/clk/SM\n", ops / (ms * 1e-3) / clk / sms);
    }
    {
        const int iters = 20000;
        const float ms = time_ms([&] { k_f2f_int<<<blocks, 256>>>(d_out, d_f, iters); });
        const double ops = (double)blocks * 256 * 8 * iters;
        printf("f32->f64 int: %6.1f conversions/clk/SM (4 integer instructions each)\n", ops / (ms * 1e-3) / clk / sms);
    }
    {
        long long* d_c; cudaMalloc(&d_c, 8);
        k_chain<<<1, 32>>>(d_out, d_f, 10000, d_c);
        long long c; cudaMemcpy(&c, d_c, 8, cudaMemcpyDeviceToHost);
        printf("chain F2F.F64.F32 -> DFMA -> F2F.F32.F64: %.1f clk per round\n", c / 10000.0);
    }
    {
        const int iters = 2000;
        const float ms = time_ms([&] { k_explog<<<blocks, 256>>>(d_out, d_d, iters); });
        const double ops = (double)blocks * 256 * iters;
        printf("logistic    : %6.2f clk per pixel per SM (exp + log + div, fp64)\n", (ms * 1e-3) * clk * sms / ops);
    }
    return 0;
}
