// phi(t) = log(1 + exp(-t)) and theta(t) = exp(-t) / (1 + exp(-t)), evaluated the way dsm.py:298-322 states them
// (h = exp(-t); log(1 + h), not log1p; h / (1 + h)), as ONE function with shared sub-expressions.
//
// Why not exp() / log() / "/" of the CUDA math library: profiles/r01f_ksolve_hotspots.txt -- the three library calls
// are 183 dependent instructions per pixel (log 101, exp 63, division 19), half of all instructions of an evaluation,
// in a warp that has nothing else to issue meanwhile.  The library versions pay for argument ranges that cannot occur
// here (negative / subnormal / infinite arguments of log, a general divisor); this version knows 1 + h >= 1 and finite,
// and splits its polynomials in two independent halves so that two FMA chains are in flight.
//
// Accuracy (tests/test_host_cpu.py::test_logistic_restatement_accuracy compiles this file for the host and compares it
// with glibc in extended precision, 4 M points): exp 1.02 ulp, log(1 + h) 0.81 ulp, h / (1 + h) 0.5 ulp (correctly rounded),
// the same class as the library functions it replaces (CUDA: exp 1 ulp, log 1 ulp, division 0.5 ulp).
//
// Every multiply-add is an explicit fma (the library is built with -fmad=false, see build.py): the value does not depend
// on the code around the call, which the line searches of the solver rely on.
#pragma once
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define SDSM_HD __host__ __device__ __forceinline__
#else
#include <cmath>
#define SDSM_HD inline
#endif

namespace sdsm {

// 64-bit constants: from the constant bank on the device (an FMA reads c[bank][offset] directly; as immediates each costs
// two extra instructions), the same table on the host.
#define LG_TABLE { 1.4426950408889634074, -6.93147180369123816490e-01, -1.90821492927058770002e-10, 1.0 / 479001600.0, 1.0 / 6227020800.0, 1.0 / 3628800.0, 1.0 / 39916800.0, 1.0 / 40320.0, 1.0 / 362880.0, 1.0 / 720.0, 1.0 / 5040.0, 1.0 / 24.0, 1.0 / 120.0, 1.0 / 6.0, 1.531383769920937332e-01, 2.222219843214978396e-01, 1.479819860511658591e-01, 1.818357216161805012e-01, 3.999999999940941908e-01, 2.857142874366239149e-01, 6.666666666666735130e-01, 1.90821492927058770002e-10, 6.93147180369123816490e-01 }
#if defined(__CUDACC__)
__constant__ double lg_const_dev[23] = LG_TABLE;
#endif
static const double lg_const_host[23] = LG_TABLE;
#if defined(__CUDA_ARCH__)
#define LGC(i) lg_const_dev[i]
#else
#define LGC(i) lg_const_host[i]
#endif

SDSM_HD double lg_fma(double a, double b, double c)
{
#if defined(__CUDA_ARCH__)
    return fma(a, b, c);
#else
    return std::fma(a, b, c);
#endif
}

SDSM_HD double lg_from_bits(uint64_t u)
{
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)u);
#else
    double d; std::memcpy(&d, &u, 8); return d;
#endif
}

SDSM_HD uint64_t lg_bits(double d)
{
#if defined(__CUDA_ARCH__)
    return (uint64_t)__double_as_longlong(d);
#else
    uint64_t u; std::memcpy(&u, &d, 8); return u;
#endif
}

// 1 / d for a normal d whose reciprocal is normal too: ~20-bit seed and two Newton steps (relative error < 2^-52.x)
SDSM_HD double lg_rcp(double d)
{
    double r;
#if defined(__CUDA_ARCH__)
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
#else
    r = lg_from_bits(lg_bits(1.0 / d) & 0xffffffff00000000ull);     // host stand-in for the hardware seed: truncated to 20 bits
#endif
    double e = lg_fma(-d, r, 1.0);
    r = lg_fma(r, e, r);
    e = lg_fma(-d, r, 1.0);
    return lg_fma(r, e, r);
}

// exp(x) for x <= 709.79 (the caller's branch), any negative x
SDSM_HD double lg_exp(double x)
{
    x = x < -750.0 ? -750.0 : x;                          // exp(-750) = 0 in double; keeps k in range
    const double magic = 6755399441055744.0;              // 1.5 * 2^52: adding it rounds to the nearest integer
    const double kd = lg_fma(x, LGC(0), magic);
    const int k = (int)(uint32_t)lg_bits(kd);
    const double kf = kd - magic;
    double r = lg_fma(kf, LGC(1), x);   // ln2 high part (low 32 bits zero: kf * hi is exact)
    r = lg_fma(kf, LGC(2), r);          // ln2 low part
    // exp(r), |r| <= 0.3466: Taylor polynomial of degree 13 (remainder < 4e-18), even and odd halves as two chains
    const double z = r * r;
    double pe = LGC(3);                         // r^12
    double po = LGC(4);                        // r^13
    pe = lg_fma(pe, z, LGC(5));  po = lg_fma(po, z, LGC(6));
    pe = lg_fma(pe, z, LGC(7));    po = lg_fma(po, z, LGC(8));
    pe = lg_fma(pe, z, LGC(9));      po = lg_fma(po, z, LGC(10));
    pe = lg_fma(pe, z, LGC(11));       po = lg_fma(po, z, LGC(12));
    pe = lg_fma(pe, z, 0.5);              po = lg_fma(po, z, LGC(13));
    // exp(r) = 1 + r + z * (pe + r * po)
    const double q = lg_fma(r, po, pe);
    const double p = 1.0 + lg_fma(z, q, r);
    // p * 2^k in two exact-or-once-rounded steps (k down to -1083: the result may be subnormal or zero)
    const int k1 = k >> 1, k2 = k - k1;
    const double s1 = lg_from_bits((uint64_t)(uint32_t)(k1 + 1023) << 52), s2 = lg_from_bits((uint64_t)(uint32_t)(k2 + 1023) << 52);
    return (p * s1) * s2;
}

// log(w) for finite w >= 1 (fdlibm's argument reduction and evaluation order, without its special cases)
SDSM_HD double lg_log_ge1(double w)
{
    const uint64_t u = lg_bits(w);
    int e = (int)(u >> 52) - 1023;
    uint64_t mb = (u & 0x000fffffffffffffull) | 0x3ff0000000000000ull;    // mantissa in [1, 2)
    if ((u & 0x000fffffffffffffull) > 0x0006a09e667f3bccull) { mb -= 0x0010000000000000ull; e += 1; }   // > sqrt(2): halve
    const double m = lg_from_bits(mb);
    const double f = m - 1.0;                              // exact, in [sqrt(1/2) - 1, sqrt(2) - 1]
    const double s = f * lg_rcp(2.0 + f);
    const double z = s * s, z2 = z * z;
    // 2 atanh(s) - 2 s = s * R(z): minimax coefficients of fdlibm's e_log.c (Lg1 .. Lg7), odd and even halves
    double t1 = lg_fma(z2, LGC(14), LGC(15));
    double t2 = lg_fma(z2, LGC(16), LGC(17));
    t1 = lg_fma(z2, t1, LGC(18));
    t2 = lg_fma(z2, t2, LGC(19));
    t1 = z2 * t1;
    t2 = lg_fma(z2, t2, LGC(20));
    const double R = lg_fma(z, t2, t1);
    const double hfsq = 0.5 * f * f;
    const double dk = (double)e;
    // log(w) = dk ln2_hi - ((hfsq - (s (hfsq + R) + dk ln2_lo)) - f)
    const double c = lg_fma(s, hfsq + R, dk * LGC(21));
    return lg_fma(dk, LGC(22), -((hfsq - c) - f));
}

// t = y * S.  phi = log(1 + exp(-t)), theta = exp(-t) / (1 + exp(-t));  exp(-t) overflowing: dsm.py:299,309,322
SDSM_HD void logistic_eval(double t, double& phi, double& theta)
{
    if (t >= -709.782712893384) {                          // -log(finfo(float64).max), dsm.py:298
        const double h = lg_exp(-t);
        const double w = 1.0 + h;
        phi = lg_log_ge1(w);
        // h / w: reciprocal, product, one correction step; for h >= 2^54 the quotient is 1 exactly (w == h)
        const double hc = h < 18014398509481984.0 ? h : 1.0, wc = h < 18014398509481984.0 ? w : 1.0;
        const double r = lg_rcp(wc);
        const double q0 = hc * r;
        theta = lg_fma(lg_fma(-q0, wc, hc), r, q0);
    } else {
        phi = -t;
        theta = 1.0;
    }
}

}  // namespace sdsm
