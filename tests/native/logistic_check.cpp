// Host check of superdsm_b200/csrc/logistic.cuh against glibc in extended precision (test infrastructure).
// Prints the worst errors in ulps of the double result: exp, log(w >= 1), h / w, and of phi / theta end to end
// relative to the reference's statement evaluated in long double FROM THE SAME h (so that the comparison is about
// log and the division, not about how an error of exp propagates).
#include "../../superdsm_b200/csrc/logistic.cuh"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>

static double ulps(double got, long double want)
{
    if (want == 0) return got == 0 ? 0 : 1e30;
    int ex; std::frexp((double)want, &ex);
    const long double ulp = std::ldexp(1.0L, ex - 53);
    return (double)(fabsl((long double)got - want) / ulp);
}

int main()
{
    std::mt19937_64 rng(12345);
    std::uniform_real_distribution<double> U(0, 1);
    double e_exp = 0, e_log = 0, e_div = 0, e_phi = 0, e_theta = 0;
    const int N = 4000000;
    for (int i = 0; i < N; ++i) {
        double t;
        const int kind = i % 8;
        if (kind < 3) t = (U(rng) - 0.5) * 80;                 // the usual range
        else if (kind < 5) t = (U(rng) - 0.5) * 4;
        else if (kind == 5) t = (U(rng) - 0.5) * 1490;          // out to the overflow / underflow ends
        else if (kind == 6) t = std::ldexp(U(rng) - 0.5, -(int)(U(rng) * 60));
        else t = 700 + U(rng) * 50;
        const double x = -t;
        if (x <= 709.782712893384) {
            const double h = sdsm::lg_exp(x);
            const long double he = expl((long double)x);
            if (he > 2.3e-308L) { const double e = ulps(h, he); if (e > e_exp) e_exp = e; }
            else if (fabsl((long double)h - he) > 4.95e-324L) { printf("subnormal exp mismatch x %.17g got %.17g want %.17Lg\n", x, h, he); return 1; }
        }
        double phi, theta;
        sdsm::logistic_eval(t, phi, theta);
        if (t >= -709.782712893384) {
            const double h = sdsm::lg_exp(-t), w = 1.0 + h;
            const double e1 = ulps(sdsm::lg_log_ge1(w), logl((long double)w));
            if (e1 > e_log) e_log = e1;
            const double e2 = ulps(theta, (long double)h / (long double)w);
            if (e2 > e_div) e_div = e2;
            const double e3 = ulps(phi, logl((long double)w));
            if (e3 > e_phi) e_phi = e3;
            if (e2 > e_theta) e_theta = e2;
        } else if (phi != -t || theta != 1.0) { printf("overflow branch wrong\n"); return 1; }
    }
    // special values
    double phi, theta;
    sdsm::logistic_eval(INFINITY, phi, theta);
    if (phi != 0 || theta != 0) { printf("t = inf wrong %g %g\n", phi, theta); return 1; }
    sdsm::logistic_eval(-INFINITY, phi, theta);
    if (!(phi == INFINITY) || theta != 1) { printf("t = -inf wrong\n"); return 1; }
    sdsm::logistic_eval(NAN, phi, theta);
    if (phi == phi) { printf("nan does not propagate\n"); return 1; }
    sdsm::logistic_eval(0.0, phi, theta);
    if (std::fabs(phi - 0.6931471805599453) > 2e-16 || theta != 0.5) { printf("t = 0 wrong %.17g %.17g\n", phi, theta); return 1; }
    sdsm::logistic_eval(-709.0, phi, theta);
    if (std::fabs(phi - 709.0) > 1e-13 || theta != 1.0) { printf("t = -709 wrong %.17g %.17g\n", phi, theta); return 1; }
    printf("max ulps: exp %.3f log %.3f div %.3f phi %.3f theta %.3f\n", e_exp, e_log, e_div, e_phi, e_theta);
    return (e_exp <= 1.05 && e_log <= 1.0 && e_div <= 0.5) ? 0 : 2;
}
