// TEST INFRASTRUCTURE ONLY: compiles the product's per-lane physics headers (billiard_ai_b200/csrc/*.cuh)
// for the HOST with g++, so that their arithmetic can be checked bit-for-bit against the oracle in the
// CPU-only test tier.  This object is never loaded by the product (which has no CPU path); it lives in
// tests/_build/.
#include <cstdint>
#define BCU_COUNT_FALLBACKS 1
#include "../billiard_ai_b200/csrc/bcu_physics.cuh"

using namespace bcu;

extern "C" {

int hh_quartic_batch(int64_t n, const double* coef5, double* outReIm8, int32_t* outN) {
    for (int64_t i = 0; i < n; ++i) {
        const double* c = coef5 + 5 * i;
        cplx r[4];
        int k = solve_quartic_roots(c[0], c[1], c[2], c[3], c[4], r);
        outN[i] = k;
        for (int j = 0; j < k; ++j) { outReIm8[8 * i + 2 * j] = r[j].re; outReIm8[8 * i + 2 * j + 1] = r[j].im; }
    }
    return 0;
}

int hh_first_root_batch(int64_t n, const double* coef5, float* outT, int32_t* outHas) {
    for (int64_t i = 0; i < n; ++i) {
        const double* c = coef5 + 5 * i;
        float t = -1.0f;
        outHas[i] = quartic_first_real_root(c[0], c[1], c[2], c[3], c[4], &t) ? 1 : 0;
        outT[i] = t;
    }
    return 0;
}

int hh_time_to_collision_batch(int64_t n, const float* p1, const float* v1, const int32_t* roll1, const float* p2,
                               const float* v2, const int32_t* roll2, float diameter, float* outTime, int32_t* outHas) {
    PhysConst pc = phys_const();
    for (int64_t i = 0; i < n; ++i) {
        bcu::v2 P1 = mk(p1[2 * i], p1[2 * i + 1]), V1 = mk(v1[2 * i], v1[2 * i + 1]);
        bcu::v2 P2 = mk(p2[2 * i], p2[2 * i + 1]), V2 = mk(v2[2 * i], v2[2 * i + 1]);
        bcu::v2 n1, a1, n2, a2;
        derive_ball(V1, roll1[i] != 0, pc, &n1, &a1);
        derive_ball(V2, roll2[i] != 0, pc, &n2, &a2);
        float t = -1.0f;
        bool has = time_to_collision(a1, V1, P1, a2, V2, P2, diameter, &t);
        outHas[i] = has ? 1 : 0;
        outTime[i] = has ? t : -1.0f;
    }
    return 0;
}

long long hh_fallback_count(int reset) { long long v = g_fallbacks; if (reset) g_fallbacks = 0; return v; }

// complex helpers, for direct checks against glibc / libgcc
void hh_csqrt(double re, double im, double* out2) { cplx r = g_csqrt(c_make(re, im)); out2[0] = r.re; out2[1] = r.im; }
void hh_clog(double re, double im, double* out2) { cplx r = g_clog(c_make(re, im)); out2[0] = r.re; out2[1] = r.im; }
void hh_cdiv(double a, double b, double c, double d, double* out2) { cplx r = c_div(c_make(a, b), c_make(c, d)); out2[0] = r.re; out2[1] = r.im; }
double hh_hypot(double x, double y) { return g_hypot(x, y); }
void hh_cpow(double re, double im, double y, double* out2) { cplx r = c_pow_real(c_make(re, im), y); out2[0] = r.re; out2[1] = r.im; }

// glibc itself, called from C so that nothing (numpy SIMD loops, ...) stands between the test and libm
void hh_glibc_batch(int which, int64_t n, const double* x, const double* y, double* out, double* out2) {
    for (int64_t i = 0; i < n; ++i) {
        switch (which) {
            case 0: out[i] = ::exp(x[i]); break;
            case 1: out[i] = ::log(x[i]); break;
            case 2: out[i] = ::log1p(x[i]); break;
            case 3: out[i] = ::pow(x[i], y[i]); break;
            case 4: out[i] = ::atan2(y[i], x[i]); break;
            case 5: ::sincos(x[i], &out[i], &out2[i]); break;
            case 6: out[i] = ::hypot(x[i], y[i]); break;
        }
    }
}

double hh_exp(double x) { return g_exp(x); }
double hh_log(double x) { return g_log(x); }
double hh_log1p(double x) { return g_log1p(x); }
double hh_pow(double x, double y) { return g_pow(x, y); }
double hh_atan2(double y, double x) { return g_atan2(y, x); }
void hh_sincos(double x, double* s, double* c) { g_sincos(x, s, c); }

void hh_unary_batch(int which, int64_t n, const double* x, const double* y, double* out, double* out2) {
    for (int64_t i = 0; i < n; ++i) {
        switch (which) {
            case 0: out[i] = g_exp(x[i]); break;
            case 1: out[i] = g_log(x[i]); break;
            case 2: out[i] = g_log1p(x[i]); break;
            case 3: out[i] = g_pow(x[i], y[i]); break;
            case 4: out[i] = g_atan2(y[i], x[i]); break;
            case 5: g_sincos(x[i], &out[i], &out2[i]); break;
            case 6: out[i] = g_hypot(x[i], y[i]); break;
        }
    }
}

}  // extern "C"
