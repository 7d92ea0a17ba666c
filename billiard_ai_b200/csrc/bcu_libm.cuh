// Real-valued elementary functions used by the quartic solver (via bcu_math.cuh).
//
// Goal: the same result bits as glibc 2.39 on an FMA-capable x86-64 host, which is what the
// reference's std::pow / std::log(complex) / std::polar calls resolve to when it runs on the
// GPU box's host CPU.  Each g_* function states which glibc routine it follows.
#pragma once

#include "bcu_common.cuh"

#include <math.h>

namespace bcu {

BCU_HD double bcu_fma(double a, double b, double c) {
#if defined(__CUDA_ARCH__)
    return __fma_rn(a, b, c);
#else
    return ::fma(a, b, c);
#endif
}

BCU_HD uint64_t d2u(double x) {
#if defined(__CUDA_ARCH__)
    return (uint64_t)__double_as_longlong(x);
#else
    uint64_t u; __builtin_memcpy(&u, &x, 8); return u;
#endif
}
BCU_HD double u2d(uint64_t u) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)u);
#else
    double x; __builtin_memcpy(&x, &u, 8); return x;
#endif
}

#if !defined(BCU_EXACT_LIBM)
// Phase-1 placeholders: the platform's own libm (glibc on the host, CUDA libdevice on the GPU).
BCU_HD double g_exp(double x) { return ::exp(x); }
BCU_HD double g_log(double x) { return ::log(x); }
BCU_HD double g_log1p(double x) { return ::log1p(x); }
BCU_HD double g_pow(double x, double y) { return ::pow(x, y); }
BCU_HD double g_atan2(double y, double x) { return ::atan2(y, x); }
BCU_HD void g_sincos(double x, double* s, double* c) {
#if defined(__CUDA_ARCH__)
    ::sincos(x, s, c);
#else
    *s = ::sin(x); *c = ::cos(x);
#endif
}
#endif

}  // namespace bcu
