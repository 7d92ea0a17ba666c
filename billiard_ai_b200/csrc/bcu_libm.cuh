// Real-valued elementary functions used by the quartic solver (via bcu_math.cuh).
//
// Goal: the same result BITS as glibc 2.39 on an FMA-capable x86-64 host -- what the reference's
// std::pow(complex<double>, double) / std::sqrt(complex<double>) calls resolve to (libstdc++ <complex> ->
// clog, csqrt, exp, pow, sincos: exactly the libm imports of the reference when built on this image).  glibc selects its
// "_fma" ifunc variants on every AVX2/FMA host; each g_* function below restates one of them operation by
// operation, INCLUDING which multiply-adds the x86-64 build fuses (read off the machine code of
// libm-2.39.a with tools/annot_disasm.py), because a fused and an un-fused evaluation differ in the last bit.
// Lookup tables are the library's own (tools/extract_libm_tables.py -> bcu_libm_tables.inc).
//
//   g_exp    <- __ieee754_exp_fma   (sysdeps/ieee754/dbl-64/e_exp.c,  Szabolcs Nagy's algorithm)
//   g_log    <- __ieee754_log_fma   (sysdeps/ieee754/dbl-64/e_log.c)
//   g_pow    <- __ieee754_pow_fma   (sysdeps/ieee754/dbl-64/e_pow.c), positive finite base
//   g_log1p  <- __log1p_fma         (sysdeps/ieee754/dbl-64/s_log1p.c)
//   g_sincos <- __sincos_fma        (sysdeps/ieee754/dbl-64/s_sincos.c + s_sin.c do_sin/do_cos), |x| < 1.44
//   g_atan2  <- __ieee754_atan2_fma (sysdeps/ieee754/dbl-64/e_atan2.c, 2.39: single-stage)
//
// Inputs outside the ranges this path produces (overflow / underflow / NaN handling, huge arguments of sincos)
// take the platform's own function (BCU_LIBM_FALLBACK): they cannot occur for finite quartic coefficients and
// are not part of the parity claim.
//
// The code is plain C++ so that tests/host_harness.cpp can compile it with g++ -ffp-contract=off and compare it
// with the host's glibc bit for bit (tests/test_libm_exact.py); the product only ever runs it on the device.
#pragma once

#include "bcu_common.cuh"

#include <math.h>

namespace bcu {

BCU_HD double bcu_fma(double a, double b, double c) {
#if defined(__CUDA_ARCH__)
    return __fma_rn(a, b, c);
#else
    return __builtin_fma(a, b, c);
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

// n / d.  The device's IEEE division leaves its inline fast path for a ~100-instruction subroutine whenever the quotient is
// zero or subnormal, and this solver divides exact zeros all the time (real inputs carried as complex numbers with a zero
// imaginary part).  0 / d for a non-zero, non-NaN d is a zero with the XOR of the signs -- returned directly.
// -DBCU_OUTLINE_DIVSQRT=1: ONE out-of-line copy of the IEEE division and of the square root per kernel instead of an inline
// expansion (~20 instructions and a slow-path call) at each of their ~18 sites in the solver -- quartic_kernel is bound by
// instruction fetch, and these expansions are 5 KB of the 39 KB a job walks through.  Same operations, same results.
#ifndef BCU_OUTLINE_DIVSQRT
#define BCU_OUTLINE_DIVSQRT 0
#endif
#if defined(__CUDA_ARCH__) && BCU_OUTLINE_DIVSQRT
BCU_HD_NOINLINE double bcu_div_raw(double n, double d) { return n / d; }
BCU_HD_NOINLINE double bcu_sqrt(double x) { return sqrt(x); }
#else
BCU_HD double bcu_div_raw(double n, double d) { return n / d; }
BCU_HD double bcu_sqrt(double x) { return sqrt(x); }
#endif
BCU_HD double bcu_div(double n, double d) {
#if defined(__CUDA_ARCH__)
    if (n == 0.0 && d == d && d != 0.0) return u2d((d2u(n) ^ d2u(d)) & 0x8000000000000000ull);
#endif
    return bcu_div_raw(n, d);
}

// ---- tables -------------------------------------------------------------------------------------------------
#include "bcu_libm_tables.inc"

#if defined(__CUDACC__)
#define BCU_DEFINE_TABLE(name, n)                                  \
    static __device__ const uint64_t d_##name[n] = {name##_INIT};  \
    static const uint64_t h_##name[n] = {name##_INIT};
#else
#define BCU_DEFINE_TABLE(name, n) static const uint64_t h_##name[n] = {name##_INIT};
#endif
BCU_DEFINE_TABLE(BCU_EXP_HEAD, 8)
BCU_DEFINE_TABLE(BCU_EXP_TAB, 256)
BCU_DEFINE_TABLE(BCU_LOG_HEAD, 18)
BCU_DEFINE_TABLE(BCU_LOG_TAB, 256)
BCU_DEFINE_TABLE(BCU_POWLOG_HEAD, 9)
BCU_DEFINE_TABLE(BCU_POWLOG_TAB, 384)
BCU_DEFINE_TABLE(BCU_SINCOS_TAB, 440)
BCU_DEFINE_TABLE(BCU_ATAN_CIJ, 1687)
#if defined(__CUDA_ARCH__)
#define BCU_TAB(name, i) (d_##name[i])
#else
#define BCU_TAB(name, i) (h_##name[i])
#endif
#define BCU_TABD(name, i) u2d(BCU_TAB(name, i))

// ---- fallbacks (unreachable on this path; see header) --------------------------------------------------------------
// Host test builds count them (BCU_COUNT_FALLBACKS) so that the CPU tier can assert that the realistic input
// distributions never leave the restated code: on the host the fallback IS glibc and would hide a gap.
#if defined(BCU_COUNT_FALLBACKS) && !defined(__CUDA_ARCH__)
static long long g_fallbacks = 0;
#define BCU_FB() (++g_fallbacks)
#else
#define BCU_FB() ((void)0)
#endif
BCU_HD double fb_exp(double x) { BCU_FB(); return ::exp(x); }
BCU_HD double fb_log(double x) { BCU_FB(); return ::log(x); }
BCU_HD double fb_pow(double x, double y) { BCU_FB(); return ::pow(x, y); }
BCU_HD double fb_log1p(double x) { BCU_FB(); return ::log1p(x); }
BCU_HD double fb_atan2(double y, double x) { BCU_FB(); return ::atan2(y, x); }
BCU_HD void fb_sincos(double x, double* s, double* c) {
    BCU_FB();
#if defined(__CUDA_ARCH__)
    ::sincos(x, s, c);
#else
    *s = ::sin(x); *c = ::cos(x);
#endif
}

// ---- exp -------------------------------------------------------------------------------------------------------
// exp_inline of e_pow.c / the body of e_exp.c: exp(x + xtail) * (-1)^(sign_bias != 0); N = 128.
// the straight-line part of exp_core (2^-54 <= |x| < 512): no branches, so several calls interleave in one block
BCU_HD double exp_core_body(double x, double xtail, bool withTail, uint64_t signBias) {
    const double InvLn2N = BCU_TABD(BCU_EXP_HEAD, 0), Shift = BCU_TABD(BCU_EXP_HEAD, 1);
    const double NegLn2hiN = BCU_TABD(BCU_EXP_HEAD, 2), NegLn2loN = BCU_TABD(BCU_EXP_HEAD, 3);
    const double C2 = BCU_TABD(BCU_EXP_HEAD, 4), C3 = BCU_TABD(BCU_EXP_HEAD, 5);
    const double C4 = BCU_TABD(BCU_EXP_HEAD, 6), C5 = BCU_TABD(BCU_EXP_HEAD, 7);
    double kd = bcu_fma(x, InvLn2N, Shift);          // z = InvLn2N * x; kd = z + Shift  (fused)
    const uint64_t ki = d2u(kd);
    kd = kd - Shift;
    double r = bcu_fma(kd, NegLn2hiN, x);            // r = x + kd*NegLn2hiN + kd*NegLn2loN
    r = bcu_fma(kd, NegLn2loN, r);
    if (withTail) r = xtail + r;                     // e_pow.c: r += xtail
    const uint32_t idx = 2 * (uint32_t)(ki & 127);
    const uint64_t top = (ki + signBias) << 45;
    const double tail = u2d(BCU_TAB(BCU_EXP_TAB, idx));
    const uint64_t sbits = BCU_TAB(BCU_EXP_TAB, idx + 1) + top;
    const double r2 = r * r;
    // tmp = tail + r + r2*(C2 + r*C3) + r2*r2*(C4 + r*C5)
    const double p23 = bcu_fma(C3, r, C2);
    const double tr = r + tail;
    const double p45 = bcu_fma(r, C5, C4);
    double tmp = bcu_fma(p23, r2, tr);
    const double r4 = r2 * r2;
    tmp = bcu_fma(r4, p45, tmp);
    const double scale = u2d(sbits);
    return bcu_fma(scale, tmp, scale);               // scale + scale*tmp
}

BCU_HD double exp_core(double x, double xtail, bool withTail, uint64_t signBias, bool* ok) {
    const uint32_t abstop = (uint32_t)(d2u(x) >> 52) & 0x7ff;
    *ok = true;
    if (abstop - 0x3c9u >= 0x3fu) {  // |x| < 2^-54 or |x| >= 512: special handling in glibc
        if (abstop - 0x3c9u >= 0x80000000u) {
            // tiny: exp(x) = 1 + x (e_exp.c) ; pow: +-1 (e_pow.c returns 1.0 + x only when rounding matters)
            const double one = 1.0;
            return signBias ? -(one + x) : one + x;
        }
        *ok = false;  // overflow / underflow / subnormal result paths
        return 0.0;
    }
    return exp_core_body(x, xtail, withTail, signBias);
}

BCU_HD double g_exp(double x) {
    bool ok;
    const double r = exp_core(x, 0.0, false, 0, &ok);
    return ok ? r : fb_exp(x);
}

// ---- log -------------------------------------------------------------------------------------------------------
BCU_HD double g_log(double x) {
    const uint64_t ix = d2u(x);
    const uint32_t top = (uint32_t)(ix >> 48);
    const uint64_t LO = 0x3fee000000000000ULL;  // asuint64(1.0 - 0x1p-4)
    const uint64_t HI = 0x3ff1090000000000ULL;  // asuint64(1.0 + 0x1.09p-4)
    if (ix - LO < HI - LO) {
        if (ix == 0x3ff0000000000000ULL) return 0.0;
        const double r = x - 1.0;
        const double B0 = BCU_TABD(BCU_LOG_HEAD, 7), B1 = BCU_TABD(BCU_LOG_HEAD, 8), B2 = BCU_TABD(BCU_LOG_HEAD, 9);
        const double B3 = BCU_TABD(BCU_LOG_HEAD, 10), B4 = BCU_TABD(BCU_LOG_HEAD, 11), B5 = BCU_TABD(BCU_LOG_HEAD, 12);
        const double B6 = BCU_TABD(BCU_LOG_HEAD, 13), B7 = BCU_TABD(BCU_LOG_HEAD, 14), B8 = BCU_TABD(BCU_LOG_HEAD, 15);
        const double B9 = BCU_TABD(BCU_LOG_HEAD, 16), B10 = BCU_TABD(BCU_LOG_HEAD, 17);
        double q1 = bcu_fma(B2, r, B1);
        double q2 = bcu_fma(B5, r, B4);
        const double r2 = r * r;
        double q3 = bcu_fma(B8, r, B7);
        q1 = bcu_fma(r2, B3, q1);
        q2 = bcu_fma(r2, B6, q2);
        const double r3 = r * r2;
        q3 = bcu_fma(r2, B9, q3);
        q3 = bcu_fma(r3, B10, q3);
        double p = bcu_fma(q3, r3, q2);
        p = bcu_fma(p, r3, q1);
        // w = r*2^27; rhi = r + w - w; rlo = r - rhi
        const double two27 = 134217728.0;
        const double rw = bcu_fma(r, two27, r);
        const double rhi = bcu_fma(-two27, r, rw);
        const double rhi2 = rhi * rhi;
        const double rlo = r - rhi;
        const double hi = bcu_fma(rhi2, B0, r);          // hi = r + rhi*rhi*B0
        const double rmh = r - hi;
        const double rpr = r + rhi;
        double lo = bcu_fma(rhi2, B0, rmh);              // lo = r - hi + w
        const double brlo = B0 * rlo;
        lo = bcu_fma(brlo, rpr, lo);                     // lo += B0*rlo*(rhi + r)
        const double y = bcu_fma(p, r3, lo);             // y = r3*(...) + lo
        return hi + y;
    }
    if (top - 0x0010u >= 0x7ff0u - 0x0010u) return fb_log(x);  // zero, subnormal, negative, inf, nan
    const uint64_t OFF = 0x3fe6000000000000ULL;
    const uint64_t tmp = ix - OFF;
    const uint32_t i = (uint32_t)(tmp >> 45) & 127;
    const int64_t k = (int64_t)tmp >> 52;
    const uint64_t iz = ix - (tmp & (0xfffULL << 52));
    const double invc = BCU_TABD(BCU_LOG_TAB, 2 * i), logc = BCU_TABD(BCU_LOG_TAB, 2 * i + 1);
    const double z = u2d(iz);
    const double Ln2hi = BCU_TABD(BCU_LOG_HEAD, 0), Ln2lo = BCU_TABD(BCU_LOG_HEAD, 1);
    const double A0 = BCU_TABD(BCU_LOG_HEAD, 2), A1 = BCU_TABD(BCU_LOG_HEAD, 3), A2 = BCU_TABD(BCU_LOG_HEAD, 4);
    const double A3 = BCU_TABD(BCU_LOG_HEAD, 5), A4 = BCU_TABD(BCU_LOG_HEAD, 6);
    const double kd = (double)(int32_t)k;
    const double w = bcu_fma(Ln2hi, kd, logc);           // w = kd*Ln2hi + logc
    const double r = bcu_fma(z, invc, -1.0);
    const double p12 = bcu_fma(A2, r, A1);
    const double hi = r + w;
    const double r2 = r * r;
    double lo = w - hi;
    lo = lo + r;
    lo = bcu_fma(kd, Ln2lo, lo);                         // lo = w - hi + r + kd*Ln2lo
    const double rr2 = r * r2;
    const double p34 = bcu_fma(r, A4, A3);
    lo = bcu_fma(r2, A0, lo);                            // lo + r2*A[0]
    const double pp = bcu_fma(p34, r2, p12);
    const double y = bcu_fma(rr2, pp, lo);
    return y + hi;
}

// ---- pow (x > 0 finite normal, y finite) --------------------------------------------------------------------------
// checkint of e_pow.c: 0 = not an integer, 1 = odd, 2 = even
BCU_HD int pow_checkint(uint64_t iy) {
    const int e = (int)(iy >> 52) & 0x7ff;
    if (e < 0x3ff) return 0;
    if (e > 0x3ff + 52) return 2;
    if (iy & ((1ULL << (0x3ff + 52 - e)) - 1)) return 0;
    if (iy & (1ULL << (0x3ff + 52 - e))) return 1;
    return 2;
}

// log_inline of e_pow.c for a positive normal ix: lhi + llo = log(x), straight-line
BCU_HD void pow_log_part(uint64_t ix, double* outHi, double* outLo) {
    const uint64_t OFF = 0x3fe6955500000000ULL;
    const uint64_t tmp = ix - OFF;
    const uint32_t i = (uint32_t)(tmp >> 45) & 127;
    const int64_t k = (int64_t)tmp >> 52;
    const uint64_t iz = ix - (tmp & (0xfffULL << 52));
    const double z = u2d(iz);
    const double kd = (double)(int32_t)k;
    const double invc = BCU_TABD(BCU_POWLOG_TAB, 3 * i), logc = BCU_TABD(BCU_POWLOG_TAB, 3 * i + 1);
    const double logctail = BCU_TABD(BCU_POWLOG_TAB, 3 * i + 2);
    const double Ln2hi = BCU_TABD(BCU_POWLOG_HEAD, 0), Ln2lo = BCU_TABD(BCU_POWLOG_HEAD, 1);
    const double A0 = BCU_TABD(BCU_POWLOG_HEAD, 2), A1 = BCU_TABD(BCU_POWLOG_HEAD, 3), A2 = BCU_TABD(BCU_POWLOG_HEAD, 4);
    const double A3 = BCU_TABD(BCU_POWLOG_HEAD, 5), A4 = BCU_TABD(BCU_POWLOG_HEAD, 6), A5 = BCU_TABD(BCU_POWLOG_HEAD, 7);
    const double A6 = BCU_TABD(BCU_POWLOG_HEAD, 8);
    const double t1 = bcu_fma(Ln2hi, kd, logc);
    const double lo1 = bcu_fma(Ln2lo, kd, logctail);
    const double r = bcu_fma(z, invc, -1.0);
    const double ar = r * A0;
    const double p12 = bcu_fma(A2, r, A1);
    const double p34 = bcu_fma(A4, r, A3);
    const double t2 = r + t1;
    const double lo2 = (t1 - t2) + r;
    const double ar2 = r * ar;
    const double ar3 = r * ar2;
    const double lo3 = bcu_fma(ar, r, -ar2);
    const double hi = t2 + ar2;
    double q = bcu_fma(r, A6, A5);
    q = bcu_fma(q, ar2, p34);
    const double lo4 = (t2 - hi) + ar2;
    q = bcu_fma(ar2, q, p12);
    double lo = lo1 + lo2;
    lo = lo + lo3;
    lo = lo + lo4;
    lo = bcu_fma(ar3, q, lo);
    const double lhi = hi + lo;
    const double llo = (hi - lhi) + lo;
    *outHi = lhi;
    *outLo = llo;
}

BCU_HD double g_pow(double x, double y) {
    uint64_t ix = d2u(x);
    const uint64_t iy = d2u(y);
    uint32_t topx = (uint32_t)(ix >> 52);
    const uint32_t topy = (uint32_t)(iy >> 52);
    uint64_t signBias = 0;
    if (topx - 0x001u >= 0x7ffu - 0x001u || (topy & 0x7ff) - 0x3beu >= 0x43eu - 0x3beu) {
        // e_pow.c special cases.  Restated: x = +-0 and finite negative normal x, both with a finite normal-range y
        // (the quartic calls pow(b, 3) / pow(c, 3) with b, c of either sign or zero); the rest falls back.
        if ((topy & 0x7ff) - 0x3beu >= 0x43eu - 0x3beu) return fb_pow(x, y);
        if (2 * ix == 0) {
            double x2 = x * x;
            if ((ix >> 63) && pow_checkint(iy) == 1) x2 = -x2;
            return (iy >> 63) ? 1 / x2 : x2;
        }
        if ((ix >> 63) && (topx & 0x7ff) - 0x001u < 0x7ffu - 0x001u) {
            const int yint = pow_checkint(iy);
            if (yint == 0) return fb_pow(x, y);           // invalid -> NaN
            if (yint == 1) signBias = 0x800ULL << 7;      // SIGN_BIAS
            ix &= 0x7fffffffffffffffULL;
            topx &= 0x7ff;
        } else {
            return fb_pow(x, y);                          // subnormal, inf, nan
        }
    }
    double lhi, llo;
    pow_log_part(ix, &lhi, &llo);
    // ehi + elo = y * (lhi + llo)
    const double ehi = y * lhi;
    double elo = bcu_fma(lhi, y, -ehi);
    elo = bcu_fma(y, llo, elo);
    bool ok;
    const double res = exp_core(ehi, elo, true, signBias, &ok);
    return ok ? res : fb_pow(x, y);
}

// pow(x0, 3.0), pow(x1, 3.0), pow(x2, 3.0) -- each exactly the operation sequence of g_pow, but the three chains sit in
// one straight-line block so that the scheduler interleaves them (the quartic's b^3, a^3, c^3: a serial chain of three
// pows is a tenth of the solver's latency).  Anything outside the ordinary range goes through g_pow one by one.
BCU_HD void g_cubes3(double x0, double x1, double x2, double* r0, double* r1, double* r2) {
    uint64_t i0 = d2u(x0), i1 = d2u(x1), i2 = d2u(x2);
    const uint32_t e0 = (uint32_t)(i0 >> 52) & 0x7ff, e1 = (uint32_t)(i1 >> 52) & 0x7ff, e2 = (uint32_t)(i2 >> 52) & 0x7ff;
    if (e0 - 1u < 0x7feu && e1 - 1u < 0x7feu && e2 - 1u < 0x7feu) {   // finite, normal, non-zero (either sign)
        const uint64_t SIGN = 0x8000000000000000ULL, BIAS = 0x800ULL << 7;   // 3 is odd: SIGN_BIAS for a negative base
        const uint64_t s0 = (i0 & SIGN) ? BIAS : 0, s1 = (i1 & SIGN) ? BIAS : 0, s2 = (i2 & SIGN) ? BIAS : 0;
        i0 &= ~SIGN; i1 &= ~SIGN; i2 &= ~SIGN;
        double h0, l0, h1, l1, h2, l2;
        pow_log_part(i0, &h0, &l0);
        pow_log_part(i1, &h1, &l1);
        pow_log_part(i2, &h2, &l2);
        const double y = 3.0;
        const double eh0 = y * h0, eh1 = y * h1, eh2 = y * h2;
        double el0 = bcu_fma(h0, y, -eh0), el1 = bcu_fma(h1, y, -eh1), el2 = bcu_fma(h2, y, -eh2);
        el0 = bcu_fma(y, l0, el0); el1 = bcu_fma(y, l1, el1); el2 = bcu_fma(y, l2, el2);
        const uint32_t a0 = (uint32_t)(d2u(eh0) >> 52) & 0x7ff, a1 = (uint32_t)(d2u(eh1) >> 52) & 0x7ff;
        const uint32_t a2 = (uint32_t)(d2u(eh2) >> 52) & 0x7ff;
        if (a0 - 0x3c9u < 0x3fu && a1 - 0x3c9u < 0x3fu && a2 - 0x3c9u < 0x3fu) {   // exp_core's ordinary range
            *r0 = exp_core_body(eh0, el0, true, s0);
            *r1 = exp_core_body(eh1, el1, true, s1);
            *r2 = exp_core_body(eh2, el2, true, s2);
            return;
        }
    }
    *r0 = g_pow(x0, 3.0); *r1 = g_pow(x1, 3.0); *r2 = g_pow(x2, 3.0);
}

// ---- log1p -----------------------------------------------------------------------------------------------------
BCU_HD double g_log1p(double x) {
    const double ln2_hi = 6.93147180369123816490e-01, ln2_lo = 1.90821492927058770002e-10;
    const double Lp1 = 6.666666666666735130e-01, Lp2 = 3.999999999940941908e-01, Lp3 = 2.857142874366239149e-01;
    const double Lp4 = 2.222219843214978396e-01, Lp5 = 1.818357216161805012e-01, Lp6 = 1.531383769920937332e-01;
    const double Lp7 = 1.479819860511658591e-01;
    const int32_t hx = (int32_t)(d2u(x) >> 32);
    const int32_t ax = hx & 0x7fffffff;
    int k = 1, hu = 0;
    double f = 0.0, c = 0.0;
    if (hx < 0x3FDA827A) {                                   // x < 0.41422
        if (ax >= 0x3ff00000) return fb_log1p(x);            // x <= -1
        if (ax < 0x3e200000) {                               // |x| < 2^-29
            if (ax < 0x3c900000) return x;                   // |x| < 2^-54 (underflow check elided)
            return bcu_fma(-(x * x), 0.5, x);                // x - x*x*0.5   (vfnmadd231sd)
        }
        if (hx > 0 || hx <= (int32_t)0xbfd2bec3) {           // -0.2929 < x < 0.41422
            k = 0; f = x; hu = 1;
        }
    } else if (hx >= 0x7ff00000) {
        return fb_log1p(x);
    }
    if (k != 0) {
        double u;
        if (hx < 0x43400000) {
            u = 1.0 + x;
            hu = (int32_t)(d2u(u) >> 32);
            k = (hu >> 20) - 1023;
            c = (k > 0) ? 1.0 - (u - x) : x - (u - 1.0);
            c = bcu_div_raw(c, u);
        } else {
            u = x;
            hu = (int32_t)(d2u(u) >> 32);
            k = (hu >> 20) - 1023;
            c = 0;
        }
        hu &= 0x000fffff;
        if (hu < 0x6a09e) {
            u = u2d((d2u(u) & 0xffffffffULL) | ((uint64_t)(uint32_t)(hu | 0x3ff00000) << 32));
        } else {
            k += 1;
            u = u2d((d2u(u) & 0xffffffffULL) | ((uint64_t)(uint32_t)(hu | 0x3fe00000) << 32));
            hu = (0x00100000 - hu) >> 2;
        }
        f = u - 1.0;
    }
    const double hfsq = (f * 0.5) * f;                       // 0.5*f*f evaluated as (f*0.5)*f
    const double kd = (double)k;
    if (hu == 0) {                                           // |f| < 2^-20
        if (f == 0.0) {
            if (k == 0) return 0.0;
            c = bcu_fma(kd, ln2_lo, c);                      // c += k*ln2_lo
            return bcu_fma(kd, ln2_hi, c);                   // k*ln2_hi + c
        }
        const double R0 = bcu_fma(-f, 0.66666666666666666, 1.0) * hfsq;   // hfsq*(1.0 - 0.666..*f)
        if (k == 0) return f - R0;
        c = bcu_fma(kd, ln2_lo, c);
        return bcu_fma(kd, ln2_hi, -((R0 - c) - f));         // k*ln2_hi - ((R - (k*ln2_lo + c)) - f)
    }
    const double s = bcu_div_raw(f, 2.0 + f);
    const double zz = s * s;
    // R = z*Lp1 + z2*(Lp2 + z*Lp3) + z4*(Lp4 + z*Lp5) + z6*(Lp6 + z*Lp7), glibc's pairing
    const double R2 = bcu_fma(Lp3, zz, Lp2);
    const double R3 = bcu_fma(Lp5, zz, Lp4);
    const double R4 = bcu_fma(Lp7, zz, Lp6);
    const double z2 = zz * zz;
    const double z4 = z2 * z2;
    const double z6 = z2 * z4;
    double R = z2 * R2;
    R = bcu_fma(zz, Lp1, R);
    R = bcu_fma(z4, R3, R);
    R = bcu_fma(z6, R4, R);
    const double sR = (R + hfsq) * s;                        // s*(hfsq + R)
    if (k == 0) return f - (hfsq - sR);
    c = bcu_fma(kd, ln2_lo, c);
    c = c + sR;
    return bcu_fma(kd, ln2_hi, -((hfsq - c) - f));           // k*ln2_hi - ((hfsq - (s*(hfsq+R) + (k*ln2_lo + c))) - f)
}

// ---- sincos (|x| < 1.44: no range reduction) ------------------------------------------------------------------------
struct SinCosTab { double sn, ssn, cs, ccs; };

BCU_HD void g_sincos(double x, double* sinx, double* cosx) {
    const double big = 52776558133248.0;                     // 0x1.8p45
    const double sn3 = -1.66666666666664880952546298448555E-01, sn5 = 8.33333214285722277379541354343671E-03;
    const double cs2 = 4.99999999999999999999950396842453E-01, cs4 = -4.16666666666664434524222570944589E-02;
    const double cs6 = 1.38888874007937613028114285595617E-03;
    const double hp0 = 1.5707963267948966, hp1 = 6.123233995736766e-17;
    const uint32_t k = (uint32_t)(d2u(x) >> 32) & 0x7fffffff;
    const double ax = fabs(x);
    if (k < 0x3e400000) { *sinx = x; *cosx = 1.0; return; }  // |x| < 2^-27
    if (k < 0x3ff70a3d) {
        // |x| < 0.855469:          sin = do_sin(x, 0) (Taylor below 0.126),  cos = do_cos(x, 0)
        // 0.855469 <= |x| < 1.44:  a + da = pi/2 - |x|;  sin = copysign(do_cos(a, da), x),  cos = do_sin(a, da)
        // (glibc's second range ends at 2.426265; the arguments of this path stay below pi/3.)
        // do_sin / do_cos are the same two instruction streams in both ranges, so they are evaluated once on the selected
        // table argument and the results are routed by selects -- the lanes of a warp fall into both ranges.
        const bool hiRange = k >= 0x3feb6000;
        const double y = hp0 - ax;
        const double a = y + hp1;
        const double da = (y - a) + hp1;
        const double A = hiRange ? fabs(a) : ax;
        const double u = A + big;
        const uint32_t j = ((uint32_t)d2u(u)) << 2;
        const double xr = A - (u - big);
        const double sn = BCU_TABD(BCU_SINCOS_TAB, j), ssn = BCU_TABD(BCU_SINCOS_TAB, j + 1);
        const double cs = BCU_TABD(BCU_SINCOS_TAB, j + 2), ccs = BCU_TABD(BCU_SINCOS_TAB, j + 3);
        double S, C;
        {   // do_sin (xr, dxs)
            const double dxs = hiRange ? ((a <= 0) ? -da : da) : ((x > 0) ? 0.0 : -0.0);
            const double xx = xr * xr;
            const double x3 = xx * xr;
            const double ps = bcu_fma(xx, sn5, sn3);
            double s1 = bcu_fma(x3, ps, dxs);
            double pc = bcu_fma(xx, cs6, cs4);
            s1 = s1 + xr;
            pc = bcu_fma(pc, xx, cs2);
            double c = xx * pc;
            c = bcu_fma(dxs, xr, c);
            double cor = bcu_fma(s1, ccs, ssn);
            cor = bcu_fma(-c, sn, cor);
            cor = bcu_fma(s1, cs, cor);
            S = cor + sn;
        }
        {   // do_cos (xr + dxc)
            const double dxc = hiRange ? da : ((x < 0) ? -0.0 : 0.0);
            const double xc = dxc + xr;
            const double xx = xc * xc;
            const double x3 = xc * xx;
            const double ps = bcu_fma(xx, sn5, sn3);
            double pc = bcu_fma(xx, cs6, cs4);
            const double s1 = bcu_fma(x3, ps, xc);
            pc = bcu_fma(pc, xx, cs2);
            double cor = bcu_fma(-ssn, s1, ccs);
            const double c = xx * pc;
            cor = bcu_fma(-c, cs, cor);
            cor = bcu_fma(-s1, sn, cor);
            C = cor + cs;
        }
        if (hiRange) {
            *sinx = copysign(C, x);
            *cosx = copysign(S, a);
        } else {
            *cosx = C;
            if (ax < 0.126) {
                // TAYLOR_SIN (xx, x, 0)
                const double s1 = -0.16666666666666666, s2 = 0.008333333333332329, s3 = -0.00019841269834414642;
                const double s4 = 2.755729806860771e-06, s5 = -2.5022014848318398e-08;
                const double xx = x * x;
                double p = bcu_fma(s5, xx, s4);
                p = bcu_fma(p, xx, s3);
                p = bcu_fma(p, xx, s2);
                p = bcu_fma(p, xx, s1);
                const double t1 = bcu_fma(p, x, -0.0);
                const double t = bcu_fma(xx, t1, 0.0);
                *sinx = x + t;
            } else {
                *sinx = copysign(S, x);
            }
        }
        return;
    }
    fb_sincos(x, sinx, cosx);
}

// ---- atan2 -----------------------------------------------------------------------------------------------------
BCU_HD double atan2_poly(double v) {
    const double d3 = -0.3333333333333333, d5 = 0.19999999999998855, d7 = -0.14285714283953163;
    const double d9 = 0.11111109821886427, d11 = -0.09090424391727987, d13 = 0.07601836584380736;
    double p = bcu_fma(d13, v, d11);
    p = bcu_fma(p, v, d9);
    p = bcu_fma(p, v, d7);
    p = bcu_fma(p, v, d5);
    return bcu_fma(p, v, d3);
}

BCU_HD double g_atan2(double y, double x) {
    const double hpi = 1.5707963267948966, hpi1 = 6.123233995736766e-17;
    const double opi = 3.141592653589793, opi1 = 1.2246467991473532e-16;
    const double inv16 = 0.0625, two52 = 4503599627370496.0, two8 = 256.0;
    const double twom500 = 3.054936363499605e-151, two500 = 3.273390607896142e+150;
    const uint64_t ux64 = d2u(x), uy64 = d2u(y);
    const int32_t ux = (int32_t)(ux64 >> 32), uy = (int32_t)(uy64 >> 32);
    const uint32_t dy = (uint32_t)uy64;
    if ((ux & 0x7ff00000) == 0x7ff00000 || (uy & 0x7ff00000) == 0x7ff00000) return fb_atan2(y, x);  // inf / nan
    if (uy == 0 && dy == 0) return (ux < 0) ? opi : 0.0;                       // y = +0
    if (uy == (int32_t)0x80000000 && dy == 0) return (ux < 0) ? -opi : -0.0;   // y = -0
    if (x == 0.0) return (y > 0) ? hpi : -hpi;
    double ax = (x < 0) ? -x : x;
    double ay = (y < 0) ? -y : y;
    const int32_t de = (uy & 0x7ff00000) - (ux & 0x7ff00000);
    if (de >= 59768832) return (y > 0) ? hpi : -hpi;
    if (de <= -59768832) {
        if (x > 0) return copysign(ay / ax, y);   // underflow signalling elided
        return (y > 0) ? opi : -opi;
    }
    if (ax < twom500 || ay < twom500) { ax = ax * two500; ay = ay * two500; }
    if (ax > two500 || ay > two500) { ax = ax * twom500; ay = ay * twom500; }
    // u = min/max, du = its correction: one stream for both orders (operands selected)
    const bool yLess = ay < ax;
    const double num = yLess ? ay : ax, den = yLess ? ax : ay;
    const double u = bcu_div_raw(num, den);
    const double du = bcu_div_raw((num - u * den) - bcu_fma(u, den, -(u * den)), den);
    double z;
    if (x > 0 && yLess) {                                 // (i) atan(ay/ax)
        if (u < inv16) {
            const double v = u * u;
            const double p = atan2_poly(v);
            const double zz = bcu_fma(u * v, p, du);
            z = u + zz;
        } else {
            const int i = (int)(bcu_fma(u, two8, two52) - two52) - 16;
            const double c0 = BCU_TABD(BCU_ATAN_CIJ, 7 * i), t1 = BCU_TABD(BCU_ATAN_CIJ, 7 * i + 1);
            const double t2 = BCU_TABD(BCU_ATAN_CIJ, 7 * i + 2), c3 = BCU_TABD(BCU_ATAN_CIJ, 7 * i + 3);
            const double c4 = BCU_TABD(BCU_ATAN_CIJ, 7 * i + 4), c5 = BCU_TABD(BCU_ATAN_CIJ, 7 * i + 5);
            const double c6 = BCU_TABD(BCU_ATAN_CIJ, 7 * i + 6);
            const double t3 = u - c0;
            const double v = du + t3;                 // EADD (t3, du, v, dv)
            const double dv = (fabs(t3) > fabs(du)) ? ((t3 - v) + du) : ((du - v) + t3);
            const double vsq = v * v;
            double q = bcu_fma(c6, v, c5);
            q = bcu_fma(q, v, c4);
            q = bcu_fma(q, v, c3);
            q = vsq * q;
            q = bcu_fma(dv, t2, q);
            const double zz = bcu_fma(v, t2, q);
            z = zz + t1;
        }
    } else {
        // (ii) x > 0, ay >= ax:  pi/2 - atan(ax/ay)      base = pi/2, sign -
        // (iii) x < 0, ay > ax:  pi/2 + atan(ax/ay)      base = pi/2, sign +
        // (iv) x < 0, ay <= ax:  pi   - atan(ay/ax)      base = pi,   sign -
        // glibc writes the three with explicit + / -; a - b == a + (-b) and fma(-q, v, c) == fma(q, -v, c) bit for bit,
        // so one stream with a sign serves all three.
        const bool plus = !(x > 0) && ay > ax;
        const bool usePi = !(x > 0) && !(ay > ax);
        const double base = usePi ? opi : hpi, base1 = usePi ? opi1 : hpi1;
        if (u < inv16) {
            const double v = u * u;
            const double p = atan2_poly(v);
            const double zz = (u * v) * p;
            const double su = plus ? u : -u, sdu = plus ? du : -du, szz = plus ? zz : -zz;
            const double t2 = base + su;              // ESUB / EADD (base, u, t2, cor); base > |u| here (u < 1/16)
            const double cor = (base - t2) + su;
            double t3 = cor + base1;
            t3 = t3 + sdu;
            t3 = t3 + szz;
            z = t3 + t2;
        } else {
            const int i = (int)(bcu_fma(u, two8, two52) - two52) - 16;
            const double c0 = BCU_TABD(BCU_ATAN_CIJ, 7 * i), c1 = BCU_TABD(BCU_ATAN_CIJ, 7 * i + 1);
            const double v = (u - c0) + du;
            double q = bcu_fma(BCU_TABD(BCU_ATAN_CIJ, 7 * i + 6), v, BCU_TABD(BCU_ATAN_CIJ, 7 * i + 5));
            q = bcu_fma(q, v, BCU_TABD(BCU_ATAN_CIJ, 7 * i + 4));
            q = bcu_fma(q, v, BCU_TABD(BCU_ATAN_CIJ, 7 * i + 3));
            q = bcu_fma(q, v, BCU_TABD(BCU_ATAN_CIJ, 7 * i + 2));
            const double zz = bcu_fma(plus ? q : -q, v, base1);
            const double t1 = base + (plus ? c1 : -c1);
            z = t1 + zz;
        }
    }
    return copysign(z, y);
}

}  // namespace bcu
