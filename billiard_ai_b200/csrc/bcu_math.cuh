// Double-precision complex arithmetic for the quartic collision-time solver, written so that
// every rounding step is the one the reference takes on x86-64 / GCC 13 / glibc 2.39:
//
//   std::complex<double> * + - /      -> GCC's inline expansions and libgcc's __divdc3
//   std::sqrt(complex<double>)        -> glibc csqrt (math/s_csqrt_template.c)
//   std::pow(complex<double>, double) -> libstdc++ <complex>: clog + exp + sincos (polar), or
//                                        real pow when the base is a positive real
//   std::log(complex<double>)         -> glibc clog (math/s_clog_template.c), incl. hypot,
//                                        log1p and __x2y2m1 branches
//
// Reference call sites: lib/billiard_physics/src/billiard_physics.cpp:470-548.
// The real-valued elementary functions (exp, log, pow, log1p, sin, cos, atan2) come from
// bcu_libm.cuh, which restates glibc 2.39's x86-64 FMA variants (the ones ifunc selects on any
// FMA-capable host).
#pragma once

#include "bcu_common.cuh"
#include "bcu_libm.cuh"

namespace bcu {

struct cplx {
    double re, im;
};

BCU_HD cplx c_make(double re, double im = 0.0) { cplx z; z.re = re; z.im = im; return z; }
BCU_HD cplx c_add(cplx a, cplx b) { return c_make(a.re + b.re, a.im + b.im); }
BCU_HD cplx c_sub(cplx a, cplx b) { return c_make(a.re - b.re, a.im - b.im); }
BCU_HD cplx c_neg(cplx a) { return c_make(-a.re, -a.im); }
// complex (op) real scalar: libstdc++ complex<double>::operator*=(double) etc. act componentwise;
// operator+=(double) touches the real part only.
BCU_HD cplx c_scale(cplx a, double s) { return c_make(a.re * s, a.im * s); }
BCU_HD cplx c_divs(cplx a, double s) { return c_make(a.re / s, a.im / s); }
BCU_HD cplx c_adds(cplx a, double s) { return c_make(a.re + s, a.im); }
BCU_HD cplx c_subs(cplx a, double s) { return c_make(a.re - s, a.im); }
// GCC tree-complex.cc expand_complex_multiplication: (ac - bd) + i(ad + bc); the NaN-recovery call
// to __muldc3 is not reproduced (a candidate with NaN roots yields no collision either way).
BCU_HD cplx c_mul(cplx a, cplx b) {
    double ac = a.re * b.re, bd = a.im * b.im, ad = a.re * b.im, bc = a.im * b.re;
    return c_make(ac - bd, ad + bc);
}
BCU_HD bool c_eq0(cplx a) { return a.re == 0.0 && a.im == 0.0; }

// libgcc2.c __divdc3 (GCC 13): Smith's algorithm with Baudin-Smith scaling.  Scalings are by
// powers of two and only trigger for |denominator| outside [2^-52 .. DBL_MAX/2], or numerators
// below DBL_MIN; kept for fidelity.
BCU_HD_NOINLINE cplx c_div(cplx n, cplx dd) {
    double a = n.re, b = n.im, c = dd.re, d = dd.im;
    const double RBIG = 1.7976931348623157e308 / 2;
    const double RMIN = 2.2250738585072014e-308;
    const double RMIN2 = 2.220446049250313e-16;
    const double RMINSCAL = 1 / 2.220446049250313e-16;
    const double RMAX2 = RBIG * RMIN2;
    // The two arms of Smith's algorithm (|c| < |d| and not) are mirror images; here they share one instruction stream
    // with the operands swapped by selects (lanes of a warp take both arms about equally often):
    //   arm 1 (|c| < |d|): ratio = c/d, denom = c*ratio + d, x = (a*ratio + b)/denom, y = (b*ratio - a)/denom
    //   arm 2            : ratio = d/c, denom = d*ratio + c, x = (b*ratio + a)/denom, y = (b - a*ratio)/denom
    // (y of arm 1 is t - a with t = b*ratio, y of arm 2 is b - t with t = a*ratio: the same subtraction, operands swapped.)
    const bool arm1 = fabs(c) < fabs(d);
    double lo = arm1 ? c : d, hi = arm1 ? d : c;   // |lo| <= |hi| (arm 2 also takes |c| == |d|)
    if (fabs(hi) >= RBIG) { a = a / 2; b = b / 2; lo = lo / 2; hi = hi / 2; }
    if (fabs(hi) < RMIN2) {
        a = a * RMINSCAL; b = b * RMINSCAL; lo = lo * RMINSCAL; hi = hi * RMINSCAL;
    } else if (((fabs(a) < RMIN) && (fabs(b) < RMAX2) && (fabs(hi) < RMAX2)) ||
               ((fabs(b) < RMIN) && (fabs(a) < RMAX2) && (fabs(hi) < RMAX2))) {
        a = a * RMINSCAL; b = b * RMINSCAL; lo = lo * RMINSCAL; hi = hi * RMINSCAL;
    }
    const double u = arm1 ? a : b, v = arm1 ? b : a;   // x = (u*ratio + v) / denom
    const double ratio = bcu_div(lo, hi);
    const double denom = (lo * ratio) + hi;
    double x, y;
    if (fabs(ratio) > RMIN) {
        const double t = v * ratio;
        x = bcu_div((u * ratio) + v, denom);
        y = bcu_div(arm1 ? (t - u) : (u - t), denom);
    } else {
        const double t = lo * bcu_div(v, hi);
        x = bcu_div((lo * bcu_div(u, hi)) + v, denom);
        y = bcu_div(arm1 ? (t - u) : (u - t), denom);
    }
    // Annex-G infinity/zero recovery (both parts NaN) is not reproduced; see c_mul.
    return c_make(x, y);
}

// glibc 2.39 sysdeps/ieee754/dbl-64/e_hypot.c, non-FMA kernel (the generic x86-64 build; there is
// no ifunc variant of hypot).  Finite inputs only.
BCU_HD double g_hypot(double x, double y) {
    x = fabs(x);
    y = fabs(y);
    double ax = x < y ? y : x;
    double ay = x < y ? x : y;
    const double SCALE = 0x1p-600, LARGE_VAL = 0x1p+511, TINY_VAL = 0x1p-459, EPS = 0x1p-54;
    double scaleBack = 1.0;
    if (ax > LARGE_VAL) {
        if (ay <= ax * EPS) return ax + ay;
        ax = ax * SCALE; ay = ay * SCALE; scaleBack = 0x1p+600;
    } else if (ay < TINY_VAL) {
        if (ax >= ay / EPS) return ax + ay;
        ax = ax / SCALE; ay = ay / SCALE; scaleBack = SCALE;
    } else if (ax * EPS >= ay) {
        return ax + ay;
    }
    double h = bcu_sqrt(ax * ax + ay * ay);
    // both correction formulas are evaluated and one is selected (no divergent branch; identical operations)
    const bool small = h <= 2.0 * ay;
    const double delta = h - (small ? ay : ax);
    const double t1a = ax * (2.0 * delta - ax);
    const double t2a = (delta - 2.0 * (ax - ay)) * delta;
    const double t1b = 2.0 * delta * (ax - 2.0 * ay);
    const double t2b = (4.0 * delta - ay) * ay + delta * delta;
    const double t1 = small ? t1a : t1b, t2 = small ? t2a : t2b;
    h -= bcu_div(t1 + t2, 2.0 * h);
    return scaleBack == 1.0 ? h : h * scaleBack;
}

// glibc math/s_csqrt_template.c, finite operands (the quartic never feeds inf; NaN propagates).
BCU_HD_NOINLINE cplx g_csqrt(cplx x) {
    const double DMIN = 2.2250738585072014e-308, DMAX = 1.7976931348623157e308;
    cplx res;
    if (x.im == 0.0) {
        // re < 0: (0, copysign(sqrt(-re), im));  otherwise: (|sqrt(re)|, copysign(0, im)).  One square root of |re| serves both
        // arms (|sqrt(-0)| = sqrt(0) = 0): the lanes of a warp take them about equally often (the two discriminants of a
        // quartic with two real roots have opposite signs).
        const double r = bcu_sqrt(fabs(x.re));
        const bool neg = x.re < 0;
        res.re = neg ? 0.0 : r;
        res.im = copysign(neg ? r : 0.0, x.im);
        return res;
    }
    if (x.re == 0.0) {
        double r;
        if (fabs(x.im) >= 2 * DMIN) r = bcu_sqrt(0.5 * fabs(x.im));
        else r = 0.5 * bcu_sqrt(2 * fabs(x.im));
        res.re = r;
        res.im = copysign(r, x.im);
        return res;
    }
    double d, r, s;
    int scale = 0;
    if (fabs(x.re) > DMAX / 4) {
        scale = 1;
        x.re = x.re * 0.25; x.im = x.im * 0.25;
    } else if (fabs(x.im) > DMAX / 4) {
        scale = 1;
        if (fabs(x.re) >= 4 * DMIN) x.re = x.re * 0.25; else x.re = 0;
        x.im = x.im * 0.25;
    } else if (fabs(x.re) < 2 * DMIN && fabs(x.im) < 2 * DMIN) {
        scale = -((53 + 1) / 2);
        x.re = x.re * 0x1p+54; x.im = x.im * 0x1p+54;
    }
    d = g_hypot(x.re, x.im);
    {
        // re > 0:  r = sqrt(0.5 (d + re)), s = 0.5 (im / r);   re < 0:  s = sqrt(0.5 (d - re)), r = |0.5 (im / s)|.
        // d - re with re < 0 is d + |re|, so both arms are  t = sqrt(0.5 (d + |re|)), u = 0.5 (im / t)  with roles swapped.
        const bool pos = x.re > 0;
        double t = bcu_sqrt(0.5 * (d + fabs(x.re))), u;
        if (scale == 1 && fabs(x.im) < 1) { u = bcu_div(x.im, t); t = t * 2.0; scale = 0; }
        else u = 0.5 * bcu_div(x.im, t);
        r = pos ? t : fabs(u);
        s = pos ? u : t;
    }
    if (scale == 1) { r = r * 2.0; s = s * 2.0; }
    else if (scale != 0) { r = r * 0x1p-27; s = s * 0x1p-27; }
    res.re = r;
    res.im = copysign(s, x.im);
    return res;
}

// glibc sysdeps/ieee754/dbl-64/x2y2m1.c: x*x + y*y - 1 with exact products and a sorted,
// error-free summation.  The exact low product word is the same whether produced by Dekker
// splitting (glibc's generic build) or by an FMA (here).
BCU_HD void x2y2m1_cswap(double& a, double& b) {   // one step of a stable insertion sort on |.|
    const bool sw = fabs(a) > fabs(b);
    const double t = a;
    a = sw ? b : a;
    b = sw ? t : b;
}
BCU_HD double g_x2y2m1(double x, double y) {
    double v0, v1, v2, v3, v4;
    v1 = x * x; v0 = bcu_fma(x, x, -v1);
    v3 = y * y; v2 = bcu_fma(y, y, -v3);
    v4 = -1.0;
    // glibc sorts by |.| ascending (qsort) and re-sorts the tail after every add_split.  An insertion sort is a fixed
    // sequence of adjacent compare-exchanges (the ones after the insertion point are no-ops), so it runs in registers
    // without a branch; strict '>' keeps equal magnitudes in their order, as the array version did.
    x2y2m1_cswap(v0, v1);
    x2y2m1_cswap(v1, v2); x2y2m1_cswap(v0, v1);
    x2y2m1_cswap(v2, v3); x2y2m1_cswap(v1, v2); x2y2m1_cswap(v0, v1);
    x2y2m1_cswap(v3, v4); x2y2m1_cswap(v2, v3); x2y2m1_cswap(v1, v2); x2y2m1_cswap(v0, v1);
    double hi;
    // i = 0: add_split(v1, v0), re-sort v1..v4
    hi = v1 + v0; v0 = (v1 - hi) + v0; v1 = hi;
    x2y2m1_cswap(v1, v2);
    x2y2m1_cswap(v2, v3); x2y2m1_cswap(v1, v2);
    x2y2m1_cswap(v3, v4); x2y2m1_cswap(v2, v3); x2y2m1_cswap(v1, v2);
    // i = 1: add_split(v2, v1), re-sort v2..v4
    hi = v2 + v1; v1 = (v2 - hi) + v1; v2 = hi;
    x2y2m1_cswap(v2, v3);
    x2y2m1_cswap(v3, v4); x2y2m1_cswap(v2, v3);
    // i = 2: add_split(v3, v2), re-sort v3..v4
    hi = v3 + v2; v2 = (v3 - hi) + v2; v3 = hi;
    x2y2m1_cswap(v3, v4);
    // i = 3: add_split(v4, v3)
    hi = v4 + v3; v3 = (v4 - hi) + v3; v4 = hi;
    return v4 + v3 + v2 + v1 + v0;
}

// glibc math/s_clog_template.c, finite non-zero operand.
BCU_HD cplx g_clog(cplx x) {
    const double DMIN = 2.2250738585072014e-308, DMAX = 1.7976931348623157e308, DEPS = 2.220446049250313e-16;
    const double LN2 = 0.693147180559945309417232121458176568;
    cplx result;
    if (x.re == 0.0 && x.im == 0.0) {
        // clog(0): imag = signbit(re) ? pi : 0 with the sign of im; real = -1/|re| = -inf
        result.im = signbit(x.re) ? 3.14159265358979323846 : 0.0;
        result.im = copysign(result.im, x.im);
        result.re = -1.0 / fabs(x.re);
        return result;
    }
    double absx = fabs(x.re), absy = fabs(x.im);
    int scale = 0;
    if (absx < absy) { double t = absx; absx = absy; absy = t; }
    if (absx > DMAX / 2) {
        scale = -1;
        absx = absx * 0.5;
        absy = (absy >= DMIN * 2 ? absy * 0.5 : 0);
    } else if (absx < DMIN && absy < DMIN) {
        scale = 53;
        absx = absx * 0x1p+53; absy = absy * 0x1p+53;
    }
    if (absx == 1 && scale == 0) {
        result.re = g_log1p(absy * absy) / 2;
    } else if (absx > 1 && absx < 2 && absy < 1 && scale == 0) {
        double d2m1 = (absx - 1) * (absx + 1);
        if (absy >= DEPS) d2m1 += absy * absy;
        result.re = g_log1p(d2m1) / 2;
    } else if (absx < 1 && absx >= 0.5 && absy < DEPS / 2 && scale == 0) {
        double d2m1 = (absx - 1) * (absx + 1);
        result.re = g_log1p(d2m1) / 2;
    } else if (absx < 1 && absx >= 0.5 && scale == 0 && absx * absx + absy * absy >= 0.5) {
        double d2m1 = g_x2y2m1(absx, absy);
        result.re = g_log1p(d2m1) / 2;
    } else {
        double d = g_hypot(absx, absy);
        result.re = g_log(d) - scale * LN2;
    }
    result.im = g_atan2(x.im, x.re);
    return result;
}

// libstdc++ <complex> std::pow(const complex<double>&, const double&) with _GLIBCXX_USE_C99_COMPLEX.
BCU_HD_NOINLINE cplx c_pow_real(cplx x, double y) {
    if (x.im == 0.0 && x.re > 0.0) return c_make(g_pow(x.re, y), 0.0);
    cplx t = g_clog(x);
    double rho = g_exp(y * t.re);
    double theta = y * t.im;
    double sn, cs;
    g_sincos(theta, &sn, &cs);
    return c_make(rho * cs, rho * sn);  // std::polar
}

// glibc csqrt is what std::sqrt(complex<double>) resolves to.
BCU_HD cplx c_sqrt(cplx x) { return g_csqrt(x); }

}  // namespace bcu
