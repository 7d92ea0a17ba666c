// Closed-form quartic for the ball-ball collision time, following
// reference lib/billiard_physics/src/billiard_physics.cpp:470-560 operation by operation
// (solveQuarticFormula, solveQuartic, complex_order + std::sort, realRoot).
#pragma once

#include "bcu_math.cuh"

namespace bcu {

// 256a^3e^3 - 192a^2bde^2 - ... (billiard_physics.cpp:478-481), evaluated left to right as written.
// Only consulted in "Spezialfall 1" (:494), so callers evaluate it lazily.
BCU_HD double quartic_discriminant(double a, double b, double c, double d, double e) {
    return (256*a*a*a*e*e*e) - (192*a*a*b*d*e*e) - (128*a*a*c*c*e*e) + (144*a*a*c*d*d*e)
            -(27*a*a*d*d*d*d) + (144*a*b*b*c*e*e) - (6*a*b*b*d*d*e) - (80*a*b*c*c*d*e) + (18*a*b*c*d*d*d)
            +(16*a*c*c*c*c*e) -(4*a*c*c*c*d*d) - (27*b*b*b*b*e*e) + (18*b*b*b*c*d*e)
            -(4*b*b*b*d*d*d) - (4*b*b*c*c*c*e) + (b*b*c*c*d*d);
}

// S of billiard_physics.cpp:500: 0.5 * sqrt(-(2/3) p + (1/(3a)) (Q + deltaZero/Q))
BCU_HD cplx quartic_S(cplx p, cplx Q, cplx deltaZero, double a) {
    cplx inner = c_add(c_scale(p, -(2.0 / 3.0)), c_scale(c_add(Q, c_div(deltaZero, Q)), bcu_div_raw(1.0, (3.0 * a))));
    return c_scale(c_sqrt(inner), 0.5);
}

// Returns the number of roots (0 or 4); roots[] in the order x0,x1 of (firstPart1, discriminant1)
// then x0,x1 of (firstPart2, discriminant2), i.e. BEFORE the reference's sort.
BCU_HD int solve_quartic_roots(double a, double b, double c, double d, double e, cplx roots[4]) {
    if (a == 0.0) return 0;

    double b3, a3, c3;   // std::pow(b, 3.0), std::pow(a, 3.0), std::pow(c, 3.0) (:484-487)
    g_cubes3(b, a, c, &b3, &a3, &c3);
    cplx p = c_make(bcu_div_raw((8 * a * c) - (3 * b * b), (8 * a * a)));
    cplx q = c_make(bcu_div_raw(b3 - (4 * a * b * c) + (8 * a * a * d), (8 * a3)));
    cplx deltaZero = c_make((c * c) - (3 * b * d) + (12 * a * e));
    cplx deltaOne = c_make((2 * c3) - (9 * b * c * d) + (27 * b * b * e) + (27 * a * d * d) - (72 * a * c * e));

    // std::pow(complex, int 3) = __complex_pow_unsigned: y = x; x *= x; y *= x
    cplx dz2 = c_mul(deltaZero, deltaZero);
    cplx powerOfDeltaZero = c_scale(c_mul(deltaZero, dz2), 4.0);
    cplx deltaOneSquared = c_mul(deltaOne, deltaOne);
    cplx QInnerSqrt = c_sqrt(c_sub(deltaOneSquared, powerOfDeltaZero));
    cplx QInnerCbrt = c_divs(c_add(deltaOne, QInnerSqrt), 2.0);
    cplx Q = c_pow_real(QInnerCbrt, (1.0 / 3.0));

    // Spezialfall 1 (:494)
    if (c_eq0(Q) && c_eq0(deltaZero)) {
        if (quartic_discriminant(a, b, c, d, e) != 0) {
            QInnerCbrt = c_divs(c_sub(deltaOne, QInnerSqrt), 2.0);
            Q = c_pow_real(QInnerCbrt, (1.0 / 3.0));
        }
    }

    // Spezialfall 2 (:500-505)
    cplx S = quartic_S(p, Q, deltaZero, a);
    if (c_eq0(S)) {
        QInnerCbrt = c_divs(c_sub(deltaOne, QInnerSqrt), 2.0);
        Q = c_pow_real(QInnerCbrt, (1.0 / 3.0));
        S = quartic_S(p, Q, deltaZero, a);
    }

    cplx shared1 = c_sub(c_mul(c_scale(S, -4.0), S), c_scale(p, 2.0));
    cplx shared2 = c_div(q, S);
    cplx disc1 = c_add(shared1, shared2);
    cplx disc2 = c_sub(shared1, shared2);

    cplx firstPartShared = c_make(bcu_div_raw(-b, (4 * a)));
    cplx first1 = c_sub(firstPartShared, S);
    cplx first2 = c_add(firstPartShared, S);

    cplx s1 = c_scale(c_sqrt(disc1), 0.5);
    cplx s2 = c_scale(c_sqrt(disc2), 0.5);
    roots[0] = c_add(first1, s1);
    roots[1] = c_sub(first1, s1);
    roots[2] = c_add(first2, s2);
    roots[3] = c_sub(first2, s2);
    return 4;
}

// realRoot(sorted roots).at(0): the near-real root (|imag| < 0.001f) that sorts first under
// complex_order (real, then imag), cast to float.  Returns false when there is none.
BCU_HD bool quartic_first_real_root(double a, double b, double c, double d, double e, float* out) {
    cplx r[4];
    if (solve_quartic_roots(a, b, c, d, e, r) == 0) return false;
    const double eps = (double)0.001f;
    bool has = false;
    double best = 0.0;
    for (int i = 0; i < 4; ++i) {
        if (fabs(r[i].im) < eps) {
            if (!has || r[i].re < best) { best = r[i].re; has = true; }
        }
    }
    if (!has) return false;
    *out = (float)best;
    return true;
}

}  // namespace bcu
