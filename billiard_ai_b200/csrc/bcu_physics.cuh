// Per-lane physics primitives: binary32, no FMA, same operation order as the reference
// (lib/billiard_physics/src/billiard_physics.cpp, glm 0.9.9.8 scalar semantics).
// Each function cites the reference lines it follows.
#pragma once

#include "bcu_common.cuh"
#include "bcu_quartic.cuh"

namespace bcu {

struct v2 {
    float x, y;
};
BCU_HD v2 mk(float x, float y) { v2 r; r.x = x; r.y = y; return r; }
BCU_HD v2 operator+(v2 a, v2 b) { return mk(a.x + b.x, a.y + b.y); }
BCU_HD v2 operator-(v2 a, v2 b) { return mk(a.x - b.x, a.y - b.y); }
BCU_HD v2 operator-(v2 a) { return mk(-a.x, -a.y); }
BCU_HD v2 operator*(v2 a, float s) { return mk(a.x * s, a.y * s); }
BCU_HD v2 operator*(float s, v2 a) { return mk(s * a.x, s * a.y); }
BCU_HD bool is_zero(v2 a) { return a.x == 0.0f && a.y == 0.0f; }
// glm::dot(vec2): tmp = a*b; tmp.x + tmp.y
BCU_HD float dot(v2 a, v2 b) { float tx = a.x * b.x, ty = a.y * b.y; return tx + ty; }
// glm::length = sqrt(dot(v,v)); glm::normalize = v * (1/sqrt(dot(v,v)))
BCU_HD float length(v2 a) { return sqrtf(dot(a, a)); }
BCU_HD v2 glm_normalize(v2 a) { return a * (1.0f / sqrtf(dot(a, a))); }
// billiard::physics::normalize (billiard_physics.cpp:637-642): zero-safe
BCU_HD v2 safe_normalize(v2 a) { return is_zero(a) ? mk(0.0f, 0.0f) : glm_normalize(a); }

// state::BallState ctor (type.cpp:176-184): velocityNormed and acceleration from velocity.
BCU_HD void derive_ball(v2 vel, bool rolling, const PhysConst& pc, v2* vn, v2* acc) {
    *vn = safe_normalize(vel);
    float len = rolling ? pc.accLen : pc.slideAccLen;
    *acc = len * (*vn);  // accelerationByNormed (:271-273): accelerationLength * velocity
}

// physics::position (:279-281), physics::accelerate (:275-277)
BCU_HD v2 advance_position(v2 acc, v2 vel, float dt, v2 pos) {
    return (0.5f * acc * dt * dt) + (vel * dt) + pos;
}
BCU_HD v2 advance_velocity(v2 acc, v2 vel, float dt) { return acc * dt + vel; }

// physics::timeToStop (:240-253)
BCU_HD float time_to_stop(v2 acc, v2 vel) {
    float tx = acc.x == 0.0f ? 0.0f : (-vel.x / acc.x);
    float ty = acc.y == 0.0f ? 0.0f : (-vel.y / acc.y);
    return tx < ty ? ty : tx;  // std::max(timeX, timeY)
}

// physics::timeToRolling (:255-257)
BCU_HD float time_to_rolling(v2 vel, const PhysConst& pc) { return length(vel) / pc.rollDenom; }

// intersection::solveQuadraticFormula (:562-582).  NOTE: the unqualified sqrt(discriminant) in the
// reference resolves to ::sqrt(double) (libstdc++ <cmath> puts no float overload in the global
// namespace), so q is rounded once from a double expression:  q = (float)(-0.5 * ((double)b +- sqrt((double)disc))).
// Returns the number of roots; x[0] <= x[1].
BCU_HD int solve_quadratic(float a, float b, float c, float x[2]) {
    float disc = b * b - 4 * a * c;
    if (disc < 0) return 0;
    if (disc == 0) { x[0] = -0.5f * b / a; return 1; }
    double sq = sqrt((double)disc);
    float q = (b > 0) ? (float)(-0.5 * ((double)b + sq)) : (float)(-0.5 * ((double)b - sq));
    float x0 = q / a;
    float x1 = c / q;
    if (x0 <= x1) { x[0] = x0; x[1] = x1; } else { x[0] = x1; x[1] = x0; }
    return 2;
}

// intersection::lineIntersectsLine (:651-723)
BCU_HD bool line_intersects_line(v2 P, v2 D, v2 Q, v2 V, float* t, float* u) {
    float detDV = D.x * V.y - D.y * V.x;
    if (detDV == 0.0f) return false;
    float detVP = V.x * P.y - V.y * P.x;
    float detQV = Q.x * V.y - Q.y * V.x;
    *t = (detVP + detQV) / detDV;
    *u = V.x != 0.0f ? (P.x + *t * D.x - Q.x) / V.x : (P.y + *t * D.y - Q.y) / V.y;
    return true;
}

// intersection::halfLineIntersectsHalfLine (:755-783): only has_value() is consumed by callers.
BCU_HD bool half_lines_intersect(v2 p1, v2 d1, v2 p2, v2 d2) {
    float t, u;
    if (!line_intersects_line(p1, d1, p2, d2, &t, &u)) return false;
    if (t < 0.0f) return false;
    if (u < 0.0f) return false;
    return true;
}

// physics::timeToCollisionWithRail (:181-238) with halfLineIntersectsLineSegment (:725-753).
BCU_HD bool time_to_rail(v2 pos, v2 vel, v2 acc, const RailGeom& rail, float* out) {
    if (is_zero(vel)) return false;
    float t, u;
    if (!line_intersects_line(pos, vel, mk(rail.sx, rail.sy), mk(rail.dx, rail.dy), &t, &u)) return false;
    if (t < 0.0f) return false;
    if (u < 0.0f || u > 1.0f) return false;
    float lambda = t;
    if (lambda < 0.0f) return false;
    float distance = length(lambda * vel);
    float velocity = length(vel);
    float a = -length(acc);
    float x[2];
    if (solve_quadratic(0.5f * a, velocity, -distance, x) == 0) return false;
    *out = x[0];
    return true;
}

// physics::timeToTarget (:383-457) with lineIntersectsCircle (:584-625).
BCU_HD bool time_to_pocket(const PocketGeom& pk, v2 pos, v2 vn, v2 vel, v2 acc, float* out) {
    if (is_zero(vel)) return false;
    v2 delta = pos - mk(pk.x, pk.y);
    float qa = 1;
    float qb = 2 * dot(vn, delta);
    float qc = dot(delta, delta) - pk.r * pk.r;
    float l2[2];
    if (solve_quadratic(qa, qb, qc, l2) == 0) return false;
    float l = l2[0];
    if (l < 0.0f) return false;
    v2 distance = l * vn;
    float a = -0.5f * length(acc);
    float b = length(vel);
    float c = -length(distance);
    float x[2];
    if (solve_quadratic(a, b, c, x) == 0) return false;
    float first = x[0];
    if (first < 0.0f) return false;
    *out = first;
    return true;
}

// physics::pointToHalfLineSquaredDistance (:54-70)
BCU_HD float point_to_half_line_sq(v2 linePoint, v2 lineDir, v2 point) {
    v2 dn = glm_normalize(lineDir);
    v2 toPoint = point - linePoint;
    float factor = dot(dn, toPoint) / length(lineDir);
    if (factor < 0) {
        v2 s = point - linePoint;
        return dot(s, s);
    }
    v2 dv = toPoint - (lineDir * factor);
    return dot(dv, dv);
}

// The geometric pre-checks of physics::timeToCollision (:287-324): returns true when the reference
// goes on to build and solve the quartic for ball 1 -> ball 2.
BCU_HD bool collision_precheck(v2 v1, v2 p1, v2 v2_, v2 p2, float diameter) {
    if (is_zero(v1)) return false;
    float v2sq = dot(v2_, v2_);
    if (v2sq > 0) {
        if (!half_lines_intersect(p1, v1, p2, v2_)) {
            v2 from1To2 = p2 - p1;
            v2 v1n = glm_normalize(v1);
            v2 v2n = glm_normalize(v2_);
            v2 offset = from1To2 - v1n * dot(from1To2, v1n);
            v2 back1 = v1n * -diameter;
            v2 back2 = v2n * -diameter;
            if (!half_lines_intersect(p1 + offset + back1, v1, p2 + back2, v2_)) {
                float velocityDistance = dot(v1n, v2n);
                float offsetSq = dot(offset, offset);
                if (velocityDistance != 1 && offsetSq > diameter * diameter) return false;
            }
        }
    } else {
        float distance = point_to_half_line_sq(p1, v1, p2);
        if (distance > diameter * diameter) return false;
    }
    return true;
}

// The moving-vs-moving arm of the same pre-checks (:292-317) with the normalised velocities passed in: a ball's stored
// velocityNormed (state::BallState, type.cpp:176-184) is glm::normalize(velocity), the very value the reference recomputes
// here, so the two square roots and divisions per call are saved without changing a bit.  Both velocities are non-zero.
BCU_HD bool collision_precheck_moving(v2 v1, v2 p1, v2 v1n, v2 v2_, v2 p2, v2 v2n, float diameter) {
    if (!half_lines_intersect(p1, v1, p2, v2_)) {
        v2 from1To2 = p2 - p1;
        v2 offset = from1To2 - v1n * dot(from1To2, v1n);
        v2 back1 = v1n * -diameter;
        v2 back2 = v2n * -diameter;
        if (!half_lines_intersect(p1 + offset + back1, v1, p2 + back2, v2_)) {
            float velocityDistance = dot(v1n, v2n);
            float offsetSq = dot(offset, offset);
            if (velocityDistance != 1 && offsetSq > diameter * diameter) return false;
        }
    }
    return true;
}

// Quantised quartic coefficients of physics::timeToCollision (:326-337).
BCU_HD void collision_coefficients(v2 a1, v2 v1, v2 p1, v2 a2, v2 v2_, v2 p2, float diameter, float co[5]) {
    float toMeter = 1.0f / 1000.0f;
    float dm = diameter * toMeter;
    v2 dA = (a1 - a2) * toMeter;
    v2 dV = (v1 - v2_) * toMeter;
    v2 dP = (p1 - p2) * toMeter;
    co[0] = floorf(0.25f * dot(dA, dA) * 1000000) / 1000000;
    co[1] = floorf(dot(dA, dV) * 1000000) / 1000000;
    co[2] = floorf((dot(dA, dP) + dot(dV, dV)) * 1000000) / 1000000;
    co[3] = floorf(2 * (dot(dV, dP)) * 1000000) / 1000000;
    co[4] = floorf((dot(dP, dP) - (dm * dm)) * 1000000) / 1000000;
}

// Tail of physics::timeToCollision (:339-380): quartic, first near-real root, negative => none.
BCU_HD bool collision_time_from_coefficients(const float co[5], float* out) {
    float t;
    if (!quartic_first_real_root((double)co[0], (double)co[1], (double)co[2], (double)co[3], (double)co[4], &t))
        return false;
    if (t < 0.0f) return false;
    *out = t;
    return true;
}

// physics::timeToCollision (:283-381), complete (used by tests and by the static-partner path).
BCU_HD bool time_to_collision(v2 a1, v2 v1, v2 p1, v2 a2, v2 v2_, v2 p2, float diameter, float* out) {
    if (!collision_precheck(v1, p1, v2_, p2, diameter)) return false;
    float co[5];
    collision_coefficients(a1, v1, p1, a2, v2_, p2, diameter, co);
    return collision_time_from_coefficients(co, out);
}

// physics::variableTopspin (:121-130)
BCU_HD v2 variable_topspin(v2 velocity, v2 vt, const PhysConst& pc) {
    v2 topspin = pc.topspinRest * velocity;
    v2 vtn = vt + topspin;
    if (dot(vtn, vtn) < dot(velocity, velocity)) return vtn;
    return length(velocity) * safe_normalize(vtn);
}

// physics::elasticCollision (:132-166), isTopspinConstant == false
BCU_HD void elastic_collision(v2 p1, v2 v1, v2 p2, v2 v2_, bool rolling1, bool rolling2, const PhysConst& pc,
                              v2* out1, v2* out2) {
    v2 z = glm_normalize(p2 - p1);
    v2 w1 = v1 * pc.lossBall;
    v2 w2 = v2_ * pc.lossBall;
    v2 v1z = dot(w1, z) * z;
    v2 v2z = dot(w2, z) * z;
    v2 v1t = w1 - v1z;
    v2 v2t = w2 - v2z;
    v2 v1tn = rolling1 ? variable_topspin(w1, v1t, pc) : v1t;
    v2 v2tn = rolling2 ? variable_topspin(w2, v2t, pc) : v2t;
    *out1 = v2z + v1tn;
    *out2 = v1z + v2tn;
}

// physics::railCollision (:168-175): glm::reflect(v, n) * energyLossFactorRail,
// reflect(I,N) = I - N * dot(N,I) * 2
BCU_HD v2 rail_collision(v2 vel, v2 normal, const PhysConst& pc) {
    v2 refl = vel - normal * dot(normal, vel) * 2.0f;
    return refl * pc.lossRail;
}

}  // namespace bcu
