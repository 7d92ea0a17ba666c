// ORACLE / TEST INFRASTRUCTURE ONLY -- never linked into, loaded by, or called from the product.
//
// Plain sequential C++ restatement of the reference's strike-evaluation path, flat arrays instead
// of unordered_map<string, Node>.  It exists next to oracle/_ref/libref_oracle.so (the reference
// itself, compiled from /root/reference) for three things the reference cannot give:
//   * near-tie detection (second-best event time at every step) for the north_star tie clause,
//   * operation counters (pair tests, quartics, rail/pocket tests) for the roofline's algorithmic work,
//   * a reference-free CPU baseline ("port") that travels to the GPU box as source.
// PINNED: tests/test_oracle_port.py checks it bit-for-bit against libref_oracle.so and against the
// reference's golden vectors (quartic KATs, filmed-shot layer structures, thesis cost logs).
//
// Float arithmetic must stay IEEE binary32 without contraction: build with -O2 -ffp-contract=off.
// The quartic deliberately goes through std::complex<double> / libm exactly as the reference does.
//
// Reference citations are to /root/reference/lib/... :
//   P = billiard_physics/src/billiard_physics.cpp, S = billiard_search/src/billiard_search.cpp,
//   T = billiard_search/src/type.cpp
#include <algorithm>
#include <atomic>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

namespace port {

struct V { float x, y; };
static inline V mk(float x, float y) { return V{x, y}; }
static inline V operator+(V a, V b) { return mk(a.x + b.x, a.y + b.y); }
static inline V operator-(V a, V b) { return mk(a.x - b.x, a.y - b.y); }
static inline V operator*(V a, float s) { return mk(a.x * s, a.y * s); }
static inline V operator*(float s, V a) { return mk(s * a.x, s * a.y); }
static inline bool zero(V a) { return a.x == 0.0f && a.y == 0.0f; }
static inline float dot(V a, V b) { float tx = a.x * b.x, ty = a.y * b.y; return tx + ty; }  // glm compute_dot<vec2>
static inline float len(V a) { return std::sqrt(dot(a, a)); }                                 // glm::length
static inline V gnorm(V a) { return a * (1.0f / std::sqrt(dot(a, a))); }                      // glm::normalize
static inline V snorm(V a) { return zero(a) ? mk(0, 0) : gnorm(a); }                          // P:637-642

// P hpp:14-25
static const float G = 9.8 * 1000;
static const float MU = 0.0142435;
static const float MU_SLIDE = 12.0f * MU;
static const float LOSS_BALL = 1 - (float)0.05;
static const float TOPSPIN_REST = 0.25f;
static const float LOSS_RAIL = 1 - 0.38f;
static inline float accLen() { return -G * MU; }             // P:259-261
static inline float slideAccLen() { return -G * MU_SLIDE; }  // P:263-265

enum Kind { MOVING = 0, COLLISION = 1, RAIL = 2, POTTING = 3, SHOT = 4, REST = 5, ABSENT = 7 };  // type.hpp:167-180
enum Ev { EV_COLLISION = 0, EV_RAIL = 1, EV_POTTING = 2, EV_REST = 3, EV_ROLLING = 4, EV_NONE = 7 };  // type.hpp:87-98
enum : uint32_t { F_COMPLETED = 1u, F_VALID = 2u, F_OVERFLOW = 4u, F_NOEVENT = 8u };

struct Rail { V s, d, n; };          // shifted start, shifted end - shifted start, normal (T:56-72)
struct Pocket { V c; float r; };

struct Table {
    std::vector<Rail> rails;
    std::vector<Pocket> pockets;
    float diameter = 0;
    int nTypes = 0, cueType = 0;
    std::vector<double> score;
};
static Table g;

struct Counters {
    std::atomic<uint64_t> events{0}, pairTests{0}, quartics{0}, railTests{0}, pocketTests{0}, quadratics{0}, nearTies{0};
};
static Counters cnt;

// ---- physics ---------------------------------------------------------------------------------
// P:562-582 (the unqualified sqrt() there is ::sqrt(double))
static int quadratic(float a, float b, float c, float x[2]) {
    float disc = b * b - 4 * a * c;
    if (disc < 0) return 0;
    if (disc == 0) { x[0] = -0.5f * b / a; return 1; }
    float q = (b > 0) ? -0.5f * (b + ::sqrt((double)disc)) : -0.5f * (b - ::sqrt((double)disc));
    float x0 = q / a, x1 = c / q;
    if (x0 <= x1) { x[0] = x0; x[1] = x1; } else { x[0] = x1; x[1] = x0; }
    return 2;
}

// P:651-723
static bool lineLine(V P, V D, V Q, V Vv, float& t, float& u) {
    float detDV = D.x * Vv.y - D.y * Vv.x;
    if (detDV == 0.0f) return false;
    float detVP = Vv.x * P.y - Vv.y * P.x;
    float detQV = Q.x * Vv.y - Q.y * Vv.x;
    t = (detVP + detQV) / detDV;
    u = Vv.x != 0.0f ? (P.x + t * D.x - Q.x) / Vv.x : (P.y + t * D.y - Q.y) / Vv.y;
    return true;
}
// P:755-783
static bool halfHalf(V p1, V d1, V p2, V d2) {
    float t, u;
    if (!lineLine(p1, d1, p2, d2, t, u)) return false;
    return !(t < 0.0f) && !(u < 0.0f);
}
// P:181-238 + P:725-753
static bool railTime(V pos, V vel, V acc, const Rail& r, float& out) {
    if (zero(vel)) return false;
    float t, u;
    if (!lineLine(pos, vel, r.s, r.d, t, u)) return false;
    if (t < 0.0f) return false;
    if (u < 0.0f || u > 1.0f) return false;
    float distance = len(t * vel), velocity = len(vel), a = -len(acc);
    float x[2];
    if (!quadratic(0.5f * a, velocity, -distance, x)) return false;
    out = x[0];
    return true;
}
// P:383-457 + P:584-625
static bool pocketTime(const Pocket& pk, V pos, V vn, V vel, V acc, float& out) {
    if (zero(vel)) return false;
    V delta = pos - pk.c;
    float b = 2 * dot(vn, delta), c = dot(delta, delta) - pk.r * pk.r;
    float l[2];
    if (!quadratic(1, b, c, l)) return false;
    if (l[0] < 0.0f) return false;
    V distance = l[0] * vn;
    float x[2];
    if (!quadratic(-0.5f * len(acc), len(vel), -len(distance), x)) return false;
    if (x[0] < 0.0f) return false;
    out = x[0];
    return true;
}
// P:54-70
static float pointHalfLineSq(V lp, V ld, V pt) {
    V dn = gnorm(ld);
    V toPoint = pt - lp;
    float factor = dot(dn, toPoint) / len(ld);
    if (factor < 0) { V s = pt - lp; return dot(s, s); }
    V dv = toPoint - (ld * factor);
    return dot(dv, dv);
}

struct complex_order {  // P:459-468
    bool operator()(const std::complex<double>& l, const std::complex<double>& r) const {
        if (l.real() != r.real()) return l.real() < r.real();
        return l.imag() < r.imag();
    }
};

// P:470-548, written with std::complex exactly as the reference so libstdc++/glibc do the same work.
static std::vector<std::complex<double>> quartic(double a, double b, double c, double d, double e) {
    using C = std::complex<double>;
    if (a == 0.0) return {};
    double discriminant = (256*a*a*a*e*e*e) - (192*a*a*b*d*e*e) - (128*a*a*c*c*e*e) + (144*a*a*c*d*d*e)
            -(27*a*a*d*d*d*d) + (144*a*b*b*c*e*e) - (6*a*b*b*d*d*e) - (80*a*b*c*c*d*e) + (18*a*b*c*d*d*d)
            +(16*a*c*c*c*c*e) -(4*a*c*c*c*d*d) - (27*b*b*b*b*e*e) + (18*b*b*b*c*d*e)
            -(4*b*b*b*d*d*d) - (4*b*b*c*c*c*e) + (b*b*c*c*d*d);
    C p = ((8 * a * c) - (3 * b * b)) / (8 * a * a);
    C q = (std::pow(b, 3) - (4 * a * b * c) + (8 * a * a * d)) / (8 * std::pow(a, 3));
    C deltaZero = (c * c) - (3 * b * d) + (12 * a * e);
    C deltaOne = (2 * std::pow(c, 3)) - (9 * b * c * d) + (27 * b * b * e) + (27 * a * d * d) - (72 * a * c * e);
    C powerOfDeltaZero = (4.0 * std::pow(deltaZero, 3));
    C deltaOneSquared = (deltaOne * deltaOne);
    C QInnerSqrt = std::sqrt(deltaOneSquared - powerOfDeltaZero);
    C QInnerCbrt = (deltaOne + QInnerSqrt) / 2.0;
    C Q = std::pow(QInnerCbrt, (1.0 / 3.0));
    if (Q == 0.0 && discriminant != 0 && deltaZero == 0.0) {
        QInnerCbrt = (deltaOne - QInnerSqrt) / 2.0;
        Q = std::pow(QInnerCbrt, (1.0 / 3.0));
    }
    C S = 0.5 * std::sqrt((-(2.0 / 3.0) * p) + ((1.0 / (3.0 * a)) * (Q + (deltaZero / Q))));
    if (S == 0.0) {
        QInnerCbrt = (deltaOne - QInnerSqrt) / 2.0;
        Q = std::pow(QInnerCbrt, (1.0 / 3.0));
        S = 0.5 * std::sqrt((-(2.0 / 3.0) * p) + ((1.0 / (3.0 * a)) * (Q + (deltaZero / Q))));
    }
    C sh1 = (-4.0 * S * S) - (2.0 * p);
    C sh2 = q / S;
    C d1 = sh1 + sh2, d2 = sh1 - sh2;
    C fps = (-b / (4 * a));
    C f1 = fps - S, f2 = fps + S;
    C s1 = 0.5 * std::sqrt(d1), s2 = 0.5 * std::sqrt(d2);
    std::vector<C> sol{f1 + s1, f1 - s1, f2 + s2, f2 - s2};
    std::sort(sol.begin(), sol.end(), complex_order());
    return sol;
}

// P:283-381
static bool collisionTime(V a1, V v1, V p1, V a2, V v2, V p2, float diameter, float& out, bool* reachedQuartic = nullptr) {
    if (reachedQuartic) *reachedQuartic = false;
    if (zero(v1)) return false;
    float v2sq = dot(v2, v2);
    if (v2sq > 0) {
        if (!halfHalf(p1, v1, p2, v2)) {
            V from1To2 = p2 - p1;
            V v1n = gnorm(v1), v2n = gnorm(v2);
            V offset = from1To2 - v1n * dot(from1To2, v1n);
            V back1 = v1n * -diameter, back2 = v2n * -diameter;
            if (!halfHalf(p1 + offset + back1, v1, p2 + back2, v2)) {
                float velocityDistance = dot(v1n, v2n);
                float offsetSq = dot(offset, offset);
                if (velocityDistance != 1 && offsetSq > diameter * diameter) return false;
            }
        }
    } else {
        float distance = pointHalfLineSq(p1, v1, p2);
        if (distance > diameter * diameter) return false;
    }
    float toMeter = 1.0f / 1000.0f;
    float dm = diameter * toMeter;
    V dA = (a1 - a2) * toMeter, dV = (v1 - v2) * toMeter, dP = (p1 - p2) * toMeter;
    float a = floorf(0.25f * dot(dA, dA) * 1000000) / 1000000;
    float b = floorf(dot(dA, dV) * 1000000) / 1000000;
    float c = floorf((dot(dA, dP) + dot(dV, dV)) * 1000000) / 1000000;
    float d = floorf(2 * (dot(dV, dP)) * 1000000) / 1000000;
    float e = floorf((dot(dP, dP) - (dm * dm)) * 1000000) / 1000000;
    if (reachedQuartic) *reachedQuartic = true;
    auto sol = quartic(a, b, c, d, e);
    static const float REAL_ROOT_EPSILON = 0.001f;  // P:550-560
    for (auto& s : sol) {
        if (std::abs(s.imag()) < REAL_ROOT_EPSILON) {
            float first = static_cast<float>(s.real());
            if (first < 0.0f) return false;
            out = first;
            return true;
        }
    }
    return false;
}

// P:121-130, 132-166
static V variableTopspin(V velocity, V vt) {
    V vtn = vt + TOPSPIN_REST * velocity;
    return dot(vtn, vtn) < dot(velocity, velocity) ? vtn : len(velocity) * snorm(vtn);
}
static void elastic(V p1, V v1, V p2, V v2, bool r1, bool r2, V& o1, V& o2) {
    V z = gnorm(p2 - p1);
    V w1 = v1 * LOSS_BALL, w2 = v2 * LOSS_BALL;
    V v1z = dot(w1, z) * z, v2z = dot(w2, z) * z;
    V v1t = w1 - v1z, v2t = w2 - v2z;
    V v1tn = r1 ? variableTopspin(w1, v1t) : v1t;
    V v2tn = r2 ? variableTopspin(w2, v2t) : v2t;
    o1 = v2z + v1tn;
    o2 = v1z + v2tn;
}

// ---- simulation state ---------------------------------------------------------------------------
struct Ball {
    V pos, vel, vn, acc;
    bool rolling;
    int kind;
    bool hasNextRolling;
    float nextRolling;
    int lastRail;
    V velBefore;  // "before" velocity of the node (for traces)
    int cause;
};

static void derive(Ball& b) {  // T:176-184
    b.vn = snorm(b.vel);
    b.acc = (b.rolling ? accLen() : slideAccLen()) * b.vn;
}
static inline bool dynamicKind(int k) { return k == SHOT || k == MOVING || k == COLLISION || k == RAIL; }  // T:404-425

struct Event { int type = EV_NONE; float time = 0; int a = -1, b = -1; };

struct Sim {
    int n;
    std::vector<Ball> balls;
    float time = 0;
    int nEvents = 0;
    uint32_t flags = 0;
    bool nearTie = false;
    float minGap = INFINITY;  // smallest (second best - best) over all steps
};

// S:1550-1573 with the per-ball order rest, rolling, static pairs, dynamic pairs, rails, pockets.
// Balls are visited in index order (the reference visits them in unordered_map order; only exact
// ties can tell the difference).  `second` receives the runner-up time for tie detection.
static Event nextEvent(const Sim& s, float& second) {
    Event best;
    second = INFINITY;
    auto consider = [&](int type, float t, int a, int b) {
        if (t != t) return;  // NaN never wins a strict '<' unless first; the first per ball (rest) is never NaN
        if (best.type == EV_NONE) { best = Event{type, t, a, b}; return; }
        if (t < best.time) {
            second = best.time;
            best = Event{type, t, a, b};
        } else if (!(best.type == EV_COLLISION && type == EV_COLLISION &&
                     ((best.a == a && best.b == b) || (best.a == b && best.b == a)))) {
            if (t < second) second = t;
        }
    };
    for (int i = 0; i < s.n; ++i) {
        const Ball& bi = s.balls[i];
        if (!dynamicKind(bi.kind)) continue;
        // S:1590-1601 -> P:240-253
        {
            float tx = bi.acc.x == 0.0f ? 0.0f : (-bi.vel.x / bi.acc.x);
            float ty = bi.acc.y == 0.0f ? 0.0f : (-bi.vel.y / bi.acc.y);
            consider(EV_REST, std::max(tx, ty), i, -1);
        }
        if (bi.hasNextRolling) consider(EV_ROLLING, bi.nextRolling, i, -1);  // S:1575-1588
        for (int pass = 0; pass < 2; ++pass) {  // static partners first, then dynamic ones (S:1562-1564)
            for (int j = 0; j < s.n; ++j) {
                if (j == i) continue;
                const Ball& bj = s.balls[j];
                bool isStatic = bj.kind == REST;  // T:427-445: only BALL_IN_REST
                bool isDyn = dynamicKind(bj.kind);
                if ((pass == 0 && !isStatic) || (pass == 1 && !isDyn)) continue;
                float t;
                bool rq;
                cnt.pairTests++;
                bool has = collisionTime(bi.acc, bi.vel, bi.pos, bj.acc, bj.vel, bj.pos, g.diameter, t, &rq);
                if (rq) cnt.quartics++;
                if (has) consider(EV_COLLISION, t, i, j);
            }
        }
        for (size_t r = 0; r < g.rails.size(); ++r) {  // S:1639-1682
            if (bi.lastRail == int(r)) continue;
            float t;
            cnt.railTests++;
            if (railTime(bi.pos, bi.vel, bi.acc, g.rails[r], t)) consider(EV_RAIL, t, i, int(r));
        }
        for (size_t p = 0; p < g.pockets.size(); ++p) {  // S:1684-1714
            float t;
            cnt.pocketTests++;
            if (pocketTime(g.pockets[p], bi.pos, bi.vn, bi.vel, bi.acc, t)) consider(EV_POTTING, t, i, int(p));
        }
    }
    return best;
}

static inline V position(V acc, V vel, float dt, V pos) { return (0.5f * acc * dt * dt) + (vel * dt) + pos; }  // P:279-281
static inline V accelerate(V acc, V vel, float dt) { return acc * dt + vel; }                                  // P:275-277

// S:1957-1992 createLayer + T:467-531 System::append
static void applyEvent(Sim& s, const Event& ev) {
    float dt = ev.time;
    float newTime = s.time + dt;
    float deltaTime = newTime - s.time;  // T:475: layer._time - previous._time
    std::vector<Ball> prev = s.balls;
    for (int i = 0; i < s.n; ++i) {
        Ball& b = s.balls[i];
        const Ball& p = prev[i];
        if (p.kind == ABSENT) continue;
        bool affected = (i == ev.a) || (ev.type == EV_COLLISION && i == ev.b);
        if (!affected) {
            if (p.kind == REST || p.kind == POTTING) {
                // unchanged node, copied with its cause (S:1966-1967)
            } else {
                b.cause = -1;
                b.vel = accelerate(p.acc, p.vel, dt);
                b.pos = position(p.acc, p.vel, dt, p.pos);
                b.velBefore = b.vel;
                b.kind = MOVING;
                derive(b);
            }
        }
    }
    if (ev.type == EV_POTTING) {  // S:1716-1746
        Ball& b = s.balls[ev.a]; const Ball& p = prev[ev.a];
        b.pos = position(p.acc, p.vel, dt, p.pos);
        b.velBefore = accelerate(p.acc, p.vel, dt);
        b.vel = mk(0, 0); b.rolling = false; b.kind = POTTING; b.cause = ev.b; derive(b);
    } else if (ev.type == EV_RAIL) {  // S:1748-1786
        Ball& b = s.balls[ev.a]; const Ball& p = prev[ev.a];
        b.pos = position(p.acc, p.vel, dt, p.pos);
        V vb = accelerate(p.acc, p.vel, dt);
        V nrm = g.rails[ev.b].n;
        b.velBefore = vb;
        b.vel = (vb - nrm * dot(nrm, vb) * 2.0f) * LOSS_RAIL;  // P:168-175, glm::reflect
        b.rolling = true; b.kind = RAIL; b.cause = ev.b; derive(b);
    } else if (ev.type == EV_REST) {  // S:1788-1808
        Ball& b = s.balls[ev.a]; const Ball& p = prev[ev.a];
        b.pos = position(p.acc, p.vel, dt, p.pos);
        b.vel = mk(0, 0); b.velBefore = b.vel; b.rolling = false; b.kind = REST; b.cause = -1; derive(b);
    } else if (ev.type == EV_ROLLING) {  // S:1810-1840
        Ball& b = s.balls[ev.a]; const Ball& p = prev[ev.a];
        b.pos = position(p.acc, p.vel, dt, p.pos);
        b.vel = accelerate(p.acc, p.vel, dt);
        b.velBefore = b.vel; b.rolling = true; b.kind = MOVING; b.cause = -1; derive(b);
    } else if (ev.type == EV_COLLISION) {  // S:1842-1905
        Ball& b1 = s.balls[ev.a]; const Ball& p1 = prev[ev.a];
        Ball& b2 = s.balls[ev.b]; const Ball& p2 = prev[ev.b];
        V pos1 = position(p1.acc, p1.vel, dt, p1.pos), vb1 = accelerate(p1.acc, p1.vel, dt);
        V pos2 = position(p2.acc, p2.vel, dt, p2.pos), vb2 = accelerate(p2.acc, p2.vel, dt);
        V o1, o2;
        elastic(pos1, vb1, pos2, vb2, p1.rolling, p2.rolling, o1, o2);
        b1.pos = pos1; b1.velBefore = vb1; b1.vel = o1; b1.kind = COLLISION; b1.cause = ev.b; derive(b1);
        b2.pos = pos2; b2.velBefore = vb2; b2.vel = o2; b2.kind = COLLISION; b2.cause = ev.a; derive(b2);
    }
    // System::append bookkeeping, T:478-527
    for (int i = 0; i < s.n; ++i) {
        Ball& b = s.balls[i];
        if (b.kind == ABSENT) continue;
        if (b.hasNextRolling) b.nextRolling = std::max(0.0f, b.nextRolling - deltaTime);
        if (b.kind == COLLISION) {
            b.lastRail = -1;
            if (!b.rolling && !zero(b.vel)) { b.hasNextRolling = true; b.nextRolling = len(b.vel) / ((3.5f) * G * MU_SLIDE); }
            else b.hasNextRolling = false;
        } else if (b.kind == SHOT) {
            b.hasNextRolling = true; b.nextRolling = len(b.vel) / ((3.5f) * G * MU_SLIDE);
        } else if (b.kind == REST || b.kind == POTTING) {
            b.hasNextRolling = false;
        } else if (b.kind == MOVING) {
            if (b.rolling) b.hasNextRolling = false;
        }
        if (b.kind == RAIL) { b.lastRail = b.cause; b.hasNextRolling = false; }
    }
    s.time = newTime;
}

static bool isFinal(const Sim& s) {  // T:392-396
    for (auto& b : s.balls) if (b.kind != ABSENT && b.kind != REST && b.kind != POTTING) return false;
    return true;
}

static Sim initSim(int n, const float* pos, const int32_t* types, float vx, float vy) {  // S:760-795
    Sim s;
    s.n = n;
    s.balls.resize(n);
    for (int i = 0; i < n; ++i) {
        Ball& b = s.balls[i];
        b = Ball{};
        b.lastRail = -1; b.cause = -1;
        if (types[i] < 0) { b.kind = ABSENT; continue; }
        b.pos = mk(pos[2 * i], pos[2 * i + 1]);
        b.rolling = false;
        if (types[i] == g.cueType) {
            b.vel = mk(vx, vy); b.kind = SHOT;
            b.hasNextRolling = true; b.nextRolling = len(b.vel) / ((3.5f) * G * MU_SLIDE);  // T:502-506
        } else {
            b.vel = mk(0, 0); b.kind = REST;
        }
        b.velBefore = b.vel;
        derive(b);
    }
    return s;
}

typedef void (*LayerSink)(void* ctx, int layer, const Sim& s, const Event& ev);

// S:1358-1471 event loop.  tieEps: near-tie threshold in seconds (north_star: 1e-9).
static void run(Sim& s, float tieEps, LayerSink sink, void* ctx) {
    if (sink) sink(ctx, 0, s, Event{});
    int eventCount = 0;
    while (!isFinal(s)) {
        float second;
        Event ev = nextEvent(s, second);
        if (ev.type == EV_NONE) { s.flags |= F_NOEVENT; return; }
        eventCount++;
        if (eventCount > 30) { s.flags |= F_OVERFLOW; return; }
        float gap = second - ev.time;
        if (gap < s.minGap) s.minGap = gap;
        if (gap < tieEps) { s.nearTie = true; cnt.nearTies++; }
        applyEvent(s, ev);
        s.nEvents = eventCount;
        cnt.events++;
        if (sink) sink(ctx, eventCount, s, ev);
    }
    s.flags |= F_COMPLETED;
}

// S:1473-1532 + billiard_snooker.cpp:1128-1153
static uint64_t simulationCost(const Sim& s, const int32_t* types, float vx, float vy, uint64_t searchCost) {
    double pottingScore = 0.0;
    for (int i = 0; i < s.n; ++i)
        if (s.balls[i].kind == POTTING) pottingScore += g.score[types[i]];
    V v = mk(vx, vy);
    float velocitySquared = dot(v, v);
    double velocityCost = velocitySquared / (5000.0f * 5000.0f);
    double pottingCost = 1.0 - std::min(pottingScore, 1.0);
    double totalCost = std::min(0.5 * pottingCost + 0.5 * velocityCost, 1.0);
    double cap = 1.0 * (double)searchCost;
    return (uint64_t)(totalCost * cap);
}
static bool validEnd(const Sim& s, const int32_t* types, uint32_t expectedMask) {
    int potted = 0;
    for (int i = 0; i < s.n; ++i) {
        if (s.balls[i].kind != POTTING) continue;
        if (types[i] == g.cueType) return false;
        if (!(expectedMask >> types[i] & 1u)) return false;
        potted++;
    }
    return potted != 0;
}

}  // namespace port

using namespace port;

extern "C" {

// railsDerived: R x {shiftedStart.xy, shiftedEnd.xy, normal.xy} as computed by the table builder
// (billiard_ai_b200 host code or ref_get_rail_derived); pockets3: P x {x, y, radius}
int port_set_table(int R, const float* railsDerived6, int P, const float* pockets3, float diameter) {
    g.rails.clear(); g.pockets.clear();
    for (int i = 0; i < R; ++i) {
        const float* r = railsDerived6 + 6 * i;
        Rail rl;
        rl.s = mk(r[0], r[1]);
        rl.d = mk(r[2], r[3]) - mk(r[0], r[1]);  // P:731: lineSegmentEndPoint - lineSegmentStartPoint
        rl.n = mk(r[4], r[5]);
        g.rails.push_back(rl);
    }
    for (int i = 0; i < P; ++i) g.pockets.push_back(Pocket{mk(pockets3[3 * i], pockets3[3 * i + 1]), pockets3[3 * i + 2]});
    g.diameter = diameter;
    return 0;
}
int port_set_rules(int nTypes, int cueType, const double* scores) {
    g.nTypes = nTypes; g.cueType = cueType; g.score.assign(scores, scores + nTypes);
    return 0;
}

struct TraceCtx { int n, maxLayers; float* times; float* rec; int32_t* events; };
static void traceSink(void* c, int layer, const Sim& s, const Event& ev) {
    TraceCtx* t = (TraceCtx*)c;
    if (layer >= t->maxLayers) return;
    t->times[layer] = s.time;
    for (int b = 0; b < t->n; ++b) {
        float* r = t->rec + (size_t(layer) * t->n + b) * 10;
        const Ball& B = s.balls[b];
        for (int k = 0; k < 10; ++k) r[k] = -1.0f;
        if (B.kind == ABSENT) continue;
        r[0] = float(B.kind); r[1] = B.pos.x; r[2] = B.pos.y; r[3] = B.vel.x; r[4] = B.vel.y;
        r[5] = B.rolling ? 1.0f : 0.0f; r[6] = B.velBefore.x; r[7] = B.velBefore.y; r[8] = float(B.cause);
        r[9] = -1.0f;  // before.isRolling is not tracked by the port
    }
    if (t->events && layer > 0) {
        int32_t* e = t->events + (layer - 1) * 4;
        int32_t tb; float tm = s.time; std::memcpy(&tb, &tm, 4);
        e[0] = ev.type; e[1] = ev.a; e[2] = ev.b; e[3] = tb;
    }
}

// Same record layout as ref_simulate_trace (oracle/ref_driver.cpp); record slot [9] is -1 here.
// Returns the layer count, or 0 when the reference's public simulate() would return nullopt.
int port_simulate_trace(int n, const float* pos, const int32_t* types, float vx, float vy, int maxLayers, float* outTimes,
                        float* outRec, float* outMinGap) {
    Sim s = initSim(n, pos, types, vx, vy);
    TraceCtx t{n, maxLayers, outTimes, outRec, nullptr};
    run(s, 1e-9f, traceSink, &t);
    if (outMinGap) *outMinGap = s.minGap;
    return (s.flags & F_COMPLETED) ? s.nEvents + 1 : 0;
}

int port_evaluate(int nStates, int n, const float* statePos, const int32_t* stateTypes, const uint32_t* expectedMask,
                  int64_t nCand, const int32_t* candState, const float* candVx, const float* candVy,
                  const uint64_t* candSearchCost, const uint64_t* candBaseCost, int nThreads, uint64_t* outCost,
                  uint32_t* outFlags, int32_t* outNEvents, uint32_t* outPotted, float* outFinalTime, float* outFinalPos,
                  int32_t* outEventLog, float* outMinGap) {
    std::atomic<int64_t> next{0};
    auto worker = [&]() {
        for (;;) {
            int64_t c0 = next.fetch_add(256);
            if (c0 >= nCand) break;
            for (int64_t c = c0; c < std::min(nCand, c0 + 256); ++c) {
                int si = candState[c];
                const float* pos = statePos + size_t(si) * n * 2;
                const int32_t* types = stateTypes + size_t(si) * n;
                Sim s = initSim(n, pos, types, candVx[c], candVy[c]);
                int32_t* ev = nullptr;
                if (outEventLog) {
                    ev = outEventLog + size_t(c) * 32 * 4;
                    std::memset(ev, 0xff, 32 * 4 * sizeof(int32_t));
                }
                struct Ctx2 { int32_t* ev; } c2{ev};
                run(s, 1e-9f,
                    ev ? +[](void* cx, int layer, const Sim& ss, const Event& e) {
                        int32_t* evp = ((Ctx2*)cx)->ev;
                        if (layer == 0 || layer > 32) return;
                        int32_t tb; float tm = ss.time; std::memcpy(&tb, &tm, 4);
                        int a = e.a, b = e.b;
                        if (e.type == EV_COLLISION && a > b) std::swap(a, b);
                        evp[(layer - 1) * 4 + 0] = e.type; evp[(layer - 1) * 4 + 1] = a;
                        evp[(layer - 1) * 4 + 2] = b; evp[(layer - 1) * 4 + 3] = tb;
                    } : (LayerSink) nullptr,
                    &c2);
                uint32_t flags = s.flags;
                if (flags & F_COMPLETED) {
                    if (validEnd(s, types, expectedMask[si])) flags |= F_VALID;
                    outCost[c] = candBaseCost[c] + simulationCost(s, types, candVx[c], candVy[c], candSearchCost[c]);
                } else {
                    outCost[c] = ~uint64_t(0);
                }
                outFlags[c] = flags;
                outNEvents[c] = s.nEvents;
                uint32_t potted = 0;
                for (int i = 0; i < n; ++i) {
                    float* fp = outFinalPos + (size_t(c) * n + i) * 2;
                    fp[0] = fp[1] = 0.0f;
                    if (s.balls[i].kind == ABSENT) continue;
                    if (s.balls[i].kind == POTTING) potted |= 1u << i;
                    fp[0] = s.balls[i].pos.x; fp[1] = s.balls[i].pos.y;
                }
                outPotted[c] = potted;
                outFinalTime[c] = s.time;
                if (outMinGap) outMinGap[c] = s.minGap;
            }
        }
    };
    if (nThreads <= 1) worker();
    else {
        std::vector<std::thread> th;
        for (int t = 0; t < nThreads; ++t) th.emplace_back(worker);
        for (auto& t : th) t.join();
    }
    return 0;
}

int port_quartic(double a, double b, double c, double d, double e, double* outReIm8) {
    auto r = quartic(a, b, c, d, e);
    for (size_t i = 0; i < r.size() && i < 4; ++i) { outReIm8[2 * i] = r[i].real(); outReIm8[2 * i + 1] = r[i].imag(); }
    return int(r.size());
}

int port_time_to_collision_batch(int64_t n, const float* p1, const float* v1, const int32_t* roll1, const float* p2,
                                 const float* v2, const int32_t* roll2, float diameter, float* outTime, int32_t* outHas) {
    for (int64_t i = 0; i < n; ++i) {
        Ball a{}, b{};
        a.pos = mk(p1[2 * i], p1[2 * i + 1]); a.vel = mk(v1[2 * i], v1[2 * i + 1]); a.rolling = roll1[i] != 0; derive(a);
        b.pos = mk(p2[2 * i], p2[2 * i + 1]); b.vel = mk(v2[2 * i], v2[2 * i + 1]); b.rolling = roll2[i] != 0; derive(b);
        float t = -1.0f;
        bool has = collisionTime(a.acc, a.vel, a.pos, b.acc, b.vel, b.pos, diameter, t);
        outHas[i] = has ? 1 : 0;
        outTime[i] = has ? t : -1.0f;
    }
    return 0;
}

// counters: events, pairTests, quartics, railTests, pocketTests, nearTies
void port_get_counters(uint64_t* out6, int reset) {
    out6[0] = cnt.events; out6[1] = cnt.pairTests; out6[2] = cnt.quartics; out6[3] = cnt.railTests;
    out6[4] = cnt.pocketTests; out6[5] = cnt.nearTies;
    if (reset) { cnt.events = 0; cnt.pairTests = 0; cnt.quartics = 0; cnt.railTests = 0; cnt.pocketTests = 0; cnt.nearTies = 0; }
}

}  // extern "C"
