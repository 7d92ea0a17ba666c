// TEST / ANALYSIS INFRASTRUCTURE ONLY (never part of the product path).
// Statistics for the quartic-pruning bound, gathered on the CPU restatement of the reference (oracle/port):
// how many ball-pair quartics a travel-distance lower bound removes, and whether any removed job would have won.
#include "../oracle/port/port_oracle.cpp"

namespace study {
struct Stats {
    uint64_t steps = 0, jobs = 0, jobsStatic = 0, jobsDyn = 0, jobsHit = 0;
    uint64_t prunedNP = 0, prunedPartial = 0;            // single pass: pruned against the non-pair minimum
    uint64_t pass1 = 0, pass2 = 0;    // two passes: nearest-bound job per strike first
    uint64_t pass1b = 0, pass2b = 0;  // two passes: nearest-bound job per dynamic ball first
    uint64_t violNP = 0, viol2 = 0, viol2b = 0;
    double maxRho = -1e30;            // max over hits of rho(t_hit): must stay <= 0
    uint64_t winnerIsPair = 0;
};
struct Job { int i, j; float L, vrel, a, vi, vj, ai, aj, sin2; bool has; float t; bool isStatic; double guard; int beyond; };

static inline float rho(const Job& jb, float T, float d) {
    T = T + 2e-3f;
    const float poly = (((T + 1.0f) * T + 1.0f) * T + 1.0f) * T + 1.0f;
    const float di = std::max(jb.vi * T, 0.5f * jb.ai * T * T), dj = std::max(jb.vj * T, 0.5f * jb.aj * T * T);
    return jb.L - d - di - dj - 1e-3f * (jb.vi + jb.vj + (jb.ai + jb.aj) * T) - 2.0f - 0.02f * poly;
}
static inline bool eligible(const Job& jb, float d) {
    if (!jb.isStatic) return jb.a >= 4e-3f && jb.sin2 >= 1e-2f && jb.vi > 0 && jb.vj > 0;
    // resting partner: only when the moving ball's polynomial reaches past the partner before its turning point (four
    // real roots); otherwise the depressed quartic is biquadratic up to quantisation noise and the closed form is noise
    return jb.L + d + 5.0f <= 0.99f * (jb.vi * jb.vi) / (2.0f * jb.ai);
}
static inline bool prune(const Job& jb, float T, float d) {
    if (!eligible(jb, d)) return false;
    return rho(jb, T, d) > 0.0f;
}

static Event nextEventStudy(const Sim& s, Stats& st) {
    float second;
    Event best = nextEvent(s, second);
    if (best.type == EV_NONE) return best;
    float Tnp = INFINITY, Trr = INFINITY;
    std::vector<float> Town(s.n, INFINITY);
    std::vector<Job> jobs;
    for (int i = 0; i < s.n; ++i) {
        const Ball& bi = s.balls[i];
        if (!dynamicKind(bi.kind)) continue;
        float tx = bi.acc.x == 0.0f ? 0.0f : (-bi.vel.x / bi.acc.x);
        float ty = bi.acc.y == 0.0f ? 0.0f : (-bi.vel.y / bi.acc.y);
        Tnp = std::min(Tnp, std::max(tx, ty)); Trr = std::min(Trr, std::max(tx, ty));
        if (bi.hasNextRolling) { Tnp = std::min(Tnp, bi.nextRolling); Trr = std::min(Trr, bi.nextRolling); }
        for (size_t r = 0; r < g.rails.size(); ++r) {
            if (bi.lastRail == int(r)) continue;
            float t;
            if (railTime(bi.pos, bi.vel, bi.acc, g.rails[r], t)) { Tnp = std::min(Tnp, t); Town[i] = std::min(Town[i], t); }
        }
        for (size_t p = 0; p < g.pockets.size(); ++p) {
            float t;
            if (pocketTime(g.pockets[p], bi.pos, bi.vn, bi.vel, bi.acc, t)) { Tnp = std::min(Tnp, t); Town[i] = std::min(Town[i], t); }
        }
        for (int j = 0; j < s.n; ++j) {
            if (j == i) continue;
            const Ball& bj = s.balls[j];
            bool isStatic = bj.kind == REST, isDyn = dynamicKind(bj.kind);
            if (!isStatic && !isDyn) continue;
            if (isDyn && j < i) continue;  // one job per dynamic pair (the pipeline posts i < j once)
            float t = 0, t2 = 0; bool rq = false, rq2 = false, has = false;
            has = collisionTime(bi.acc, bi.vel, bi.pos, bj.acc, bj.vel, bj.pos, g.diameter, t, &rq);
            if (isDyn) {
                bool has2 = collisionTime(bj.acc, bj.vel, bj.pos, bi.acc, bi.vel, bi.pos, g.diameter, t2, &rq2);
                if (!rq && rq2) { has = has2; t = t2; }
                rq = rq || rq2;
            }
            if (!rq) continue;
            Job jb;
            jb.i = i; jb.j = j; jb.isStatic = isStatic; jb.has = has; jb.t = t;
            jb.L = len(bj.pos - bi.pos);
            jb.vrel = len(bi.vel) + len(bj.vel); jb.vi = len(bi.vel); jb.vj = len(bj.vel); jb.ai = len(bi.acc); jb.aj = len(bj.acc);
            V dA = (bi.acc - bj.acc) * (1.0f / 1000.0f);
            jb.a = 0.25f * dot(dA, dA);
            { V dV = (bi.vel - bj.vel) * (1.0f / 1000.0f); float cr = dA.x * dV.y - dA.y * dV.x; jb.sin2 = cr * cr / (dot(dA, dA) * dot(dV, dV)); }
            jb.guard = -1; jb.beyond = 0;
            if (isStatic) {   // the kernel's beyond_turning_point() quantities
                V tp = bj.pos - bi.pos; V dn = bi.vn;
                float along = dot(dn, tp);
                float factor = along / len(bi.vel);
                V dv = tp - bi.vel * factor;
                float perpSq = dot(dv, dv), dd = g.diameter * g.diameter;
                float hSq = dd - perpSq;
                float stop = (jb.vi * jb.vi) / (2.0f * jb.ai);
                if (factor >= 0 && hSq >= 25.0f && along - std::sqrt(hSq) > 1.05f * stop + 5.0f) {
                    jb.beyond = 1;
                    float toMeter = 1.0f / 1000.0f, dm = g.diameter * toMeter;
                    V dA = (bi.acc - bj.acc) * toMeter, dV = (bi.vel - bj.vel) * toMeter, dP = (bi.pos - bj.pos) * toMeter;
                    double a = floorf(0.25f * dot(dA, dA) * 1000000) / 1000000, b = floorf(dot(dA, dV) * 1000000) / 1000000;
                    double c = floorf((dot(dA, dP) + dot(dV, dV)) * 1000000) / 1000000, d = floorf(2 * (dot(dV, dP)) * 1000000) / 1000000;
                    double pp = (8 * a * c - 3 * b * b) / (8 * a * a), qq = (b * b * b - 4 * a * b * c + 8 * a * a * d) / (8 * a * a * a);
                    double S = std::fabs(qq) * jb.ai / (8.0 * std::sqrt(hSq));
                    jb.guard = (pp > 0) ? S * S / (1e-12 * (1.0 + pp)) : -2;
                }
            }
            jobs.push_back(jb);
        }
    }
    const float d = g.diameter;
    st.steps++;
    st.jobs += jobs.size();
    if (best.type == EV_COLLISION) st.winnerIsPair++;
    for (auto& jb : jobs) {
        if (jb.beyond && getenv("STUDY_BEYOND")) {
            static uint64_t nB = 0, nHit = 0, nGuardOk = 0; static double minHitFree = 1e300;
            nB++; if (jb.guard >= 1.0) nGuardOk++;
            if (jb.has) { nHit++; fprintf(stderr, "BEYOND-HIT guard %.3g t %.6g vi %g ai %g L %g a %g\n", jb.guard, jb.t, jb.vi, jb.ai, jb.L, jb.a); }
            if (nB % 200000 == 0) fprintf(stderr, "beyond %llu guard-ok %llu hits %llu\n", (unsigned long long)nB, (unsigned long long)nGuardOk, (unsigned long long)nHit);
        }
    }
    // single pass
    for (auto& jb : jobs) {
        (jb.isStatic ? st.jobsStatic : st.jobsDyn)++;
        if (jb.has && eligible(jb, d) && rho(jb, jb.t, d) > -1.5f && getenv("STUDY_VERBOSE")) {
            const Ball& bi = s.balls[jb.i]; const Ball& bj = s.balls[jb.j];
            fprintf(stderr, "rho %.3f t %.6g L %.3f vrel %.3f a %.6g static %d  i(%g,%g v %g,%g acc %g,%g) j(%g,%g v %g,%g acc %g,%g) Tnp %g\n",
                    rho(jb, jb.t, d), jb.t, jb.L, jb.vrel, jb.a, (int)jb.isStatic, bi.pos.x, bi.pos.y, bi.vel.x, bi.vel.y, bi.acc.x, bi.acc.y,
                    bj.pos.x, bj.pos.y, bj.vel.x, bj.vel.y, bj.acc.x, bj.acc.y, Tnp);
        }
        if (jb.has) { st.jobsHit++; st.maxRho = std::max(st.maxRho, (double)(eligible(jb, d) ? rho(jb, jb.t, d) : -1e30)); }
        if (prune(jb, std::min(Trr, Town[jb.i]), d)) st.prunedPartial++;
        if (getenv("STUDY_BREAKDOWN")) {
            static uint64_t c[8] = {0};
            static uint64_t nn = 0;
            const float Tp = std::min(Trr, Town[jb.i]);
            const bool boundOk = rho(jb, Tp, d) > 0;
            if (jb.isStatic) { c[0]++; if (boundOk) c[1]++; if (boundOk && eligible(jb, d)) c[2]++; }
            else { c[3]++; if (boundOk) c[4]++; if (boundOk && jb.a >= 4e-3f) c[5]++; if (boundOk && jb.sin2 >= 1e-2f) c[6]++; if (boundOk && eligible(jb, d)) c[7]++; }
            if (++nn % 100000 == 0) fprintf(stderr, "static %llu bound %llu elig %llu | dyn %llu bound %llu a-ok %llu sin-ok %llu elig %llu\n",
                (unsigned long long)c[0], (unsigned long long)c[1], (unsigned long long)c[2], (unsigned long long)c[3], (unsigned long long)c[4], (unsigned long long)c[5], (unsigned long long)c[6], (unsigned long long)c[7]);
        }
        if (prune(jb, Tnp, d)) { st.prunedNP++; if (jb.has && jb.t < Tnp) st.violNP++; }
    }
    // two passes, per strike
    if (!jobs.empty()) {
        size_t k0 = 0;
        auto ts = [&](const Job& jb) { return (jb.L - d) / jb.vrel; };
        for (size_t k = 1; k < jobs.size(); ++k) if (ts(jobs[k]) < ts(jobs[k0])) k0 = k;
        float T1 = Tnp;
        st.pass1++;   // only if not pruned by Tnp, actually
        if (jobs[k0].has) T1 = std::min(T1, jobs[k0].t);
        for (size_t k = 0; k < jobs.size(); ++k) {
            if (k == k0) continue;
            if (prune(jobs[k], T1, d)) { if (jobs[k].has && jobs[k].t < T1) st.viol2++; }
            else st.pass2++;
        }
        // per dynamic ball
        std::vector<int> firstOf(s.n, -1);
        for (size_t k = 0; k < jobs.size(); ++k) {
            int i = jobs[k].i;
            if (firstOf[i] < 0 || ts(jobs[k]) < ts(jobs[firstOf[i]])) firstOf[i] = (int)k;
        }
        float T1b = Tnp;
        std::vector<char> inP1(jobs.size(), 0);
        for (int i = 0; i < s.n; ++i) if (firstOf[i] >= 0) {
            inP1[firstOf[i]] = 1; st.pass1b++;
            if (jobs[firstOf[i]].has) T1b = std::min(T1b, jobs[firstOf[i]].t);
        }
        for (size_t k = 0; k < jobs.size(); ++k) {
            if (inP1[k]) continue;
            if (prune(jobs[k], T1b, d)) { if (jobs[k].has && jobs[k].t < T1b) st.viol2b++; }
            else st.pass2b++;
        }
    }
    return best;
}
}  // namespace study

extern "C" int study_run(int n, const float* pos, const int32_t* types, int64_t nCand, const float* vx, const float* vy,
                         double* out) {
    study::Stats st;
    for (int64_t c = 0; c < nCand; ++c) {
        Sim s = initSim(n, pos, types, vx[c], vy[c]);
        int eventCount = 0;
        while (!isFinal(s)) {
            Event ev = study::nextEventStudy(s, st);
            if (ev.type == EV_NONE) break;
            if (++eventCount > 30) break;
            applyEvent(s, ev);
        }
    }
    double v[] = {(double)st.steps, (double)st.jobs, (double)st.jobsStatic, (double)st.jobsDyn, (double)st.jobsHit,
                  (double)st.prunedNP, (double)st.pass1, (double)st.pass2, (double)st.pass1b, (double)st.pass2b,
                  (double)st.violNP, (double)st.viol2, (double)st.viol2b, st.maxRho, (double)st.winnerIsPair, (double)st.prunedPartial};
    for (int k = 0; k < 16; ++k) out[k] = v[k];
    return 0;
}
