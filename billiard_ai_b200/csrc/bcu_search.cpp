// bcu_search_levels: the level-synchronous multi-break search as ONE call of the C ABI (include/billiard_cuda.h).
//
// The reference searches break after break best-first (lib/billiard_search/src/billiard_search.cpp:74-98; expand :224-244):
// a finished, valid simulation either is a solution or -- depth search on, fewer than BREAKS simulations on its path --
// seeds the next break: the balls at rest form the next State, `_rules._nextSearch` names the types to pot next and
// `_rules._modifyState` puts potted colours back (:1407-1462); solutions leave the manager in ascending cost order
// (lib/process/include/process/process.hpp:19-25, 188-198).  Here the same step is taken for a whole beam at once:
//
//   level L:  beam states x (directions x speeds)  --bcu_evaluate_grid-->  cost = state's accumulated cost + simulationCost,
//             validity against the state's searched types  --bcu_level_select: device top-k on (cost, id), merged across
//             the ranks of the engine's communicator-->  the k best strikes with their end states  --rules, one host pass
//             over the k records-->  level L + 1 states / solutions
//
// Host C++ on top of the engine's own entry points: no device code here, and nothing of it runs per candidate.  The rules
// between breaks are data (`bcu_break_rules`): the reference's std::function rules (type.hpp:324-332) cannot cross a C
// ABI, so the rule sets the path needs are named -- the Configuration's default (single break), "repeat" (pool layouts),
// snooker (billiard_snooker.cpp:792-822, 1097-1126 through bcu_snooker_*).  The C++ mirror (billiard_ai_b200/host/)
// calls the caller's own std::function rules instead.
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../include/billiard_cuda.h"

namespace {

struct BeamState {
    std::vector<float> pos;         // [nBalls][2]
    std::vector<int32_t> types;     // [nBalls], -1 = no ball
    std::vector<int32_t> searched;  // Search::_types of the break played on this state
    uint64_t cost = 0;              // accumulated cost of the path that led here
    int parent = -1;                // state of the previous level the strike was played on
    int dir = -1, speed = -1;       // that strike
};

struct Found {
    uint64_t cost;
    int breaks;
    int state;           // state of level breaks - 1 the last strike was played on
    int dir, speed;      // the last strike
    std::vector<float> endPos;
    std::vector<uint8_t> endPresent;
};

uint32_t mask_of(const std::vector<int32_t>& typeCodes) {
    uint32_t m = 0;
    for (int32_t t : typeCodes) if (t >= 0 && t < 32) m |= 1u << t;
    return m;
}

double ms_since(std::chrono::steady_clock::time_point t0) {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

}  // namespace

extern "C" int bcu_engine_rank_info(const bcu_engine* e, int32_t* outRank, int32_t* outRanks);   // bcu_api.cu

extern "C" int bcu_search_levels(bcu_engine* e, int32_t nBalls, const float* pos, const int32_t* types,
                                 const bcu_search_params* sp, const bcu_break_rules* rules, bcu_search_result* out) {
    if (!e || !pos || !types || !sp || !rules || !out) return BCU_ERR_INVALID;
    if (nBalls < 1 || nBalls > BCU_MAX_BALLS || sp->levels < 1 || sp->levels > BCU_MAX_SEARCH_LEVELS || sp->beam < 1 || sp->beam > 1024 ||
        sp->n_dirs < 0 || (sp->n_dirs > 0 && (sp->n_speeds < 1 || !sp->dirs || !sp->speeds)) || sp->n_search_types < 0 ||
        (sp->n_search_types > 0 && !sp->search_types) || out->capacity < 0 ||
        (out->capacity > 0 && (!out->cost || !out->breaks)) ||
        (rules->kind != BCU_BREAK_RULES_SINGLE && rules->kind != BCU_BREAK_RULES_REPEAT && rules->kind != BCU_BREAK_RULES_SNOOKER))
        return BCU_ERR_INVALID;
    const auto t0 = std::chrono::steady_clock::now();
    int32_t rank = 0, world = 1;
    bcu_engine_rank_info(e, &rank, &world);
    const int NB = nBalls, L = sp->levels, K = sp->beam;
    // candidates of a state: the direction x speed grid, or -- n_dirs == 0 -- the strikes the reference's planner finds for it
    // (bcu_set_planner; every rank plans the whole beam -- the plan is deterministic -- and simulates its contiguous share of it)
    const bool planned = sp->n_dirs == 0;
    const int64_t perState = planned ? 0 : (int64_t)sp->n_dirs * sp->n_speeds;
    std::vector<int32_t> planState;
    std::vector<float> planVx, planVy;

    std::vector<std::vector<BeamState>> levels(1);
    {
        BeamState root;
        root.pos.assign(pos, pos + (size_t)NB * 2);
        root.types.assign(types, types + NB);
        root.searched.assign(sp->search_types, sp->search_types + sp->n_search_types);
        levels[0].push_back(std::move(root));
    }
    std::vector<Found> found;
    std::vector<std::pair<float, float>> planVel;   // velocities of the selected planned strikes, all levels
    std::vector<float> sPos, bTime, bPos;
    std::vector<int32_t> sTypes;
    std::vector<uint32_t> sMask, bPotted;
    std::vector<uint64_t> sCost, bCost, bId;
    bCost.resize(K); bId.resize(K); bPotted.resize(K); bTime.resize(K); bPos.resize((size_t)K * NB * 2);
    out->count = 0; out->total = 0; out->levels_run = 0; out->stopped = BCU_SEARCH_EXHAUSTED;

    for (int level = 0; level < L; ++level) {
        std::vector<BeamState>& states = levels[level];
        if (states.empty()) break;
        // the reference's manager ends a search when enough solutions are final, when its time is up, or when a newer request
        // replaces it (process.hpp:188-198, 82-83); all three are looked at between levels
        if (sp->cancel && *sp->cancel) { out->stopped = BCU_SEARCH_CANCELLED; break; }
        if (sp->min_solutions > 0 && (int64_t)found.size() >= sp->min_solutions) { out->stopped = BCU_SEARCH_ENOUGH; break; }
        if (sp->max_duration_ms > 0 && ms_since(t0) >= (double)sp->max_duration_ms) { out->stopped = BCU_SEARCH_TIMEOUT; break; }
        const auto tLevel = std::chrono::steady_clock::now();
        const int nS = (int)states.size();
        sPos.resize((size_t)nS * NB * 2); sTypes.resize((size_t)nS * NB); sMask.resize(nS); sCost.resize(nS);
        for (int i = 0; i < nS; ++i) {
            std::memcpy(&sPos[(size_t)i * NB * 2], states[i].pos.data(), (size_t)NB * 2 * sizeof(float));
            std::memcpy(&sTypes[(size_t)i * NB], states[i].types.data(), (size_t)NB * sizeof(int32_t));
            sMask[i] = mask_of(states[i].searched);
            sCost[i] = states[i].cost;
        }
        int rc = bcu_set_states(e, nS, NB, sPos.data(), sTypes.data(), sMask.data(), sCost.data());
        if (rc) return rc;
        // this rank's contiguous share of the level's candidate ids; the merged beam is identical on every rank
        int64_t total = (int64_t)nS * perState;
        if (planned) {
            rc = bcu_plan(e, &total);
            if (rc) return rc;
            planState.resize((size_t)total); planVx.resize((size_t)total); planVy.resize((size_t)total);
            rc = bcu_fetch_plan(e, 0, total, planState.data(), planVx.data(), planVy.data(), nullptr, nullptr);
            if (rc) return rc;
            const int64_t per = (total + world - 1) / world;
            const int64_t lo = std::min(total, (int64_t)rank * per), hi = std::min(total, lo + per);
            rc = total > 0 ? bcu_evaluate_plan_range(e, lo, hi - lo, nullptr) : BCU_OK;
            if (rc) return rc;
        } else {
            const int64_t per = (total + world - 1) / world;
            const int64_t lo = std::min(total, (int64_t)rank * per), hi = std::min(total, lo + per);
            rc = bcu_evaluate_grid(e, sp->n_dirs, sp->n_speeds, sp->dirs, sp->speeds, sp->search_cost, lo, hi - lo);
            if (rc) return rc;
        }
        float kernelMs = 0.0f;
        int32_t nBeam = 0;
        if (!planned || total > 0) {
            if (out->stats) bcu_last_kernel_ms(e, &kernelMs);
            rc = bcu_level_select(e, K);
            if (rc) return rc;
            rc = bcu_fetch_beam(e, &nBeam, bCost.data(), bId.data(), bPotted.data(), bTime.data(), bPos.data());
            if (rc) return rc;
        }
        const auto tRules = std::chrono::steady_clock::now();

        std::vector<BeamState> next;
        std::vector<uint8_t> present(NB), present2(NB);
        std::vector<float> rest((size_t)NB * 2), pos2((size_t)NB * 2);
        for (int b = 0; b < nBeam; ++b) {
            const int64_t cid = (int64_t)bId[b];
            int si, k, j;   // planned strikes: k = index into this level's plan (its velocity is kept below), j = -1
            if (planned) {
                si = planState[(size_t)cid]; k = (int)planVel.size(); j = -1;
                planVel.push_back({planVx[(size_t)cid], planVy[(size_t)cid]});
            } else {
                si = (int)(cid / perState);
                const int64_t rem = cid - (int64_t)si * perState;
                k = (int)(rem / sp->n_speeds); j = (int)(rem - (int64_t)k * sp->n_speeds);
            }
            const BeamState& parent = states[si];
            // the balls at rest in the final layer form the next State (:1411-1426)
            for (int i = 0; i < NB; ++i) {
                const bool on = parent.types[i] >= 0 && !((bPotted[b] >> i) & 1u);
                present[i] = on ? 1 : 0;
                rest[2 * i] = on ? bPos[((size_t)b * NB + i) * 2] : 0.0f;
                rest[2 * i + 1] = on ? bPos[((size_t)b * NB + i) * 2 + 1] : 0.0f;
            }
            Found f{bCost[b], level + 1, si, k, j, rest, present};
            if (level + 1 >= L) {   // simCount == BREAKS (:1407)
                found.push_back(std::move(f));
                continue;
            }
            std::vector<int32_t> search;
            bool ok = true;
            if (rules->kind == BCU_BREAK_RULES_SINGLE) {          // Configuration's default rules (type.hpp:325-328)
                std::memcpy(pos2.data(), rest.data(), rest.size() * sizeof(float));
                present2 = present;
            } else if (rules->kind == BCU_BREAK_RULES_REPEAT) {
                search = parent.searched;
                std::memcpy(pos2.data(), rest.data(), rest.size() * sizeof(float));
                present2 = present;
            } else {
                int32_t outTypes[8], cnt = 0, okFlag = 0;
                rc = bcu_snooker_next_search(&rules->snooker, NB, parent.types.data(), present.data(), (int32_t)parent.searched.size(),
                                             parent.searched.data(), outTypes, &cnt);
                if (rc) return rc;
                search.assign(outTypes, outTypes + cnt);
                rc = bcu_snooker_state_after_break(&rules->snooker, NB, rest.data(), parent.types.data(), present.data(), pos2.data(),
                                                   present2.data(), &okFlag);
                if (rc) return rc;
                ok = okFlag != 0;
            }
            if (!ok) continue;   // "could not replace all balls" (:1432-1435): the candidate is dropped
            bool hasTypeToSearch = false;   // :1437-1443
            for (int i = 0; i < NB && !hasTypeToSearch; ++i)
                if (present2[i]) hasTypeToSearch = std::find(search.begin(), search.end(), parent.types[i]) != search.end();
            if (hasTypeToSearch) {
                BeamState s;
                s.pos.resize((size_t)NB * 2); s.types.resize(NB);
                for (int i = 0; i < NB; ++i) {
                    s.pos[2 * i] = present2[i] ? pos2[2 * i] : 0.0f;
                    s.pos[2 * i + 1] = present2[i] ? pos2[2 * i + 1] : 0.0f;
                    s.types[i] = present2[i] ? parent.types[i] : -1;
                }
                s.searched = std::move(search);
                s.cost = bCost[b];
                s.parent = si; s.dir = k; s.speed = j;
                next.push_back(std::move(s));
            } else {
                f.endPos = pos2; f.endPresent = present2;
                found.push_back(std::move(f));
            }
        }
        if (out->stats) {
            bcu_level_stats& st = out->stats[level];
            st.states = nS; st.candidates = total; st.beam = nBeam; st.kernel_ms = kernelMs;
            st.rules_ms = (float)ms_since(tRules); st.wall_ms = (float)ms_since(tLevel);
        }
        out->levels_run = level + 1;
        levels.push_back(std::move(next));
    }

    // ascending by cost, equal costs in discovery order (the reference pops its solution queue by cost, process.hpp:19-25)
    std::vector<int> order(found.size());
    for (size_t i = 0; i < order.size(); ++i) order[i] = (int)i;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return found[a].cost < found[b].cost; });
    out->total = (int32_t)found.size();
    const int n = (int)std::min<size_t>(found.size(), (size_t)out->capacity);
    for (int r = 0; r < n; ++r) {
        const Found& f = found[order[r]];
        out->cost[r] = f.cost;
        out->breaks[r] = f.breaks;
        if (out->end_pos) std::memcpy(out->end_pos + (size_t)r * NB * 2, f.endPos.data(), (size_t)NB * 2 * sizeof(float));
        if (out->end_present) std::memcpy(out->end_present + (size_t)r * NB, f.endPresent.data(), NB);
        for (int l = 0; l < L; ++l) {
            if (out->dir_index) out->dir_index[(size_t)r * L + l] = -1;
            if (out->speed_index) out->speed_index[(size_t)r * L + l] = -1;
            if (out->strike) { out->strike[((size_t)r * L + l) * 2] = 0.0f; out->strike[((size_t)r * L + l) * 2 + 1] = 0.0f; }
            if (out->frame_types) for (int i = 0; i < NB; ++i) out->frame_types[((size_t)r * L + l) * NB + i] = -1;
            if (out->frame_pos) std::memset(out->frame_pos + ((size_t)r * L + l) * NB * 2, 0, (size_t)NB * 2 * sizeof(float));
        }
        // walk the path back: the strike of break l was played on a state of level l
        int si = f.state, k = f.dir, j = f.speed;
        for (int l = f.breaks - 1; l >= 0; --l) {
            const BeamState& s = levels[l][si];
            if (out->dir_index) out->dir_index[(size_t)r * L + l] = planned ? -1 : k;
            if (out->speed_index) out->speed_index[(size_t)r * L + l] = planned ? -1 : j;
            if (out->strike && planned) {
                out->strike[((size_t)r * L + l) * 2] = planVel[(size_t)k].first;
                out->strike[((size_t)r * L + l) * 2 + 1] = planVel[(size_t)k].second;
            } else if (out->strike) {   // as the kernel decodes a grid candidate: one binary32 multiply per component
                out->strike[((size_t)r * L + l) * 2] = sp->speeds[j] * sp->dirs[2 * k];
                out->strike[((size_t)r * L + l) * 2 + 1] = sp->speeds[j] * sp->dirs[2 * k + 1];
            }
            if (out->frame_pos) std::memcpy(out->frame_pos + ((size_t)r * L + l) * NB * 2, s.pos.data(), (size_t)NB * 2 * sizeof(float));
            if (out->frame_types) std::memcpy(out->frame_types + ((size_t)r * L + l) * NB, s.types.data(), (size_t)NB * sizeof(int32_t));
            k = s.dir; j = s.speed; si = s.parent;
        }
    }
    out->count = n;
    out->wall_ms = (float)ms_since(t0);
    return BCU_OK;
}
