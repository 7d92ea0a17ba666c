// TEST INFRASTRUCTURE: the planner restatement (billiard_ai_b200/csrc/bcu_planner.cuh, compiled here FOR THE HOST) against the
// reference's own planner.  Built by oracle/Makefile (target `planner`) against the reference's headers and linked with
// oracle/_ref/libref_oracle.so; no device involved, so the CPU tier runs it.
//
//   reference:   a best-first loop over billiard::search::expand() in which only SEARCH nodes are expanded, run until none
//                is left; every SIMULATION node that appears is recorded (BALL_SHOT velocity, _searchCost, _cost)
//   restatement: bcu::plan_root() for every search pocket, tables built by the product's own plan_table_build() from
//                the product's own PlannerTables::from(Configuration)
// The two multisets of (velocity bits, searchCost, cost) must be identical.
#include <billiard_search/billiard_search.hpp>

#include <algorithm>
#include <array>
#include <cstdio>
#include <cstring>
#include <queue>

#include "../billiard_ai_b200/csrc/bcu_planner.cuh"
#include "../billiard_ai_b200/host/billiard_cuda_search.hpp"

using namespace billiard::search;

extern "C" const void* ref_get_configuration(void);

namespace {
using Strike = std::array<uint64_t, 4>;   // vx bits, vy bits, searchCost, cost

uint64_t bitsOf(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }

std::vector<Strike> referencePlan(const State& state, const Search& search, const Configuration& conf, size_t* expansions) {
    auto ss = std::make_shared<SearchState>(SearchState{conf});
    ss->_ball._accelerationLength = billiard::physics::accelerationLength();
    ss->_ball._slideAccelerationLength = billiard::physics::slideAccelerationLength();
    std::vector<std::shared_ptr<SearchNode>> open = addPocketsAsInitialTargets(state, search, conf);
    std::vector<Strike> out;
    size_t n = 0;
    while (!open.empty()) {
        auto node = open.back();
        open.pop_back();
        ++n;
        for (auto& child : expand(node, ss)) {
            if (child->_type == SearchNodeType::SIMULATION) {
                for (auto& kv : child->asSimulation()->_simulation.firstLayer()._nodes)
                    if (kv.second._type == node::NodeType::BALL_SHOT)
                        out.push_back({bitsOf(kv.second.after()->_velocity.x), bitsOf(kv.second.after()->_velocity.y), child->_searchCost, child->_cost});
            } else {
                open.push_back(child);
            }
        }
    }
    if (expansions) *expansions = n;
    return out;
}

struct Collect {
    std::vector<Strike>* out;
    void operator()(float vx, float vy, uint64_t searchCost, uint64_t cost) { out->push_back({bitsOf(vx), bitsOf(vy), searchCost, cost}); }
};
}  // namespace

// out: [0] reference strikes, [1] restatement strikes, [2] in the reference only, [3] in the restatement only,
// [4] reference expansions.  outStrikes (may be null): up to maxOut reference strikes {vx, vy} / {searchCost, cost}.
extern "C" int planner_check(int n, const float* pos, const char* const* typeNames, int nTypes, const int32_t* types, int nSearchTypes,
                             const int32_t* searchTypes, int velocityMultiply, uint64_t baseCost, int64_t* out, int64_t maxOut,
                             float* outVel, uint64_t* outCost) {
    std::vector<std::string> names(typeNames, typeNames + nTypes);
    Configuration conf = *static_cast<const Configuration*>(ref_get_configuration());
    conf._velocityMultiplyEnabled = velocityMultiply != 0;
    conf._depthSearchEnabled = false;
    std::vector<Ball> balls;
    for (int i = 0; i < n; ++i)
        if (types[i] >= 0) balls.emplace_back(glm::vec2{pos[2 * i], pos[2 * i + 1]}, names[types[i]], "b" + std::to_string(i));
    std::vector<std::string> searched;
    uint32_t mask = 0;
    for (int i = 0; i < nSearchTypes; ++i) { searched.push_back(names[searchTypes[i]]); mask |= 1u << searchTypes[i]; }
    size_t expansions = 0;
    auto want = referencePlan(State(balls), Search{"", searched}, conf, &expansions);
    if (baseCost) for (auto& s : want) s[3] += baseCost;   // the reference's roots of a later break carry the accumulated cost (:1447-1451)

    // the product's tables: rails in the iteration order of the Configuration's map (as the host mirror's Engine passes them)
    auto tables = cuda::PlannerTables::from(conf);
    std::vector<float> derived;
    for (auto& kv : conf._table._rails) {
        const Rail& r = kv.second;
        derived.insert(derived.end(), {r._normal.x, r._normal.y, r._shiftedStart.x, r._shiftedStart.y, r._shiftedEnd.x, r._shiftedEnd.y});
    }
    int cue = -1;
    for (int i = 0; i < nTypes; ++i) if (names[i] == "WHITE") cue = i;
    bcu::PlanTable tb;
    if (!bcu::plan_table_build(&tb, tables.desc.n_rails, derived.data(), tables.desc.rail_start, tables.desc.rail_order, tables.desc.n_set_order,
                               tables.desc.set_order, tables.desc.n_search_pockets, tables.desc.search_pockets, conf._ball._radius,
                               tables.desc.minimal_pocket_velocity, tables.desc.diagonal_length_squared, tables.desc.velocity_multiply, cue))
        return -1;
    std::vector<int8_t> t8(n);
    for (int i = 0; i < n; ++i) t8[i] = (int8_t)types[i];
    bcu::PlanState st{n, pos, t8.data(), mask};
    std::vector<Strike> got;
    Collect collect{&got};
    const bcu::PhysConst pc = bcu::phys_const();
    for (int p = 0; p < tb.nPockets; ++p) bcu::plan_root(tb, st, p, baseCost, pc, collect);

    for (int64_t i = 0; i < maxOut && i < (int64_t)want.size() && outVel && outCost; ++i) {
        uint32_t a = (uint32_t)want[i][0], b = (uint32_t)want[i][1];
        std::memcpy(&outVel[2 * i], &a, 4); std::memcpy(&outVel[2 * i + 1], &b, 4);
        outCost[2 * i] = want[i][2]; outCost[2 * i + 1] = want[i][3];
    }
    std::sort(want.begin(), want.end());
    std::sort(got.begin(), got.end());
    std::vector<Strike> onlyRef, onlyGot;
    std::set_difference(want.begin(), want.end(), got.begin(), got.end(), std::back_inserter(onlyRef));
    std::set_difference(got.begin(), got.end(), want.begin(), want.end(), std::back_inserter(onlyGot));
    out[0] = (int64_t)want.size(); out[1] = (int64_t)got.size(); out[2] = (int64_t)onlyRef.size(); out[3] = (int64_t)onlyGot.size();
    out[4] = (int64_t)expansions;
    if (getenv("PLANNER_CHECK_VERBOSE")) {
        for (size_t i = 0; i < onlyRef.size() && i < 8; ++i) { float x, y; uint32_t a = (uint32_t)onlyRef[i][0], b = (uint32_t)onlyRef[i][1]; std::memcpy(&x, &a, 4); std::memcpy(&y, &b, 4);
            std::fprintf(stderr, "ref only: v=(%.9g, %.9g) searchCost=%llu cost=%llu\n", x, y, (unsigned long long)onlyRef[i][2], (unsigned long long)onlyRef[i][3]); }
        for (size_t i = 0; i < onlyGot.size() && i < 8; ++i) { float x, y; uint32_t a = (uint32_t)onlyGot[i][0], b = (uint32_t)onlyGot[i][1]; std::memcpy(&x, &a, 4); std::memcpy(&y, &b, 4);
            std::fprintf(stderr, "got only: v=(%.9g, %.9g) searchCost=%llu cost=%llu\n", x, y, (unsigned long long)onlyGot[i][2], (unsigned long long)onlyGot[i][3]); }
    }
    return 0;
}

// the two iteration orders the device planner takes as data, as THIS reference build has them (for callers that hand the
// engine its rails in another order than the Configuration's map: the Python tests).  railIdsInCallerOrder: '\n'-joined.
extern "C" int planner_orders(const char* railIdsInCallerOrder, int32_t* outRailOrder, int32_t* outSetOrder, int32_t* outNSet) {
    const Configuration& conf = *static_cast<const Configuration*>(ref_get_configuration());
    std::vector<std::string> ids;
    std::string cur;
    for (const char* p = railIdsInCallerOrder; *p; ++p) { if (*p == '\n') { ids.push_back(cur); cur.clear(); } else cur.push_back(*p); }
    if (!cur.empty()) ids.push_back(cur);
    auto indexOf = [&](const std::string& id) { for (size_t i = 0; i < ids.size(); ++i) if (ids[i] == id) return (int32_t)i; return (int32_t)-1; };
    std::unordered_set<std::string> railIds;
    int n = 0;
    for (auto& kv : conf._table._rails) {
        outRailOrder[n++] = indexOf(kv.first);
        if (kv.second._normal.x != 0.0f && kv.second._normal.y != 0.0f) continue;
        railIds.insert(kv.first);
    }
    int m = 0;
    for (auto& id : railIds) outSetOrder[m++] = indexOf(id);
    *outNSet = m;
    return n;
}
