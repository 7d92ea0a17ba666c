// TEST INFRASTRUCTURE: checks the C++ host mirror (billiard_ai_b200/host/billiard_cuda_search.hpp) against the reference.
// Compiled in the build container against the reference's own headers (oracle/Makefile target `mirror`), linked with
// oracle/_ref/libref_oracle.so (the reference) and billiard_ai_b200/libbilliard_cuda.so (the product); the binary
// travels to the GPU box.  For each strike it calls
//     billiard::search::simulate(state, velocity, config)              -- the reference, CPU
//     billiard::search::cuda::Engine::simulate(state, velocity)        -- the drop-in, B200
// and compares the two node::System objects node by node (types, ball types, before/after BallState bits, causes,
// layer times, _nextRolling / _lastRailCollisions bookkeeping).
#include <billiard_search/billiard_search.hpp>

#include <chrono>
#include <cstdio>
#include <cstring>
#include <queue>
#include <thread>

#include "../billiard_ai_b200/host/billiard_cuda_search.hpp"

using namespace billiard::search;

extern "C" const void* ref_get_configuration(void);

static bool sameBits(float a, float b) { return std::memcmp(&a, &b, 4) == 0; }
static bool sameState(const state::BallState& a, const state::BallState& b) {
    return sameBits(a._position.x, b._position.x) && sameBits(a._position.y, b._position.y) &&
           sameBits(a._velocity.x, b._velocity.x) && sameBits(a._velocity.y, b._velocity.y) &&
           sameBits(a._acceleration.x, b._acceleration.x) && sameBits(a._acceleration.y, b._acceleration.y) &&
           sameBits(a._velocityNormed.x, b._velocityNormed.x) && sameBits(a._velocityNormed.y, b._velocityNormed.y) &&
           a._isRolling == b._isRolling;
}

static int compareSystems(const node::System& a, const node::System& b) {
    int diff = 0;
    if (a._layers.size() != b._layers.size()) return 1000;
    for (size_t l = 0; l < a._layers.size(); ++l) {
        const auto& la = a._layers[l];
        const auto& lb = b._layers[l];
        if (!sameBits(la._time, lb._time) || la._nodes.size() != lb._nodes.size()) diff++;
        if (la._dynamic.size() != lb._dynamic.size() || la._static.size() != lb._static.size()) diff++;
        for (auto& kv : la._nodes) {
            auto it = lb._nodes.find(kv.first);
            if (it == lb._nodes.end()) { diff++; continue; }
            const node::Node& na = kv.second;
            const node::Node& nb = it->second;
            if (na._type != nb._type || na._ballType != nb._ballType) { diff++; continue; }
            if (!sameState(*na.after(), *nb.after()) || !sameState(*na.before(), *nb.before())) diff++;
            auto ia = na.getInvolvedId(), ib = nb.getInvolvedId();
            if (na._type == node::BALL_COLLISION) {
                // the reference's {ball1, ball2} order depends on hash-map iteration; compare as a set
                if (!((ia.first == ib.first && ia.second == ib.second) || (ia.first == ib.second && ia.second == ib.first))) diff++;
            } else if (ia != ib) diff++;
        }
    }
    if (a._nextRolling.size() != b._nextRolling.size() || a._lastRailCollisions != b._lastRailCollisions) diff++;
    for (auto& kv : a._nextRolling) {
        auto it = b._nextRolling.find(kv.first);
        if (it == b._nextRolling.end() || !sameBits(it->second, kv.second)) diff++;
    }
    return diff;
}

// returns the number of strikes whose Systems differ (or where only one side returned nullopt); -1 on engine failure
extern "C" int mirror_check(int n, const float* pos, const char* const* typeNames, int nTypes, const int32_t* types,
                            int nStrikes, const float* vx, const float* vy, int* outCompared, int* outNullopt) {
    const Configuration& conf = *static_cast<const Configuration*>(ref_get_configuration());
    std::vector<std::string> names(typeNames, typeNames + nTypes);
    try {
        cuda::Engine engine(conf, cuda::RulesTable::fromConfiguration(conf, names));
        std::vector<Ball> balls;
        for (int i = 0; i < n; ++i) balls.emplace_back(glm::vec2{pos[2 * i], pos[2 * i + 1]}, names[types[i]], "b" + std::to_string(i));
        State state(balls);
        int bad = 0, compared = 0, nullopts = 0;
        for (int s = 0; s < nStrikes; ++s) {
            const glm::vec2 v{vx[s], vy[s]};
            auto want = simulate(state, v, conf);
            auto got = engine.simulate(state, v);
            if (want.has_value() != got.has_value()) { bad++; continue; }
            if (!want) { nullopts++; continue; }
            compared++;
            if (compareSystems(*want, *got) != 0) bad++;
        }
        // the same strikes in one batch (Engine::simulateAll -> bcu_trace_strikes), two copies of the state
        std::vector<std::pair<int, glm::vec2>> strikes;
        for (int s = 0; s < nStrikes; ++s) strikes.push_back({s & 1, glm::vec2{vx[s], vy[s]}});
        auto all = engine.simulateAll({state, state}, strikes);
        if ((int)all.size() != nStrikes) return -1;
        for (int s = 0; s < nStrikes; ++s) {
            auto want = simulate(state, strikes[s].second, conf);
            if (want.has_value() != all[s].has_value()) { bad++; continue; }
            if (want && compareSystems(*want, *all[s]) != 0) bad++;
        }
        *outCompared = compared;
        *outNullopt = nullopts;
        return bad;
    } catch (const std::exception& e) {
        std::fprintf(stderr, "mirror_check: %s\n", e.what());
        return -1;
    }
}

// ---- search(): the drop-in against the reference's own best-first search run to exhaustion ---------------------------------
// The reference's search() is timing dependent (worker threads, a 100 ms sync period, a 10 s budget), so the comparison is
// made against what it converges to: a deterministic best-first loop over the reference's own expand() -- planner AND its CPU
// simulate() -- until no node is left, solutions ascending by cost (process.hpp:19-25), each mapped with allSimulations()
// (type.hpp:525-537) exactly as mapToSolution does (billiard_search.cpp:185-222).
namespace billiard::snooker {   // oracle/_ref/libref_rules.so: the reference's own rules between breaks (billiard_snooker.cpp:792-822, 1097-1126)
billiard::search::Search nextSearch(const billiard::search::State& state, const std::vector<std::string>& previousTypes);
billiard::search::State stateAfterBreak(const billiard::search::State& state, const std::unordered_map<std::string, std::string>& ids);
}  // namespace billiard::snooker

namespace {
struct RefSolution {
    uint64_t cost;
    std::vector<node::System> systems;
    bool used = false;
};

std::vector<RefSolution> referenceSearch(const State& state, const Search& search, const Configuration& conf, size_t* outExpansions) {
    auto ss = std::make_shared<SearchState>(SearchState{conf});
    ss->_ball._accelerationLength = billiard::physics::accelerationLength();
    ss->_ball._slideAccelerationLength = billiard::physics::slideAccelerationLength();
    struct Item { std::shared_ptr<SearchNode> node; uint64_t order; };
    auto later = [](const Item& a, const Item& b) { return a.node->_cost != b.node->_cost ? a.node->_cost > b.node->_cost : a.order > b.order; };
    std::priority_queue<Item, std::vector<Item>, decltype(later)> open(later);
    uint64_t order = 0;
    for (auto& nd : addPocketsAsInitialTargets(state, search, conf)) open.push({nd, order++});
    std::vector<Item> found;
    size_t expansions = 0;
    while (!open.empty()) {
        auto item = open.top();
        open.pop();
        expansions++;
        for (auto& e : expand(item.node, ss)) {
            if (e->_isSolution) found.push_back({e, order++});
            else open.push({e, order++});
        }
    }
    std::sort(found.begin(), found.end(), [](const Item& a, const Item& b) { return a.node->_cost != b.node->_cost ? a.node->_cost < b.node->_cost : a.order < b.order; });
    std::vector<RefSolution> out;
    for (auto& f : found) out.push_back(RefSolution{f.node->_cost, f.node->allSimulations()});
    if (outExpansions) *outExpansions = expansions;
    return out;
}

Configuration configured(int depth, int velocityMultiply, int rulesKind) {
    Configuration conf = *static_cast<const Configuration*>(ref_get_configuration());   // table, ball, validEndState, scores
    conf._depthSearchEnabled = depth != 0;
    conf._velocityMultiplyEnabled = velocityMultiply != 0;
    if (rulesKind == 1) {          // "repeat": the same types again, the state as it is
        conf._rules._nextSearch = [](const State&, const std::vector<std::string>& previous) { return Search{"", previous}; };
        conf._rules._modifyState = [](const State& s, const std::unordered_map<std::string, std::string>&) { return s; };
    } else if (rulesKind == 2) {   // the reference's snooker rules (searchConfig set through refrules_configure)
        conf._rules._nextSearch = billiard::snooker::nextSearch;
        conf._rules._modifyState = billiard::snooker::stateAfterBreak;
    }                              // 0: the Configuration's defaults (type.hpp:325-328)
    return conf;
}

State stateOf(int n, const float* pos, const std::vector<std::string>& names, const int32_t* types) {
    std::vector<Ball> balls;
    for (int i = 0; i < n; ++i) balls.emplace_back(glm::vec2{pos[2 * i], pos[2 * i + 1]}, names[types[i]], "b" + std::to_string(i));
    return State(balls);
}
}  // namespace

// out: [0] reference solutions, [1] solutions returned, [2] cost mismatches among the first min(solutions, both) entries,
// [3] returned solutions whose Systems match no reference solution of the same cost, [4] reference expansions,
// [5] drop-in expansions, [6] drop-in simulations, [7] drop-in batches, [8] ended by enoughSolutions, [9] by noWork.
// outCosts [2][maxCosts]: reference costs, returned costs.  outMs [2]: reference, drop-in.  Returns 0, or -1 on failure.
extern "C" int mirror_search_check(int n, const float* pos, const char* const* typeNames, int nTypes, const int32_t* types,
                                   int nSearchTypes, const int32_t* searchTypes, int solutions, int depth, int velocityMultiply,
                                   int rulesKind, int batch, int64_t* out, int maxCosts, uint64_t* outCosts, double* outMs) {
    std::vector<std::string> names(typeNames, typeNames + nTypes);
    try {
        const Configuration conf = configured(depth, velocityMultiply, rulesKind);
        const State state = stateOf(n, pos, names, types);
        std::vector<std::string> searched;
        for (int i = 0; i < nSearchTypes; ++i) searched.push_back(names[searchTypes[i]]);
        const Search search{"", searched};
        auto t0 = std::chrono::steady_clock::now();
        size_t refExpansions = 0;
        auto want = referenceSearch(state, search, conf, &refExpansions);
        outMs[0] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();

        cuda::SearchOptions options;
        if (batch > 0) options.batch = (size_t)batch;
        if (batch < 0) options.devicePlanner = true;   // batch = -1: the planner runs on the device as well
        cuda::SearchFront front{options};
        t0 = std::chrono::steady_clock::now();
        auto got = front.process(state, search, (uint16_t)solutions, conf).get();
        outMs[1] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        const cuda::SearchReport report = front.lastReport();
        if (report.costs.size() != got.size()) return -1;

        int64_t costMismatch = 0, systemMismatch = 0;
        const size_t k = std::min<size_t>({(size_t)solutions, want.size(), got.size()});
        if (got.size() < std::min<size_t>((size_t)solutions, want.size())) costMismatch += 1000;   // too few solutions returned
        for (size_t i = 0; i < k; ++i) costMismatch += want[i].cost != report.costs[i];
        for (size_t i = 1; i < got.size(); ++i) costMismatch += report.costs[i] < report.costs[i - 1];   // ascending
        for (size_t i = 0; i < got.size(); ++i) {
            bool matched = false;
            for (auto& w : want) {
                if (w.used || w.cost != report.costs[i] || w.systems.size() != got[i].size()) continue;
                bool same = true;
                for (size_t b = 0; b < got[i].size() && same; ++b) same = compareSystems(w.systems[b], got[i][b]) == 0;
                if (same) { w.used = true; matched = true; break; }
            }
            systemMismatch += !matched;
        }
        out[0] = (int64_t)want.size(); out[1] = (int64_t)got.size(); out[2] = costMismatch; out[3] = systemMismatch;
        out[4] = (int64_t)refExpansions; out[5] = (int64_t)report.expansions; out[6] = (int64_t)report.simulations;
        out[7] = (int64_t)report.batches; out[8] = report.enoughSolutions; out[9] = report.noWork;
        for (int i = 0; i < maxCosts; ++i) {
            outCosts[i] = i < (int)want.size() ? want[i].cost : ~0ull;
            outCosts[maxCosts + i] = i < (int)got.size() ? report.costs[i] : ~0ull;
        }
        return 0;
    } catch (const std::exception& e) {
        std::fprintf(stderr, "mirror_search_check: %s\n", e.what());
        return -1;
    }
}

// The manager's behaviour around a request (test/process/src/process_test.cpp:29-132): a newer request replaces the one in
// flight (its future breaks its promise, as the reference's does), an invalid search and `solutions` = 0 return at once,
// the time budget ends a request with what it has, and the free function search() returns ascending results.
// out: [0] first future broke its promise, [1] second future's solutions, [2] invalid search returned empty, [3] budget run
// ended by noTimeLeft, [4] ms of the budget run, [5] solutions of search(), [6] solutions = 0 returned empty,
// [7] searchOnly() returned the same solutions as nodes.
extern "C" int mirror_front_check(int n, const float* pos, const char* const* typeNames, int nTypes, const int32_t* types,
                                  int nSearchTypes, const int32_t* searchTypes, int64_t* out) {
    std::vector<std::string> names(typeNames, typeNames + nTypes);
    try {
        const Configuration conf = configured(1, 1, 1);
        const State state = stateOf(n, pos, names, types);
        std::vector<std::string> searched;
        for (int i = 0; i < nSearchTypes; ++i) searched.push_back(names[searchTypes[i]]);
        const Search search{"", searched};
        {
            cuda::SearchFront front{};
            auto first = front.process(state, search, 65535, conf);    // exhaustive two-break search: seconds of planning
            std::this_thread::sleep_for(std::chrono::milliseconds(30));
            auto second = front.process(state, search, 1, conf);
            bool broken = false;
            try { first.get(); } catch (const std::future_error& fe) { broken = fe.code() == std::future_errc::broken_promise; }
            out[0] = broken;
            out[1] = (int64_t)second.get().size();
            out[2] = front.process(state, Search{"", {"WHITE"}}, 5, conf).get().empty();   // isValidSearch (:100-120)
            out[6] = front.process(state, search, 0, conf).get().empty();
        }
        {
            cuda::SearchOptions options;
            options.maxDurationMs = 20;
            cuda::SearchFront front{options};
            const auto t0 = std::chrono::steady_clock::now();
            front.process(state, search, 65535, conf).get();
            out[4] = (int64_t)std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            out[3] = front.lastReport().noTimeLeft;
        }
        auto systems = cuda::search(state, search, 3, conf).get();
        out[5] = (int64_t)systems.size();
        // searchOnly(): the same solutions as SearchNodes (cost ascending, Systems on their paths)
        auto nodes = cuda::searchOnly(state, search, 3, conf).get();
        bool same = nodes.size() == systems.size();
        for (size_t i = 0; same && i < nodes.size(); ++i) {
            auto sims = nodes[i]->allSimulations();
            same = nodes[i]->_isSolution && sims.size() == systems[i].size() && (i == 0 || nodes[i - 1]->_cost <= nodes[i]->_cost);
            for (size_t b = 0; same && b < sims.size(); ++b) same = compareSystems(sims[b], systems[i][b]) == 0;
        }
        out[7] = same;
        return 0;
    } catch (const std::exception& e) {
        std::fprintf(stderr, "mirror_front_check: %s\n", e.what());
        return -1;
    }
}

// The reference side alone (no device): costs of every solution of the exhaustive best-first search, ascending.
// Used by the CPU tier to pin the multi-break oracle on the thesis' two-break list (04_suche.tex:1114-1246).
extern "C" int64_t mirror_reference_search(int n, const float* pos, const char* const* typeNames, int nTypes, const int32_t* types,
                                           int nSearchTypes, const int32_t* searchTypes, int depth, int velocityMultiply, int rulesKind,
                                           int64_t maxOut, uint64_t* outCosts, int32_t* outBreaks) {
    std::vector<std::string> names(typeNames, typeNames + nTypes);
    const Configuration conf = configured(depth, velocityMultiply, rulesKind);
    const State state = stateOf(n, pos, names, types);
    std::vector<std::string> searched;
    for (int i = 0; i < nSearchTypes; ++i) searched.push_back(names[searchTypes[i]]);
    auto found = referenceSearch(state, Search{"", searched}, conf, nullptr);
    for (int64_t i = 0; i < (int64_t)found.size() && i < maxOut; ++i) {
        outCosts[i] = found[i].cost;
        outBreaks[i] = (int32_t)found[i].systems.size();
    }
    return (int64_t)found.size();
}
