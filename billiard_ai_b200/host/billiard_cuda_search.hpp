// Host-side C++ mirror of the reference's entry points for the strike-evaluation path, on top of the C ABI
// (include/billiard_cuda.h).  Header-only and written AGAINST THE REFERENCE'S OWN HEADERS: it is compiled inside the
// reference tree (it includes <billiard_search/type.hpp> and <billiard_physics/billiard_physics.hpp> from there), so
// no reference type is copied or re-declared here.
//
//   billiard::search::cuda::Engine::simulate(state, velocity)
//        == billiard::search::simulate(const State&, const glm::vec2&, const Configuration&)
//           lib/billiard_search/include/billiard_search/billiard_search.hpp:17-18, src/billiard_search.cpp:141-175
//        same argument meaning, same result type (std::optional<node::System>), same "nullopt when more than
//        MAX_EVENTS events or no event is found" behaviour; the System is rebuilt layer by layer from the device
//        trace with the reference's own Node / Layer / System::append, so callers (unity_adapter.cpp:548-568,
//        test/billiard_search/src/simulation_vs_reality.cpp:86) consume it unchanged.
//   Engine::evaluate(state, expectedTypes, candidates)
//        == one SIMULATION node per candidate: simulate(node, state, mustBeValid) + simulationCost + rules
//           (billiard_search.cpp:1358-1471, 1473-1532), returning cost / validity / end state per candidate.
//
//   billiard::search::cuda::search(state, search, solutions, config)
//        == billiard::search::search(const State&, const Search&, uint16_t, const Configuration&)
//           billiard_search.hpp:9-12, src/billiard_search.cpp:74-98: same arguments, same result type
//           (std::future<std::vector<std::vector<node::System>>>: outer vector ascending by cost, inner vector one System
//           per break, type.hpp:525-537), same ways to end (enough solutions / 10 000 ms / no work left,
//           process.hpp:188-198), a newer request replaces the one in flight (process.hpp:82-83).  The geometric planner
//           stays the reference's own expand() on the host (billiard_search.cpp:224-244); every SIMULATION node it emits
//           is evaluated on the device in batches instead of one by one on the CPU (:1358-1471), the rules between
//           breaks are the Configuration's own std::functions (:1407-1462).  See SearchFront below.
//
// Errors: a failing bcu_* call throws std::runtime_error with bcu_last_error(); there is no CPU fallback.
#pragma once

#include <billiard_physics/billiard_physics.hpp>
#include <billiard_search/type.hpp>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <future>
#include <memory>
#include <mutex>
#include <optional>
#include <queue>
#include <stdexcept>
#include <string>
#include <thread>
#include <unordered_map>
#include <unordered_set>
#include <utility>
#include <vector>

#include "billiard_cuda.h"

namespace billiard::search::cuda {

// Rules as data: the reference's std::function rules (type.hpp:324-332) cannot run on a device.  One entry per
// ball type: name, score = _scoreForPottedBall(name).  "WHITE" must be present (it is the cue ball,
// billiard_search.cpp:148-150).
struct RulesTable {
    std::vector<std::string> typeNames;
    std::vector<double> scores;

    static RulesTable fromConfiguration(const Configuration& config, const std::vector<std::string>& typeNames) {
        RulesTable t;
        t.typeNames = typeNames;
        for (auto& n : typeNames) t.scores.push_back(config._rules._scoreForPottedBall(n));
        return t;
    }
    int code(const std::string& name) const {
        for (size_t i = 0; i < typeNames.size(); ++i) if (typeNames[i] == name) return (int)i;
        return -1;
    }
};

struct Candidate {
    glm::vec2 velocity;
    uint64_t searchCost;  // SearchNode::_searchCost of the planner's node (type.hpp:572-575)
    uint64_t baseCost;    // SearchNode::_cost accumulated so far
};

struct Evaluation {
    uint64_t cost;        // parent cost + simulationCost; UINT64_MAX when the reference would discard the candidate
    bool completed, valid;
    int events;
    float finalTime;
    std::vector<glm::vec2> finalPositions;  // per ball of the State, in its order
    std::vector<bool> potted;
};

class Engine {
public:
    Engine(const Configuration& config, RulesTable rules, int device = 0) : _rules(std::move(rules)) {
        std::vector<float> rails, pockets;
        for (auto& kv : config._table._rails) {  // the reference iterates this very map (billiard_search.cpp:1648)
            const Rail& r = kv.second;
            rails.insert(rails.end(), {r._normal.x, r._normal.y, r._shiftedStart.x, r._shiftedStart.y,
                                       r._shiftedEnd.x, r._shiftedEnd.y});
            _railIds.push_back(r._id);
        }
        for (auto& p : config._table._pockets) {
            pockets.insert(pockets.end(), {p._position.x, p._position.y, p._radius});
            _pocketIds.push_back(p._id);
        }
        bcu_table_desc td{(int32_t)_railIds.size(), (int32_t)_pocketIds.size(), config._ball._radius, nullptr,
                          pockets.data(), rails.data()};
        const int cue = _rules.code("WHITE");
        if (cue < 0) throw std::runtime_error("billiard::search::cuda: rules table lacks the WHITE type");
        bcu_rules_desc rd{(int32_t)_rules.typeNames.size(), cue, _rules.scores.data()};
        const int rc = bcu_create(&td, &rd, device, &_e);
        if (rc != BCU_OK) throw std::runtime_error("bcu_create failed (status " + std::to_string(rc) + "): no CUDA device? There is no CPU fallback.");
    }
    ~Engine() { if (_e) bcu_destroy(_e); }
    Engine(const Engine&) = delete;
    Engine& operator=(const Engine&) = delete;

    // == billiard::search::simulate(state, velocity, config)
    std::optional<node::System> simulate(const State& state, const glm::vec2& velocity) {
        const int n = (int)state._balls.size();
        std::vector<float> pos;
        std::vector<int32_t> types;
        flatten(state, pos, types);
        std::vector<float> times(BCU_MAX_LAYERS), rec((size_t)BCU_MAX_LAYERS * n * BCU_TRACE_REC);
        int32_t layers = 0;
        check(bcu_simulate_trace(_e, n, pos.data(), types.data(), velocity.x, velocity.y, &layers, times.data(), rec.data()),
              "bcu_simulate_trace");
        return systemFromTrace(state, layers, times.data(), rec.data());
    }

    // The strikes of several solutions at once -- one node::System per (state, velocity) pair, as the inner vectors of
    // billiard::search::search()'s result hold one System per break (billiard_search.hpp:9-12, type.hpp:525-537).
    // All states must have the same number of balls (the frame's slots; one launch, one copy back).
    std::vector<std::optional<node::System>> simulateAll(const std::vector<State>& states,
                                                         const std::vector<std::pair<int, glm::vec2>>& strikes) {
        std::vector<std::optional<node::System>> out;
        if (states.empty() || strikes.empty()) return out;
        const int n = (int)states[0]._balls.size();
        std::vector<float> pos;
        std::vector<int32_t> types;
        for (auto& st : states) {
            if ((int)st._balls.size() != n) throw std::runtime_error("billiard::search::cuda: states of different sizes in one batch");
            flatten(st, pos, types);
        }
        std::vector<uint32_t> mask(states.size(), 0xffffffffu);
        check(bcu_set_states(_e, (int32_t)states.size(), n, pos.data(), types.data(), mask.data(), nullptr), "bcu_set_states");
        const int32_t m = (int32_t)strikes.size();
        std::vector<int32_t> sidx(m), layers(m);
        std::vector<float> vx(m), vy(m), times((size_t)m * BCU_MAX_LAYERS), rec((size_t)m * BCU_MAX_LAYERS * n * BCU_TRACE_REC);
        for (int32_t i = 0; i < m; ++i) { sidx[i] = strikes[i].first; vx[i] = strikes[i].second.x; vy[i] = strikes[i].second.y; }
        check(bcu_trace_strikes(_e, m, sidx.data(), vx.data(), vy.data(), layers.data(), times.data(), rec.data()), "bcu_trace_strikes");
        for (int32_t i = 0; i < m; ++i)
            out.push_back(systemFromTrace(states[sidx[i]], layers[i], &times[(size_t)i * BCU_MAX_LAYERS],
                                          &rec[(size_t)i * BCU_MAX_LAYERS * n * BCU_TRACE_REC]));
        return out;
    }

    // device trace (include/billiard_cuda.h: bcu_simulate_trace) -> the reference's node::System
    // `slots` > 0: the trace was made on a frame padded to that many ball slots (slots past the State's balls are empty)
    std::optional<node::System> systemFromTrace(const State& state, int layers, const float* times, const float* rec, int slots = 0) {
        const int n = slots > 0 ? slots : (int)state._balls.size();
        if (layers == 0) return std::nullopt;  // billiard_search.cpp:170-174
        const float al = physics::accelerationLength(), sl = physics::slideAccelerationLength();
        node::System system;
        std::unordered_map<std::string, node::Node> previous;
        for (int l = 0; l < layers; ++l) {
            std::unordered_map<std::string, node::Node> nodes;
            for (int b = 0; b < n; ++b) {
                const float* r = &rec[((size_t)l * n + b) * BCU_TRACE_REC];
                if (r[0] < 0.0f || b >= (int)state._balls.size()) continue;  // no such ball
                const Ball& ball = state._balls[b];
                {
                    // createLayer copies the nodes of static balls (at rest / potted) unchanged (billiard_search.cpp:1965-1966)
                    const auto staticType = (node::NodeType)(int)r[0];
                    auto it = previous.find(ball._id);
                    if (it != previous.end() && it->second._type == staticType &&
                        (staticType == node::BALL_POTTING || staticType == node::BALL_IN_REST)) {
                        nodes.insert(*it);
                        continue;
                    }
                }
                const bool afterRolling = r[5] != 0.0f;
                const bool beforeRolling = l > 0 ? rec[((size_t)(l - 1) * n + b) * BCU_TRACE_REC + 5] != 0.0f : false;
                const state::BallState after{glm::vec2{r[1], r[2]}, glm::vec2{r[3], r[4]}, al, sl, afterRolling};
                const state::BallState before{glm::vec2{r[1], r[2]}, glm::vec2{r[6], r[7]}, al, sl, beforeRolling};
                const int cause = (int)r[8];
                const auto type = (node::NodeType)(int)r[0];
                switch (type) {
                    case node::BALL_MOVING:
                        nodes.insert({ball._id, node::Node{type, ball._type, node::BallMovingNode{before, after}}});
                        break;
                    case node::BALL_COLLISION:
                        nodes.insert({ball._id, node::Node{type, ball._type, node::BallCollisionNode{
                            before, after, event::BallCollision{ball._id, state._balls[cause]._id}}}});
                        break;
                    case node::BALL_RAIL_COLLISION:
                        nodes.insert({ball._id, node::Node{type, ball._type, node::BallRailCollisionNode{
                            before, after, event::BallRailCollision{ball._id, _railIds[cause]}}}});
                        break;
                    case node::BALL_POTTING:
                        nodes.insert({ball._id, node::Node{type, ball._type, node::BallPottingNode{
                            before, after, event::BallPotting{ball._id, _pocketIds[cause]}}}});
                        break;
                    case node::BALL_SHOT:
                        nodes.insert({ball._id, node::Node{type, ball._type, node::BallShotNode{after}}});
                        break;
                    case node::BALL_IN_REST:
                        nodes.insert({ball._id, node::Node{type, ball._type, node::BallInRestNode{after}}});
                        break;
                }
            }
            system.append(node::Layer{times[l], nodes});  // the reference's own bookkeeping (type.cpp:467-531)
            previous = std::move(nodes);
        }
        return system;
    }

    // One SIMULATION node per candidate on one table state.
    std::vector<Evaluation> evaluate(const State& state, const std::vector<std::string>& expectedTypes,
                                     const std::vector<Candidate>& candidates) {
        const int n = (int)state._balls.size();
        std::vector<float> pos;
        std::vector<int32_t> types;
        flatten(state, pos, types);
        uint32_t mask = 0;
        for (auto& t : expectedTypes) { const int c = _rules.code(t); if (c >= 0) mask |= 1u << c; }
        check(bcu_set_states(_e, 1, n, pos.data(), types.data(), &mask, nullptr), "bcu_set_states");
        const int64_t m = (int64_t)candidates.size();
        std::vector<int32_t> sidx(m, 0);
        std::vector<float> vx(m), vy(m);
        std::vector<uint64_t> sc(m), bc(m), cost(m);
        std::vector<uint32_t> flags(m), potted(m);
        std::vector<float> ftime(m), fpos((size_t)m * n * 2);
        for (int64_t i = 0; i < m; ++i) {
            vx[i] = candidates[i].velocity.x; vy[i] = candidates[i].velocity.y;
            sc[i] = candidates[i].searchCost; bc[i] = candidates[i].baseCost;
        }
        bcu_results out{cost.data(), flags.data(), potted.data(), ftime.data(), fpos.data(), nullptr};
        check(bcu_evaluate(_e, m, sidx.data(), vx.data(), vy.data(), sc.data(), bc.data(), &out), "bcu_evaluate");
        std::vector<Evaluation> res(m);
        for (int64_t i = 0; i < m; ++i) {
            Evaluation& ev = res[i];
            ev.cost = cost[i];
            ev.completed = flags[i] & BCU_F_COMPLETED;
            ev.valid = flags[i] & BCU_F_VALID;
            ev.events = (int)BCU_N_EVENTS(flags[i]);
            ev.finalTime = ftime[i];
            for (int b = 0; b < n; ++b) {
                ev.finalPositions.push_back(glm::vec2{fpos[((size_t)i * n + b) * 2], fpos[((size_t)i * n + b) * 2 + 1]});
                ev.potted.push_back((potted[i] >> b) & 1u);
            }
        }
        return res;
    }

    bcu_engine* handle() { return _e; }
    const RulesTable& rules() const { return _rules; }
    void checked(int rc, const char* what) const { check(rc, what); }

private:
    void flatten(const State& state, std::vector<float>& pos, std::vector<int32_t>& types) const {
        if (state._balls.size() > BCU_MAX_BALLS) throw std::runtime_error("billiard::search::cuda: more than 32 balls");
        for (auto& b : state._balls) {
            pos.push_back(b._position.x); pos.push_back(b._position.y);
            types.push_back(_rules.code(b._type));
        }
    }
    void check(int rc, const char* what) const {
        if (rc != BCU_OK) throw std::runtime_error(std::string(what) + ": " + bcu_last_error(_e));
    }

    bcu_engine* _e = nullptr;
    RulesTable _rules;
    std::vector<std::string> _railIds, _pocketIds;
};

}  // namespace billiard::search::cuda

// ---- tables of the device planner (bcu_set_planner) from a Configuration ------------------------------------------------
namespace billiard::search::cuda {

// The POD arrays behind bcu_planner_desc.  Rails are in the iteration order of Configuration::_table._rails -- the order
// Engine's constructor hands them to bcu_create -- so rail_order is the identity; set_order is read off a real
// std::unordered_set<std::string> filled exactly as railCombinations fills its own (billiard_search.cpp:871-880), so it is
// the reference's order on this standard library.
struct PlannerTables {
    std::vector<float> railStart, searchPockets;
    std::vector<int32_t> setOrder;
    bcu_planner_desc desc{};

    static PlannerTables from(const Configuration& config) {
        PlannerTables t;
        std::unordered_map<std::string, int32_t> indexOf;
        std::unordered_set<std::string> railIds;
        for (auto& kv : config._table._rails) {
            const Rail& r = kv.second;
            indexOf.insert({kv.first, (int32_t)indexOf.size()});
            t.railStart.push_back(r._start.x); t.railStart.push_back(r._start.y);
            if (r._normal.x != 0.0f && r._normal.y != 0.0f) continue;   // ignoreRailInSearch (:860-863)
            railIds.insert(kv.first);
        }
        for (auto& id : railIds) t.setOrder.push_back(indexOf.at(id));
        for (auto& p : config._searchPockets)
            t.searchPockets.insert(t.searchPockets.end(), {p._pottingPoint.x, p._pottingPoint.y, p._normal.x, p._normal.y});
        t.fill(config);
        return t;
    }
    void fill(const Configuration& config) {
        desc.n_rails = (int32_t)(railStart.size() / 2);
        desc.rail_start = railStart.data();
        desc.rail_order = nullptr;
        desc.n_set_order = (int32_t)setOrder.size();
        desc.set_order = setOrder.data();
        desc.n_search_pockets = (int32_t)(searchPockets.size() / 4);
        desc.search_pockets = searchPockets.data();
        desc.minimal_pocket_velocity = config._table.minimalPocketVelocity;
        desc.diagonal_length_squared = config._table.diagonalLengthSquared;
        desc.velocity_multiply = config._velocityMultiplyEnabled ? 1 : 0;
    }
    PlannerTables() = default;
    PlannerTables(const PlannerTables&) = delete;             // desc points into the vectors
    PlannerTables& operator=(const PlannerTables&) = delete;
    PlannerTables(PlannerTables&& o) noexcept : railStart(std::move(o.railStart)), searchPockets(std::move(o.searchPockets)),
                                                setOrder(std::move(o.setOrder)), desc(o.desc) {
        desc.rail_start = railStart.data(); desc.set_order = setOrder.data(); desc.search_pockets = searchPockets.data();
    }
};

}  // namespace billiard::search::cuda

// ---- the search() drop-in ---------------------------------------------------------------------------------------------
// Internals of the reference's billiard_search.cpp with external linkage (declared there at :42-49, defined :100-120,
// 224-297): the planner and the request check are the reference's own code, called, not restated.
namespace billiard::search {
std::vector<std::shared_ptr<SearchNode>> expand(const std::shared_ptr<SearchNode>& input, const std::shared_ptr<SearchState>& config);
std::vector<std::shared_ptr<SearchNode>> addPocketsAsInitialTargets(const State& state, const Search& search, const Configuration& config);
bool isValidSearch(const State& state, const Search& search);
}  // namespace billiard::search

namespace billiard::search::cuda {

struct SearchOptions {
    int device = 0;
    int breaks = 2;                  // BREAKS, billiard_search.cpp:15
    uint64_t maxDurationMs = 10000;  // ProcessManager::process default, process.hpp:80
    size_t batch = 32;               // SIMULATION nodes collected before a device launch: the planner runs ahead this far, so
                                     // a request for few solutions plans (and simulates) at most this much more than the
                                     // reference's best-first order would before "enough solutions" ends it
    int prune = 0;                   // bcu_set_prune mode (0: every quartic the reference solves is solved)
    bool devicePlanner = false;      // plan on the device as well (bcu_set_planner / bcu_plan / bcu_evaluate_plan): break by break,
                                     // every state of a break planned and simulated in one call each, nothing of the
                                     // reference's planner on the host.  Exhaustive per break (what the reference's best-first
                                     // search converges to); `solutions` does not end it early, the time budget and a newer
                                     // request do (between breaks).
};

struct SearchReport {   // what the last finished request did (diagnosis / tests)
    size_t expansions = 0, simulations = 0, batches = 0, solutions = 0;
    bool enoughSolutions = false, noTimeLeft = false, noWork = false, cancelled = false;
    double ms = 0.0;
    std::vector<uint64_t> costs;   // accumulated cost of every solution returned, in the order of the result
};

// == process::ProcessManager<std::shared_ptr<SearchNode>, SearchState, std::vector<node::System>> for this path
// (lib/process/include/process/process.hpp): process() returns a future and replaces the request in flight (:79-98: the
// replaced request's promise is dropped, so its future reports std::future_errc::broken_promise exactly as the
// reference's does); one worker thread pops nodes in ascending cost (:19-25), SEARCH nodes are expanded by the reference's
// planner, SIMULATION nodes are collected and evaluated on the device `batch` at a time; the request ends when enough
// solutions are final, its time is up or no work is left (:188-198), and all solutions found come back ascending.
// "Enough" is exact here: the request ends once `solutions` solutions are final AND no open node is cheaper than the
// last of them (costs only grow along a path), so the first `solutions` entries are the cheapest that exist -- what the
// reference's best-first order approaches but, with worker threads syncing every 100 ms, does not guarantee.
class SearchFront {
public:
    using Result = std::vector<std::vector<node::System>>;
    using NodeResult = std::vector<std::shared_ptr<SearchNode>>;

    explicit SearchFront(SearchOptions options = {}) : _options(options), _worker([this] { work(); }) {}
    ~SearchFront() {
        {
            std::lock_guard<std::mutex> guard(_lock);
            _exit = true;
            _cancel = true;
        }
        _wake.notify_all();
        _worker.join();
    }
    SearchFront(const SearchFront&) = delete;
    SearchFront& operator=(const SearchFront&) = delete;

    std::future<Result> process(const State& state, const Search& search, uint16_t solutions, const Configuration& config) {
        auto request = std::make_unique<Request>(Request{state, search, solutions, config, std::promise<Result>{}});
        std::future<Result> future = request->result.get_future();
        {
            std::lock_guard<std::mutex> guard(_lock);
            _pending = std::move(request);   // a request still waiting is dropped here (broken promise), one in flight is cancelled
            _cancel = true;
        }
        _wake.notify_all();
        return future;
    }

    // == billiard::search::searchOnly (billiard_search.hpp:13-15, src :122-139): the solution SearchNodes themselves, ascending
    // by cost; each node's path carries its SIMULATION nodes with their Systems filled in (host-planner mode only)
    std::future<NodeResult> processNodes(const State& state, const Search& search, uint16_t solutions, const Configuration& config) {
        auto request = std::make_unique<Request>(Request{state, search, solutions, config, std::promise<Result>{}, true, std::promise<NodeResult>{}});
        std::future<NodeResult> future = request->nodes.get_future();
        {
            std::lock_guard<std::mutex> guard(_lock);
            _pending = std::move(request);
            _cancel = true;
        }
        _wake.notify_all();
        return future;
    }

    SearchReport lastReport() {
        std::lock_guard<std::mutex> guard(_lock);
        return _report;
    }

private:
    struct Request {
        State state;
        Search search;
        uint16_t solutions;
        Configuration config;
        std::promise<Result> result;
        bool wantNodes = false;       // searchOnly(): the solution nodes instead of their Systems
        std::promise<NodeResult> nodes;
    };
    struct Open {
        std::shared_ptr<SearchNode> node;
        uint64_t order;
    };
    struct Later {   // process::Comparator (ascending _cost), creation order breaks ties: deterministic
        bool operator()(const Open& a, const Open& b) const {
            return a.node->_cost != b.node->_cost ? a.node->_cost > b.node->_cost : a.order > b.order;
        }
    };
    struct Frame {   // the table one break is played on: the _state of the break's root SEARCH node, in a fixed slot order
        State state{std::vector<Ball>{}};
        std::vector<std::string> searched;   // getTypeOf(search, state), billiard_search.cpp:1994-2000
    };

    void work() {
        for (;;) {
            std::unique_ptr<Request> request;
            {
                std::unique_lock<std::mutex> guard(_lock);
                _wake.wait(guard, [this] { return _exit || _pending; });
                if (_exit) return;
                request = std::move(_pending);
                _cancel = false;
            }
            try {
                SearchReport report;
                _solutionNodes.clear();
                Result result = run(*request, report);
                std::lock_guard<std::mutex> guard(_lock);
                _report = report;
                if (!report.cancelled) {
                    if (request->wantNodes) request->nodes.set_value(std::move(_solutionNodes));
                    else request->result.set_value(std::move(result));
                }
                // cancelled: the promise dies with the request -> broken_promise, as in the reference
            } catch (...) {
                if (request->wantNodes) request->nodes.set_exception(std::current_exception());
                else request->result.set_exception(std::current_exception());
            }
        }
    }

    bool cancelled() {
        std::lock_guard<std::mutex> guard(_lock);
        return _cancel;
    }

    // the engine is kept across requests and rebuilt when the table or the ball types change
    Engine& engineFor(const Configuration& config, const std::vector<std::string>& typeNames) {
        std::vector<float> key{config._ball._radius};
        for (auto& kv : config._table._rails) {
            const Rail& r = kv.second;
            key.insert(key.end(), {r._normal.x, r._normal.y, r._shiftedStart.x, r._shiftedStart.y, r._shiftedEnd.x, r._shiftedEnd.y});
        }
        for (auto& p : config._table._pockets) key.insert(key.end(), {p._position.x, p._position.y, p._radius});
        RulesTable rules = RulesTable::fromConfiguration(config, typeNames);
        for (double sc : rules.scores) key.push_back((float)sc);
        if (!_engine || key != _engineKey || typeNames != _engine->rules().typeNames) {
            _engine.reset();
            _engine = std::make_unique<Engine>(config, rules, _options.device);
            _engine->checked(bcu_set_prune(_engine->handle(), _options.prune), "bcu_set_prune");
            _engineKey = key;
        }
        return *_engine;
    }

    static std::vector<std::string> typeNamesOf(const State& state) {
        std::vector<std::string> names{"WHITE"};
        for (auto& b : state._balls)
            if (std::find(names.begin(), names.end(), b._type) == names.end()) names.push_back(b._type);
        std::sort(names.begin() + 1, names.end());
        return names;
    }

    static const SearchNode* breakRoot(const SearchNode* simulation) {
        const SearchNode* node = simulation->_parent.get();
        while (node->_parent && node->_parent->_type == SearchNodeType::SEARCH) node = node->_parent.get();
        return node;
    }

    Result run(Request& rq, SearchReport& report) {
        using clock = std::chrono::steady_clock;
        const auto t0 = clock::now();
        auto elapsedMs = [&] { return std::chrono::duration<double, std::milli>(clock::now() - t0).count(); };
        Result result;
        if (!isValidSearch(rq.state, rq.search)) return result;   // billiard_search.cpp:78-84
        const Configuration& config = rq.config;
        Engine& engine = engineFor(config, typeNamesOf(rq.state));
        const RulesTable& rules = engine.rules();
        if (_options.devicePlanner) return runPlanned(rq, report, engine);
        auto searchState = std::make_shared<SearchState>(SearchState{config});   // :92-94
        searchState->_ball._accelerationLength = physics::accelerationLength();
        searchState->_ball._slideAccelerationLength = physics::slideAccelerationLength();

        std::priority_queue<Open, std::vector<Open>, Later> open;
        uint64_t order = 0;
        for (auto& node : addPocketsAsInitialTargets(rq.state, rq.search, config)) open.push(Open{node, order++});
        std::vector<Open> solutions;   // SIMULATION nodes with _isSolution, _cost = accumulated cost
        std::unordered_map<const SearchNode*, std::shared_ptr<Frame>> frames;   // by break root
        auto frameOf = [&](const SearchNode* root) -> std::shared_ptr<Frame> {
            auto it = frames.find(root);
            if (it != frames.end()) return it->second;
            auto frame = std::make_shared<Frame>();
            auto asSearch = root->asSearch();
            for (auto& kv : asSearch->_state) frame->state._balls.push_back(kv.second);
            std::sort(frame->state._balls.begin(), frame->state._balls.end(), [](const Ball& a, const Ball& b) { return a._id < b._id; });
            if (!asSearch->_search._types.empty()) frame->searched = asSearch->_search._types;
            else if (asSearch->_state.count(asSearch->_search._id)) frame->searched = {asSearch->_state.at(asSearch->_search._id)._type};
            frames.insert({root, frame});
            return frame;
        };
        auto kthCost = [&]() -> uint64_t {   // cost of the `solutions`-th cheapest solution so far
            if (rq.solutions == 0) return 0ull;   // the reference's `_solutions.size() >= _minimalSolutions` holds at once
            if (solutions.size() < rq.solutions) return ~0ull;
            std::vector<uint64_t> costs;
            for (auto& s : solutions) costs.push_back(s.node->_cost);
            std::nth_element(costs.begin(), costs.begin() + (rq.solutions - 1), costs.end());
            return costs[rq.solutions - 1];
        };

        std::vector<std::shared_ptr<SearchNode>> batch;
        for (;;) {
            // ---- the planner runs ahead: SEARCH nodes are expanded in ascending cost, SIMULATION nodes collected ------------
            batch.clear();
            const uint64_t bound = kthCost();
            while (!open.empty() && batch.size() < _options.batch) {
                if ((report.expansions & 63) == 0) {
                    if (cancelled()) { report.cancelled = true; break; }
                    if (elapsedMs() >= (double)_options.maxDurationMs) { report.noTimeLeft = true; break; }
                }
                if (open.top().node->_cost >= bound) { report.enoughSolutions = true; break; }
                auto node = open.top().node;
                open.pop();
                report.expansions++;
                if (node->_type == SearchNodeType::SEARCH) {
                    for (auto& child : expand(node, searchState)) open.push(Open{child, order++});
                } else {
                    batch.push_back(node);
                }
            }
            if (report.cancelled) return result;
            if (batch.empty()) {
                report.noWork = open.empty() && !report.noTimeLeft && !report.enoughSolutions;
                break;
            }
            // ---- one device launch for the batch (billiard_search.cpp:1358-1405 per node) -------------------------------------
            std::vector<std::shared_ptr<Frame>> batchFrames;
            std::unordered_map<const Frame*, int> slotOf;
            std::vector<int32_t> stateIdx;
            int slots = 0;
            for (auto& node : batch) {
                auto frame = frameOf(breakRoot(node.get()));
                auto it = slotOf.find(frame.get());
                if (it == slotOf.end()) {
                    it = slotOf.insert({frame.get(), (int)batchFrames.size()}).first;
                    batchFrames.push_back(frame);
                    slots = std::max(slots, (int)frame->state._balls.size());
                }
                stateIdx.push_back(it->second);
            }
            if (slots > BCU_MAX_BALLS) throw std::runtime_error("billiard::search::cuda: more than 32 balls");
            const size_t nF = batchFrames.size(), m = batch.size();
            std::vector<float> pos(nF * slots * 2, 0.0f);
            std::vector<int32_t> types(nF * slots, -1);
            std::vector<uint32_t> mask(nF, 0u);
            for (size_t f = 0; f < nF; ++f) {
                const Frame& frame = *batchFrames[f];
                for (size_t b = 0; b < frame.state._balls.size(); ++b) {
                    pos[(f * slots + b) * 2] = frame.state._balls[b]._position.x;
                    pos[(f * slots + b) * 2 + 1] = frame.state._balls[b]._position.y;
                    types[f * slots + b] = rules.code(frame.state._balls[b]._type);
                }
                for (auto& t : frame.searched) { const int c = rules.code(t); if (c >= 0) mask[f] |= 1u << c; }
            }
            std::vector<float> vx(m), vy(m), ftime(m), fpos(m * slots * 2);
            std::vector<uint64_t> sc(m), bc(m), cost(m);
            std::vector<uint32_t> flags(m), potted(m);
            for (size_t i = 0; i < m; ++i) {
                // the strike is the BALL_SHOT node of the input layer the planner made (prepareForSimulation :797-826)
                for (auto& kv : batch[i]->asSimulation()->_simulation.firstLayer()._nodes)
                    if (kv.second._type == node::NodeType::BALL_SHOT) { vx[i] = kv.second.after()->_velocity.x; vy[i] = kv.second.after()->_velocity.y; }
                sc[i] = batch[i]->_searchCost;
                bc[i] = batch[i]->_cost;
            }
            bcu_engine* e = engine.handle();
            engine.checked(bcu_set_states(e, (int32_t)nF, slots, pos.data(), types.data(), mask.data(), nullptr), "bcu_set_states");
            bcu_results out{cost.data(), flags.data(), potted.data(), ftime.data(), fpos.data(), nullptr};
            engine.checked(bcu_evaluate(e, (int64_t)m, stateIdx.data(), vx.data(), vy.data(), sc.data(), bc.data(), &out), "bcu_evaluate");
            report.simulations += m;
            report.batches++;
            // ---- results, node by node (:1388-1462) --------------------------------------------------------------------------
            for (size_t i = 0; i < m; ++i) {
                if (!(flags[i] & BCU_F_COMPLETED) || !(flags[i] & BCU_F_VALID)) continue;   // discarded / "End state is not valid"
                auto& node = batch[i];
                const Frame& frame = *batchFrames[stateIdx[i]];
                node->asSimulation()->_cost = (double)(cost[i] - node->_cost);   // simCost (:1403-1405)
                int simCount = 0;
                for (const SearchNode* p = node.get(); p; p = p->_parent.get()) simCount += p->_type == SearchNodeType::SIMULATION;
                bool solution = true;
                if (config._depthSearchEnabled && simCount < _options.breaks) {
                    std::vector<Ball> rest;   // the balls at rest in the final layer, with the types of the frame (:1410-1426)
                    std::unordered_map<std::string, std::string> idsByType;
                    for (size_t b = 0; b < frame.state._balls.size(); ++b) {
                        const Ball& ball = frame.state._balls[b];
                        if (!((potted[i] >> b) & 1u))
                            rest.emplace_back(glm::vec2{fpos[(i * slots + b) * 2], fpos[(i * slots + b) * 2 + 1]}, ball._type, ball._id);
                        if (!idsByType.count(ball._type)) idsByType.insert({ball._type, ball._id});
                    }
                    State nextState{rest};
                    Search nextSearch = config._rules._nextSearch(nextState, frame.searched);
                    State modified = config._rules._modifyState(nextState, idsByType);
                    if (modified._balls.empty()) continue;   // "could not replace all balls!" (:1432-1435)
                    bool hasTypeToSearch = false;
                    for (auto& ball : modified._balls)
                        if ((hasTypeToSearch = std::count(nextSearch._types.begin(), nextSearch._types.end(), ball._type) > 0)) break;
                    if (hasTypeToSearch) {
                        solution = false;
                        for (auto& root : addPocketsAsInitialTargets(modified, nextSearch, config)) {
                            root->_parent = node;
                            root->_cost = cost[i];
                            root->_searchCost = 0;
                            open.push(Open{root, order++});
                        }
                    }
                }
                if (solution) {
                    node->_isSolution = true;
                    node->_cost = cost[i];
                    solutions.push_back(Open{node, order++});
                }
            }
            if (cancelled()) { report.cancelled = true; return result; }
            if (elapsedMs() >= (double)_options.maxDurationMs) { report.noTimeLeft = true; break; }
        }

        // ---- mapToSolution (:185-222): one node::System per break of every solution, ascending ------------------------------
        std::sort(solutions.begin(), solutions.end(), [](const Open& a, const Open& b) {
            return a.node->_cost != b.node->_cost ? a.node->_cost < b.node->_cost : a.order < b.order;
        });
        report.solutions = solutions.size();
        for (auto& s : solutions) report.costs.push_back(s.node->_cost);
        if (!solutions.empty()) {
            // all strikes of all solutions are traced in one device call; a SIMULATION node shared by several solutions once
            std::vector<SearchNode*> simNodes;
            std::unordered_map<SearchNode*, int> seen;
            for (auto& s : solutions)
                for (SearchNode* p = s.node.get(); p; p = p->_parent.get())
                    if (p->_type == SearchNodeType::SIMULATION && !seen.count(p)) { seen.insert({p, (int)simNodes.size()}); simNodes.push_back(p); }
            std::vector<std::shared_ptr<Frame>> traceFrames;
            std::unordered_map<const Frame*, int> slotOf;
            std::vector<int32_t> stateIdx;
            int slots = 0;
            for (SearchNode* p : simNodes) {
                auto frame = frameOf(breakRoot(p));
                auto it = slotOf.find(frame.get());
                if (it == slotOf.end()) {
                    it = slotOf.insert({frame.get(), (int)traceFrames.size()}).first;
                    traceFrames.push_back(frame);
                    slots = std::max(slots, (int)frame->state._balls.size());
                }
                stateIdx.push_back(it->second);
            }
            const size_t nF = traceFrames.size();
            std::vector<float> pos(nF * slots * 2, 0.0f);
            std::vector<int32_t> types(nF * slots, -1);
            std::vector<uint32_t> mask(nF, 0xffffffffu);
            for (size_t f = 0; f < nF; ++f)
                for (size_t b = 0; b < traceFrames[f]->state._balls.size(); ++b) {
                    pos[(f * slots + b) * 2] = traceFrames[f]->state._balls[b]._position.x;
                    pos[(f * slots + b) * 2 + 1] = traceFrames[f]->state._balls[b]._position.y;
                    types[f * slots + b] = rules.code(traceFrames[f]->state._balls[b]._type);
                }
            bcu_engine* e = engine.handle();
            engine.checked(bcu_set_states(e, (int32_t)nF, slots, pos.data(), types.data(), mask.data(), nullptr), "bcu_set_states");
            const size_t recPer = (size_t)BCU_MAX_LAYERS * slots * BCU_TRACE_REC;
            for (size_t first = 0; first < simNodes.size(); first += 4096) {   // bcu_trace_strikes takes up to 65 536; 4096 bounds the host buffers
                const size_t m = std::min<size_t>(4096, simNodes.size() - first);
                std::vector<float> vx(m), vy(m), times(m * BCU_MAX_LAYERS), rec(m * recPer);
                std::vector<int32_t> layers(m);
                for (size_t i = 0; i < m; ++i)
                    for (auto& kv : simNodes[first + i]->asSimulation()->_simulation.firstLayer()._nodes)
                        if (kv.second._type == node::NodeType::BALL_SHOT) { vx[i] = kv.second.after()->_velocity.x; vy[i] = kv.second.after()->_velocity.y; }
                engine.checked(bcu_trace_strikes(e, (int32_t)m, stateIdx.data() + first, vx.data(), vy.data(), layers.data(), times.data(),
                                                 rec.data()), "bcu_trace_strikes");
                for (size_t i = 0; i < m; ++i) {
                    auto system = engine.systemFromTrace(traceFrames[stateIdx[first + i]]->state, layers[i], &times[i * BCU_MAX_LAYERS],
                                                         &rec[i * recPer], slots);
                    if (!system) throw std::runtime_error("billiard::search::cuda: a solution's strike did not re-simulate");
                    auto simulation = simNodes[first + i]->asSimulation();
                    system->_id = simulation->_simulation._id;
                    simulation->_simulation = std::move(*system);
                }
            }
            for (auto& s : solutions) result.push_back(s.node->allSimulations());   // type.hpp:525-537
            for (auto& s : solutions) _solutionNodes.push_back(s.node);
        }
        report.ms = elapsedMs();
        return result;
    }

    // The same search with the planner on the device: level-synchronous over breaks.  A strike of break b is one entry of
    // `strikes` (its frame, velocity, accumulated cost, the strike of break b - 1 it follows).
    Result runPlanned(Request& rq, SearchReport& report, Engine& engine) {
        using clock = std::chrono::steady_clock;
        const auto t0 = clock::now();
        auto elapsedMs = [&] { return std::chrono::duration<double, std::milli>(clock::now() - t0).count(); };
        const Configuration& config = rq.config;
        const RulesTable& rules = engine.rules();
        bcu_engine* e = engine.handle();
        Result result;
        PlannerTables tables = PlannerTables::from(config);
        engine.checked(bcu_set_planner(e, &tables.desc), "bcu_set_planner");
        struct PlannedFrame { State state{std::vector<Ball>{}}; std::vector<std::string> searched; uint64_t cost = 0; int parentStrike = -1; };
        struct PlannedStrike { int frame; glm::vec2 velocity; uint64_t cost; int parent; };
        std::vector<PlannedFrame> allFrames;       // every frame of every break (index = id)
        std::vector<PlannedStrike> strikes;        // every valid strike
        std::vector<std::pair<uint64_t, int>> solutions;   // (cost, last strike)
        std::vector<int> level;                    // frames of the current break
        {
            PlannedFrame root;
            root.state = rq.state;
            if (!rq.search._types.empty()) root.searched = rq.search._types;
            else for (auto& b : rq.state._balls) if (b._id == rq.search._id) root.searched = {b._type};   // getTypeOf (:1994-2000)
            allFrames.push_back(root);
            level.push_back(0);
        }
        const int breaks = config._depthSearchEnabled ? _options.breaks : 1;
        auto upload = [&](const std::vector<int>& frames, int& slots, bool withMask) {
            slots = 0;
            for (int f : frames) slots = std::max(slots, (int)allFrames[f].state._balls.size());
            if (slots > BCU_MAX_BALLS) throw std::runtime_error("billiard::search::cuda: more than 32 balls");
            const size_t nF = frames.size();
            std::vector<float> pos(nF * slots * 2, 0.0f);
            std::vector<int32_t> types(nF * slots, -1);
            std::vector<uint32_t> mask(nF, withMask ? 0u : 0xffffffffu);
            std::vector<uint64_t> base(nF, 0);
            for (size_t i = 0; i < nF; ++i) {
                const PlannedFrame& fr = allFrames[frames[i]];
                for (size_t b = 0; b < fr.state._balls.size(); ++b) {
                    pos[(i * slots + b) * 2] = fr.state._balls[b]._position.x;
                    pos[(i * slots + b) * 2 + 1] = fr.state._balls[b]._position.y;
                    types[i * slots + b] = rules.code(fr.state._balls[b]._type);
                }
                if (withMask) for (auto& t : fr.searched) { const int c = rules.code(t); if (c >= 0) mask[i] |= 1u << c; }
                base[i] = fr.cost;
            }
            engine.checked(bcu_set_states(e, (int32_t)nF, slots, pos.data(), types.data(), mask.data(), base.data()), "bcu_set_states");
        };
        for (int brk = 0; brk < breaks && !level.empty(); ++brk) {
            if (cancelled()) { report.cancelled = true; return result; }
            if (elapsedMs() >= (double)_options.maxDurationMs) { report.noTimeLeft = true; break; }
            int slots = 0;
            upload(level, slots, true);
            int64_t n = 0;
            engine.checked(bcu_plan(e, &n), "bcu_plan");
            report.batches++;
            if (n == 0) { level.clear(); break; }
            std::vector<int32_t> sidx((size_t)n);
            std::vector<float> vx((size_t)n), vy((size_t)n), ftime((size_t)n), fpos((size_t)n * slots * 2);
            std::vector<uint64_t> cost((size_t)n);
            std::vector<uint32_t> flags((size_t)n), potted((size_t)n);
            engine.checked(bcu_fetch_plan(e, 0, n, sidx.data(), vx.data(), vy.data(), nullptr, nullptr), "bcu_fetch_plan");
            bcu_results out{cost.data(), flags.data(), potted.data(), ftime.data(), fpos.data(), nullptr};
            engine.checked(bcu_evaluate_plan(e, &out), "bcu_evaluate_plan");
            report.simulations += (size_t)n;
            std::vector<int> next;
            for (int64_t i = 0; i < n; ++i) {
                if (!(flags[i] & BCU_F_COMPLETED) || !(flags[i] & BCU_F_VALID)) continue;
                const int frameId = level[sidx[i]];
                const PlannedFrame& fr = allFrames[frameId];
                const int strikeId = (int)strikes.size();
                strikes.push_back(PlannedStrike{frameId, glm::vec2{vx[i], vy[i]}, cost[i], fr.parentStrike});
                bool solution = true;
                if (config._depthSearchEnabled && brk + 1 < breaks) {   // simCount < BREAKS (:1407)
                    std::vector<Ball> rest;
                    std::unordered_map<std::string, std::string> idsByType;
                    for (size_t b = 0; b < fr.state._balls.size(); ++b) {
                        const Ball& ball = fr.state._balls[b];
                        if (!((potted[i] >> b) & 1u))
                            rest.emplace_back(glm::vec2{fpos[((size_t)i * slots + b) * 2], fpos[((size_t)i * slots + b) * 2 + 1]}, ball._type, ball._id);
                        if (!idsByType.count(ball._type)) idsByType.insert({ball._type, ball._id});
                    }
                    State nextState{rest};
                    Search nextSearch = config._rules._nextSearch(nextState, fr.searched);
                    State modified = config._rules._modifyState(nextState, idsByType);
                    if (modified._balls.empty()) { strikes.pop_back(); continue; }   // "could not replace all balls!" (:1432-1435)
                    bool hasTypeToSearch = false;
                    for (auto& ball : modified._balls)
                        if ((hasTypeToSearch = std::count(nextSearch._types.begin(), nextSearch._types.end(), ball._type) > 0)) break;
                    if (hasTypeToSearch) {
                        solution = false;
                        PlannedFrame nf;
                        nf.state = modified;
                        std::sort(nf.state._balls.begin(), nf.state._balls.end(), [](const Ball& a, const Ball& b) { return a._id < b._id; });
                        nf.searched = nextSearch._types;
                        nf.cost = cost[i];
                        nf.parentStrike = strikeId;
                        next.push_back((int)allFrames.size());
                        allFrames.push_back(std::move(nf));
                    }
                }
                if (solution) solutions.push_back({cost[i], strikeId});
            }
            level = std::move(next);
        }
        report.noWork = !report.noTimeLeft;
        std::stable_sort(solutions.begin(), solutions.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
        report.solutions = solutions.size();
        for (auto& s : solutions) report.costs.push_back(s.first);
        if (!solutions.empty()) {
            // every strike on the path of a solution, traced once (one frame table, one bcu_trace_strikes call per 4096)
            std::vector<int> needed;
            std::unordered_map<int, int> slotOfStrike;
            for (auto& s : solutions)
                for (int k = s.second; k >= 0; k = strikes[k].parent)
                    if (!slotOfStrike.count(k)) { slotOfStrike.insert({k, (int)needed.size()}); needed.push_back(k); }
            std::vector<int> frames;
            std::unordered_map<int, int> slotOfFrame;
            for (int k : needed)
                if (!slotOfFrame.count(strikes[k].frame)) { slotOfFrame.insert({strikes[k].frame, (int)frames.size()}); frames.push_back(strikes[k].frame); }
            int slots = 0;
            upload(frames, slots, false);
            const size_t recPer = (size_t)BCU_MAX_LAYERS * slots * BCU_TRACE_REC;
            std::vector<std::optional<node::System>> systems(needed.size());
            for (size_t first = 0; first < needed.size(); first += 4096) {
                const size_t m = std::min<size_t>(4096, needed.size() - first);
                std::vector<int32_t> sidx(m), layers(m);
                std::vector<float> vx(m), vy(m), times(m * BCU_MAX_LAYERS), rec(m * recPer);
                for (size_t i = 0; i < m; ++i) {
                    const PlannedStrike& st = strikes[needed[first + i]];
                    sidx[i] = slotOfFrame.at(st.frame); vx[i] = st.velocity.x; vy[i] = st.velocity.y;
                }
                engine.checked(bcu_trace_strikes(e, (int32_t)m, sidx.data(), vx.data(), vy.data(), layers.data(), times.data(), rec.data()),
                               "bcu_trace_strikes");
                for (size_t i = 0; i < m; ++i)
                    systems[first + i] = engine.systemFromTrace(allFrames[strikes[needed[first + i]].frame].state, layers[i],
                                                                &times[i * BCU_MAX_LAYERS], &rec[i * recPer], slots);
            }
            for (auto& s : solutions) {
                std::vector<node::System> path;
                for (int k = s.second; k >= 0; k = strikes[k].parent) {
                    auto& sys = systems[slotOfStrike.at(k)];
                    if (!sys) throw std::runtime_error("billiard::search::cuda: a solution's strike did not re-simulate");
                    path.push_back(*sys);
                }
                std::reverse(path.begin(), path.end());   // allSimulations(): first break first (type.hpp:525-537)
                result.push_back(std::move(path));
            }
        }
        report.ms = elapsedMs();
        return result;
    }

    SearchOptions _options;
    std::mutex _lock;
    std::condition_variable _wake;
    std::unique_ptr<Request> _pending;
    bool _cancel = false, _exit = false;
    SearchReport _report;
    std::unique_ptr<Engine> _engine;
    std::vector<float> _engineKey;
    NodeResult _solutionNodes;   // of the request being worked on (worker thread only)
    std::thread _worker;   // last: started when everything above exists
};

// == billiard::search::searchOnly(state, search, solutions, config), billiard_search.hpp:13-15
inline std::future<std::vector<std::shared_ptr<SearchNode>>> searchOnly(const State& state, const Search& search, uint16_t solutions,
                                                                        const Configuration& config) {
    static SearchFront front{};
    return front.processNodes(state, search, solutions, config);
}

// == billiard::search::search(state, search, solutions, config), billiard_search.hpp:9-12.  Like the reference's, the
// manager behind it is a function-local static that lives until the process exits (billiard_search.cpp:86-90).
inline std::future<std::vector<std::vector<node::System>>> search(const State& state, const Search& search, uint16_t solutions,
                                                                   const Configuration& config) {
    static SearchFront front{};
    return front.process(state, search, solutions, config);
}

}  // namespace billiard::search::cuda
