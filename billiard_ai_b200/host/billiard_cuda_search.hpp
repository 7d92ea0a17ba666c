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
// Errors: a failing bcu_* call throws std::runtime_error with bcu_last_error(); there is no CPU fallback.
#pragma once

#include <billiard_physics/billiard_physics.hpp>
#include <billiard_search/type.hpp>

#include <cstdint>
#include <optional>
#include <stdexcept>
#include <string>
#include <unordered_map>
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
    std::optional<node::System> systemFromTrace(const State& state, int layers, const float* times, const float* rec) {
        const int n = (int)state._balls.size();
        if (layers == 0) return std::nullopt;  // billiard_search.cpp:170-174
        const float al = physics::accelerationLength(), sl = physics::slideAccelerationLength();
        node::System system;
        std::unordered_map<std::string, node::Node> previous;
        for (int l = 0; l < layers; ++l) {
            std::unordered_map<std::string, node::Node> nodes;
            for (int b = 0; b < n; ++b) {
                const float* r = &rec[((size_t)l * n + b) * BCU_TRACE_REC];
                const Ball& ball = state._balls[b];
                if (r[0] < 0.0f) continue;  // no such ball
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
