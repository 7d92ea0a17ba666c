// ORACLE / TEST INFRASTRUCTURE ONLY -- never linked into, loaded by, or called from the product.
//
// C-ABI driver around the UNMODIFIED reference sources, compiled where they lie under
// /root/reference (see oracle/Makefile): lib/billiard_physics/src/billiard_physics.cpp,
// lib/billiard_search/src/{billiard_search,type}.cpp, lib/billiard_debug/src/billiard_debug.cpp.
// The only stand-in is oracle/glm_standin (glm 0.9.9.8 is not vendored in the reference).
//
// What it exposes (all POD, for ctypes / bench.py):
//   ref_set_table / ref_set_rules  -- build billiard::search::Configuration the way
//                                     test/billiard_search/src/util.cpp:24-80 does
//   ref_simulate_trace             -- public billiard::search::simulate (billiard_search.cpp:141-175),
//                                     full per-layer dump
//   ref_evaluate                   -- internal simulate(node,state,mustBeValid) (:1358-1471) +
//                                     simulationCost (:1473-1532) + rules, batched, multi-threaded
//   ref_plan                       -- deterministic best-first over the reference's expand() (:224-244)
//                                     to exhaustion: the reference's own candidate list + solution costs
//   ref_quartic / ref_time_to_collision -- physics known-answer entry points
//
// The snooker rules (lib/billiard_snooker/src/billiard_snooker.cpp:750-1153) cannot be compiled
// here (their translation unit needs OpenCV C++ headers), so validEndState (:1128-1153) and
// scoreForPottedBall (:787-790) are restated below as lambdas over caller-supplied tables.
#include <billiard_physics/billiard_physics.hpp>
#include <billiard_search/billiard_search.hpp>

#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstring>
#include <queue>
#include <string>
#include <thread>
#include <vector>

namespace billiard::search {
// external-linkage internals of billiard_search.cpp (declared there at :42-49)
std::vector<std::shared_ptr<SearchNode>> expand(const std::shared_ptr<SearchNode>& input,
                                                const std::shared_ptr<SearchState>& config);
std::vector<std::shared_ptr<SearchNode>> simulate(const std::shared_ptr<SearchNode>& input,
                                                  const std::shared_ptr<SearchState>& config, bool mustBeValid);
std::vector<std::shared_ptr<SearchNode>> addPocketsAsInitialTargets(const State& state, const Search& search,
                                                                    const Configuration& config);
node::Layer toInputLayer(const std::shared_ptr<SearchNode>& input, const std::string& cueBallId, const glm::vec2 force,
                         float accelerationLength, float slideAccelerationLength);
}  // namespace billiard::search

using namespace billiard;
using namespace billiard::search;

namespace {

struct Ctx {
    Configuration conf;
    std::vector<std::string> railIds;    // index -> id
    std::vector<std::string> pocketIds;  // index -> id
    std::vector<std::string> typeNames;  // type code -> name ("WHITE" must be one of them)
    std::vector<double> typeScore;       // type code -> score in [0,1]
};
Ctx g;

std::vector<std::string> splitLines(const char* s) {
    std::vector<std::string> out;
    std::string cur;
    for (const char* p = s; *p; ++p) {
        if (*p == '\n') { out.push_back(cur); cur.clear(); } else cur.push_back(*p);
    }
    if (!cur.empty()) out.push_back(cur);
    return out;
}

int indexOf(const std::vector<std::string>& v, const std::string& s) {
    auto it = std::find(v.begin(), v.end(), s);
    return it == v.end() ? -1 : int(it - v.begin());
}

std::string ballId(int i) { return "b" + std::to_string(i); }
int ballIndex(const std::string& id) { return id.empty() ? -1 : std::atoi(id.c_str() + 1); }

std::shared_ptr<SearchState> makeSearchState(const Configuration& conf) {
    // billiard_search.cpp:93-95
    auto ss = std::make_shared<SearchState>(SearchState{conf});
    ss->_ball._accelerationLength = physics::accelerationLength();
    ss->_ball._slideAccelerationLength = physics::slideAccelerationLength();
    return ss;
}

State makeState(int n, const float* pos, const int32_t* types) {
    std::vector<Ball> balls;
    for (int i = 0; i < n; ++i) {
        if (types[i] < 0) continue;  // absent ball
        balls.emplace_back(glm::vec2{pos[2 * i], pos[2 * i + 1]}, g.typeNames.at(types[i]), ballId(i));
    }
    return State(balls);
}

// Per-layer dump record, one per (layer, ball): 10 floats.
//  [0] node type (node::NodeType as float, -1 = ball absent)
//  [1..2] after.position  [3..4] after.velocity  [5] after.isRolling
//  [6..7] before.velocity [8] cause partner: other ball index / rail index / pocket index, else -1
//  [9] before.isRolling
constexpr int REC = 10;

void dumpSystem(const node::System& sys, int n, int maxLayers, int32_t* outNLayers, float* outTimes, float* outRec) {
    int L = std::min<int>(int(sys._layers.size()), maxLayers);
    *outNLayers = int(sys._layers.size());
    for (int l = 0; l < L; ++l) {
        const auto& layer = sys._layers[l];
        outTimes[l] = layer._time;
        for (int b = 0; b < n; ++b) {
            float* r = outRec + (size_t(l) * n + b) * REC;
            for (int k = 0; k < REC; ++k) r[k] = -1.0f;
            auto it = layer._nodes.find(ballId(b));
            if (it == layer._nodes.end()) continue;
            const auto& node = it->second;
            auto after = node.after();
            auto before = node.before();
            r[0] = float(int(node._type));
            r[1] = after->_position.x; r[2] = after->_position.y;
            r[3] = after->_velocity.x; r[4] = after->_velocity.y;
            r[5] = after->_isRolling ? 1.0f : 0.0f;
            r[6] = before->_velocity.x; r[7] = before->_velocity.y;
            r[9] = before->_isRolling ? 1.0f : 0.0f;
            r[8] = -1.0f;
            if (auto c = node.toBallCollision()) {
                int a = ballIndex(c->_cause._ball1), bb = ballIndex(c->_cause._ball2);
                r[8] = float(a == b ? bb : a);
            } else if (auto rc = node.toBallRailCollision()) {
                r[8] = float(indexOf(g.railIds, rc->_cause._rail));
            } else if (auto p = node.toPotted()) {
                r[8] = float(indexOf(g.pocketIds, p->_cause._pocket));
            }
        }
    }
}

}  // namespace

extern "C" {

// rails6: R x {start.x, start.y, end.x, end.y, shiftDirection.x, shiftDirection.y}
// pockets7: P x {position.x, position.y, radius, normal.x, normal.y, pottingPoint.x, pottingPoint.y}
// ids are '\n'-joined.  Mirrors test/billiard_search/src/util.cpp:24-80 (+ _searchPockets and
// minimalPocketVelocity, which that helper forgets; see type.cpp:43-54, configuration.json:156).
int ref_set_table(int R, const char* railIds, const float* rails6, int P, const char* pocketIds, const float* pockets7,
                  float ballRadius, float minimalPocketVelocity, float innerLength, float innerWidth) {
    g.conf = Configuration{};
    g.railIds = splitLines(railIds);
    g.pocketIds = splitLines(pocketIds);
    if (int(g.railIds.size()) != R || int(g.pocketIds.size()) != P) return -1;
    std::vector<Pocket> pockets;
    for (int i = 0; i < P; ++i) {
        const float* p = pockets7 + 7 * i;
        pockets.emplace_back(Pocket{g.pocketIds[i], PocketType::CORNER, glm::vec2{p[0], p[1]},
                                    std::vector<glm::vec2>{glm::vec2{p[5], p[6]}}, glm::vec2{p[3], p[4]}, p[2]});
    }
    std::unordered_map<std::string, Rail> rails;
    for (int i = 0; i < R; ++i) {
        const float* r = rails6 + 6 * i;
        rails.insert(std::make_pair(
            g.railIds[i], Rail{g.railIds[i], glm::vec2{r[0], r[1]}, glm::vec2{r[2], r[3]}, glm::vec2{r[4], r[5]}, ballRadius}));
    }
    g.conf._ball._radius = ballRadius;
    g.conf._ball._diameter = ballRadius * 2;
    g.conf._ball._diameterSquared = (g.conf._ball._radius * 2) * (g.conf._ball._radius * 2);
    g.conf._table._pockets = pockets;
    g.conf._table._rails = rails;
    g.conf._table.minimalPocketVelocity = minimalPocketVelocity;
    g.conf._table.diagonalLengthSquared = innerLength * innerLength + innerWidth * innerWidth;
    g.conf._searchPockets = buildSearchPocketsOfPockets(pockets);
    g.conf._depthSearchEnabled = false;
    g.conf._velocityMultiplyEnabled = true;
    return 0;
}

// Derived rail geometry as the reference computed it (type.cpp:56-81): per rail
// {normal.x, normal.y, shiftedStart.x, shiftedStart.y, shiftedEnd.x, shiftedEnd.y}
int ref_get_rail_derived(float* out6) {
    for (size_t i = 0; i < g.railIds.size(); ++i) {
        const Rail& r = g.conf._table._rails.at(g.railIds[i]);
        float* o = out6 + 6 * i;
        o[0] = r._normal.x; o[1] = r._normal.y;
        o[2] = r._shiftedStart.x; o[3] = r._shiftedStart.y;
        o[4] = r._shiftedEnd.x; o[5] = r._shiftedEnd.y;
    }
    return int(g.railIds.size());
}

// typeNames '\n'-joined; scores[i] = score for potting a ball of type i
// (billiard_snooker.cpp:751-759, 787-790: points / 7.0).
int ref_set_rules(const char* typeNames, const double* scores) {
    g.typeNames = splitLines(typeNames);
    g.typeScore.assign(scores, scores + g.typeNames.size());
    // restated billiard_snooker.cpp:1128-1153
    g.conf._rules._isValidEndState = [](const std::vector<std::string>& expectedTypes, const node::Layer& layer) {
        int potted = 0;
        for (auto& node : layer._nodes) {
            if (node.second._type == node::NodeType::BALL_POTTING) {
                if (node.second._ballType == "WHITE") return false;
                if (std::find(expectedTypes.begin(), expectedTypes.end(), node.second._ballType) == expectedTypes.end())
                    return false;
                potted++;
            }
        }
        return potted != 0;
    };
    // restated billiard_snooker.cpp:779-790
    g.conf._rules._scoreForPottedBall = [](const std::string& ballType) {
        int i = indexOf(g.typeNames, ballType);
        return i < 0 ? 0.0 : g.typeScore[i];
    };
    return 0;
}

// The Configuration built by ref_set_table / ref_set_rules, for test programs compiled against the reference's
// headers (tests/host_mirror_check.cpp).
const void* ref_get_configuration(void) { return &g.conf; }

void ref_set_toggles(int depthSearchEnabled, int velocityMultiplyEnabled) {
    g.conf._depthSearchEnabled = depthSearchEnabled != 0;
    g.conf._velocityMultiplyEnabled = velocityMultiplyEnabled != 0;
}

// Public simulate() (billiard_search.cpp:141-175).  Returns number of layers (0 = nullopt).
int ref_simulate_trace(int n, const float* pos, const int32_t* types, float vx, float vy, int maxLayers, float* outTimes,
                       float* outRec) {
    State st = makeState(n, pos, types);
    auto sys = simulate(st, glm::vec2{vx, vy}, g.conf);
    if (!sys) return 0;
    int32_t nl = 0;
    dumpSystem(*sys, n, maxLayers, &nl, outTimes, outRec);
    return nl;
}

enum : uint32_t { F_COMPLETED = 1u, F_VALID = 2u, F_OVERFLOW = 4u, F_NOEVENT = 8u };

// Batched candidate evaluation through the reference's internal simulate(node, state, mustBeValid=false)
// followed by the rules' end-state check.  Inputs mirror the product's C ABI (include/billiard_cuda.h).
//   states: nStates x n balls: pos[2], type code (-1 = absent); expectedMask[s]: bit t set = type t searched
//   candidates: stateIdx, vx, vy, searchCost, baseCost
// Outputs per candidate: cost (= baseCost + simulationCost; meaningful when COMPLETED), flags, nEvents,
//   pottedMask (bit b = ball b in BALL_POTTING at the end), finalTime, finalPos[n][2] (after-state of last layer).
// Optional trace of ONE candidate set is available via ref_simulate_trace.
int ref_evaluate(int nStates, int n, const float* statePos, const int32_t* stateTypes, const uint32_t* expectedMask,
                 int64_t nCand, const int32_t* candState, const float* candVx, const float* candVy,
                 const uint64_t* candSearchCost, const uint64_t* candBaseCost, int nThreads, uint64_t* outCost,
                 uint32_t* outFlags, int32_t* outNEvents, uint32_t* outPotted, float* outFinalTime, float* outFinalPos,
                 int32_t* outEventLog /* optional: nCand x 32 x 4 {type,a,b,timebits} */) {
    Configuration conf = g.conf;
    conf._depthSearchEnabled = false;  // one break per call; levels are driven by the caller
    auto ss = makeSearchState(conf);
    std::atomic<int64_t> next{0};
    auto worker = [&]() {
        const int64_t CH = 64;
        for (;;) {
            int64_t c0 = next.fetch_add(CH);
            if (c0 >= nCand) break;
            for (int64_t c = c0; c < std::min(nCand, c0 + CH); ++c) {
                int s = candState[c];
                const float* pos = statePos + size_t(s) * n * 2;
                const int32_t* types = stateTypes + size_t(s) * n;
                std::unordered_map<std::string, Ball> ballState;
                std::string cueId;
                for (int i = 0; i < n; ++i) {
                    if (types[i] < 0) continue;
                    Ball b{glm::vec2{pos[2 * i], pos[2 * i + 1]}, g.typeNames.at(types[i]), ballId(i)};
                    ballState.insert({b._id, b});
                    if (b._type == "WHITE") cueId = b._id;
                }
                std::vector<std::string> expected;
                for (size_t t = 0; t < g.typeNames.size(); ++t)
                    if (expectedMask[s] >> t & 1u) expected.push_back(g.typeNames[t]);

                auto parent = SearchNode::search();
                parent->asSearch()->_state = ballState;
                parent->asSearch()->_search = Search{"", expected};
                parent->_cost = candBaseCost[c];
                parent->_searchCost = candSearchCost[c];
                auto layer0 = toInputLayer(parent, cueId, glm::vec2{candVx[c], candVy[c]}, ss->_ball._accelerationLength,
                                           ss->_ball._slideAccelerationLength);
                auto simNode = SearchNode::simulation(parent);
                simNode->asSimulation()->_simulation.append(layer0);
                auto out = simulate(simNode, ss, false);

                const node::System& sys = simNode->asSimulation()->_simulation;
                uint32_t flags = 0;
                int nEvents = int(sys._layers.size()) - 1;
                if (!out.empty()) {
                    flags |= F_COMPLETED;
                    if (conf._rules._isValidEndState(expected, sys.lastLayer())) flags |= F_VALID;
                    outCost[c] = simNode->_cost;
                } else {
                    // distinguish the two discard paths of billiard_search.cpp:1376-1379 / 1464-1467
                    flags |= (nEvents >= 30) ? F_OVERFLOW : F_NOEVENT;
                    outCost[c] = ~uint64_t(0);
                }
                outFlags[c] = flags;
                outNEvents[c] = nEvents;
                uint32_t potted = 0;
                const auto& last = sys.lastLayer();
                for (int i = 0; i < n; ++i) {
                    float* fp = outFinalPos + (size_t(c) * n + i) * 2;
                    fp[0] = fp[1] = 0.0f;
                    auto it = last._nodes.find(ballId(i));
                    if (it == last._nodes.end()) continue;
                    if (it->second._type == node::NodeType::BALL_POTTING) potted |= 1u << i;
                    auto a = it->second.after();
                    fp[0] = a->_position.x; fp[1] = a->_position.y;
                }
                outPotted[c] = potted;
                outFinalTime[c] = last._time;
                if (outEventLog) {
                    int32_t* ev = outEventLog + size_t(c) * 32 * 4;
                    std::memset(ev, 0xff, 32 * 4 * sizeof(int32_t));
                    for (int l = 1; l < int(sys._layers.size()) && l < 32; ++l) {
                        const auto& L = sys._layers[l];
                        const auto& Pv = sys._layers[l - 1];
                        int type = -1, a = -1, b = -1;
                        for (int i = 0; i < n; ++i) {
                            auto it = L._nodes.find(ballId(i));
                            if (it == L._nodes.end()) continue;
                            const auto& nd = it->second;
                            auto prevType = Pv._nodes.at(ballId(i))._type;
                            if (nd._type == node::NodeType::BALL_COLLISION) {
                                type = event::EventType::BALL_COLLISION;
                                if (a < 0) a = i; else b = i;
                            } else if (nd._type == node::NodeType::BALL_RAIL_COLLISION) {
                                type = event::EventType::BALL_RAIL_COLLISION; a = i;
                                b = indexOf(g.railIds, nd.toBallRailCollision()->_cause._rail);
                            } else if (nd._type == node::NodeType::BALL_POTTING && prevType != node::NodeType::BALL_POTTING) {
                                type = event::EventType::BALL_POTTING; a = i;
                                b = indexOf(g.pocketIds, nd.toPotted()->_cause._pocket);
                            } else if (nd._type == node::NodeType::BALL_IN_REST && prevType != node::NodeType::BALL_IN_REST) {
                                type = event::EventType::BALL_IN_REST; a = i;
                            } else if (nd._type == node::NodeType::BALL_MOVING &&
                                       nd.before()->_isRolling != nd.after()->_isRolling) {
                                type = event::EventType::BALL_ROLLING; a = i;
                            }
                        }
                        float t = L._time;
                        int32_t tb;
                        std::memcpy(&tb, &t, 4);
                        ev[(l - 1) * 4 + 0] = type; ev[(l - 1) * 4 + 1] = a; ev[(l - 1) * 4 + 2] = b; ev[(l - 1) * 4 + 3] = tb;
                    }
                }
            }
        }
    };
    if (nThreads <= 1) {
        worker();
    } else {
        std::vector<std::thread> th;
        for (int t = 0; t < nThreads; ++t) th.emplace_back(worker);
        for (auto& t : th) t.join();
    }
    return 0;
}

// Deterministic best-first search over the reference's expand() to exhaustion (no ProcessManager, no
// wall clock): heap on (cost, creation order).  Records every SIMULATION node expanded.
//   out per simulation: vx, vy, searchCost, baseCost (node->_cost before simulating), resultCost (~0 if none)
// Single break unless depthSearchEnabled was set through ref_set_toggles.
int64_t ref_plan(int n, const float* pos, const int32_t* types, uint32_t expectedMask, int64_t maxOut, float* outVx,
                 float* outVy, uint64_t* outSearchCost, uint64_t* outBaseCost, uint64_t* outResultCost,
                 int32_t* outSimId, int64_t* outExpansions) {
    State st = makeState(n, pos, types);
    std::vector<std::string> expected;
    for (size_t t = 0; t < g.typeNames.size(); ++t)
        if (expectedMask >> t & 1u) expected.push_back(g.typeNames[t]);
    auto ss = makeSearchState(g.conf);
    struct Item { std::shared_ptr<SearchNode> node; uint64_t order; };
    auto cmp = [](const Item& a, const Item& b) {
        if (a.node->_cost != b.node->_cost) return a.node->_cost > b.node->_cost;
        return a.order > b.order;
    };
    std::priority_queue<Item, std::vector<Item>, decltype(cmp)> open(cmp);
    uint64_t order = 0;
    for (auto& nd : addPocketsAsInitialTargets(st, Search{"", expected}, g.conf)) open.push({nd, order++});
    int64_t nOut = 0, expansions = 0;
    while (!open.empty()) {
        auto item = open.top();
        open.pop();
        expansions++;
        auto nd = item.node;
        int64_t slot = -1;
        if (nd->_type == SearchNodeType::SIMULATION) {
            if (nOut < maxOut) {
                slot = nOut;
                const auto& first = nd->asSimulation()->_simulation.firstLayer();
                for (auto& kv : first._nodes) {
                    if (kv.second._type == node::NodeType::BALL_SHOT) {
                        auto a = kv.second.after();
                        outVx[slot] = a->_velocity.x; outVy[slot] = a->_velocity.y;
                    }
                }
                outSearchCost[slot] = nd->_searchCost;
                outBaseCost[slot] = nd->_cost;
                outResultCost[slot] = ~uint64_t(0);
                outSimId[slot] = int32_t(nd->asSimulation()->_simulation._id);
            }
            nOut++;
        }
        for (auto& e : expand(nd, ss)) {
            if (e->_isSolution) {
                if (slot >= 0) outResultCost[slot] = e->_cost;
            } else {
                open.push({e, order++});
            }
        }
    }
    if (outExpansions) *outExpansions = expansions;
    return nOut;
}

// Physics known-answer entry points (test/billiard_physics/src/billiard_physics_test.cpp:84-163).
int ref_quartic(double a, double b, double c, double d, double e, double* outReIm8) {
    auto roots = physics::intersection::solveQuarticFormula(a, b, c, d, e);
    for (size_t i = 0; i < roots.size() && i < 4; ++i) { outReIm8[2 * i] = roots[i].real(); outReIm8[2 * i + 1] = roots[i].imag(); }
    return int(roots.size());
}

int ref_time_to_collision(const float* a1, const float* v1, const float* p1, const float* a2, const float* v2,
                          const float* p2, float diameter, float* outTime) {
    auto t = physics::timeToCollision({a1[0], a1[1]}, {v1[0], v1[1]}, {p1[0], p1[1]}, {a2[0], a2[1]}, {v2[0], v2[1]},
                                      {p2[0], p2[1]}, diameter);
    if (!t) return 0;
    *outTime = *t;
    return 1;
}

// timeToCollision over a batch of ball pairs given (position, velocity, isRolling) -- builds the
// BallState (type.cpp:176-184) so the accelerations are the reference's.
int ref_time_to_collision_batch(int64_t n, const float* p1, const float* v1, const int32_t* roll1, const float* p2,
                                const float* v2, const int32_t* roll2, float diameter, float* outTime, int32_t* outHas) {
    float al = physics::accelerationLength(), sl = physics::slideAccelerationLength();
    for (int64_t i = 0; i < n; ++i) {
        state::BallState s1{glm::vec2{p1[2 * i], p1[2 * i + 1]}, glm::vec2{v1[2 * i], v1[2 * i + 1]}, al, sl, roll1[i] != 0};
        state::BallState s2{glm::vec2{p2[2 * i], p2[2 * i + 1]}, glm::vec2{v2[2 * i], v2[2 * i + 1]}, al, sl, roll2[i] != 0};
        auto t = physics::timeToCollision(s1._acceleration, s1._velocity, s1._position, s2._acceleration, s2._velocity,
                                          s2._position, diameter);
        outHas[i] = t ? 1 : 0;
        outTime[i] = t ? *t : -1.0f;
    }
    return 0;
}

int ref_quartic_batch(int64_t n, const double* coef5, double* outReIm8, int32_t* outN) {
    for (int64_t i = 0; i < n; ++i) {
        const double* c = coef5 + 5 * i;
        outN[i] = ref_quartic(c[0], c[1], c[2], c[3], c[4], outReIm8 + 8 * i);
    }
    return 0;
}

// Restates the test helper calculateInitialVelocity (test/billiard_search/src/simulation_vs_reality.cpp:3202-3223)
// on top of the reference's own solveQuadraticFormula / nonNegative; used only to generate fixtures.
void ref_calc_initial_velocity(float dx, float dy, int frameDuration, float* out2) {
    glm::vec2 traveledDistance{dx, dy};
    float duration = ((float) frameDuration) / 30.0f;
    const float g_ = physics::gravitationalAcceleration;
    const float mu_r = physics::frictionCoefficient;
    const float mu_g = physics::slideFrictionCoefficient;
    float s = glm::length(traveledDistance);
    float a = (-mu_r - 5.0f * mu_g)/(49.0f/2.0f * g_ * mu_g * mu_g) + 12.0f/(49.0f * g_ * mu_g);
    float b = (mu_r * duration)/(7.0f/2.0f * mu_g) + (5.0f * duration)/7.0f;
    float c = -0.5f * g_ * mu_r * duration * duration - s;
    auto result = physics::nonNegative(physics::intersection::solveQuadraticFormula(a, b, c));
    if (result.empty()) { out2[0] = 0; out2[1] = 0; return; }
    glm::vec2 velocity = result[0] * glm::normalize(traveledDistance);
    out2[0] = velocity.x; out2[1] = velocity.y;
}

}  // extern "C"
