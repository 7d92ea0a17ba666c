// TEST INFRASTRUCTURE: checks the C++ host mirror (billiard_ai_b200/host/billiard_cuda_search.hpp) against the reference.
// Compiled in the build container against the reference's own headers (oracle/Makefile target `mirror`), linked with
// oracle/_ref/libref_oracle.so (the reference) and billiard_ai_b200/libbilliard_cuda.so (the product); the binary
// travels to the GPU box.  For each strike it calls
//     billiard::search::simulate(state, velocity, config)              -- the reference, CPU
//     billiard::search::cuda::Engine::simulate(state, velocity)        -- the drop-in, B200
// and compares the two node::System objects node by node (types, ball types, before/after BallState bits, causes,
// layer times, _nextRolling / _lastRailCollisions bookkeeping).
#include <billiard_search/billiard_search.hpp>

#include <cstdio>
#include <cstring>

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
