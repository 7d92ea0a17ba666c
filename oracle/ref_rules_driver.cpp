// ORACLE / TEST INFRASTRUCTURE ONLY -- never linked into or called by the product.
//
// The reference's rules between breaks, UNMODIFIED: billiard::snooker::nextSearch / stateAfterBreak and their helpers are
// lines 749-1126 of /root/reference/lib/billiard_snooker/src/billiard_snooker.cpp.  That translation unit as a whole needs
// OpenCV (camera-side code in the same file), which this image lacks, so oracle/Makefile cuts exactly those lines (and the
// Spot / SnookerSearchConfig declarations, billiard_snooker.hpp:124-135) into oracle/_ref/*.inc at build time -- generated,
// git-ignored, never committed -- and this driver compiles them inside the reference's own namespace against the
// reference's own billiard_search / billiard_physics headers.  Nothing of the rules is restated here: the flat C interface
// below only converts between slot arrays and the reference's State / Search types.
#include <billiard_debug/billiard_debug.hpp>
#include <billiard_physics/billiard_physics.hpp>
#include <billiard_search/type.hpp>

#include <algorithm>
#include <cstdint>
#include <optional>
#include <ostream>
#include <string>
#include <unordered_map>
#include <vector>

#define EXPORT_BILLIARD_SNOOKER_LIB
#define SPOT_REPLACE_RADIUS_MULTIPLIER 5   // billiard_snooker.hpp:15

namespace billiard::snooker {
#include "_ref/snooker_decls.inc"
SnookerSearchConfig _searchConfig;         // billiard_snooker.cpp:76
#include "_ref/snooker_rules.inc"
}  // namespace billiard::snooker

namespace {
const char* TYPE_NAMES[] = {"WHITE", "RED", "YELLOW", "GREEN", "BROWN", "BLUE", "PINK", "BLACK"};   // repo-wide type codes
const int N_TYPES = 8;
std::string nameOf(int code) { return code >= 0 && code < N_TYPES ? TYPE_NAMES[code] : "UNKNOWN"; }
int codeOf(const std::string& name) {
    for (int i = 0; i < N_TYPES; ++i) if (name == TYPE_NAMES[i]) return i;
    return -1;
}
billiard::search::State toState(int n, const float* pos, const int32_t* types, const uint8_t* present) {
    std::vector<billiard::search::Ball> balls;
    for (int i = 0; i < n; ++i)
        if (present[i] && types[i] >= 0)
            balls.push_back(billiard::search::Ball{glm::vec2{pos[2 * i], pos[2 * i + 1]}, nameOf(types[i]), std::to_string(i)});
    return billiard::search::State{balls};
}
}  // namespace

extern "C" {

// spots: YELLOW, GREEN, BROWN, BLUE, PINK, BLACK
int refrules_configure(const float* spots12, const float* headRailDir2, float diameterSquared, float radius,
                       const float* minimum2, const float* maximum2) {
    billiard::snooker::SnookerSearchConfig c;
    for (int k = 0; k < 6; ++k) c.spots[TYPE_NAMES[2 + k]] = billiard::snooker::Spot{glm::vec2{spots12[2 * k], spots12[2 * k + 1]}};
    c.headRailDirection = glm::vec2{headRailDir2[0], headRailDir2[1]};
    c.diameterSquared = diameterSquared;
    c.radius = radius;
    c._minimum = glm::vec2{minimum2[0], minimum2[1]};
    c._maximum = glm::vec2{maximum2[0], maximum2[1]};
    billiard::snooker::_searchConfig = c;
    return 0;
}

int refrules_next_search(int n, const float* pos, const int32_t* types, const uint8_t* present, int nPrevious,
                         const int32_t* previousTypes, int32_t* outTypes) {
    std::vector<std::string> prev;
    for (int i = 0; i < nPrevious; ++i) prev.push_back(nameOf(previousTypes[i]));
    auto search = billiard::snooker::nextSearch(toState(n, pos, types, present), prev);
    int k = 0;
    for (auto& t : search._types) outTypes[k++] = codeOf(t);
    return k;
}

// ids: every type that has a slot in the frame maps to its lowest slot (the reference keys ids by type, one id each)
int refrules_state_after_break(int n, const float* pos, const int32_t* types, const uint8_t* present, float* outPos,
                               uint8_t* outPresent) {
    std::unordered_map<std::string, std::string> ids;
    for (int i = 0; i < n; ++i)
        if (types[i] >= 0 && !ids.count(nameOf(types[i]))) ids.insert({nameOf(types[i]), std::to_string(i)});
    auto next = billiard::snooker::stateAfterBreak(toState(n, pos, types, present), ids);
    for (int i = 0; i < n; ++i) { outPos[2 * i] = pos[2 * i]; outPos[2 * i + 1] = pos[2 * i + 1]; outPresent[i] = 0; }
    if (next._balls.empty()) return 0;
    for (auto& b : next._balls) {
        const int i = std::stoi(b._id);
        outPos[2 * i] = b._position.x; outPos[2 * i + 1] = b._position.y;
        outPresent[i] = 1;
    }
    return 1;
}

}  // extern "C"
