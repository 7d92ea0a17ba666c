// Rules between breaks (host side of the multi-break search): which ball types the next break may pot, and the table
// state the next break starts from (potted colours go back on their spots).  Plain host C++ behind the C ABI -- the
// reference runs these once per finished simulation on the CPU as well:
//
//   bcu_snooker_next_search        <- billiard::snooker::nextSearch       lib/billiard_snooker/src/billiard_snooker.cpp:792-822
//   bcu_snooker_state_after_break  <- billiard::snooker::stateAfterBreak  :1097-1126, replace :1053-1095,
//                                     replaceInArea :969-1046, expandBySpace :867-907, tryToSetBallAtPositions :934-967
//
// Ball types are small integer codes, a ball is a slot (index) of the state table; the reference's strings and ids map to
// them one to one (a ball's id is its slot).  All arithmetic is binary32 in the reference's operation order (glm scalar
// semantics, see bcu_physics.cuh), so re-spotted positions come out bit-identical.
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../include/billiard_cuda.h"

namespace {

struct V2 {
    float x, y;
};
inline V2 operator+(V2 a, V2 b) { return V2{a.x + b.x, a.y + b.y}; }
inline V2 operator-(V2 a, V2 b) { return V2{a.x - b.x, a.y - b.y}; }
inline V2 operator-(V2 a) { return V2{-a.x, -a.y}; }
inline V2 operator*(float s, V2 a) { return V2{s * a.x, s * a.y}; }
inline bool operator==(V2 a, V2 b) { return a.x == b.x && a.y == b.y; }
inline float dot(V2 a, V2 b) { float tx = a.x * b.x, ty = a.y * b.y; return tx + ty; }   // glm::dot(vec2)

// order of billiard_snooker.cpp:773-780 (typesToReplace) and of the :761-770 successor chain
enum { YELLOW = 0, GREEN, BROWN, BLUE, PINK, BLACK, N_COLORS };

int color_index(const bcu_snooker_config* c, int type) {
    for (int i = 0; i < N_COLORS; ++i) if (c->color_types[i] == type) return i;
    return -1;
}

// ---- replaceInArea (:969-1046) ---------------------------------------------------------------------------------------
enum Action { EXPAND_1, EXPAND_2, EXPAND_3, EXPAND_4, NONE };
inline Action revert(Action a) {   // :842-847
    return a == EXPAND_1 ? EXPAND_4 : a == EXPAND_4 ? EXPAND_1 : a == EXPAND_2 ? EXPAND_3 : EXPAND_2;
}
struct Place {
    V2 position;
    Action action;
};
using Places = std::vector<Place>;

// expandBySpace (:867-907)
Places expand_by_space(V2 spot, float searchRadiusSquared, const Places& fringe, V2 space, V2 minimum, V2 maximum, float radius,
                       Action action) {
    Places out;
    const float minX = minimum.x + radius, minY = minimum.y + radius, maxX = maximum.x - radius, maxY = maximum.y - radius;
    for (const Place& p : fringe) {
        if (p.action == NONE || action != revert(p.action)) {
            const V2 next = p.position + space;
            if (next.x < minX || next.y < minY || next.x > maxX || next.y > maxY) continue;
            const V2 distance = next - spot;
            if (dot(distance, distance) < searchRadiusSquared) out.push_back(Place{next, action});
        }
    }
    return out;
}
// vector operator- (:909-932): elements of `from` whose position is not in `subtract` (ReplacePosition::operator== compares
// positions only, :853-855)
Places minus(const Places& from, const Places& subtract) {
    Places out;
    for (const Place& p : from) {
        bool found = false;
        for (const Place& q : subtract) if (q.position == p.position) { found = true; break; }
        if (!found) out.push_back(p);
    }
    return out;
}

struct StateView {
    int n;
    const float* pos;
    const uint8_t* present;
};
bool too_close(const StateView& st, V2 where, float diameterSquared) {
    for (int i = 0; i < st.n; ++i) {
        if (!st.present[i]) continue;
        const V2 s = V2{st.pos[2 * i], st.pos[2 * i + 1]} - where;
        if (dot(s, s) < diameterSquared) return true;
    }
    return false;
}
// tryToSetBallAtPositions (:934-967)
bool try_places(const Places& places, float diameterSquared, const StateView& st, V2* out) {
    for (const Place& p : places)
        if (!too_close(st, p.position, diameterSquared)) { *out = p.position; return true; }
    return false;
}

bool replace_in_area(const bcu_snooker_config* c, V2 spot, const StateView& st, V2* out) {
    const float searchRadius = c->radius * 5;   // SPOT_REPLACE_RADIUS_MULTIPLIER, billiard_snooker.hpp:15
    const V2 head = V2{c->head_rail_dir[0], c->head_rail_dir[1]};
    const V2 perpDir = V2{-head.y, head.x};     // physics::perp, billiard_physics.cpp:177-179
    const float step = c->radius / 2.0f;
    const V2 space1 = step * head, space2 = step * perpDir, space3 = step * -perpDir, space4 = step * -head;
    const V2 minimum = V2{c->minimum[0], c->minimum[1]}, maximum = V2{c->maximum[0], c->maximum[1]};
    Places fringe{Place{spot, NONE}};
    const float searchRadiusSquared = searchRadius * searchRadius;
    while (!fringe.empty()) {
        const Places n1 = minus(expand_by_space(spot, searchRadiusSquared, fringe, space1, minimum, maximum, c->radius, EXPAND_1), fringe);
        if (try_places(n1, c->diameter_squared, st, out)) return true;
        const Places n2 = minus(minus(expand_by_space(spot, searchRadiusSquared, fringe, space2, minimum, maximum, c->radius, EXPAND_2), fringe), n1);
        if (try_places(n2, c->diameter_squared, st, out)) return true;
        const Places n3 = minus(minus(minus(expand_by_space(spot, searchRadiusSquared, fringe, space3, minimum, maximum, c->radius, EXPAND_3), fringe), n1), n2);
        if (try_places(n3, c->diameter_squared, st, out)) return true;
        const Places n4 = minus(minus(minus(minus(expand_by_space(spot, searchRadiusSquared, fringe, space4, minimum, maximum, c->radius, EXPAND_4), fringe), n1), n2), n3);
        if (try_places(n4, c->diameter_squared, st, out)) return true;
        fringe.clear();
        fringe.insert(fringe.end(), n1.begin(), n1.end());
        fringe.insert(fringe.end(), n2.begin(), n2.end());
        fringe.insert(fringe.end(), n3.begin(), n3.end());
        fringe.insert(fringe.end(), n4.begin(), n4.end());
    }
    return false;
}

}  // namespace

extern "C" {

int bcu_snooker_next_search(const bcu_snooker_config* c, int32_t nBalls, const int32_t* types, const uint8_t* present,
                            int32_t nPrevious, const int32_t* previousTypes, int32_t* outTypes, int32_t* outCount) {
    if (!c || nBalls < 0 || (nBalls > 0 && (!types || !present)) || nPrevious < 0 || (nPrevious > 0 && !previousTypes) ||
        !outTypes || !outCount)
        return BCU_ERR_INVALID;
    *outCount = 0;
    if (nPrevious == 0) return BCU_OK;                                      // :796-798
    int reds = 0;                                                            // (uint8_t in the reference: <= 32 balls here)
    for (int i = 0; i < nBalls; ++i) reds += (present[i] && types[i] == c->red_type) ? 1 : 0;
    bool previousHadRed = false;
    for (int i = 0; i < nPrevious; ++i) previousHadRed |= previousTypes[i] == c->red_type;
    // nextSearchType[previousTypes.at(0)] (:761-770); anything that is not a colour -- RED, UNKNOWN, a missing key -- maps to
    // the empty type name, which no ball carries: code -1
    auto successor = [&](int type) {
        if (type == c->red_type) return (int)c->color_types[YELLOW];
        const int k = color_index(c, type);
        return (k >= 0 && k + 1 < N_COLORS) ? (int)c->color_types[k + 1] : -1;
    };
    if (previousHadRed) {
        if (reds > 0) {                                                      // :806-808: all colours, BLACK first
            for (int k = N_COLORS - 1; k >= 0; --k) outTypes[(*outCount)++] = c->color_types[k];
        } else {                                                             // :809-811
            outTypes[(*outCount)++] = successor(previousTypes[0]);
        }
    } else {
        if (reds > 0) outTypes[(*outCount)++] = c->red_type;                  // :814-816
        else outTypes[(*outCount)++] = successor(previousTypes[0]);          // :817-819
    }
    return BCU_OK;
}

int bcu_snooker_state_after_break(const bcu_snooker_config* c, int32_t nBalls, const float* pos, const int32_t* types,
                                  const uint8_t* present, float* outPos, uint8_t* outPresent, int32_t* outOk) {
    if (!c || nBalls < 0 || (nBalls > 0 && (!pos || !types || !present || !outPos || !outPresent)) || !outOk)
        return BCU_ERR_INVALID;
    std::memcpy(outPos, pos, (size_t)nBalls * 2 * sizeof(float));
    std::memcpy(outPresent, present, (size_t)nBalls);
    *outOk = 1;
    bool hasRed = false;                                                      // :1099-1110
    for (int i = 0; i < nBalls; ++i) hasRed |= present[i] && types[i] == c->red_type;
    if (!hasRed) return BCU_OK;
    const StateView st{nBalls, pos, present};   // every check is made against the state BEFORE any ball of this call is re-spotted
    for (int k = 0; k < N_COLORS; ++k) {                                     // typesToReplace order (:773-780)
        const int type = c->color_types[k];
        int slot = -1;                                                        // ids.count(type): the ball exists in this frame
        bool onTable = false;                                                 // replace :1057-1062
        for (int i = 0; i < nBalls; ++i) {
            if (types[i] != type) continue;
            if (slot < 0) slot = i;
            onTable |= present[i] != 0;
        }
        if (slot < 0 || onTable) continue;
        bool placed = false;
        V2 where{0, 0};
        for (int sk = k; sk < N_COLORS; ++sk) {                               // own spot, then the next higher ones (:1064-1087)
            const V2 spot = V2{c->spots[sk][0], c->spots[sk][1]};
            if (!too_close(st, spot, c->diameter_squared)) { where = spot; placed = true; break; }
        }
        if (!placed) placed = replace_in_area(c, V2{c->spots[k][0], c->spots[k][1]}, st, &where);   // :1089-1094
        if (!placed) {                                                        // :1118-1120: empty state
            *outOk = 0;
            std::memset(outPresent, 0, (size_t)nBalls);
            return BCU_OK;
        }
        outPos[2 * slot] = where.x; outPos[2 * slot + 1] = where.y;
        outPresent[slot] = 1;
    }
    return BCU_OK;
}

}  // extern "C"
