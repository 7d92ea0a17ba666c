// The geometric planner of the reference -- which strikes are worth simulating on a table state -- restated as plain
// host + device code, so that the reference's OWN candidate generator runs per beam state on the device,
// next to the simulation, without a host round trip.  Reference: lib/billiard_search/src/billiard_search.cpp
//
//   expandSearchNode :299-310         expandSearchNodeByBalls :312-348     pocketCollision :446-480
//   ballCollision :497-540            enoughSpace :563-601                 expandSearchNodeByBall :623-650
//   calculateMinimalVelocity :652-758 prepareForSimulation :797-826        expandSearchNodeByBanks :828-840
//   canReflect / reflect :846-858     railCombinations :865-927            railTargets :936-1034
//   expandByRailPositions :1063-1098  expandSearchNodeByBankBall :1100-1140  searchCost :1158-1347
//   and lib/billiard_physics/src/billiard_physics.cpp: pointToLineSegmentSquaredDistance :30-52, startVelocity :72-88,
//   elasticCollisionTargetPosition :90-94, elasticCollisionReverse :96-103, railCollisionReverse :105-107.
//
// The reference searches best-first over a tree of SEARCH nodes -- pocket -> ball to pot -> ball that hits it -> ... ->
// cue ball, each link direct or over up to BANK_INDIRECTION cushions (mirror method) -- and turns every path that ends at the
// cue ball into SIMULATION nodes (prepareForSimulation: the minimal strike velocity and, optionally, two faster ones).  The
// tree is bounded (MAX_SEARCH_DEPTH = 3), so the same set of strikes comes out of a plain enumeration: that is what
// plan_root() does for one (pocket, ball to pot) pair, depth first, with the reference's feasibility tests, its cost
// function and its velocity back-propagation operation for operation (binary32 / binary64 exactly where the reference
// uses them; glm semantics as in bcu_physics.cuh).  Two things the reference leaves to its hash containers are inputs here:
// the iteration order of Configuration::_table._rails (railTargets takes the FIRST rail hit in that order) and the
// iteration order of the std::unordered_set of rail ids in railCombinations (it decides which cushion follows which in a
// multi-cushion combination); the host side reads both off the real containers (host mirror) or takes them as data.
//
// Strikes are emitted through a functor: emit(vx, vy, searchCost, cost) = the BALL_SHOT velocity of one SIMULATION node,
// its SearchNode::_searchCost and its accumulated SearchNode::_cost.  The order of emission is not the reference's
// (best-first); the SET is (tests/test_planner.py compares it with the reference's expand() run to exhaustion).
#pragma once

#include "bcu_physics.cuh"

namespace bcu {

constexpr int PLAN_MAX_SEARCH_DEPTH = 3;        // MAX_SEARCH_DEPTH, billiard_search.cpp:32
constexpr int PLAN_BANK_INDIRECTION = 3;        // BANK_INDIRECTION :16
constexpr int PLAN_FORWARD_SEARCHES = 2;        // FORWARD_SEARCHES :17
constexpr int PLAN_MAX_EVENTS = 2 + PLAN_BANK_INDIRECTION;   // target event, cushions, BALL_KICK

enum PlanEventType : int { PE_BALL_COLLISION = 0, PE_RAIL_COLLISION = 1, PE_POCKET_COLLISION = 2, PE_BALL_KICK = 3 };   // type.hpp:399-404

struct PlanRail {
    float sx, sy, ex, ey;      // _shiftedStart, _shiftedEnd
    float nx, ny;              // _normal
    float ox, oy;              // _start (unshifted: canReflect :852-858)
    float m00, m11, m20, m21;  // _reflectionMatrix: diag(s.x, s.y) and the translation column (type.cpp:73-80)
    int ignore;                // ignoreRailInSearch :860-863: a rail that is neither horizontal nor vertical
};
struct PlanPocket {
    float px, py;              // SearchPocket::_pottingPoint
    float nx, ny;              // SearchPocket::_normal
};
struct PlanTable {
    int nRails, nPockets, nSetOrder;
    PlanRail rails[MAX_RAILS];          // in the iteration order of Configuration::_table._rails
    int setOrder[MAX_RAILS];            // iteration order of railCombinations' unordered_set of non-ignored rail ids (rail indices)
    PlanPocket pockets[MAX_POCKETS * 2];// Configuration::_searchPockets
    float radius, diameter, diameterSq;
    float minimalPocketVelocity;
    float diagonalLengthSquared;
    double maxHitAngleCos;              // MAX_POSSIBLE_HIT_ANGLE = std::cos(PI / 180.0f * 87.0f) :37, evaluated on the host
    int velocityMultiply;               // Configuration::_velocityMultiplyEnabled
    int cueType;
};

struct PlanEvent {
    int type;
    float x, y;   // _targetPosition
    int id;       // rail / pocket index (ball index for BALL_COLLISION; unused)
};
struct PlanNode {
    PlanEvent ev[PLAN_MAX_EVENTS];
    int nEv;
    int ball;                  // _ballId; -1 for the pocket root
    int pocket;                // root only
    bool root;                 // _action == NONE
    uint64_t cost, searchCost;
};
struct PlanState {
    int nBalls;
    const float* pos;          // [nBalls][2]
    const int8_t* type;        // [nBalls], < 0: no ball
    uint32_t searchMask;       // Search::_types as a type mask
};

BCU_HD v2 plan_pos(const PlanState& st, int b) { return mk(st.pos[2 * b], st.pos[2 * b + 1]); }

// physics::pointToLineSegmentSquaredDistance (billiard_physics.cpp:30-52)
BCU_HD float plan_point_segment_sq(v2 p1, v2 p2, v2 point) {
    const v2 dir = p2 - p1;
    const v2 dn = glm_normalize(dir);
    const v2 toPoint = point - p1;
    const float factor = dot(dn, toPoint) / length(dir);
    if (factor < 0) { const v2 s = point - p1; return dot(s, s); }
    if (factor > 1.0) { const v2 e = point - p2; return dot(e, e); }
    const v2 dv = toPoint - (dir * factor);
    return dot(dv, dv);
}
// physics::startVelocity (:72-88)
BCU_HD v2 plan_start_velocity(v2 target, float distance, const PhysConst& pc) {
    const v2 n = glm_normalize(target);
    const float sq = dot(target, target);
    const float velocity = sqrtf((49.0f * pc.muSlide * (sq + 2 * pc.g * pc.mu * distance)) /
                                 ((49.0f * pc.muSlide) + 24.0f * (pc.mu - pc.muSlide)));
    return velocity * n;
}
// physics::elasticCollisionTargetPosition (:90-94)
BCU_HD v2 plan_collision_target(v2 targetBallPos, v2 targetDir, float radius) { return targetBallPos - 2 * radius * glm_normalize(targetDir); }
// physics::elasticCollisionReverse (:96-103); energyAdditionFactorBall = 1 / energyLossFactorBall
BCU_HD v2 plan_collision_reverse(v2 targetVel, v2 originVel, const PhysConst& pc) {
    const v2 on = glm_normalize(originVel);
    const float denom = dot(on, targetVel);
    const float addBall = 1 / pc.lossBall;
    return (dot(targetVel, targetVel) / denom) * on * addBall;
}
// physics::railCollisionReverse (:105-107); energyAdditionFactorRail = 1 / energyLossFactorRail
BCU_HD v2 plan_rail_reverse(v2 vel, v2 newDir, const PhysConst& pc) {
    const float addRail = 1 / pc.lossRail;
    return glm_normalize(newDir) * length(vel * addRail);
}
// reflect (:850-852): rail._reflectionMatrix * vec3(position, 1)
BCU_HD v2 plan_reflect(v2 p, const PlanRail& r) {
    return mk(r.m00 * p.x + 0.0f * p.y + r.m20 * 1.0f, 0.0f * p.x + r.m11 * p.y + r.m21 * 1.0f);
}
// canReflect (:854-861)
BCU_HD bool plan_can_reflect(v2 p, const PlanRail& r) {
    const v2 moved = p - mk(r.ox, r.oy);
    const v2 refl = mk(moved.x * r.nx, moved.y * r.ny);
    const float px = refl.x >= 0.0f ? 1.0f : 0.0f, py = refl.y >= 0.0f ? 1.0f : 0.0f;
    return dot(mk(px, py), mk(px, py)) == 2 && (r.nx == 0 || refl.x > 0) && (r.ny == 0 || refl.y > 0);
}
// intersection::halfLineIntersectsLineSegment (:725-753): t of the half-line, or false
BCU_HD bool plan_half_line_segment(v2 p, v2 d, v2 s, v2 e, float* tOut) {
    float t, u;
    if (!line_intersects_line(p, d, s, e - s, &t, &u)) return false;
    if (t < 0.0f) return false;
    if (u < 0.0f || u > 1.0f) return false;
    *tOut = t;
    return true;
}

// collidesOnTheWay (:388-423): any ball of the state other than `ball` and `targetBall` within (diameter + 10 mm) of the segment
BCU_HD bool plan_collides_on_the_way(const PlanTable& tb, const PlanState& st, v2 from, int ball, int targetBall, v2 target) {
    const float spacing = 10.0f;
    const float minSq = tb.diameterSq + 2 * tb.diameter * spacing + spacing * spacing;
    for (int o = 0; o < st.nBalls; ++o) {
        if (st.type[o] < 0 || o == ball || o == targetBall) continue;
        if (plan_point_segment_sq(from, target, plan_pos(st, o)) <= minSq) return true;
    }
    return false;
}

// direction in which the parent's ball has to leave (secondLastEvent - lastEvent, :482-495)
BCU_HD v2 plan_parent_vector(const PlanNode& parent) {
    const PlanEvent& last = parent.ev[parent.nEv - 1];
    const PlanEvent& second = parent.ev[parent.nEv - 2];
    return mk(second.x, second.y) - mk(last.x, last.y);
}
BCU_HD v2 plan_parent_target(const PlanTable& tb, const PlanNode& parent) {   // elasticCollisionTargetPositionFromParent
    const PlanEvent& last = parent.ev[parent.nEv - 1];
    return plan_collision_target(mk(last.x, last.y), glm_normalize(plan_parent_vector(parent)), tb.radius);
}

// enoughSpace (:563-601): room behind the cue ball for the cue
BCU_HD bool plan_enough_space(const PlanTable& tb, const PlanState& st, const PlanNode& parent, int ball) {
    const float lengthS = tb.radius + tb.radius;
    const float minDistance = 10 + tb.radius;
    const float minSq = minDistance * minDistance;
    const v2 s = -lengthS * glm_normalize(plan_parent_vector(parent));
    const v2 bp = plan_pos(st, ball);
    const v2 targetPoint = bp + s;
    for (int o = 0; o < st.nBalls; ++o) {
        if (st.type[o] < 0 || o == ball) continue;
        if (plan_point_segment_sq(bp, targetPoint, plan_pos(st, o)) < minSq) return false;
    }
    return true;
}

// searchCost (:1158-1347) -> the integer step cost
BCU_HD uint64_t plan_search_cost(const PlanTable& tb, const PlanNode& node, bool ballIsCue, const PlanNode* parent) {
    const double maxDistanceSquared = tb.diagonalLengthSquared;
    double totalDistance = 0.0, totalAngle = 0.0, totalIndirection = 0.0;
    const bool haveParent = parent != nullptr && parent->nEv > 0;
    for (int i = 0; i < node.nEv; ++i) {
        const PlanEvent& ev = node.ev[i];
        if (i > 0) {
            const v2 way = mk(node.ev[i - 1].x, node.ev[i - 1].y) - mk(ev.x, ev.y);
            const double distanceCost = dot(way, way) / maxDistanceSquared;
            double parentWeight = 1.0;
            if (haveParent) {
                const v2 tv = plan_parent_vector(*parent);
                const float parentDistanceCost = (float)(dot(tv, tv) / maxDistanceSquared);
                parentWeight = parentDistanceCost;
            }
            totalDistance += ballIsCue ? distanceCost * parentWeight : distanceCost;
        }
        if (ev.type == PE_BALL_KICK) continue;
        const v2 previousPosition = mk(node.ev[i + 1].x, node.ev[i + 1].y);
        const v2 targetDirection = glm_normalize(mk(ev.x, ev.y) - previousPosition);
        double cosTheta = 0.0;
        if (ev.type == PE_POCKET_COLLISION) {
            const PlanPocket& pk = tb.pockets[ev.id];
            cosTheta = dot(mk(pk.nx, pk.ny), -targetDirection);
        } else if (ev.type == PE_RAIL_COLLISION) {
            const PlanRail& r = tb.rails[ev.id];
            cosTheta = dot(mk(r.nx, r.ny), -targetDirection);
        } else {
            cosTheta = dot(glm_normalize(plan_parent_vector(*parent)), targetDirection);
        }
        const double angleCost = -1 * cosTheta + 1;
        // the bezier weighting: t * coeff * points with glm::vec4 / mat4x4 / mat2x4 in binary32 (:1283-1296)
        const float t0 = (float)(angleCost * angleCost * angleCost), t1 = (float)(angleCost * angleCost), t2 = (float)angleCost, t3 = 1.0f;
        const float c0 = -1.0f * t0 + 3.0f * t1 + -3.0f * t2 + 1.0f * t3;   // row vector * coeff: dot(coeff[col], t)
        const float c1 = 3.0f * t0 + -6.0f * t1 + 3.0f * t2 + 0.0f * t3;
        const float c2 = -3.0f * t0 + 3.0f * t1 + 0.0f * t2 + 0.0f * t3;
        const float c3 = 1.0f * t0 + 0.0f * t1 + 0.0f * t2 + 0.0f * t3;
        const float ry = c0 * 0.0f + c1 * 0.0f + c2 * 1.0f + c3 * 1.0f;     // * points, y column {0, 0, 1, 1}
        totalAngle += (double)ry;
        if (ev.type == PE_BALL_COLLISION) totalIndirection += 1.0 / ((double)2);                         // SEARCH_MAX_BALL_COLLISIONS
        else if (ev.type == PE_RAIL_COLLISION) totalIndirection += 1.0 / ((double)PLAN_BANK_INDIRECTION) * 2;   // SEARCH_RAIL_COLLISION_WEIGHT
    }
    const double total = totalDistance + totalAngle + totalIndirection;
    return (uint64_t)(total * 100000);   // COST_FLOAT_TO_INT_FACTOR
}

// calculateMinimalVelocity (:652-758) over the path pocket -> ... -> cue ball (`path[0]` is the first node below the pocket)
BCU_HD v2 plan_minimal_velocity(const PlanTable& tb, const PlanNode* const* path, int nPath, const PhysConst& pc) {
    v2 minimal = mk(0.0f, 0.0f);
    const PlanEvent* previous = nullptr;
    const float maxSq = 5000.0f * 5000.0f;
    for (int n = 0; n < nPath; ++n) {
        const PlanNode& node = *path[n];
        for (int i = 0; i < node.nEv; ++i) {
            const PlanEvent& ev = node.ev[i];
            if (previous) {
                const v2 cur = mk(ev.x, ev.y), tgt = mk(previous->x, previous->y);
                if (ev.type == PE_BALL_COLLISION) {
                    // handled with the next event
                } else if (previous->type == PE_BALL_COLLISION) {
                    const v2 toTarget = tgt - cur;
                    const v2 origin = glm_normalize(toTarget);
                    minimal = plan_collision_reverse(minimal, origin, pc);
                    const float distance = length(toTarget);
                    minimal = plan_start_velocity(minimal, distance, pc);
                } else if (previous->type == PE_POCKET_COLLISION) {
                    const v2 toTarget = tgt - cur;
                    const float distance = length(toTarget);
                    const v2 finalVelocity = tb.minimalPocketVelocity * glm_normalize(toTarget);
                    minimal = plan_start_velocity(finalVelocity, distance, pc);
                } else if (previous->type == PE_RAIL_COLLISION) {
                    const v2 toTarget = tgt - cur;
                    const float distance = length(toTarget);
                    minimal = plan_rail_reverse(minimal, toTarget, pc);
                    minimal = plan_start_velocity(minimal, distance, pc);
                } else {
                    const v2 toTarget = tgt - cur;
                    const float distance = length(toTarget);
                    minimal = plan_start_velocity(minimal, distance, pc);
                }
                if (dot(minimal, minimal) > maxSq) break;   // (leaves `previous` where it was, as the reference does)
            }
            previous = &ev;
        }
    }
    return minimal;
}

// prepareForSimulation (:797-826)
template <class Emit>
BCU_HD void plan_emit_strikes(const PlanTable& tb, const PlanNode* const* path, int nPath, const PlanNode& leaf, const PhysConst& pc, Emit& emit) {
    const v2 minimal = plan_minimal_velocity(tb, path, nPath, pc);
    const v2 mn = glm_normalize(minimal);
    const float step = 2000.0f / PLAN_FORWARD_SEARCHES;   // VELOCITY_STEP
    for (int i = 0; i <= (tb.velocityMultiply ? PLAN_FORWARD_SEARCHES : 0); ++i) {
        const v2 v = minimal + (i * step * mn);
        if (dot(v, v) > 5000.0f * 5000.0f) break;
        emit(v.x, v.y, leaf.searchCost, leaf.cost);
    }
}

// railCombinations (:865-927): the mirrored targets of `target` over `depth` cushions.  out[i] = mirrored target, or flagged invalid.
struct PlanCombos {
    int n;
    v2 target[MAX_RAILS];
};
BCU_HD void plan_rail_combinations(const PlanTable& tb, int depth, v2 target, PlanCombos* out) {
    v2 refl[MAX_RAILS];
    bool ok[MAX_RAILS];
    int n = 0;
    for (int r = 0; r < tb.nRails; ++r) {
        if (tb.rails[r].ignore) continue;
        refl[n] = plan_reflect(target, tb.rails[r]);
        ok[n] = plan_can_reflect(target, tb.rails[r]);
        ++n;
    }
    for (int reflection = 1; reflection < depth; ++reflection) {
        for (int i = 0; i < tb.nSetOrder && i < n; ++i) {
            const PlanRail& rail = tb.rails[tb.setOrder[i]];
            if (ok[i] && plan_can_reflect(refl[i], rail)) refl[i] = plan_reflect(refl[i], rail);
            else ok[i] = false;
        }
    }
    out->n = 0;
    for (int i = 0; i < n; ++i) {
        if (!ok[i]) continue;
        bool already = false;
        for (int j = 0; j < out->n; ++j) {
            const float epsilon = (float)0.0001;
            if (fabsf(out->target[j].x - refl[i].x) <= epsilon && fabsf(out->target[j].y - refl[i].y) <= epsilon) { already = true; break; }
        }
        if (!already) out->target[out->n++] = refl[i];
    }
}

// railTargets (:936-1034) + expandByRailPositions (:1063-1098): the cushion points of one combination, walked back from the
// ball; fills node.ev and returns false when the path is not possible
BCU_HD bool plan_bank_node(const PlanTable& tb, const PlanState& st, const PlanNode& parent, int ball, v2 reflected, int depth, PlanNode* node) {
    v2 start = plan_pos(st, ball);
    int excluded = ball;
    int previousRail = -1;
    v2 tPos[PLAN_BANK_INDIRECTION];
    int tRail[PLAN_BANK_INDIRECTION];
    int nT = 0;
    for (int k = 0; k < depth; ++k) {
        const v2 direction = reflected - start;
        int collided = -1;
        v2 target = mk(0.0f, 0.0f);
        for (int r = 0; r < tb.nRails; ++r) {
            const PlanRail& rail = tb.rails[r];
            if (rail.ignore) continue;
            float t;
            if (plan_half_line_segment(start, direction, mk(rail.sx, rail.sy), mk(rail.ex, rail.ey), &t) && previousRail != r) {
                target = start + t * direction;
                reflected = plan_reflect(reflected, rail);
                collided = r;
                previousRail = r;
                break;
            }
        }
        if (collided < 0) return false;
        if (plan_collides_on_the_way(tb, st, start, excluded, -1, target)) return false;
        tPos[nT] = target; tRail[nT] = collided; ++nT;
        start = target;
        excluded = -1;
    }
    v2 targetPosition;
    if (parent.root) {
        targetPosition = mk(tb.pockets[parent.pocket].px, tb.pockets[parent.pocket].py);
    } else {
        const v2 dir = glm_normalize(plan_parent_vector(parent));
        const PlanEvent& last = parent.ev[parent.nEv - 1];
        targetPosition = plan_collision_target(mk(last.x, last.y), dir, tb.radius);
        const v2 ballToTarget = targetPosition - start;
        const float cosTheta = dot(dir, glm_normalize(ballToTarget));
        if (cosTheta <= tb.maxHitAngleCos) return false;
    }
    if (plan_collides_on_the_way(tb, st, start, parent.root ? -1 : parent.ball, -1, targetPosition)) return false;
    node->nEv = 0;
    if (parent.root) node->ev[node->nEv++] = PlanEvent{PE_POCKET_COLLISION, targetPosition.x, targetPosition.y, parent.pocket};
    else {
        const v2 tp = plan_parent_target(tb, parent);
        node->ev[node->nEv++] = PlanEvent{PE_BALL_COLLISION, tp.x, tp.y, parent.ball};
    }
    for (int i = nT - 1; i >= 0; --i) node->ev[node->nEv++] = PlanEvent{PE_RAIL_COLLISION, tPos[i].x, tPos[i].y, tRail[i]};
    const v2 bp = plan_pos(st, ball);
    node->ev[node->nEv++] = PlanEvent{PE_BALL_KICK, bp.x, bp.y, -1};
    return true;
}

// expandBallIfPossible (:603-621) with pocketCollision (:446-480) / ballCollision (:497-540): the direct link
BCU_HD bool plan_direct_node(const PlanTable& tb, const PlanState& st, const PlanNode& parent, int ball, bool ballIsCue, PlanNode* node) {
    if (ballIsCue && !plan_enough_space(tb, st, parent, ball)) return false;
    const v2 bp = plan_pos(st, ball);
    node->nEv = 0;
    if (parent.root) {
        const v2 tp = mk(tb.pockets[parent.pocket].px, tb.pockets[parent.pocket].py);
        if (plan_collides_on_the_way(tb, st, bp, ball, -1, tp)) return false;
        node->ev[node->nEv++] = PlanEvent{PE_POCKET_COLLISION, tp.x, tp.y, parent.pocket};
    } else {
        const v2 dir = glm_normalize(plan_parent_vector(parent));
        const PlanEvent& last = parent.ev[parent.nEv - 1];
        const v2 tp = plan_collision_target(mk(last.x, last.y), dir, tb.radius);
        const v2 ballToTarget = tp - bp;
        const float cosTheta = dot(dir, glm_normalize(ballToTarget));
        if (cosTheta <= tb.maxHitAngleCos) return false;   // only possible to hit the ball from "behind"
        if (plan_collides_on_the_way(tb, st, bp, ball, parent.ball, tp)) return false;
        node->ev[node->nEv++] = PlanEvent{PE_BALL_COLLISION, tp.x, tp.y, parent.ball};
    }
    node->ev[node->nEv++] = PlanEvent{PE_BALL_KICK, bp.x, bp.y, -1};
    return true;
}

// One search node below `parent` for `ball`: the direct link, or the link over `depth` cushions whose mirrored target is
// `reflected` (one rail combination).  Fills the cost fields (expandSearchNodeByBall :623-650, ...ByBankBall :1100-1140).
BCU_HD bool plan_make_node(const PlanTable& tb, const PlanState& st, const PlanNode& parent, int ball, bool direct, v2 reflected, int depth,
                           PlanNode* node) {
    const bool ballIsCue = st.type[ball] == tb.cueType;
    const bool made = direct ? plan_direct_node(tb, st, parent, ball, ballIsCue, node)
                             : plan_bank_node(tb, st, parent, ball, reflected, depth, node);
    if (!made) return false;
    node->ball = ball; node->pocket = -1; node->root = false;
    const uint64_t step = plan_search_cost(tb, *node, ballIsCue, &parent);
    node->cost = parent.cost + step;
    node->searchCost = parent.searchCost + step;
    return true;
}

// Everything below one search node (expandSearchNode :299-310): every unused ball (at the pocket root: every ball of a searched
// type), direct and over the cushions, depth first until the cue ball closes a path.  LEVEL = _searchDepth of the children;
// a child that is not the cue ball is expanded only while LEVEL < MAX_SEARCH_DEPTH (:224-244), so the recursion is three
// nested instantiations.  A PlanFilter restricts the enumeration to one slice of the tree -- the unit of parallel work on the
// device: the first link (ball, number of cushions, combination) and the ball of the second link.
struct PlanFilter {
    int ball = -1, depth = -1, combo = -1;   // children of this node: only this (ball, cushions, combination); -1 = all
    int nextBall = -1, nextDepth = -1;       // grandchildren: only this ball / this number of cushions; -1 = all
};
template <int LEVEL, class Emit>
BCU_HD void plan_children(const PlanTable& tb, const PlanState& st, const PlanNode& parent, const PlanNode* const* path, int nPath,
                          uint32_t usedMask, const PhysConst& pc, Emit& emit, const PlanFilter& filter) {
    PlanCombos combos;
    v2 bankTarget;
    if (parent.root) bankTarget = mk(tb.pockets[parent.pocket].px, tb.pockets[parent.pocket].py);
    else bankTarget = plan_parent_target(tb, parent);
    PlanFilter below;
    below.ball = filter.nextBall;
    below.depth = filter.nextDepth;
    for (int depth = 0; depth <= PLAN_BANK_INDIRECTION; ++depth) {   // 0 = direct (expandSearchNodeByBalls), 1.. = banks
        if (filter.depth >= 0 && depth != filter.depth) continue;
        if (depth > 0) plan_rail_combinations(tb, depth, bankTarget, &combos);
        const int nVariants = depth == 0 ? 1 : combos.n;
        for (int c = 0; c < nVariants; ++c) {
            if (filter.combo >= 0 && c != filter.combo) continue;
            for (int b = 0; b < st.nBalls; ++b) {
                if (filter.ball >= 0 && b != filter.ball) continue;
                if (st.type[b] < 0 || ((usedMask >> b) & 1u)) continue;
                if (parent.root && !((st.searchMask >> st.type[b]) & 1u)) continue;   // :319-322, :1052-1053
                // a node of the last level is never expanded (:224-244), so only the cue ball can contribute there: the reference
                // builds the other nodes and drops them, which is most of its work
                if (LEVEL == PLAN_MAX_SEARCH_DEPTH && st.type[b] != tb.cueType) continue;
                PlanNode node;
                if (!plan_make_node(tb, st, parent, b, depth == 0, depth == 0 ? mk(0.0f, 0.0f) : combos.target[c], depth, &node)) continue;
                const PlanNode* sub[PLAN_MAX_SEARCH_DEPTH];
                for (int i = 0; i < nPath; ++i) sub[i] = path[i];
                sub[nPath] = &node;
                if (st.type[b] == tb.cueType) {
                    plan_emit_strikes(tb, sub, nPath + 1, node, pc, emit);
                } else if constexpr (LEVEL < PLAN_MAX_SEARCH_DEPTH) {
                    plan_children<LEVEL + 1>(tb, st, node, sub, nPath + 1, usedMask | (1u << b), pc, emit, below);
                }
            }
        }
    }
}

// One pocket root of one state (addPocketsAsInitialTargets :246-271): cost = the state's accumulated cost, searchCost = 0.
template <class Emit>
BCU_HD void plan_root(const PlanTable& tb, const PlanState& st, int pocket, uint64_t baseCost, const PhysConst& pc, Emit& emit,
                      const PlanFilter& filter = PlanFilter()) {
    PlanNode root;
    root.nEv = 0; root.ball = -1; root.pocket = pocket; root.root = true;
    root.cost = baseCost; root.searchCost = 0;
    const PlanNode* path[PLAN_MAX_SEARCH_DEPTH];
    path[0] = nullptr;
    plan_children<1>(tb, st, root, path, 0, 0u, pc, emit, filter);
}

// ---- host side: the planner's tables from the caller's description (bcu_set_planner; tests build the same tables) -------
// derived6 [nRails][6] = Rail::_normal, _shiftedStart, _shiftedEnd; railStart [nRails][2] = Rail::_start; railOrder [nRails]
// (may be null: as given) = the iteration order of Configuration::_table._rails as indices into these arrays; setOrder
// [nSet] = the iteration order of railCombinations' unordered_set of rail ids, as indices into these arrays; searchPockets
// [nPockets][4] = SearchPocket::_pottingPoint, _normal.  Returns false on a bad description.
inline bool plan_table_build(PlanTable* tb, int nRails, const float* derived6, const float* railStart, const int* railOrder, int nSet,
                             const int* setOrder, int nPockets, const float* searchPockets, float radius, float minimalPocketVelocity,
                             float diagonalLengthSquared, int velocityMultiply, int cueType) {
    if (nRails < 0 || nRails > MAX_RAILS || nPockets < 0 || nPockets > MAX_POCKETS * 2 || nSet < 0 || nSet > nRails) return false;
    *tb = PlanTable{};
    int slotOf[MAX_RAILS];
    for (int i = 0; i < nRails; ++i) {
        const int r = railOrder ? railOrder[i] : i;
        if (r < 0 || r >= nRails) return false;
        slotOf[r] = i;
        const float* d = derived6 + 6 * r;
        PlanRail& pr = tb->rails[i];
        pr.nx = d[0]; pr.ny = d[1]; pr.sx = d[2]; pr.sy = d[3]; pr.ex = d[4]; pr.ey = d[5];
        pr.ox = railStart[2 * r]; pr.oy = railStart[2 * r + 1];
        // Rail::calculateReflectionMatrix(_shiftedStart, _normal), type.cpp:73-80: s = abs(normal) * -2 + 1
        const float sxm = fabsf(pr.nx) * -2.0f + 1.0f, sym = fabsf(pr.ny) * -2.0f + 1.0f;
        pr.m00 = sxm; pr.m11 = sym;
        pr.m20 = pr.sx - sxm * pr.sx; pr.m21 = pr.sy - sym * pr.sy;
        pr.ignore = pr.nx != 0.0f && pr.ny != 0.0f;
    }
    tb->nRails = nRails;
    int nonIgnored = 0;
    for (int i = 0; i < nRails; ++i) nonIgnored += !tb->rails[i].ignore;
    if (nSet != nonIgnored) return false;   // the set holds exactly the rails that take part in the search
    for (int i = 0; i < nSet; ++i) {
        if (setOrder[i] < 0 || setOrder[i] >= nRails) return false;
        tb->setOrder[i] = slotOf[setOrder[i]];
        if (tb->rails[tb->setOrder[i]].ignore) return false;
    }
    tb->nSetOrder = nSet;
    for (int i = 0; i < nPockets; ++i) {
        tb->pockets[i].px = searchPockets[4 * i]; tb->pockets[i].py = searchPockets[4 * i + 1];
        tb->pockets[i].nx = searchPockets[4 * i + 2]; tb->pockets[i].ny = searchPockets[4 * i + 3];
    }
    tb->nPockets = nPockets;
    tb->radius = radius;
    tb->diameter = radius * 2;
    tb->diameterSq = (radius * 2) * (radius * 2);
    tb->minimalPocketVelocity = minimalPocketVelocity;
    tb->diagonalLengthSquared = diagonalLengthSquared;
    tb->maxHitAngleCos = cos(3.14159265358979323846 / 180.0f * 87.0f);   // MAX_POSSIBLE_HIT_ANGLE, billiard_search.cpp:36-37
    tb->velocityMultiply = velocityMultiply;
    tb->cueType = cueType;
    return true;
}

}  // namespace bcu
