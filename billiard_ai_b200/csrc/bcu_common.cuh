// Shared definitions for the billiard strike-evaluation kernels (sm_100a).
//
// Everything on this path is IEEE binary32 / binary64 with NO implicit contraction: the
// translation units are compiled with -fmad=false (device) and -ffp-contract=off (host test
// builds).  Explicit fused operations are written as bcu_fma() where the reference's libm
// binary uses one.
#pragma once

#include <stdint.h>

#if defined(__CUDACC__)
#define BCU_HD __host__ __device__ __forceinline__
#define BCU_D __device__ __forceinline__
// large routines called from several places of the quartic: one out-of-line copy each keeps the kernel's instruction
// footprint inside the instruction cache (ncu on the fully inlined build: "no_instruction" was the top stall)
#define BCU_HD_NOINLINE inline __host__ __device__ __noinline__
#else
#define BCU_HD inline
#define BCU_D inline
#define BCU_HD_NOINLINE inline
#endif

namespace bcu {

// ---- limits ---------------------------------------------------------------------------------
constexpr int MAX_BALLS = 32;    // one 32-bit mask per property
constexpr int MAX_RAILS = 32;
constexpr int MAX_POCKETS = 8;
constexpr int MAX_EVENTS = 30;   // reference billiard_search.cpp:20  #define MAX_EVENTS 30
constexpr int MAX_TYPES = 16;

// ---- node kinds: numerically equal to reference node::NodeType (type.hpp:167-180) -----------
enum NodeKind : int {
    K_MOVING = 0,
    K_COLLISION = 1,
    K_RAIL = 2,
    K_POTTING = 3,
    K_SHOT = 4,
    K_REST = 5,
    K_ABSENT = 7,
};

// ---- event types: numerically equal to reference event::EventType (type.hpp:87-98) ----------
enum EventKind : int {
    E_BALL_COLLISION = 0,
    E_RAIL_COLLISION = 1,
    E_POTTING = 2,
    E_REST = 3,
    E_ROLLING = 4,
    E_NONE = 7,
};

// ---- result flags (same values as oracle/ref_driver.cpp) ------------------------------------
enum ResultFlags : uint32_t {
    F_COMPLETED = 1u,  // final layer reached within MAX_EVENTS
    F_VALID = 2u,      // rules: valid end state (billiard_snooker.cpp:1128-1153)
    F_OVERFLOW = 4u,   // more than MAX_EVENTS events  (billiard_search.cpp:1376-1379)
    F_NOEVENT = 8u,    // no next event although not final (billiard_search.cpp:1464-1467)
};

// ---- physics constants, reference billiard_physics.hpp:14-25 --------------------------------
// Evaluated exactly as the reference's `const float` initialisers are (double literal -> float,
// float products in float).
struct PhysConst {
    float g;          // gravitationalAcceleration = 9.8 * 1000
    float mu;         // frictionCoefficient = 0.0142435
    float muSlide;    // 12.0f * mu
    float lossBall;   // energyLossFactorBall = 1 - 0.05
    float lossRail;   // energyLossFactorRail = 1 - 0.38f
    float topspinRest;
    float accLen;       // accelerationLength()      = -g * mu
    float slideAccLen;  // slideAccelerationLength() = -g * muSlide
    float rollDenom;    // (3.5f) * g * muSlide  (timeToRolling)
};

BCU_HD PhysConst phys_const() {
    PhysConst c;
    c.g = (float)(9.8 * 1000);
    c.mu = (float)0.0142435;
    c.muSlide = 12.0f * c.mu;
    c.lossBall = 1 - (float)0.05;  // `1 - energyLossByBall`: int - float -> float
    c.lossRail = 1 - 0.38f;
    c.topspinRest = 0.25f;
    c.accLen = -c.g * c.mu;
    c.slideAccLen = -c.g * c.muSlide;
    c.rollDenom = (3.5f) * c.g * c.muSlide;
    return c;
}

// ---- table geometry as the kernels consume it ----------------------------------------------
struct RailGeom {
    float sx, sy;   // shifted start (type.cpp:67-72)
    float dx, dy;   // shifted end - shifted start (direction passed to lineIntersectsLine)
    float nx, ny;   // normal = normalize(perp(end - start))  (type.cpp:63)
};
struct PocketGeom {
    float x, y, r, r2;  // r2 = r*r
};

struct TableGeom {
    int nRails, nPockets;
    float diameter;      // Configuration::_ball._diameter (= radius*2)
    float diameterSq;    // diameter*diameter (computed as the reference does at each use)
    RailGeom rails[MAX_RAILS];
    PocketGeom pockets[MAX_POCKETS];
};

// ---- rules as data (std::function rules cannot run on device) -------------------
struct RulesData {
    int nTypes;
    int cueType;                 // type code of "WHITE"
    double score[MAX_TYPES];     // scoreForPottedBall per type code
};

}  // namespace bcu
