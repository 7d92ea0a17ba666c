// Kernel-side parameter blocks shared between bcu_kernels.cu and the C-ABI layer (bcu_api.cu).
#pragma once

#include "bcu_common.cuh"

namespace bcu {

// Where a launch takes its candidates from.
struct CandidateSource {
    // explicit candidates (SoA, device pointers); used when gridDirs == 0
    const int32_t* stateIdx;
    const float* vx;
    const float* vy;
    const uint64_t* searchCost;
    const uint64_t* baseCost;
    // direction x speed grid per beam state; used when gridDirs > 0:
    //   candidate c -> state s = c / (gridDirs*gridSpeeds), dir k = (c / gridSpeeds) % gridDirs, speed j = c % gridSpeeds
    //   velocity = speedTable[j] * (dirTable[2k], dirTable[2k+1])  (one float multiply per component)
    int32_t gridDirs, gridSpeeds;
    const float* dirTable;     // [gridDirs][2] unit vectors, computed on the host
    const float* speedTable;   // [gridSpeeds]
    uint64_t gridSearchCost;   // constant searchCost of synthetic candidates
    const uint64_t* stateBaseCost;  // [nStates] accumulated cost of the beam state (may be null = 0)
    int64_t firstCandidate;    // global id of this launch's first candidate (multi-GPU shards)
};

struct StateTable {
    int32_t nStates, nBalls;
    const float* pos;             // [nStates][nBalls][2]
    const int8_t* type;           // [nStates][nBalls], -1 = absent
    const uint32_t* expectedMask; // [nStates] bit t = type t is a legal pot
};

struct Results {
    uint64_t* cost;       // baseCost + simulationCost, ~0 when not COMPLETED
    uint32_t* flags;      // ResultFlags | nEvents << 8
    uint32_t* potted;     // bit b = ball b potted at the end
    float* finalTime;
    float* finalPos;      // [nCand][nBalls][2]  (may be null)
    int4* eventLog;       // [nCand][32] {type, a, b, layer time bits} (may be null)
    // traced strikes (bcu_simulate_trace, bcu_trace_strikes): per layer, per ball 10 floats as oracle/ref_driver.cpp
    float* traceTimes;    // [nCand][32]
    float* traceRec;      // [nCand][32][nBalls][10]
    // work counters (may be null): [0] events [1] pair items [2] quartics [3] rail items [4] pocket items
    unsigned long long* counters;
};

struct LaunchParams {
    int64_t nCand;
    const TableGeom* table;    // device copies owned by the engine (nothing table- or rules-specific is process-global)
    const RulesData* rules;
    CandidateSource src;
    StateTable states;
    Results out;
};

// ---- K1 v3, the event-synchronous batch pipeline (bcu_pipeline.cu) ----------------------------------------------
constexpr int64_t PIPE_MAX_STRIKES = 1 << 20;   // per launch; larger batches are chunked by the caller
constexpr int PIPE_MAX_STREAMS = 8;
constexpr int PIPE_JOBS_PER_STRIKE = 8;          // initial quartic-queue capacity per strike and round; on overflow the
                                                 // launch raises a flag and the host doubles the queue and re-runs it
// one layer's per-ball state, ball-major [nBalls][n] (the merged layout keeps two: layer k - 1 is read while layer k is written)
struct StateBuf {
    float2* pos;
    float2* vel;
    float2* vn;
    float2* acc;
    float* nextRolling;
    uint32_t* meta;
};
struct PipeWorkspace {
    void* base;
    int64_t n;                 // strikes this layout was made for
    size_t dynCap, jobCap;
    // per ball, ball-major [nBalls][n]: threads of consecutive strikes touch consecutive words
    float2* pos;
    float2* vel;
    float2* vn;                // velocityNormed (state::BallState, type.cpp:176-184)
    float2* acc;
    float* nextRolling;        // < 0: none  (System::_nextRolling, type.cpp:478-527)
    uint32_t* meta;            // kind | rolling << 8 | (lastRail + 1) << 16 | (type + 1) << 24
    uint32_t* lineCand;        // rail (bits 0-23) and pocket (24-31) candidates of the ball's current line of motion (bcu_items.cuh)
    // per strike
    unsigned long long* key;   // [3][n] (ordered time bits << 32 | evaluation-order code), by round % 3
    float* T;                  // [2][n] layer time, by round & 1
    uint32_t* dynMask;         // [n]
    uint32_t* restMask;        // [n]
    // stream-compacted lists, by round & 1
    uint32_t* live;            // [2][n] strike ids
    uint32_t* dyn;             // [2][dynCap] strike << 5 | ball
    uint32_t* counts;          // [rounds + 1][4] {live, dynamic balls, quartic jobs, -}; last row: {-, -, audit failures, queue overflow}
    // quartic jobs of the round
    float* jobCo;              // [5][jobCap]
    uint32_t* jobCode;         // [jobCap] (strike << 5 | ball) << 6 | partner is dynamic << 5 | partner; bit 31 = audit mark
    // constants of the launch BY VALUE (the table and the ball count do not change between launches of a captured graph: the
    // ball count is part of its key, the table belongs to the engine).  Reading them through the LaunchParams block in device
    // memory was two dependent memory round trips at the start of every kernel of the chain (params -> table pointer -> rails).
    const TableGeom* table;
    int nBalls, nRails, nPockets;
    float diameter;
    int pdlEarly;              // kernels signal cudaTriggerProgrammaticLaunchCompletion() at once (launch_pipeline)
    int pruneMode;             // reach bound on ball pairs (bcu_items.cuh): 0 off, 1 on, 2 audit (solve everything, count failures)
    // merged layout (bcu_round.cuh): double-buffered layer state, masks and "ball that stopped in the last event"
    StateBuf buf[2];           // buf[0] is the single buffer of the other layouts (pos, vel, ... above)
    uint32_t* maskDyn[2];
    uint32_t* maskRest[2];
    uint32_t* stopped[2];      // [n] ball that came to rest / was potted in the strike's last event (0xff: none)
    int merged;                // layout 2
    int fused;                 // layout of a round: 1 = step + quartic (bcu_step.cuh; one key / T buffer), 0 = items + quartic + apply
};
// Per-engine launch configuration of the pipeline (no process-wide state: two engines on two devices or threads each
// own one).  Filled by bcu_create from the BCU_* environment; the occupancy figures are computed on first use for the
// engine's device and the shared-memory size of its ball count.
struct PipeLaunchCfg {
    int occItems = 0, occQuartic = 0, occApply = 0;
    size_t occSmem = ~(size_t)0;
    int occStep = 0;
    int occItemsDense = 0, occQuarticWide = 0;
    // parts of this many strikes and more run items_kernel_dense / quartic_kernel_wide (BCU_DENSE_MIN sets both,
    // BCU_DENSE_MIN_ITEMS / BCU_DENSE_MIN_QUARTIC one of them)
    int64_t denseMinItems = 65536, denseMinQuartic = 65536;
    int layout = 0;            // BCU_LAYOUT: 0 = the three-kernel round (default), 1 = fused step kernel (bcu_step.cuh: bit-exact,
                               // measured 3-4 x slower: groups of lanes per strike leave most lanes idle), 2 = merged round
                               // kernel (bcu_round.cuh: apply folded into items, two kernels per round)
    int occRound = 0;
    int stepLanes = 8;         // BCU_STEP_LANES: lanes of a warp that own one strike in step_kernel (4, 8 or 16)
    int gridPct = 100;         // BCU_GRID_PCT (experiment): grids of the round kernels as a percentage of one resident wave
    int pdlMode = 1;           // BCU_PDL: 0 off, 1 programmatic dependent launch, 2 + early trigger
    bool roundTiming = false;  // BCU_ROUND_TIMING (diagnosis): CUDA events around every kernel, times on stderr
    bool quarticTwice = false; // BCU_QUARTIC_TWICE (diagnosis, with BCU_ROUND_TIMING): every quartic kernel is launched a second time
};
size_t pipeline_workspace_bytes(int64_t n, int nBalls, int jobsPerStrike, PipeWorkspace* layout);
const uint32_t* pipeline_overflow_flag(const PipeWorkspace& w);
void pipeline_upload_constants(cudaStream_t stream);
cudaError_t launch_pipeline(const LaunchParams* dParams, int nBalls, int nRails, int nPockets, bool withCounters,
                            const PipeWorkspace& w, int sms, cudaStream_t stream, bool resetOverflow, PipeLaunchCfg& cfg);
int pipeline_launch_count(const PipeLaunchCfg& cfg);

void upload_constants(cudaStream_t stream);   // physics constants only (identical for every engine)
cudaError_t launch_trace(const LaunchParams& p, cudaStream_t stream);

// top-k selection on (cost, candidate id) among VALID candidates; see bcu_topk.cu
cudaError_t launch_topk(const uint64_t* cost, const uint32_t* flags, int64_t nCand, int64_t firstCandidate, int k,
                        uint64_t* outKeyCost, uint64_t* outKeyId, int32_t* outCount, void* scratch, size_t scratchBytes,
                        cudaStream_t stream);
size_t topk_scratch_bytes(int64_t nCand, int k);

// beam records and their multi-GPU merge; see bcu_merge.cu
size_t record_block_bytes(int k, int nBalls);
cudaError_t launch_pack_records(const uint64_t* keyCost, const uint64_t* keyId, int k, int64_t firstCandidate,
                                const uint32_t* potted, const float* finalTime, const float* finalPos, int nBalls,
                                void* out, cudaStream_t stream);
cudaError_t launch_merge_records(const void* gathered, int G, int k, int nBalls, void* out, cudaStream_t stream);
// the reference's planner on the device; see bcu_plan.cu / bcu_planner.cuh
struct PlanTable;
int64_t plan_work_items(const PlanTable& tb, int nStates, int nBalls);
int64_t plan_blocks(const PlanTable& tb, int nStates, int nBalls);
cudaError_t launch_plan_count(const PlanTable* dTable, const PlanTable& hTable, const StateTable& states, const uint64_t* stateBase,
                              uint32_t* itemCount, uint32_t* blockSum, unsigned long long* blockOffset, unsigned long long* total,
                              cudaStream_t stream);
cudaError_t launch_plan_write(const PlanTable* dTable, const PlanTable& hTable, const StateTable& states, const uint64_t* stateBase,
                              const uint32_t* itemCount, const unsigned long long* blockOffset, int32_t* stateIdx, float* vx, float* vy,
                              uint64_t* searchCost, uint64_t* cost, cudaStream_t stream);
// measurement helpers
int launch_issue_rate(int kind, int ctas, int iters, double* sink, cudaStream_t stream);
cudaError_t launch_flush(void* buf, size_t bytes, cudaStream_t stream);

}  // namespace bcu
