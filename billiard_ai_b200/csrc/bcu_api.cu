// C-ABI layer of libbilliard_cuda.so (see include/billiard_cuda.h).  Host-side plumbing only:
// derives the rail geometry the way the reference's Rail ctor does, owns device buffers, launches
// the kernels of bcu_kernels.cu / bcu_topk.cu on one stream and copies results back.
// There is no CPU path in this library.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/billiard_cuda.h"
#include "bcu_kernels.cuh"
#include "bcu_planner.cuh"

using namespace bcu;

namespace {

template <class T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = n + n / 4 + 64;
        cudaError_t e = cudaMalloc(&p, want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

}  // namespace

struct PipeGraphKey {   // everything the captured launch sequence depends on
    void* ws; void* params; int64_t per; int G; int nBalls; int jobsPerStrike; bool counters; int pruneMode;
    bool operator==(const PipeGraphKey& o) const {
        return ws == o.ws && params == o.params && per == o.per && G == o.G && nBalls == o.nBalls &&
               jobsPerStrike == o.jobsPerStrike && counters == o.counters && pruneMode == o.pruneMode;
    }
};

struct bcu_engine {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t evStart = nullptr, evStop = nullptr;
    std::string err;
    TableGeom table;
    RulesData rules;
    std::vector<float> railDerived;
    int64_t launches = 0;
    bool countersOn = false;

    // states
    int nStates = 0, nBalls = 0;
    DevBuf<float> dStatePos;
    DevBuf<int8_t> dStateType;
    DevBuf<uint32_t> dExpected;
    DevBuf<uint64_t> dStateBase;
    bool hasStateBase = false;
    // explicit candidates
    DevBuf<int32_t> dCandState;
    DevBuf<float> dVx, dVy;
    DevBuf<uint64_t> dSearchCost, dBaseCost;
    // grid
    DevBuf<float> dDirs, dSpeeds;
    // results
    DevBuf<uint64_t> dCost;
    DevBuf<uint32_t> dFlags, dPotted;
    DevBuf<float> dFinalTime, dFinalPos;
    DevBuf<int4> dEventLog;
    DevBuf<float> dTraceTimes, dTraceRec;
    DevBuf<unsigned long long> dCounters;
    DevBuf<TableGeom> dTable;        // this engine's table geometry and rules, read by the kernels through LaunchParams
    DevBuf<RulesData> dRules;
    DevBuf<unsigned char> dPipeWs;   // workspace of the batch pipeline (bcu_pipeline.cu)
    int sms = 148;
    int jobsPerStrike = PIPE_JOBS_PER_STRIKE;
    int pruneMode = 0;                  // reach bound on ball pairs (bcu_set_prune): 0 off (default: every quartic the
                                        // reference solves is solved), 1 on (opt-in: audited, not proven), 2 audit
    PipeLaunchCfg pipeCfg;              // occupancy-sized grids, PDL mode, round timing: per engine, nothing process-wide
    bool selectTiming = false;          // BCU_SELECT_TIMING (diagnosis)
    int8_t* hTypeStage = nullptr;       // pinned staging of bcu_set_states' type codes (no per-call synchronisation)
    size_t hTypeStageCap = 0;
    cudaEvent_t evStage = nullptr;      // the last copy out of hTypeStage
    // single-strike trace (bcu_simulate_trace): its own one-state table, so a trace never clobbers the uploaded beam
    DevBuf<float> dTrPos;
    DevBuf<int8_t> dTrType;
    DevBuf<uint32_t> dTrExpected;
    uint32_t* hOverflow = nullptr;   // pinned, one word per pipeline stream
    cudaStream_t copyStream = nullptr;   // bcu_evaluate_level: result copies overlap the selection kernels
    cudaEvent_t evResults = nullptr;
    bool beamStaged = false;         // hBeam already holds the current beam block (bcu_evaluate_level copies it ahead)
    void* hBeam = nullptr;           // pinned staging for bcu_fetch_beam
    size_t hBeamCap = 0;
    const void* beamBlock = nullptr; // the record block bcu_fetch_beam reads (local block on one rank, else the merged one)
    int pipeStreams = 2;             // BCU_STREAMS: parts of a launch in flight at once (measured: 2 best at 65 536 strikes)
    cudaStream_t pipeStream[PIPE_MAX_STREAMS] = {};
    cudaEvent_t pipeDone[PIPE_MAX_STREAMS] = {};
    cudaEvent_t pipeFork = nullptr;
    DevBuf<LaunchParams> dPipeParams;   // one parameter block per stream, read by the kernels
    bool useGraph = true;               // BCU_NO_GRAPH / BCU_ROUND_TIMING: plain stream launches
    // captured launch sequences, most recently used first: a multi-break search alternates between a few launch shapes
    // (level sizes), and re-capturing 190 kernel nodes costs more than a small level takes to run
    static constexpr int PIPE_GRAPH_CACHE = 8;
    std::vector<std::pair<PipeGraphKey, cudaGraphExec_t>> pipeGraphs;
    // top-k
    DevBuf<unsigned char> dTopkScratch;
    DevBuf<uint64_t> dTopCost, dTopId;
    DevBuf<int32_t> dTopCount;
    DevBuf<uint64_t> dGatherIds;
    DevBuf<float> dGatherPos, dGatherTime;
    DevBuf<uint32_t> dGatherPotted;

    // beam record blocks (bcu_merge.cu): local top-k, gathered from all ranks, merged
    DevBuf<unsigned char> dBlockLocal, dBlockAll, dBlockMerged;
    int beamK = 0;
    bool haveBeam = false;
    // user timer + measurement helpers
    cudaEvent_t evT0 = nullptr, evT1 = nullptr;
    DevBuf<unsigned char> dFlush;
    DevBuf<double> dSink;
    // NCCL (loaded lazily with dlopen; only needed for n_ranks > 1)
    void* ncclLib = nullptr;
    void* ncclComm = nullptr;
    int nRanks = 1, rank = 0;
    int64_t idBase = 0;  // global id of candidate 0 of the next bcu_evaluate (multi-GPU shards of explicit lists)

    // the reference's planner on the device (bcu_set_planner / bcu_plan)
    bool havePlanner = false;
    PlanTable planTable{};
    DevBuf<PlanTable> dPlanTable;
    DevBuf<uint32_t> dPlanItemCount, dPlanBlockSum;
    DevBuf<unsigned long long> dPlanBlockOffset, dPlanTotal;
    int64_t planCount = -1;          // strikes of the last bcu_plan (in dCandState / dVx / dVy / dSearchCost / dBaseCost)
    int planStates = 0;              // nStates the plan was made for

    LaunchParams last{};
    bool haveLast = false;
    int64_t lastFirst = 0;
};

// ---- NCCL through dlopen: the library has no link-time dependency on it -----------------------------------
namespace {
struct NcclApi {
    struct Uid { char b[128]; };  // ncclUniqueId (NCCL_UNIQUE_ID_BYTES = 128), passed by value
    void* lib = nullptr;
    int (*GetUniqueId)(void*) = nullptr;
    int (*CommInitRank)(void**, int, Uid, int) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
// The one piece of process-wide state: the function table of the dlopen'ed NCCL library (a shared object is loaded once per
// process whatever we do).  Filled once under a lock; read-only afterwards.
NcclApi g_nccl;
std::mutex g_ncclLock;
bool load_nccl(const char* path, std::string* err) {
    std::lock_guard<std::mutex> guard(g_ncclLock);
    if (g_nccl.lib) return true;
    const char* cands[] = {path, getenv("BCU_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    void* h = nullptr;
    for (const char* c : cands) {
        if (!c || !*c) continue;
        h = dlopen(c, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) { *err = std::string("cannot dlopen NCCL: ") + dlerror(); return false; }
    g_nccl.GetUniqueId = (decltype(g_nccl.GetUniqueId))dlsym(h, "ncclGetUniqueId");
    g_nccl.CommInitRank = (decltype(g_nccl.CommInitRank))dlsym(h, "ncclCommInitRank");
    g_nccl.AllGather = (decltype(g_nccl.AllGather))dlsym(h, "ncclAllGather");
    g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy))dlsym(h, "ncclCommDestroy");
    g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))dlsym(h, "ncclGetErrorString");
    if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllGather || !g_nccl.CommDestroy) {
        *err = "NCCL library lacks required symbols";
        return false;
    }
    g_nccl.lib = h;
    return true;
}
}  // namespace

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t _e = (call);                                                                   \
        if (_e != cudaSuccess) {                                                                   \
            e->err = std::string(#call) + ": " + cudaGetErrorString(_e);                           \
            return BCU_ERR_CUDA;                                                                   \
        }                                                                                          \
    } while (0)

static int fail(bcu_engine* e, int code, const char* msg) {
    if (e) e->err = msg;
    return code;
}

// search::Rail ctor (reference type.cpp:56-72), binary32:
//   normal = glm::normalize(perp(end - start)), perp(v) = (-v.y, v.x); shifted = p + shiftDirection * ballRadius
static void derive_rail(const float* r6, float radius, RailGeom* g, float* derived6) {
    const float sx = r6[0], sy = r6[1], ex = r6[2], ey = r6[3], hx = r6[4], hy = r6[5];
    const float dx = ex - sx, dy = ey - sy;
    const float px = -dy, py = dx;
    const float tx = px * px, ty = py * py;
    const float inv = 1.0f / std::sqrt(tx + ty);
    const float nx = px * inv, ny = py * inv;
    const float ssx = sx + hx * radius, ssy = sy + hy * radius;
    const float sex = ex + hx * radius, sey = ey + hy * radius;
    g->sx = ssx; g->sy = ssy;
    g->dx = sex - ssx; g->dy = sey - ssy;  // billiard_physics.cpp:731-734: end - start, taken at each call
    g->nx = nx; g->ny = ny;
    derived6[0] = nx; derived6[1] = ny; derived6[2] = ssx; derived6[3] = ssy; derived6[4] = sex; derived6[5] = sey;
}

extern "C" {

const char* bcu_version(void) { return "billiard_cuda 0.1 (sm_100a)"; }

int bcu_create(const bcu_table_desc* table, const bcu_rules_desc* rules, int device, bcu_engine** out) {
    if (!out) return BCU_ERR_INVALID;
    *out = nullptr;
    if (!table || !rules || (!table->rails && !table->rails_derived) || !table->pockets || !rules->score) return BCU_ERR_INVALID;
    if (table->n_rails < 0 || table->n_rails > MAX_RAILS || table->n_pockets < 0 || table->n_pockets > MAX_POCKETS ||
        rules->n_types < 1 || rules->n_types > MAX_TYPES)
        return BCU_ERR_INVALID;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0 || device < 0 || device >= count) {
        cudaGetLastError();
        return BCU_ERR_NO_DEVICE;  // no CPU fallback
    }
    bcu_engine* e = new bcu_engine();
    e->device = device;
    if (cudaSetDevice(device) != cudaSuccess) { delete e; return BCU_ERR_NO_DEVICE; }
    std::memset(&e->table, 0, sizeof(e->table));
    std::memset(&e->rules, 0, sizeof(e->rules));
    e->table.nRails = table->n_rails;
    e->table.nPockets = table->n_pockets;
    e->table.diameter = table->ball_radius * 2;  // util.cpp:65: conf._ball._diameter = ballRadius * 2
    e->table.diameterSq = e->table.diameter * e->table.diameter;
    e->railDerived.resize((size_t)table->n_rails * 6);
    for (int i = 0; i < table->n_rails; ++i) {
        if (table->rails_derived) {
            const float* d = table->rails_derived + 6 * i;
            RailGeom& g = e->table.rails[i];
            g.nx = d[0]; g.ny = d[1]; g.sx = d[2]; g.sy = d[3];
            g.dx = d[4] - d[2]; g.dy = d[5] - d[3];  // billiard_physics.cpp:731-734
            std::memcpy(&e->railDerived[6 * i], d, 6 * sizeof(float));
        } else {
            derive_rail(table->rails + 6 * i, table->ball_radius, &e->table.rails[i], &e->railDerived[6 * i]);
        }
    }
    for (int i = 0; i < table->n_pockets; ++i) {
        const float* p = table->pockets + 3 * i;
        e->table.pockets[i].x = p[0]; e->table.pockets[i].y = p[1]; e->table.pockets[i].r = p[2];
        e->table.pockets[i].r2 = p[2] * p[2];
    }
    e->rules.nTypes = rules->n_types;
    e->rules.cueType = rules->cue_type;
    for (int i = 0; i < rules->n_types; ++i) e->rules.score[i] = rules->score[i];
    if (cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&e->evStart) != cudaSuccess || cudaEventCreate(&e->evStop) != cudaSuccess ||
        cudaEventCreate(&e->evT0) != cudaSuccess || cudaEventCreate(&e->evT1) != cudaSuccess) {
        delete e;
        return BCU_ERR_CUDA;
    }
    upload_constants(e->stream);
    if (e->dTable.reserve(1) != cudaSuccess || e->dRules.reserve(1) != cudaSuccess ||
        cudaMemcpyAsync(e->dTable.p, &e->table, sizeof(TableGeom), cudaMemcpyHostToDevice, e->stream) != cudaSuccess ||
        cudaMemcpyAsync(e->dRules.p, &e->rules, sizeof(RulesData), cudaMemcpyHostToDevice, e->stream) != cudaSuccess ||
        cudaStreamSynchronize(e->stream) != cudaSuccess) {
        delete e;
        return BCU_ERR_CUDA;
    }
    cudaDeviceGetAttribute(&e->sms, cudaDevAttrMultiProcessorCount, device);
    if (getenv("BCU_NO_GRAPH") || getenv("BCU_ROUND_TIMING")) e->useGraph = false;
    e->pipeCfg.roundTiming = getenv("BCU_ROUND_TIMING") != nullptr;
    e->pipeCfg.quarticTwice = getenv("BCU_QUARTIC_TWICE") != nullptr;
    if (const char* pv = getenv("BCU_PDL")) e->pipeCfg.pdlMode = atoi(pv);
    if (const char* gv = getenv("BCU_GRID_PCT")) { const int v = atoi(gv); e->pipeCfg.gridPct = v < 5 ? 5 : v > 400 ? 400 : v; }
    if (const char* dv = getenv("BCU_DENSE_MIN")) { const long long v = atoll(dv); e->pipeCfg.denseMinItems = e->pipeCfg.denseMinQuartic = v < 0 ? 0 : v; }
    if (const char* dv = getenv("BCU_DENSE_MIN_ITEMS")) { const long long v = atoll(dv); e->pipeCfg.denseMinItems = v < 0 ? 0 : v; }
    if (const char* dv = getenv("BCU_DENSE_MIN_QUARTIC")) { const long long v = atoll(dv); e->pipeCfg.denseMinQuartic = v < 0 ? 0 : v; }
    if (const char* lv = getenv("BCU_LAYOUT")) { const int v = atoi(lv); e->pipeCfg.layout = v == 1 ? 1 : v == 2 ? 2 : 0; }
    if (const char* lv = getenv("BCU_STEP_LANES")) e->pipeCfg.stepLanes = atoi(lv);
    e->selectTiming = getenv("BCU_SELECT_TIMING") != nullptr;
    if (const char* jv = getenv("BCU_JOBS_PER_STRIKE")) {   // initial quartic-queue capacity (tests force the overflow path)
        const int v = atoi(jv);
        e->jobsPerStrike = v < 1 ? 1 : v > 1024 ? 1024 : v;
    }
    if (const char* pv = getenv("BCU_PRUNE")) {
        const int v = atoi(pv);
        e->pruneMode = v < 0 ? 0 : v > 2 ? 2 : v;
    }
    if (const char* sv = getenv("BCU_STREAMS")) {
        const int v = atoi(sv);
        e->pipeStreams = v < 1 ? 1 : v > PIPE_MAX_STREAMS ? PIPE_MAX_STREAMS : v;
    }
    if (cudaGetLastError() != cudaSuccess) { delete e; return BCU_ERR_CUDA; }
    *out = e;
    return BCU_OK;
}

void bcu_destroy(bcu_engine* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    cudaStreamSynchronize(e->stream);
    e->dStatePos.release(); e->dStateType.release(); e->dExpected.release(); e->dStateBase.release();
    e->dCandState.release(); e->dVx.release(); e->dVy.release(); e->dSearchCost.release(); e->dBaseCost.release();
    e->dDirs.release(); e->dSpeeds.release();
    e->dCost.release(); e->dFlags.release(); e->dPotted.release(); e->dFinalTime.release(); e->dFinalPos.release();
    e->dEventLog.release(); e->dTraceTimes.release(); e->dTraceRec.release(); e->dCounters.release(); e->dPipeWs.release();
    e->dTable.release(); e->dRules.release();
    e->dTrPos.release(); e->dTrType.release(); e->dTrExpected.release();
    e->dPlanTable.release(); e->dPlanItemCount.release(); e->dPlanBlockSum.release(); e->dPlanBlockOffset.release(); e->dPlanTotal.release();
    if (e->hTypeStage) cudaFreeHost(e->hTypeStage);
    if (e->evStage) cudaEventDestroy(e->evStage);
    e->dTopkScratch.release(); e->dTopCost.release(); e->dTopId.release(); e->dTopCount.release();
    e->dGatherIds.release(); e->dGatherPos.release(); e->dGatherTime.release(); e->dGatherPotted.release();
    e->dBlockLocal.release(); e->dBlockAll.release(); e->dBlockMerged.release(); e->dFlush.release(); e->dSink.release();
    if (e->ncclComm && g_nccl.CommDestroy) g_nccl.CommDestroy(e->ncclComm);
    if (e->hOverflow) cudaFreeHost(e->hOverflow);
    if (e->hBeam) cudaFreeHost(e->hBeam);
    if (e->copyStream) cudaStreamDestroy(e->copyStream);
    if (e->evResults) cudaEventDestroy(e->evResults);
    for (int g = 0; g < PIPE_MAX_STREAMS; ++g) {
        if (e->pipeDone[g]) cudaEventDestroy(e->pipeDone[g]);
        if (e->pipeStream[g]) cudaStreamDestroy(e->pipeStream[g]);
    }
    if (e->pipeFork) cudaEventDestroy(e->pipeFork);
    for (auto& kv : e->pipeGraphs) cudaGraphExecDestroy(kv.second);
    e->dPipeParams.release();
    if (e->evT0) cudaEventDestroy(e->evT0);
    if (e->evT1) cudaEventDestroy(e->evT1);
    if (e->evStart) cudaEventDestroy(e->evStart);
    if (e->evStop) cudaEventDestroy(e->evStop);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
}

int bcu_host_alloc(size_t bytes, void** out) {
    if (!out || bytes == 0) return BCU_ERR_INVALID;
    return cudaHostAlloc(out, bytes, cudaHostAllocDefault) == cudaSuccess ? BCU_OK : BCU_ERR_CUDA;
}
int bcu_host_free(void* p) { return cudaFreeHost(p) == cudaSuccess ? BCU_OK : BCU_ERR_CUDA; }

const char* bcu_last_error(const bcu_engine* e) { return e ? e->err.c_str() : "null engine"; }

int bcu_get_rail_derived(const bcu_engine* e, float* out6) {
    if (!e || !out6) return BCU_ERR_INVALID;
    std::memcpy(out6, e->railDerived.data(), e->railDerived.size() * sizeof(float));
    return BCU_OK;
}

int bcu_set_states(bcu_engine* e, int32_t nStates, int32_t nBalls, const float* pos, const int32_t* types,
                   const uint32_t* expectedMask, const uint64_t* baseCost) {
    if (!e) return BCU_ERR_INVALID;
    if (nStates < 1 || nBalls < 1 || nBalls > MAX_BALLS || !pos || !types || !expectedMask)
        return fail(e, BCU_ERR_INVALID, "bcu_set_states: bad arguments");
    CK(cudaSetDevice(e->device));
    const size_t nb = (size_t)nStates * nBalls;
    // The 8-bit type codes are built in page-locked staging owned by the engine, so the call does not have to wait for its
    // copies (copies from the caller's pageable arrays are staged by the runtime before cudaMemcpyAsync returns).
    if (!e->evStage) CK(cudaEventCreateWithFlags(&e->evStage, cudaEventDisableTiming));
    else CK(cudaEventSynchronize(e->evStage));   // the previous upload has left the staging (normally long ago)
    if (e->hTypeStageCap < nb) {
        if (e->hTypeStage) cudaFreeHost(e->hTypeStage);
        e->hTypeStage = nullptr; e->hTypeStageCap = 0;
        CK(cudaMallocHost(&e->hTypeStage, nb + nb / 4 + 64));
        e->hTypeStageCap = nb + nb / 4 + 64;
    }
    int8_t* t8 = e->hTypeStage;
    for (size_t i = 0; i < nb; ++i) {
        if (types[i] >= e->rules.nTypes) return fail(e, BCU_ERR_INVALID, "bcu_set_states: type code out of range");
        t8[i] = (int8_t)(types[i] < 0 ? -1 : types[i]);
    }
    CK(e->dStatePos.reserve(nb * 2));
    CK(e->dStateType.reserve(nb));
    CK(e->dExpected.reserve(nStates));
    CK(cudaMemcpyAsync(e->dStatePos.p, pos, nb * 2 * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(e->dStateType.p, t8, nb, cudaMemcpyHostToDevice, e->stream));
    CK(cudaEventRecord(e->evStage, e->stream));
    CK(cudaMemcpyAsync(e->dExpected.p, expectedMask, nStates * sizeof(uint32_t), cudaMemcpyHostToDevice, e->stream));
    e->hasStateBase = baseCost != nullptr;
    if (baseCost) {
        CK(e->dStateBase.reserve(nStates));
        CK(cudaMemcpyAsync(e->dStateBase.p, baseCost, nStates * sizeof(uint64_t), cudaMemcpyHostToDevice, e->stream));
    }
    e->nStates = nStates;
    e->nBalls = nBalls;
    e->haveLast = false;
    e->planCount = -1;
    return BCU_OK;
}

static int reserve_results(bcu_engine* e, int64_t n, bool wantPos, bool wantLog) {
    if (n < 1) n = 1;   // an empty shard still gets valid (unused) result pointers
    CK(e->dCost.reserve(n)); CK(e->dFlags.reserve(n)); CK(e->dPotted.reserve(n)); CK(e->dFinalTime.reserve(n));
    if (wantPos) CK(e->dFinalPos.reserve((size_t)n * e->nBalls * 2));
    if (wantLog) CK(e->dEventLog.reserve((size_t)n * 32));
    CK(e->dCounters.reserve(8));
    return BCU_OK;
}

static void fill_common(bcu_engine* e, LaunchParams& p, int64_t n, bool wantPos, bool wantLog) {
    std::memset(&p, 0, sizeof(p));
    p.nCand = n;
    p.table = e->dTable.p; p.rules = e->dRules.p;
    p.states.nStates = e->nStates; p.states.nBalls = e->nBalls;
    p.states.pos = e->dStatePos.p; p.states.type = e->dStateType.p; p.states.expectedMask = e->dExpected.p;
    p.out.cost = e->dCost.p; p.out.flags = e->dFlags.p; p.out.potted = e->dPotted.p; p.out.finalTime = e->dFinalTime.p;
    p.out.finalPos = wantPos ? e->dFinalPos.p : nullptr;
    p.out.eventLog = wantLog ? e->dEventLog.p : nullptr;
    p.out.counters = e->countersOn ? e->dCounters.p : nullptr;
}

static int enqueue_copy_back(bcu_engine* e, int64_t offset, int64_t n, const bcu_results* out, cudaStream_t stream);
static int reserve_beam_staging(bcu_engine* e, size_t bytes) {
    if (e->hBeamCap < bytes) {
        CK(cudaStreamSynchronize(e->stream));
        if (e->hBeam) cudaFreeHost(e->hBeam);
        e->hBeam = nullptr; e->hBeamCap = 0;
        CK(cudaMallocHost(&e->hBeam, bytes));
        e->hBeamCap = bytes;
    }
    return BCU_OK;
}
extern "C" int bcu_level_select(bcu_engine* e, int32_t k);

// `out` (batch path, single wave): the device->host copies of the results are queued right behind the launch, before the
// host waits for it, so the transfer starts the moment the last kernel ends (a retry after a job-queue overflow copies again)
// `selectK` > 0 (one rank): bcu_level_select is queued behind the launch as well, and the result copies move to a second
// stream so that they overlap the selection kernels; *selected tells the caller that the beam is on its way.
static int run(bcu_engine* e, const LaunchParams& p, bool trace, const bcu_results* out = nullptr, bool* copied = nullptr,
               int selectK = 0, bool* selected = nullptr) {
    if (p.out.counters) CK(cudaMemsetAsync(p.out.counters, 0, 8 * sizeof(unsigned long long), e->stream));
    if (p.out.eventLog && trace) CK(cudaMemsetAsync(p.out.eventLog, 0xff, (size_t)p.nCand * 32 * sizeof(int4), e->stream));
    CK(cudaEventRecord(e->evStart, e->stream));
    if (trace) {
        CK(launch_trace(p, e->stream));
        e->launches++;
    } else {
        // batch path: the event-synchronous pipeline.  A launch is cut into parts that run on separate streams: every
        // part is a chain of ~100 short kernels whose rounds are latency-bound, so independent parts fill each other's
        // gaps (a part is still large enough to keep the stream-compacted kernels dense).  The whole fork / chains / join
        // sequence is captured once into a CUDA graph (kernel parameters live in device memory, grids are persistent),
        // so a level costs one graph launch instead of several hundred kernel launches.
        const int nBalls = p.states.nBalls;
        if (p.nCand == 0) {   // an empty shard (more ranks than candidates): nothing to launch, the timer still closes
            CK(cudaEventRecord(e->evStop, e->stream));
            return BCU_OK;
        }
        int64_t per = (p.nCand + e->pipeStreams - 1) / e->pipeStreams;
        if (per < 8192) per = 8192;
        if (per > PIPE_MAX_STRIKES) per = PIPE_MAX_STRIKES;
        per = (per + 255) & ~(int64_t)255;
        const int64_t nParts = (p.nCand + per - 1) / per;
        const int G = (int)(nParts < e->pipeStreams ? nParts : e->pipeStreams);
        const bool withCounters = p.out.counters != nullptr;
        if (!e->hOverflow) CK(cudaMallocHost(&e->hOverflow, PIPE_MAX_STREAMS * sizeof(uint32_t)));
        for (int g = 0; g < G; ++g) {
            if (!e->pipeStream[g]) {
                CK(cudaStreamCreateWithFlags(&e->pipeStream[g], cudaStreamNonBlocking));
                CK(cudaEventCreateWithFlags(&e->pipeDone[g], cudaEventDisableTiming));
            }
        }
        if (!e->pipeFork) CK(cudaEventCreateWithFlags(&e->pipeFork, cudaEventDisableTiming));
        CK(e->dPipeParams.reserve(PIPE_MAX_STREAMS));
        for (int64_t wave = 0; wave < nParts; wave += G) {
            for (;;) {
                const size_t wsBytes = pipeline_workspace_bytes(per, nBalls, e->jobsPerStrike, nullptr);
                {   // a captured graph must not outlive the allocation its memset / copy / kernel nodes point into: a later
                    // allocation at the same address is a different allocation to the driver (cuGraphLaunch crashed on it)
                    const size_t capBefore = e->dPipeWs.cap;
                    CK(e->dPipeWs.reserve(wsBytes * G));
                    if (e->dPipeWs.cap != capBefore) {
                        for (auto& kv : e->pipeGraphs) cudaGraphExecDestroy(kv.second);
                        e->pipeGraphs.clear();
                    }
                }
                // this wave's parameter blocks
                LaunchParams hp[PIPE_MAX_STREAMS];
                for (int g = 0; g < G; ++g) {
                    const int64_t off = (wave + g) * per;
                    LaunchParams q = p;
                    q.nCand = off >= p.nCand ? 0 : (p.nCand - off < per ? p.nCand - off : per);
                    if (q.nCand > 0) {
                        if (q.src.gridDirs > 0) {
                            q.src.firstCandidate += off;
                        } else {
                            q.src.stateIdx += off; q.src.vx += off; q.src.vy += off; q.src.searchCost += off; q.src.baseCost += off;
                        }
                        q.out.cost += off; q.out.flags += off; q.out.potted += off; q.out.finalTime += off;
                        if (q.out.finalPos) q.out.finalPos += (size_t)off * nBalls * 2;
                        if (q.out.eventLog) q.out.eventLog += (size_t)off * 32;
                    }
                    hp[g] = q;
                }
                CK(cudaMemcpyAsync(e->dPipeParams.p, hp, G * sizeof(LaunchParams), cudaMemcpyHostToDevice, e->stream));
                if (p.out.eventLog) {   // -1 padding of this wave's logs (again on a retry: a run that overflowed logged
                                        // events of strikes that went wrong)
                    const int64_t off = wave * per;
                    const int64_t cnt = p.nCand - off < (int64_t)G * per ? p.nCand - off : (int64_t)G * per;
                    CK(cudaMemsetAsync(p.out.eventLog + (size_t)off * 32, 0xff, (size_t)cnt * 32 * sizeof(int4), e->stream));
                }
                PipeGraphKey key{e->dPipeWs.p, e->dPipeParams.p, per, G, nBalls, e->jobsPerStrike, withCounters, e->pruneMode};
                const bool useGraph = e->useGraph;
                cudaGraphExec_t graphExec = nullptr;
                for (size_t gi = 0; useGraph && gi < e->pipeGraphs.size(); ++gi) {
                    if (e->pipeGraphs[gi].first == key) {
                        auto hit = e->pipeGraphs[gi];
                        e->pipeGraphs.erase(e->pipeGraphs.begin() + gi);
                        e->pipeGraphs.insert(e->pipeGraphs.begin(), hit);
                        graphExec = hit.second;
                        break;
                    }
                }
                if (!useGraph || !graphExec) {
                    if (useGraph) CK(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
                    CK(cudaEventRecord(e->pipeFork, e->stream));
                    for (int g = 0; g < G; ++g) {
                        PipeWorkspace ws{};
                        ws.base = e->dPipeWs.p + wsBytes * g;
                        pipeline_workspace_bytes(per, nBalls, e->jobsPerStrike, &ws);
                        ws.pruneMode = e->pruneMode;
                        ws.table = e->dTable.p; ws.nBalls = nBalls; ws.nRails = e->table.nRails; ws.nPockets = e->table.nPockets;
                        ws.diameter = e->table.diameter;
                        CK(cudaStreamWaitEvent(e->pipeStream[g], e->pipeFork, 0));
                        CK(launch_pipeline(e->dPipeParams.p + g, nBalls, e->table.nRails, e->table.nPockets, withCounters, ws, e->sms, e->pipeStream[g], true, e->pipeCfg));
                        // the quartic job queue is sized for the common case; a round that posts more jobs raises a flag
                        CK(cudaMemcpyAsync(e->hOverflow + g, pipeline_overflow_flag(ws), sizeof(uint32_t), cudaMemcpyDeviceToHost,
                                           e->pipeStream[g]));
                        CK(cudaEventRecord(e->pipeDone[g], e->pipeStream[g]));
                        CK(cudaStreamWaitEvent(e->stream, e->pipeDone[g], 0));
                    }
                    if (useGraph) {
                        cudaGraph_t graph = nullptr;
                        CK(cudaStreamEndCapture(e->stream, &graph));
                        CK(cudaGraphInstantiate(&graphExec, graph, 0));
                        cudaGraphDestroy(graph);
                        if ((int)e->pipeGraphs.size() >= bcu_engine::PIPE_GRAPH_CACHE) {
                            cudaGraphExecDestroy(e->pipeGraphs.back().second);
                            e->pipeGraphs.pop_back();
                        }
                        e->pipeGraphs.insert(e->pipeGraphs.begin(), {key, graphExec});
                    }
                }
                if (useGraph) CK(cudaGraphLaunch(graphExec, e->stream));
                e->launches += (int64_t)G * pipeline_launch_count(e->pipeCfg);
                CK(cudaEventRecord(e->evStop, e->stream));
                bool onCopyStream = false;
                if (out && nParts <= G) {
                    cudaStream_t cs = e->stream;
                    if (selectK > 0 && e->nRanks == 1) {
                        if (!e->copyStream) {
                            CK(cudaStreamCreateWithFlags(&e->copyStream, cudaStreamNonBlocking));
                            CK(cudaEventCreateWithFlags(&e->evResults, cudaEventDisableTiming));
                        }
                        CK(cudaEventRecord(e->evResults, e->stream));
                        CK(cudaStreamWaitEvent(e->copyStream, e->evResults, 0));
                        cs = e->copyStream;
                        onCopyStream = true;
                    }
                    int rc = enqueue_copy_back(e, 0, p.nCand, out, cs);
                    if (rc) return rc;
                    if (copied) *copied = true;
                    if (onCopyStream) {
                        rc = bcu_level_select(e, selectK);
                        if (rc) return rc;
                        // the beam block follows into page-locked staging: bcu_fetch_beam only copies host memory
                        const size_t bb = record_block_bytes(selectK, e->nBalls);
                        rc = reserve_beam_staging(e, bb);
                        if (rc) return rc;
                        CK(cudaMemcpyAsync(e->hBeam, e->beamBlock, bb, cudaMemcpyDeviceToHost, e->stream));
                        e->beamStaged = true;
                        if (selected) *selected = true;
                    }
                }
                CK(cudaStreamSynchronize(e->stream));
                if (onCopyStream) CK(cudaStreamSynchronize(e->copyStream));
                bool overflow = false;
                for (int g = 0; g < G; ++g) overflow |= e->hOverflow[g] != 0;
                if (!overflow) break;
                if (e->jobsPerStrike >= 1024) return fail(e, BCU_ERR_STATE, "quartic job queue overflow");
                e->jobsPerStrike *= 2;
            }
        }
    }
    if (trace) CK(cudaEventRecord(e->evStop, e->stream));
    return BCU_OK;
}

static int enqueue_copy_back(bcu_engine* e, int64_t offset, int64_t n, const bcu_results* out, cudaStream_t stream) {
    if (!out) return BCU_OK;
    if (out->cost) CK(cudaMemcpyAsync(out->cost, e->dCost.p + offset, n * sizeof(uint64_t), cudaMemcpyDeviceToHost, stream));
    if (out->flags) CK(cudaMemcpyAsync(out->flags, e->dFlags.p + offset, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
    if (out->potted) CK(cudaMemcpyAsync(out->potted, e->dPotted.p + offset, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
    if (out->final_time) CK(cudaMemcpyAsync(out->final_time, e->dFinalTime.p + offset, n * sizeof(float), cudaMemcpyDeviceToHost, stream));
    if (out->final_pos) {
        if (!e->last.out.finalPos) return fail(e, BCU_ERR_STATE, "final_pos was not produced by the last launch");
        CK(cudaMemcpyAsync(out->final_pos, e->dFinalPos.p + (size_t)offset * e->nBalls * 2,
                           (size_t)n * e->nBalls * 2 * sizeof(float), cudaMemcpyDeviceToHost, stream));
    }
    if (out->event_log) {
        if (!e->last.out.eventLog) return fail(e, BCU_ERR_STATE, "event_log was not produced by the last launch");
        CK(cudaMemcpyAsync(out->event_log, e->dEventLog.p + (size_t)offset * 32, (size_t)n * 32 * sizeof(int4),
                           cudaMemcpyDeviceToHost, stream));
    }
    return BCU_OK;
}

static int copy_back(bcu_engine* e, int64_t offset, int64_t n, const bcu_results* out) {
    int rc = enqueue_copy_back(e, offset, n, out, e->stream);
    if (rc) return rc;
    CK(cudaStreamSynchronize(e->stream));
    return BCU_OK;
}

static int evaluate_impl(bcu_engine* e, int64_t n, const int32_t* stateIdx, const float* vx, const float* vy,
                         const uint64_t* searchCost, const uint64_t* baseCost, const bcu_results* out, int selectK) {
    if (!e) return BCU_ERR_INVALID;
    if (e->nStates == 0) return fail(e, BCU_ERR_STATE, "bcu_evaluate: call bcu_set_states first");
    if (n < 0 || (n > 0 && (!stateIdx || !vx || !vy || !searchCost || !baseCost)))
        return fail(e, BCU_ERR_INVALID, "bcu_evaluate: bad arguments");
    if (n == 0) return BCU_OK;
    CK(cudaSetDevice(e->device));
    CK(e->dCandState.reserve(n)); CK(e->dVx.reserve(n)); CK(e->dVy.reserve(n));
    CK(e->dSearchCost.reserve(n)); CK(e->dBaseCost.reserve(n));
    e->planCount = -1;   // the candidate arrays now hold the caller's strikes
    const bool wantPos = out && out->final_pos, wantLog = out && out->event_log;
    // final positions are always produced on the device (the beam records need them); they are copied out on request
    int rc = reserve_results(e, n, true, wantLog);
    if (rc) return rc;
    CK(cudaMemcpyAsync(e->dCandState.p, stateIdx, n * sizeof(int32_t), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(e->dVx.p, vx, n * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(e->dVy.p, vy, n * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(e->dSearchCost.p, searchCost, n * sizeof(uint64_t), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(e->dBaseCost.p, baseCost, n * sizeof(uint64_t), cudaMemcpyHostToDevice, e->stream));
    // argument check while the copies are in flight (nothing has been launched on them yet)
    int32_t lowest = 0, highest = 0;
    for (int64_t i = 0; i < n; ++i) { const int32_t s = stateIdx[i]; lowest = s < lowest ? s : lowest; highest = s > highest ? s : highest; }
    if (lowest < 0 || highest >= e->nStates) {
        cudaStreamSynchronize(e->stream);
        return fail(e, BCU_ERR_INVALID, "bcu_evaluate: state index out of range");
    }
    LaunchParams p;
    fill_common(e, p, n, true, wantLog);
    (void)wantPos;
    p.src.stateIdx = e->dCandState.p; p.src.vx = e->dVx.p; p.src.vy = e->dVy.p;
    p.src.searchCost = e->dSearchCost.p; p.src.baseCost = e->dBaseCost.p;
    e->last = p; e->haveLast = true; e->lastFirst = e->idBase;   // (the copies and the selection look at the launch)
    bool copied = false, selected = false;
    rc = run(e, p, false, out, &copied, selectK, &selected);
    if (rc) { e->haveLast = false; return rc; }
    if (!copied) { rc = copy_back(e, 0, n, out); if (rc) return rc; }
    if (selectK > 0 && !selected) return bcu_level_select(e, selectK);
    return BCU_OK;
}

int bcu_evaluate(bcu_engine* e, int64_t n, const int32_t* stateIdx, const float* vx, const float* vy,
                 const uint64_t* searchCost, const uint64_t* baseCost, const bcu_results* out) {
    return evaluate_impl(e, n, stateIdx, vx, vy, searchCost, baseCost, out, 0);
}

int bcu_evaluate_level(bcu_engine* e, int64_t n, const int32_t* stateIdx, const float* vx, const float* vy,
                       const uint64_t* searchCost, const uint64_t* baseCost, const bcu_results* out, int32_t k) {
    if (k < 1 || k > 1024) return e ? fail(e, BCU_ERR_INVALID, "bcu_evaluate_level: k must be in 1..1024") : BCU_ERR_INVALID;
    return evaluate_impl(e, n, stateIdx, vx, vy, searchCost, baseCost, out, k);
}

int bcu_evaluate_grid(bcu_engine* e, int32_t nDirs, int32_t nSpeeds, const float* dirs, const float* speeds,
                      uint64_t searchCost, int64_t first, int64_t count) {
    if (!e) return BCU_ERR_INVALID;
    if (e->nStates == 0) return fail(e, BCU_ERR_STATE, "bcu_evaluate_grid: call bcu_set_states first");
    if (nDirs < 1 || nSpeeds < 1 || !dirs || !speeds) return fail(e, BCU_ERR_INVALID, "bcu_evaluate_grid: bad arguments");
    if ((int64_t)nDirs * nSpeeds > INT32_MAX) return fail(e, BCU_ERR_INVALID, "bcu_evaluate_grid: n_dirs * n_speeds exceeds 2^31 - 1");
    const int64_t total = (int64_t)e->nStates * nDirs * nSpeeds;
    if (count < 0) count = total - first;
    if (first < 0 || count < 0 || first + count > total) return fail(e, BCU_ERR_INVALID, "bcu_evaluate_grid: shard out of range");
    CK(cudaSetDevice(e->device));
    CK(e->dDirs.reserve((size_t)nDirs * 2)); CK(e->dSpeeds.reserve(nSpeeds));
    CK(cudaMemcpyAsync(e->dDirs.p, dirs, (size_t)nDirs * 2 * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(e->dSpeeds.p, speeds, nSpeeds * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    int rc = reserve_results(e, count, true, false);
    if (rc) return rc;
    LaunchParams p;
    fill_common(e, p, count, true, false);
    p.src.gridDirs = nDirs; p.src.gridSpeeds = nSpeeds; p.src.dirTable = e->dDirs.p; p.src.speedTable = e->dSpeeds.p;
    p.src.gridSearchCost = searchCost;
    p.src.stateBaseCost = e->hasStateBase ? e->dStateBase.p : nullptr;
    p.src.firstCandidate = first;
    rc = run(e, p, false);
    if (rc) return rc;
    e->last = p; e->haveLast = true; e->lastFirst = first;
    CK(cudaStreamSynchronize(e->stream));
    return BCU_OK;
}

int bcu_rerun(bcu_engine* e) {
    if (!e) return BCU_ERR_INVALID;
    if (!e->haveLast) return fail(e, BCU_ERR_STATE, "bcu_rerun: nothing to re-run");
    CK(cudaSetDevice(e->device));
    int rc = run(e, e->last, false);
    if (rc) return rc;
    CK(cudaStreamSynchronize(e->stream));
    return BCU_OK;
}

int bcu_fetch_results(bcu_engine* e, int64_t offset, int64_t n, const bcu_results* out) {
    if (!e) return BCU_ERR_INVALID;
    if (!e->haveLast) return fail(e, BCU_ERR_STATE, "bcu_fetch_results: no launch yet");
    if (offset < 0 || n < 0 || offset + n > e->last.nCand) return fail(e, BCU_ERR_INVALID, "bcu_fetch_results: range");
    CK(cudaSetDevice(e->device));
    return copy_back(e, offset, n, out);
}

int bcu_topk(bcu_engine* e, int32_t k, uint64_t* outCost, uint64_t* outId, int32_t* outCount) {
    if (!e) return BCU_ERR_INVALID;
    if (!e->haveLast) return fail(e, BCU_ERR_STATE, "bcu_topk: no launch yet");
    if (k < 1 || k > 1024 || !outCost || !outId || !outCount) return fail(e, BCU_ERR_INVALID, "bcu_topk: bad arguments");
    CK(cudaSetDevice(e->device));
    const int64_t n = e->last.nCand;
    const size_t sb = topk_scratch_bytes(n, k);
    CK(e->dTopkScratch.reserve(sb)); CK(e->dTopCost.reserve(k)); CK(e->dTopId.reserve(k)); CK(e->dTopCount.reserve(1));
    CK(launch_topk(e->dCost.p, e->dFlags.p, n, e->lastFirst, k, e->dTopCost.p, e->dTopId.p,
                   e->dTopCount.p, e->dTopkScratch.p, e->dTopkScratch.cap, e->stream));
    e->launches += 4;
    int32_t valid = 0;
    CK(cudaMemcpyAsync(outCost, e->dTopCost.p, k * sizeof(uint64_t), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaMemcpyAsync(outId, e->dTopId.p, k * sizeof(uint64_t), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaMemcpyAsync(&valid, e->dTopCount.p, sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    *outCount = valid < k ? valid : k;
    return BCU_OK;
}

int bcu_simulate_trace(bcu_engine* e, int32_t nBalls, const float* pos, const int32_t* types, float vx, float vy,
                       int32_t* outLayers, float* outTimes, float* outRec) {
    if (!e || !pos || !types || !outLayers || !outTimes || !outRec) return BCU_ERR_INVALID;
    if (nBalls < 1 || nBalls > MAX_BALLS) return fail(e, BCU_ERR_INVALID, "bcu_simulate_trace: bad ball count");
    CK(cudaSetDevice(e->device));
    // the strike's one-state table lives in its own buffers: the beam uploaded with bcu_set_states and the last launch
    // (bcu_rerun / bcu_fetch_results / bcu_level_select) stay as they are -- except the per-strike result arrays, see below
    const uint32_t expected = 0xffffffffu;
    int8_t t8[MAX_BALLS];
    for (int i = 0; i < nBalls; ++i) {
        if (types[i] >= e->rules.nTypes) return fail(e, BCU_ERR_INVALID, "bcu_simulate_trace: type code out of range");
        t8[i] = (int8_t)(types[i] < 0 ? -1 : types[i]);
    }
    CK(e->dTrPos.reserve((size_t)nBalls * 2)); CK(e->dTrType.reserve(nBalls)); CK(e->dTrExpected.reserve(1));
    CK(cudaMemcpyAsync(e->dTrPos.p, pos, (size_t)nBalls * 2 * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(e->dTrType.p, t8, nBalls, cudaMemcpyHostToDevice, e->stream));   // (pageable source: staged before return)
    CK(cudaMemcpyAsync(e->dTrExpected.p, &expected, sizeof(expected), cudaMemcpyHostToDevice, e->stream));
    int rc;
    const int savedBalls = e->nBalls;
    e->nBalls = nBalls;   // reserve_results / fill_common size the result arrays by it
    const int32_t s0 = 0;
    const uint64_t sc = 0, bc = 0;
    CK(e->dCandState.reserve(1)); CK(e->dVx.reserve(1)); CK(e->dVy.reserve(1)); CK(e->dSearchCost.reserve(1)); CK(e->dBaseCost.reserve(1));
    e->planCount = -1;
    rc = reserve_results(e, 1, true, false);
    if (rc) { e->nBalls = savedBalls; return rc; }
    e->haveLast = false;   // the per-strike result arrays now hold the traced strike
    if (e->dTraceTimes.reserve(BCU_MAX_LAYERS) != cudaSuccess ||
        e->dTraceRec.reserve((size_t)BCU_MAX_LAYERS * nBalls * BCU_TRACE_REC) != cudaSuccess) {
        e->nBalls = savedBalls;
        return fail(e, BCU_ERR_CUDA, "bcu_simulate_trace: out of device memory");
    }
    CK(cudaMemcpyAsync(e->dCandState.p, &s0, sizeof(s0), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(e->dVx.p, &vx, sizeof(float), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(e->dVy.p, &vy, sizeof(float), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(e->dSearchCost.p, &sc, sizeof(sc), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(e->dBaseCost.p, &bc, sizeof(bc), cudaMemcpyHostToDevice, e->stream));
    LaunchParams p;
    fill_common(e, p, 1, true, false);
    p.states.nStates = 1; p.states.nBalls = nBalls;
    p.states.pos = e->dTrPos.p; p.states.type = e->dTrType.p; p.states.expectedMask = e->dTrExpected.p;
    p.src.stateIdx = e->dCandState.p; p.src.vx = e->dVx.p; p.src.vy = e->dVy.p;
    p.src.searchCost = e->dSearchCost.p; p.src.baseCost = e->dBaseCost.p;
    p.out.traceTimes = e->dTraceTimes.p; p.out.traceRec = e->dTraceRec.p;
    rc = run(e, p, true);
    e->nBalls = savedBalls;
    if (rc) return rc;
    uint32_t flags = 0;
    CK(cudaMemcpyAsync(&flags, e->dFlags.p, sizeof(flags), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaMemcpyAsync(outTimes, e->dTraceTimes.p, BCU_MAX_LAYERS * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaMemcpyAsync(outRec, e->dTraceRec.p, (size_t)BCU_MAX_LAYERS * nBalls * BCU_TRACE_REC * sizeof(float),
                       cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    e->haveLast = false;
    *outLayers = (flags & BCU_F_COMPLETED) ? (int32_t)BCU_N_EVENTS(flags) + 1 : 0;
    return BCU_OK;
}

int bcu_trace_strikes(bcu_engine* e, int32_t n, const int32_t* stateIdx, const float* vx, const float* vy,
                      int32_t* outLayers, float* outTimes, float* outRec) {
    if (!e) return BCU_ERR_INVALID;
    if (e->nStates == 0) return fail(e, BCU_ERR_STATE, "bcu_trace_strikes: call bcu_set_states first");
    if (n < 0 || n > 65536 || (n > 0 && (!stateIdx || !vx || !vy || !outLayers || !outTimes || !outRec)))
        return fail(e, BCU_ERR_INVALID, "bcu_trace_strikes: bad arguments");
    if (n == 0) return BCU_OK;
    for (int32_t i = 0; i < n; ++i)
        if (stateIdx[i] < 0 || stateIdx[i] >= e->nStates) return fail(e, BCU_ERR_INVALID, "bcu_trace_strikes: state index out of range");
    CK(cudaSetDevice(e->device));
    const int NB = e->nBalls;
    const size_t recPer = (size_t)BCU_MAX_LAYERS * NB * BCU_TRACE_REC;
    CK(e->dCandState.reserve(n)); CK(e->dVx.reserve(n)); CK(e->dVy.reserve(n)); CK(e->dSearchCost.reserve(n)); CK(e->dBaseCost.reserve(n));
    e->planCount = -1;
    int rc = reserve_results(e, n, true, false);
    if (rc) return rc;
    CK(e->dTraceTimes.reserve((size_t)n * BCU_MAX_LAYERS)); CK(e->dTraceRec.reserve((size_t)n * recPer));
    CK(cudaMemcpyAsync(e->dCandState.p, stateIdx, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(e->dVx.p, vx, (size_t)n * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(e->dVy.p, vy, (size_t)n * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemsetAsync(e->dSearchCost.p, 0, (size_t)n * sizeof(uint64_t), e->stream));
    CK(cudaMemsetAsync(e->dBaseCost.p, 0, (size_t)n * sizeof(uint64_t), e->stream));
    LaunchParams p;
    fill_common(e, p, n, true, false);
    p.src.stateIdx = e->dCandState.p; p.src.vx = e->dVx.p; p.src.vy = e->dVy.p;
    p.src.searchCost = e->dSearchCost.p; p.src.baseCost = e->dBaseCost.p;
    p.out.traceTimes = e->dTraceTimes.p; p.out.traceRec = e->dTraceRec.p;
    rc = run(e, p, true);
    if (rc) return rc;
    std::vector<uint32_t> flags(n);
    CK(cudaMemcpyAsync(flags.data(), e->dFlags.p, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaMemcpyAsync(outTimes, e->dTraceTimes.p, (size_t)n * BCU_MAX_LAYERS * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaMemcpyAsync(outRec, e->dTraceRec.p, (size_t)n * recPer * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    e->haveLast = false;   // the per-strike result arrays now hold the traced strikes (the beam records are untouched)
    for (int32_t i = 0; i < n; ++i) outLayers[i] = (flags[i] & BCU_F_COMPLETED) ? (int32_t)BCU_N_EVENTS(flags[i]) + 1 : 0;
    return BCU_OK;
}

int bcu_gather(bcu_engine* e, int32_t n, const uint64_t* ids, float* outPos, uint32_t* outPotted, float* outTime) {
    if (!e) return BCU_ERR_INVALID;
    if (!e->haveLast || !e->last.out.finalPos) return fail(e, BCU_ERR_STATE, "bcu_gather: no launch with final positions");
    if (n < 0 || (n > 0 && !ids)) return fail(e, BCU_ERR_INVALID, "bcu_gather: bad arguments");
    if (n == 0) return BCU_OK;
    CK(cudaSetDevice(e->device));
    const int NB = e->nBalls;
    for (int32_t i = 0; i < n; ++i) {
        const int64_t local = (int64_t)ids[i] - e->lastFirst;
        if (local < 0 || local >= e->last.nCand) return fail(e, BCU_ERR_INVALID, "bcu_gather: id not in the last launch");
    }
    // one kernel gathers the n records into a record block (the beam's own pack kernel), three copies bring it back
    const size_t bb = record_block_bytes(n, NB);
    CK(e->dGatherIds.reserve(n)); CK(e->dGatherPos.reserve((bb + 3) / 4));
    CK(cudaMemcpyAsync(e->dGatherIds.p, ids, (size_t)n * sizeof(uint64_t), cudaMemcpyHostToDevice, e->stream));
    CK(launch_pack_records(e->dGatherIds.p, e->dGatherIds.p, n, e->lastFirst, e->dPotted.p, e->dFinalTime.p, e->dFinalPos.p, NB,
                           e->dGatherPos.p, e->stream));
    e->launches += 1;
    const char* blk = (const char*)e->dGatherPos.p;
    if (outPotted) CK(cudaMemcpyAsync(outPotted, blk + (size_t)n * 16, (size_t)n * 4, cudaMemcpyDeviceToHost, e->stream));
    if (outTime) CK(cudaMemcpyAsync(outTime, blk + (size_t)n * 20, (size_t)n * 4, cudaMemcpyDeviceToHost, e->stream));
    if (outPos) CK(cudaMemcpyAsync(outPos, blk + (size_t)n * 24, (size_t)n * NB * 8, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return BCU_OK;
}

// ---- one break level, device resident: evaluate -> top-k -> record block -> (allgather) -> merge -----------------
static int topk_device(bcu_engine* e, int32_t k) {
    const int64_t n = e->last.nCand;
    const size_t sb = topk_scratch_bytes(n, k);
    CK(e->dTopkScratch.reserve(sb)); CK(e->dTopCost.reserve(k)); CK(e->dTopId.reserve(k)); CK(e->dTopCount.reserve(1));
    CK(launch_topk(e->dCost.p, e->dFlags.p, n, e->lastFirst, k, e->dTopCost.p, e->dTopId.p,
                   e->dTopCount.p, e->dTopkScratch.p, e->dTopkScratch.cap, e->stream));
    e->launches += 4;
    return BCU_OK;
}

int bcu_level_select(bcu_engine* e, int32_t k) {
    if (!e) return BCU_ERR_INVALID;
    if (!e->haveLast || !e->last.out.finalPos) return fail(e, BCU_ERR_STATE, "bcu_level_select: no grid launch to select from");
    if (k < 1 || k > 1024) return fail(e, BCU_ERR_INVALID, "bcu_level_select: k must be in 1..1024");
    CK(cudaSetDevice(e->device));
    // BCU_SELECT_TIMING=1 (diagnosis only): CUDA events between the phases, times on stderr
    const bool timing = e->selectTiming;
    cudaEvent_t tv[5] = {};
    int nt = 0;
    auto mark = [&]() { if (timing && nt < 5) { cudaEventCreate(&tv[nt]); cudaEventRecord(tv[nt], e->stream); ++nt; } };
    mark();
    int rc = topk_device(e, k);
    if (rc) return rc;
    const size_t bb = record_block_bytes(k, e->nBalls);
    CK(e->dBlockLocal.reserve(bb)); CK(e->dBlockAll.reserve(bb * e->nRanks)); CK(e->dBlockMerged.reserve(bb));
    CK(launch_pack_records(e->dTopCost.p, e->dTopId.p, k, e->lastFirst, e->dPotted.p, e->dFinalTime.p, e->dFinalPos.p,
                           e->nBalls, e->dBlockLocal.p, e->stream));
    e->launches += 1;
    mark();
    const void* gathered = e->dBlockLocal.p;
    if (e->nRanks > 1) {
        // the ONE collective of the path: per-GPU top-k records, ~ k * (24 + 8 N) bytes per rank
        int nr = g_nccl.AllGather(e->dBlockLocal.p, e->dBlockAll.p, bb, /*ncclInt8*/ 0, e->ncclComm, e->stream);
        if (nr != 0) {
            e->err = std::string("ncclAllGather: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(nr) : "error");
            return BCU_ERR_CUDA;
        }
        gathered = e->dBlockAll.p;
    }
    mark();
    if (e->nRanks > 1) {
        CK(launch_merge_records(gathered, e->nRanks, k, e->nBalls, e->dBlockMerged.p, e->stream));
        e->launches += 2;
        e->beamBlock = e->dBlockMerged.p;
    } else {
        e->beamBlock = e->dBlockLocal.p;   // one rank: the packed local block is the beam
    }
    mark();
    if (timing) {
        cudaStreamSynchronize(e->stream);
        float a = 0, b = 0, c = 0;
        cudaEventElapsedTime(&a, tv[0], tv[1]); cudaEventElapsedTime(&b, tv[1], tv[2]); cudaEventElapsedTime(&c, tv[2], tv[3]);
        fprintf(stderr, "[bcu select] rank %d: top-k + pack %.1f us, allgather %.1f us, merge %.1f us\n", e->rank, a * 1e3f, b * 1e3f, c * 1e3f);
        for (int i = 0; i < nt; ++i) cudaEventDestroy(tv[i]);
    }
    e->beamK = k;
    e->haveBeam = true;
    e->beamStaged = false;
    return BCU_OK;  // asynchronous: bcu_fetch_beam / bcu_sync wait for it
}

int bcu_level_rerun(bcu_engine* e, int32_t k) {
    if (!e) return BCU_ERR_INVALID;
    if (!e->haveLast) return fail(e, BCU_ERR_STATE, "bcu_level_rerun: nothing to re-run");
    CK(cudaSetDevice(e->device));
    int rc = run(e, e->last, false);
    if (rc) return rc;
    return bcu_level_select(e, k);
}

int bcu_fetch_beam(bcu_engine* e, int32_t* outCount, uint64_t* outCost, uint64_t* outId, uint32_t* outPotted,
                   float* outTime, float* outPos) {
    if (!e || !outCount) return BCU_ERR_INVALID;
    if (!e->haveBeam) return fail(e, BCU_ERR_STATE, "bcu_fetch_beam: no bcu_level_select yet");
    CK(cudaSetDevice(e->device));
    const int k = e->beamK, NB = e->nBalls;
    // one copy of the whole record block into page-locked staging (the caller's arrays are usually pageable: five small
    // copies into them would each be staged and synchronised by the driver), then plain host copies
    const size_t bb = record_block_bytes(k, NB), need = outPos ? bb : (size_t)k * 24;
    if (!e->beamStaged) {
        int rc = reserve_beam_staging(e, bb);
        if (rc) return rc;
        CK(cudaMemcpyAsync(e->hBeam, e->beamBlock, need, cudaMemcpyDeviceToHost, e->stream));
    }
    CK(cudaStreamSynchronize(e->stream));
    const char* base = (const char*)e->hBeam;
    const uint64_t* ids = (const uint64_t*)(base + (size_t)k * 8);
    int cnt = 0;
    while (cnt < k && ids[cnt] != ~0ull) cnt++;
    if (outCost) std::memcpy(outCost, base, (size_t)k * 8);
    if (outId) std::memcpy(outId, ids, (size_t)k * 8);
    if (outPotted) std::memcpy(outPotted, base + (size_t)k * 16, (size_t)k * 4);
    if (outTime) std::memcpy(outTime, base + (size_t)k * 20, (size_t)k * 4);
    if (outPos) std::memcpy(outPos, base + (size_t)k * 24, (size_t)k * NB * 8);
    *outCount = cnt;
    return BCU_OK;
}

int bcu_set_id_base(bcu_engine* e, int64_t idBase) {
    if (!e || idBase < 0) return BCU_ERR_INVALID;
    e->idBase = idBase;
    return BCU_OK;
}

int bcu_sync(bcu_engine* e) {
    if (!e) return BCU_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    CK(cudaStreamSynchronize(e->stream));
    return BCU_OK;
}

// ---- multi-GPU: one NCCL communicator per engine (one process per GPU) ------------------------------------------
int bcu_comm_unique_id(const char* ncclLibPath, void* out128) {
    std::string err;
    if (!out128) return BCU_ERR_INVALID;
    if (!load_nccl(ncclLibPath, &err)) return BCU_ERR_CUDA;
    return g_nccl.GetUniqueId(out128) == 0 ? BCU_OK : BCU_ERR_CUDA;
}

int bcu_comm_init(bcu_engine* e, int32_t nRanks, int32_t rank, const void* uniqueId128, const char* ncclLibPath) {
    if (!e) return BCU_ERR_INVALID;
    if (nRanks < 1 || rank < 0 || rank >= nRanks) return fail(e, BCU_ERR_INVALID, "bcu_comm_init: bad rank / size");
    if (nRanks == 1) { e->nRanks = 1; e->rank = 0; return BCU_OK; }
    if (!uniqueId128) return fail(e, BCU_ERR_INVALID, "bcu_comm_init: unique id required");
    if (!load_nccl(ncclLibPath, &e->err)) return BCU_ERR_CUDA;
    CK(cudaSetDevice(e->device));
    NcclApi::Uid uid;
    std::memcpy(uid.b, uniqueId128, 128);
    int nr = g_nccl.CommInitRank(&e->ncclComm, nRanks, uid, rank);
    if (nr != 0) {
        e->err = std::string("ncclCommInitRank: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(nr) : "error");
        return BCU_ERR_CUDA;
    }
    e->nRanks = nRanks;
    e->rank = rank;
    return BCU_OK;
}

// ---- measurement hooks --------------------------------------------------------------------------------------
int bcu_timer_begin(bcu_engine* e) {
    if (!e) return BCU_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    CK(cudaEventRecord(e->evT0, e->stream));
    return BCU_OK;
}

int bcu_timer_end(bcu_engine* e, float* outMs) {
    if (!e || !outMs) return BCU_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    CK(cudaEventRecord(e->evT1, e->stream));
    CK(cudaEventSynchronize(e->evT1));
    CK(cudaEventElapsedTime(outMs, e->evT0, e->evT1));
    return BCU_OK;
}

int bcu_flush_l2(bcu_engine* e) {
    if (!e) return BCU_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    const size_t bytes = (size_t)256 << 20;  // > 126 MB L2
    CK(e->dFlush.reserve(bytes));
    CK(launch_flush(e->dFlush.p, bytes, e->stream));
    return BCU_OK;
}

int bcu_issue_rate(bcu_engine* e, int32_t kind, double* outGinstPerSec) {
    if (!e || !outGinstPerSec || kind < 0 || kind > 3) return BCU_ERR_INVALID;
    CK(cudaSetDevice(e->device));
    CK(e->dSink.reserve(1));
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, e->device));
    const int ctas = sms * 8, iters = 1 << 15;
    launch_issue_rate(kind, ctas, 1 << 10, e->dSink.p, e->stream);  // warm-up
    float best = 1e30f;
    int per = 0;
    for (int rep = 0; rep < 3; ++rep) {
        CK(cudaEventRecord(e->evT0, e->stream));
        per = launch_issue_rate(kind, ctas, iters, e->dSink.p, e->stream);
        CK(cudaEventRecord(e->evT1, e->stream));
        CK(cudaEventSynchronize(e->evT1));
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, e->evT0, e->evT1));
        if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    const double inst = (double)ctas * 256.0 * (double)iters * per;
    *outGinstPerSec = inst / (best * 1e-3) / 1e9;
    return BCU_OK;
}

int bcu_last_kernel_ms(bcu_engine* e, float* outMs) {
    if (!e || !outMs) return BCU_ERR_INVALID;
    CK(cudaEventSynchronize(e->evStop));
    CK(cudaEventElapsedTime(outMs, e->evStart, e->evStop));
    return BCU_OK;
}

int bcu_enable_counters(bcu_engine* e, int enable) {
    if (!e) return BCU_ERR_INVALID;
    e->countersOn = enable != 0;
    if (e->haveLast) e->last.out.counters = e->countersOn ? e->dCounters.p : nullptr;
    return BCU_OK;
}

int bcu_set_prune(bcu_engine* e, int mode) {
    if (!e || mode < 0 || mode > 2) return BCU_ERR_INVALID;
    e->pruneMode = mode;
    return BCU_OK;
}

int bcu_get_prune(const bcu_engine* e) { return e ? e->pruneMode : -1; }

// Test hook: the multi-GPU merge kernel on host-supplied record blocks (what ncclAllGather delivers on a multi-GPU run),
// so that merge_records_kernel is covered on a single GPU.
int bcu_test_merge_records(bcu_engine* e, int32_t nBlocks, int32_t k, int32_t nBalls, const void* blocks, void* outBlock) {
    if (!e) return BCU_ERR_INVALID;
    if (nBlocks < 1 || nBlocks > 64 || k < 1 || k > 1024 || nBalls < 1 || nBalls > MAX_BALLS || !blocks || !outBlock)
        return fail(e, BCU_ERR_INVALID, "bcu_test_merge_records: bad arguments");
    CK(cudaSetDevice(e->device));
    const size_t bb = record_block_bytes(k, nBalls);
    DevBuf<unsigned char> dIn, dOut;
    cudaError_t ce;
    if ((ce = dIn.reserve(bb * nBlocks)) != cudaSuccess || (ce = dOut.reserve(bb)) != cudaSuccess) {
        dIn.release(); dOut.release();
        e->err = std::string("bcu_test_merge_records: ") + cudaGetErrorString(ce);
        return BCU_ERR_CUDA;
    }
    ce = cudaMemcpyAsync(dIn.p, blocks, bb * nBlocks, cudaMemcpyHostToDevice, e->stream);
    if (ce == cudaSuccess) ce = launch_merge_records(dIn.p, nBlocks, k, nBalls, dOut.p, e->stream);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(outBlock, dOut.p, bb, cudaMemcpyDeviceToHost, e->stream);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(e->stream);
    e->launches += 2;
    dIn.release(); dOut.release();
    if (ce != cudaSuccess) { e->err = std::string("bcu_test_merge_records: ") + cudaGetErrorString(ce); return BCU_ERR_CUDA; }
    return BCU_OK;
}

int bcu_get_prune_counters(bcu_engine* e, uint64_t* out2) {
    if (!e || !out2) return BCU_ERR_INVALID;
    if (!e->dCounters.p) return fail(e, BCU_ERR_STATE, "bcu_get_prune_counters: no launch yet");
    unsigned long long h[8];
    CK(cudaMemcpyAsync(h, e->dCounters.p, sizeof(h), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    out2[0] = h[5]; out2[1] = h[6];
    return BCU_OK;
}

int bcu_get_counters(bcu_engine* e, uint64_t* out5) {
    if (!e || !out5) return BCU_ERR_INVALID;
    if (!e->dCounters.p) return fail(e, BCU_ERR_STATE, "bcu_get_counters: no launch yet");
    unsigned long long h[8];
    CK(cudaMemcpyAsync(h, e->dCounters.p, sizeof(h), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    for (int i = 0; i < 5; ++i) out5[i] = h[i];
    return BCU_OK;
}

int64_t bcu_launch_count(const bcu_engine* e) { return e ? e->launches : 0; }

// ---- the reference's geometric planner on the device (bcu_planner.cuh, bcu_plan.cu) --------------------------------------
int bcu_set_planner(bcu_engine* e, const bcu_planner_desc* d) {
    if (!e) return BCU_ERR_INVALID;
    if (!d || d->n_rails != e->table.nRails || !d->rail_start || (d->n_set_order > 0 && !d->set_order) || d->n_search_pockets < 1 ||
        !d->search_pockets)
        return fail(e, BCU_ERR_INVALID, "bcu_set_planner: bad description (the rails must be those of bcu_create)");
    CK(cudaSetDevice(e->device));
    PlanTable tb;
    if (!plan_table_build(&tb, d->n_rails, e->railDerived.data(), d->rail_start, d->rail_order, d->n_set_order, d->set_order,
                          d->n_search_pockets, d->search_pockets, e->table.diameter / 2, d->minimal_pocket_velocity,
                          d->diagonal_length_squared, d->velocity_multiply, e->rules.cueType))
        return fail(e, BCU_ERR_INVALID, "bcu_set_planner: inconsistent rail orders / counts");
    // (the ball radius: bcu_create keeps diameter = radius * 2, an exact doubling, so diameter / 2 is the radius bit for bit)
    e->planTable = tb;
    CK(e->dPlanTable.reserve(1));
    CK(cudaMemcpyAsync(e->dPlanTable.p, &e->planTable, sizeof(PlanTable), cudaMemcpyHostToDevice, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    e->havePlanner = true;
    e->planCount = -1;
    return BCU_OK;
}

int bcu_plan(bcu_engine* e, int64_t* outCount) {
    if (!e || !outCount) return BCU_ERR_INVALID;
    if (!e->havePlanner) return fail(e, BCU_ERR_STATE, "bcu_plan: call bcu_set_planner first");
    if (e->nStates == 0) return fail(e, BCU_ERR_STATE, "bcu_plan: call bcu_set_states first");
    CK(cudaSetDevice(e->device));
    StateTable st{};
    st.nStates = e->nStates; st.nBalls = e->nBalls;
    st.pos = e->dStatePos.p; st.type = e->dStateType.p; st.expectedMask = e->dExpected.p;
    const uint64_t* base = e->hasStateBase ? e->dStateBase.p : nullptr;
    const int64_t items = plan_work_items(e->planTable, e->nStates, e->nBalls), blocks = plan_blocks(e->planTable, e->nStates, e->nBalls);
    CK(e->dPlanItemCount.reserve((size_t)(items > 0 ? items : 1))); CK(e->dPlanBlockSum.reserve((size_t)(blocks > 0 ? blocks : 1)));
    CK(e->dPlanBlockOffset.reserve((size_t)(blocks > 0 ? blocks : 1))); CK(e->dPlanTotal.reserve(1));
    CK(launch_plan_count(e->dPlanTable.p, e->planTable, st, base, e->dPlanItemCount.p, e->dPlanBlockSum.p, e->dPlanBlockOffset.p,
                         e->dPlanTotal.p, e->stream));
    e->launches += 2;
    unsigned long long total = 0;
    CK(cudaMemcpyAsync(&total, e->dPlanTotal.p, sizeof(total), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    const size_t n = (size_t)(total > 0 ? total : 1);
    CK(e->dCandState.reserve(n)); CK(e->dVx.reserve(n)); CK(e->dVy.reserve(n)); CK(e->dSearchCost.reserve(n)); CK(e->dBaseCost.reserve(n));
    if (total > 0) {
        CK(launch_plan_write(e->dPlanTable.p, e->planTable, st, base, e->dPlanItemCount.p, e->dPlanBlockOffset.p, e->dCandState.p, e->dVx.p,
                             e->dVy.p, e->dSearchCost.p, e->dBaseCost.p, e->stream));
        e->launches += 1;
    }
    e->planCount = (int64_t)total;
    e->planStates = e->nStates;
    e->haveLast = false;
    *outCount = (int64_t)total;
    return BCU_OK;
}

int bcu_fetch_plan(bcu_engine* e, int64_t offset, int64_t n, int32_t* stateIdx, float* vx, float* vy, uint64_t* searchCost, uint64_t* cost) {
    if (!e) return BCU_ERR_INVALID;
    if (e->planCount < 0) return fail(e, BCU_ERR_STATE, "bcu_fetch_plan: no plan (bcu_plan; bcu_set_states / bcu_evaluate discard it)");
    if (offset < 0 || n < 0 || offset + n > e->planCount) return fail(e, BCU_ERR_INVALID, "bcu_fetch_plan: range");
    if (n == 0) return BCU_OK;
    CK(cudaSetDevice(e->device));
    if (stateIdx) CK(cudaMemcpyAsync(stateIdx, e->dCandState.p + offset, n * sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
    if (vx) CK(cudaMemcpyAsync(vx, e->dVx.p + offset, n * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    if (vy) CK(cudaMemcpyAsync(vy, e->dVy.p + offset, n * sizeof(float), cudaMemcpyDeviceToHost, e->stream));
    if (searchCost) CK(cudaMemcpyAsync(searchCost, e->dSearchCost.p + offset, n * sizeof(uint64_t), cudaMemcpyDeviceToHost, e->stream));
    if (cost) CK(cudaMemcpyAsync(cost, e->dBaseCost.p + offset, n * sizeof(uint64_t), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return BCU_OK;
}

int bcu_evaluate_plan(bcu_engine* e, const bcu_results* out) { return bcu_evaluate_plan_range(e, 0, -1, out); }

int bcu_evaluate_plan_range(bcu_engine* e, int64_t first, int64_t count, const bcu_results* out) {
    if (!e) return BCU_ERR_INVALID;
    if (e->planCount < 0) return fail(e, BCU_ERR_STATE, "bcu_evaluate_plan: no plan (bcu_plan; bcu_set_states / bcu_evaluate discard it)");
    if (count < 0) count = e->planCount - first;
    if (first < 0 || count < 0 || first + count > e->planCount) return fail(e, BCU_ERR_INVALID, "bcu_evaluate_plan_range: range");
    const int64_t n = count;
    CK(cudaSetDevice(e->device));
    const bool wantLog = out && out->event_log;
    int rc = reserve_results(e, n, true, wantLog);
    if (rc) return rc;
    LaunchParams p;
    fill_common(e, p, n, true, wantLog);
    p.src.stateIdx = e->dCandState.p + first; p.src.vx = e->dVx.p + first; p.src.vy = e->dVy.p + first;
    p.src.searchCost = e->dSearchCost.p + first; p.src.baseCost = e->dBaseCost.p + first;
    e->last = p; e->haveLast = true; e->lastFirst = first;
    bool copied = false;
    rc = run(e, p, false, out, &copied);
    if (rc) { e->haveLast = false; return rc; }
    if (!copied) { rc = copy_back(e, 0, n, out); if (rc) return rc; }
    return BCU_OK;
}

int bcu_engine_rank_info(const bcu_engine* e, int32_t* outRank, int32_t* outRanks) {
    if (!e) return BCU_ERR_INVALID;
    if (outRank) *outRank = e->rank;
    if (outRanks) *outRanks = e->nRanks;
    return BCU_OK;
}

}  // extern "C"
