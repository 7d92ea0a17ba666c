/*
 * billiard_cuda.h -- thin C ABI of the B200 strike-evaluation engine (libbilliard_cuda.so).
 *
 * This is the drop-in boundary below the reference's C++ entry points
 *     billiard::search::simulate(State, vec2, Configuration)      lib/billiard_search/include/billiard_search/billiard_search.hpp:17-18
 *     billiard::search::search(State, Search, solutions, Config)  lib/billiard_search/include/billiard_search/billiard_search.hpp:9-12
 * whose hot loop (billiard_search.cpp:1358-1471 simulate, :1550-1573 nextEvent, :1957-1992 createLayer,
 * :1473-1532 simulationCost, rules billiard_snooker.cpp:787-790, 1128-1153) it replaces.
 * The C++ mirror of those entry points that calls this ABI lives in billiard_ai_b200/host/.
 *
 * Conventions: POD only, caller-owned buffers, one CUDA stream per engine, every call returns a bcu_status (0 = OK).
 * Thread-compatible: all state lives in the engine (launch configuration, occupancy figures, staging buffers), so
 * different engines may be used from different threads and on different devices at the same time; one engine must not
 * be used from two threads at once.  The only process-wide state is the function table of the dlopen'ed NCCL library
 * (filled once under a lock).  There is NO CPU fallback: if no CUDA device is usable, bcu_create fails with
 * BCU_ERR_NO_DEVICE.
 */
#ifndef BILLIARD_CUDA_H
#define BILLIARD_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BCU_MAX_BALLS 32
#define BCU_MAX_RAILS 32
#define BCU_MAX_POCKETS 8
#define BCU_MAX_TYPES 16
#define BCU_MAX_LAYERS 32 /* MAX_EVENTS (30) + the input layer, rounded up; billiard_search.cpp:20 */
#define BCU_TRACE_REC 10

typedef enum {
    BCU_OK = 0,
    BCU_ERR_NO_DEVICE = 1,   /* no CUDA device / driver: there is no CPU fallback */
    BCU_ERR_INVALID = 2,     /* bad argument */
    BCU_ERR_CUDA = 3,        /* a CUDA runtime call failed; see bcu_last_error */
    BCU_ERR_STATE = 4        /* call sequence error (e.g. evaluate before set_states) */
} bcu_status;

/* result flags, per candidate (low byte of bcu_results.flags; bits 8..15 = number of events) */
#define BCU_F_COMPLETED 1u /* all balls at rest / potted within MAX_EVENTS                         */
#define BCU_F_VALID 2u     /* rules: valid end state (billiard_snooker.cpp:1128-1153)              */
#define BCU_F_OVERFLOW 4u  /* > MAX_EVENTS events: the reference discards it (billiard_search.cpp:1376-1379) */
#define BCU_F_NOEVENT 8u   /* no next event found (billiard_search.cpp:1464-1467)                  */
#define BCU_N_EVENTS(flags) (((flags) >> 8) & 0xffu)

/* node kinds in traces == billiard::search::node::NodeType (type.hpp:167-180) */
/* event types in logs  == billiard::search::event::EventType (type.hpp:87-98)  */

typedef struct bcu_engine bcu_engine;

/* Table geometry: what search::Configuration::_table/_ball hold (type.hpp:301-322).
 * rails: n_rails x {start.x, start.y, end.x, end.y, shiftDirection.x, shiftDirection.y} -- the Rail ctor
 * arguments (type.hpp:69-71); normal and shifted end points are derived as type.cpp:56-72 does.
 * pockets: n_pockets x {position.x, position.y, radius} (type.hpp:46-55). */
typedef struct {
    int32_t n_rails;
    int32_t n_pockets;
    float ball_radius;
    const float* rails;
    const float* pockets;
    /* Optional (may be NULL): n_rails x {normal.xy, shiftedStart.xy, shiftedEnd.xy} = Rail::_normal, _shiftedStart,
     * _shiftedEnd as the reference's Rail objects already hold them (type.hpp:75-77).  When given it is used
     * instead of deriving the geometry from `rails` (which may then be NULL). */
    const float* rails_derived;
} bcu_table_desc;

/* Rules as data (the reference's std::function rules, type.hpp:324-332, cannot run on a device):
 * score[t] = _scoreForPottedBall(type t); cue_type = type code of "WHITE".
 * Valid end state = billiard::snooker::validEndState with the per-state expected-type mask. */
typedef struct {
    int32_t n_types;
    int32_t cue_type;
    const double* score;
} bcu_rules_desc;

/* Host result buffers; any pointer may be NULL (that output is then not copied back). */
typedef struct {
    uint64_t* cost;      /* [n] parent cost + simulationCost; UINT64_MAX when not COMPLETED */
    uint32_t* flags;     /* [n] */
    uint32_t* potted;    /* [n] bit b: ball b is potted in the final layer */
    float* final_time;   /* [n] time of the final layer */
    float* final_pos;    /* [n][n_balls][2] */
    int32_t* event_log;  /* [n][32][4] {event type, ball a, ball b | rail | pocket, layer time bits}, -1 padded */
} bcu_results;

/* ---- lifecycle ---------------------------------------------------------------------------------- */
int bcu_create(const bcu_table_desc* table, const bcu_rules_desc* rules, int device, bcu_engine** out);
void bcu_destroy(bcu_engine* e);
const char* bcu_last_error(const bcu_engine* e);
const char* bcu_version(void);
/* derived rail geometry as uploaded: n_rails x {normal.xy, shiftedStart.xy, shiftedEnd.xy} */
int bcu_get_rail_derived(const bcu_engine* e, float* out6);

/* Page-locked host memory for candidate / result buffers (cudaHostAlloc): copies from and to such buffers run at full
 * PCIe speed and asynchronously; any host pointer is accepted by the calls below, pinned ones are just faster. */
int bcu_host_alloc(size_t bytes, void** out);
int bcu_host_free(void* p);

/* ---- table states (the beam) ------------------------------------------------------------------------
 * pos [n_states][n_balls][2]; types [n_states][n_balls] (type code, -1 = no ball);
 * expected_mask [n_states] (bit t: potting type t is legal = the Search::_types of that state);
 * base_cost [n_states] accumulated cost of the state's parent chain (may be NULL = 0). */
int bcu_set_states(bcu_engine* e, int32_t n_states, int32_t n_balls, const float* pos, const int32_t* types,
                   const uint32_t* expected_mask, const uint64_t* base_cost);

/* ---- evaluation of explicit candidates (reference: one SIMULATION node each, billiard_search.cpp:797-826) --
 * Host buffers; host->device and device->host copies are part of the call. */
int bcu_evaluate(bcu_engine* e, int64_t n, const int32_t* state_idx, const float* vx, const float* vy,
                 const uint64_t* search_cost, const uint64_t* base_cost, const bcu_results* out);

/* ---- evaluation of a direction x speed grid per state (synthetic configs 2-5) -------------
 * candidate id c = (state * n_dirs + dir) * n_speeds + speed; velocity = speeds[speed] * dirs[dir].
 * dirs [n_dirs][2] unit vectors and speeds [n_speeds] are host arrays (uploaded once per call).
 * Candidates [first, first+count) are evaluated (count < 0: all) -- the multi-GPU shard of this rank.
 * Results stay on the device; fetch them with bcu_fetch_results / bcu_topk. */
int bcu_evaluate_grid(bcu_engine* e, int32_t n_dirs, int32_t n_speeds, const float* dirs, const float* speeds,
                      uint64_t search_cost, int64_t first, int64_t count);
/* Re-run the last bcu_evaluate_grid launch with everything already resident (timing loops). */
int bcu_rerun(bcu_engine* e);
/* Copy results of the last launch (device -> host). */
int bcu_fetch_results(bcu_engine* e, int64_t offset, int64_t n, const bcu_results* out);

/* ---- block-level top-k on (cost, candidate id) among VALID candidates of the last launch ------------
 * out_cost/out_id [k] ascending; *out_count = number of valid entries written (<= k). */
int bcu_topk(bcu_engine* e, int32_t k, uint64_t* out_cost, uint64_t* out_id, int32_t* out_count);
/* ---- the same selection without any allocation ------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Engine-free: the caller passes DEVICE
 * pointers on the current device -- d_cost [n] u64 and d_flags [n] u32 (BCU_F_VALID marks a candidate that takes part)
 * in, d_out_cost [k] / d_out_id [k] ascending by (cost, id) and padded with ~0, *d_out_count = number of valid candidates
 * (may exceed k) out -- and the stream (a cudaStream_t; NULL = the default stream).  ids are first_candidate + index.
 * Kernels are queued on the stream; the call does not synchronise and calls no cudaMalloc / cudaFree or any allocating
 * routine.  The scratch of the selection (2 687 232 bytes) is a file-scope __device__ array that every GPU which loads the
 * library holds until exit, which is why sizes are limited: n <= BCU_TOPK_STATIC_MAX_CANDIDATES, k <= BCU_TOPK_STATIC_MAX_K;
 * anything larger is refused with BCU_ERR_INVALID and a message in err (optional, err_cap bytes).  All calls on a device
 * share that array: calls ordered on one stream are fine, calls that overlap on the GPU race. */
#define BCU_TOPK_STATIC_MAX_CANDIDATES ((int64_t)1 << 22)
#define BCU_TOPK_STATIC_MAX_K 1024
int bcu_topk_static(const uint64_t* d_cost, const uint32_t* d_flags, int64_t n, int64_t first_candidate, int32_t k,
                    uint64_t* d_out_cost, uint64_t* d_out_id, int32_t* d_out_count, void* stream, char* err, int32_t err_cap);
/* Final positions / potted mask of selected candidates of the last launch (ids are global ids). */
int bcu_gather(bcu_engine* e, int32_t n, const uint64_t* ids, float* out_pos, uint32_t* out_potted, float* out_time);

/* ---- single strike with the full layer trace (reference public simulate(), billiard_search.cpp:141-175) --
 * pos [n_balls][2], types [n_balls]; out_times [32]; out_rec [32][n_balls][10]:
 *   {node kind, after.pos.xy, after.vel.xy, after.isRolling, before.vel.xy, cause index, -1}
 * *out_layers = number of layers, 0 when the reference would return nullopt.
 * The strike's state is uploaded into buffers of its own: the states of the last bcu_set_states stay in place.  The
 * per-strike result arrays of the last launch are overwritten (bcu_fetch_results / bcu_topk / bcu_rerun need a new
 * launch afterwards); a beam selected with bcu_level_select stays valid. */
int bcu_simulate_trace(bcu_engine* e, int32_t n_balls, const float* pos, const int32_t* types, float vx, float vy,
                       int32_t* out_layers, float* out_times, float* out_rec);

/* The same for n strikes at once on the states of the last bcu_set_states (the reference returns one node::System per
 * break of every solution, billiard_search.hpp:9-12 / type.hpp:525-537: trace the strikes of the best solutions in one
 * call).  state_idx / vx / vy [n], n <= 65536; out_layers [n] (0: the reference would return nullopt),
 * out_times [n][32], out_rec [n][32][n_balls][10] as above.  Overwrites the per-strike results of the last launch
 * (bcu_fetch_results / bcu_topk need a new launch afterwards); the beam of bcu_level_select stays valid. */
int bcu_trace_strikes(bcu_engine* e, int32_t n, const int32_t* state_idx, const float* vx, const float* vy,
                      int32_t* out_layers, float* out_times, float* out_rec);

/* ---- one break level, device resident (billiard_search.cpp:224-244 expand + process.hpp:19-25 pop order) ----
 * After bcu_evaluate_grid: select the k best VALID candidates by (cost, id), pack their end states into a record
 * block, all-gather the blocks of all ranks (one ncclAllGather; skipped for a single rank) and merge them into
 * the global top-k -- identical on every rank.  Asynchronous on the engine's stream. */
int bcu_level_select(bcu_engine* e, int32_t k);
/* bcu_evaluate followed by bcu_level_select(k) as ONE call: the selection is queued right behind the simulation and the
 * per-strike result copies (out, may be NULL) run on a second stream next to it.  Returns when the results are in the
 * caller's buffers; the beam is fetched with bcu_fetch_beam as usual.  Same arguments and errors as bcu_evaluate. */
int bcu_evaluate_level(bcu_engine* e, int64_t n, const int32_t* state_idx, const float* vx, const float* vy,
                       const uint64_t* search_cost, const uint64_t* base_cost, const bcu_results* out, int32_t k);
/* Re-run the last grid launch and select again, everything resident (timing loops). */
int bcu_level_rerun(bcu_engine* e, int32_t k);
/* Merged beam, device -> host: *out_count valid entries (ascending); arrays sized for k (any may be NULL):
 * out_pos [k][n_balls][2] = rest positions after the strike (potted balls keep their last position; see out_potted). */
int bcu_fetch_beam(bcu_engine* e, int32_t* out_count, uint64_t* out_cost, uint64_t* out_id, uint32_t* out_potted,
                   float* out_time, float* out_pos);
int bcu_sync(bcu_engine* e);
/* Global id of candidate 0 of subsequent bcu_evaluate calls (a rank's shard of an explicit candidate list);
 * ids only matter as the tie-break of the top-k order. */
int bcu_set_id_base(bcu_engine* e, int64_t id_base);

/* ---- multi-GPU: one process per GPU, one NCCL communicator per engine -------------------------------------
 * NCCL is dlopen'ed (nccl_lib_path, else $BCU_NCCL_LIB, else libnccl.so.2); rank 0 creates the 128-byte id and
 * the caller distributes it (torch.distributed / MPI / a file). */
int bcu_comm_unique_id(const char* nccl_lib_path, void* out128);
int bcu_comm_init(bcu_engine* e, int32_t n_ranks, int32_t rank, const void* unique_id128, const char* nccl_lib_path);

/* ---- rules between breaks (host side of the multi-break search; no device work) ------------------------------
 * The reference calls Configuration::_rules._nextSearch / _modifyState once per finished simulation that goes on to a
 * further break (billiard_search.cpp:1407-1462); for snooker these are billiard::snooker::nextSearch
 * (lib/billiard_snooker/src/billiard_snooker.cpp:792-822) and stateAfterBreak (:1097-1126, with replace :1053-1095 and
 * replaceInArea :969-1046).  Ball types are integer codes, a ball is a slot of the state table (its id); `present[i]`
 * says whether slot i is on the table, types[i] < 0 marks a slot that is not part of the frame any more. */
typedef struct bcu_snooker_config {      /* billiard::snooker::SnookerSearchConfig, billiard_snooker.hpp:128-136 */
    int32_t red_type;
    int32_t color_types[6];              /* YELLOW, GREEN, BROWN, BLUE, PINK, BLACK (typesToReplace, :773-780) */
    float spots[6][2];                   /* their spots, same order */
    float head_rail_dir[2];
    float diameter_squared, radius;
    float minimum[2], maximum[2];        /* table rectangle */
} bcu_snooker_config;
/* nextSearch: previous_types = the types the finished break searched for, in the order of the reference's list;
 * out_types receives up to 6 type codes (-1 = the reference's empty type name, which matches no ball). */
int bcu_snooker_next_search(const bcu_snooker_config* cfg, int32_t n_balls, const int32_t* types, const uint8_t* present,
                            int32_t n_previous, const int32_t* previous_types, int32_t* out_types, int32_t* out_count);
/* stateAfterBreak: potted colours return to the table while a red is left (own spot, else the next higher free spot,
 * else the nearest free place within 5 radii, searched in radius/2 steps).  out_ok = 0 reproduces the reference's
 * empty State ("could not replace all balls": the candidate is dropped). */
int bcu_snooker_state_after_break(const bcu_snooker_config* cfg, int32_t n_balls, const float* pos, const int32_t* types,
                                  const uint8_t* present, float* out_pos, uint8_t* out_present, int32_t* out_ok);

/* ---- the multi-break search as one call (billiard::search::search, billiard_search.hpp:9-12) -----------------------
 * Level-synchronous restatement of the reference's best-first search over breaks (billiard_search.cpp:74-98, 224-244,
 * next-break seeding :1407-1462): per level every beam state x (n_dirs x n_speeds) grid strike is simulated, the `beam`
 * best valid strikes are selected on the device (merged across the ranks of bcu_comm_init: every rank calls this function
 * with the same arguments and evaluates its contiguous share of the candidate ids), and ONE host pass over the selected
 * records applies the rules between breaks: empty state => dropped, no ball of a searched type left => solution, else a
 * state of the next level.  Strikes of the last level (simCount == BREAKS) are solutions. */
#define BCU_MAX_SEARCH_LEVELS 8
/* The rules between breaks (Configuration::_rules._nextSearch / _modifyState, type.hpp:324-332) as data: */
#define BCU_BREAK_RULES_SINGLE 0   /* the Configuration's defaults: nextSearch = Search{} (nothing left to search: every valid
                                      strike is a solution after one break), modifyState = identity (type.hpp:325-328) */
#define BCU_BREAK_RULES_REPEAT 1   /* not in the reference (it ships snooker rules only): search the same types again, state as
                                      it is -- the pool layouts of BASELINE configs 1 and 3 */
#define BCU_BREAK_RULES_SNOOKER 2  /* billiard::snooker::nextSearch / stateAfterBreak (billiard_snooker.cpp:792-822, 1097-1126) */
typedef struct bcu_break_rules {
    int32_t kind;
    bcu_snooker_config snooker;    /* kind BCU_BREAK_RULES_SNOOKER */
} bcu_break_rules;
typedef struct bcu_search_params {
    int32_t levels;                /* BREAKS (billiard_search.cpp:15), 1..BCU_MAX_SEARCH_LEVELS */
    int32_t beam;                  /* strikes kept per level, 1..1024 */
    int32_t n_dirs, n_speeds;      /* the candidate grid of every state (as bcu_evaluate_grid); n_dirs = 0: the strikes the
                                      reference's planner finds for the state instead (bcu_set_planner first; planned on every
                                      rank, simulated in shares; dir_index / speed_index of the result are -1) */
    const float* dirs;             /* [n_dirs][2] */
    const float* speeds;           /* [n_speeds] */
    uint64_t search_cost;          /* SearchNode::_searchCost of every strike */
    int32_t n_search_types;        /* Search::_types of the first break */
    const int32_t* search_types;
    /* the manager's three ways to end a search (process.hpp:188-198), looked at between levels: */
    int32_t min_solutions;         /* > 0: stop once this many solutions are final (`solutions` of search()); 0: all levels */
    uint32_t max_duration_ms;      /* > 0: wall-clock budget (the reference: 10 000 ms); 0: none */
    const volatile int32_t* cancel;/* may be NULL; set non-zero by another thread to abandon the search (a newer request,
                                      process.hpp:82-83, 261-268) */
} bcu_search_params;
typedef struct bcu_level_stats {
    int32_t states, beam;          /* beam states evaluated, strikes selected */
    int64_t candidates;            /* strikes of the level, all ranks */
    float kernel_ms;               /* the level's simulation launch on this rank (CUDA events) */
    float rules_ms, wall_ms;       /* host pass over the beam; the whole level */
} bcu_level_stats;
#define BCU_SEARCH_EXHAUSTED 0     /* all levels run / no state left */
#define BCU_SEARCH_ENOUGH 1        /* min_solutions reached */
#define BCU_SEARCH_TIMEOUT 2
#define BCU_SEARCH_CANCELLED 3
typedef struct bcu_search_result {
    int32_t capacity;              /* in: room for this many solutions in the arrays below */
    int32_t count;                 /* out: solutions written, ascending by cost (equal costs in discovery order) */
    int32_t total;                 /* out: solutions found */
    int32_t levels_run;            /* out */
    int32_t stopped;               /* out: BCU_SEARCH_* */
    float wall_ms;                 /* out: the whole call */
    uint64_t* cost;                /* [capacity] accumulated cost of the solution */
    int32_t* breaks;               /* [capacity] strikes on its path */
    /* any of the following may be NULL; per break b < breaks (entries past `breaks` are -1 / 0): */
    int32_t* dir_index;            /* [capacity][levels] */
    int32_t* speed_index;          /* [capacity][levels] */
    float* strike;                 /* [capacity][levels][2] the strike's velocity */
    float* frame_pos;              /* [capacity][levels][n_balls][2] the table the strike was played on ... */
    int32_t* frame_types;          /* [capacity][levels][n_balls] ... and its types (-1 = no ball): bcu_trace_strikes input */
    float* end_pos;                /* [capacity][n_balls][2] table after the last break (rules applied when they ended the path) */
    uint8_t* end_present;          /* [capacity][n_balls] */
    bcu_level_stats* stats;        /* [levels], may be NULL */
} bcu_search_result;
int bcu_search_levels(bcu_engine* e, int32_t n_balls, const float* pos, const int32_t* types, const bcu_search_params* params,
                      const bcu_break_rules* rules, bcu_search_result* out);
/* rank / number of ranks of the engine's communicator (0 / 1 without bcu_comm_init) */
int bcu_engine_rank_info(const bcu_engine* e, int32_t* out_rank, int32_t* out_ranks);

/* ---- the reference's geometric planner on the device --------------------------------------------------
 * Which strikes are worth simulating on a table state: the reference's best-first expansion of SEARCH nodes -- pocket ->
 * ball to pot -> ball that hits it -> ... -> cue ball, every link direct or over up to three cushions (billiard_search.cpp:
 * 299-1347: expandSearchNode*, railCombinations / railTargets, searchCost, calculateMinimalVelocity, prepareForSimulation) --
 * as an exhaustive enumeration on the device, one thread per (state, pocket, first ball, first link).  It emits exactly the
 * SIMULATION nodes the reference's expand() creates when run to exhaustion (the same velocities bit for bit, the same
 * _searchCost and accumulated _cost), for every state of the last bcu_set_states at once: the states' expected_mask is the
 * Search::_types of the break, their base_cost the cost of the pocket roots.  The planned strikes stay on the device and are
 * simulated by bcu_evaluate_plan like explicit candidates (ids = position in the plan: ascending by state, then by pocket,
 * first ball, first link -- deterministic).
 * Two iteration orders the reference leaves to its hash containers are inputs: rail_order and set_order (below); the C++
 * mirror reads them off the real containers (billiard_ai_b200/host/billiard_cuda_search.hpp: PlannerTables). */
typedef struct bcu_planner_desc {
    int32_t n_rails;               /* the rails of bcu_table_desc, same order */
    const float* rail_start;       /* [n_rails][2] Rail::_start (unshifted; canReflect, billiard_search.cpp:854-861) */
    const int32_t* rail_order;     /* [n_rails] or NULL (= as given): iteration order of Configuration::_table._rails
                                      (railTargets takes the first rail hit in this order, :958-978) */
    int32_t n_set_order;           /* rails that take part in the search (horizontal / vertical ones, :860-863) */
    const int32_t* set_order;      /* [n_set_order] iteration order of railCombinations' std::unordered_set of rail ids (:871-897) */
    int32_t n_search_pockets;      /* Configuration::_searchPockets */
    const float* search_pockets;   /* [n_search_pockets][4] _pottingPoint.xy, _normal.xy */
    float minimal_pocket_velocity; /* Configuration::_table.minimalPocketVelocity */
    float diagonal_length_squared; /* Configuration::_table.diagonalLengthSquared */
    int32_t velocity_multiply;     /* Configuration::_velocityMultiplyEnabled: also the two faster strikes of a path (:809) */
} bcu_planner_desc;
int bcu_set_planner(bcu_engine* e, const bcu_planner_desc* desc);
/* plans every state of the last bcu_set_states; *out_count = strikes planned (all states) */
int bcu_plan(bcu_engine* e, int64_t* out_count);
/* the plan, device -> host: any array may be NULL; cost = SearchNode::_cost of the SIMULATION node (the state's base cost
 * plus the search cost of its path), search_cost = SearchNode::_searchCost */
int bcu_fetch_plan(bcu_engine* e, int64_t offset, int64_t n, int32_t* state_idx, float* vx, float* vy, uint64_t* search_cost,
                   uint64_t* cost);
/* simulate the planned strikes (as bcu_evaluate on them: results, bcu_topk, bcu_level_select, bcu_gather afterwards) */
int bcu_evaluate_plan(bcu_engine* e, const bcu_results* out);
/* the same for strikes [first, first + count) of the plan (count < 0: to the end) -- a rank's contiguous share when the plan
 * (identical on every rank: it is deterministic) is simulated on several GPUs; candidate ids stay positions in the whole plan */
int bcu_evaluate_plan_range(bcu_engine* e, int64_t first, int64_t count, const bcu_results* out);

/* ---- measurement hooks -------------------------------------------------------------------------------- */
/* CUDA-event timer on the engine's stream (the stream every kernel of this library is launched on) */
int bcu_timer_begin(bcu_engine* e);
int bcu_timer_end(bcu_engine* e, float* out_ms);
/* writes a 256 MB buffer (> L2) on the engine's stream */
int bcu_flush_l2(bcu_engine* e);
/* vector-pipe issue rate in 1e9 thread-instructions/s: kind 0 DADD+DMUL, 1 DFMA, 2 FADD+FMUL, 3 FFMA */
int bcu_issue_rate(bcu_engine* e, int32_t kind, double* out_ginst_per_s);
/* ms of the last kernel launch (CUDA events on the engine's stream) */
int bcu_last_kernel_ms(bcu_engine* e, float* out_ms);
/* work counters of the last launch when enabled: {events, pair items, quartics, rail items, pocket items} */
int bcu_enable_counters(bcu_engine* e, int enable);
int bcu_get_counters(bcu_engine* e, uint64_t* out5);
/* Reach bound on ball pairs: a pair whose centre distance exceeds what the two balls can cover
 * before the strike's best event time so far cannot produce the next event (the reference would solve its quartic,
 * billiard_physics.cpp:283-381, and discard the result in min(), billiard_search.cpp:2002-2013), so it is not solved.
 * mode 0 (DEFAULT): off -- every test the reference makes is made and every quartic it solves is solved; parity with the
 *         reference follows from construction.  (One shortcut is taken in every mode because it is the reference's own test
 *         read off before the division: a rail whose intersection parameter lambda = N / D has N and D of opposite signs
 *         and |N| > 1e-10 |D| is rejected by `lambda < 0` whatever the rounding, so lambda is not computed for it.)
 * mode 1: on (opt-in, for throughput).  The bound is a theorem about the roots of the collision polynomial plus an
 *         EMPIRICAL characterisation of where the reference's closed-form solver returns them:
 *         audited on > 8e8 random strikes with no differing result, not proven.  Also enables the division-free rail /
 *         pocket / partner filters and the per-line candidate cache, which are conservative by construction.
 * mode 2: audit -- everything is solved and the jobs the bound would have skipped are checked: out2 = {pairs skipped or
 *         marked, marked jobs whose result would have mattered}; the second number must be 0.
 * Counters need bcu_enable_counters(e, 1).  $BCU_PRUNE overrides the default at bcu_create. */
int bcu_set_prune(bcu_engine* e, int mode);
int bcu_get_prune(const bcu_engine* e);
int bcu_get_prune_counters(bcu_engine* e, uint64_t* out2);
/* Test hook: runs the multi-GPU merge kernel (bcu_level_select's step after ncclAllGather) on n_blocks record blocks
 * supplied by the host -- each k entries for n_balls balls, layout
 *   cost u64[k] | id u64[k] | potted u32[k] | time f32[k] | pos f32[k][n_balls][2], entries past the valid count cost = id = ~0,
 * every block ascending by (cost, id) -- and returns the merged block (same layout).  Lets a one-GPU test cover the
 * kernel that otherwise only runs with more than one rank. */
int bcu_test_merge_records(bcu_engine* e, int32_t n_blocks, int32_t k, int32_t n_balls, const void* blocks, void* out_block);
/* number of kernels launched by this engine so far */
int64_t bcu_launch_count(const bcu_engine* e);

#ifdef __cplusplus
}
#endif
#endif /* BILLIARD_CUDA_H */
