// Tree pool layout and the per-slot search state (arenas in HBM; DESIGN.md "data layout").
//
// Per game slot g the pool holds two half-arenas of node_cap nodes each; node ids are
// half-local u32.  A node is ONE 32-byte record (one DRAM sector): the statistics its parent's
// PUCT select reads (Plays, Value, prior) next to its own header (parent, first child, action,
// cached network value).  Children of one expansion are allocated contiguously in ascending
// action order, so a select is one coalesced run of 32-byte loads (one lane per child) that
// also brings in the header of whichever child wins -- one dependent memory latency per tree
// level.  Child STATES are never stored: the descent re-derives them by applying actions to
// the root state held in registers.  At MoveRoot the new root's subtree is copied breadth-first
// into the other half (children stay contiguous and ordered), which bounds a slot's memory by
// one move's growth plus the carried subtree instead of the whole game.
//
// Reference objects restated: MCTS.Node (MCTS.py:7-98) -> Node; MCTS.Root and the _findLeaf
// locals (MCTS.py:120, DynamicMCTS.py:22-32) -> Slot.
#pragma once
#include "common.cuh"

namespace bb {

constexpr uint32_t NONE = 0xFFFFFFFFu;
constexpr int PATH_LEN = 64;        // descent path kept in registers / the slot's path row

struct __align__(16) NodeStat {   // read by the PARENT's select
  uint32_t plays;                 // Node.Plays
  float value;                    // Node.Value, f32 accumulation
  double prior;                   // Parent.Priors[action] (f64, A.9)
};

struct __align__(16) NodeHdr {
  uint32_t parent;                // Node.Parent (NONE for the root of the current half)
  uint32_t child_base;            // first child id; NONE = Children is None
  uint32_t meta;                  // action:12 | nchild:10 | player-1:1 | has_value:1
  float eval;                     // raw network value of this state (cached getEvaluation)
};

struct __align__(32) Node {
  NodeStat s;
  NodeHdr h;
};

constexpr uint32_t META_HAS_VALUE = 1u << 23;
constexpr uint32_t META_PENDING = 1u << 24;   // virtual-loss mode: an evaluation of this node's state is in flight (asked by another context)
__host__ __device__ __forceinline__ uint32_t meta_pack(int action, int nchild, int player, bool has_value) {
  return (uint32_t)action | ((uint32_t)nchild << 12) | ((uint32_t)(player - 1) << 22) | (has_value ? META_HAS_VALUE : 0u);
}
__host__ __device__ __forceinline__ int meta_action(uint32_t m) { return m & 0xFFF; }
__host__ __device__ __forceinline__ int meta_nchild(uint32_t m) { return (m >> 12) & 0x3FF; }
__host__ __device__ __forceinline__ int meta_player(uint32_t m) { return 1 + ((m >> 22) & 1); }

template <class G>
struct Slot {
  typename G::State root_state;   // Root.State
  typename G::State cur_state;    // state of `cur` during a descent
  uint32_t root, cur;             // NONE root = "Root is None"
  int32_t depth, last_action;     // _findLeaf locals
  uint32_t n_nodes;               // bump pointer of the current half
  uint32_t end_plays;             // endPlays (MCTS.py:297)
  uint32_t root_plays;            // Root.Plays, mirrored in the slot (saves a dependent load per simulation)
  int32_t phase, status;          // BB_PH_*, BB_* error code
  int32_t ply;                    // moves played
  int32_t root_parent_player;     // Player of Root.Parent.State (0: Root has no parent), A.13
  int32_t winner;                 // BB_WIN_* after GAME_OVER
  uint32_t req_idx;               // index of the open evaluator request
  uint32_t pool_used;             // record pool bump pointer
  int32_t moves_budget;           // moves this slot may still play (<0: unlimited)
  uint32_t game_no;               // games this slot has finished (refill mode)
  uint32_t half;                  // which half-arena is live
  uint32_t cp_head, cp_n;         // subtree copy in progress: next node to scan, nodes written
  uint32_t cp_src;                // old id of the new root
  uint32_t row;                   // record / uniform row of the game this slot is playing (= slot id unless n_rows > n_games)
  uint32_t rng_pos;               // rollout evaluator: next unread word of the slot's MT19937 word stream
};

// Virtual-loss mode (bb_config.contexts = W >= 1; DESIGN.md "x1"): W simulations of one game are in flight at a time.  A
// context is the _findLeaf locals of one of them -- parked here while its evaluation is out; Slot.cur / cur_state /
// depth / last_action / req_idx are the working copy of the context whose turn it is.
enum { CTX_IDLE = 0, CTX_WAIT_EXPAND = 1, CTX_WAIT_LEAF = 2, CTX_RETRY = 3 };
template <class G>
struct Ctx {
  typename G::State cur_state;
  uint32_t cur;
  int32_t depth, last_action;
  uint32_t req_idx;
  int32_t st;                     // CTX_*
};

// numpy pairwise-sum plan for a dense vector of length A, built on the host.
constexpr int PW_MAX_LEAVES = 64;
constexpr int PW_MAX_OPS = 2 * PW_MAX_LEAVES;
struct PairwisePlan {
  int n_leaves, n_ops;
  int16_t leaf_lo[PW_MAX_LEAVES], leaf_n[PW_MAX_LEAVES];
  int8_t ops[PW_MAX_OPS];         // >=0: push leaf sum; -1: add the two top entries
  // the same tree by levels (pairwise_sum_warp): value slots 0 .. n_leaves-1 are the leaf sums, slot n_leaves + t the t-th
  // addition; additions sorted by height, lvl_off[h] .. lvl_off[h + 1] are the ones whose operands are ready after level h - 1
  int n_levels, root;
  uint8_t lvl_off[16];
  uint8_t add_dst[PW_MAX_LEAVES], add_l[PW_MAX_LEAVES], add_r[PW_MAX_LEAVES];
};

struct Counters {                 // mirrors bb_counters
  unsigned long long sims, evals, nodes, expansions, select_steps, ticks, kernel_launches, moves, games_done;
  unsigned long long cache_lookups, cache_hits;     // evaluation cache (streaming mode): evaluations asked of it / answered by it
};

// Per-tick control block.  Request / activity counters are double-buffered by tick parity so a
// tick is exactly two kernels (advance, evaluator) with no memsets in between: the last CTA of
// the advance kernel clears the other parity and publishes tick+1.
struct TickCtl {
  unsigned tick;                  // ticks completed
  unsigned blocks_done;           // CTA ticket of the running advance kernel
  int req[2];                     // evaluator requests posted in tick t -> req[t & 1]
  int active[2];                  // slots still searching after tick t  -> active[t & 1]
};

template <class G>
struct EngineDev {                // passed to kernels by value
  Slot<G>* slots;
  Node* nodes;                    // [n_games][2][node_cap]
  uint32_t* path;                 // [n_games][PATH_LEN] descent paths of blocked slots
  int n_games;                    // slots served by this launch (a lane of the engine, or all of them)
  int g0;                         // first global slot id of the lane
  int n_total;                    // slots in the whole engine
  uint32_t node_cap;
  int search, max_depth, max_plies;
  int step_budget;                // tree levels a slot may descend per tick before it yields (keeps ticks short)
  int compact_rounds;             // subtree-copy rounds (32 scanned nodes each) a slot in BB_PH_COMPACT runs per tick
  int prefetch;                   // 1: a select asks L2 for every child's own child block (BB_PREFETCH=1; off by default)
  double c;                       // ExplorationRate
  // evaluator mailbox
  TickCtl* ctl;
  int* req_slot;
  typename G::State* req_state;
  float* resp_policy;             // [n_games][A], indexed by request (tick mode) or by slot (streaming mode)
  float* resp_value;
  // streaming mode (q != nullptr): ticket ring + per-slot request states + per-slot answer sequence numbers
  StreamCtl* q;
  unsigned long long* q_ent;      // [q_mask + 1] ring entry of ticket t (entry t & q_mask): low word t, high word slot | RING_PUBLISHED
  typename G::State* q_state;     // [n_games] the state a slot asks about (a slot has one open request at a time)
  unsigned* resp_seq;             // [n_games]
  unsigned q_mask;
  uint2* vl_hdr;                  // [n_games] virtual-loss mode: x = the context whose turn it is (strict cyclic order), y = contexts with a simulation in progress
  Ctx<G>* ctx;                    // [n_games][vl] parked contexts (virtual-loss mode); mailboxes (q_state, resp_*, path rows) are indexed g * vl + context
  int vl;                         // contexts per slot; 0 = exact mode (one simulation in flight per game, the reference's order)
  uint32_t* ecache;               // evaluation cache (common.cuh); nullptr = off
  uint32_t ecache_mask;
  // self-play
  int autoplay;                   // 1: choose / move root / end games on the device
  int refill;                     // 1: a finished game's slot starts the next game (steady-state stepping)
  int n_rows;                     // > 0: bb_selfplay_games -- n_rows whole games on n_total slots; a slot that ends a game takes the
  unsigned* next_row;             //      next unplayed row from this counter (records and uniforms are indexed by row)
  int play_limit;
  int temp_mode;                  // 1: temp == 1 (sample), 0: temp == 0 (PUCT argmax)
  const double* uniforms;         // [n_games][max_plies]
  typename G::State* rec_states;  // [n_games][max_plies+1]
  uint32_t* rec_off;              // [n_games][max_plies+2]
  uint16_t* pool_actions;         // [n_games][pool_cap]
  uint32_t* pool_plays;
  uint16_t* chosen;               // [n_games][max_plies]
  int32_t* n_moves;
  int32_t* winners;
  uint32_t pool_cap;
  Counters* counters;
  // rollout evaluator (BB_EVAL_ROLLOUT: the base-class GetPriors / SampleValue of MCTS.py:346-383).  Priors are the
  // constant 1, so a node's f64 `prior` field holds Node.Value instead (Python 0 / 0.5 / 1 sums, exact: A.9) and the
  // f32 `value` field is unused.  Rollout draws come from a per-slot stream of raw 32-bit MT19937 outputs supplied by
  // the host (np.random's own stream, so that the reference's games replay bit for bit).
  int rollout;
  const uint32_t* rng_words;      // [n_games][rng_cap]
  uint32_t rng_cap;

  __device__ __forceinline__ Node* half(int g, uint32_t h) const { return nodes + ((size_t)g * 2 + h) * node_cap; }
};

}  // namespace bb
