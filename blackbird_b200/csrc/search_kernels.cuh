// Search kernels: the per-tick "advance every game until it blocks on the evaluator" kernel
// (SURVEY.md 7.4 lock-step formulation), move begin / choose / root motion.
#pragma once
#include "search_dev.cuh"

namespace bb {

enum { ST_TOP = 0, ST_EXPAND = 1, ST_SELECT = 2, ST_LEAF = 3 };

template <class G>
__device__ __forceinline__ void post_request(const EngineDev<G>& E, Slot<G>& s, int g) {
  int idx = 0;
  if (lane_id() == 0) {
    idx = atomicAdd(E.req_count, 1);
    E.req_slot[idx] = g;
    E.req_state[idx] = s.cur_state;
  }
  s.req_idx = (uint32_t)__shfl_sync(0xFFFFFFFFu, idx, 0);
}

template <class G>
__device__ __forceinline__ void start_descent(Slot<G>& s) {
  s.cur = s.root;
  s.cur_state = s.root_state;
  s.depth = 0;
  s.last_action = -1;
}

// FindMove prologue (MCTS.py:184-186, 297): make a Root if there is none, fix endPlays.
template <class G>
__device__ __forceinline__ void begin_move(const EngineDev<G>& E, Slot<G>& s, NodeHdr* hdr, NodeStat* stat, int play_limit) {
  if (s.root == NONE) {
    if (lane_id() == 0) {
      NodeHdr h;
      h.parent = NONE;
      h.child_base = NONE;
      h.meta = meta_pack(0, 0, G::player(s.root_state), false);
      h.eval = 0.f;
      hdr[0] = h;
      NodeStat st = {0u, 0.f, 0.0};
      stat[0] = st;
    }
    __syncwarp();
    s.root = 0;
    s.n_nodes = 1;
    s.root_parent_player = 0;
  }
  s.end_plays = stat[s.root].plays + (uint32_t)play_limit;
  s.phase = BB_PH_RUN;
  start_descent(s);
}

// Record the searched position, then state = nextState; MoveRoot(state); winner = state.Winner()
// (Blackbird.py:246-253 and, at game end, :256-258).
template <class G>
__device__ void play_move(const EngineDev<G>& E, Slot<G>& s, int g, NodeHdr* hdr, NodeStat* stat, int child_j,
                          unsigned long long* games_done) {
  const int lane = lane_id();
  const NodeHdr rh = hdr[s.root];
  const int k = meta_nchild(rh.meta);
  const size_t rbase = (size_t)g * (E.max_plies + 1);
  const size_t obase = (size_t)g * (E.max_plies + 2);
  const size_t pbase = (size_t)g * E.pool_cap;
  if (E.rec_states) {
    if (s.pool_used + (uint32_t)k > E.pool_cap) { s.phase = BB_PH_ERROR; s.status = BB_ENOMEM; return; }
    for (int j = lane; j < k; j += 32) {
      E.pool_actions[pbase + s.pool_used + j] = (uint16_t)meta_action(hdr[rh.child_base + j].meta);
      E.pool_plays[pbase + s.pool_used + j] = stat[rh.child_base + j].plays;
    }
    if (lane == 0) {
      E.rec_states[rbase + s.ply] = s.root_state;
      E.rec_off[obase + s.ply] = s.pool_used;
      E.rec_off[obase + s.ply + 1] = s.pool_used + (uint32_t)k;
    }
    s.pool_used += (uint32_t)k;
  }
  const uint32_t child = rh.child_base + (uint32_t)child_j;
  const int action = meta_action(hdr[child].meta);
  if (E.chosen && lane == 0) E.chosen[(size_t)g * E.max_plies + s.ply] = (uint16_t)action;
  s.root_parent_player = G::player(s.root_state);
  G::apply(s.root_state, action);
  s.root = child;
  s.ply += 1;
  const int w = G::winner(s.root_state, -1);                 // lastAction is always None (Blackbird.py:242,253)
  if (w != WIN_NONE || s.ply >= E.max_plies) {
    if (lane == 0) {
      if (E.rec_states) {
        E.rec_states[rbase + s.ply] = s.root_state;          // terminal record, zero policy (:256-258)
        E.rec_off[obase + s.ply + 1] = s.pool_used;
      }
      if (E.n_moves) E.n_moves[g] = s.ply;
      if (E.winners) E.winners[g] = w;
      *games_done += 1;
    }
    s.winner = w;
    s.phase = BB_PH_GAME_OVER;
  } else {
    s.phase = BB_PH_MOVE_READY;
  }
}

template <class G>
__global__ void __launch_bounds__(128) k_advance(EngineDev<G> E) {
  const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = lane_id();
  if (g >= E.n_games) return;
  Slot<G> s = E.slots[g];
  if (s.phase != BB_PH_RUN && s.phase != BB_PH_WAIT_EXPAND && s.phase != BB_PH_WAIT_LEAF) return;
  NodeHdr* hdr = E.hdr + (size_t)g * E.node_cap;
  NodeStat* stat = E.stat + (size_t)g * E.node_cap;
  unsigned long long n_sims = 0, n_nodes = 0, n_exp = 0, n_sel = 0, n_moves = 0, n_done = 0, n_evals = 0;

  int step = ST_TOP;
  if (s.phase != BB_PH_RUN) {                                  // an evaluation came back for s.cur
    if (lane == 0) {
      NodeHdr h = hdr[s.cur];
      h.eval = E.resp_value[s.req_idx];
      h.meta |= META_HAS_VALUE;
      hdr[s.cur] = h;
    }
    __syncwarp();
    step = s.phase == BB_PH_WAIT_EXPAND ? ST_EXPAND : ST_LEAF;
    s.phase = BB_PH_RUN;
  }

  while (s.phase == BB_PH_RUN) {
    if (step == ST_TOP) {
      if (E.search == BB_FIXED && s.depth == E.max_depth) { step = ST_LEAF; continue; }     // FixedMCTS.py:23
      const NodeHdr h = hdr[s.cur];
      if (h.child_base == NONE) {                               // Children is None
        if (G::winner(s.cur_state, s.last_action) != WIN_NONE) { step = ST_LEAF; continue; } // :24-26
        post_request(E, s, g);                                  // GetPriors(state) of this node
        n_evals += 1;
        s.phase = BB_PH_WAIT_EXPAND;
        break;
      }
      step = meta_nchild(h.meta) == 0 ? ST_LEAF : ST_SELECT;    // np.sum(LegalActions) == 0
    } else if (step == ST_EXPAND) {
      const int k = expand_node(E, s, hdr, stat, s.cur, E.resp_policy + (size_t)s.req_idx * G::A);
      if (k < 0) { s.phase = BB_PH_ERROR; s.status = BB_ENOMEM; break; }
      n_exp += 1;
      n_nodes += k;
      step = (E.search == BB_DYNAMIC || k == 0) ? ST_LEAF : ST_SELECT;                      // DynamicMCTS.py:27-28
    } else if (step == ST_SELECT) {
      const NodeHdr h = hdr[s.cur];
      double u;
      const int j = puct_select(stat + h.child_base, meta_nchild(h.meta), E.c, &u);
      const int a = j == 0x7FFFFFFF ? 0 : meta_action(hdr[h.child_base + j].meta);
      // MCTS.py:341: the dense argmax lands on an illegal index when no legal U is positive
      if (j == 0x7FFFFFFF || (!(u > 0.0) && j < a)) { s.phase = BB_PH_ERROR; s.status = BB_EASSERT; break; }
      G::apply(s.cur_state, a);
      s.cur = h.child_base + (uint32_t)j;
      s.last_action = a;
      s.depth += 1;
      n_sel += 1;
      step = ST_TOP;
    } else {                                                    // ST_LEAF: SampleValue + _backProp
      const NodeHdr h = hdr[s.cur];
      if (!(h.meta & META_HAS_VALUE)) {
        post_request(E, s, g);                                  // getEvaluation(state) of the leaf
        n_evals += 1;
        s.phase = BB_PH_WAIT_LEAF;
        break;
      }
      backup(E, s, hdr, stat, h.eval);
      n_sims += 1;
      if (stat[s.root].plays >= s.end_plays) {                  // MCTS.py:298-299
        s.phase = BB_PH_MOVE_READY;
        if (E.autoplay) {
          const NodeHdr rh = hdr[s.root];
          const int k = meta_nchild(rh.meta);
          int j;
          if (E.temp_mode == 1) {
            const int row = (int)(((unsigned)g + s.game_no) % (unsigned)E.n_games);   // refill: next game, next row
            j = sample_child(stat + rh.child_base, k, E.uniforms[(size_t)row * E.max_plies + s.ply]);
          } else {
            double u;
            j = puct_select(stat + rh.child_base, k, E.c, &u);
          }
          play_move(E, s, g, hdr, stat, j, &n_done);
          n_moves += 1;
          if (s.moves_budget > 0) s.moves_budget -= 1;
          if (s.phase == BB_PH_GAME_OVER && E.refill) {         // Model.Game(); DropRoot() for the next game
            s.game_no += 1;
            s.root_state = G::initial();
            s.root = NONE;
            s.ply = 0;
            s.pool_used = 0;
            s.winner = WIN_NONE;
            s.phase = BB_PH_MOVE_READY;
          }
          if (s.phase == BB_PH_MOVE_READY && s.moves_budget != 0) begin_move(E, s, hdr, stat, E.play_limit);
        }
      } else {
        start_descent(s);
      }
      step = ST_TOP;
    }
  }
  if (lane == 0) {
    E.slots[g] = s;
    if (s.phase == BB_PH_RUN || s.phase == BB_PH_WAIT_EXPAND || s.phase == BB_PH_WAIT_LEAF) atomicAdd(E.active, 1);
    Counters* c = E.counters;
    if (n_sims) atomicAdd(&c->sims, n_sims);
    if (n_evals) atomicAdd(&c->evals, n_evals);
    if (n_nodes) atomicAdd(&c->nodes, n_nodes);
    if (n_exp) atomicAdd(&c->expansions, n_exp);
    if (n_sel) atomicAdd(&c->select_steps, n_sel);
    if (n_moves) atomicAdd(&c->moves, n_moves);
    if (n_done) atomicAdd(&c->games_done, n_done);
  }
}

template <class G>
__global__ void k_reset(EngineDev<G> E, const typename G::State* init) {
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= E.n_games) return;
  Slot<G> s;
  s.root_state = init ? init[g] : G::initial();
  s.cur_state = s.root_state;
  s.root = NONE;
  s.cur = NONE;
  s.depth = 0;
  s.last_action = -1;
  s.n_nodes = 0;
  s.end_plays = 0;
  s.phase = BB_PH_IDLE;
  s.status = BB_OK;
  s.ply = 0;
  s.root_parent_player = 0;
  s.winner = WIN_NONE;
  s.req_idx = 0;
  s.pool_used = 0;
  s.moves_budget = -1;
  s.game_no = 0;
  E.slots[g] = s;
  if (E.n_moves) E.n_moves[g] = 0;
  if (E.winners) E.winners[g] = WIN_NONE;
  if (E.rec_off) E.rec_off[(size_t)g * (E.max_plies + 2)] = 0;
}

template <class G>
__global__ void k_drop_root(EngineDev<G> E) {
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= E.n_games) return;
  E.slots[g].root = NONE;
  E.slots[g].n_nodes = 0;
  E.slots[g].root_parent_player = 0;
}

template <class G>
__global__ void k_begin_move(EngineDev<G> E, int play_limit, int moves_budget) {
  const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (g >= E.n_games) return;
  Slot<G> s = E.slots[g];
  if (s.phase == BB_PH_GAME_OVER || s.phase == BB_PH_ERROR) return;
  s.moves_budget = moves_budget;
  begin_move(E, s, E.hdr + (size_t)g * E.node_cap, E.stat + (size_t)g * E.node_cap, play_limit);
  if (lane_id() == 0) E.slots[g] = s;
}

// _selectAction(exploring=False): child index -> action per slot (temp 1: uniforms; temp 0: PUCT)
template <class G>
__global__ void k_choose(EngineDev<G> E, int temp_mode, const double* uniforms, int32_t* actions) {
  const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (g >= E.n_games) return;
  const Slot<G> s = E.slots[g];
  if (s.phase != BB_PH_MOVE_READY) { if (lane_id() == 0) actions[g] = -1; return; }
  NodeHdr* hdr = E.hdr + (size_t)g * E.node_cap;
  NodeStat* stat = E.stat + (size_t)g * E.node_cap;
  const NodeHdr rh = hdr[s.root];
  const int k = meta_nchild(rh.meta);
  int j;
  if (temp_mode == 1) j = sample_child(stat + rh.child_base, k, uniforms[g]);
  else { double u; j = puct_select(stat + rh.child_base, k, E.c, &u); }
  if (lane_id() == 0) actions[g] = meta_action(hdr[rh.child_base + j].meta);
}

template <class G>
__global__ void k_move_root(EngineDev<G> E, const int32_t* actions) {
  const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (g >= E.n_games) return;
  Slot<G> s = E.slots[g];
  if (s.phase != BB_PH_MOVE_READY || actions[g] < 0) return;
  NodeHdr* hdr = E.hdr + (size_t)g * E.node_cap;
  NodeStat* stat = E.stat + (size_t)g * E.node_cap;
  const NodeHdr rh = hdr[s.root];
  const int k = meta_nchild(rh.meta);
  int j = -1;
  for (int i = 0; i < k; ++i)
    if (meta_action(hdr[rh.child_base + i].meta) == actions[g]) { j = i; break; }
  if (j < 0) { s.phase = BB_PH_ERROR; s.status = BB_EILLEGAL; }
  else {
    unsigned long long done = 0;
    play_move(E, s, g, hdr, stat, j, &done);
    if (lane_id() == 0 && done) atomicAdd(&E.counters->games_done, done);
    if (lane_id() == 0) atomicAdd(&E.counters->moves, 1ull);
  }
  if (lane_id() == 0) E.slots[g] = s;
}

}  // namespace bb
