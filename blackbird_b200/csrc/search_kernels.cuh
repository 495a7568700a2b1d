// Search kernels: the per-tick "advance every game until it blocks on the evaluator" kernel
// (SURVEY.md 7.4 lock-step formulation), move begin / choose / root motion.
#pragma once
#include "search_dev.cuh"

namespace bb {

enum { ST_TOP = 0, ST_EXPAND = 1, ST_SELECT = 2, ST_LEAF = 3 };

__device__ __forceinline__ bool phase_busy(int ph) {
  return ph == BB_PH_RUN || ph == BB_PH_WAIT_EXPAND || ph == BB_PH_WAIT_LEAF || ph == BB_PH_COMPACT;
}

template <class G>
__device__ __forceinline__ void post_request(const EngineDev<G>& E, Slot<G>& s, int g, unsigned parity) {
  int idx = 0;
  if (E.q) {                                                  // streaming mode: park the state, take a ticket, publish its ring entry
    if (lane_id() == 0) {
      E.q_state[g] = s.cur_state;
      idx = (int)atomicAdd(&E.q->tail, 1u);
      volatile unsigned long long* ent = reinterpret_cast<volatile unsigned long long*>(&E.q_ent[(unsigned)idx & E.q_mask]);
      unsigned long long was = *ent;                          // the entry's previous ticket (idx - ring) must have been read by its pipeline
      fence_acq_rel();                                        // state before entry
      while ((unsigned)(was >> 32) & RING_PUBLISHED) { __nanosleep(100); was = *ent; }
      *ent = ((unsigned long long)((unsigned)g | RING_PUBLISHED) << 32) | (unsigned)idx;
    }
    s.req_idx = (uint32_t)__shfl_sync(0xFFFFFFFFu, idx, 0);
    return;
  }
  if (lane_id() == 0) {
    idx = atomicAdd(&E.ctl->req[parity], 1);
    E.req_slot[idx] = g;
    E.req_state[idx] = s.cur_state;
  }
  s.req_idx = (uint32_t)__shfl_sync(0xFFFFFFFFu, idx, 0);
}

template <class G>
__device__ __forceinline__ void start_descent(Slot<G>& s, Path& path) {
  s.cur = s.root;
  s.cur_state = s.root_state;
  s.depth = 0;
  s.last_action = -1;
  path.set(0, s.root, G::player(s.root_state));
}

// FindMove prologue (MCTS.py:184-186, 297): make a Root if there is none, fix endPlays.
template <class G>
__device__ __forceinline__ void begin_move(const EngineDev<G>& E, Slot<G>& s, Node* nodes, Path& path, int play_limit) {
  uint32_t root_plays;
  if (s.root == NONE) {
    if (lane_id() == 0) {
      Node n;
      n.s.plays = 0u;
      n.s.value = 0.f;
      n.s.prior = 0.0;
      n.h.parent = NONE;
      n.h.child_base = NONE;
      n.h.meta = meta_pack(0, 0, G::player(s.root_state), false);
      n.h.eval = 0.f;
      nodes[0] = n;
    }
    __syncwarp();
    s.root = 0;
    s.n_nodes = 1;
    s.root_parent_player = 0;
    root_plays = 0;
  } else {
    root_plays = nodes[s.root].s.plays;
  }
  s.root_plays = root_plays;
  s.end_plays = root_plays + (uint32_t)play_limit;
  s.phase = BB_PH_RUN;
  start_descent(s, path);
}

// Record the searched position, then state = nextState; MoveRoot(state); winner = state.Winner()
// (Blackbird.py:246-253 and, at game end, :256-258).  A game that goes on enters BB_PH_COMPACT:
// the chosen child's subtree is copied into the other half-arena (subtree_copy).
template <class G>
__device__ void play_move(const EngineDev<G>& E, Slot<G>& s, int g, Node* nodes, int child_j, unsigned long long* games_done) {
  const int lane = lane_id();
  const NodeHdr rh = nodes[s.root].h;
  const int k = meta_nchild(rh.meta);
  const size_t row = E.n_rows ? (size_t)s.row : (size_t)g;   // records belong to the game, not to the slot that plays it
  const size_t rbase = row * (E.max_plies + 1);
  const size_t obase = row * (E.max_plies + 2);
  const size_t pbase = row * E.pool_cap;
  if (E.rec_states) {
    if (s.pool_used + (uint32_t)k > E.pool_cap) { s.phase = BB_PH_ERROR; s.status = BB_ENOMEM; return; }
    for (int j = lane; j < k; j += 32) {
      const Node kid = nodes[rh.child_base + j];
      E.pool_actions[pbase + s.pool_used + j] = (uint16_t)meta_action(kid.h.meta);
      E.pool_plays[pbase + s.pool_used + j] = kid.s.plays;
    }
    if (lane == 0) {
      E.rec_states[rbase + s.ply] = s.root_state;
      E.rec_off[obase + s.ply] = s.pool_used;
      E.rec_off[obase + s.ply + 1] = s.pool_used + (uint32_t)k;
    }
    s.pool_used += (uint32_t)k;
  }
  const uint32_t child = rh.child_base + (uint32_t)child_j;
  const int action = meta_action(nodes[child].h.meta);
  if (E.chosen && lane == 0) E.chosen[row * E.max_plies + s.ply] = (uint16_t)action;
  s.root_parent_player = G::player(s.root_state);
  apply_child<G>(s.root_state, action);
  s.ply += 1;
  const int w = winner_w<G>(s.root_state, -1);               // lastAction is always None (Blackbird.py:242,253)
  if (w != WIN_NONE || s.ply >= E.max_plies) {
    if (lane == 0) {
      if (E.rec_states) {
        E.rec_states[rbase + s.ply] = s.root_state;          // terminal record, zero policy (:256-258)
        E.rec_off[obase + s.ply + 1] = s.pool_used;
      }
      if (E.n_moves) E.n_moves[row] = s.ply;
      if (E.winners) E.winners[row] = w;
      *games_done += 1;
    }
    s.root = child;
    s.winner = w;
    s.phase = BB_PH_GAME_OVER;
  } else {
    s.cp_src = child;
    s.cp_n = 0;
    s.cp_head = 0;
    s.phase = BB_PH_COMPACT;
  }
}

// BB_PH_COMPACT -> BB_PH_MOVE_READY: swap halves once the copy is complete
template <class G>
__device__ __forceinline__ void run_compaction(const EngineDev<G>& E, Slot<G>& s, int g, WarpScratch& ws, int rounds) {
  const bool done = subtree_copy(E, s, E.half(g, s.half), E.half(g, s.half ^ 1u), ws, rounds);
  if (done && s.phase == BB_PH_COMPACT) {
    s.half ^= 1u;
    s.root = 0;
    s.n_nodes = s.cp_n;
    s.phase = BB_PH_MOVE_READY;
  }
}


// Search warps per SM in streaming mode and their register cap.  An SM's register file is four 16 K-register
// quarters (one per scheduler); a CTA's warps are dealt over them and allocations are per warp, so what must fit is
// ceil(warps / 4) * registers * 32 of this kernel plus the same figure of the network kernel (3 * 56 * 32) <= 16384.
// 12 warps of 112 registers beat 8 warps of 128 (the search is latency-bound: more warps, not more registers).
#ifndef BB_STREAM_WARPS
#define BB_STREAM_WARPS 12
#endif
#ifndef BB_STREAM_MAXREG
#define BB_STREAM_MAXREG 112
#endif
constexpr int STREAM_WARPS_PER_CTA = BB_STREAM_WARPS;
#ifndef BB_ADV_MIN_CTAS
#define BB_ADV_MIN_CTAS 1
#endif

#ifdef BB_SEARCH_TIMING
#define ST_T0() long long st_t = clock64()
#define ST_ADD(i) do { const long long st_n = clock64(); st_acc[i] += st_n - st_t; st_cnt[i] += 1; st_t = st_n; } while (0)
#else
#define ST_T0()
#define ST_ADD(i)
#endif
// RO: the rollout evaluator (BB_EVAL_ROLLOUT) -- expansion needs no evaluation (constant priors), every leaf visit plays a
// fresh random rollout (MCTS.py:360-383), values accumulate exactly; a slot never waits for an evaluator.
template <class G, bool RO = false>
__device__ int advance_slot(const EngineDev<G>& E, int g, unsigned par, WarpScratch& ws, uint32_t* ticket = nullptr) {
  const int lane = lane_id();
#ifdef BB_SEARCH_TIMING
  long long st_acc[12] = {0};
  int st_cnt[12] = {0};
#endif
  ST_T0();
  Slot<G> s = E.slots[g];
  if (!phase_busy(s.phase)) return s.phase;
#ifdef BB_ADV_TIMING
  const long long dbg_t0 = clock64();
  const int dbg_phase0 = s.phase;
  int dbg_iters = 0, dbg_compact = 0;
#endif
  unsigned long long n_sims = 0, n_nodes = 0, n_exp = 0, n_sel = 0, n_moves = 0, n_done = 0, n_evals = 0, n_look = 0, n_hit = 0;
  constexpr bool EC = !RO && G::A <= ECACHE_MAX_ACTIONS;     // evaluation cache: narrow games, network evaluator
  bool pol_cached = false;                                    // the pending expansion takes its policy from ws.cpol (cache hit)
  uint32_t* prow = E.path + (size_t)g * PATH_LEN;
  Path path;
  Node* nodes = E.half(g, s.half);
  NodeHdr ch;
  ch.parent = NONE; ch.child_base = NONE; ch.meta = 0; ch.eval = 0.f;
  int step = ST_TOP;
  if (s.phase == BB_PH_WAIT_EXPAND || s.phase == BB_PH_WAIT_LEAF) {   // an evaluation came back for s.cur
    path.load(prow);
    BB_ASSERT(s.cur < s.n_nodes && (E.q || s.req_idx < (uint32_t)E.n_games));
    ch = nodes[s.cur].h;
    ch.eval = E.q ? __ldcg(&E.resp_value[g]) : E.resp_value[s.req_idx];   // streaming mode answers by slot
    ch.meta |= META_HAS_VALUE;
    if (lane == 0) nodes[s.cur].h = ch;
    step = s.phase == BB_PH_WAIT_EXPAND ? ST_EXPAND : ST_LEAF;
    s.phase = BB_PH_RUN;
  } else if (s.phase == BB_PH_RUN) {                                   // fresh from k_begin_move
    path.load(prow);
    BB_ASSERT(s.cur < s.n_nodes);
    ch = nodes[s.cur].h;
  }

  int budget = E.step_budget;
  ST_ADD(0);                                                  // slot, path, header, answer loads
  for (;;) {
#ifdef BB_ADV_TIMING
    dbg_iters++;
    if (s.phase == BB_PH_COMPACT) dbg_compact++;
#endif
    if (s.phase == BB_PH_COMPACT) {
      budget -= 2 * E.compact_rounds;                        // a copy round costs about two tree levels of latency
      run_compaction(E, s, g, ws, E.compact_rounds);
      if (s.phase != BB_PH_MOVE_READY || !E.autoplay || s.moves_budget == 0) break;
      nodes = E.half(g, s.half);
      begin_move(E, s, nodes, path, E.play_limit);
      ch = nodes[s.cur].h;
      step = ST_TOP;
    }
    if (s.phase != BB_PH_RUN) break;
    if (step == ST_TOP) {
      if (E.search == BB_FIXED && s.depth == E.max_depth) { step = ST_LEAF; continue; }     // FixedMCTS.py:23
      if (ch.child_base == NONE) {                              // Children is None
        if (winner_w<G>(s.cur_state, s.last_action) != WIN_NONE) { step = ST_LEAF; continue; } // :24-26
        if (RO) { step = ST_EXPAND; continue; }                 // AddChildren with GetPriors == ones: nothing to ask
        if constexpr (EC) if (E.ecache) {                       // the memo of GetPriors / SampleValue (Blackbird.py:350,372)
          float v;
          n_look += 1;
          if (ecache_lookup(E, s.cur_state, ws, &v)) {
            n_hit += 1;
            ch.eval = v;
            ch.meta |= META_HAS_VALUE;
            if (lane == 0) nodes[s.cur].h = ch;
            pol_cached = true;
            step = ST_EXPAND;
            continue;
          }
        }
        ST_ADD(6);
        post_request(E, s, g, par);                             // GetPriors(state) of this node
        ST_ADD(1);
        n_evals += 1;
        s.phase = BB_PH_WAIT_EXPAND;
        break;
      }
      step = meta_nchild(ch.meta) == 0 ? ST_LEAF : ST_SELECT;   // np.sum(LegalActions) == 0
    } else if (step == ST_EXPAND) {
      ST_ADD(7);
      const int k = expand_node<G, RO>(E, s, nodes, s.cur, &ch, RO ? nullptr : pol_cached ? ws.cpol : E.resp_policy + (size_t)(E.q ? (uint32_t)g : s.req_idx) * G::A, ws,
                                       pol_cached);
      pol_cached = false;
      ST_ADD(2);
      if (k < 0) { s.phase = BB_PH_ERROR; s.status = BB_ENOMEM; break; }
      n_exp += 1;
      n_nodes += k;
      step = (E.search == BB_DYNAMIC || k == 0) ? ST_LEAF : ST_SELECT;                      // DynamicMCTS.py:27-28
    } else if (step == ST_SELECT) {
      double u;
      NodeHdr chosen = ch;
      ST_ADD(8);
#ifdef BB_DEBUG
      if (ch.child_base + (uint32_t)meta_nchild(ch.meta) > s.n_nodes && lane == 0)
        printf("DBG g=%d cur=%u depth=%d cb=%u nch=%d meta=%x n_nodes=%u ply=%d half=%u root=%u rp=%u cp_n=%u cp_head=%u\n", g, s.cur, s.depth,
               ch.child_base, meta_nchild(ch.meta), ch.meta, s.n_nodes, s.ply, s.half, s.root, s.root_plays, s.cp_n, s.cp_head);
#endif
      BB_ASSERT(ch.child_base + (uint32_t)meta_nchild(ch.meta) <= s.n_nodes);
      const int j = puct_select<SelectChunks<G>::N, RO>(nodes + ch.child_base, meta_nchild(ch.meta), E.c, &u, &chosen, E.prefetch ? nodes : nullptr);
      const int a = j == 0x7FFFFFFF ? 0 : meta_action(chosen.meta);
      // MCTS.py:341: the dense argmax lands on an illegal index when no legal U is positive
      if (j == 0x7FFFFFFF || (!(u > 0.0) && j < a)) { s.phase = BB_PH_ERROR; s.status = BB_EASSERT; break; }
      apply_child<G>(s.cur_state, a);
      s.cur = ch.child_base + (uint32_t)j;
      ch = chosen;
      s.last_action = a;
      s.depth += 1;
      path.set(s.depth, s.cur, meta_player(ch.meta));
      n_sel += 1;
      budget -= 1;
      step = ST_TOP;
      ST_ADD(3);
    } else {                                                    // ST_LEAF: SampleValue + _backProp
      if (RO) {
        const int halves = rollout_value(E, s, g);
        if (halves < 0) { s.phase = BB_PH_ERROR; break; }
        backup_rollout(s, nodes, path, halves);
      } else {
      if constexpr (EC) if (E.ecache && !(ch.meta & META_HAS_VALUE)) {
        float v;
        n_look += 1;
        if (ecache_lookup(E, s.cur_state, ws, &v)) {
          n_hit += 1;
          ch.eval = v;
          ch.meta |= META_HAS_VALUE;
          if (lane == 0) nodes[s.cur].h = ch;
        }
      }
      if (!(ch.meta & META_HAS_VALUE)) {
        ST_ADD(9);
        post_request(E, s, g, par);                             // getEvaluation(state) of the leaf
        ST_ADD(1);
        n_evals += 1;
        s.phase = BB_PH_WAIT_LEAF;
        break;
      }
      ST_ADD(10);
      backup(s, nodes, path, ch.eval);
      ST_ADD(4);
      }
      s.root_plays += 1;
      n_sims += 1;
      if (s.root_plays >= s.end_plays) {                        // MCTS.py:298-299
        s.phase = BB_PH_MOVE_READY;
        if (E.autoplay) {
          BB_ASSERT(s.root < s.n_nodes);
          const NodeHdr rh = nodes[s.root].h;
          const int k = meta_nchild(rh.meta);
          BB_ASSERT(rh.child_base + (uint32_t)k <= s.n_nodes && s.ply < E.max_plies);
          int j;
          if (E.temp_mode == 1) {
            const int row = E.n_rows ? (int)s.row : (int)(((unsigned)g + s.game_no) % (unsigned)E.n_total);   // refill: next game, next row
            j = sample_child(nodes + rh.child_base, k, E.uniforms[(size_t)row * E.max_plies + s.ply]);
          } else {
            double u;
            NodeHdr dummy;
            dummy.meta = 0;
            j = puct_select<SelectChunks<G>::N, RO>(nodes + rh.child_base, k, E.c, &u, &dummy);
            // MCTS.py:341: the dense argmax lands on an illegal index when no legal U is positive
            if (j != 0x7FFFFFFF && !(u > 0.0) && j < meta_action(dummy.meta)) j = 0x7FFFFFFF;
          }
          // a root without children (no legal move and no winner: a stalemated DragonChess position) or without a
          // candidate (every prior NaN): the reference trips `assert root.LegalActions[choice] == 1` (MCTS.py:341)
          if (rh.child_base == NONE || k == 0 || j < 0 || j >= k) { s.phase = BB_PH_ERROR; s.status = BB_EASSERT; break; }
          play_move(E, s, g, nodes, j, &n_done);
          n_moves += 1;
          if (s.moves_budget > 0) s.moves_budget -= 1;
          bool again = s.phase == BB_PH_GAME_OVER && E.refill;
          if (s.phase == BB_PH_GAME_OVER && E.n_rows) {         // whole-game mode with more games than slots: take the next unplayed row
            unsigned r = 0;
            if (lane == 0) r = atomicAdd(E.next_row, 1u);
            r = __shfl_sync(0xFFFFFFFFu, r, 0);
            again = r < (unsigned)E.n_rows;
            if (again) {
              s.row = r;
              if (lane == 0) E.rec_off[(size_t)r * (E.max_plies + 2)] = 0;
            }
          }
          if (again) {                                          // Model.Game(); DropRoot() for the next game
            s.game_no += 1;
            s.root_state = G::initial();
            s.root = NONE;
            s.n_nodes = 0;
            s.ply = 0;
            s.pool_used = 0;
            s.winner = WIN_NONE;
            s.phase = BB_PH_MOVE_READY;
            if (s.moves_budget != 0) {
              begin_move(E, s, nodes, path, E.play_limit);
              ch = nodes[s.cur].h;
            }
          }
        }
      } else {
        start_descent(s, path);
        ch = nodes[s.cur].h;
      }
      step = ST_TOP;
      if (budget <= 0 && s.phase == BB_PH_RUN) { path.store(prow); break; }   // yield: next tick resumes at the root
    }
  }
  if (s.phase == BB_PH_WAIT_EXPAND || s.phase == BB_PH_WAIT_LEAF) path.store(prow);
#ifdef BB_ADV_TIMING
  {
    const long long dt = clock64() - dbg_t0;
    const unsigned tk = E.ctl->tick;
    if (lane == 0 && (tk % 997) == 5 && (dt > 30000 || (g & 63) == 0))
      printf("ADV tick %u g %d cyc %lld ph %d->%d iters %d compact %d sel %llu sims %llu exp %llu evals %llu moves %llu depth %d nodes %u\n", tk, g, dt,
             dbg_phase0, s.phase, dbg_iters, dbg_compact, n_sel, n_sims, n_exp, n_evals, n_moves, s.depth, s.n_nodes);
  }
#endif
  if (ticket) *ticket = s.req_idx;
#ifdef BB_SEARCH_TIMING
  ST_ADD(11);
  if (lane == 0 && E.q) for (int i = 0; i < 12; ++i) { atomicAdd(&E.q->sdbg[i], (unsigned long long)st_acc[i]); atomicAdd(&E.q->sdbg[12 + i], (unsigned long long)st_cnt[i]); }
#endif
  if (lane == 0) {
    E.slots[g] = s;
    if (!E.q && phase_busy(s.phase)) atomicAdd(&E.ctl->active[par], 1);
    Counters* c = E.counters;
    if (n_sims) atomicAdd(&c->sims, n_sims);
    if (n_evals) atomicAdd(&c->evals, n_evals);
    if (n_nodes) atomicAdd(&c->nodes, n_nodes);
    if (n_exp) atomicAdd(&c->expansions, n_exp);
    if (n_sel) atomicAdd(&c->select_steps, n_sel);
    if (n_moves) atomicAdd(&c->moves, n_moves);
    if (n_done) atomicAdd(&c->games_done, n_done);
    if (n_look) atomicAdd(&c->cache_lookups, n_look);
    if (n_hit) atomicAdd(&c->cache_hits, n_hit);
  }
  __syncwarp();
  return s.phase;
}


// ---- virtual-loss mode (bb_config.contexts = W; streaming mode only) -----------------------------------------------------
// W simulations of a game are in flight at a time.  A simulation counts its visit at every node ON THE WAY DOWN (the root at the
// start of the descent, each child when it is selected) and adds its value on the way up, so the next descent already sees the
// visits of the ones still waiting for the network and is steered elsewhere -- the "virtual loss": a visit without a value
// lowers Value / Plays.  Everything else is the exact mode's code (select, expansion, the value transform).
// Order is fixed, so results do not depend on when answers arrive: contexts take turns in strict cyclic order; a context runs
// until it must wait (its evaluation is posted, or it met a node whose evaluation another context has asked for: META_PENDING)
// or its simulation ends; then the NEXT context's turn comes, and a context that waits for an answer is waited for.  Between
// two turns of a context every other context has had one, so a pending node it met has been resolved when it looks again.
// With W = 1 the visits counted on the way down are never read before the values arrive: the records equal the exact mode's
// bit for bit (tests).  W > 1 is NOT the reference's search: reported separately, never the bench line.
template <class G>
__device__ int advance_slot_vl(const EngineDev<G>& E, int g, WarpScratch& ws, uint32_t* ticket, uint32_t* mailbox) {
  const int lane = lane_id();
  const int W = E.vl;
  Slot<G> s = E.slots[g];
  if (!phase_busy(s.phase)) return s.phase;
  uint32_t ctx_cur, ctx_busy;                                  // kept beside the slot (the exact mode's Slot record stays as it was)
  { const uint2 h = E.vl_hdr[g]; ctx_cur = h.x; ctx_busy = h.y; }
  unsigned long long n_sims = 0, n_nodes = 0, n_exp = 0, n_sel = 0, n_moves = 0, n_done = 0, n_evals = 0, n_look = 0, n_hit = 0;
  constexpr bool EC = G::A <= ECACHE_MAX_ACTIONS;
  Path path;
  Node* nodes = E.half(g, s.half);
  NodeHdr ch;
  ch.parent = NONE; ch.child_base = NONE; ch.meta = 0; ch.eval = 0.f;
  if (s.phase == BB_PH_WAIT_EXPAND || s.phase == BB_PH_WAIT_LEAF) s.phase = BB_PH_RUN;   // "the context whose turn it is waits": the caller has seen its answer
  int budget = E.step_budget * (W > 4 ? 4 : W);
  for (;;) {
    if (s.phase == BB_PH_COMPACT) {
      budget -= 2 * E.compact_rounds;
      run_compaction(E, s, g, ws, E.compact_rounds);
      if (s.phase != BB_PH_MOVE_READY || !E.autoplay || s.moves_budget == 0) break;
      nodes = E.half(g, s.half);
      begin_move(E, s, nodes, path, E.play_limit);
    }
    if (s.phase != BB_PH_RUN) break;
    // ---- the turn of context c ---------------------------------------------------------------------------------------
    const uint32_t c = ctx_cur, mb = (uint32_t)g * (uint32_t)W + c;
    Ctx<G>* cx = E.ctx + mb;
    uint32_t* prow = E.path + (size_t)mb * PATH_LEN;
    const int cst = cx->st;
    int step = ST_TOP;
    bool cached = false;                                        // the evaluation this context parks on was answered by the memo
    float cached_v = 0.f;
    if (cst == CTX_WAIT_EXPAND || cst == CTX_WAIT_LEAF) {
      const uint32_t tk = cx->req_idx;
      if (*reinterpret_cast<volatile unsigned*>(&E.resp_seq[mb]) != tk + 1u) {       // not answered yet: the slot waits for THIS context
        s.req_idx = tk;
        s.phase = BB_PH_WAIT_LEAF;
        break;
      }
      fence_acq_rel();
      s.cur_state = cx->cur_state; s.cur = cx->cur; s.depth = cx->depth; s.last_action = cx->last_action;
      path.load(prow);
      ch = nodes[s.cur].h;
      ch.eval = __ldcg(&E.resp_value[mb]);
      ch.meta = (ch.meta | META_HAS_VALUE) & ~META_PENDING;
      if (lane == 0) nodes[s.cur].h = ch;
      step = cst == CTX_WAIT_EXPAND ? ST_EXPAND : ST_LEAF;
    } else if (cst == CTX_RETRY) {
      s.cur_state = cx->cur_state; s.cur = cx->cur; s.depth = cx->depth; s.last_action = cx->last_action;
      path.load(prow);
      ch = nodes[s.cur].h;
      BB_ASSERT(!(ch.meta & META_PENDING));
    } else {                                                      // idle: start the next simulation, if the move has any left
      if (s.root_plays >= s.end_plays) {
        if (ctx_busy != 0) { ctx_cur = (c + 1u) % (uint32_t)W; continue; }
        // ---- every simulation of the move has come back: MCTS.py:298-299, then the self-play step as in the exact mode ----
        ctx_cur = 0;
        s.phase = BB_PH_MOVE_READY;
        if (!E.autoplay) break;
        BB_ASSERT(s.root < s.n_nodes);
        const NodeHdr rh = nodes[s.root].h;
        const int k = meta_nchild(rh.meta);
        BB_ASSERT(rh.child_base + (uint32_t)k <= s.n_nodes && s.ply < E.max_plies);
        int j;
        if (E.temp_mode == 1) {
          const int row = E.n_rows ? (int)s.row : (int)(((unsigned)g + s.game_no) % (unsigned)E.n_total);
          j = sample_child(nodes + rh.child_base, k, E.uniforms[(size_t)row * E.max_plies + s.ply]);
        } else {
          double u;
          NodeHdr dummy;
          dummy.meta = 0;
          j = puct_select<SelectChunks<G>::N, false>(nodes + rh.child_base, k, E.c, &u, &dummy);
          if (j != 0x7FFFFFFF && !(u > 0.0) && j < meta_action(dummy.meta)) j = 0x7FFFFFFF;
        }
        if (rh.child_base == NONE || k == 0 || j < 0 || j >= k) { s.phase = BB_PH_ERROR; s.status = BB_EASSERT; break; }
        play_move(E, s, g, nodes, j, &n_done);
        n_moves += 1;
        if (s.moves_budget > 0) s.moves_budget -= 1;
        bool again = s.phase == BB_PH_GAME_OVER && E.refill;
        if (s.phase == BB_PH_GAME_OVER && E.n_rows) {
          unsigned r = 0;
          if (lane == 0) r = atomicAdd(E.next_row, 1u);
          r = __shfl_sync(0xFFFFFFFFu, r, 0);
          again = r < (unsigned)E.n_rows;
          if (again) {
            s.row = r;
            if (lane == 0) E.rec_off[(size_t)r * (E.max_plies + 2)] = 0;
          }
        }
        if (again) {
          s.game_no += 1;
          s.root_state = G::initial();
          s.root = NONE;
          s.n_nodes = 0;
          s.ply = 0;
          s.pool_used = 0;
          s.winner = WIN_NONE;
          s.phase = BB_PH_MOVE_READY;
          if (s.moves_budget != 0) begin_move(E, s, nodes, path, E.play_limit);
        }
        if (budget <= 0) break;
        continue;
      }
      start_descent(s, path);
      if (lane == 0) nodes[s.root].s.plays += 1;               // the visit, on the way down
      s.root_plays += 1;
      ctx_busy += 1;
      ch = nodes[s.root].h;
    }
    // ---- run the context until its simulation ends or it must wait -----------------------------------------------------
    int park = -1;                                              // CTX_* to park the context with; CTX_IDLE = simulation finished
    while (park < 0) {
      if (step == ST_TOP) {
        if (E.search == BB_FIXED && s.depth == E.max_depth) { step = ST_LEAF; continue; }
        if (ch.child_base == NONE) {
          if (winner_w<G>(s.cur_state, s.last_action) != WIN_NONE) { step = ST_LEAF; continue; }
          if (ch.meta & META_PENDING) { park = CTX_RETRY; break; }
          // A memoised answer is taken exactly like a computed one -- the context parks, the node is pending, the answer sits in
          // the mailbox when the turn comes back -- so that WHEN another game's answer reached the cache cannot change the order
          // of events in this game (the memo is filled by all games: hit or miss depends on timing).
          if constexpr (EC) if (E.ecache) {
            n_look += 1;
            if (ecache_lookup(E, s.cur_state, ws, &cached_v)) { n_hit += 1; cached = true; }
          }
          park = CTX_WAIT_EXPAND;
          break;
        }
        step = meta_nchild(ch.meta) == 0 ? ST_LEAF : ST_SELECT;
      } else if (step == ST_EXPAND) {
        const int k = expand_node<G, false>(E, s, nodes, s.cur, &ch, E.resp_policy + (size_t)mb * G::A, ws, false);
        if (k < 0) { s.phase = BB_PH_ERROR; s.status = BB_ENOMEM; break; }
        n_exp += 1;
        n_nodes += k;
        step = (E.search == BB_DYNAMIC || k == 0) ? ST_LEAF : ST_SELECT;
      } else if (step == ST_SELECT) {
        double u;
        NodeHdr chosen = ch;
        BB_ASSERT(ch.child_base + (uint32_t)meta_nchild(ch.meta) <= s.n_nodes);
        const int j = puct_select<SelectChunks<G>::N, false>(nodes + ch.child_base, meta_nchild(ch.meta), E.c, &u, &chosen);
        const int a = j == 0x7FFFFFFF ? 0 : meta_action(chosen.meta);
        if (j == 0x7FFFFFFF || (!(u > 0.0) && j < a)) { s.phase = BB_PH_ERROR; s.status = BB_EASSERT; break; }
        apply_child<G>(s.cur_state, a);
        s.cur = ch.child_base + (uint32_t)j;
        ch = chosen;
        if (lane == 0) nodes[s.cur].s.plays += 1;              // the visit, on the way down
        __syncwarp();
        s.last_action = a;
        s.depth += 1;
        path.set(s.depth, s.cur, meta_player(ch.meta));
        n_sel += 1;
        budget -= 1;
        step = ST_TOP;
      } else {                                                  // ST_LEAF
        if (!(ch.meta & META_HAS_VALUE)) {
          if (ch.meta & META_PENDING) { park = CTX_RETRY; break; }
          if constexpr (EC) if (E.ecache) {
            n_look += 1;
            if (ecache_lookup(E, s.cur_state, ws, &cached_v)) { n_hit += 1; cached = true; }
          }
          park = CTX_WAIT_LEAF;
          break;
        }
        backup<G, true>(s, nodes, path, ch.eval);
        n_sims += 1;
        ctx_busy -= 1;
        park = CTX_IDLE;
      }
    }
    if (s.phase == BB_PH_ERROR) break;
    // ---- park the context; the next one's turn ---------------------------------------------------------------------------
    if ((park == CTX_WAIT_EXPAND || park == CTX_WAIT_LEAF) && cached) {
      if constexpr (EC) if (lane < G::A) E.resp_policy[(size_t)mb * G::A + lane] = ws.cpol[lane];
      if (lane == 0) {
        ch.meta |= META_PENDING;
        nodes[s.cur].h = ch;
        E.resp_value[mb] = cached_v;
        E.resp_seq[mb] = 0xFFFFFFFFu;                          // "ticket" 0xFFFFFFFE: answered
      }
      s.req_idx = 0xFFFFFFFEu;
      __syncwarp();
    } else if (park == CTX_WAIT_EXPAND || park == CTX_WAIT_LEAF) {
      int idx = 0;
      if (lane == 0) {
        ch.meta |= META_PENDING;
        nodes[s.cur].h = ch;
        E.q_state[mb] = s.cur_state;
        idx = (int)atomicAdd(&E.q->tail, 1u);
        volatile unsigned long long* ent = reinterpret_cast<volatile unsigned long long*>(&E.q_ent[(unsigned)idx & E.q_mask]);
        unsigned long long was = *ent;
        fence_acq_rel();
        while ((unsigned)(was >> 32) & RING_PUBLISHED) { __nanosleep(100); was = *ent; }
        *ent = ((unsigned long long)(mb | RING_PUBLISHED) << 32) | (unsigned)idx;
      }
      s.req_idx = (uint32_t)__shfl_sync(0xFFFFFFFFu, idx, 0);
      n_evals += 1;
    }
    if (park != CTX_IDLE) {
      path.store(prow);
      if (lane == 0) { cx->cur_state = s.cur_state; cx->cur = s.cur; cx->depth = s.depth; cx->last_action = s.last_action; cx->req_idx = s.req_idx; }
    }
    if (lane == 0) cx->st = park;
    __syncwarp();
    ctx_cur = (c + 1u) % (uint32_t)W;
    if (budget <= 0 && park == CTX_IDLE) break;                 // yield between simulations (keeps a pass over the warp's slots short)
  }
  if (ticket) *ticket = s.req_idx;
  if (mailbox) *mailbox = (uint32_t)g * (uint32_t)W + ctx_cur;
  if (lane == 0) {
    E.slots[g] = s;
    E.vl_hdr[g] = make_uint2(ctx_cur, ctx_busy);
    Counters* c = E.counters;
    if (n_sims) atomicAdd(&c->sims, n_sims);
    if (n_evals) atomicAdd(&c->evals, n_evals);
    if (n_nodes) atomicAdd(&c->nodes, n_nodes);
    if (n_exp) atomicAdd(&c->expansions, n_exp);
    if (n_sel) atomicAdd(&c->select_steps, n_sel);
    if (n_moves) atomicAdd(&c->moves, n_moves);
    if (n_done) atomicAdd(&c->games_done, n_done);
    if (n_look) atomicAdd(&c->cache_lookups, n_look);
    if (n_hit) atomicAdd(&c->cache_hits, n_hit);
  }
  __syncwarp();
  return s.phase;
}

// ---- streaming mode: persistent search kernel ------------------------------------------------------
// Every warp owns the slots g = w, w + W, w + 2W ... (W = warps of the grid) for the whole launch and
// round-robins over them: a slot whose answer has not arrived is skipped, every other busy slot is
// advanced until it blocks again (or its step budget ends).  Nothing here is a barrier between slots, so a
// slot costs one network batch of latency per evaluation instead of one engine-wide tick.  Results do not
// depend on the interleaving: each slot's tree is private and the evaluator is a pure function.
constexpr int STREAM_MAX_OWN = 64;      // slots per warp (65536 slots on 148 SMs x 12 warps need 37)
constexpr unsigned STREAM_WATCHDOG_SPINS = 4u << 20;   // idle polls (~0.5 us each) before a warp gives up

template <class G, bool VL = false>
__global__ void __maxnreg__(BB_STREAM_MAXREG) k_search_stream(EngineDev<G> E) {   // <= 128 registers: one CTA fits an SM NEXT TO the network CTA
  __shared__ WarpScratch s_ws[STREAM_WARPS_PER_CTA];
  __shared__ uint32_t s_ticket[STREAM_WARPS_PER_CTA][STREAM_MAX_OWN];
  __shared__ uint8_t s_state[STREAM_WARPS_PER_CTA][STREAM_MAX_OWN];   // 0 runnable, 1 waiting for an answer, 2 finished
  __shared__ uint32_t s_mbox[VL ? STREAM_WARPS_PER_CTA : 1][VL ? STREAM_MAX_OWN : 1];   // virtual-loss mode: the mailbox (slot x contexts + context) the slot waits on
  const int wl = threadIdx.x >> 5, lane = lane_id();
  const int W = gridDim.x * STREAM_WARPS_PER_CTA;
  const int w = blockIdx.x * STREAM_WARPS_PER_CTA + wl;
  int n_own = 0;
  for (int g = w; g < E.n_games && n_own < STREAM_MAX_OWN; g += W) {
    if (lane == 0) {
      const int ph = E.slots[g].phase;
      // a slot left waiting by an earlier launch keeps its ticket (the answer is in flight or already there)
      s_state[wl][n_own] = (ph == BB_PH_WAIT_EXPAND || ph == BB_PH_WAIT_LEAF) ? 1 : (phase_busy(ph) ? 0 : 2);
      s_ticket[wl][n_own] = E.slots[g].req_idx;
      if (VL) s_mbox[wl][n_own] = (uint32_t)g * (uint32_t)E.vl + E.vl_hdr[g].x;
    }
    ++n_own;
  }
  __syncwarp();
  volatile unsigned* stop = &E.q->stop;
  const volatile unsigned long long* moves = reinterpret_cast<const volatile unsigned long long*>(&E.counters->moves);
  const unsigned long long goal = E.q->goal_moves;
  unsigned idle = 0;
  const long long t_in = clock64();
  long long t_busy = 0;
  unsigned calls = 0;
  for (;;) {
    bool progressed = false, any_busy = false;
    for (int j = 0; j < n_own; ++j) {
      const int st = s_state[wl][j];
      if (st == 2) continue;
      const int g = w + j * W;
      if (st == 1) {
        const unsigned have = *reinterpret_cast<volatile unsigned*>(&E.resp_seq[VL ? s_mbox[wl][j] : (uint32_t)g]);
        if (have != s_ticket[wl][j] + 1u) { any_busy = true; continue; }
        fence_acq_rel();                                     // acquire: the answer was written before its sequence number
      }
      uint32_t ticket = 0;
      const long long t0 = clock64();
      uint32_t mbox = 0;
      const int ph = VL ? advance_slot_vl(E, g, s_ws[wl], &ticket, &mbox) : advance_slot(E, g, 0u, s_ws[wl], &ticket);
      t_busy += clock64() - t0;
      ++calls;
      progressed = true;
      int nst = 2;
      if (ph == BB_PH_WAIT_EXPAND || ph == BB_PH_WAIT_LEAF) nst = 1;
      else if (phase_busy(ph)) nst = 0;
      if (lane == 0) { s_state[wl][j] = (uint8_t)nst; s_ticket[wl][j] = ticket; if (VL) s_mbox[wl][j] = mbox; }
      __syncwarp();
      any_busy |= nst != 2;
    }
    if (!any_busy) break;
    if (goal && lane == 0 && *moves >= goal) *stop = 1u;
    __syncwarp();
    if (*stop) break;
    if (progressed) idle = 0;
    else {
      __nanosleep(400);
      if (++idle > STREAM_WATCHDOG_SPINS) {                  // no answer for seconds: never hang the device, report instead
        if (lane == 0) {
          int j0 = 0;
          while (j0 < n_own && s_state[wl][j0] != 1) ++j0;
          printf("[bbengine] streaming watchdog: warp %d waited too long (slot %d ticket %u, answer seq %u, head %u tail %u)\n", w,
                 w + j0 * W, j0 < n_own ? s_ticket[wl][j0] : 0u, j0 < n_own ? E.resp_seq[VL ? s_mbox[wl][j0] : (uint32_t)(w + j0 * W)] : 0u, E.q->head, E.q->tail);
          *stop = 2u;
        }
        break;
      }
    }
  }
  if (lane == 0) {
    atomicAdd(&E.q->cyc_search_busy, (unsigned long long)t_busy);
    atomicAdd(&E.q->cyc_search_total, (unsigned long long)(clock64() - t_in));
    atomicAdd(&E.q->search_calls, (unsigned long long)calls);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(&E.q->search_done, 1u);
  }
}

// One tick.  `par` = tick parity (host-tracked): requests / activity of this tick accumulate in
// ctl->req[par] / ctl->active[par]; the last CTA to finish clears the other parity for the next tick.
template <class G, bool RO = false>
__global__ void __launch_bounds__(32 * WARPS_PER_CTA, BB_ADV_MIN_CTAS) k_advance(EngineDev<G> E, unsigned par) {
  __shared__ WarpScratch s_ws[WARPS_PER_CTA];
  const int gl = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;   // lane-local slot
  if (gl < E.n_games) advance_slot<G, RO>(E, E.g0 + gl, par, s_ws[threadIdx.x >> 5]);
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned ticket = atomicAdd(&E.ctl->blocks_done, 1u);
    if (ticket == gridDim.x - 1) {
      E.ctl->req[par ^ 1u] = 0;
      E.ctl->active[par ^ 1u] = 0;
      E.ctl->blocks_done = 0;
      E.ctl->tick += 1;
    }
  }
}

template <class G>
__global__ void k_reset(EngineDev<G> E, const typename G::State* init) {
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g == 0) {
    E.ctl->blocks_done = 0;
    E.ctl->req[0] = E.ctl->req[1] = 0;
    E.ctl->active[0] = E.ctl->active[1] = 0;
  }
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
  s.root_plays = 0;
  s.phase = BB_PH_IDLE;
  s.status = BB_OK;
  s.ply = 0;
  s.root_parent_player = 0;
  s.winner = WIN_NONE;
  s.req_idx = 0;
  s.pool_used = 0;
  s.moves_budget = -1;
  s.game_no = 0;
  s.half = 0;
  s.cp_head = s.cp_n = s.cp_src = 0;
  s.row = (uint32_t)g;
  s.rng_pos = 0;
  if (E.vl) E.vl_hdr[g] = make_uint2(0u, 0u);
  for (int c = 0; c < E.vl; ++c) E.ctx[(size_t)g * E.vl + c].st = CTX_IDLE;
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

// Slots that are in the middle of a search keep going (only their move budget is updated).
template <class G>
__global__ void __launch_bounds__(32 * WARPS_PER_CTA) k_begin_move(EngineDev<G> E, int play_limit, int moves_budget) {
  const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (g >= E.n_games) return;
  Slot<G> s = E.slots[g];
  if (s.phase == BB_PH_GAME_OVER || s.phase == BB_PH_ERROR) return;
  s.moves_budget = moves_budget;
  if (!phase_busy(s.phase)) {
    Path path;
    begin_move(E, s, E.half(g, s.half), path, play_limit);
    if (!E.vl) path.store(E.path + (size_t)g * PATH_LEN);    // (virtual-loss mode: path rows belong to contexts, which start their own descents)
  }
  if (lane_id() == 0) E.slots[g] = s;
}

// _selectAction(exploring=False): child index -> action per slot (temp 1: uniforms; temp 0: PUCT)
template <class G>
__global__ void __launch_bounds__(32 * WARPS_PER_CTA) k_choose(EngineDev<G> E, int temp_mode, const double* uniforms, int32_t* actions) {
  const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (g >= E.n_games) return;
  const Slot<G> s = E.slots[g];
  if (s.phase != BB_PH_MOVE_READY || s.root == NONE) { if (lane_id() == 0) actions[g] = -1; return; }
  const Node* nodes = E.half(g, s.half);
  const NodeHdr rh = nodes[s.root].h;
  const int k = meta_nchild(rh.meta);
  if (rh.child_base == NONE || k == 0) { if (lane_id() == 0) actions[g] = -1; return; }
  int j;
  NodeHdr chosen;
  if (temp_mode == 1) { j = sample_child(nodes + rh.child_base, k, uniforms[g]); chosen = nodes[rh.child_base + j].h; }
  else {
    double u;
    j = E.rollout ? puct_select<SelectChunks<G>::N, true>(nodes + rh.child_base, k, E.c, &u, &chosen)
                  : puct_select<SelectChunks<G>::N, false>(nodes + rh.child_base, k, E.c, &u, &chosen);
  }
  if (lane_id() == 0) actions[g] = meta_action(chosen.meta);
}

// state = nextState; MoveRoot(state) by action, with the whole subtree copy done in this launch
template <class G>
__global__ void __launch_bounds__(32 * WARPS_PER_CTA) k_move_root(EngineDev<G> E, const int32_t* actions) {
  __shared__ WarpScratch s_ws[WARPS_PER_CTA];
  const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (g >= E.n_games) return;
  Slot<G> s = E.slots[g];
  if (s.phase != BB_PH_MOVE_READY || actions[g] < 0 || s.root == NONE) return;
  Node* nodes = E.half(g, s.half);
  const NodeHdr rh = nodes[s.root].h;
  const int k = rh.child_base == NONE ? 0 : meta_nchild(rh.meta);
  int j = -1;
  for (int i = 0; i < k; ++i)
    if (meta_action(nodes[rh.child_base + i].h.meta) == actions[g]) { j = i; break; }
  if (j < 0) { s.phase = BB_PH_ERROR; s.status = BB_EILLEGAL; }
  else {
    unsigned long long done = 0;
    play_move(E, s, g, nodes, j, &done);
    while (s.phase == BB_PH_COMPACT) run_compaction(E, s, g, s_ws[threadIdx.x >> 5], 1 << 20);
    if (lane_id() == 0 && done) atomicAdd(&E.counters->games_done, done);
    if (lane_id() == 0) atomicAdd(&E.counters->moves, 1ull);
  }
  if (lane_id() == 0) E.slots[g] = s;
}

// rollout evaluator: rewind every slot's word cursor (a new stream was uploaded) / read the cursors back
template <class G>
__global__ void k_rng_rewind(EngineDev<G> E) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g < E.n_games) E.slots[g].rng_pos = 0;
}
template <class G>
__global__ void k_rng_used(EngineDev<G> E, uint32_t* used) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g < E.n_games) used[g] = E.slots[g].rng_pos;
}

}  // namespace bb
