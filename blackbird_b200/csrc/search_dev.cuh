// Warp-cooperative search steps: PUCT select, expand, backup, final move choice, subtree copy.
// One warp per game slot; all control variables are warp-uniform.  IEEE arithmetic is spelled out
// with _rn intrinsics so nvcc cannot contract to FMA (numpy does not fuse).
#pragma once
#include "tree.cuh"
#ifdef BB_DEBUG
#include <assert.h>
#define BB_ASSERT(c) assert(c)
#else
#define BB_ASSERT(c) ((void)0)
#endif

namespace bb {

__constant__ PairwisePlan c_plan[3];   // indexed by G::ID

constexpr int WARPS_PER_CTA = 4;       // k_advance block = 128 threads
constexpr int STAGE_MAX = 128;         // children staged in shared memory per expansion

struct WarpScratch {                   // per-warp shared memory
  double prior[STAGE_MAX];             // unnormalised priors of the children being created
  uint16_t action[STAGE_MAX];
  uint32_t scan[33];                   // subtree copy: exclusive scan of child counts
  uint32_t obase[32];                  // subtree copy: old child_base per scanned node
  float cpol[16];                      // policy row of an evaluation-cache hit
};

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

// j-th set bit of m (j < popc(m))
__device__ __forceinline__ int nth_bit(uint64_t m, int j) {
  for (int i = 0; i < j; ++i) m &= m - 1;
  return ctz64(m);
}

__device__ __forceinline__ NodeHdr shfl_hdr(const NodeHdr& h, int src) {
  NodeHdr r;
  r.parent = __shfl_sync(0xFFFFFFFFu, h.parent, src);
  r.child_base = __shfl_sync(0xFFFFFFFFu, h.child_base, src);
  r.meta = __shfl_sync(0xFFFFFFFFu, h.meta, src);
  r.eval = __shfl_sync(0xFFFFFFFFu, h.eval, src);
  return r;
}

// ---- a23: _selectAction(exploring=True), MCTS.py:327-334 --------------------------------------
// U_i = W_i + ((c * P_i) * sqrt(1 + sum N)) / (1 + N_i) in f64, W_i = f32(Value)/f32(Plays) widened;
// first maximal index.  Children are stored in ascending action order, so "first" is preserved.
// Lane j loads child j's whole 32-byte record, so the winner's header (its own child block, action,
// cached value) arrives with the same memory round trip.  Returns the child index (warp-uniform);
// *best_u receives U_j and *chosen the winner's header.
// RO (rollout evaluator): priors are the constant 1 and the f64 field holds Node.Value: W_i = Value / Plays in f64.
// pf_nodes != nullptr (narrow games): every lane also asks L2 for ITS child's own child block before the PUCT
// arithmetic starts, so the next level's dependent load -- one DRAM round trip of ~1.3 k cycles, the search's critical
// path (5.4 levels per descent) -- finds the winner's children on their way in whichever child wins.  DRAM is ~2 %
// busy in this kernel; the extra traffic (all k child blocks instead of one) is what that idle bandwidth is spent on.
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
template <int CHUNKS, bool RO = false>
__device__ __forceinline__ int puct_select(const Node* kids, int k, double c, double* best_u, NodeHdr* chosen, const Node* pf_nodes = nullptr) {
  const int lane = lane_id();
  Node r[CHUNKS];
  uint32_t tot = 0;
#pragma unroll
  for (int ci = 0; ci < CHUNKS; ++ci) {
    const int j = ci * 32 + lane;
    if (j < k) {
      r[ci] = kids[j];
      tot += r[ci].s.plays;
    }
  }
  if (CHUNKS == 1 && pf_nodes && lane < k && r[0].h.child_base != NONE) {
    const char* b = reinterpret_cast<const char*>(pf_nodes + r[0].h.child_base);
    const int bytes = meta_nchild(r[0].h.meta) * (int)sizeof(Node);
    prefetch_l2(b);
    if (bytes > 128) prefetch_l2(b + 128);
    if (((reinterpret_cast<uintptr_t>(b) & 127) + bytes) > 256) prefetch_l2(b + bytes - 1);
  }
  for (int j = CHUNKS * 32 + lane; j < k; j += 32) tot += kids[j].s.plays;
  tot = __reduce_add_sync(0xFFFFFFFFu, tot);                 // one REDUX instead of five shuffle rounds
  const double t = __dsqrt_rn(__dadd_rn(1.0, (double)tot));   // exact integer sum (builtin sum of f64 ints)
  double bu = -1.0;
  int bj = 0x7FFFFFFF;
  auto score = [&](const NodeStat& s, int j) {
    double u;
    if (RO) {
      const double wr = s.plays > 0 ? __ddiv_rn(s.prior, (double)s.plays) : 0.0;   // exact Python Value / Plays (A.9)
      u = __dadd_rn(wr, __ddiv_rn(__dmul_rn(c, t), __dadd_rn(1.0, (double)s.plays)));   // c * 1.0 == c
    } else {
      float wr = s.plays > 0 ? __fdiv_rn(s.value, (float)s.plays) : 0.f;          // Node.WinRate, f32 (A.9)
      u = __dadd_rn((double)wr, __ddiv_rn(__dmul_rn(__dmul_rn(c, s.prior), t), __dadd_rn(1.0, (double)s.plays)));
    }
    if (u > bu) { bu = u; bj = j; }
  };
#pragma unroll
  for (int ci = 0; ci < CHUNKS; ++ci) {
    const int j = ci * 32 + lane;
    if (j < k) score(r[ci].s, j);
  }
  for (int j = CHUNKS * 32 + lane; j < k; j += 32) score(kids[j].s, j);
  // Warp argmax by (value, then smallest index) with three REDUX instructions instead of five rounds of 64-bit
  // shuffles and compares: map the double to an unsigned key of the same order (-0.0 folded into +0.0 first, so
  // equal values get equal keys), take the maximum of the high words, then of the low words among the lanes that
  // hold it, then the smallest index among the lanes that hold both.  The winning value is rebuilt from the key.
  {
    const long long b = __double_as_longlong(__dadd_rn(bu, 0.0));
    const unsigned long long key = (unsigned long long)(b ^ ((b >> 63) | (long long)0x8000000000000000ull));
    const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
    const unsigned mhi = __reduce_max_sync(0xFFFFFFFFu, hi);
    const unsigned mlo = __reduce_max_sync(0xFFFFFFFFu, hi == mhi ? lo : 0u);
    bj = (int)__reduce_min_sync(0xFFFFFFFFu, (hi == mhi && lo == mlo) ? (unsigned)bj : 0x7FFFFFFFu);
    const unsigned long long wk = ((unsigned long long)mhi << 32) | mlo;
    bu = __longlong_as_double((long long)((wk >> 63) ? (wk ^ 0x8000000000000000ull) : ~wk));
  }
  *best_u = bu;
  if (bj != 0x7FFFFFFF) {
    const int chunk = bj >> 5;
    if (chunk < CHUNKS) {
      NodeHdr mine = r[0].h;
#pragma unroll
      for (int ci = 1; ci < CHUNKS; ++ci)
        if (ci == chunk) mine = r[ci].h;
      *chosen = shfl_hdr(mine, bj & 31);
    } else {
      *chosen = kids[bj].h;
    }
  }
  return bj;
}

// Rules calls of the descent, in their warp-cooperative / unchecked forms where a game has them (games.cuh: DragonChess).
template <class G>
__device__ __forceinline__ int winner_w(const typename G::State& s, int last) {
  if constexpr (G::NROWS > 1) return G::winner_warp(s, last); else return G::winner(s, last);
}
template <class G>
__device__ __forceinline__ void apply_child(typename G::State& s, int a) {      // a is the action of a tree child: legal by construction
  if constexpr (G::NROWS > 1) G::apply_trusted(s, a); else G::apply(s, a);
}
template <class G>
__device__ __forceinline__ void occ_w(const typename G::State& s, uint64_t* own, uint64_t* foe) {
  if constexpr (G::NROWS > 1) G::occ_warp(s, own, foe); else { *own = 0; *foe = 0; }
}
template <class G>
__device__ __forceinline__ uint64_t legal_row_w(const typename G::State& s, int row, uint64_t own, uint64_t foe) {
  if constexpr (G::NROWS > 1) return G::legal_row_occ(s, row, own, foe); else return G::legal_row(s, row);
}

template <class G>
struct SelectChunks { static constexpr int N = G::A <= 32 ? 1 : 4; };

// ---- a26: priors = f32 policy * f64 mask, / np.sum(...) with numpy's pairwise order (A.11) ------
// One lane walks the k sparse (action, numerator) pairs in ascending action order through the
// host-built plan; zeros at illegal indices are identities and are skipped.  get(j, &action)
// returns the j-th numerator.
template <class G, class Get>
__device__ double pairwise_sum_sparse(int k, Get get) {
  const PairwisePlan& P = c_plan[G::ID];
  constexpr int NL = G::A <= 128 ? 1 : PW_MAX_LEAVES;     // A <= 128 is a single leaf
  double leaf_sum[NL];
  int j = 0;
  int a = 0;
  double v = 0.0;
  if (k > 0) v = get(0, &a);
  for (int L = 0; L < P.n_leaves; ++L) {
    const int lo = P.leaf_lo[L], n = P.leaf_n[L];
    const int nfull = n < 8 ? 0 : n - (n & 7);
    if (!(j < k && a < lo + n)) { leaf_sum[L] = 0.0; continue; }   // no legal action in this leaf: a sum of zeros (most leaves of a 4032-wide row)
    double r[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    double res = 0.0;
    bool combined = false;
    while (j < k && a < lo + n) {
      const int i = a - lo;
      if (i < nfull) {
        r[i & 7] = __dadd_rn(r[i & 7], v);
      } else {
        if (!combined) {
          res = nfull ? __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                                  __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7]))) : 0.0;
          combined = true;
        }
        res = __dadd_rn(res, v);
      }
      ++j;
      if (j < k) v = get(j, &a);
    }
    if (!combined)
      res = nfull ? __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                              __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7]))) : 0.0;
    leaf_sum[L] = res;
  }
  if (NL == 1) return leaf_sum[0];
  double stack[16];
  int sp = 0;
  for (int i = 0; i < P.n_ops; ++i) {
    int op = P.ops[i];
    if (op >= 0) stack[sp++] = leaf_sum[op];
    else { sp--; stack[sp - 1] = __dadd_rn(stack[sp - 1], stack[sp]); }
  }
  return stack[0];
}

// The same sum by the whole warp, for the staged children of a wide game (ws.action / ws.prior, k <= STAGE_MAX): leaf L of
// the plan belongs to lane L % 32, which walks the k pairs and adds the ones in its range in numpy's order (eight
// interleaved accumulators over the leaf's multiple-of-eight head, their fixed tree, then the tail one by one); the leaf
// sums are then combined level by level of the plan's tree (PairwisePlan::add_*), one addition per lane.  Same additions on the
// same operands as pairwise_sum_sparse -- which, on one lane with its accumulators, leaf sums and stack in local memory, took 46 k
// cycles of a 61 k-cycle DragonChess expansion.  Overwrites ws.prior (the caller keeps the numerators in registers).
template <class G>
__device__ double pairwise_sum_warp(int k, WarpScratch& ws) {
  const PairwisePlan& P = c_plan[G::ID];
  const int lane = lane_id();
  double mine[2] = {0.0, 0.0};
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int L = lane + 32 * t;
    if (L < P.n_leaves) {
      const int lo = P.leaf_lo[L], n = P.leaf_n[L];
      const int nfull = n < 8 ? 0 : n - (n & 7);
      double r0 = 0, r1 = 0, r2 = 0, r3 = 0, r4 = 0, r5 = 0, r6 = 0, r7 = 0;
      bool any = false, has_tail = false;
      // pass 1: the head goes to accumulator i & 7; the tail elements are summed in order on top of the combined head below
      for (int j = 0; j < k; ++j) {
        const int i = (int)ws.action[j] - lo;
        if (i < 0 || i >= n) continue;
        any = true;
        if (i < nfull) {
          const double v = ws.prior[j];
          const int c = i & 7;
          r0 = c == 0 ? __dadd_rn(r0, v) : r0; r1 = c == 1 ? __dadd_rn(r1, v) : r1; r2 = c == 2 ? __dadd_rn(r2, v) : r2; r3 = c == 3 ? __dadd_rn(r3, v) : r3;
          r4 = c == 4 ? __dadd_rn(r4, v) : r4; r5 = c == 5 ? __dadd_rn(r5, v) : r5; r6 = c == 6 ? __dadd_rn(r6, v) : r6; r7 = c == 7 ? __dadd_rn(r7, v) : r7;
        } else {
          has_tail = true;
        }
      }
      if (any) {
        double res = nfull ? __dadd_rn(__dadd_rn(__dadd_rn(r0, r1), __dadd_rn(r2, r3)), __dadd_rn(__dadd_rn(r4, r5), __dadd_rn(r6, r7))) : 0.0;
        if (has_tail)
          for (int j = 0; j < k; ++j) {
            const int i = (int)ws.action[j] - lo;
            if (i >= nfull && i < n) res = __dadd_rn(res, ws.prior[j]);
          }
        mine[t] = res;
      }
    }
  }
  __syncwarp();                                               // every lane has read the numerators
#pragma unroll
  for (int t = 0; t < 2; ++t)
    if (lane + 32 * t < P.n_leaves) ws.prior[lane + 32 * t] = mine[t];
  __syncwarp();
  for (int lv = 0; lv < P.n_levels; ++lv) {                   // the plan's additions, one tree level at a time (a + b is commutative: operand order is free)
    for (int q = P.lvl_off[lv] + lane; q < P.lvl_off[lv + 1]; q += 32) ws.prior[P.add_dst[q]] = __dadd_rn(ws.prior[P.add_l[q]], ws.prior[P.add_r[q]]);
    __syncwarp();
  }
  const double total = ws.prior[P.root];
  __syncwarp();
  return total;
}

// ---- evaluation cache lookup (common.cuh): true = hit, policy in ws.cpol[0..A), raw value in *value -----------------------
template <class G>
__device__ __forceinline__ bool ecache_lookup(const EngineDev<G>& E, const typename G::State& st, WarpScratch& ws, float* value) {
  static_assert(sizeof(typename G::State) <= 16, "the key is at most four words");
  uint32_t k[4] = {0u, 0u, 0u, 0u};
  memcpy(k, &st, sizeof(typename G::State));
  uint32_t idx, tag;
  ecache_hash(k, E.ecache_mask, &idx, &tag);
  const uint32_t* ent = E.ecache + (size_t)idx * ECACHE_WORDS;
  const int lane = lane_id();
  const uint32_t w = lane < ECACHE_WORDS ? __ldcg(ent + lane) : 0u;             // one 64-byte transaction, through L2
  const uint32_t want = lane < 4 ? (lane == 0 ? k[0] : lane == 1 ? k[1] : lane == 2 ? k[2] : k[3]) : tag;
  if (__ballot_sync(0xFFFFFFFFu, (lane < 4 || lane == 14) && w != want)) return false;
  fence_acq_rel();                                                              // the second look at the tag comes after the payload reads
  const uint32_t again = lane == 14 ? *reinterpret_cast<const volatile uint32_t*>(ent + 14) : tag;
  if (__ballot_sync(0xFFFFFFFFu, again != tag)) return false;                   // a writer took the entry meanwhile
  if (lane >= 4 && lane < 4 + G::A) ws.cpol[lane - 4] = __uint_as_float(w);
  *value = __uint_as_float(__shfl_sync(0xFFFFFFFFu, w, 13));
  __syncwarp();
  return true;
}

// ---- a22: AddChildren (MCTS.py:122-139), lazily evaluated ---------------------------------------
// Creates the k children of `node` (ascending action id) with Plays 0, Value 0 and
// prior = f64(policy[a]) / sum.  Updates *node_hdr (and its copy in HBM).  Returns k, or -1 when
// the half-arena is full.
// RO: priors are GetPriors' constant ones (MCTS.py:346-358) times the legal mask -- nothing to read or normalise; the
// f64 field starts as Node.Value = 0.
template <class G, bool RO = false>
__device__ int expand_node(const EngineDev<G>& E, Slot<G>& s, Node* nodes, uint32_t node, NodeHdr* node_hdr,
                           const float* policy, WarpScratch& ws, bool policy_in_smem = false) {
  const int lane = lane_id();
  const uint32_t base = s.n_nodes;
  const int child_player = G::next_player(s.cur_state);
  int k;
  if (G::NROWS == 1) {
    // small action spaces: the legal mask is warp-uniform, lane j creates the j-th child
    const uint64_t m = G::legal_row(s.cur_state, 0);
    k = popc64(m);
    if (base + (uint32_t)k > E.node_cap) return -1;
    int a = 0;
    double p = 0.0;
    if (lane < k) {
      a = nth_bit(m, lane);
      if (!RO) {
        p = (double)(policy_in_smem ? policy[a] : __ldcg(policy + a));   // f32 policy * 1.0 (f64 mask); L2 load: written by another SM in streaming mode
        ws.prior[lane] = p;
        ws.action[lane] = (uint16_t)a;
      }
    }
    __syncwarp();
    double sum = 1.0;
    if (!RO) {
      if constexpr (G::A < 8) {
        // fewer than eight actions: numpy's pairwise sum is its plain loop (res = 0.; res += a[i]), zeros at illegal
        // indices are identities -- the k numerators in ascending action order, straight from their lanes' registers
        sum = 0.0;
        for (int j = 0; j < k; ++j) sum = __dadd_rn(sum, __shfl_sync(0xFFFFFFFFu, p, j));
      } else {
        if (lane == 0)
          sum = pairwise_sum_sparse<G>(k, [&](int j, int* act) { *act = ws.action[j]; return ws.prior[j]; });
        sum = __shfl_sync(0xFFFFFFFFu, sum, 0);
      }
    }
    if (lane < k) {
      Node n;
      n.s.plays = 0u;
      n.s.value = 0.f;
      n.s.prior = RO ? 0.0 : __ddiv_rn(p, sum);
      n.h.parent = node;
      n.h.child_base = NONE;
      n.h.meta = meta_pack(a, 0, child_player, false);
      n.h.eval = 0.f;
      nodes[base + lane] = n;
    }
  } else {
    constexpr int RPL = (G::NROWS + 31) / 32;             // action rows per lane
    uint64_t m[RPL];
    int cnt = 0;
    uint64_t own, foe;
#ifdef BB_SEARCH_TIMING
    long long xt0 = clock64(), xt1 = 0, xt2 = 0, xt3 = 0;
#endif
    occ_w<G>(s.cur_state, &own, &foe);                    // occupancy once per position, by the warp
#pragma unroll
    for (int r = 0; r < RPL; ++r) {
      int row = lane * RPL + r;
      m[r] = row < G::NROWS ? legal_row_w<G>(s.cur_state, row, own, foe) : 0ull;
      cnt += popc64(m[r]);
    }
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (lane >= o) incl += v;
    }
    k = __shfl_sync(0xFFFFFFFFu, incl, 31);
    if (base + (uint32_t)k > E.node_cap) return -1;
    int j = incl - cnt;
#ifdef BB_SEARCH_TIMING
    xt1 = clock64();
#endif
    if (k <= STAGE_MAX) {
      // Every lane lists the actions of its rows (ascending, at its scan offset), then the WARP creates the children: child i
      // by lane i % 32, so the policy reads of an expansion are in flight together (the row-by-row loop below pays one L2
      // round trip per move of the busiest lane -- a queen's 20-odd moves one after the other) and a node is written once,
      // with its normalised prior.
#pragma unroll
      for (int r = 0; r < RPL; ++r) {
        uint64_t mm = m[r];
        const int row = lane * RPL + r;
        while (mm) {
          const int bit = ctz64(mm);
          mm &= mm - 1;
          ws.action[j++] = (uint16_t)G::action_of(row, bit);
        }
      }
      __syncwarp();
      constexpr int CPL = STAGE_MAX / 32;                   // children per lane
      double pr[CPL];
#pragma unroll
      for (int q = 0; q < CPL; ++q) {
        const int i = q * 32 + lane;
        pr[q] = 0.0;
        if (!RO && i < k) { pr[q] = (double)__ldcg(policy + ws.action[i]); ws.prior[i] = pr[q]; }
      }
      __syncwarp();
#ifdef BB_SEARCH_TIMING
      xt2 = clock64();
#endif
      double sum = 1.0;
      if (!RO) {
        sum = pairwise_sum_warp<G>(k, ws);
      }
#ifdef BB_SEARCH_TIMING
      xt3 = clock64();
#endif
#pragma unroll
      for (int q = 0; q < CPL; ++q) {
        const int i = q * 32 + lane;
        if (i < k) {
          Node n;
          n.s.plays = 0u;
          n.s.value = 0.f;
          n.s.prior = RO ? 0.0 : __ddiv_rn(pr[q], sum);
          n.h.parent = node;
          n.h.child_base = NONE;
          n.h.meta = meta_pack((int)ws.action[i], 0, child_player, false);
          n.h.eval = 0.f;
          nodes[base + i] = n;
        }
      }
    } else {
#pragma unroll
    for (int r = 0; r < RPL; ++r) {
      uint64_t mm = m[r];
      const int row = lane * RPL + r;
      while (mm) {
        int bit = ctz64(mm);
        mm &= mm - 1;
        int a = G::action_of(row, bit);
        Node n;
        n.s.plays = 0u;
        n.s.value = 0.f;
        n.s.prior = RO ? 0.0 : (double)__ldcg(policy + a);
        n.h.parent = node;
        n.h.child_base = NONE;
        n.h.meta = meta_pack(a, 0, child_player, false);
        n.h.eval = 0.f;
        nodes[base + j] = n;
        ++j;
      }
    }
    __syncwarp();
    if (!RO) {
      double sum = 0.0;
      if (lane == 0)
        sum = pairwise_sum_sparse<G>(k, [&](int q, int* act) { *act = meta_action(nodes[base + q].h.meta); return nodes[base + q].s.prior; });
      sum = __shfl_sync(0xFFFFFFFFu, sum, 0);
      for (int i = lane; i < k; i += 32) nodes[base + i].s.prior = __ddiv_rn(nodes[base + i].s.prior, sum);
    }
    }
#ifdef BB_SEARCH_TIMING
    if (lane == 0 && E.q && xt3) {
      const long long xt4 = clock64();
      atomicAdd(&E.q->xdbg[0], (unsigned long long)(xt1 - xt0)); atomicAdd(&E.q->xdbg[1], (unsigned long long)(xt2 - xt1));
      atomicAdd(&E.q->xdbg[2], (unsigned long long)(xt3 - xt2)); atomicAdd(&E.q->xdbg[3], (unsigned long long)(xt4 - xt3));
      atomicAdd(&E.q->xdbg[4], 1ull);
    }
#endif
  }
  node_hdr->child_base = base;
  node_hdr->meta = (node_hdr->meta & ~(0x3FFu << 12)) | ((uint32_t)k << 12);
  if (lane == 0) nodes[node].h = *node_hdr;
  s.n_nodes = base + (uint32_t)k;
  __syncwarp();
  return k;
}

// ---- descent path: warp-distributed array, entry d = node id | (player to move at it - 1) << 31 --
struct Path {
  uint32_t r[PATH_LEN / 32];           // lane l holds entries l, 32 + l, ...
  __device__ __forceinline__ void set(int d, uint32_t node, int player) {
    if (d < PATH_LEN && lane_id() == (d & 31)) {
#pragma unroll
      for (int c = 0; c < PATH_LEN / 32; ++c)
        if (c == (d >> 5)) r[c] = node | ((uint32_t)(player - 1) << 31);
    }
  }
  __device__ __forceinline__ void load(const uint32_t* row) {
#pragma unroll
    for (int c = 0; c < PATH_LEN / 32; ++c) r[c] = row[c * 32 + lane_id()];
  }
  __device__ __forceinline__ void store(uint32_t* row) const {
#pragma unroll
    for (int c = 0; c < PATH_LEN / 32; ++c) row[c * 32 + lane_id()] = r[c];
  }
};

// ---- a27 + a28: SampleValue transform (Blackbird.py:365-368) and _backProp (MCTS.py:238-258) -----
// Every node on the path gets Plays += 1 and (if it has a parent) Value += v or 1 - v depending on
// whether the PARENT's side to move is playerForValue.  The path is in registers, so all the
// read-modify-writes are independent and issue in one memory round trip (lane d owns depth d).
// Ancestors above Root are never read again in self-play (A.13) and are not kept.
// VO (virtual-loss mode): values only -- the descent has already counted the visit at every node it passed.
template <class G, bool VO = false>
__device__ void backup(const Slot<G>& s, Node* nodes, const Path& path, float eval) {
  const int lane = lane_id();
  float v = __fmul_rn(__fadd_rn(eval, 1.0f), 0.5f);         // (value + 1) * 0.5 in f32
  const int p = G::prev(s.cur_state);                        // playerForValue = PreviousPlayer (0 = None)
  if (G::player(s.cur_state) != p) v = __fsub_rn(1.0f, v);
  const float flip = __fsub_rn(1.0f, v);
  int d = s.depth;
  if (d >= PATH_LEN) {                                       // deeper than the register path: walk parents
    if (lane == 0) {
      uint32_t m = s.cur;
      while (d >= PATH_LEN) {
        const uint32_t parent = nodes[m].h.parent;
        const int pp = meta_player(nodes[parent].h.meta);
        NodeStat st = nodes[m].s;
        if (!VO) nodes[m].s.plays = st.plays + 1;
        nodes[m].s.value = __fadd_rn(st.value, pp == p ? v : flip);
        m = parent;
        --d;
      }
    }
    d = PATH_LEN - 1;
    __syncwarp();
  }
  uint32_t carry = 0;                                        // entry of depth 32*c - 1
#pragma unroll
  for (int c = 0; c < PATH_LEN / 32; ++c) {
    const uint32_t e = path.r[c];
    uint32_t up = __shfl_up_sync(0xFFFFFFFFu, e, 1);
    if (lane == 0) up = carry;
    carry = __shfl_sync(0xFFFFFFFFu, e, 31);
    const int idx = c * 32 + lane;
    if (idx <= d) {
      const int parent_player = idx == 0 ? s.root_parent_player : 1 + (int)(up >> 31);
      BB_ASSERT((e & 0x7FFFFFFFu) < s.n_nodes);
      uint2* q = reinterpret_cast<uint2*>(&nodes[e & 0x7FFFFFFFu].s);
      uint2 w = *q;                                          // (plays, value)
      if (!VO) w.x += 1;
      if (parent_player != 0) w.y = __float_as_uint(__fadd_rn(__uint_as_float(w.y), parent_player == p ? v : flip));
      *q = w;
    }
  }
  __syncwarp();
}

// ---- rollout evaluator: MCTS.SampleValue (MCTS.py:360-383) and its exact-valued _backProp --------------
constexpr int ROLLOUT_MAX_PLIES = 8192;   // DragonChess has no draw and no move cap (A.20); random playouts end after a few hundred plies

// One masked-rejection draw of np.random.choice(actions) from the slot's MT19937 word stream (legacy RandomState.randint:
// no draw at all when a single action is legal).  Returns -1 when the stream is exhausted.
template <class G>
__device__ __forceinline__ int rng_choice(const EngineDev<G>& E, Slot<G>& s, int g, int n) {
  const uint32_t rng = (uint32_t)n - 1u;
  if (rng == 0u) return 0;
  uint32_t mask = rng;
  mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4; mask |= mask >> 8; mask |= mask >> 16;
  for (;;) {
    if (s.rng_pos >= E.rng_cap) return -1;
    const uint32_t v = E.rng_words[(size_t)g * E.rng_cap + s.rng_pos++] & mask;
    if (v <= rng) return (int)v;
  }
}

// -> Python 0 / 0.5 / 1 in halves (0, 1, 2), or -1 on an error (s.status set).  Warp-uniform: every lane plays the same game.
template <class G>
__device__ int rollout_value(const EngineDev<G>& E, Slot<G>& s, int g) {
  typename G::State st = s.cur_state;
  const int player = G::prev(st);                              // SampleValue(leaf.State, leaf.State.PreviousPlayer); 0 = None
  int w = G::winner(st, -1);                                   // rolloutState.Winner()
  for (int plies = 0; w == WIN_NONE; ++plies) {
    int a;
    if (G::NROWS == 1) {
      const uint64_t m = G::legal_row(st, 0);
      const int n = popc64(m);
      if (n == 0) { s.status = BB_EASSERT; return -1; }         // np.random.choice([]) raises in the reference
      const int idx = rng_choice(E, s, g, n);
      if (idx < 0) { s.status = BB_ESTATE; return -1; }
      a = G::action_of(0, nth_bit(m, idx));
    } else {                                                    // wide action spaces: the lanes share the rows
      constexpr int RPL = (G::NROWS + 31) / 32;
      const int lane = lane_id();
      uint64_t m[RPL];
      int cnt = 0;
#pragma unroll
      for (int r = 0; r < RPL; ++r) {
        const int row = lane * RPL + r;
        m[r] = row < G::NROWS ? G::legal_row(st, row) : 0ull;
        cnt += popc64(m[r]);
      }
      int incl = cnt;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += v;
      }
      const int n = __shfl_sync(0xFFFFFFFFu, incl, 31);
      if (n == 0) { s.status = BB_EASSERT; return -1; }
      const int idx = rng_choice(E, s, g, n);
      if (idx < 0) { s.status = BB_ESTATE; return -1; }
      int mine = -1;
      int j = idx - (incl - cnt);                               // position inside this lane's rows
      if (j >= 0 && j < cnt) {
#pragma unroll
        for (int r = 0; r < RPL; ++r) {
          const int c = popc64(m[r]);
          if (j >= 0 && j < c) mine = G::action_of(lane * RPL + r, nth_bit(m[r], j));
          j -= c;
        }
      }
      const unsigned who = __ballot_sync(0xFFFFFFFFu, mine >= 0);
      a = __shfl_sync(0xFFFFFFFFu, mine, __ffs(who) - 1);
    }
    G::apply(st, a);                                            // _applyAction: Copy + ApplyAction
    w = G::winner(st, a);                                       // rolloutState.Winner(action)
    if (plies >= ROLLOUT_MAX_PLIES) { s.status = BB_ESTATE; return -1; }
  }
  if (w == WIN_DRAW || w == WIN_FLOAT0) return 1;               // `winner == 0` holds for int 0 and np.float64(0.0)
  return player == w ? 2 : 0;                                   // int(player == winner); player None -> 0
}

// _backProp with exact values: Plays += 1 and (if the node has a parent) Value += v or 1 - v, Value kept in the f64 field
template <class G>
__device__ void backup_rollout(const Slot<G>& s, Node* nodes, const Path& path, int halves) {
  const int lane = lane_id();
  const double v = 0.5 * (double)halves, flip = 1.0 - v;
  const int p = G::prev(s.cur_state);
  int d = s.depth;
  if (d >= PATH_LEN) {
    if (lane == 0) {
      uint32_t m = s.cur;
      while (d >= PATH_LEN) {
        const uint32_t parent = nodes[m].h.parent;
        const int pp = meta_player(nodes[parent].h.meta);
        nodes[m].s.plays += 1;
        nodes[m].s.prior = __dadd_rn(nodes[m].s.prior, pp == p ? v : flip);
        m = parent;
        --d;
      }
    }
    d = PATH_LEN - 1;
    __syncwarp();
  }
  uint32_t carry = 0;
#pragma unroll
  for (int c = 0; c < PATH_LEN / 32; ++c) {
    const uint32_t e = path.r[c];
    uint32_t up = __shfl_up_sync(0xFFFFFFFFu, e, 1);
    if (lane == 0) up = carry;
    carry = __shfl_sync(0xFFFFFFFFu, e, 31);
    const int idx = c * 32 + lane;
    if (idx <= d) {
      const int parent_player = idx == 0 ? s.root_parent_player : 1 + (int)(up >> 31);
      NodeStat* q = &nodes[e & 0x7FFFFFFFu].s;
      NodeStat w = *q;
      w.plays += 1;
      if (parent_player != 0) w.prior = __dadd_rn(w.prior, parent_player == p ? v : flip);
      *q = w;
    }
  }
  __syncwarp();
}

// ---- a31: _selectAction(exploring=False), MCTS.py:335-338, temp == 1 -----------------------------
// p_i = plays_i / sum(plays); np.random.choice == searchsorted(cumsum(p)/cumsum(p)[-1], u, 'right') (A.15)
__device__ __forceinline__ int sample_child(const Node* kids, int k, double u) {
  int pick = 0;
  if (lane_id() == 0) {
    double tot = 0.0;
    for (int j = 0; j < k; ++j) tot = __dadd_rn(tot, (double)kids[j].s.plays);
    double last = 0.0;
    for (int j = 0; j < k; ++j) last = __dadd_rn(last, __ddiv_rn((double)kids[j].s.plays, tot));
    double run = 0.0;
    pick = k - 1;
    for (int j = 0; j < k; ++j) {
      run = __dadd_rn(run, __ddiv_rn((double)kids[j].s.plays, tot));
      if (__ddiv_rn(run, last) > u) { pick = j; break; }
    }
  }
  return __shfl_sync(0xFFFFFFFFu, pick, 0);
}

// ---- a32: MoveRoot (MCTS.py:260-282) as a breadth-first subtree copy into the other half ---------
// dst[0] = the new root; then, scanning dst in order, every node that has children gets a fresh
// contiguous block at the end of dst and its children are copied there (ascending action order
// is kept, statistics verbatim).  Up to `rounds` scans of 32 nodes per call so one tick's latency
// stays bounded; returns true when the copy is complete.
template <class G>
__device__ bool subtree_copy(const EngineDev<G>& E, Slot<G>& s, const Node* __restrict__ src, Node* __restrict__ dst,
                             WarpScratch& ws, int rounds) {
  const int lane = lane_id();
  if (s.cp_n == 0) {
    if (lane == 0) {
      BB_ASSERT(s.cp_src < s.n_nodes);
      Node r = src[s.cp_src];
      r.h.parent = NONE;
      dst[0] = r;
    }
    s.cp_n = 1;
    s.cp_head = 0;
    __syncwarp();
  }
  for (int it = 0; it < rounds && s.cp_head < s.cp_n; ++it) {
    const uint32_t i = s.cp_head + lane;
    const bool valid = i < s.cp_n;
    NodeHdr h;
    h.child_base = NONE;
    h.meta = 0;
    if (valid) h = dst[i].h;
    const uint32_t k = (valid && h.child_base != NONE) ? (uint32_t)meta_nchild(h.meta) : 0u;
    uint32_t incl = k;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (lane >= o) incl += v;
    }
    const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, 31);
    if (s.cp_n + total > E.node_cap) { s.phase = BB_PH_ERROR; s.status = BB_ENOMEM; return true; }
    ws.scan[lane] = incl - k;
    ws.obase[lane] = h.child_base;
    if (lane == 31) ws.scan[32] = total;
    if (k) {                                                 // re-point this node at its new block
      h.child_base = s.cp_n + (incl - k);
      dst[i].h = h;
    }
    __syncwarp();
    for (uint32_t e = lane; e < total; e += 32) {            // flattened copy: every load independent
      int lo = 0, hi = 31;                                   // last owner with scan[owner] <= e and k > 0
      while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (ws.scan[mid] <= e) lo = mid; else hi = mid - 1;
      }
      const uint32_t off = e - ws.scan[lo];
      BB_ASSERT(ws.obase[lo] + off < s.n_nodes && off < 1024 && s.cp_n + e < E.node_cap);
      Node n = src[ws.obase[lo] + off];
      n.h.parent = s.cp_head + (uint32_t)lo;
      dst[s.cp_n + e] = n;
    }
    __syncwarp();
    s.cp_head = min(s.cp_head + 32u, s.cp_n);                // only nodes that existed at the start of this round
    s.cp_n += total;
  }
  return s.cp_head >= s.cp_n;
}

}  // namespace bb
