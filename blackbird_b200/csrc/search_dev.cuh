// Warp-cooperative search steps: PUCT select, expand, backup, final move choice, root motion.
// One warp per game slot; all control variables are warp-uniform.  IEEE arithmetic is spelled out
// with _rn intrinsics so nvcc cannot contract to FMA (numpy does not fuse).
#pragma once
#include "tree.cuh"

namespace bb {

__constant__ PairwisePlan c_plan[3];   // indexed by G::ID

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

// j-th set bit of m (j < popc(m))
__device__ __forceinline__ int nth_bit(uint64_t m, int j) {
  for (int i = 0; i < j; ++i) m &= m - 1;
  return ctz64(m);
}

// ---- a23: _selectAction(exploring=True), MCTS.py:327-334 --------------------------------------
// U_i = W_i + ((c * P_i) * sqrt(1 + sum N)) / (1 + N_i) in f64, W_i = f32(Value)/f32(Plays) widened;
// first maximal index.  Children are stored in ascending action order, so "first" is preserved.
// Returns child index j (warp-uniform); *best_u receives U_j.
__device__ __forceinline__ int puct_select(const NodeStat* kids, int k, double c, double* best_u) {
  const int lane = lane_id();
  NodeStat first = {0u, 0.f, 0.0};
  uint32_t tot = 0;
  for (int j = lane; j < k; j += 32) {
    NodeStat s = kids[j];
    if (j < 32) first = s;
    tot += s.plays;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) tot += __shfl_xor_sync(0xFFFFFFFFu, tot, o);
  const double t = __dsqrt_rn(__dadd_rn(1.0, (double)tot));   // exact integer sum (builtin sum of f64 ints)
  double bu = -1.0;
  int bj = 0x7FFFFFFF;
  for (int j = lane; j < k; j += 32) {
    NodeStat s = (j < 32) ? first : kids[j];
    float wr = s.plays > 0 ? __fdiv_rn(s.value, (float)s.plays) : 0.f;          // Node.WinRate, f32 (A.9)
    double u = __dadd_rn((double)wr, __ddiv_rn(__dmul_rn(__dmul_rn(c, s.prior), t), __dadd_rn(1.0, (double)s.plays)));
    if (u > bu) { bu = u; bj = j; }
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    double ou = __shfl_xor_sync(0xFFFFFFFFu, bu, o);
    int oj = __shfl_xor_sync(0xFFFFFFFFu, bj, o);
    if (ou > bu || (ou == bu && oj < bj)) { bu = ou; bj = oj; }
  }
  *best_u = bu;
  return bj;
}

// ---- a26: priors = f32 policy * f64 mask, / np.sum(...) with numpy's pairwise order (A.11) ------
// Lane 0 walks the k sparse (action, numerator) pairs in ascending action order through the
// host-built plan; zeros at illegal indices are identities and are skipped.
template <class G>
__device__ double pairwise_sum_sparse(const NodeHdr* khdr, const NodeStat* kstat, int k) {
  const PairwisePlan& P = c_plan[G::ID];
  double leaf_sum[PW_MAX_LEAVES];
  int j = 0;
  for (int L = 0; L < P.n_leaves; ++L) {
    const int lo = P.leaf_lo[L], n = P.leaf_n[L];
    const int nfull = n < 8 ? 0 : n - (n & 7);
    double r[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    double res = 0.0;
    bool combined = false;
    while (j < k) {
      int a = meta_action(khdr[j].meta);
      if (a >= lo + n) break;
      double v = kstat[j].prior;
      int i = a - lo;
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
    }
    if (!combined)
      res = nfull ? __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                              __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7]))) : 0.0;
    leaf_sum[L] = res;
  }
  double stack[16];
  int sp = 0;
  for (int i = 0; i < P.n_ops; ++i) {
    int op = P.ops[i];
    if (op >= 0) stack[sp++] = leaf_sum[op];
    else { sp--; stack[sp - 1] = __dadd_rn(stack[sp - 1], stack[sp]); }
  }
  return stack[0];
}

// ---- a22: AddChildren (MCTS.py:122-139), lazily evaluated ---------------------------------------
// Creates the k children of `node` (ascending action id) with Plays 0, Value 0 and
// prior = f64(policy[a]) / sum.  Returns k, or -1 when the arena is full.
template <class G>
__device__ int expand_node(const EngineDev<G>& E, Slot<G>& s, NodeHdr* hdr, NodeStat* stat, uint32_t node,
                           const float* policy) {
  const int lane = lane_id();
  constexpr int RPL = (G::NROWS + 31) / 32;             // action rows per lane
  uint64_t m[RPL];
  int cnt = 0;
#pragma unroll
  for (int r = 0; r < RPL; ++r) {
    int row = lane * RPL + r;
    m[r] = row < G::NROWS ? G::legal_row(s.cur_state, row) : 0ull;
    cnt += popc64(m[r]);
  }
  int incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if (lane >= o) incl += v;
  }
  const int k = __shfl_sync(0xFFFFFFFFu, incl, 31);
  const uint32_t base = s.n_nodes;
  if (base + (uint32_t)k > E.node_cap) return -1;
  int j = incl - cnt;
#pragma unroll
  for (int r = 0; r < RPL; ++r) {
    uint64_t mm = m[r];
    const int row = lane * RPL + r;
    while (mm) {
      int bit = ctz64(mm);
      mm &= mm - 1;
      int a = G::action_of(row, bit);
      NodeHdr h;
      h.parent = node;
      h.child_base = NONE;
      h.meta = meta_pack(a, 0, G::next_player(s.cur_state), false);
      h.eval = 0.f;
      hdr[base + j] = h;
      NodeStat st;
      st.plays = 0u;
      st.value = 0.f;
      st.prior = (double)policy[a];                       // f32 policy * 1.0 (f64 mask)
      stat[base + j] = st;
      ++j;
    }
  }
  __syncwarp();
  double sum = 0.0;
  if (lane == 0) sum = pairwise_sum_sparse<G>(hdr + base, stat + base, k);
  sum = __shfl_sync(0xFFFFFFFFu, sum, 0);
  for (int i = lane; i < k; i += 32) stat[base + i].prior = __ddiv_rn(stat[base + i].prior, sum);
  if (lane == 0) {
    NodeHdr h = hdr[node];
    h.child_base = base;
    h.meta = (h.meta & ~(0x3FFu << 12)) | ((uint32_t)k << 12);
    hdr[node] = h;
  }
  s.n_nodes = base + (uint32_t)k;
  __syncwarp();
  return k;
}

// ---- a27 + a28: SampleValue transform (Blackbird.py:365-368) and _backProp (MCTS.py:238-258) -----
template <class G>
__device__ void backup(const EngineDev<G>& E, const Slot<G>& s, NodeHdr* hdr, NodeStat* stat, float eval) {
  if (lane_id() == 0) {
    float v = __fmul_rn(__fadd_rn(eval, 1.0f), 0.5f);       // (value + 1) * 0.5 in f32
    const int p = G::prev(s.cur_state);                      // playerForValue = PreviousPlayer (0 = None)
    if (G::player(s.cur_state) != p) v = __fsub_rn(1.0f, v);
    const float flip = __fsub_rn(1.0f, v);
    uint32_t m = s.cur;
    for (;;) {
      NodeStat st = stat[m];
      st.plays += 1;
      int parent_player;
      uint32_t parent = NONE;
      if (m == s.root) parent_player = s.root_parent_player;   // ancestors above Root are never read again (A.13)
      else { parent = hdr[m].parent; parent_player = meta_player(hdr[parent].meta); }
      if (parent_player != 0) st.value = __fadd_rn(st.value, parent_player == p ? v : flip);
      stat[m].plays = st.plays;
      stat[m].value = st.value;
      if (m == s.root) break;
      m = parent;
    }
  }
  __syncwarp();
}

// ---- a31: _selectAction(exploring=False), MCTS.py:335-338, temp == 1 -----------------------------
// p_i = plays_i / sum(plays); np.random.choice == searchsorted(cumsum(p)/cumsum(p)[-1], u, 'right') (A.15)
__device__ __forceinline__ int sample_child(const NodeStat* kids, int k, double u) {
  int pick = 0;
  if (lane_id() == 0) {
    double tot = 0.0;
    for (int j = 0; j < k; ++j) tot = __dadd_rn(tot, (double)kids[j].plays);
    double last = 0.0;
    for (int j = 0; j < k; ++j) last = __dadd_rn(last, __ddiv_rn((double)kids[j].plays, tot));
    double run = 0.0;
    pick = k - 1;
    for (int j = 0; j < k; ++j) {
      run = __dadd_rn(run, __ddiv_rn((double)kids[j].plays, tot));
      if (__ddiv_rn(run, last) > u) { pick = j; break; }
    }
  }
  return __shfl_sync(0xFFFFFFFFu, pick, 0);
}

}  // namespace bb
