// Engine host side: owns the tree pool in HBM, runs the tick loop (advance kernel + evaluator),
// and implements the search / self-play / record entry points of include/bbengine.h.
#include <chrono>
#include <cstdlib>
#include <string.h>
#include <algorithm>
#include <vector>

#include "search_kernels.cuh"
#include "records.cuh"

namespace bb {

static void plan_rec(PairwisePlan& P, int lo, int n) {        // numpy pairwise_sum recursion (A.11)
  if (n <= 128) {
    int id = P.n_leaves++;
    P.leaf_lo[id] = (int16_t)lo;
    P.leaf_n[id] = (int16_t)n;
    P.ops[P.n_ops++] = (int8_t)id;
    return;
  }
  int n2 = n / 2;
  n2 -= n2 % 8;
  plan_rec(P, lo, n2);
  plan_rec(P, lo + n2, n - n2);
  P.ops[P.n_ops++] = -1;
}

static void plan_levels(PairwisePlan& P) {                   // the additions of P.ops grouped by height
  int stack[32], sp = 0, n_int = 0;
  int l[PW_MAX_LEAVES], r[PW_MAX_LEAVES], h[2 * PW_MAX_LEAVES] = {0};
  for (int i = 0; i < P.n_ops; ++i) {
    if (P.ops[i] >= 0) { stack[sp++] = P.ops[i]; continue; }
    const int b = stack[--sp], a = stack[--sp], t = n_int++;
    l[t] = a; r[t] = b;
    h[P.n_leaves + t] = 1 + (h[a] > h[b] ? h[a] : h[b]);
    stack[sp++] = P.n_leaves + t;
  }
  P.root = stack[0];
  int hmax = 0;
  for (int t = 0; t < n_int; ++t) hmax = h[P.n_leaves + t] > hmax ? h[P.n_leaves + t] : hmax;
  P.n_levels = hmax;
  int q = 0;
  for (int lv = 1; lv <= hmax; ++lv) {
    P.lvl_off[lv - 1] = (uint8_t)q;
    for (int t = 0; t < n_int; ++t)
      if (h[P.n_leaves + t] == lv) { P.add_dst[q] = (uint8_t)(P.n_leaves + t); P.add_l[q] = (uint8_t)l[t]; P.add_r[q] = (uint8_t)r[t]; ++q; }
  }
  P.lvl_off[hmax] = (uint8_t)q;
}

template <class G>
__global__ void k_root_stats(EngineDev<G> E, int kmax, int32_t* n_children, int32_t* actions, uint32_t* plays,
                             float* child_value, double* child_prior, uint32_t* root_plays, double* root_value,
                             int32_t* root_has_parent) {
  const int g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (g >= E.n_games) return;
  const Slot<G> s = E.slots[g];
  const int lane = lane_id();
  if (s.root == NONE) { if (lane == 0) { n_children[g] = -1; root_plays[g] = 0; root_value[g] = 0; root_has_parent[g] = 0; } return; }
  const Node* nodes = E.half(g, s.half);
  const Node rn = nodes[s.root];
  const int k = rn.h.child_base == NONE ? -1 : meta_nchild(rn.h.meta);
  if (lane == 0) {
    n_children[g] = k;
    root_plays[g] = rn.s.plays;
    root_value[g] = E.rollout ? rn.s.prior : (double)rn.s.value;      // rollout evaluator: Node.Value lives in the f64 field
    root_has_parent[g] = s.root_parent_player != 0;
  }
  for (int j = lane; j < k && j < kmax; j += 32) {
    const size_t o = (size_t)g * kmax + j;
    const Node kid = nodes[rn.h.child_base + j];
    actions[o] = meta_action(kid.h.meta);
    plays[o] = kid.s.plays;
    child_value[o] = E.rollout ? (float)kid.s.prior : kid.s.value;     // multiples of 0.5 below 2^23: exact in f32
    child_prior[o] = E.rollout ? 1.0 : kid.s.prior;
  }
}

template <class G>
__global__ void k_gather_slots(EngineDev<G> E, typename G::State* states, int32_t* phases, int32_t* status) {
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= E.n_games) return;
  if (states) states[g] = E.slots[g].root_state;
  if (phases) phases[g] = E.slots[g].phase;
  if (status) status[g] = E.slots[g].status;
}

// streaming mode: (re)arm the control block for one launch pair; tickets (head / tail) carry over
__global__ void k_stream_arm(StreamCtl* q, unsigned search_ctas, unsigned servers, unsigned long long goal) {
  q->stop = 0;
  q->head = q->tail;              // tickets claimed past the tail by the previous launch were never issued
  q->search_done = 0;
  q->search_ctas = search_ctas;
  q->servers = servers;
  q->goal_moves = goal;
}

// Do two kernels on two streams really run at the same time?  A waits (up to ~50 ms) for a flag that only B sets.  Under a
// profiler that serialises kernel launches (ncu kernel replay) B starts after A has given up: the persistent
// search / network pair cannot work there (each side would wait for the other until its watchdog fires).
__global__ void k_probe_wait(volatile unsigned* flag, unsigned* seen) {
  const long long t0 = clock64();
  unsigned v = 0;
  while (!(v = *flag) && clock64() - t0 < 100000000ll) __nanosleep(1000);
  *seen = v;
}
__global__ void k_probe_set(volatile unsigned* flag) { *flag = 1u; }

// ---- allocation-free record serialization (bb_records_serialize_static) ---------------------------------------------------
// The per-game byte offsets that k_pack_records reads are scratch memory: bb_records_serialize builds them on the host and
// keeps them in a cudaMalloc'ed buffer.  This variant keeps them in a file-scope device array instead, so the call
// allocates nothing.  Every GPU that loads the library holds the array (512 KB) until the process exits; every call on
// a device shares it, so two calls that overlap on the GPU race.  Nothing in it has to start at zero: the offsets kernel
// rewrites every word a call reads (offsets [0, rows], then the totals) before the pack kernel runs.
constexpr int REC_STATIC_MAX_ROWS = 65536;                     // games (record rows) per call
static __device__ __align__(256) unsigned long long g_rec_static[REC_STATIC_MAX_ROWS + 8];   // [rows + 1] offsets | [MAX + 1] records | [MAX + 2] bytes

// one CTA: exclusive scan of the games' byte sizes -> scratch (and the caller's game_off / rec_bytes, when given)
template <class G>
__global__ void __launch_bounds__(1024) k_record_offsets(EngineDev<G> E, int rows, unsigned long long* __restrict__ scratch,
                                                          unsigned long long* __restrict__ game_off_out, uint32_t* __restrict__ rec_bytes_out) {
  using F = RecFmt<G>;
  __shared__ unsigned long long s_part[1024], s_recs[1024];
  const int tid = threadIdx.x, per = (rows + 1023) / 1024;
  const int lo = min(rows, tid * per), hi = min(rows, lo + per);
  unsigned long long bytes = 0, recs = 0;
  for (int g = lo; g < hi; ++g) {
    const int w = E.winners[g];
    const unsigned sz = (unsigned)F::SIZE0 + ((w == WIN_P1 || w == WIN_P2) ? 5u : 0u);
    bytes += (unsigned long long)(E.n_moves[g] + 1) * sz;
    recs += (unsigned long long)(E.n_moves[g] + 1);
  }
  s_part[tid] = bytes;
  s_recs[tid] = recs;
  __syncthreads();
  if (tid == 0) {                                               // 1024 partial sums: a serial pass is a few microseconds
    unsigned long long run = 0, nrec = 0;
    for (int i = 0; i < 1024; ++i) { const unsigned long long b = s_part[i]; s_part[i] = run; run += b; nrec += s_recs[i]; }
    scratch[rows] = run;
    scratch[REC_STATIC_MAX_ROWS + 1] = nrec;
    scratch[REC_STATIC_MAX_ROWS + 2] = run;
    if (game_off_out) game_off_out[rows] = run;
  }
  __syncthreads();
  unsigned long long off = s_part[tid];
  for (int g = lo; g < hi; ++g) {
    const int w = E.winners[g];
    const unsigned sz = (unsigned)F::SIZE0 + ((w == WIN_P1 || w == WIN_P2) ? 5u : 0u);
    scratch[g] = off;
    if (game_off_out) game_off_out[g] = off;
    if (rec_bytes_out) rec_bytes_out[g] = sz;
    off += (unsigned long long)(E.n_moves[g] + 1) * sz;
  }
}

// device address of g_rec_static on `device` (which must be current): resolved once per device.  In host code the array's
// name is only a stand-in for the symbol, not an address a kernel can use.
static int rec_static_ptr(int device, unsigned long long** out) {
  static unsigned long long* ptrs[64] = {nullptr};
  if (device < 0 || device >= 64) { set_error("bb_records_serialize_static: device %d out of range", device); return BB_EINVAL; }
  if (!ptrs[device]) {
    unsigned long long* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_rec_static);
    if (e != cudaSuccess || !p) { set_error("cudaGetSymbolAddress(g_rec_static) failed on device %d: %s", device, cudaGetErrorString(e)); return BB_ECUDA; }
    ptrs[device] = p;
  }
  *out = ptrs[device];
  return BB_OK;
}

}  // namespace bb

using namespace bb;

struct bb_engine {
  bb_config cfg;
  virtual ~bb_engine() {}
  virtual int net_load(const bb_net_desc*, const float*, size_t) = 0;
  virtual int net_fwd(const void*, int, float*, float*, cudaStream_t) = 0;
  virtual int reset(const void* states_dev) = 0;
  virtual int drop_root() = 0;
  virtual int begin_move(int play_limit) = 0;
  virtual int run(int* pending) = 0;
  virtual int root_stats(int, int32_t*, int32_t*, uint32_t*, float*, double*, uint32_t*, double*, int32_t*) = 0;
  virtual int choose(double, const double*, int32_t*) = 0;
  virtual int move_root(const int32_t*) = 0;
  virtual int states(void*) = 0;
  virtual int phases(int32_t*, int32_t*) = 0;
  virtual int eval_pending(int32_t*, const int32_t**, const void**) = 0;
  virtual int eval_fetch(void*, int32_t*) = 0;
  virtual int eval_submit(const float*, const float*) = 0;
  virtual int selfplay(int, double, const double*) = 0;
  virtual int selfplay_games(int, int, double, const double*) = 0;
  virtual int records_rows() = 0;
  virtual int step(int, int, double, const double*) = 0;
  virtual int stream_moves(long long, int, double, const double*) = 0;
  virtual void* stream_handle() = 0;
  virtual int profile(int, int) = 0;
  virtual int profile_read(double*, double*, int*) = 0;
  virtual int pool_cap() = 0;
  virtual int records_export(int32_t*, int32_t*, void*, uint32_t*, uint16_t*, uint32_t*, uint16_t*) = 0;
  virtual int records_device(void**, void**, void**, void**, void**, void**, void**) = 0;
  virtual int counters(bb_counters*, int) = 0;
  virtual int n_lanes() = 0;
  virtual int streaming() = 0;
  virtual int stream_stats(uint64_t*) = 0;
  virtual int stream_answers(void*, float*, float*, uint32_t*) = 0;
  virtual int records_serialize(void*, uint64_t, uint64_t*, uint64_t*, uint64_t*, uint32_t*) = 0;
  virtual int records_serialize_static(void*, uint64_t, uint64_t*, uint32_t*, uint64_t*, uint64_t*) = 0;
  virtual int set_rng(const uint32_t*, uint32_t) = 0;
  virtual int rng_used(uint32_t*) = 0;
};

template <class G>
struct EngineT : bb_engine {
  using State = typename G::State;
  EngineDev<G> d{};
  NetDev* net = nullptr;
  cudaStream_t stream = nullptr;    // lane 0's stream; whole-engine kernels and copies run here while the lanes idle
  std::vector<void*> allocs;
  double* uniforms_dev = nullptr;
  unsigned* next_row_dev = nullptr;
  size_t rec_cap_rows = 0, rows_last = 0;      // record / uniform rows allocated; rows of the last self-play call
  struct Snap { TickCtl ctl; unsigned long long moves; unsigned par; };
  // A lane is a contiguous range of game slots with its own stream, tick control block, evaluator
  // mailbox and counters.  Lanes tick independently, so one lane's advance kernel (a latency-bound
  // pointer chase whose length is set by its slowest warp) overlaps the other lane's network kernel.
  struct Lane {
    EngineDev<G> d{};
    cudaStream_t stream = nullptr;
    unsigned par = 0;               // parity of the next tick
    Snap* snap = nullptr;           // pinned: two snapshots of the tick control block + move counter
    cudaEvent_t snap_ev[2] = {nullptr, nullptr};
    std::vector<cudaEvent_t> prof_ev;
    size_t prof_used = 0;
  };
  std::vector<Lane> lanes;
  TickCtl* ctl_dev = nullptr;
  Counters* counters_dev = nullptr;
  // streaming mode (built-in tensor-core network only): request ring + persistent kernel pair
  StreamCtl* q_dev = nullptr;
  uint32_t* ecache_dev = nullptr;   // evaluation cache of streaming mode (answers of the CURRENT weights only)
  size_t ecache_entries = 0;
  cudaStream_t net_stream = nullptr;
  cudaEvent_t st_ev[4] = {nullptr, nullptr, nullptr, nullptr};
  double st_search_ms = 0, st_net_ms = 0;
  unsigned long long st_runs = 0, st_batches0 = 0;
  int last_active = 0, last_req = 0;
  void* scratch = nullptr;         // device scratch for exports
  size_t scratch_bytes = 0;
  unsigned long long host_ticks = 0, host_launches = 0;
  bool prof = false;               // per-tick CUDA-event timing of the two kernels (bench only)

  template <class T>
  int alloc(T** p, size_t count) {
    void* q = nullptr;
    cudaError_t e = cudaMalloc(&q, count * sizeof(T));
    if (e != cudaSuccess) { set_error("cudaMalloc of %zu bytes failed: %s", count * sizeof(T), cudaGetErrorString(e)); return BB_ENOMEM; }
    allocs.push_back(q);
    *p = (T*)q;
    return BB_OK;
  }
  int ensure_scratch(size_t bytes) {
    if (bytes <= scratch_bytes) return BB_OK;
    if (scratch) cudaFree(scratch);
    scratch = nullptr;
    scratch_bytes = 0;
    BB_CUDA(cudaMalloc(&scratch, bytes));
    scratch_bytes = bytes;
    return BB_OK;
  }

  int init(const bb_config& c) {
    cfg = c;
    BB_CUDA(cudaSetDevice(c.device));
    int n_lanes = c.lanes;
    if (const char* e = getenv("BB_LANES")) n_lanes = atoi(e);
    if (n_lanes <= 0) n_lanes = c.n_games >= 512 ? 2 : 1;
    if (c.evaluator == BB_EVAL_EXTERNAL) n_lanes = 1;            // the caller serves one mailbox
    if (c.evaluator == BB_EVAL_ROLLOUT) n_lanes = 1;             // no evaluator kernel to overlap with
    if (n_lanes > c.n_games) n_lanes = c.n_games;
    if (n_lanes > 8) n_lanes = 8;
    lanes.resize(n_lanes);
    for (Lane& L : lanes) BB_CUDA(cudaStreamCreateWithFlags(&L.stream, cudaStreamNonBlocking));
    stream = lanes[0].stream;
    PairwisePlan P{};
    plan_rec(P, 0, G::A);
    plan_levels(P);
    BB_CUDA(cudaMemcpyToSymbol(c_plan, &P, sizeof P, sizeof(PairwisePlan) * G::ID));
    const size_t n = c.n_games, cap = c.node_cap;
    const size_t nw = n * (size_t)(c.contexts > 0 ? c.contexts : 1);   // evaluator mailboxes: one per slot, or per (slot, context) in virtual-loss mode
    d.vl = c.contexts;
    d.ctx = nullptr;
    d.n_games = c.n_games;
    d.g0 = 0;
    d.n_total = c.n_games;
    d.node_cap = c.node_cap;
    d.search = c.search;
    d.max_depth = c.max_depth;
    d.max_plies = c.max_plies;
    d.c = c.exploration_rate;
    d.rollout = c.evaluator == BB_EVAL_ROLLOUT ? 1 : 0;
    d.rng_words = nullptr;
    d.rng_cap = 0;
    {
      const char* sb = getenv("BB_STEP_BUDGET");               // tuning knob; results do not depend on it
      d.step_budget = sb ? atoi(sb) : (c.search == BB_FIXED ? c.max_depth + 6 : 24);
      const char* cr = getenv("BB_COMPACT_ROUNDS");
      d.compact_rounds = cr ? atoi(cr) : 8;
      if (d.compact_rounds < 1) d.compact_rounds = 1;
      const char* pf = getenv("BB_PREFETCH");
      d.prefetch = pf ? atoi(pf) : 0;             // measured neutral while the network kernel is the limiter (profiles/r02_summary.md): off by default
    }
    d.pool_cap = (uint32_t)c.max_plies * (uint32_t)(G::A < 64 ? G::A : 64);
    int rc;
#define A_(p, cnt) if ((rc = alloc(&p, cnt)) != BB_OK) return rc
    A_(d.slots, n);
    A_(d.nodes, n * 2 * cap);
    A_(d.path, nw * PATH_LEN);
    d.vl_hdr = nullptr;
    if (c.contexts > 0) { A_(d.ctx, nw); A_(d.vl_hdr, n); }
    A_(ctl_dev, lanes.size());
    A_(d.req_slot, n);
    A_(d.req_state, n);
    A_(d.resp_policy, nw * G::A);
    A_(d.resp_value, nw);
    A_(d.rec_states, n * (c.max_plies + 1));
    A_(d.rec_off, n * (c.max_plies + 2));
    A_(d.pool_actions, n * d.pool_cap);
    A_(d.pool_plays, n * d.pool_cap);
    A_(d.chosen, n * c.max_plies);
    A_(d.n_moves, n);
    A_(d.winners, n);
    A_(counters_dev, lanes.size());
    A_(uniforms_dev, n * c.max_plies);
    A_(next_row_dev, 1);
    rec_cap_rows = rows_last = n;
    d.n_rows = 0;
    d.next_row = next_row_dev;
    {
      size_t qn = RING_MIN_ENTRIES;                                // power of two, >= 4 tickets per slot (see StreamCtl)
      while (qn < 4 * nw) qn <<= 1;
      d.q_mask = (unsigned)(qn - 1);
      A_(q_dev, 1);
      A_(d.q_ent, qn);
      A_(d.q_state, nw);
      A_(d.resp_seq, nw);
      BB_CUDA(cudaMemset(q_dev, 0, sizeof(StreamCtl)));
      BB_CUDA(cudaMemset(d.q_ent, 0, qn * sizeof(unsigned long long)));
      BB_CUDA(cudaMemset(d.resp_seq, 0, nw * sizeof(unsigned)));
      d.q = nullptr;                                               // set only in the copy handed to the streaming kernels
      // evaluation cache (common.cuh): 2^22 entries of 64 bytes unless BB_EVAL_CACHE says otherwise (log2 entries; 0 = off)
      d.ecache = nullptr;
      d.ecache_mask = 0;
      if (G::A <= ECACHE_MAX_ACTIONS && c.evaluator == BB_EVAL_NET) {
        int lg = 22;
        if (const char* ec = getenv("BB_EVAL_CACHE")) lg = atoi(ec);
        if (lg > 28) lg = 28;
        if (lg > 0) {
          A_(ecache_dev, ((size_t)1 << lg) * ECACHE_WORDS);
          ecache_entries = (size_t)1 << lg;
          BB_CUDA(cudaMemset(ecache_dev, 0, ecache_entries * ECACHE_WORDS * sizeof(uint32_t)));
        }
      }
      BB_CUDA(cudaStreamCreateWithFlags(&net_stream, cudaStreamNonBlocking));
      for (int i = 0; i < 4; ++i) BB_CUDA(cudaEventCreate(&st_ev[i]));
    }
#undef A_
    d.ctl = ctl_dev;
    d.counters = counters_dev;
    BB_CUDA(cudaMemset(counters_dev, 0, lanes.size() * sizeof(Counters)));
    BB_CUDA(cudaMemset(d.rec_off, 0, n * (c.max_plies + 2) * sizeof(uint32_t)));
    BB_CUDA(cudaMemset(ctl_dev, 0, lanes.size() * sizeof(TickCtl)));
    d.uniforms = uniforms_dev;
    const int per = (c.n_games + n_lanes - 1) / n_lanes;
    for (int l = 0; l < n_lanes; ++l) {
      Lane& L = lanes[l];
      BB_CUDA(cudaMallocHost(&L.snap, 2 * sizeof(Snap)));
      for (int i = 0; i < 2; ++i) BB_CUDA(cudaEventCreateWithFlags(&L.snap_ev[i], cudaEventDisableTiming));
      L.d = d;
      L.d.g0 = l * per;
      L.d.n_games = std::min(per, c.n_games - l * per);
      L.d.ctl = ctl_dev + l;
      L.d.counters = counters_dev + l;
      L.d.req_slot = d.req_slot + L.d.g0;                          // the lane's own mailbox
      L.d.req_state = d.req_state + L.d.g0;
      L.d.resp_policy = d.resp_policy + (size_t)L.d.g0 * G::A;
      L.d.resp_value = d.resp_value + L.d.g0;
    }
    d.uniforms = uniforms_dev;
    // Kernels that should share an SM must agree on its shared-memory / L1 split: an SM is reconfigured only when
    // it is idle, so a search CTA with the default (small) carve-out could never join an SM that runs a network
    // CTA (which needs ~100 KB and therefore the large carve-out).  Same preference for all three.
    BB_CUDA(cudaFuncSetAttribute(k_search_stream<G>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
    BB_CUDA(cudaFuncSetAttribute(k_search_stream<G, true>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
    BB_CUDA(cudaFuncSetAttribute(k_advance<G>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
    return reset(nullptr);
  }
  ~EngineT() override {
    cudaSetDevice(cfg.device);
    for (Lane& L : lanes) if (L.stream) cudaStreamSynchronize(L.stream);
    for (void* p : allocs) cudaFree(p);
    if (scratch) cudaFree(scratch);
    for (Lane& L : lanes) {
      for (cudaEvent_t e : L.prof_ev) cudaEventDestroy(e);
      if (L.snap) cudaFreeHost(L.snap);
      for (int i = 0; i < 2; ++i) if (L.snap_ev[i]) cudaEventDestroy(L.snap_ev[i]);
    }
    if (net) net_destroy(net);
    for (Lane& L : lanes) if (L.stream) cudaStreamDestroy(L.stream);
    if (net_stream) cudaStreamDestroy(net_stream);
    for (int i = 0; i < 4; ++i) if (st_ev[i]) cudaEventDestroy(st_ev[i]);
  }

  void set_mode(int autoplay, int refill, int play_limit, int temp_mode) {
    d.autoplay = autoplay; d.refill = refill; d.play_limit = play_limit; d.temp_mode = temp_mode;
    for (Lane& L : lanes) { L.d.autoplay = autoplay; L.d.refill = refill; L.d.play_limit = play_limit; L.d.temp_mode = temp_mode; }
  }
  unsigned warp_grid() const { return grid_for((long long)d.n_games * 32, 32 * WARPS_PER_CTA); }

  int net_load(const bb_net_desc* nd, const float* w, size_t nfl) override {
    BB_CUDA(cudaSetDevice(cfg.device));
    stream_fits = -1;
    if (ecache_dev) {                                             // new weights (or noise settings): every memoised answer is stale
      int rc0 = sync_lanes();
      if (rc0) return rc0;
      BB_CUDA(cudaMemset(ecache_dev, 0, ecache_entries * ECACHE_WORDS * sizeof(uint32_t)));
    }
    if (net) { net_destroy(net); net = nullptr; }
    return net_create(G::ID, nd, w, nfl, &net);
  }
  int net_fwd(const void* st, int n, float* pol, float* val, cudaStream_t s) override {
    if (!net) { set_error("bb_net_forward: no network loaded (bb_net_load)"); return BB_ESTATE; }
    return net_forward(net, st, n, nullptr, pol, val, s);
  }
  int reset(const void* states_dev) override {
    BB_CUDA(cudaSetDevice(cfg.device));
    int rc = sync_lanes();
    if (rc) return rc;
    k_reset<G><<<grid_for(d.n_games, 128), 128, 0, stream>>>(d, (const State*)states_dev);
    BB_CUDA(cudaGetLastError());
    BB_CUDA(cudaMemsetAsync(ctl_dev, 0, lanes.size() * sizeof(TickCtl), stream));
    BB_CUDA(cudaMemsetAsync(q_dev, 0, sizeof(StreamCtl), stream));
    BB_CUDA(cudaMemsetAsync(d.q_ent, 0, ((size_t)d.q_mask + 1) * sizeof(unsigned long long), stream));
    BB_CUDA(cudaMemsetAsync(d.resp_seq, 0, (size_t)d.n_games * (size_t)(d.vl > 0 ? d.vl : 1) * sizeof(unsigned), stream));
    for (Lane& L : lanes) L.par = 0;
    BB_CUDA(cudaStreamSynchronize(stream));
    return BB_OK;
  }
  int drop_root() override {
    k_drop_root<G><<<grid_for(d.n_games, 128), 128, 0, stream>>>(d);
    BB_CUDA(cudaGetLastError());
    return BB_OK;
  }
  int begin_move(int play_limit) override { return begin_move_budget(play_limit, -1); }
  int begin_move_budget(int play_limit, int moves_budget) {
    if (play_limit <= 0) { set_error("play_limit must be > 0"); return BB_EINVAL; }
    k_begin_move<G><<<warp_grid(), 32 * WARPS_PER_CTA, 0, stream>>>(d, play_limit, moves_budget);
    BB_CUDA(cudaGetLastError());
    return BB_OK;
  }

  // one tick of one lane = advance kernel + evaluator kernel; request counters alternate by tick parity
  int tick(Lane& L) {
    if (d.vl > 0) { set_error("virtual-loss mode (bb_config.contexts) runs in streaming mode only: load a bf16 network of the template width (bb_net_load, precision 1)"); return BB_EINVAL; }
    const bool timed = prof && L.prof_used + 3 <= L.prof_ev.size();
    if (timed) cudaEventRecord(L.prof_ev[L.prof_used], L.stream);
    if (cfg.evaluator == BB_EVAL_ROLLOUT)
      k_advance<G, true><<<grid_for((long long)L.d.n_games * 32, 32 * WARPS_PER_CTA), 32 * WARPS_PER_CTA, 0, L.stream>>>(L.d, L.par);
    else
      k_advance<G><<<grid_for((long long)L.d.n_games * 32, 32 * WARPS_PER_CTA), 32 * WARPS_PER_CTA, 0, L.stream>>>(L.d, L.par);
    BB_CUDA(cudaGetLastError());
    if (timed) cudaEventRecord(L.prof_ev[L.prof_used + 1], L.stream);
    host_ticks++;
    host_launches++;
    const int* count = &L.d.ctl->req[L.par];
    if (cfg.evaluator == BB_EVAL_KAT) {
      int rc = launch_kat(G::ID, L.d.req_state, L.d.n_games, count, L.d.resp_policy, L.d.resp_value, L.stream);
      if (rc) return rc;
      host_launches++;
    } else if (cfg.evaluator == BB_EVAL_NET) {
      if (!net) { set_error("evaluator is BB_EVAL_NET but no network is loaded"); return BB_ESTATE; }
      int rc = net_forward(net, L.d.req_state, L.d.n_games, count, L.d.resp_policy, L.d.resp_value, L.stream, 1);   // the search reads legal entries only
      if (rc) return rc;
      host_launches++;
    }
    if (timed) { cudaEventRecord(L.prof_ev[L.prof_used + 2], L.stream); L.prof_used += 3; }
    L.par ^= 1u;
    return BB_OK;
  }
  int profile(int enable, int max_ticks) override {
    if (use_stream()) {                                         // nothing to arm: the launch pair is always timed
      StreamCtl h{};
      BB_CUDA(cudaMemcpy(&h, q_dev, sizeof h, cudaMemcpyDeviceToHost));
      st_batches0 = h.batches;
      st_search_ms = st_net_ms = 0;
      st_runs = 0;
      return BB_OK;
    }
    prof = enable != 0;
    for (Lane& L : lanes) {
      L.prof_used = 0;
      while (prof && L.prof_ev.size() < (size_t)max_ticks * 3) {
        cudaEvent_t e;
        BB_CUDA(cudaEventCreate(&e));
        L.prof_ev.push_back(e);
      }
    }
    return BB_OK;
  }
  int profile_read(double* adv_ms, double* eval_ms, int* ticks) override {
    if (use_stream()) {                                         // streaming mode: the two persistent kernels, launches as "ticks" -> batches served
      StreamCtl h{};
      BB_CUDA(cudaMemcpy(&h, q_dev, sizeof h, cudaMemcpyDeviceToHost));
      if (adv_ms) *adv_ms = st_search_ms;
      if (eval_ms) *eval_ms = st_net_ms;
      if (ticks) *ticks = (int)(h.batches - st_batches0);
      return BB_OK;
    }
    double a = 0, b = 0;
    size_t n = 0;
    for (Lane& L : lanes) {
      BB_CUDA(cudaStreamSynchronize(L.stream));
      for (size_t i = 0; i + 2 < L.prof_used; i += 3) {
        float t0 = 0, t1 = 0;
        cudaEventElapsedTime(&t0, L.prof_ev[i], L.prof_ev[i + 1]);
        cudaEventElapsedTime(&t1, L.prof_ev[i + 1], L.prof_ev[i + 2]);
        a += t0;
        b += t1;
      }
      n += L.prof_used / 3;
    }
    if (adv_ms) *adv_ms = a;
    if (eval_ms) *eval_ms = b;
    if (ticks) *ticks = (int)n;
    return BB_OK;
  }
  // snapshot of a lane's control block after the ticks enqueued so far (async; wait on snap_ev[i])
  int post_snapshot(Lane& L, int i) {
    L.snap[i].par = L.par ^ 1u;                                // parity of the last enqueued tick
    BB_CUDA(cudaMemcpyAsync(&L.snap[i].ctl, L.d.ctl, sizeof(TickCtl), cudaMemcpyDeviceToHost, L.stream));
    BB_CUDA(cudaMemcpyAsync(&L.snap[i].moves, &L.d.counters->moves, sizeof(unsigned long long), cudaMemcpyDeviceToHost, L.stream));
    BB_CUDA(cudaEventRecord(L.snap_ev[i], L.stream));
    return BB_OK;
  }
  // wait for snapshot i of every lane; totals over lanes
  int wait_snapshots(int i, unsigned long long* moves) {
    last_active = last_req = 0;
    unsigned long long m = 0;
    for (Lane& L : lanes) {
      BB_CUDA(cudaEventSynchronize(L.snap_ev[i]));
      last_active += L.snap[i].ctl.active[L.snap[i].par];
      last_req += L.snap[i].ctl.req[L.snap[i].par];
      m += L.snap[i].moves;
    }
    if (moves) *moves = m;
    return BB_OK;
  }
  int burst_all(int burst, int snap_i) {
    int rc;
    for (int t = 0; t < burst; ++t)
      for (Lane& L : lanes) if ((rc = tick(L))) return rc;         // interleaved so the streams fill evenly
    for (Lane& L : lanes) if ((rc = post_snapshot(L, snap_i))) return rc;
    return BB_OK;
  }
  // Tick all lanes until done(active slots, moves played) holds.  Bursts are double-buffered: the
  // next burst is already queued while the host waits for the previous snapshot, so the GPU never
  // idles on the host round trip (the price is one surplus burst of near-empty ticks at the end).
  template <class Done>
  int tick_until(Done done) {
    const int burst = 16;
    int rc, i = 0;
    if ((rc = burst_all(burst, 0))) return rc;
    for (;;) {
      if ((rc = burst_all(burst, i ^ 1))) return rc;
      unsigned long long moves = 0;
      if ((rc = wait_snapshots(i, &moves))) return rc;
      if (done(last_active, moves)) break;
      i ^= 1;
    }
    return wait_snapshots(i ^ 1, nullptr);
  }
  int sync_lanes() {
    for (Lane& L : lanes) BB_CUDA(cudaStreamSynchronize(L.stream));
    return BB_OK;
  }
  // Streaming mode is used whenever the evaluator is the built-in tensor-core network (BB_STREAM=0 falls back
  // to ticks).  Results are identical in both modes (tests/test_gpu_network.py).
  bool stack_sized = false;         // run_stream has reserved the search kernel's local-memory frame
  int stream_fits = -1;             // the two persistent kernels must share an SM, or the search would spin on a network that never starts
  bool use_stream() {
    if (cfg.evaluator != BB_EVAL_NET || !net || !net_is_tc(net)) return false;
    const char* e = getenv("BB_STREAM");
    if (e && atoi(e) == 0) return false;
    // every slot must be owned by a warp of ONE co-resident wave of search CTAs (a second wave would start only
    // after the first has left, and its slots would sit idle for whole games)
    if ((long long)d.n_games > (long long)net_sm_count(net) * STREAM_WARPS_PER_CTA * STREAM_MAX_OWN) return false;
    if (stream_fits < 0) {
      int nregs = 0, nsmem = 0;
      cudaFuncAttributes fa{};
      stream_fits = 0;
      if (net_tc_footprint(net, &nregs, &nsmem) == BB_OK &&
          (d.vl > 0 ? cudaFuncGetAttributes(&fa, k_search_stream<G, true>) : cudaFuncGetAttributes(&fa, k_search_stream<G>)) == cudaSuccess) {
        int sregs = ((fa.numRegs * 32 + 255) / 256 * 256) * ((STREAM_WARPS_PER_CTA + 3) / 4);   // per register-file quarter
        if (getenv("BB_STREAM_FORCE")) sregs = 0;                 // experiments only: a wrong guess ends in the watchdog, not a hang
        stream_fits = (nregs + sregs <= 16384 && nsmem + (int)fa.sharedSizeBytes + 1024 <= 227 * 1024) ? 1 : 0;
        if (getenv("BB_DEBUG_OCC"))
          fprintf(stderr, "[bbengine] streaming: network CTA %d regs per SM quarter %d B smem, search CTA %d regs per quarter %zu B smem -> %s\n", nregs, nsmem, sregs,
                  fa.sharedSizeBytes, stream_fits ? "co-resident" : "do NOT fit one SM, tick mode");
      }
    }
    return stream_fits == 1;
  }
  // One launch pair: the persistent network kernel (one CTA per SM) and the persistent search kernel
  // (one CTA of STREAM_WARPS_PER_CTA warps per SM; both fit an SM's register file together) run until every slot has left the
  // busy phases, or until the engine-wide move counter reaches goal_moves.
  int run_stream(unsigned long long goal_moves) {
    int rc = sync_lanes();
    if (rc) return rc;
    EngineDev<G> e = d;
    // streaming mode has no tick to keep short: a slot goes on through cache hits for several simulations before it yields
    // (measured: +1 - 3 % over the tick-mode budget; results do not depend on it)
    if (!getenv("BB_STEP_BUDGET")) e.step_budget = d.search == BB_FIXED ? 4 * d.max_depth + 8 : 48;
    e.q = q_dev;
    e.ecache = ecache_dev;
    e.ecache_mask = ecache_dev ? (uint32_t)(ecache_entries - 1) : 0u;
    const int sms = net_sm_count(net);
    int grid = (d.n_games + STREAM_WARPS_PER_CTA - 1) / STREAM_WARPS_PER_CTA;
    if (grid > sms) grid = sms;
    if (const char* g = getenv("BB_STREAM_CTAS")) grid = atoi(g) > 0 ? atoi(g) : grid;     // tuning knob
    int div = sms;                                                                          // a claim takes open / div tickets (1 .. NP); div = pipelines serving
    if (const char* dv = getenv("BB_STREAM_DIV")) div = atoi(dv) > 0 ? atoi(dv) : div;
    if (!stack_sized) {
      // A launch whose per-thread local memory exceeds what the context has reserved makes the driver grow the reservation,
      // which waits for every running kernel -- here the persistent network kernel, which in turn waits for the search kernel:
      // the pair would sit until the watchdog.  Reserve the search kernel's frame while nothing runs.
      cudaFuncAttributes fa{};
      BB_CUDA(d.vl > 0 ? cudaFuncGetAttributes(&fa, k_search_stream<G, true>) : cudaFuncGetAttributes(&fa, k_search_stream<G>));
      size_t cur = 0;
      BB_CUDA(cudaDeviceGetLimit(&cur, cudaLimitStackSize));
      if (fa.localSizeBytes > cur) BB_CUDA(cudaDeviceSetLimit(cudaLimitStackSize, (fa.localSizeBytes + 255) & ~(size_t)255));
      stack_sized = true;
    }
    k_stream_arm<<<1, 1, 0, stream>>>(q_dev, (unsigned)grid, (unsigned)div, goal_moves);
    BB_CUDA(cudaGetLastError());
    BB_CUDA(cudaStreamSynchronize(stream));
    NetServe sv{q_dev, d.q_ent, d.q_state, d.resp_seq, d.q_mask, e.ecache, e.ecache_mask, 1};
    // the network kernel goes first so that its CTAs are resident before the search CTAs spin on it
    BB_CUDA(cudaEventRecord(st_ev[2], net_stream));
    if ((rc = net_serve(net, sv, d.resp_policy, d.resp_value, net_stream))) return rc;
    BB_CUDA(cudaEventRecord(st_ev[3], net_stream));
    BB_CUDA(cudaEventRecord(st_ev[0], stream));
    if (d.vl > 0) k_search_stream<G, true><<<grid, 32 * STREAM_WARPS_PER_CTA, 0, stream>>>(e);
    else k_search_stream<G><<<grid, 32 * STREAM_WARPS_PER_CTA, 0, stream>>>(e);
    BB_CUDA(cudaGetLastError());
    BB_CUDA(cudaEventRecord(st_ev[1], stream));
    host_launches += 3;
    BB_CUDA(cudaStreamSynchronize(stream));
    BB_CUDA(cudaStreamSynchronize(net_stream));
    float a = 0, b = 0;
    cudaEventElapsedTime(&a, st_ev[0], st_ev[1]);
    cudaEventElapsedTime(&b, st_ev[2], st_ev[3]);
    st_search_ms += a;
    st_net_ms += b;
    st_runs += 1;
    StreamCtl h{};
    BB_CUDA(cudaMemcpy(&h, q_dev, sizeof h, cudaMemcpyDeviceToHost));
    if (getenv("BB_DEBUG_STREAM") && h.batches)
      fprintf(stderr, "[bbengine] streaming: %llu batches; per batch: wait %.0f, tower %.0f, heads+publish %.0f cycles; search warps busy %.3f, %.0f cycles per advance call\n", h.batches,
              (double)h.cyc_claim / h.batches, (double)h.cyc_tower / h.batches, (double)h.cyc_heads / h.batches,
              (double)h.cyc_search_busy / (double)(h.cyc_search_total ? h.cyc_search_total : 1), (double)h.cyc_search_busy / (double)(h.search_calls ? h.search_calls : 1));
#ifdef BB_NET_TIMING_TOWER
    if (getenv("BB_DEBUG_STREAM") && h.batches)
      fprintf(stderr, "[bbengine] network pipeline per batch: MMA warp waits for ready %.0f, issues %.0f; epilogue warp waits for full %.0f, works %.0f; tile-layers %.1f (per tile-layer: wait %.0f issue %.0f | wait %.0f work %.0f)\n",
              (double)h.ndbg[0] / h.batches, (double)h.ndbg[1] / h.batches, (double)h.ndbg[2] / h.batches, (double)h.ndbg[3] / h.batches, (double)h.ndbg[5] / h.batches,
              (double)h.ndbg[0] / h.ndbg[5], (double)h.ndbg[1] / h.ndbg[5], (double)h.ndbg[2] / h.ndbg[5], (double)h.ndbg[3] / h.ndbg[5]);
    if (getenv("BB_DEBUG_STREAM") && h.batches)
      fprintf(stderr, "[bbengine] epilogue per tile-layer: tcgen05.ld + wait %.0f, bias/skip/relu/pack/store %.0f, fences + arrive %.0f\n", (double)h.ndbg[4] / h.ndbg[5],
              (double)h.ndbg[6] / h.ndbg[5], ((double)h.ndbg[3] - h.ndbg[4] - h.ndbg[6]) / h.ndbg[5]);
#endif
#ifdef BB_NET_TIMING_HEADS
    if (getenv("BB_DEBUG_STREAM") && h.hdbg[4])
      fprintf(stderr, "[bbengine] legal-only policy head per position: moves + list %.0f, logits %.0f, soft-max + noise %.0f, stores %.0f cycles\n",
              (double)h.hdbg[0] / h.hdbg[4], (double)h.hdbg[1] / h.hdbg[4], (double)h.hdbg[2] / h.hdbg[4], (double)h.hdbg[3] / h.hdbg[4]);
    if (getenv("BB_DEBUG_STREAM") && h.batches)
      fprintf(stderr, "[bbengine] heads phase per batch: sums + sync %.0f, heads_finish %.0f, fence %.0f, sync %.0f, publish + last sync %.0f\n", (double)h.ndbg[0] / h.batches,
              (double)h.ndbg[1] / h.batches, (double)h.ndbg[2] / h.batches, (double)h.ndbg[3] / h.batches, (double)h.ndbg[7] / h.batches);
#endif
#ifdef BB_SEARCH_TIMING
    if (getenv("BB_DEBUG_STREAM") && h.xdbg[4])
      fprintf(stderr, "[bbengine] expand_node (wide games) per call: legal rows + scan %.0f, list + policy reads %.0f, prior sum %.0f, node writes %.0f cycles\n",
              (double)h.xdbg[0] / h.xdbg[4], (double)h.xdbg[1] / h.xdbg[4], (double)h.xdbg[2] / h.xdbg[4], (double)h.xdbg[3] / h.xdbg[4]);
    if (getenv("BB_DEBUG_STREAM") && h.search_calls) {
      const char* nm[12] = {"load", "post_request", "expand", "select level", "backup", "(unused)", "top -> post (winner test, cache miss)", "-> expand (cache hit / answer)",
                            "-> select", "leaf -> post (cache miss)", "leaf -> backup", "exit"};
      for (int i = 0; i < 12; ++i)
        if (h.sdbg[12 + i])
          fprintf(stderr, "[bbengine] advance_slot %-38s %8.0f cycles per call (%.2f per call, %.0f cycles each)\n", nm[i], (double)h.sdbg[i] / h.search_calls,
                  (double)h.sdbg[12 + i] / h.search_calls, (double)h.sdbg[i] / h.sdbg[12 + i]);
    }
#endif
    if (h.stop == 2u) { set_error("streaming watchdog fired: a search warp or the network kernel waited for seconds (see stderr)"); return BB_ESTATE; }
    return BB_OK;
  }
  int run(int* pending) override {
    BB_CUDA(cudaSetDevice(cfg.device));
    int rc;
    BB_CUDA(cudaStreamSynchronize(stream));                     // whole-engine kernels (begin_move ...) before the lanes tick
    if (cfg.evaluator == BB_EVAL_EXTERNAL) {
      do {                                                       // a tick can end on the step budget with nothing to ask
        if ((rc = tick(lanes[0]))) return rc;
        if ((rc = post_snapshot(lanes[0], 0))) return rc;
        if ((rc = wait_snapshots(0, nullptr))) return rc;
      } while (last_req == 0 && last_active > 0);
      if (pending) *pending = last_req;
      return BB_OK;
    }
    if (use_stream()) rc = run_stream(0);
    else rc = tick_until([&](int active, unsigned long long) { return active == 0; });
    if (pending) *pending = 0;
    return rc;
  }

  int root_stats(int kmax, int32_t* nch, int32_t* act, uint32_t* pl, float* cv, double* cp, uint32_t* rp, double* rv,
                 int32_t* hp) override {
    const size_t n = d.n_games, nk = n * kmax;
    const size_t bytes = n * (4 + 4 + 8 + 4) + nk * (4 + 4 + 4 + 8) + 64;
    int rc = ensure_scratch(bytes);
    if (rc) return rc;
    char* p = (char*)scratch;
    double* d_cp = (double*)p; p += nk * 8;
    double* d_rv = (double*)p; p += n * 8;
    int32_t* d_act = (int32_t*)p; p += nk * 4;
    uint32_t* d_pl = (uint32_t*)p; p += nk * 4;
    float* d_cv = (float*)p; p += nk * 4;
    int32_t* d_nch = (int32_t*)p; p += n * 4;
    uint32_t* d_rp = (uint32_t*)p; p += n * 4;
    int32_t* d_hp = (int32_t*)p; p += n * 4;
    BB_CUDA(cudaMemsetAsync(scratch, 0, bytes, stream));
    k_root_stats<G><<<warp_grid(), 32 * WARPS_PER_CTA, 0, stream>>>(d, kmax, d_nch, d_act, d_pl, d_cv, d_cp, d_rp, d_rv, d_hp);
    BB_CUDA(cudaGetLastError());
#define D2H(dst, src, b) if (dst) BB_CUDA(cudaMemcpyAsync(dst, src, b, cudaMemcpyDeviceToHost, stream))
    D2H(nch, d_nch, n * 4); D2H(act, d_act, nk * 4); D2H(pl, d_pl, nk * 4); D2H(cv, d_cv, nk * 4);
    D2H(cp, d_cp, nk * 8); D2H(rp, d_rp, n * 4); D2H(rv, d_rv, n * 8); D2H(hp, d_hp, n * 4);
    BB_CUDA(cudaStreamSynchronize(stream));
    return BB_OK;
  }
  int choose(double temp, const double* u_host, int32_t* actions_host) override {
    if (temp != 1.0 && temp != 0.0) { set_error("device move choice supports temp 1 (sampled) and 0 (PUCT argmax); got %g", temp); return BB_EINVAL; }
    const size_t n = d.n_games;
    int rc = ensure_scratch(n * 16);
    if (rc) return rc;
    double* d_u = (double*)scratch;
    int32_t* d_a = (int32_t*)((char*)scratch + n * 8);
    if (temp == 1.0) {
      if (!u_host) { set_error("temp 1 needs one uniform per slot"); return BB_EINVAL; }
      BB_CUDA(cudaMemcpyAsync(d_u, u_host, n * 8, cudaMemcpyHostToDevice, stream));
    }
    k_choose<G><<<warp_grid(), 32 * WARPS_PER_CTA, 0, stream>>>(d, temp == 1.0 ? 1 : 0, d_u, d_a);
    BB_CUDA(cudaGetLastError());
    BB_CUDA(cudaMemcpyAsync(actions_host, d_a, n * 4, cudaMemcpyDeviceToHost, stream));
    BB_CUDA(cudaStreamSynchronize(stream));
    return BB_OK;
  }
  int move_root(const int32_t* actions_host) override {
    const size_t n = d.n_games;
    int rc = ensure_scratch(n * 4);
    if (rc) return rc;
    BB_CUDA(cudaMemcpyAsync(scratch, actions_host, n * 4, cudaMemcpyHostToDevice, stream));
    EngineDev<G> e = d;
    e.autoplay = 0;
    k_move_root<G><<<warp_grid(), 32 * WARPS_PER_CTA, 0, stream>>>(e, (const int32_t*)scratch);
    BB_CUDA(cudaGetLastError());
    BB_CUDA(cudaStreamSynchronize(stream));
    return BB_OK;
  }
  int states(void* out) override {
    const size_t n = d.n_games;
    int rc = ensure_scratch(n * sizeof(State));
    if (rc) return rc;
    k_gather_slots<G><<<grid_for(n, 128), 128, 0, stream>>>(d, (State*)scratch, nullptr, nullptr);
    BB_CUDA(cudaGetLastError());
    BB_CUDA(cudaMemcpyAsync(out, scratch, n * sizeof(State), cudaMemcpyDeviceToHost, stream));
    BB_CUDA(cudaStreamSynchronize(stream));
    return BB_OK;
  }
  int phases(int32_t* ph, int32_t* st) override {
    const size_t n = d.n_games;
    int rc = ensure_scratch(n * 8);
    if (rc) return rc;
    int32_t* a = (int32_t*)scratch;
    int32_t* b = a + n;
    k_gather_slots<G><<<grid_for(n, 128), 128, 0, stream>>>(d, nullptr, a, b);
    BB_CUDA(cudaGetLastError());
    if (ph) BB_CUDA(cudaMemcpyAsync(ph, a, n * 4, cudaMemcpyDeviceToHost, stream));
    if (st) BB_CUDA(cudaMemcpyAsync(st, b, n * 4, cudaMemcpyDeviceToHost, stream));
    BB_CUDA(cudaStreamSynchronize(stream));
    return BB_OK;
  }
  int eval_pending(int32_t* n_host, const int32_t** slots, const void** sts) override {
    *n_host = last_req;
    if (slots) *slots = d.req_slot;
    if (sts) *sts = d.req_state;
    return BB_OK;
  }
  int eval_fetch(void* sts, int32_t* slots) override {
    int n = last_req;
    if (n <= 0) return BB_OK;
    if (sts) BB_CUDA(cudaMemcpyAsync(sts, d.req_state, (size_t)n * sizeof(State), cudaMemcpyDeviceToDevice, stream));
    if (slots) BB_CUDA(cudaMemcpyAsync(slots, d.req_slot, (size_t)n * 4, cudaMemcpyDeviceToDevice, stream));
    BB_CUDA(cudaStreamSynchronize(stream));
    return BB_OK;
  }
  int eval_submit(const float* pol, const float* val) override {
    int n = last_req;
    if (n <= 0) return BB_OK;
    BB_CUDA(cudaStreamSynchronize(stream));
    BB_CUDA(cudaMemcpyAsync(d.resp_policy, pol, (size_t)n * G::A * 4, cudaMemcpyDeviceToDevice, stream));
    BB_CUDA(cudaMemcpyAsync(d.resp_value, val, (size_t)n * 4, cudaMemcpyDeviceToDevice, stream));
    return BB_OK;
  }

  int selfplay(int play_limit, double temp, const double* u_host) override { return selfplay_games(d.n_games, play_limit, temp, u_host); }
  // record and uniform buffers for `rows` games (they grow, never shrink; the old ones are freed with the engine)
  int ensure_rows(size_t rows) {
    if (rows <= rec_cap_rows) return BB_OK;
    int rc = sync_lanes();
    if (rc) return rc;
    const size_t mp = d.max_plies;
#define A_(p, cnt) if ((rc = alloc(&p, cnt)) != BB_OK) return rc
    A_(d.rec_states, rows * (mp + 1));
    A_(d.rec_off, rows * (mp + 2));
    A_(d.pool_actions, rows * d.pool_cap);
    A_(d.pool_plays, rows * d.pool_cap);
    A_(d.chosen, rows * mp);
    A_(d.n_moves, rows);
    A_(d.winners, rows);
    A_(uniforms_dev, rows * mp);
#undef A_
    BB_CUDA(cudaMemset(d.rec_off, 0, rows * (mp + 2) * sizeof(uint32_t)));
    d.uniforms = uniforms_dev;
    for (Lane& L : lanes) {
      L.d.rec_states = d.rec_states; L.d.rec_off = d.rec_off; L.d.pool_actions = d.pool_actions; L.d.pool_plays = d.pool_plays;
      L.d.chosen = d.chosen; L.d.n_moves = d.n_moves; L.d.winners = d.winners; L.d.uniforms = uniforms_dev;
    }
    rec_cap_rows = rows;
    return BB_OK;
  }
  // n_rows whole games on the engine's n_games slots (n_rows >= n_games): a slot that ends a game takes the next
  // unplayed row, so the GPU stays full until the last games run out.  Game r draws its moves from uniforms row r
  // and its records land in row r, whichever slot plays it.
  int selfplay_games(int n_rows, int play_limit, double temp, const double* u_host) override {
    if (temp != 1.0 && temp != 0.0) { set_error("bb_selfplay_run supports temp 1 and 0 on the device; got %g", temp); return BB_EINVAL; }
    if (cfg.evaluator == BB_EVAL_EXTERNAL) { set_error("bb_selfplay_run needs a device evaluator"); return BB_EINVAL; }
    if (n_rows < d.n_games) { set_error("bb_selfplay_games: n_rows %d < n_games %d (create a smaller engine)", n_rows, d.n_games); return BB_EINVAL; }
    BB_CUDA(cudaSetDevice(cfg.device));
    int rc = ensure_rows((size_t)n_rows);
    if (rc) return rc;
    if (temp == 1.0) {
      if (!u_host) { set_error("temp 1 needs uniforms[n_rows][max_plies]"); return BB_EINVAL; }
      BB_CUDA(cudaMemcpyAsync(uniforms_dev, u_host, (size_t)n_rows * d.max_plies * 8, cudaMemcpyHostToDevice, stream));
    }
    rc = reset(nullptr);
    if (rc) return rc;
    const int multi = n_rows > d.n_games ? n_rows : 0;
    const unsigned first_free = (unsigned)d.n_games;
    BB_CUDA(cudaMemcpyAsync(next_row_dev, &first_free, sizeof first_free, cudaMemcpyHostToDevice, stream));
    BB_CUDA(cudaStreamSynchronize(stream));
    d.n_rows = multi;
    for (Lane& L : lanes) L.d.n_rows = multi;
    set_mode(1, 0, play_limit, temp == 1.0 ? 1 : 0);
    rc = begin_move(play_limit);
    if (!rc) rc = run(nullptr);
    set_mode(0, 0, play_limit, 0);
    d.n_rows = 0;
    for (Lane& L : lanes) L.d.n_rows = 0;
    rows_last = (size_t)n_rows;
    return rc;
  }
  int records_rows() override { return (int)rows_last; }
  // steady-state stepping: every slot plays n_moves moves (finished games restart in place)
  int step(int n_moves, int play_limit, double temp, const double* u_host) override {
    if (temp != 1.0 && temp != 0.0) { set_error("bb_selfplay_step supports temp 1 and 0; got %g", temp); return BB_EINVAL; }
    if (cfg.evaluator == BB_EVAL_EXTERNAL) { set_error("bb_selfplay_step needs a device evaluator"); return BB_EINVAL; }
    if (n_moves <= 0) { set_error("n_moves must be > 0"); return BB_EINVAL; }
    BB_CUDA(cudaSetDevice(cfg.device));
    if (u_host) BB_CUDA(cudaMemcpyAsync(uniforms_dev, u_host, (size_t)d.n_games * d.max_plies * 8, cudaMemcpyHostToDevice, stream));
    set_mode(1, 1, play_limit, temp == 1.0 ? 1 : 0);
    int rc = begin_move_budget(play_limit, n_moves);
    if (!rc) rc = run(nullptr);
    set_mode(0, 0, play_limit, 0);
    return rc;
  }
  // steady state without a per-move barrier: slots keep playing (and restarting) until the engine
  // as a whole has played target_moves further moves; slots are left mid-search
  int stream_moves(long long target_moves, int play_limit, double temp, const double* u_host) override {
    if (temp != 1.0 && temp != 0.0) { set_error("bb_selfplay_stream supports temp 1 and 0; got %g", temp); return BB_EINVAL; }
    if (cfg.evaluator == BB_EVAL_EXTERNAL) { set_error("bb_selfplay_stream needs a device evaluator"); return BB_EINVAL; }
    if (target_moves <= 0) { set_error("target_moves must be > 0"); return BB_EINVAL; }
    BB_CUDA(cudaSetDevice(cfg.device));
    if (u_host) BB_CUDA(cudaMemcpyAsync(uniforms_dev, u_host, (size_t)d.n_games * d.max_plies * 8, cudaMemcpyHostToDevice, stream));
    set_mode(1, 1, play_limit, temp == 1.0 ? 1 : 0);
    int rc = begin_move_budget(play_limit, -1);
    if (!rc) rc = sync_lanes();
    unsigned long long start = 0;
    for (Lane& L : lanes) if (!rc) rc = post_snapshot(L, 0);
    if (!rc) rc = wait_snapshots(0, &start);
    if (!rc) {
      const unsigned long long goal = start + (unsigned long long)target_moves;
      if (use_stream()) rc = run_stream(goal);
      else rc = tick_until([&](int active, unsigned long long moves) { return moves >= goal || active == 0; });
    }
    set_mode(0, 0, play_limit, 0);
    return rc;
  }
  void* stream_handle() override { return (void*)stream; }
  int pool_cap() override { return (int)d.pool_cap; }
  int records_export(int32_t* nm, int32_t* win, void* st, uint32_t* off, uint16_t* pa, uint32_t* pp, uint16_t* ch) override {
    const size_t n = rows_last, mp = d.max_plies;
    D2H(nm, d.n_moves, n * 4); D2H(win, d.winners, n * 4); D2H(st, d.rec_states, n * (mp + 1) * sizeof(State));
    D2H(off, d.rec_off, n * (mp + 2) * 4); D2H(pa, d.pool_actions, n * d.pool_cap * 2);
    D2H(pp, d.pool_plays, n * d.pool_cap * 4); D2H(ch, d.chosen, n * mp * 2);
    BB_CUDA(cudaStreamSynchronize(stream));
    return BB_OK;
  }
#undef D2H
  int records_device(void** nm, void** win, void** st, void** off, void** pa, void** pp, void** ch) override {
    if (nm) *nm = d.n_moves;
    if (win) *win = d.winners;
    if (st) *st = d.rec_states;
    if (off) *off = d.rec_off;
    if (pa) *pa = d.pool_actions;
    if (pp) *pp = d.pool_plays;
    if (ch) *ch = d.chosen;
    return BB_OK;
  }
  // rollout evaluator: one stream of raw 32-bit MT19937 outputs per slot (words_host [n_games][words_per_slot]); cursors rewound
  uint32_t* rng_dev = nullptr;
  size_t rng_alloc = 0;
  int set_rng(const uint32_t* words_host, uint32_t words_per_slot) override {
    if (cfg.evaluator != BB_EVAL_ROLLOUT) { set_error("bb_search_set_rng: the engine's evaluator is not BB_EVAL_ROLLOUT"); return BB_EINVAL; }
    if (!words_host || words_per_slot == 0) { set_error("bb_search_set_rng: empty stream"); return BB_EINVAL; }
    BB_CUDA(cudaSetDevice(cfg.device));
    int rc = sync_lanes();
    if (rc) return rc;
    const size_t need = (size_t)d.n_games * words_per_slot;
    if (need > rng_alloc) {
      uint32_t* p = nullptr;
      if ((rc = alloc(&p, need))) return rc;
      rng_dev = p;
      rng_alloc = need;
    }
    BB_CUDA(cudaMemcpyAsync(rng_dev, words_host, need * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
    d.rng_words = rng_dev;
    d.rng_cap = words_per_slot;
    for (Lane& L : lanes) { L.d.rng_words = rng_dev; L.d.rng_cap = words_per_slot; }
    k_rng_rewind<G><<<grid_for(d.n_games, 128), 128, 0, stream>>>(d);
    BB_CUDA(cudaGetLastError());
    BB_CUDA(cudaStreamSynchronize(stream));
    return BB_OK;
  }
  int rng_used(uint32_t* used_host) override {
    if (cfg.evaluator != BB_EVAL_ROLLOUT) { set_error("bb_search_rng_used: the engine's evaluator is not BB_EVAL_ROLLOUT"); return BB_EINVAL; }
    BB_CUDA(cudaSetDevice(cfg.device));
    int rc = sync_lanes();
    if (rc) return rc;
    if ((rc = ensure_scratch((size_t)d.n_games * 4))) return rc;
    k_rng_used<G><<<grid_for(d.n_games, 128), 128, 0, stream>>>(d, (uint32_t*)scratch);
    BB_CUDA(cudaGetLastError());
    BB_CUDA(cudaMemcpyAsync(used_host, scratch, (size_t)d.n_games * 4, cudaMemcpyDeviceToHost, stream));
    BB_CUDA(cudaStreamSynchronize(stream));
    return BB_OK;
  }
  // ExampleState.SerializeState + labels for every record of the last self-play call, on the device (records.cuh)
  int records_serialize(void* out_dev, uint64_t cap, uint64_t* n_records, uint64_t* n_bytes, uint64_t* game_off_host, uint32_t* rec_bytes_host) override {
    using F = RecFmt<G>;
    BB_CUDA(cudaSetDevice(cfg.device));
    int rc = sync_lanes();
    if (rc) return rc;
    const size_t rows = rows_last;
    std::vector<int32_t> nm(rows), win(rows);
    BB_CUDA(cudaMemcpy(nm.data(), d.n_moves, rows * 4, cudaMemcpyDeviceToHost));
    BB_CUDA(cudaMemcpy(win.data(), d.winners, rows * 4, cudaMemcpyDeviceToHost));
    std::vector<unsigned long long> off(rows + 1);
    unsigned long long bytes = 0, recs = 0;
    for (size_t g = 0; g < rows; ++g) {
      const uint32_t sz = (uint32_t)F::SIZE0 + ((win[g] == WIN_P1 || win[g] == WIN_P2) ? 5u : 0u);
      off[g] = bytes;
      bytes += (unsigned long long)(nm[g] + 1) * sz;
      recs += (unsigned long long)(nm[g] + 1);
      if (rec_bytes_host) rec_bytes_host[g] = sz;
    }
    off[rows] = bytes;
    if (n_records) *n_records = recs;
    if (n_bytes) *n_bytes = bytes;
    if (game_off_host) memcpy(game_off_host, off.data(), (rows + 1) * sizeof(unsigned long long));
    if (!out_dev) return BB_OK;                                // size query
    if (cap < bytes) { set_error("bb_records_serialize: buffer of %llu bytes, %llu needed", (unsigned long long)cap, bytes); return BB_EINVAL; }
    if (rows == 0) return BB_OK;
    if ((rc = ensure_scratch((rows + 1) * sizeof(unsigned long long)))) return rc;
    BB_CUDA(cudaMemcpyAsync(scratch, off.data(), (rows + 1) * sizeof(unsigned long long), cudaMemcpyHostToDevice, stream));
    const size_t smem = pack_records_smem<G>();
    k_pack_records<G><<<(unsigned)rows, 128, smem, stream>>>(d, (int)rows, (const unsigned long long*)scratch, (uint8_t*)out_dev);
    BB_CUDA(cudaGetLastError());
    BB_CUDA(cudaStreamSynchronize(stream));
    host_launches++;
    return BB_OK;
  }
  // the same serialization without any allocation: offsets in the file-scope device array g_rec_static (see there)
  int records_serialize_static(void* out_dev, uint64_t cap, uint64_t* game_off_dev, uint32_t* rec_bytes_dev, uint64_t* n_records, uint64_t* n_bytes) override {
    if (!out_dev) { set_error("bb_records_serialize_static: null output buffer"); return BB_EINVAL; }
    const size_t rows = rows_last;
    if (rows > (size_t)REC_STATIC_MAX_ROWS) {
      set_error("bb_records_serialize_static: %zu game rows, the static scratch array holds at most %d (use bb_records_serialize)", rows, REC_STATIC_MAX_ROWS);
      return BB_EINVAL;
    }
    BB_CUDA(cudaSetDevice(cfg.device));
    unsigned long long* scratch_static = nullptr;
    int rc = rec_static_ptr(cfg.device, &scratch_static);
    if (rc) return rc;
    if ((rc = sync_lanes())) return rc;
    k_record_offsets<G><<<1, 1024, 0, stream>>>(d, (int)rows, scratch_static, (unsigned long long*)game_off_dev, rec_bytes_dev);
    BB_CUDA(cudaGetLastError());
    unsigned long long totals[2] = {0, 0};                      // records, bytes
    BB_CUDA(cudaMemcpyAsync(totals, scratch_static + REC_STATIC_MAX_ROWS + 1, sizeof totals, cudaMemcpyDeviceToHost, stream));
    BB_CUDA(cudaStreamSynchronize(stream));
    if (n_records) *n_records = totals[0];
    if (n_bytes) *n_bytes = totals[1];
    if (cap < totals[1]) { set_error("bb_records_serialize_static: buffer of %llu bytes, %llu needed", (unsigned long long)cap, totals[1]); return BB_EINVAL; }
    if (rows == 0) return BB_OK;
    k_pack_records<G><<<(unsigned)rows, 128, pack_records_smem<G>(), stream>>>(d, (int)rows, scratch_static, (uint8_t*)out_dev);
    BB_CUDA(cudaGetLastError());
    BB_CUDA(cudaStreamSynchronize(stream));
    host_launches += 2;
    return BB_OK;
  }
  int counters(bb_counters* out, int do_reset) override {
    std::vector<Counters> c(lanes.size());
    int rc = sync_lanes();
    if (rc) return rc;
    BB_CUDA(cudaMemcpy(c.data(), counters_dev, c.size() * sizeof(Counters), cudaMemcpyDeviceToHost));
    memset(out, 0, sizeof *out);
    for (const Counters& x : c) {
      out->sims += x.sims; out->evals += x.evals; out->nodes += x.nodes; out->expansions += x.expansions;
      out->select_steps += x.select_steps; out->moves += x.moves; out->games_done += x.games_done;
      out->cache_lookups += x.cache_lookups; out->cache_hits += x.cache_hits;
    }
    out->ticks = host_ticks / lanes.size(); out->kernel_launches = host_launches;
    if (do_reset) { BB_CUDA(cudaMemset(counters_dev, 0, c.size() * sizeof(Counters))); host_ticks = host_launches = 0; }
    return BB_OK;
  }
  // streaming mode keeps, per slot, the last state it asked about and the network's answer to it
  int stream_answers(void* st, float* pol, float* val, uint32_t* seq) override {
    if (!use_stream()) { set_error("bb_stream_last_answers: the engine is not in streaming mode"); return BB_ESTATE; }
    int rc = sync_lanes();
    if (rc) return rc;
    const size_t n = d.n_games;
    if (st) BB_CUDA(cudaMemcpy(st, d.q_state, n * sizeof(State), cudaMemcpyDeviceToHost));
    if (pol) BB_CUDA(cudaMemcpy(pol, d.resp_policy, n * G::A * sizeof(float), cudaMemcpyDeviceToHost));
    if (val) BB_CUDA(cudaMemcpy(val, d.resp_value, n * sizeof(float), cudaMemcpyDeviceToHost));
    if (seq) BB_CUDA(cudaMemcpy(seq, d.resp_seq, n * sizeof(unsigned), cudaMemcpyDeviceToHost));
    return BB_OK;
  }
  int n_lanes() override { return (int)lanes.size(); }
  int streaming() override { return use_stream() ? 1 : 0; }
  int stream_stats(uint64_t* out) override {
    for (int i = 0; i < 8; ++i) out[i] = 0;
    if (!use_stream() || !q_dev) return BB_OK;
    StreamCtl h{};
    BB_CUDA(cudaMemcpy(&h, q_dev, sizeof h, cudaMemcpyDeviceToHost));
    out[0] = h.batches; out[1] = h.cyc_claim; out[2] = h.cyc_tower; out[3] = h.cyc_heads;
    out[4] = (unsigned long long)net_pipes(net) * (unsigned long long)net_sm_count(net);
    out[5] = h.cyc_search_busy; out[6] = h.cyc_search_total; out[7] = h.search_calls;
    return BB_OK;
  }
};

// ------------------------------------------------------------------------------------------
// C-ABI (include/bbengine.h)
// ------------------------------------------------------------------------------------------
#define BB_NEED(e) if (!(e)) { bb::set_error("null engine handle"); return BB_EINVAL; }

extern "C" {

int bb_engine_create(const bb_config* cfg, bb_engine** out) {
  if (!cfg || !out) { set_error("bb_engine_create: null argument"); return BB_EINVAL; }
  *out = nullptr;
  if (cfg->n_games <= 0 || cfg->node_cap < 2 || cfg->max_plies <= 0) { set_error("n_games, node_cap, max_plies must be positive"); return BB_EINVAL; }
  if (cfg->search != BB_FIXED && cfg->search != BB_DYNAMIC) { set_error("unknown search kind %d", cfg->search); return BB_EINVAL; }
  if (cfg->search == BB_FIXED && cfg->max_depth <= 0) { set_error("MaxDepth for MCTS must be > 0."); return BB_EINVAL; }   // FixedMCTS.py:15-16
  if (cfg->evaluator < BB_EVAL_KAT || cfg->evaluator > BB_EVAL_ROLLOUT) { set_error("unknown evaluator %d", cfg->evaluator); return BB_EINVAL; }
  if ((cfg->value_f64 != 0) != (cfg->evaluator == BB_EVAL_ROLLOUT)) {
    set_error("value_f64 goes with BB_EVAL_ROLLOUT and only with it: rollout values are exact Python numbers (f64), network values accumulate in f32 (SURVEY.md A.9)");
    return BB_EINVAL;
  }
  if (cfg->contexts < 0 || cfg->contexts > 64) { set_error("contexts must be in [0, 64] (got %d)", cfg->contexts); return BB_EINVAL; }
  if (cfg->contexts > 0 && cfg->evaluator != BB_EVAL_NET) { set_error("virtual-loss mode (contexts > 0) needs the built-in network evaluator (BB_EVAL_NET)"); return BB_EINVAL; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { set_error("no CUDA device: the engine has no CPU fallback"); return BB_ECUDA; }
  if (cfg->device < 0 || cfg->device >= ndev) { set_error("device %d out of range (%d devices)", cfg->device, ndev); return BB_EINVAL; }
  int rc = BB_EINVAL;
  BB_DISPATCH_GAME(cfg->game, {
    auto* e = new EngineT<G>();
    rc = e->init(*cfg);
    if (rc != BB_OK) { delete e; return rc; }
    *out = e;
  });
  return rc;
}
int bb_engine_destroy(bb_engine* e) { delete e; return BB_OK; }

int bb_net_load(bb_engine* e, const bb_net_desc* d, const float* w, size_t n) { BB_NEED(e); return e->net_load(d, w, n); }
int bb_net_forward(bb_engine* e, const void* st, int n, float* pol, float* val, void* stream) {
  BB_NEED(e);
  return e->net_fwd(st, n, pol, val, (cudaStream_t)stream);
}
int bb_search_reset(bb_engine* e, const void* states_dev) { BB_NEED(e); return e->reset(states_dev); }
int bb_search_drop_root(bb_engine* e) { BB_NEED(e); return e->drop_root(); }
int bb_search_begin_move(bb_engine* e, int play_limit) { BB_NEED(e); return e->begin_move(play_limit); }
int bb_search_run(bb_engine* e, int* pending) { BB_NEED(e); return e->run(pending); }
int bb_search_root_stats(bb_engine* e, int kmax, int32_t* nch, int32_t* act, uint32_t* pl, float* cv, double* cp,
                         uint32_t* rp, double* rv, int32_t* hp) {
  BB_NEED(e);
  return e->root_stats(kmax, nch, act, pl, cv, cp, rp, rv, hp);
}
int bb_search_choose(bb_engine* e, double temp, const double* u, int32_t* a) { BB_NEED(e); return e->choose(temp, u, a); }
int bb_search_move_root(bb_engine* e, const int32_t* a) { BB_NEED(e); return e->move_root(a); }
int bb_search_set_rng(bb_engine* e, const uint32_t* words, uint32_t n) { BB_NEED(e); return e->set_rng(words, n); }
int bb_search_rng_used(bb_engine* e, uint32_t* used) { BB_NEED(e); return e->rng_used(used); }
int bb_search_states(bb_engine* e, void* out) { BB_NEED(e); return e->states(out); }
int bb_search_phases(bb_engine* e, int32_t* ph, int32_t* st) { BB_NEED(e); return e->phases(ph, st); }
int bb_eval_pending(bb_engine* e, int32_t* n, const int32_t** slots, const void** st) { BB_NEED(e); return e->eval_pending(n, slots, st); }
int bb_eval_fetch(bb_engine* e, void* st, int32_t* slots) { BB_NEED(e); return e->eval_fetch(st, slots); }
int bb_eval_submit(bb_engine* e, const float* pol, const float* val) { BB_NEED(e); return e->eval_submit(pol, val); }
int bb_selfplay_run(bb_engine* e, int play_limit, double temp, const double* u) { BB_NEED(e); return e->selfplay(play_limit, temp, u); }
int bb_selfplay_games(bb_engine* e, int n_rows, int play_limit, double temp, const double* u) { BB_NEED(e); return e->selfplay_games(n_rows, play_limit, temp, u); }
int bb_records_rows(bb_engine* e) { BB_NEED(e); return e->records_rows(); }
int bb_selfplay_step(bb_engine* e, int n_moves, int play_limit, double temp, const double* u) {
  BB_NEED(e);
  return e->step(n_moves, play_limit, temp, u);
}
int bb_selfplay_stream(bb_engine* e, long long target_moves, int play_limit, double temp, const double* u) {
  BB_NEED(e);
  return e->stream_moves(target_moves, play_limit, temp, u);
}
int bb_profile_enable(bb_engine* e, int enable, int max_ticks) { BB_NEED(e); return e->profile(enable, max_ticks); }
int bb_profile_read(bb_engine* e, double* adv_ms, double* eval_ms, int* ticks) { BB_NEED(e); return e->profile_read(adv_ms, eval_ms, ticks); }
void* bb_engine_stream(bb_engine* e) { return e ? e->stream_handle() : nullptr; }
int bb_records_pool_cap(bb_engine* e) { BB_NEED(e); return e->pool_cap(); }
int bb_records_export(bb_engine* e, int32_t* nm, int32_t* win, void* st, uint32_t* off, uint16_t* pa, uint32_t* pp, uint16_t* ch) {
  BB_NEED(e);
  return e->records_export(nm, win, st, off, pa, pp, ch);
}
int bb_records_device(bb_engine* e, void** nm, void** win, void** st, void** off, void** pa, void** pp, void** ch) {
  BB_NEED(e);
  return e->records_device(nm, win, st, off, pa, pp, ch);
}
int bb_records_serialize(bb_engine* e, void* out_dev, uint64_t cap, uint64_t* n_records, uint64_t* n_bytes, uint64_t* game_off_host, uint32_t* rec_bytes_host) {
  BB_NEED(e);
  return e->records_serialize(out_dev, cap, n_records, n_bytes, game_off_host, rec_bytes_host);
}
int bb_records_serialize_static(bb_engine* e, void* out_dev, uint64_t cap, uint64_t* game_off_dev, uint32_t* rec_bytes_dev, uint64_t* n_records, uint64_t* n_bytes) {
  BB_NEED(e);
  return e->records_serialize_static(out_dev, cap, game_off_dev, rec_bytes_dev, n_records, n_bytes);
}
int bb_records_static_max_rows(void) { return REC_STATIC_MAX_ROWS; }
int bb_counters_read(bb_engine* e, bb_counters* out, int reset) { BB_NEED(e); return e->counters(out, reset); }
int bb_engine_lanes(bb_engine* e) { BB_NEED(e); return e->n_lanes(); }
int bb_engine_streaming(bb_engine* e) { BB_NEED(e); return e->streaming(); }
int bb_stream_stats(bb_engine* e, uint64_t out[8]) { BB_NEED(e); if (!out) return BB_EINVAL; return e->stream_stats(out); }
int bb_stream_last_answers(bb_engine* e, void* st, float* pol, float* val, uint32_t* seq) { BB_NEED(e); return e->stream_answers(st, pol, val, seq); }
int bb_debug_kernels_concurrent(int device) {
  if (cudaSetDevice(device) != cudaSuccess) { set_error("cudaSetDevice(%d) failed", device); return BB_ECUDA; }
  unsigned* buf = nullptr;
  cudaStream_t a = nullptr, b = nullptr;
  BB_CUDA(cudaMalloc(&buf, 2 * sizeof(unsigned)));
  BB_CUDA(cudaMemset(buf, 0, 2 * sizeof(unsigned)));
  BB_CUDA(cudaStreamCreateWithFlags(&a, cudaStreamNonBlocking));
  BB_CUDA(cudaStreamCreateWithFlags(&b, cudaStreamNonBlocking));
  k_probe_wait<<<1, 1, 0, a>>>(buf, buf + 1);
  k_probe_set<<<1, 1, 0, b>>>(buf);
  unsigned seen = 0;
  cudaError_t e1 = cudaStreamSynchronize(a), e2 = cudaStreamSynchronize(b);
  if (e1 == cudaSuccess && e2 == cudaSuccess) e1 = cudaMemcpy(&seen, buf + 1, sizeof seen, cudaMemcpyDeviceToHost);
  cudaStreamDestroy(a);
  cudaStreamDestroy(b);
  cudaFree(buf);
  if (e1 != cudaSuccess || e2 != cudaSuccess) { set_error("concurrency probe failed: %s", cudaGetErrorString(e1 != cudaSuccess ? e1 : e2)); return BB_ECUDA; }
  return seen ? 1 : 0;
}

}  // extern "C"
