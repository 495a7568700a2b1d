// Policy-value network forward (NetworkFactory.py:22-183 as run by Network.getPolicy /
// getEvaluation, Network.py:48-64), fp32 SIMT version: the whole tower for a tile of positions
// stays in shared memory (weights resident, activations ping-pong), input planes are decoded
// from the packed game state in the kernel, BN is folded into the convolutions on the host, the
// heads (1x1 conv -> BN -> ReLU -> per-cell dense -> sum over H,W) are reduced in a fixed order
// so results are run-to-run deterministic.
//
// Persistent grid (one CTA per SM), 256 threads, 2 output cells x 16 filters per thread.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "common.cuh"
#include "noise.cuh"

namespace bb {

constexpr int NET_F = 16;          // tower width this kernel is specialised for (template shape)
constexpr int NET_THREADS = 256;
constexpr int NET_MAX_DENSE = 64;

template <class G> struct NetTile;
template <> struct NetTile<TicTacToe> { static constexpr int P = 40; };
template <> struct NetTile<Connect4> { static constexpr int P = 12; };
template <> struct NetTile<DragonChess> { static constexpr int P = 8; };

struct NetParams {
  const float* tower_w;    // [layer][tap][ci][16] (stem: ci padded to CIN_PAD), BN folded
  const float* tower_b;    // [layer][16]
  const float* head;       // value conv w[16], b | policy conv w[16][2], b[2] | d1 w[D], b[D] | d2 w[D], b
  const float* pol_w;      // [2][A]
  const float* pol_b;      // [A]
  int layers;              // 1 + 2*blocks
  int dense;
  float eps, alpha;
  unsigned long long seed;
  unsigned long long* call_counter;
};

struct TcParams {
  const uint4* w;          // [KB_IN*9 + (layers-1)*9] weight blocks of 512 bytes (bf16, canonical K-major layout)
  int w_blocks;
};

// tower widths other than 16 (net_wide.cuh): layer-by-layer over activations in HBM
struct WideDev {
  bool on = false;
  int F = 0, layers = 0, ksteps0 = 0;      // ksteps0: 16-channel steps of the input planes
  const float* w32 = nullptr;              // fp32: [layer][tap][ci][F] (stem ci = cin), BN folded
  const float* bias = nullptr;             // [layer][F]
  std::vector<size_t> w32_off, wbf_off;    // per layer: float / byte offsets
  const uint8_t* wbf = nullptr;            // bf16 operand layout (net_wide.cuh)
  const float* head_conv = nullptr;        // 3 F + 3
  uint8_t *a0 = nullptr, *ax = nullptr, *ay = nullptr;      // bf16 activation tiles of one chunk: input planes, X, Y
  float *f0 = nullptr, *fx = nullptr, *fy = nullptr;        // fp32 twin
};

struct NetDev {
  int game;
  bb_net_desc desc;
  NetParams p;
  std::vector<void*> allocs;
  int sm_count;
  bool use_tc = false;     // bf16 tcgen05 kernel instead of the fp32 SIMT one (bb_net_desc.precision)
  TcParams tc{};
  WideDev wide;
};

template <class G>
struct NetGeom {
  static constexpr int P = NetTile<G>::P;
  static constexpr int HW = G::H * G::W;
  static constexpr int PW = G::W + 2, PH = G::H + 2, CELLS = PW * PH;
  static constexpr int CIN_PAD = (G::C + 3) / 4 * 4;
  static constexpr int NG0 = CIN_PAD / 4 > 4 ? CIN_PAD / 4 : 4;       // channel groups in buffer 0
  static constexpr int ITEMS = P * HW;
  static constexpr int IPT = (ITEMS + NET_THREADS - 1) / NET_THREADS; // items per thread
  static_assert(IPT <= 2, "tile too large");
  static constexpr int X0_FLOATS = NG0 * P * CELLS * 4;
  static constexpr int X1_FLOATS = 4 * P * CELLS * 4;
  static constexpr int STEM_W = 9 * CIN_PAD * NET_F;
  static constexpr int LAYER_W = 9 * NET_F * NET_F;
};

// 3x3 'same' convolution over the tile: in[cg][cell][4] -> out[cg][cell][4], 2 items per thread.
template <class G, int CIG, bool SKIP>
__device__ __forceinline__ void conv3x3(const float* __restrict__ in, float* __restrict__ out, const float* __restrict__ w,
                                        const float* __restrict__ bias, const int* cell, const bool* live) {
  using Ge = NetGeom<G>;
  constexpr int PLANE = Ge::P * Ge::CELLS;     // cells per channel group
  float acc[2][NET_F];
#pragma unroll
  for (int t = 0; t < 2; ++t)
#pragma unroll
    for (int co = 0; co < NET_F; ++co) acc[t][co] = bias[co];
#pragma unroll 1
  for (int tap = 0; tap < 9; ++tap) {
    const int off = (tap / 3 - 1) * Ge::PW + (tap % 3 - 1);
#pragma unroll
    for (int cg = 0; cg < CIG; ++cg) {
      const float4 a0 = *reinterpret_cast<const float4*>(in + ((size_t)cg * PLANE + cell[0] + off) * 4);
      const float4 a1 = *reinterpret_cast<const float4*>(in + ((size_t)cg * PLANE + cell[1] + off) * 4);
      const float av0[4] = {a0.x, a0.y, a0.z, a0.w}, av1[4] = {a1.x, a1.y, a1.z, a1.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4* wp = reinterpret_cast<const float4*>(w + ((tap * CIG * 4) + cg * 4 + q) * NET_F);
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          const float4 ww = wp[v];
          acc[0][v * 4 + 0] = fmaf(av0[q], ww.x, acc[0][v * 4 + 0]);
          acc[0][v * 4 + 1] = fmaf(av0[q], ww.y, acc[0][v * 4 + 1]);
          acc[0][v * 4 + 2] = fmaf(av0[q], ww.z, acc[0][v * 4 + 2]);
          acc[0][v * 4 + 3] = fmaf(av0[q], ww.w, acc[0][v * 4 + 3]);
          acc[1][v * 4 + 0] = fmaf(av1[q], ww.x, acc[1][v * 4 + 0]);
          acc[1][v * 4 + 1] = fmaf(av1[q], ww.y, acc[1][v * 4 + 1]);
          acc[1][v * 4 + 2] = fmaf(av1[q], ww.z, acc[1][v * 4 + 2]);
          acc[1][v * 4 + 3] = fmaf(av1[q], ww.w, acc[1][v * 4 + 3]);
        }
      }
    }
  }
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    if (!live[t]) continue;
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      float4* o = reinterpret_cast<float4*>(out + ((size_t)v * PLANE + cell[t]) * 4);
      float4 r = make_float4(acc[t][v * 4], acc[t][v * 4 + 1], acc[t][v * 4 + 2], acc[t][v * 4 + 3]);
      if (SKIP) { const float4 s = *o; r.x += s.x; r.y += s.y; r.z += s.z; r.w += s.w; }
      r.x = fmaxf(r.x, 0.f); r.y = fmaxf(r.y, 0.f); r.z = fmaxf(r.z, 0.f); r.w = fmaxf(r.w, 0.f);
      *o = r;
    }
  }
}

// A team = the threads that work on one batch together: the whole CTA in the fp32 kernel, one of the two
// pipelines of a CTA in the tensor-core kernel (named barrier `bar`, threads numbered tid = 0 .. n-1).
struct Team {
  int tid, n, bar;
  __device__ __forceinline__ void sync() const {           // literal barrier ids: a register id makes ptxas reserve all 16 barriers
    if (bar == 1) asm volatile("bar.sync 1, %0;" ::"r"(n) : "memory");
    else if (bar == 2) asm volatile("bar.sync 2, %0;" ::"r"(n) : "memory");
    else if (bar == 3) asm volatile("bar.sync 3, %0;" ::"r"(n) : "memory");
    else asm volatile("bar.sync 0, %0;" ::"r"(n) : "memory");
  }
};

__device__ __forceinline__ float block_reduce(float v, float* red, bool is_max, const Team& tm) {
  // fixed-order tree: deterministic
  const int lane = tm.tid & 31, wid = tm.tid >> 5;
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    float t = __shfl_xor_sync(0xFFFFFFFFu, v, o);
    v = is_max ? fmaxf(v, t) : v + t;
  }
  tm.sync();
  if (lane == 0) red[wid] = v;
  tm.sync();
  float r = red[0];
  for (int i = 1; i < (tm.n >> 5); ++i) r = is_max ? fmaxf(r, red[i]) : r + red[i];
  return r;
}

// Value and policy heads from the per-position sums over H,W of the three head feature maps
// (hsum[p*4 + {0: value, 1,2: policy}]): value = tanh(dense(relu(dense(1->D)+sum))), policy =
// softmax(dense(2->A)+sum), noise mix, renormalise (NetworkFactory.py:115-182).  Block-wide.
// FAST (the bf16 tensor-core kernel only; its tolerance is 2e-2): wide heads without noise take three passes per
// position (logits + max with batched non-coherent weight loads, ex2.approx + sum, scaled store) instead of four
// with two IEEE divisions per action.
// serve mode: where the narrow heads also store their answers for the evaluation cache (common.cuh); ent[p] = entry index + 1
// of batch position p when this pipeline holds the entry's lock, 0 otherwise
struct CacheOut {
  uint32_t* base = nullptr;
  const uint32_t* ent = nullptr;
  const uint8_t* states = nullptr;        // the batch's staged packed states
  int stride = 0, state_bytes = 0;
  int legal_only = 0;                     // wide heads: softmax over the LEGAL actions of the staged state only (see heads_finish)
#ifdef BB_NET_TIMING_HEADS
  unsigned long long* dbg = nullptr;      // legal-only head: cycles in moves + list, logits, soft-max + noise, stores; positions
#endif
};

template <class G, bool FAST = false>
__device__ void heads_finish(const NetParams& np, const float* hd, const float* pol_w, const float* pol_b, const float* hsum,
                             float* red, int base, int n, int P, unsigned long long call, float* __restrict__ policy,
                             float* __restrict__ value, const Team& tm, const int* out_idx = nullptr, float* zbuf = nullptr,
                             long long noise_base = -1, const CacheOut co = CacheOut()) {
  constexpr int HW = G::H * G::W;
  const int tid = tm.tid, nthreads = tm.n;
  // noise element index of position p: serve mode keys it on the full 32-bit ticket (noise_base), launch mode on the batch row
  const unsigned long long nb = noise_base >= 0 ? (unsigned long long)noise_base : (unsigned long long)base;
  const int D = np.dense;                               // hd / pol_w / pol_b: np.head / np.pol_w / np.pol_b or copies in shared memory
  const float* d1w = hd + 17 + 34;
  const float* d1b = d1w + D;
  const float* d2w = d1b + D;
  const float d2b = d2w[D];
  if (G::A <= 32) {
    // Narrow heads: GP lanes per position (8 for up to 8 actions, 16 up to 16, else 32), all positions of the batch side by
    // side in one or two warps.  Lane j of a group owns action j (logit, softmax through fixed-order xor trees of width GP,
    // its own noise draw, renormalisation) and the dense units j, j + GP, ... of the value head.  Round 1 gave a whole
    // position to ONE thread (7 expf, 14 divisions, a tanh and a 16-step dependent loop in a row: most of the cycles a
    // serve-mode batch spent after its tower); a whole WARP per position was measured too and lost -- the SM is short of
    // issue slots, not of lanes.
    constexpr int GP = G::A <= 8 ? 8 : G::A <= 16 ? 16 : 32;
    for (int pb = 0; pb < P; pb += nthreads / GP)         // a team of 160 threads covers 20 Connect4 or 10 TicTacToe positions per pass
    if ((tid & ~31) < (P - pb) * GP) {                    // whole warps only: the shuffles below need all 32 lanes
      const int p = pb + tid / GP, j = tid % GP;
      const bool live = p < P && base + p < n;
      const int pc = live ? p : 0;
      uint32_t* cent = nullptr;                          // this position's evaluation-cache entry, when this pipeline owns it
      if (G::A <= ECACHE_MAX_ACTIONS && co.base && live && co.ent[p]) cent = co.base + (size_t)(co.ent[p] - 1u) * ECACHE_WORDS;
      const size_t orow = out_idx ? (size_t)out_idx[pc] : (size_t)(base + pc);
      // value: dense(1->D)+sum, ReLU, dense(D->1), tanh
      const float sv = hsum[pc * 4];
      float part = 0.f;
      for (int d = j; d < D; d += GP) part = fmaf(fmaxf(fmaf(sv, d1w[d], (float)HW * d1b[d]), 0.f), d2w[d], part);
#pragma unroll
      for (int o = GP / 2; o; o >>= 1) part += __shfl_xor_sync(0xFFFFFFFFu, part, o, GP);
      if (live && j == 0) {
        const float v = tanhf(part + d2b);
        value[orow] = v;
        if (cent) {
          cent[13] = __float_as_uint(v);
          const uint32_t* st = reinterpret_cast<const uint32_t*>(co.states + p * co.stride);
#pragma unroll
          for (int i = 0; i < 4; ++i) cent[i] = i * 4 < co.state_bytes ? st[i] : 0u;
        }
      }
      // policy: dense(2->A)+sum, softmax, noise mix, renormalise (NetworkFactory.py:165-182)
      const bool act = j < G::A;
      const int a = act ? j : 0;
      const float s0 = hsum[pc * 4 + 1], s1 = hsum[pc * 4 + 2];
      const float z = act ? fmaf(s0, pol_w[a], fmaf(s1, pol_w[G::A + a], (float)HW * pol_b[a])) : -INFINITY;
      float m = z;
#pragma unroll
      for (int o = GP / 2; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xFFFFFFFFu, m, o, GP));
      const float e = act ? expf(z - m) : 0.f;
      float sum = e;
#pragma unroll
      for (int o = GP / 2; o; o >>= 1) sum += __shfl_xor_sync(0xFFFFFFFFu, sum, o, GP);
      float pa = e / sum;
      if (np.eps > 0.f && act && live)
        pa = (1.f - np.eps) * pa + np.eps * beta_noise(np.alpha, np.seed ^ (call * 0x632BE59BD9B4E019ull), (nb + p) * G::A + a);
      float tot = pa;
#pragma unroll
      for (int o = GP / 2; o; o >>= 1) tot += __shfl_xor_sync(0xFFFFFFFFu, tot, o, GP);
      if (live && act) {
        const float pf = pa / tot;
        policy[orow * G::A + a] = pf;
        if (G::A <= ECACHE_MAX_ACTIONS && cent) cent[4 + a] = __float_as_uint(pf);
      }
    }
  } else {
    if (tid < P && base + tid < n) {                     // value: dense(1->D)+sum, ReLU, dense(D->1), tanh
      const float sv = hsum[tid * 4];
      float o = d2b;
      for (int d = 0; d < D; ++d) o = fmaf(fmaxf(fmaf(sv, d1w[d], (float)HW * d1b[d]), 0.f), d2w[d], o);
      value[out_idx ? out_idx[tid] : base + tid] = tanhf(o);
    }
    if constexpr (FAST && G::NROWS > 1) {
      if (co.legal_only && zbuf) {
        // The search only ever reads the policy at the LEGAL actions of the position it asked about (GetPriors:
        // getPolicy(state) * LegalActions() / sum, Blackbird.py:386-387), and the engine holds that position: generate
        // its legal moves here, evaluate the <= ~100 logits that matter instead of all 4032, soft-max over those, and
        // store them at their action index (the other entries of the row are left alone).  The normaliser cancels in
        // GetPriors' division, so the priors are the reference's up to fp32 rounding of two different partial sums --
        // well inside the bf16 kernel's tolerance; the fp32 kernel keeps the dense head.
        // A WARP per position (the positions of a batch side by side): the occupancy by two ballots per
        // colour, lane l lists the moves of squares 2l and 2l + 1 at its scan offset, then the <= 256 logits, their soft-max
        // and the noise mix with xor-tree reductions -- no block barrier anywhere.  The block-wide form below took the
        // positions one after the other with six named barriers each: 24 k cycles of a 49 k-cycle DragonChess batch.
        const int wq = tid >> 5, lane = tid & 31, nw = nthreads >> 5;
        constexpr int CAP = 256, QPL = CAP / 32;
        int* list = reinterpret_cast<int*>(zbuf) + wq * CAP;           // (A floats of scratch: room for 15 warps)
        for (int p = wq; p < P && base + p < n; p += nw) {
#ifdef BB_NET_TIMING_HEADS
          const long long h0 = clock64();
#endif
          typename G::State st;
          memcpy(&st, co.states + p * co.stride, sizeof(st));
          uint64_t own, foe;
          G::occ_warp(st, &own, &foe);
          const uint64_t m0 = G::legal_row_occ(st, 2 * lane, own, foe), m1 = G::legal_row_occ(st, 2 * lane + 1, own, foe);
          const int cnt = popc64(m0) + popc64(m1);
          int incl = cnt;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += v;
          }
          const int k = min(__shfl_sync(0xFFFFFFFFu, incl, 31), CAP);
          int j = incl - cnt;
          for (uint64_t mm = m0; mm && j < CAP; mm &= mm - 1) list[j++] = G::action_of(2 * lane, ctz64(mm));
          for (uint64_t mm = m1; mm && j < CAP; mm &= mm - 1) list[j++] = G::action_of(2 * lane + 1, ctz64(mm));
          __syncwarp();
#ifdef BB_NET_TIMING_HEADS
          const long long h1 = clock64();
#endif
          const float s0 = hsum[p * 4 + 1], s1 = hsum[p * 4 + 2];
          float* out = policy + (out_idx ? (size_t)out_idx[p] : (size_t)(base + p)) * G::A;
          float z[QPL];
          int act[QPL];
          float mx = -INFINITY;
#pragma unroll
          for (int q = 0; q < QPL; ++q) {
            const int i = q * 32 + lane;
            act[q] = i < k ? list[i] : 0;
            z[q] = i < k ? fmaf(s0, __ldg(pol_w + act[q]), fmaf(s1, __ldg(pol_w + G::A + act[q]), (float)HW * __ldg(pol_b + act[q]))) : -INFINITY;
            mx = fmaxf(mx, z[q]);
          }
#pragma unroll
          for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, o));
#ifdef BB_NET_TIMING_HEADS
          const long long h2 = clock64();
#endif
          float sum = 0.f;
#pragma unroll
          for (int q = 0; q < QPL; ++q) { z[q] = q * 32 + lane < k ? __expf(z[q] - mx) : 0.f; sum += z[q]; }
#pragma unroll
          for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xFFFFFFFFu, sum, o);
          if (np.eps > 0.f) {                                           // (1 - eps) softmax + eps noise, renormalised -- over the legal actions
            float tot = 0.f;
#pragma unroll
            for (int q = 0; q < QPL; ++q)
              if (q * 32 + lane < k) {
                z[q] = (1.f - np.eps) * (z[q] / sum) + np.eps * beta_noise(np.alpha, np.seed ^ (call * 0x632BE59BD9B4E019ull), (nb + p) * G::A + act[q]);
                tot += z[q];
              }
#pragma unroll
            for (int o = 16; o; o >>= 1) tot += __shfl_xor_sync(0xFFFFFFFFu, tot, o);
            sum = tot;
          }
          const float inv = 1.f / sum;
#ifdef BB_NET_TIMING_HEADS
          const long long h3 = clock64();
#endif
#pragma unroll
          for (int q = 0; q < QPL; ++q)
            if (q * 32 + lane < k) out[act[q]] = z[q] * inv;
          __syncwarp();                                                 // the list is rewritten by this warp's next position
#ifdef BB_NET_TIMING_HEADS
          if (co.dbg && lane == 0) {
            const long long h4 = clock64();
            atomicAdd(co.dbg + 0, (unsigned long long)(h1 - h0)); atomicAdd(co.dbg + 1, (unsigned long long)(h2 - h1));
            atomicAdd(co.dbg + 2, (unsigned long long)(h3 - h2)); atomicAdd(co.dbg + 3, (unsigned long long)(h4 - h3)); atomicAdd(co.dbg + 4, 1ull);
          }
#endif
        }
        return;
      }
    }
    for (int p = 0; p < P && base + p < n; ++p) {      // block-wide over the 4032 actions
      const float s0 = hsum[p * 4 + 1], s1 = hsum[p * 4 + 2];
      float* out = policy + (out_idx ? (size_t)out_idx[p] : (size_t)(base + p)) * G::A;
      // intermediate values live in shared memory when the caller has room for A floats (every thread touches
      // only its own actions, so no barrier beyond the reductions'); otherwise in the output row itself.
      // Same operations in the same order either way: the results are bit-identical.
      float* z_ = (zbuf && G::A > 32) ? zbuf : out;
      float m = -INFINITY;
      if (FAST && !(np.eps > 0.f)) {
        // the head weights sit in L2 (48 KB): non-coherent loads, four actions in flight per thread, so the pass costs
        // ~7 dependent round trips instead of one per action
#pragma unroll 4
        for (int a = tid; a < G::A; a += nthreads) {
          const float z = fmaf(s0, __ldg(pol_w + a), fmaf(s1, __ldg(pol_w + G::A + a), (float)HW * __ldg(pol_b + a)));
          z_[a] = z;
          m = fmaxf(m, z);
        }
        m = block_reduce(m, red, true, tm);
        float sum = 0.f;
#pragma unroll 4
        for (int a = tid; a < G::A; a += nthreads) { const float e = __expf(z_[a] - m); z_[a] = e; sum += e; }
        sum = block_reduce(sum, red, false, tm);
        const float inv = 1.f / sum;
#pragma unroll 4
        for (int a = tid; a < G::A; a += nthreads) out[a] = z_[a] * inv;
        continue;
      }
      for (int a = tid; a < G::A; a += nthreads) {
        float z = fmaf(s0, pol_w[a], fmaf(s1, pol_w[G::A + a], (float)HW * pol_b[a]));
        z_[a] = z;
        m = fmaxf(m, z);
      }
      m = block_reduce(m, red, true, tm);
      float sum = 0.f;
      for (int a = tid; a < G::A; a += nthreads) { float e = expf(z_[a] - m); z_[a] = e; sum += e; }
      sum = block_reduce(sum, red, false, tm);
      float tot = 0.f;
      for (int a = tid; a < G::A; a += nthreads) {
        float pa = z_[a] / sum;
        if (np.eps > 0.f)
          pa = (1.f - np.eps) * pa + np.eps * beta_noise(np.alpha, np.seed ^ (call * 0x632BE59BD9B4E019ull), (nb + p) * G::A + a);
        z_[a] = pa;
        tot += pa;
      }
      tot = block_reduce(tot, red, false, tm);
      for (int a = tid; a < G::A; a += nthreads) out[a] = z_[a] / tot;
    }
  }
}

}  // namespace bb
#include "net_tc.cuh"
#include "net_wide.cuh"
namespace bb {

template <class G>
__global__ void __launch_bounds__(NET_THREADS, 1)
k_net_forward(NetParams np, const typename G::State* __restrict__ states, int n_max, const int* __restrict__ count_dev,
              float* __restrict__ policy, float* __restrict__ value) {
  using Ge = NetGeom<G>;
  extern __shared__ __align__(16) float smem[];
  float* sw = smem;                                         // tower weights
  float* sb = sw + Ge::STEM_W + (np.layers - 1) * Ge::LAYER_W;
  float* x0 = sb + np.layers * NET_F;
  float* x1 = x0 + Ge::X0_FLOATS;
  float* red = x1 + Ge::X1_FLOATS;                          // 8 floats block-reduce scratch
  float* hsum = red + 8;                                    // [P][3] head sums
  float* hcell = hsum + Ge::P * 4;                          // [3][ITEMS] per-cell head activations
  typename G::State* sst = reinterpret_cast<typename G::State*>(hcell + 3 * Ge::ITEMS + (Ge::ITEMS & 1));
  const int tid = threadIdx.x;
  const int n = count_dev ? min(n_max, *count_dev) : n_max;
  const int n_tiles = (n + Ge::P - 1) / Ge::P;
  if ((int)blockIdx.x >= n_tiles) return;

  const int wtot = Ge::STEM_W + (np.layers - 1) * Ge::LAYER_W;
  for (int i = tid * 4; i < wtot; i += NET_THREADS * 4)
    *reinterpret_cast<float4*>(sw + i) = *reinterpret_cast<const float4*>(np.tower_w + i);
  for (int i = tid; i < np.layers * NET_F; i += NET_THREADS) sb[i] = np.tower_b[i];
  for (int i = tid; i < Ge::X0_FLOATS + Ge::X1_FLOATS; i += NET_THREADS) x0[i] = 0.f;   // zero halos once

  int cell[2], pos[2], hw[2];
  bool live[2];
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int item = tid + t * NET_THREADS;
    live[t] = item < Ge::ITEMS;
    const int it = live[t] ? item : 0;
    pos[t] = it / Ge::HW;
    hw[t] = it % Ge::HW;
    cell[t] = pos[t] * Ge::CELLS + (hw[t] / G::W + 1) * Ge::PW + (hw[t] % G::W + 1);
  }
  constexpr int PLANE = Ge::P * Ge::CELLS;
  const float* hd = np.head;
  unsigned long long call = 0;
  if (np.eps > 0.f) call = *np.call_counter;

  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int base = tile * Ge::P;
    __syncthreads();
    if (tid < Ge::P) sst[tid] = states[min(base + tid, n - 1)];
    __syncthreads();
    // decode input planes (AsInputArray) straight from the packed state
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      if (!live[t]) continue;
      const typename G::State s = sst[pos[t]];
#pragma unroll
      for (int cg = 0; cg < Ge::CIN_PAD / 4; ++cg) {
        float v[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) v[q] = (cg * 4 + q < G::C) ? (float)G::plane(s, hw[t], cg * 4 + q) : 0.f;
        *reinterpret_cast<float4*>(x0 + ((size_t)cg * PLANE + cell[t]) * 4) = make_float4(v[0], v[1], v[2], v[3]);
      }
    }
    __syncthreads();
    conv3x3<G, Ge::CIN_PAD / 4, false>(x0, x1, sw, sb, cell, live);                         // stem
    const float* lw = sw + Ge::STEM_W;
    for (int l = 1; l < np.layers; l += 2) {
      __syncthreads();
      conv3x3<G, 4, false>(x1, x0, lw, sb + l * NET_F, cell, live);                          // conv_1 + BN + ReLU
      __syncthreads();
      conv3x3<G, 4, true>(x0, x1, lw + Ge::LAYER_W, sb + (l + 1) * NET_F, cell, live);       // conv_2 + BN + skip + ReLU
      lw += 2 * Ge::LAYER_W;
    }
    __syncthreads();
    // heads: 1x1 convs (1 value filter, 2 policy filters) + BN + ReLU per cell
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      if (!live[t]) continue;
      float v = hd[16], p0 = hd[17 + 32], p1 = hd[17 + 33];
#pragma unroll
      for (int cg = 0; cg < 4; ++cg) {
        const float4 a = *reinterpret_cast<const float4*>(x1 + ((size_t)cg * PLANE + cell[t]) * 4);
        const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int ci = cg * 4 + q;
          v = fmaf(av[q], hd[ci], v);
          p0 = fmaf(av[q], hd[17 + ci * 2], p0);
          p1 = fmaf(av[q], hd[17 + ci * 2 + 1], p1);
        }
      }
      const int item = tid + t * NET_THREADS;
      hcell[item] = fmaxf(v, 0.f);
      hcell[Ge::ITEMS + item] = fmaxf(p0, 0.f);
      hcell[2 * Ge::ITEMS + item] = fmaxf(p1, 0.f);
    }
    __syncthreads();
    if (tid < Ge::P * 3) {                                   // sum over H,W in a fixed order
      const int p = tid / 3, k = tid % 3;
      float s = 0.f;
      for (int c = 0; c < Ge::HW; ++c) s += hcell[k * Ge::ITEMS + p * Ge::HW + c];
      hsum[p * 4 + k] = s;
    }
    __syncthreads();
    // hcell (consumed by the sums above) doubles as the noise scratch of the narrow heads: P * A <= 3 * ITEMS floats
    heads_finish<G>(np, np.head, np.pol_w, np.pol_b, hsum, red, base, n, Ge::P, call, policy, value, Team{(int)threadIdx.x, (int)blockDim.x, 0},
                    nullptr, G::A <= 32 ? hcell : nullptr);
  }
}

__global__ void k_bump(unsigned long long* c) { *c += 1; }

template <class G>
static size_t net_smem_bytes(int layers) {
  using Ge = NetGeom<G>;
  size_t fl = Ge::STEM_W + (size_t)(layers - 1) * Ge::LAYER_W + (size_t)layers * NET_F + Ge::X0_FLOATS + Ge::X1_FLOATS + 8 + Ge::P * 4 + 3 * Ge::ITEMS + 2;
  return fl * 4 + Ge::P * sizeof(typename G::State) + 16;
}

template <class G>
static int net_launch(NetDev* net, const void* states, int n, const int* count_dev, float* policy, float* value, cudaStream_t stream) {
  using Ge = NetGeom<G>;
  const size_t smem = net_smem_bytes<G>(net->p.layers);
  static bool configured[3] = {false, false, false};
  if (!configured[G::ID]) {
    BB_CUDA(cudaFuncSetAttribute(k_net_forward<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured[G::ID] = true;
  }
  const int tiles = (n + Ge::P - 1) / Ge::P;
  const int grid = tiles < net->sm_count ? tiles : net->sm_count;
  k_net_forward<G><<<grid, NET_THREADS, smem, stream>>>(net->p, (const typename G::State*)states, n, count_dev, policy, value);
  BB_CUDA(cudaGetLastError());
  if (net->p.eps > 0.f) k_bump<<<1, 1, 0, stream>>>(net->p.call_counter);
  return BB_OK;
}

template <class G>
static int net_launch_tc(NetDev* net, const void* states, int n, const int* count_dev, float* policy, float* value, cudaStream_t stream, int legal_only) {
  using Ge = TcGeom<G>;
  const TcSmem<G> L(net->tc.w_blocks, net->p.layers);
  static bool configured[3] = {false, false, false};
  static int ctas_per_sm[3] = {1, 1, 1};
  if (!configured[G::ID]) {
    BB_CUDA(cudaFuncSetAttribute(k_net_forward_tc<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, L.total));
    BB_CUDA(cudaFuncSetAttribute(k_net_forward_tc<G>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
    int occ = 1;
    BB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_net_forward_tc<G>, Ge::THREADS, L.total));
    ctas_per_sm[G::ID] = occ < 1 ? 1 : (occ > 2 ? 2 : occ);
    if (getenv("BB_DEBUG_OCC")) {
      cudaFuncAttributes fa;
      cudaFuncGetAttributes(&fa, k_net_forward_tc<G>);
      int o2 = 0, o3 = 0;
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o2, k_net_forward_tc<G>, Ge::THREADS, 40000);
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o3, k_net_forward_tc<G>, 256, L.total);
      fprintf(stderr, "[bbengine] regs %d static smem %zu local %zu maxthreads %d; occ(smem 40000)=%d occ(256 thr)=%d\n", fa.numRegs,
              fa.sharedSizeBytes, fa.localSizeBytes, fa.maxThreadsPerBlock, o2, o3);
    }
    if (getenv("BB_DEBUG_OCC")) fprintf(stderr, "[bbengine] k_net_forward_tc<%d>: %d threads, %d B smem, occupancy %d CTAs/SM\n", G::ID, Ge::THREADS, L.total, occ);
    if (const char* e = getenv("BB_TC_CTAS_PER_SM")) ctas_per_sm[G::ID] = atoi(e) > 0 ? atoi(e) : 1;   // tuning knob
    configured[G::ID] = true;
  }
  // the kernel spreads the (device-side) request count evenly over the grid, so launch one full wave
  const int cap = net->sm_count * ctas_per_sm[G::ID];
  const int want = (n + Ge::PIPES - 1) / Ge::PIPES;             // every CTA runs PIPES batches side by side
  const int grid = want < cap ? want : cap;
  NetServe launch{};
  launch.legal_only = legal_only;
  k_net_forward_tc<G><<<grid, Ge::THREADS, L.total, stream>>>(net->p, net->tc, (const typename G::State*)states, n, count_dev, policy, value, launch);
  BB_CUDA(cudaGetLastError());
  if (net->p.eps > 0.f) k_bump<<<1, 1, 0, stream>>>(net->p.call_counter);
  return BB_OK;
}

template <class G>
static int net_serve_tc(NetDev* net, const NetServe& sv, float* policy, float* value, cudaStream_t stream) {
  using Ge = TcGeom<G>;
  const TcSmem<G> L(net->tc.w_blocks, net->p.layers);
  BB_CUDA(cudaFuncSetAttribute(k_net_forward_tc<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, L.total));
  BB_CUDA(cudaFuncSetAttribute(k_net_forward_tc<G>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
  // a new noise epoch per persistent launch: inside a launch the noise is keyed on the 32-bit ticket
  if (net->p.eps > 0.f) k_bump<<<1, 1, 0, stream>>>(net->p.call_counter);
  // one CTA per SM (a kernel that allocates TMEM never gets more), all resident for the whole launch
  k_net_forward_tc<G><<<net->sm_count, Ge::THREADS, L.total, stream>>>(net->p, net->tc, nullptr, 0, nullptr, policy, value, sv);
  BB_CUDA(cudaGetLastError());
  return BB_OK;
}

// ---- other tower widths (net_wide.cuh) ---------------------------------------------------------------------------------
template <int F>
static int wide_conv_launch(const WideLayer& L, const WideGeo& geo, int n_chunk, int chunk_base, const int* count_dev, int sm_count, cudaStream_t stream) {
  using C = WideCfg<F>;
  static bool configured = false;
  if (!configured) {
    BB_CUDA(cudaFuncSetAttribute(k_wide_conv<F>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
    configured = true;
  }
  const int tiles = (n_chunk + geo.PPT - 1) / geo.PPT, groups = (tiles + C::TPG - 1) / C::TPG;
  k_wide_conv<F><<<groups < sm_count ? groups : sm_count, C::THREADS, C::SMEM, stream>>>(L, geo, n_chunk, chunk_base, count_dev);
  BB_CUDA(cudaGetLastError());
  return BB_OK;
}

template <class G>
static int net_launch_wide(NetDev* net, const void* states, int n, const int* count_dev, float* policy, float* value, cudaStream_t stream, int legal_only) {
  using Tg = TcGeom<G>;
  using State = typename G::State;
  WideDev& wd = net->wide;
  const int F = wd.F, HW = G::H * G::W;
  const WideGeo geo = wide_geo<G>();
  const size_t hsmem = wide_heads_smem<G>(F);
  static bool configured[3][2] = {{false, false}, {false, false}, {false, false}};
  if (!configured[G::ID][net->use_tc]) {
    const int hmax = (int)wide_heads_smem<G>(1024);                  // the widest tower bb_net_load accepts
    if (net->use_tc) BB_CUDA(cudaFuncSetAttribute(k_wide_heads<G, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, hmax));
    else BB_CUDA(cudaFuncSetAttribute(k_wide_heads<G, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, hmax));
    configured[G::ID][net->use_tc] = true;
  }
  for (int base = 0; base < n; base += WIDE_CHUNK) {
    const int nc = n - base < WIDE_CHUNK ? n - base : WIDE_CHUNK;
    const unsigned cell_blocks = (unsigned)(((size_t)nc * HW + 255) / 256);
    const void* last = nullptr;
    if (net->use_tc) {
      k_wide_stage<G><<<cell_blocks, 256, 0, stream>>>((const State*)states, wd.a0, nc, base, count_dev);
      for (int l = 0; l < wd.layers; ++l) {
        const bool odd = l & 1;
        WideLayer L;
        L.in = l == 0 ? wd.a0 : (odd ? wd.ax : wd.ay);
        L.out = odd ? wd.ay : wd.ax;
        L.skip = (!odd && l > 0) ? wd.ax : nullptr;                  // conv_2 + BN + block input (NetworkFactory.py:95-98)
        L.w = wd.wbf + wd.wbf_off[l];
        L.bias = wd.bias + (size_t)l * F;
        L.ksteps = l == 0 ? wd.ksteps0 : F / 16;
        int rc = F == 32 ? wide_conv_launch<32>(L, geo, nc, base, count_dev, net->sm_count, stream)
               : F == 64 ? wide_conv_launch<64>(L, geo, nc, base, count_dev, net->sm_count, stream)
               : F == 128 ? wide_conv_launch<128>(L, geo, nc, base, count_dev, net->sm_count, stream)
                          : wide_conv_launch<256>(L, geo, nc, base, count_dev, net->sm_count, stream);
        if (rc != BB_OK) return rc;
        last = L.out;
      }
      WideHeadW hw{wd.head_conv};
      k_wide_heads<G, true><<<(nc + WIDE_HP - 1) / WIDE_HP, 256, hsmem, stream>>>(net->p, hw, F, last, (const State*)states, nc, base, count_dev, policy, value, legal_only);
    } else {
      k_wide32_stage<G><<<cell_blocks, 256, 0, stream>>>((const State*)states, wd.f0, nc, base, count_dev);
      for (int l = 0; l < wd.layers; ++l) {
        const bool odd = l & 1;
        const float* in = l == 0 ? wd.f0 : (odd ? wd.fx : wd.fy);
        float* out = odd ? wd.fy : wd.fx;
        const size_t total = (size_t)nc * HW * F;
        const unsigned blocks = (unsigned)((total + 255) / 256 < 65535 * 16 ? (total + 255) / 256 : 65535 * 16);
        k_wide32_conv<<<blocks, 256, 0, stream>>>(in, wd.w32 + wd.w32_off[l], wd.bias + (size_t)l * F, (!odd && l > 0) ? wd.fx : nullptr, out, G::H, G::W,
                                                    l == 0 ? G::C : F, F, nc, base, count_dev);
        last = out;
      }
      WideHeadW hw{wd.head_conv};
      k_wide_heads<G, false><<<(nc + WIDE_HP - 1) / WIDE_HP, 256, hsmem, stream>>>(net->p, hw, F, last, (const State*)states, nc, base, count_dev, policy, value, 0);
    }
    BB_CUDA(cudaGetLastError());
  }
  (void)sizeof(Tg);
#ifdef BB_WIDE_TIMING
  if (getenv("BB_DEBUG_WIDE") && net->use_tc) {
    unsigned long long h[8];
    cudaStreamSynchronize(stream);
    cudaMemcpyFromSymbol(h, g_wide_dbg, sizeof h);
    if (h[4]) fprintf(stderr, "[bbengine] k_wide_conv<%d> MMA warp per tile group: waits for tiles %.0f, for the accumulator %.0f, for weight stages %.0f of %.0f cycles (%llu groups)\n", F,
                      (double)h[0] / h[4], (double)h[1] / h[4], (double)h[2] / h[4], (double)h[3] / h[4], h[4]);
    unsigned long long z[8] = {0};
    cudaMemcpyToSymbol(g_wide_dbg, z, sizeof z);
  }
#endif
  if (net->p.eps > 0.f) k_bump<<<1, 1, 0, stream>>>(net->p.call_counter);
  return BB_OK;
}

int net_serve(NetDev* net, const NetServe& sv, float* policy, float* value, cudaStream_t stream) {
  if (!net->use_tc) { set_error("streaming mode needs the tensor-core network kernel (precision bf16)"); return BB_EINVAL; }
  BB_DISPATCH_GAME(net->game, return net_serve_tc<G>(net, sv, policy, value, stream));
  return BB_EINVAL;
}
template <class G>
static void tc_footprint(NetDev* net, int* regs, int* smem) {
  cudaFuncAttributes fa{};
  cudaFuncGetAttributes(&fa, k_net_forward_tc<G>);
  // an SM's register file is four 16 K files, one per scheduler; a CTA's warps are dealt round-robin, and per-warp
  // allocations are rounded up to 256 registers: what counts is the fullest quarter (measured: 12 search warps of
  // 128 registers + 10 network warps of 48 add up to 64.5 K of 64 K and still do not share an SM; 12 x 104 next to
  // 15 x 48 = 3 * 3328 + 4 * 1536 = 16128 per quarter do)
  *regs = ((fa.numRegs * 32 + 255) / 256 * 256) * ((TcGeom<G>::THREADS / 32 + 3) / 4);
  *smem = TcSmem<G>(net->tc.w_blocks, net->p.layers).total + 1024;
}
// registers (in the fullest of the SM's four register files) and shared memory one CTA of the tensor-core kernel holds on its SM
int net_pipes(NetDev* net) {                       // pipelines per CTA of the tensor-core kernel: the servers of the request ring
  BB_DISPATCH_GAME(net->game, return TcGeom<G>::PIPES);
  return 1;
}
int net_tc_footprint(NetDev* net, int* regs, int* smem) {
  BB_DISPATCH_GAME(net->game, tc_footprint<G>(net, regs, smem); return BB_OK);
  return BB_EINVAL;
}
int net_sm_count(NetDev* net) { return net->sm_count; }
bool net_is_tc(NetDev* net) { return net->use_tc && !net->wide.on; }   // the persistent (streaming) form exists for the 16-filter kernel; wider towers run tick by tick

int net_forward(NetDev* net, const void* states, int n, const int* count_dev, float* policy, float* value, cudaStream_t stream, int legal_only) {
  if (n <= 0) return BB_OK;
  if (net->wide.on) { BB_DISPATCH_GAME(net->game, return net_launch_wide<G>(net, states, n, count_dev, policy, value, stream, legal_only)); }
  if (net->use_tc) { BB_DISPATCH_GAME(net->game, return net_launch_tc<G>(net, states, n, count_dev, policy, value, stream, legal_only)); }
  else { BB_DISPATCH_GAME(net->game, return net_launch<G>(net, states, n, count_dev, policy, value, stream)); }
  return BB_EINVAL;
}

// Host side of bb_net_load for tower widths other than 16 (same flat input order as below)
template <class G>
static int net_build_wide(NetDev* net, const bb_net_desc* d, const float* w, size_t n_floats) {
  using Tg = TcGeom<G>;
  const int F = d->filters, D = d->dense, blocks = d->blocks, cin = d->cin, A = G::A, HW = G::H * G::W;
  const bool tc = d->precision == 1;
  if (F < 1 || F > 1024) { set_error("filters must be in [1, 1024] (got %d)", F); return BB_EINVAL; }
  if (tc && F != 32 && F != 64 && F != 128 && F != 256) { set_error("the tensor-core network kernels are built for filters = 16, 32, 64, 128 and 256 (got %d); precision fp32 takes any width", F); return BB_EINVAL; }
  if (cin != G::C) { set_error("cin %d does not match the game's %d input planes", cin, G::C); return BB_EINVAL; }
  if (D < 1 || D > NET_MAX_DENSE || blocks < 0) { set_error("dense must be in [1,%d], blocks >= 0", NET_MAX_DENSE); return BB_EINVAL; }
  auto conv_sz = [](int k, int ci, int co) { return (size_t)k * k * ci * co + 5 * (size_t)co; };
  const size_t need = conv_sz(3, cin, F) + 2 * (size_t)blocks * conv_sz(3, F, F) + conv_sz(1, F, 1) + D + D + D + 1 +
                      conv_sz(1, F, 2) + 2 * (size_t)A + A;
  if (n_floats != need) { set_error("bb_net_load: expected %zu floats for this shape, got %zu", need, n_floats); return BB_EINVAL; }
  const int layers = 1 + 2 * blocks;
  WideDev& wd = net->wide;
  wd.on = true; wd.F = F; wd.layers = layers; wd.ksteps0 = Tg::KB_IN;
  const float* p = w;
  auto fold = [&](int k, int ci, int co, float* wout, float* bout) {          // wout[tap][ci][co]
    const float* cw = p; p += (size_t)k * k * ci * co;
    const float *b = p, *g = p + co, *be = p + 2 * co, *mu = p + 3 * co, *var = p + 4 * co;
    p += 5 * (size_t)co;
    for (int o = 0; o < co; ++o) {
      const double s = (double)g[o] / sqrt((double)var[o] + 1e-3);
      bout[o] = (float)(((double)b[o] - (double)mu[o]) * s + (double)be[o]);
      for (size_t t = 0; t < (size_t)k * k * ci; ++t) wout[t * co + o] = (float)((double)cw[t * co + o] * s);
    }
  };
  size_t off = 0;
  for (int l = 0; l < layers; ++l) { wd.w32_off.push_back(off); off += (size_t)9 * (l == 0 ? cin : F) * F; }
  std::vector<float> tw(off), tb((size_t)layers * F);
  for (int l = 0; l < layers; ++l) fold(3, l == 0 ? cin : F, F, tw.data() + wd.w32_off[l], tb.data() + (size_t)l * F);
  std::vector<float> hconv(3 * (size_t)F + 3), head(51 + 3 * (size_t)D + 1, 0.f);
  fold(1, F, 1, hconv.data(), hconv.data() + F);                          // value conv: w[F], b
  memcpy(head.data() + 51, p, sizeof(float) * (3 * (size_t)D + 1));       // d1.w, d1.b, d2.w, d2.b where heads_finish looks for them
  p += 3 * (size_t)D + 1;
  fold(1, F, 2, hconv.data() + F + 1, hconv.data() + F + 1 + 2 * (size_t)F);   // policy conv: w[F][2], b[2]
  const float* pw = p;
  const float* pb = p + 2 * (size_t)A;
  auto upload = [&](const void* src, size_t bytes, const void** dst) -> int {
    void* q = nullptr;
    BB_CUDA(cudaMalloc(&q, bytes));
    net->allocs.push_back(q);
    BB_CUDA(cudaMemcpy(q, src, bytes, cudaMemcpyHostToDevice));
    *dst = q;
    return BB_OK;
  };
  auto zeroed = [&](size_t bytes, void** dst) -> int {
    BB_CUDA(cudaMalloc(dst, bytes));
    net->allocs.push_back(*dst);
    BB_CUDA(cudaMemset(*dst, 0, bytes));
    return BB_OK;
  };
  int rc;
  if ((rc = upload(tb.data(), tb.size() * 4, (const void**)&wd.bias))) return rc;
  if ((rc = upload(hconv.data(), hconv.size() * 4, (const void**)&wd.head_conv))) return rc;
  if ((rc = upload(head.data(), head.size() * 4, (const void**)&net->p.head))) return rc;
  if ((rc = upload(pw, 2 * (size_t)A * 4, (const void**)&net->p.pol_w))) return rc;
  if ((rc = upload(pb, (size_t)A * 4, (const void**)&net->p.pol_b))) return rc;
  net->use_tc = tc;
  if (tc) {
    // bf16 B operands: layer l, k-step ks, tap t (k-step outer: net_wide.cuh streams the activations k-step by k-step): N = F rows x K = 16,
    // element (n = cout, k) at byte (k / 8) * F * 16 + n * 16 + (k % 8) * 2
    auto bf16_rn = [](float f) -> uint16_t {
      uint32_t u;
      memcpy(&u, &f, 4);
      if ((u & 0x7F800000u) == 0x7F800000u) return (uint16_t)(u >> 16);
      u += 0x7FFFu + ((u >> 16) & 1u);
      return (uint16_t)(u >> 16);
    };
    size_t boff = 0;
    for (int l = 0; l < layers; ++l) { wd.wbf_off.push_back(boff); boff += (size_t)9 * (l == 0 ? wd.ksteps0 : F / 16) * F * 32; }
    std::vector<uint16_t> wb(boff / 2, 0);
    for (int l = 0; l < layers; ++l) {
      const int ci_n = l == 0 ? cin : F, ks_n = l == 0 ? wd.ksteps0 : F / 16;
      const float* lw = tw.data() + wd.w32_off[l];
      uint16_t* dst = wb.data() + wd.wbf_off[l] / 2;
      for (int t = 0; t < 9; ++t)
        for (int ks = 0; ks < ks_n; ++ks)
          for (int k = 0; k < 16; ++k)
            for (int o = 0; o < F; ++o) {
              const int ci = ks * 16 + k;
              const float v = ci < ci_n ? lw[((size_t)t * ci_n + ci) * F + o] : 0.f;
              const int nm = F <= 128 ? F : F / 2, hp = o / nm, on = o % nm;       // passes of net_wide.cuh (WideCfg::NS): [pass][k-step][tap][2][nm][8]
              dst[((size_t)((hp * ks_n + ks) * 9 + t) * nm * 32) / 2 + (size_t)(k / 8) * nm * 8 + (size_t)on * 8 + (k % 8)] = bf16_rn(v);
            }
    }
    if ((rc = upload(wb.data(), wb.size() * 2, (const void**)&wd.wbf))) return rc;
    const size_t tiles = (WIDE_CHUNK + Tg::PPT - 1) / Tg::PPT;
    if ((rc = zeroed(tiles * wd.ksteps0 * 2 * WIDE_PLANE, (void**)&wd.a0))) return rc;       // padding rows stay zero for good
    if ((rc = zeroed(tiles * (size_t)(F / 8) * WIDE_PLANE, (void**)&wd.ax))) return rc;
    if ((rc = zeroed(tiles * (size_t)(F / 8) * WIDE_PLANE, (void**)&wd.ay))) return rc;
  } else {
    if ((rc = upload(tw.data(), tw.size() * 4, (const void**)&wd.w32))) return rc;
    if ((rc = zeroed((size_t)WIDE_CHUNK * HW * cin * 4, (void**)&wd.f0))) return rc;
    if ((rc = zeroed((size_t)WIDE_CHUNK * HW * F * 4, (void**)&wd.fx))) return rc;
    if ((rc = zeroed((size_t)WIDE_CHUNK * HW * F * 4, (void**)&wd.fy))) return rc;
  }
  void* ctr = nullptr;
  BB_CUDA(cudaMalloc(&ctr, 8));
  BB_CUDA(cudaMemset(ctr, 0, 8));
  net->allocs.push_back(ctr);
  net->p.call_counter = (unsigned long long*)ctr;
  net->p.layers = layers;
  net->p.dense = D;
  net->p.eps = d->epsilon;
  net->p.alpha = d->alpha;
  net->p.seed = d->noise_seed;
  return BB_OK;
}

// Host side of bb_net_load: fold BN (training=False, eps 1e-3) into each convolution and lay the
// tower out as [layer][tap][ci][16].  Flat input order (oracle/network_np.py init_weights):
//   conv_bn := w[k,k,ci,co] b[co] gamma[co] beta[co] mean[co] var[co]
//   stem, block{i}.c1, block{i}.c2 ..., value.conv, value.d1.w[1,D] .b[D], value.d2.w[D,1] .b[1],
//   policy.conv, policy.d.w[2,A] .b[A]
template <class G>
static int net_build(NetDev* net, const bb_net_desc* d, const float* w, size_t n_floats) {
  using Ge = NetGeom<G>;
  const int F = d->filters, D = d->dense, blocks = d->blocks, cin = d->cin, A = G::A;
  if (F != NET_F) return net_build_wide<G>(net, d, w, n_floats);
  if (cin != G::C) { set_error("cin %d does not match the game's %d input planes", cin, G::C); return BB_EINVAL; }
  if (D < 1 || D > NET_MAX_DENSE || blocks < 0) { set_error("dense must be in [1,%d], blocks >= 0", NET_MAX_DENSE); return BB_EINVAL; }
  auto conv_sz = [](int k, int ci, int co) { return (size_t)k * k * ci * co + 5 * (size_t)co; };
  const size_t need = conv_sz(3, cin, F) + 2 * (size_t)blocks * conv_sz(3, F, F) + conv_sz(1, F, 1) + D + D + D + 1 +
                      conv_sz(1, F, 2) + 2 * (size_t)A + A;
  if (n_floats != need) { set_error("bb_net_load: expected %zu floats for this shape, got %zu", need, n_floats); return BB_EINVAL; }
  const int layers = 1 + 2 * blocks;
  std::vector<float> tw(Ge::STEM_W + (size_t)(layers - 1) * Ge::LAYER_W, 0.f), tb((size_t)layers * NET_F);
  const float* p = w;
  auto fold = [&](int k, int ci, int co, int ci_pad, float* wout, float* bout) {
    const float* cw = p; p += (size_t)k * k * ci * co;
    const float *b = p, *g = p + co, *be = p + 2 * co, *mu = p + 3 * co, *var = p + 4 * co;
    p += 5 * (size_t)co;
    for (int o = 0; o < co; ++o) {
      const double s = (double)g[o] / sqrt((double)var[o] + 1e-3);
      bout[o] = (float)(((double)b[o] - (double)mu[o]) * s + (double)be[o]);
      for (int t = 0; t < k * k; ++t)
        for (int i = 0; i < ci; ++i) wout[((size_t)t * ci_pad + i) * co + o] = (float)((double)cw[((size_t)t * ci + i) * co + o] * s);
    }
  };
  fold(3, cin, F, Ge::CIN_PAD, tw.data(), tb.data());
  for (int l = 1; l < layers; ++l) fold(3, F, F, F, tw.data() + Ge::STEM_W + (size_t)(l - 1) * Ge::LAYER_W, tb.data() + (size_t)l * NET_F);
  std::vector<float> head(17 + 34 + 3 * (size_t)D + 1);
  fold(1, F, 1, F, head.data(), head.data() + 16);                       // value conv: w[16], b
  memcpy(head.data() + 17 + 34, p, sizeof(float) * (3 * (size_t)D + 1)); // d1.w, d1.b, d2.w, d2.b
  p += 3 * (size_t)D + 1;
  fold(1, F, 2, F, head.data() + 17, head.data() + 17 + 32);             // policy conv: w[16][2], b[2]
  const float* pw = p;
  const float* pb = p + 2 * (size_t)A;
  auto upload = [&](const float* src, size_t cnt, const float** dst) -> int {
    void* q = nullptr;
    BB_CUDA(cudaMalloc(&q, cnt * sizeof(float)));
    net->allocs.push_back(q);
    BB_CUDA(cudaMemcpy(q, src, cnt * sizeof(float), cudaMemcpyHostToDevice));
    *dst = (const float*)q;
    return BB_OK;
  };
  int rc;
  if ((rc = upload(tw.data(), tw.size(), &net->p.tower_w))) return rc;
  if ((rc = upload(tb.data(), tb.size(), &net->p.tower_b))) return rc;
  if ((rc = upload(head.data(), head.size(), &net->p.head))) return rc;
  net->use_tc = d->precision == 1;
  if (net->use_tc) {
    // bf16 B operands: per (layer, k-block, tap) one 16x16 block, element (n = cout, k = cin) at byte
    // (k / 8) * 256 + n * 16 + (k % 8) * 2  (K-major, SWIZZLE_NONE: 8x8 core matrices of 128 bytes)
    using Tg = TcGeom<G>;
    const int blocks_n = Tg::KB_IN * 9 + (layers - 1) * 9;
    std::vector<uint16_t> wb((size_t)blocks_n * 256, 0);
    auto bf16_rn = [](float f) -> uint16_t {
      uint32_t u;
      memcpy(&u, &f, 4);
      if ((u & 0x7F800000u) == 0x7F800000u) return (uint16_t)(u >> 16);
      u += 0x7FFFu + ((u >> 16) & 1u);
      return (uint16_t)(u >> 16);
    };
    for (int l = 0; l < layers; ++l) {
      const int kbs = l == 0 ? Tg::KB_IN : 1, ci_n = l == 0 ? cin : F, ci_pad = l == 0 ? Ge::CIN_PAD : F;
      const float* lw = l == 0 ? tw.data() : tw.data() + Ge::STEM_W + (size_t)(l - 1) * Ge::LAYER_W;
      const int blk0 = l == 0 ? 0 : Tg::KB_IN * 9 + (l - 1) * 9;
      for (int kb = 0; kb < kbs; ++kb)
        for (int t = 0; t < 9; ++t)
          for (int k = 0; k < 16; ++k)
            for (int o = 0; o < 16; ++o) {
              const int ci = kb * 16 + k;
              const float v = ci < ci_n ? lw[((size_t)t * ci_pad + ci) * F + o] : 0.f;
              wb[(size_t)(blk0 + kb * 9 + t) * 256 + (k / 8) * 128 + o * 8 + (k % 8)] = bf16_rn(v);
            }
    }
    void* q = nullptr;
    BB_CUDA(cudaMalloc(&q, wb.size() * 2));
    net->allocs.push_back(q);
    BB_CUDA(cudaMemcpy(q, wb.data(), wb.size() * 2, cudaMemcpyHostToDevice));
    net->tc.w = (const uint4*)q;
    net->tc.w_blocks = blocks_n;
    if (TcSmem<G>(blocks_n, layers).total > 227 * 1024) { set_error("network too deep for the smem-resident tensor-core kernel (%d blocks)", blocks); return BB_EINVAL; }
  }
  if ((rc = upload(pw, 2 * (size_t)A, &net->p.pol_w))) return rc;
  if ((rc = upload(pb, (size_t)A, &net->p.pol_b))) return rc;
  void* ctr = nullptr;
  BB_CUDA(cudaMalloc(&ctr, 8));
  BB_CUDA(cudaMemset(ctr, 0, 8));
  net->allocs.push_back(ctr);
  net->p.call_counter = (unsigned long long*)ctr;
  net->p.layers = layers;
  net->p.dense = D;
  net->p.eps = d->epsilon;
  net->p.alpha = d->alpha;
  net->p.seed = d->noise_seed;
  if (net_smem_bytes<G>(layers) > 227 * 1024) { set_error("network too deep for the smem-resident kernel (%d blocks)", blocks); return BB_EINVAL; }
  return BB_OK;
}

__global__ void k_beta_fill(float alpha, unsigned long long key, int n, float* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = beta_noise(alpha, key, (unsigned long long)i);
}

int net_create(int game, const bb_net_desc* d, const float* w, size_t n_floats, NetDev** out) {
  if (!d || !w || !out) { set_error("bb_net_load: null argument"); return BB_EINVAL; }
  NetDev* net = new NetDev();
  net->game = game;
  net->desc = *d;
  int dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&net->sm_count, cudaDevAttrMultiProcessorCount, dev);
  int rc = BB_EINVAL;
  switch (game) {
    case BB_TICTACTOE: rc = net_build<TicTacToe>(net, d, w, n_floats); break;
    case BB_CONNECT4: rc = net_build<Connect4>(net, d, w, n_floats); break;
    case BB_DRAGONCHESS: rc = net_build<DragonChess>(net, d, w, n_floats); break;
    default: set_error("unknown game id %d", game);
  }
  if (rc != BB_OK) { net_destroy(net); return rc; }
  *out = net;
  return BB_OK;
}

void net_destroy(NetDev* net) {
  if (!net) return;
  for (void* p : net->allocs) cudaFree(p);
  delete net;
}

}  // namespace bb

extern "C" int bb_debug_beta_noise(float alpha, uint64_t seed, int n, float* out_dev, void* stream) {
  if (!(alpha > 0.f && alpha < 1.f) || n < 0 || !out_dev) { bb::set_error("bb_debug_beta_noise: alpha in (0,1), n >= 0"); return BB_EINVAL; }
  if (n == 0) return BB_OK;
  bb::k_beta_fill<<<bb::grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(alpha, seed, n, out_dev);
  BB_CUDA(cudaGetLastError());
  return BB_OK;
}
