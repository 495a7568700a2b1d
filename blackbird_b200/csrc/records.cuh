// Device-side ExampleState.SerializeState (Blackbird.py:68-81; proto/state.proto:3-9) with the outcome labels of
// GenerateTrainingSamples (Blackbird.py:260-264): the engine's raw record buffers (packed root states + sparse child
// visit counts per move) become the final protobuf bytes on the GPU, game after game in one compact stream, so the
// host copies (and the ranks all-gather) live bytes only and no per-record Python work is left.
//
//   State { float mctsEval = 1; bytes mctsPolicy = 2; bytes boardEncoding = 3; bytes boardDims = 4; bytes policyDims = 5; }
//
// proto3, fields in number order; mctsEval is omitted when it is 0.0 (draws and games cut by the ply cap);
// mctsPolicy = Node.ChildProbability() as raw little-endian f64[A] (plays / builtin-sum(plays), MCTS.py:66-68; zeros for
// the terminal record, Blackbird.py:256); boardEncoding = AsInputArray() int8 [1,H,W,C]; boardDims = int8 [1,H,W,C];
// policyDims = int8 [A] (DragonChess: 4032 wraps to 0xC0).  Connect4: 200 bytes with a label, 195 without.
#pragma once
#include "tree.cuh"

namespace bb {

__host__ __device__ constexpr int varint_len(unsigned v) { return v < (1u << 7) ? 1 : v < (1u << 14) ? 2 : v < (1u << 21) ? 3 : 4; }

template <class G>
struct RecFmt {
  static constexpr int PB = 8 * G::A;                         // mctsPolicy payload
  static constexpr int LB = G::H * G::W * G::C;               // boardEncoding payload
  static constexpr int HDR2 = 1 + varint_len(PB);             // tag 0x12 + length
  static constexpr int HDR3 = 1 + varint_len(LB);             // tag 0x1a + length
  static constexpr int TAIL = 6 + 3;                          // 22 04 01 H W C | 2a 01 A
  static constexpr int SIZE0 = HDR2 + PB + HDR3 + LB + TAIL;  // record without mctsEval; with it: + 5
};

__device__ __forceinline__ int put_varint(uint8_t* p, unsigned v) {
  int n = 0;
  while (v >= 0x80u) { p[n++] = (uint8_t)(v | 0x80u); v >>= 7; }
  p[n++] = (uint8_t)v;
  return n;
}

// One CTA per game row.  game_off[row] = first byte of the game's records in `out`; a game's records all have the same
// size (the label is per game up to its sign), so record r starts at game_off[row] + r * size.
template <class G>
__global__ void __launch_bounds__(128) k_pack_records(EngineDev<G> E, int rows, const unsigned long long* __restrict__ game_off,
                                                      uint8_t* __restrict__ out) {
  using F = RecFmt<G>;
  extern __shared__ __align__(16) uint8_t rec_smem[];
  double* pol = reinterpret_cast<double*>(rec_smem);          // [A]
  uint8_t* hdr = rec_smem + F::PB;                            // 5 (mctsEval) + HDR2 | HDR3 | TAIL
  __shared__ unsigned long long s_total;
  const int row = blockIdx.x, tid = threadIdx.x;
  if (row >= rows) return;
  const int nm = E.n_moves[row], winner = E.winners[row];
  const size_t rbase = (size_t)row * (E.max_plies + 1), obase = (size_t)row * (E.max_plies + 2), pbase = (size_t)row * E.pool_cap;
  const bool labelled = winner == WIN_P1 || winner == WIN_P2;  // draw (0) and "cut by the ply cap" (None) give 0.0: field omitted
  const int h1 = labelled ? 5 : 0, size = F::SIZE0 + h1;
  if (tid == 0) {                                              // the constant parts of this game's records
    uint8_t* p = hdr + 5;
    p[0] = 0x12; put_varint(p + 1, (unsigned)F::PB);
    p = hdr + 5 + F::HDR2;
    p[0] = 0x1a; put_varint(p + 1, (unsigned)F::LB);
    p = hdr + 5 + F::HDR2 + F::HDR3;
    p[0] = 0x22; p[1] = 4; p[2] = 1; p[3] = (uint8_t)G::H; p[4] = (uint8_t)G::W; p[5] = (uint8_t)G::C;
    p[6] = 0x2a; p[7] = 1; p[8] = (uint8_t)(G::A & 0xFF);
    hdr[0] = 0x0d; hdr[1] = 0; hdr[2] = 0; hdr[3] = 0x80;      // +-1.0f little endian: 00 00 80 3f / bf
  }
  for (int i = tid; i < G::A; i += blockDim.x) pol[i] = 0.0;
  __syncthreads();
  uint8_t* dst = out + game_off[row];
  for (int r = 0; r <= nm; ++r) {
    const typename G::State st = E.rec_states[rbase + r];
    const uint32_t lo = E.rec_off[obase + r], hi = r < nm ? E.rec_off[obase + r + 1] : lo;   // terminal record: zero policy
    if (tid == 0) {
      unsigned long long t = 0;
      for (uint32_t j = lo; j < hi; ++j) t += E.pool_plays[pbase + j];                        // exact: integers (builtin sum of f64 ints)
      s_total = t;
      hdr[4] = G::player(st) == winner ? 0x3f : 0xbf;                                        // Blackbird.py:260-264
    }
    __syncthreads();
    const double total = (double)s_total;
    if (s_total > 0)
      for (uint32_t j = lo + tid; j < hi; j += blockDim.x)
        pol[E.pool_actions[pbase + j]] = __ddiv_rn((double)E.pool_plays[pbase + j], total);
    __syncthreads();
    for (int i = tid; i < size; i += blockDim.x) {
      int k = i;
      uint8_t b;
      if (k < h1) b = hdr[k];
      else if ((k -= h1) < F::HDR2) b = hdr[5 + k];
      else if ((k -= F::HDR2) < F::PB) b = rec_smem[k];
      else if ((k -= F::PB) < F::HDR3) b = hdr[5 + F::HDR2 + k];
      else if ((k -= F::HDR3) < F::LB) b = (uint8_t)(int8_t)G::plane(st, k / G::C, k % G::C);
      else b = hdr[5 + F::HDR2 + F::HDR3 + (k - F::LB)];
      dst[i] = b;
    }
    __syncthreads();
    if (s_total > 0)
      for (uint32_t j = lo + tid; j < hi; j += blockDim.x) pol[E.pool_actions[pbase + j]] = 0.0;
    __syncthreads();
    dst += size;
  }
}

template <class G>
static inline size_t pack_records_smem() { return (size_t)RecFmt<G>::PB + 5 + RecFmt<G>::HDR2 + RecFmt<G>::HDR3 + RecFmt<G>::TAIL + 16; }

}  // namespace bb
