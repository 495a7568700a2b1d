// Shared helpers for libbbengine.so translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

#include "../../include/bbengine.h"
#include "games.cuh"

namespace bb {

void set_error(const char* fmt, ...);

#define BB_CUDA(call)                                                                     \
  do {                                                                                    \
    cudaError_t _e = (call);                                                              \
    if (_e != cudaSuccess) {                                                              \
      bb::set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__, cudaGetErrorString(_e)); \
      return BB_ECUDA;                                                                    \
    }                                                                                     \
  } while (0)

#define BB_DISPATCH_GAME(game, ...)                                                       \
  switch (game) {                                                                         \
    case BB_TICTACTOE: { using G = bb::TicTacToe; __VA_ARGS__; } break;                   \
    case BB_CONNECT4: { using G = bb::Connect4; __VA_ARGS__; } break;                     \
    case BB_DRAGONCHESS: { using G = bb::DragonChess; __VA_ARGS__; } break;               \
    default: bb::set_error("unknown game id %d", (int)(game)); return BB_EINVAL;          \
  }

// Appendix B.2 evaluator over n (or min(n, *count_dev)) packed states (rules.cu)
int launch_kat(int game, const void* states, int n, const int* count_dev, float* policy, float* value, cudaStream_t stream);

// Streaming (asynchronous) mode: a persistent search kernel and a persistent network kernel run side by
// side and talk through a request ring in global memory, so a slot that blocks on the evaluator gets its
// answer one network batch later instead of one whole tick later (DESIGN.md "streaming engine").
// Ticket i lives in ring entry i & mask: one 64-bit word, low half = i, high half = slot | RING_PUBLISHED, written
// with one store once the slot's state is in q_state[slot]; resp_seq[slot] == i + 1 publishes the answer.  The
// network pipeline that serves ticket i clears the PUBLISHED bit as soon as it has read the entry, and a search warp
// publishes ticket i + ring only after it has seen that bit clear: open tickets need not be contiguous (a pipeline
// can sit on ticket i for a whole batch while the other slots issue ring-many newer ones), so without that
// hand-back a late ticket's entry is overwritten and its slot waits for ever (seen with 64 slots on a 64-entry
// ring).  The ring is also much larger than the slot count (>= 8192 entries), so the wait is never taken in
// practice.  The state itself lives per slot, where nothing can overwrite it while the request is open.
constexpr unsigned RING_PUBLISHED = 0x80000000u;
constexpr unsigned RING_MIN_ENTRIES = 8192u;
struct StreamCtl {
  unsigned tail;                  // tickets handed out to the search warps
  unsigned head;                  // tickets claimed by the network CTAs
  unsigned stop;                  // search warps leave after their current pass
  unsigned search_ctas;           // grid of the running search kernel
  unsigned search_done;           // search CTAs that have left
  unsigned servers;               // grid of the running network kernel (claims are sized as open / servers)
  unsigned long long goal_moves;  // stop once the engine-wide move counter reaches this (0 = no goal)
  unsigned long long batches;     // network batches served (measurement)
  unsigned long long cyc_claim, cyc_tower, cyc_heads;   // SM cycles the network CTAs spent waiting for work / in the tower / in the heads (measurement)
  unsigned long long cyc_search_busy, cyc_search_total; // search warps: cycles inside advance_slot / resident (measurement)
  unsigned long long search_calls;                      // advance_slot calls (measurement)
#ifdef BB_SEARCH_TIMING
  unsigned long long sdbg[24];
  unsigned long long xdbg[8];     // expand_node: legal rows + scan, list + policy reads, prior sum, node writes; calls
#endif
#ifdef BB_NET_TIMING_HEADS
  unsigned long long hdbg[8];     // legal-only policy head (heads_finish): phase cycles, positions
#endif
#if defined(BB_NET_TIMING_TOWER) || defined(BB_NET_TIMING_HEADS)
  unsigned long long ndbg[8];     // MMA warp: cycles waiting for `ready`, issuing; epilogue warp 0: waiting for `full`, working; staging; tile-layers
#endif
};

// Evaluation cache (streaming mode, action counts up to 9).  The reference memoises both evaluator hooks per state --
// Model.SampleValue and Model.GetPriors are functools.lru_cache(maxsize=4096) methods (Blackbird.py:350,372), shared by
// all games a Model plays -- and Connect4 searches are full of transpositions: 45 % of the nodes one 400-simulation game
// evaluates repeat a state it has evaluated before, and two games share 39 % of their states (measured with the oracle).
// Here the memo is ONE direct-mapped table in HBM shared by every game slot of the engine: 64-byte entries
//   words 0-3 packed state (key) | 4..12 policy f32[A] | 13 value f32 | 14 tag | 15 unused
// tag: 0 empty, 1 locked by a writer, otherwise the upper hash bits of the key (>= 2).  Writers are the network pipelines:
// atomicExch(tag, LOCKED) when the batch is staged (whoever finds it locked already does not insert), fence, payload
// stores with the answers, the batch's publication fence, tag store.  Readers are the search warps: one coalesced
// 64-byte read (lane l = word l), key and tag compared, fence, tag read again -- unchanged means no writer touched
// the payload in between.  With a pure evaluator (epsilon 0) the memo is transparent, so results stay bit-identical
// (streaming == tick mode tests); with policy noise it freezes the first noisy answer of a state while the entry lives,
// which is what the reference's lru_cache does to TensorFlow's noise.
constexpr int ECACHE_WORDS = 16;
constexpr uint32_t ECACHE_LOCKED = 1u;
constexpr int ECACHE_MAX_ACTIONS = 9;
__host__ __device__ __forceinline__ void ecache_hash(const uint32_t k[4], uint32_t mask, uint32_t* idx, uint32_t* tag) {
  unsigned long long h = ((unsigned long long)k[1] << 32 | k[0]) * 0x9E3779B97F4A7C15ull;
  h ^= ((unsigned long long)k[3] << 32 | k[2]) * 0xC2B2AE3D27D4EB4Full + (h >> 29);
  h = (h ^ (h >> 32)) * 0xD6E8FEB86659FD93ull;
  h ^= h >> 32;
  *idx = (uint32_t)h & mask;
  const uint32_t t = (uint32_t)(h >> 32);
  *tag = t < 2u ? t + 2u : t;
}

// Built-in network (net.cu)
struct NetDev;
struct NetServe {                 // the request ring as the persistent network kernel sees it
  StreamCtl* q;
  unsigned long long* q_ent;      // ticket ring (see StreamCtl)
  const void* q_state;            // [n_games] request states, by slot
  unsigned* resp_seq;
  unsigned q_mask;
  uint32_t* ecache;               // evaluation cache (nullptr: off), ecache_mask + 1 entries of ECACHE_WORDS words
  uint32_t ecache_mask;
  int legal_only;                 // wide policy heads: answer at the legal actions of the asked position only (engine-internal calls)
};
// persistent network kernel for streaming mode: serves the ring until the search kernel has left and the ring is empty;
// answers go to policy[slot][A] / value[slot]
int net_serve(NetDev* net, const NetServe& sv, float* policy, float* value, cudaStream_t stream);
int net_sm_count(NetDev* net);
int net_pipes(NetDev* net);
int net_tc_footprint(NetDev* net, int* regs, int* smem);
bool net_is_tc(NetDev* net);
int net_create(int game, const bb_net_desc* d, const float* w_host, size_t n_floats, NetDev** out);
void net_destroy(NetDev* net);
// legal_only (wide heads of the bf16 kernel): the caller reads the policy at the legal actions of each state only (the search does)
int net_forward(NetDev* net, const void* states_dev, int n, const int* count_dev, float* policy, float* value,
                cudaStream_t stream, int legal_only = 0);

// Release / acquire fence at device scope.  __threadfence() is membar.gl = fence.sc.gpu, the sequentially consistent fence (~1.2 k
// cycles measured in the serve-mode publication); every hand-off in the streaming engine is a message-passing pattern
// (payload stores, fence, flag store | flag load, fence, payload loads) that needs no store -> load ordering.
__device__ __forceinline__ void fence_acq_rel() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

static inline unsigned grid_for(long long n, int block) { return (unsigned)((n + block - 1) / block); }

// Appendix B.2 integer evaluator hash over the flattened [1,H,W,C] planes.
template <class G>
__device__ __forceinline__ uint32_t kat_hash(const typename G::State& s) {
  long long acc = 0;
  for (int cell = 0; cell < G::H * G::W; ++cell) {
#pragma unroll
    for (int ch = 0; ch < G::C; ++ch) acc += (long long)(cell * G::C + ch + 1) * G::plane(s, cell, ch);
  }
  long long m = acc % 1000003ll;
  if (m < 0) m += 1000003ll;
  return (uint32_t)m;
}

}  // namespace bb
