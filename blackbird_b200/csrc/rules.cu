// Batched game-rule kernels + their C-ABI entry points (include/bbengine.h, "rule kernels").
// One thread per (state, action-row); HBM-bound byte work, coalesced over states.
#include <stdarg.h>

#include "common.cuh"

namespace bb {

static thread_local std::string g_error;
void set_error(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_error = buf;
}

template <class G>
__global__ void k_init(typename G::State* s, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) s[i] = G::initial();
}

template <class G>
__global__ void k_legal(const typename G::State* s, int n, double* out) {
  long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)n * G::NROWS) return;
  int i = (int)(t / G::NROWS), row = (int)(t % G::NROWS);
  typename G::State st = s[i];
  uint64_t m = G::legal_row(st, row);
  double* o = out + (size_t)i * G::A;
  if (G::NROWS == 1) {
    for (int a = 0; a < G::A; ++a) o[a] = (double)((m >> a) & 1ull);
  } else {
    for (int to = 0; to < 64; ++to)
      if (to != row) o[G::action_of(row, to)] = (double)((m >> to) & 1ull);
  }
}

template <class G>
__global__ void k_apply(typename G::State* s, const int32_t* actions, int n, uint8_t* ok) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  typename G::State st = s[i];
  bool good = G::apply(st, actions[i]);
  if (good) s[i] = st;
  ok[i] = good ? 1 : 0;
}

template <class G>
__global__ void k_winner(const typename G::State* s, const int32_t* prev, int n, int32_t* out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[i] = G::winner(s[i], prev ? prev[i] : -1);
}

template <class G>
__global__ void k_planes(const typename G::State* s, int n, int8_t* out) {
  constexpr int PL = G::H * G::W * G::C;
  long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)n * PL) return;
  int i = (int)(t / PL), j = (int)(t % PL);
  out[t] = (int8_t)G::plane(s[i], j / G::C, j % G::C);
}

__device__ __forceinline__ bool same(const TicTacToe::State& a, const TicTacToe::State& b) {
  return ((a.v ^ b.v) & 0x7FFFFu) == 0;                          // board + Player, not PreviousPlayer
}
__device__ __forceinline__ bool same(const Connect4::State& a, const Connect4::State& b) {
  const uint64_t M = Connect4::BOARD | (1ull << 56);
  return ((a.x ^ b.x) & Connect4::BOARD) == 0 && ((a.o ^ b.o) & M) == 0;
}
__device__ __forceinline__ bool same(const DragonChess::State& a, const DragonChess::State& b) {
  bool e = ((a.meta ^ b.meta) & 0x1Fu) == 0;                     // castle flags + Player (A.21)
#pragma unroll
  for (int r = 0; r < 8; ++r) e &= a.row[r] == b.row[r];
  return e;
}
template <class G>
__global__ void k_equal(const typename G::State* a, const typename G::State* b, int n, uint8_t* out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = same(a[i], b[i]) ? 1 : 0;
}

__device__ __forceinline__ void pack_cells(TicTacToe::State& s, const int8_t* c, int pl, int pv, int) {
  s.v = 0;
  for (int k = 0; k < 9; ++k) s.v |= (c[k] == 1 ? 1u << k : 0u) | (c[k] == 2 ? 1u << (k + 9) : 0u);
  TicTacToe::set_meta(s, pl, pv);
}
__device__ __forceinline__ void pack_cells(Connect4::State& s, const int8_t* c, int pl, int pv, int) {
  s.x = s.o = 0;
  for (int k = 0; k < 42; ++k) {
    uint64_t bit = 1ull << ((k % 7) * 7 + k / 7);
    if (c[k] == 1) s.x |= bit;
    if (c[k] == 2) s.o |= bit;
  }
  Connect4::set_meta(s, pl, pv);
}
__device__ __forceinline__ void pack_cells(DragonChess::State& s, const int8_t* c, int pl, int pv, int castle) {
  for (int r = 0; r < 8; ++r) {
    uint32_t w = 0;
    for (int col = 0; col < 8; ++col) {
      int v = c[r * 8 + col];
      uint32_t nib = v > 0 ? (uint32_t)v : (v < 0 ? (uint32_t)(8 - v) : 0u);
      w |= nib << (4 * col);
    }
    s.row[r] = w;
  }
  s.meta = (uint32_t)(castle & 0xF) | ((uint32_t)(pl - 1) << 4) | ((uint32_t)pv << 5);
  s.pad = 0;
}
template <class G>
__global__ void k_pack(const int8_t* cells, const uint8_t* pl, const uint8_t* pv, const uint8_t* castle,
                       int n, typename G::State* out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  typename G::State s;
  pack_cells(s, cells + (size_t)i * G::H * G::W, pl[i], pv ? pv[i] : 0, castle ? castle[i] : 0);
  out[i] = s;
}

__device__ __forceinline__ int cell_code(const TicTacToe::State& s, int k) {
  return ((s.v >> k) & 1) ? 1 : (((s.v >> (k + 9)) & 1) ? 2 : 0);
}
__device__ __forceinline__ int cell_code(const Connect4::State& s, int k) {
  int bit = (k % 7) * 7 + k / 7;
  return ((s.x >> bit) & 1) ? 1 : (((s.o >> bit) & 1) ? 2 : 0);
}
__device__ __forceinline__ int cell_code(const DragonChess::State& s, int k) {
  int v = DragonChess::at(s, k >> 3, k & 7);
  return v >= 9 ? -(v - 8) : v;
}
__device__ __forceinline__ int castle_bits(const TicTacToe::State&) { return 0; }
__device__ __forceinline__ int castle_bits(const Connect4::State&) { return 0; }
__device__ __forceinline__ int castle_bits(const DragonChess::State& s) { return s.meta & 0xF; }
template <class G>
__global__ void k_unpack(const typename G::State* s, int n, int8_t* cells, uint8_t* pl, uint8_t* pv, uint8_t* castle) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  typename G::State st = s[i];
  if (cells)
    for (int k = 0; k < G::H * G::W; ++k) cells[(size_t)i * G::H * G::W + k] = (int8_t)cell_code(st, k);
  if (pl) pl[i] = (uint8_t)G::player(st);
  if (pv) pv[i] = (uint8_t)G::prev(st);
  if (castle) castle[i] = (uint8_t)castle_bits(st);
}

// Appendix B.2 integer evaluator: policy f32[A] = f32(n_i)/f32(sum n), value = f32((h%1009)*2-1008)/f32(1009)
template <class G>
__global__ void k_kat(const typename G::State* s, int n, const int* count_dev, float* policy, float* value) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (count_dev) n = min(n, *count_dev);
  if (i >= n) return;
  typename G::State st = s[i];
  unsigned long long h = kat_hash<G>(st);
  long long total = 0;
  for (int a = 0; a < G::A; ++a) total += 1 + (long long)((h * (unsigned long long)(a + 1)) % 97ull);
  float denom = (float)total;
  float* p = policy + (size_t)i * G::A;
  for (int a = 0; a < G::A; ++a) p[a] = __fdiv_rn((float)(1 + (int)((h * (unsigned long long)(a + 1)) % 97ull)), denom);
  value[i] = __fdiv_rn((float)((long long)(h % 1009ull) * 2 - 1008), 1009.0f);
}

int launch_kat(int game, const void* states, int n, const int* count_dev, float* policy, float* value, cudaStream_t stream) {
  if (n <= 0) return BB_OK;
  BB_DISPATCH_GAME(game, (k_kat<G><<<grid_for(n, 128), 128, 0, stream>>>((const typename G::State*)states, n, count_dev,
                                                                        policy, value)));
  BB_CUDA(cudaGetLastError());
  return BB_OK;
}

}  // namespace bb

using namespace bb;

extern "C" {

const char* bb_last_error(void) { return g_error.c_str(); }
int bb_abi_version(void) { return 1; }

int bb_game_state_bytes(int game) { BB_DISPATCH_GAME(game, return (int)sizeof(typename G::State)); return BB_EINVAL; }
int bb_game_actions(int game) { BB_DISPATCH_GAME(game, return G::A); return BB_EINVAL; }
int bb_game_plane_shape(int game, int* h, int* w, int* c) {
  BB_DISPATCH_GAME(game, { *h = G::H; *w = G::W; *c = G::C; });
  return BB_OK;
}

#define BB_LAUNCH_CHECK() BB_CUDA(cudaGetLastError())

int bb_game_init(int game, void* states, int n, void* stream) {
  if (n <= 0) return BB_OK;
  BB_DISPATCH_GAME(game, (k_init<G><<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>((typename G::State*)states, n)));
  BB_LAUNCH_CHECK();
  return BB_OK;
}
int bb_game_legal(int game, const void* states, int n, double* out, void* stream) {
  if (n <= 0) return BB_OK;
  BB_DISPATCH_GAME(game, (k_legal<G><<<grid_for((long long)n * G::NROWS, 256), 256, 0, (cudaStream_t)stream>>>(
                             (const typename G::State*)states, n, out)));
  BB_LAUNCH_CHECK();
  return BB_OK;
}
int bb_game_apply(int game, void* states, const int32_t* actions, int n, uint8_t* ok, void* stream) {
  if (n <= 0) return BB_OK;
  BB_DISPATCH_GAME(game, (k_apply<G><<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>((typename G::State*)states, actions, n, ok)));
  BB_LAUNCH_CHECK();
  return BB_OK;
}
int bb_game_winner(int game, const void* states, const int32_t* prev, int n, int32_t* out, void* stream) {
  if (n <= 0) return BB_OK;
  BB_DISPATCH_GAME(game, (k_winner<G><<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>((const typename G::State*)states, prev, n, out)));
  BB_LAUNCH_CHECK();
  return BB_OK;
}
int bb_game_planes(int game, const void* states, int n, int8_t* out, void* stream) {
  if (n <= 0) return BB_OK;
  BB_DISPATCH_GAME(game, (k_planes<G><<<grid_for((long long)n * G::H * G::W * G::C, 256), 256, 0, (cudaStream_t)stream>>>(
                             (const typename G::State*)states, n, out)));
  BB_LAUNCH_CHECK();
  return BB_OK;
}
int bb_game_equal(int game, const void* a, const void* b, int n, uint8_t* out, void* stream) {
  if (n <= 0) return BB_OK;
  BB_DISPATCH_GAME(game, (k_equal<G><<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(
                             (const typename G::State*)a, (const typename G::State*)b, n, out)));
  BB_LAUNCH_CHECK();
  return BB_OK;
}
int bb_game_pack(int game, const int8_t* cells, const uint8_t* player, const uint8_t* prev, const uint8_t* castle,
                 int n, void* states, void* stream) {
  if (n <= 0) return BB_OK;
  BB_DISPATCH_GAME(game, (k_pack<G><<<grid_for(n, 128), 128, 0, (cudaStream_t)stream>>>(cells, player, prev, castle, n,
                                                                                      (typename G::State*)states)));
  BB_LAUNCH_CHECK();
  return BB_OK;
}
int bb_game_unpack(int game, const void* states, int n, int8_t* cells, uint8_t* player, uint8_t* prev, uint8_t* castle,
                   void* stream) {
  if (n <= 0) return BB_OK;
  BB_DISPATCH_GAME(game, (k_unpack<G><<<grid_for(n, 128), 128, 0, (cudaStream_t)stream>>>((const typename G::State*)states, n,
                                                                                        cells, player, prev, castle)));
  BB_LAUNCH_CHECK();
  return BB_OK;
}
int bb_kat_forward(int game, const void* states, int n, float* policy, float* value, void* stream) {
  return launch_kat(game, states, n, nullptr, policy, value, (cudaStream_t)stream);
}

}  // extern "C"
