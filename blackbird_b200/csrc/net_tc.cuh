// Policy-value network forward on the 5th-generation tensor cores (tcgen05 + TMEM), bf16 operands,
// fp32 accumulation.  Included by net.cu (shares NetParams / heads_finish with the fp32 kernel).
//
// The tower (NetworkFactory.py:40-103: conv3x3+BN+ReLU stem, `blocks` residual blocks) is nine
// implicit GEMMs per batch of NP positions, all operands resident in shared memory:
//
//   rows    = board cells in a padded raster: a position is (H+1) lines of (W+1) cells, line 0 and
//             the last column are zero padding shared with the neighbouring line / position, so
//             the 3x3 neighbour (dy,dx) of a row is simply row + dy*(W+1) + dx and always lands on
//             a real cell or on a zero row (no bounds tests, no im2col copy);
//   A tile  = 128 consecutive rows x 16 input channels, bf16, K-major, SWIZZLE_NONE: the buffer is
//             [k-half][row][8 channels] so a row is 16 bytes and every tap of the convolution is
//             the SAME buffer addressed through a descriptor whose start address is shifted by
//             (dy*(W+1)+dx)*16 bytes;
//   B tile  = one tap's 16x16 weight block (BN folded on the host), bf16, K-major;
//   D tile  = 128 rows x 16 output channels, fp32, in TMEM (16 columns per tile).
//
// One thread issues the 9 (x k-blocks) tcgen05.mma of every tile of the layer and commits each tile
// to its own mbarrier; all 8 warps then run the epilogue tile by tile while later tiles are still
// in the tensor pipe: tcgen05.ld (32 lanes x 16 columns per warp), + bias, (+ skip), ReLU, pack to
// bf16, two 16-byte st.shared per row into the next layer's A buffer.  The last layer's epilogue
// feeds the heads from the fp32 registers (1x1 convs -> per-cell ReLU -> fixed-order sums).
#pragma once
#include <cuda_bf16.h>

namespace bb {

constexpr int TC_THREADS = 256;
constexpr int TC_MARGIN = 16;            // zero rows in front of / behind the raster (|tap offset| <= W+2)

template <class G> struct TcBatch;
template <> struct TcBatch<TicTacToe> { static constexpr int NP = 32; };     // 32 * 16 rows = 4 tiles
template <> struct TcBatch<Connect4> { static constexpr int NP = 16; };      // 16 * 56 rows = 7 tiles
template <> struct TcBatch<DragonChess> { static constexpr int NP = 8; };    // 8 * 81 rows  = 6 tiles (5.06)

template <class G>
struct TcGeom {
  static constexpr int NP = TcBatch<G>::NP;
  static constexpr int LW = G::W + 1;                       // line width
  static constexpr int RPP = (G::H + 1) * LW;               // rows per position
  static constexpr int R = NP * RPP;                        // raster rows of a batch
  static constexpr int TILES = (R + 127) / 128;
  static constexpr int ROWS = TILES * 128 + 2 * TC_MARGIN;  // rows allocated per k-half plane
  static constexpr int PLANE = ROWS * 16;                   // bytes per k-half plane
  static constexpr int BUF = 2 * PLANE;                     // one 16-channel activation buffer
  static constexpr int KB_IN = (G::C + 15) / 16;            // 16-channel blocks of the input planes
  static constexpr int TMEM_COLS = TILES * 16 <= 32 ? 32 : (TILES * 16 <= 64 ? 64 : (TILES * 16 <= 128 ? 128 : 256));
  static_assert(LW + 1 <= TC_MARGIN, "tap offsets must stay inside the margin");
  static_assert(TILES * 16 <= 256, "too many tiles for TMEM");
  static_assert(TILES <= 8, "one mbarrier per tile");
};

// ---- PTX wrappers ------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem], M=128, N=16, K=16, bf16 x bf16 -> fp32
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// all MMAs issued so far by this thread -> one arrival on `bar` when they have completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// this warp's 32 TMEM lanes x 16 consecutive fp32 columns -> 16 registers per thread
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
      "tcgen05.wait::ld.sync.aligned;\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// Shared-memory matrix descriptor, K-major, SWIZZLE_NONE (cute::UMMA::SmemDescriptor bit layout):
// [0,14) start >> 4 | [16,30) leading-dim byte offset >> 4 (between the two 8-element K halves)
// | [32,46) stride byte offset >> 4 (between 8-row groups) | [46,48) version = 1 | [61,64) layout = 0
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): D=f32 [4,6)=1, A=bf16 [7,10)=1, B=bf16 [10,13)=1,
// A and B K-major (bits 15,16 = 0), N>>3 at [17,23), M>>4 at [24,29)
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float2 unpack_bf16(uint32_t u) {
  __nv_bfloat162 h = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(h);
}

template <class G>
struct TcSmem {                                  // byte offsets into dynamic shared memory
  using Ge = TcGeom<G>;
  int w, bias, x, y, in, hsum, red, sst, mbar, tmem, total;
  __host__ __device__ TcSmem(int w_blocks, int layers) {
    int o = 0;
    w = o; o += w_blocks * 512;
    bias = o; o += layers * 16 * 4;
    o = (o + 127) & ~127;
    x = o; o += Ge::BUF;
    y = o; o += Ge::BUF;
    if (Ge::KB_IN == 1) in = y; else { in = o; o += Ge::KB_IN * Ge::BUF; }
    hsum = o; o += Ge::NP * 4 * 4;
    red = o; o += 16 * 4;
    sst = o; o += Ge::NP * (int)sizeof(typename G::State);
    o = (o + 7) & ~7;
    mbar = o; o += 8 * 8;
    tmem = o; o += 16;
    total = o;
  }
};

template <class G>
__global__ void __launch_bounds__(TC_THREADS)
k_net_forward_tc(NetParams np, TcParams tp, const typename G::State* __restrict__ states, int n_max,
                 const int* __restrict__ count_dev, float* __restrict__ policy, float* __restrict__ value) {
  using Ge = TcGeom<G>;
  using State = typename G::State;
  extern __shared__ __align__(128) uint8_t tc_smem[];
  const int n = count_dev ? min(n_max, *count_dev) : n_max;
  const int n_batches = (n + Ge::NP - 1) / Ge::NP;
  if ((int)blockIdx.x >= n_batches) return;                 // before any allocation
  const TcSmem<G> L(tp.w_blocks, np.layers);
  uint8_t* sW = tc_smem + L.w;
  float* sBias = reinterpret_cast<float*>(tc_smem + L.bias);
  uint8_t* bufX = tc_smem + L.x;
  uint8_t* bufY = tc_smem + L.y;
  uint8_t* bufIn = tc_smem + L.in;
  float* hsum = reinterpret_cast<float*>(tc_smem + L.hsum);
  float* red = reinterpret_cast<float*>(tc_smem + L.red);
  State* sst = reinterpret_cast<State*>(tc_smem + L.sst);
  uint64_t* mbar = reinterpret_cast<uint64_t*>(tc_smem + L.mbar);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tc_smem + L.tmem);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int q = warp & 3, hsel = warp >> 2;                 // TMEM lane quarter of this warp; tile parity it serves

  // one-time setup: weights + biases resident, activation buffers (and their zero padding) cleared
  for (int i = tid; i < tp.w_blocks * 32; i += TC_THREADS) reinterpret_cast<uint4*>(sW)[i] = tp.w[i];
  for (int i = tid; i < np.layers * 16; i += TC_THREADS) sBias[i] = np.tower_b[i];
  {
    const int zero_bytes = L.hsum - L.x;                    // X, Y (and the separate input buffer)
    for (int i = tid; i < zero_bytes / 16; i += TC_THREADS) reinterpret_cast<uint4*>(bufX)[i] = make_uint4(0, 0, 0, 0);
  }
  __syncwarp();
  if (warp == 0) tmem_alloc(tmem_slot, Ge::TMEM_COLS);
  if (tid == 0) {
    for (int t = 0; t < Ge::TILES; ++t) mbar_init(&mbar[t], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  tc_fence_before();
  fence_async_smem();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  constexpr uint32_t IDESC = umma_idesc_bf16(128, 16);
  const float* hd = np.head;
  unsigned long long call = 0;
  if (np.eps > 0.f) call = *np.call_counter;
  uint32_t phase = 0;

  for (int batch = blockIdx.x; batch < n_batches; batch += gridDim.x) {
    const int base = batch * Ge::NP;
    __syncthreads();
    if (tid < Ge::NP) sst[tid] = states[min(base + tid, n - 1)];
    __syncthreads();
    // input planes (AsInputArray) decoded straight from the packed states into the raster
    for (int row = tid; row < Ge::R; row += TC_THREADS) {
      const int p = row / Ge::RPP, l = row % Ge::RPP, line = l / Ge::LW, col = l % Ge::LW;
      if (line >= 1 && col < G::W) {
        const int cell = (line - 1) * G::W + col;
        const State s = sst[p];
#pragma unroll
        for (int kb = 0; kb < Ge::KB_IN; ++kb)
#pragma unroll
          for (int kh = 0; kh < 2; ++kh) {
            uint32_t w4[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int c0 = kb * 16 + kh * 8 + 2 * i;
              const float a = c0 < G::C ? (float)G::plane(s, cell, c0) : 0.f;
              const float b = c0 + 1 < G::C ? (float)G::plane(s, cell, c0 + 1) : 0.f;
              w4[i] = pack_bf16(a, b);
            }
            *reinterpret_cast<uint4*>(bufIn + kb * Ge::BUF + kh * Ge::PLANE + (row + TC_MARGIN) * 16) =
                make_uint4(w4[0], w4[1], w4[2], w4[3]);
          }
      }
    }
    fence_async_smem();
    __syncthreads();

    for (int layer = 0; layer < np.layers; ++layer) {
      const bool odd = layer & 1;
      uint8_t* in = layer == 0 ? bufIn : (odd ? bufX : bufY);
      uint8_t* out = odd ? bufY : bufX;
      const bool skip = !odd && layer > 0;                  // conv_2 + BN + block input (NetworkFactory.py:95-98)
      const bool last = layer == np.layers - 1;
      if (tid == 0) {                                       // the whole layer's MMAs, one commit per tile
        tc_fence_after();
        const int kblocks = layer == 0 ? Ge::KB_IN : 1;
        const int wblk0 = layer == 0 ? 0 : Ge::KB_IN * 9 + (layer - 1) * 9;
        const uint32_t in_addr = smem_u32(in), w_addr = smem_u32(sW);
        for (int t = 0; t < Ge::TILES; ++t) {
          for (int kb = 0; kb < kblocks; ++kb)
#pragma unroll
            for (int tap = 0; tap < 9; ++tap) {
              const int off = (tap / 3 - 1) * Ge::LW + (tap % 3 - 1);
              const uint64_t adesc = umma_desc(in_addr + kb * Ge::BUF + (t * 128 + TC_MARGIN + off) * 16, Ge::PLANE, 128);
              const uint64_t bdesc = umma_desc(w_addr + (wblk0 + kb * 9 + tap) * 512, 256, 128);
              umma_bf16(tmem_base + t * 16, adesc, bdesc, IDESC, (kb | tap) != 0);
            }
          umma_commit(&mbar[t]);
        }
      }
      __syncwarp();
      const float* bias = sBias + layer * 16;
      for (int t = hsel; t < Ge::TILES; t += 2) {
        mbar_wait(&mbar[t], phase);
        tc_fence_after();
        float acc[16];
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + t * 16, acc);
        const int row = t * 128 + q * 32 + lane;
        const int l = row % Ge::RPP, line = l / Ge::LW, col = l % Ge::LW;
        if (row < Ge::R && line >= 1 && col < G::W) {
          uint8_t* slot = out + (row + TC_MARGIN) * 16;
#pragma unroll
          for (int c = 0; c < 16; ++c) acc[c] += bias[c];
          if (skip) {
            const uint4 s0 = *reinterpret_cast<const uint4*>(slot), s1 = *reinterpret_cast<const uint4*>(slot + Ge::PLANE);
            const uint32_t sw[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float2 f = unpack_bf16(sw[i]);
              acc[2 * i] += f.x;
              acc[2 * i + 1] += f.y;
            }
          }
#pragma unroll
          for (int c = 0; c < 16; ++c) acc[c] = fmaxf(acc[c], 0.f);
          if (!last) {
            *reinterpret_cast<uint4*>(slot) = make_uint4(pack_bf16(acc[0], acc[1]), pack_bf16(acc[2], acc[3]),
                                                         pack_bf16(acc[4], acc[5]), pack_bf16(acc[6], acc[7]));
            *reinterpret_cast<uint4*>(slot + Ge::PLANE) = make_uint4(pack_bf16(acc[8], acc[9]), pack_bf16(acc[10], acc[11]),
                                                                     pack_bf16(acc[12], acc[13]), pack_bf16(acc[14], acc[15]));
          } else {
            // heads from the fp32 registers: 1x1 convs (1 value + 2 policy filters) + BN + ReLU; the three
            // per-cell values replace this row's (now dead) slot in X
            float v = hd[16], p0 = hd[17 + 32], p1 = hd[17 + 33];
#pragma unroll
            for (int c = 0; c < 16; ++c) {
              v = fmaf(acc[c], hd[c], v);
              p0 = fmaf(acc[c], hd[17 + c * 2], p0);
              p1 = fmaf(acc[c], hd[17 + c * 2 + 1], p1);
            }
            *reinterpret_cast<float4*>(slot) = make_float4(fmaxf(v, 0.f), fmaxf(p0, 0.f), fmaxf(p1, 0.f), 0.f);
          }
        }
      }
      phase ^= 1u;
      tc_fence_before();
      fence_async_smem();
      __syncthreads();
    }
    // sum over H,W in a fixed order, then the dense heads
    if (tid < Ge::NP * 3) {
      const int p = tid / 3, k = tid % 3;
      float s = 0.f;
      for (int y = 0; y < G::H; ++y)
        for (int x = 0; x < G::W; ++x)
          s += reinterpret_cast<const float*>(bufX + (p * Ge::RPP + (y + 1) * Ge::LW + x + TC_MARGIN) * 16)[k];
      hsum[p * 4 + k] = s;
    }
    __syncthreads();
    heads_finish<G>(np, hsum, red, base, n, Ge::NP, call, policy, value);
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, Ge::TMEM_COLS);
}

}  // namespace bb
