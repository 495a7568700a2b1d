// Policy-value network forward for tower widths other than the template's 16 filters (NetworkFactory.py:46-103 takes
// `filters` from the YAML; SURVEY.md 8d row 5a lists 64 and 256).  Included by net.cu after net_tc.cuh.
//
// At 16 filters the whole tower lives in one CTA's shared memory (net_tc.cuh).  At 64 / 256 filters (32 and 128 are built too) one 3x3 layer is
// 74 KB / 1.18 MB of bf16 weights, so the tower runs LAYER BY LAYER: activations stay in HBM / L2 between layers in exactly
// the layout the tensor cores read them in, weights stream through a shared-memory ring.
//
//   activations  act[tile][F / 8][160 rows][8 channels] bf16: a tile is the same padded raster as in net_tc.cuh (PPT whole
//                positions of (H+1) x (W+1) rows, zero line 0 / last column, 16 zero margin rows either side), stored
//                K-major as F / 8 planes of 16-byte rows -- the canonical SWIZZLE_NONE operand layout.  A tile is one
//                contiguous block: ONE cp.async.bulk brings it into shared memory, and the nine taps of the convolution
//                are the same block read through descriptors whose start address is shifted by tap_offset x 16 bytes.
//                Padding rows are zeroed once at allocation and never written.
//   weights      w[layer][pass][k-step][tap][2][N][8] bf16 (BN folded), N = 64 (F = 64, one pass) or 128 (F = 256, two passes
//                over a tile group, each producing half of the output channels): a (k-step, tap) is an N x K16 operand of
//                N x 32 bytes; a ring stage is one bulk copy of three taps (6 / 12 KB).
//   accumulator  128 rows x F fp32 columns of TMEM per tile; a CTA works on a GROUP of tiles at a time (4 at F = 64, 2 at
//                F = 256) so that every weight stage is used by all tiles of the group before it is released.  Only four
//                k-steps of the group's activations are in shared memory at a time (a 40 / 80 KB ring); the rest is the weight ring.
//   kernel       k_wide_conv<F>: warp 0 = TMA producer (k-step by k-step: the group's two planes per tile, then the k-step's
//                weight stages through a 7-8 stage ring), warp 1 = MMA issuer (one elected lane; M128 N F K16, kind::f16, one
//                asm block per (k-step, tap) covering the group's tiles), warps 2-5 = epilogue (tcgen05.ld, bias, skip rows
//                prefetched from HBM into registers before the accumulator is waited for, ReLU, bf16, 16-byte stores that are
//                512 contiguous bytes per warp).  Accumulators are double-buffered at both widths (2 x 256 TMEM columns): the
//                epilogue of one pass runs under the MMAs of the next.
//   heads        k_wide_heads<G>: 1x1 head convolutions from the last activations, fixed-order sums, then heads_finish.
//
// The fp32 twin (precision 0) is plain SIMT over NHWC fp32 activations: it is the 1e-5 parity path, not a fast one.
#pragma once

namespace bb {

constexpr int WIDE_CHUNK = 4096;         // positions per pass of the layer-by-layer pipeline (activation buffers are sized for this)
constexpr int WIDE_ROWS = 128 + 2 * TC_MARGIN;   // rows of a tile's plane, margins included
constexpr int WIDE_PLANE = WIDE_ROWS * 16;       // bytes

struct WideGeo {                         // the raster of TcGeom<G>, as run-time numbers (the convolution kernel is not per game)
  int H, W, LW, RPP, PPT, HW;
};
template <class G>
inline WideGeo wide_geo() { return WideGeo{G::H, G::W, TcGeom<G>::LW, TcGeom<G>::RPP, TcGeom<G>::PPT, G::H * G::W}; }

template <int F>
struct WideCfg {
  static_assert(F == 32 || F == 64 || F == 128 || F == 256, "tower widths of the layer-by-layer kernel");
  static constexpr int TPG = F <= 64 ? 4 : 2;                        // tiles per group
  static constexpr int NS = F <= 128 ? 1 : 2;                        // passes over a group, each producing F / NS output channels (N of the MMA)
  static constexpr int NM = F / NS;                                  // output channels per pass: 32 / 64 / 128 / 128
  static constexpr int NACC = 2;                                     // accumulator sets: the epilogue of one pass overlaps the MMAs of the next
  static constexpr int ACC_COLS = TPG * NM;                          // TMEM columns of one accumulator set
  static constexpr int TMEM_COLS = NACC * ACC_COLS <= 256 ? 256 : 512;
  static_assert(NACC * ACC_COLS <= 512, "TMEM has 512 columns");
  static constexpr int KS_MAX = F / 16;                              // k-steps of a full-width layer
  static constexpr int TPS = 3;                                      // taps per weight-ring stage
  static constexpr int STAGE = TPS * NM * 32;                        // bytes per stage: 6 KB / 12 KB
  static constexpr int NST = F <= 64 ? 16 : 14;                      // weight-ring stages (96 KB / 168 KB: ~4.7 k-steps of a pass ahead at F = 256)
  static constexpr int TILE_BYTES = (F / 8) * WIDE_PLANE;            // one tile of F-channel activations (in HBM)
  static constexpr int NKA = 4;                                      // activation-ring slots: one k-step (two planes) of every tile of the group each
  static constexpr int A_SLOT = TPG * 2 * WIDE_PLANE;                // bytes per slot: 20 KB / 10 KB
  static constexpr int A_BYTES = NKA * A_SLOT;
  static constexpr int THREADS = 192;
  // shared memory: [A tiles][ring][bias F floats][barriers]
  static constexpr int OFF_RING = A_BYTES;
  static constexpr int OFF_BIAS = OFF_RING + NST * STAGE;
  static constexpr int OFF_BAR = OFF_BIAS + F * 4;
  static constexpr int N_BAR = 2 * NKA + 2 * NACC + 2 * NST;         // a_full / a_free per activation slot; d_full / d_free per accumulator set; w_full / w_free per stage
  static constexpr int OFF_TMEM = OFF_BAR + N_BAR * 8;
  static constexpr int SMEM = OFF_TMEM + 16;
  static_assert(SMEM <= 227 * 1024, "shared memory");
};

struct WideLayer {
  const uint8_t* in;       // input activations (tiles of 2 * ksteps planes)
  uint8_t* out;            // output activations (tiles of F / 8 planes)
  const uint8_t* skip;     // block input to add before the ReLU (may alias out), or null
  const uint8_t* w;        // this layer's weights [pass][k-step][tap][2][F / passes][8]
  const float* bias;       // [F]
  int ksteps;              // 16-channel steps of the input (stem: ceil(cin / 16); else F / 16)
};

#ifdef BB_WIDE_TIMING
__device__ unsigned long long g_wide_dbg[8];   // MMA warp: cycles waiting for tiles, for the accumulator, for weight stages, total, groups
#define WT(x) x
#else
#define WT(x)
#endif

// this warp's 32 TMEM lanes x 16 columns, no wait (the caller batches loads and waits once)
__device__ __forceinline__ void tmem_ld16_nowait(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// One (k-step, tap) of a tile group in ONE asm block: the same weight operand against the A operand of up to four tiles
// (descriptor low words differ by the tile stride, accumulators by F columns).  The issuing thread's instruction count
// per MMA is what paces the tensor pipe when descriptors are rebuilt in C (~110 cycles per MMA measured; N = 64 needs 32).
template <int NM, int TPG>
__device__ __forceinline__ void umma_group(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t a_step, uint32_t b_lo, uint32_t b_hi, uint32_t idesc,
                                           uint32_t accumulate, int nt) {
  if (TPG == 4) {
    asm volatile(
        "{\n"
        ".reg .pred pz, pe, p1, p2, p3;\n"
        ".reg .b32 al, dd;\n"
        ".reg .b64 ad, bd;\n"
        "elect.sync _|pe, 0xFFFFFFFF;\n"
        "setp.ne.u32 pz, %7, 0;\n"
        "setp.gt.and.s32 p1, %8, 1, pe;\n"
        "setp.gt.and.s32 p2, %8, 2, pe;\n"
        "setp.gt.and.s32 p3, %8, 3, pe;\n"
        "mov.b64 bd, {%4, %5};\n"
        "mov.b64 ad, {%1, %2};\n"
        "@pe tcgen05.mma.cta_group::1.kind::f16 [%0], ad, bd, %6, pz;\n"
        "add.u32 al, %1, %3;\n  mov.b64 ad, {al, %2};\n  add.u32 dd, %0, %9;\n"
        "@p1 tcgen05.mma.cta_group::1.kind::f16 [dd], ad, bd, %6, pz;\n"
        "add.u32 al, al, %3;\n  mov.b64 ad, {al, %2};\n  add.u32 dd, dd, %9;\n"
        "@p2 tcgen05.mma.cta_group::1.kind::f16 [dd], ad, bd, %6, pz;\n"
        "add.u32 al, al, %3;\n  mov.b64 ad, {al, %2};\n  add.u32 dd, dd, %9;\n"
        "@p3 tcgen05.mma.cta_group::1.kind::f16 [dd], ad, bd, %6, pz;\n"
        "}\n" ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(a_step), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate), "r"(nt), "n"(NM)
        : "memory");
  } else {
    asm volatile(
        "{\n"
        ".reg .pred pz, pe, p1;\n"
        ".reg .b32 al, dd;\n"
        ".reg .b64 ad, bd;\n"
        "elect.sync _|pe, 0xFFFFFFFF;\n"
        "setp.ne.u32 pz, %7, 0;\n"
        "setp.gt.and.s32 p1, %8, 1, pe;\n"
        "mov.b64 bd, {%4, %5};\n"
        "mov.b64 ad, {%1, %2};\n"
        "@pe tcgen05.mma.cta_group::1.kind::f16 [%0], ad, bd, %6, pz;\n"
        "add.u32 al, %1, %3;\n  mov.b64 ad, {al, %2};\n  add.u32 dd, %0, %9;\n"
        "@p1 tcgen05.mma.cta_group::1.kind::f16 [dd], ad, bd, %6, pz;\n"
        "}\n" ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(a_step), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate), "r"(nt), "n"(NM)
        : "memory");
  }
}

// One 3x3 convolution layer (+ folded BN, + skip, ReLU) over the live tiles of a chunk.  count_dev (may be null): device-
// side number of positions of the WHOLE call; this chunk holds positions [chunk_base, chunk_base + n_chunk).
// Loop order is k-step outer, tap inner: the two planes of a k-step are needed for its nine taps and then never again, so the
// group's tiles STREAM through shared memory k-step by k-step (a_full / a_free per k-step: the next group's planes land as
// soon as this group's MMAs on them have completed, and the first MMAs of a group start when its first 10 KB have arrived).
template <int F>
__global__ void __launch_bounds__(192, 1) k_wide_conv(WideLayer L, WideGeo geo, int n_chunk, int chunk_base, const int* __restrict__ count_dev) {
  using C = WideCfg<F>;
  extern __shared__ __align__(128) uint8_t wsm[];
  uint8_t* sA = wsm;
  uint8_t* sRing = wsm + C::OFF_RING;
  float* sBias = reinterpret_cast<float*>(wsm + C::OFF_BIAS);
  uint64_t* bars = reinterpret_cast<uint64_t*>(wsm + C::OFF_BAR);
  uint64_t *a_full = bars, *a_free = bars + C::NKA, *d_full = bars + 2 * C::NKA, *d_free = d_full + C::NACC;
  uint64_t *w_full = d_free + C::NACC, *w_free = w_full + C::NST;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wsm + C::OFF_TMEM);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  int cnt = n_chunk;
  if (count_dev) { const int live = *count_dev - chunk_base; cnt = live < cnt ? live : cnt; }
  if (cnt <= 0) return;                                              // uniform: nothing allocated yet
  const int n_tiles = (cnt + geo.PPT - 1) / geo.PPT, n_groups = (n_tiles + C::TPG - 1) / C::TPG;
  if ((int)blockIdx.x >= n_groups) return;

  const int ksteps = L.ksteps;
  constexpr int SPK = 9 / C::TPS;                                    // weight stages per k-step
  const uint32_t in_tile_bytes = (uint32_t)ksteps * 2u * WIDE_PLANE;

  if (warp == 1) {
    if (lane == 0) {
      for (int k = 0; k < C::NKA; ++k) { mbar_init(&a_full[k], 1); mbar_init(&a_free[k], 1); }
      for (int b = 0; b < C::NACC; ++b) { mbar_init(&d_full[b], 1); mbar_init(&d_free[b], 4); }
      for (int s = 0; s < C::NST; ++s) { mbar_init(&w_full[s], 1); mbar_init(&w_free[s], 1); }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    tmem_alloc(tmem_slot, C::TMEM_COLS);
  } else if (warp >= 2) {
    for (int i = threadIdx.x - 64; i < F; i += 128) sBias[i] = L.bias[i];
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ---- producer: per k-step the group's two planes per tile, then the k-step's weight stages -------------------------
    if (lane == 0) {
      unsigned wseq = 0, aseq = 0;
      for (int g = blockIdx.x; g < n_groups; g += gridDim.x)
      for (int h = 0; h < C::NS; ++h) {                              // a pass streams the group's activations again and its own half of the weights
        const int t0 = g * C::TPG, nt = min(C::TPG, n_tiles - t0);
        for (int ks = 0; ks < ksteps; ++ks, ++aseq) {
          const unsigned sl = aseq % C::NKA;
          mbar_wait(&a_free[sl], ((aseq / C::NKA) & 1u) ^ 1u);       // the MMAs that read this slot last have completed (passes at first use)
          asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&a_full[sl])), "r"((uint32_t)nt * 2u * WIDE_PLANE) : "memory");
          for (int i = 0; i < nt; ++i)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                             smem_u32(sA + sl * C::A_SLOT + i * 2 * WIDE_PLANE)),
                         "l"(L.in + (size_t)(t0 + i) * in_tile_bytes + (size_t)ks * 2 * WIDE_PLANE), "r"(2u * WIDE_PLANE), "r"(smem_u32(&a_full[sl]))
                         : "memory");
          for (int c = 0; c < SPK; ++c, ++wseq) {
            const unsigned s = wseq % C::NST;
            mbar_wait(&w_free[s], ((wseq / C::NST) & 1u) ^ 1u);
            bulk_g2s(sRing + s * C::STAGE, L.w + (size_t)((h * ksteps + ks) * SPK + c) * C::STAGE, C::STAGE, &w_full[s]);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ---- MMA issuer (warp-uniform control flow, one elected lane issues) -------------------------------------------
    constexpr uint32_t IDESC = umma_idesc_bf16(128, C::NM);
    unsigned wseq = 0, it = 0;
    WT(long long wa = 0; long long wd = 0; long long ww = 0; long long ngr = 0; const long long t_in = clock64();)
    const uint64_t adesc0 = umma_desc(smem_u32(sA) + TC_MARGIN * 16, WIDE_PLANE, 128);
    const uint32_t a_hi = (uint32_t)(adesc0 >> 32), a_lo0 = (uint32_t)adesc0, a_step = (2 * WIDE_PLANE) >> 4;
    unsigned aseq = 0;
    const uint64_t bdesc0 = umma_desc(smem_u32(sRing), C::NM * 16, 128);
    const uint32_t b_hi = (uint32_t)(bdesc0 >> 32), b_lo0 = (uint32_t)bdesc0;
    for (int g = blockIdx.x; g < n_groups; g += gridDim.x)
    for (int h = 0; h < C::NS; ++h, ++it) {
      const unsigned b = it % C::NACC, use = it / C::NACC;
      const int t0 = g * C::TPG, nt = min(C::TPG, n_tiles - t0);
      WT(const long long q1 = clock64();)
      mbar_wait(&d_free[b], (use & 1u) ^ 1u);                        // the epilogue has drained this accumulator set (passes at first use)
      WT(wd += clock64() - q1; ngr += 1;)
      const uint32_t dcol = tmem_base + b * C::ACC_COLS;
      for (int ks = 0; ks < ksteps; ++ks, ++aseq) {
        const unsigned sl = aseq % C::NKA;
        WT(const long long q0 = clock64();)
        mbar_wait(&a_full[sl], (aseq / C::NKA) & 1u);
        WT(wa += clock64() - q0;)
        tc_fence_after();
        const uint32_t a_ks = a_lo0 + (uint32_t)((sl * C::A_SLOT) >> 4);
        for (int c = 0; c < SPK; ++c, ++wseq) {
          const unsigned s = wseq % C::NST;
          WT(const long long q2 = clock64();)
          mbar_wait(&w_full[s], (wseq / C::NST) & 1u);
          WT(ww += clock64() - q2;)
          tc_fence_after();
#pragma unroll
          for (int tt = 0; tt < C::TPS; ++tt) {
            const int tap = c * C::TPS + tt;
            const int toff = (tap / 3 - 1) * geo.LW + (tap % 3 - 1);
            umma_group<C::NM, C::TPG>(dcol, a_ks + (uint32_t)toff, a_hi, a_step, b_lo0 + (uint32_t)((s * C::STAGE + tt * C::NM * 32) >> 4), b_hi, IDESC,
                                  (ks | tap) ? 1u : 0u, nt);
          }
          umma_commit_elect(&w_free[s]);
        }
        umma_commit_elect(&a_free[sl]);
      }
      umma_commit_elect(&d_full[b]);
    }
    WT(if (lane == 0 && L.ksteps > 2) { atomicAdd(&g_wide_dbg[0], (unsigned long long)wa); atomicAdd(&g_wide_dbg[1], (unsigned long long)wd);
         atomicAdd(&g_wide_dbg[2], (unsigned long long)ww); atomicAdd(&g_wide_dbg[3], (unsigned long long)(clock64() - t_in)); atomicAdd(&g_wide_dbg[4], (unsigned long long)ngr); })
  } else {
    // ---- epilogue: TMEM lane quarter q = warp % 4 ------------------------------------------------------------------
    const int q = warp & 3, rt = q * 32 + lane;
    const int pl = rt / geo.RPP, l = rt % geo.RPP, line = l / geo.LW, col = l % geo.LW;
    const bool cell = pl < geo.PPT && line >= 1 && col < geo.W;     // this row is a board cell (in every tile)
    constexpr int NPL = C::NM / 8;                                   // planes one pass writes per tile
    constexpr int SKR = NPL <= 8 ? NPL : 16;                         // skip planes held in registers ahead of their use
    unsigned it = 0;
    for (int g = blockIdx.x; g < n_groups; g += gridDim.x)
    for (int h = 0; h < C::NS; ++h, ++it) {
      const unsigned b = it % C::NACC, use = it / C::NACC;
      const int t0 = g * C::TPG, nt = min(C::TPG, n_tiles - t0);
      // The skip rows do not depend on this layer's MMAs: the first ones are asked for BEFORE the wait for the accumulator, and
      // every register that is consumed is refilled with the plane SKR further on (of this tile or the next), so the HBM / L2
      // latency of the skip reads hides behind the tensor work instead of being paid once per 32 channels.
      uint4 sk[SKR];
      const size_t row_off = (size_t)(rt + TC_MARGIN) * 16;
      const bool has_skip = L.skip != nullptr;
      auto skip_ptr = [&](int lin) {                                 // lin = tile-in-group * NPL + plane
        return L.skip + (size_t)(t0 + lin / NPL) * C::TILE_BYTES + (size_t)(h * NPL + lin % NPL) * WIDE_PLANE + row_off;
      };
      if (has_skip && cell) {
#pragma unroll
        for (int j = 0; j < SKR; ++j) sk[j] = __ldcg(reinterpret_cast<const uint4*>(skip_ptr(j)));
      }
      mbar_wait(&d_full[b], use & 1u);
      tc_fence_after();
      for (int i = 0; i < nt; ++i) {
        const bool ok = cell && (t0 + i) * geo.PPT + pl < cnt;
        uint8_t* orow = L.out + (size_t)(t0 + i) * C::TILE_BYTES + (size_t)(h * NPL) * WIDE_PLANE + row_off;
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + b * C::ACC_COLS + i * C::NM;
        const float* bias_h = sBias + h * C::NM;
#pragma unroll
        for (int cb = 0; cb < C::NM / 32; ++cb) {                        // 32 channels (four planes) per step: two loads in flight
          uint32_t r[32];
          tmem_ld16_nowait(taddr + cb * 32, r);
          tmem_ld16_nowait(taddr + cb * 32 + 16, r + 16);
          tmem_ld_wait();
#pragma unroll
          for (int hh = 0; hh < 4; ++hh) {                           // plane cb * 4 + hh of this pass: channels cb * 32 + hh * 8 ..
            const int plane = cb * 4 + hh;
            float v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = __uint_as_float(r[hh * 8 + k]) + bias_h[cb * 32 + hh * 8 + k];
            if (has_skip && cell) {
              const uint4 skv = sk[plane % SKR];
              const uint32_t sw[4] = {skv.x, skv.y, skv.z, skv.w};
#pragma unroll
              for (int k = 0; k < 4; ++k) { const float2 f = unpack_bf16(sw[k]); v[2 * k] += f.x; v[2 * k + 1] += f.y; }
              const int nxt = i * NPL + plane + SKR;                 // refill: SKR planes further on, possibly in the next tile of the group
              if (nxt < nt * NPL) sk[plane % SKR] = __ldcg(reinterpret_cast<const uint4*>(skip_ptr(nxt)));
            }
            if (ok)
              *reinterpret_cast<uint4*>(orow + (size_t)plane * WIDE_PLANE) =
                  make_uint4(pack_relu_bf16(v[0], v[1]), pack_relu_bf16(v[2], v[3]), pack_relu_bf16(v[4], v[5]), pack_relu_bf16(v[6], v[7]));
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&d_free[b]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, C::TMEM_COLS); }
}

// ---- input planes -> the stem's operand tiles (AsInputArray, decoded from the packed states) ----------------------------
template <class G>
__global__ void k_wide_stage(const typename G::State* __restrict__ states, uint8_t* __restrict__ act0, int n_chunk, int chunk_base,
                             const int* __restrict__ count_dev) {
  using Ge = TcGeom<G>;
  constexpr int HW = G::H * G::W, PLANES = Ge::KB_IN * 2;
  int cnt = n_chunk;
  if (count_dev) { const int live = *count_dev - chunk_base; cnt = live < cnt ? live : cnt; }
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= cnt * HW) return;
  const int p = i / HW, cell = i % HW;
  const typename G::State s = states[chunk_base + p];
  const int tile = p / Ge::PPT, rt = (p % Ge::PPT) * Ge::RPP + (cell / G::W + 1) * Ge::LW + cell % G::W;
  uint8_t* dst = act0 + (size_t)tile * PLANES * WIDE_PLANE + (size_t)(rt + TC_MARGIN) * 16;
#pragma unroll
  for (int pl = 0; pl < PLANES; ++pl) {
    uint32_t w4[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int c0 = pl * 8 + 2 * k;
      const float a = c0 < G::C ? (float)G::plane(s, cell, c0) : 0.f;
      const float b = c0 + 1 < G::C ? (float)G::plane(s, cell, c0 + 1) : 0.f;
      w4[k] = pack_bf16(a, b);
    }
    *reinterpret_cast<uint4*>(dst + (size_t)pl * WIDE_PLANE) = make_uint4(w4[0], w4[1], w4[2], w4[3]);
  }
}
template <class G>
__global__ void k_wide32_stage(const typename G::State* __restrict__ states, float* __restrict__ x, int n_chunk, int chunk_base,
                               const int* __restrict__ count_dev) {
  constexpr int HW = G::H * G::W;
  int cnt = n_chunk;
  if (count_dev) { const int live = *count_dev - chunk_base; cnt = live < cnt ? live : cnt; }
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= cnt * HW) return;
  const typename G::State s = states[chunk_base + i / HW];
  for (int c = 0; c < G::C; ++c) x[(size_t)i * G::C + c] = (float)G::plane(s, i % HW, c);
}

// ---- fp32 twin: NHWC activations, a thread per (cell, output channel) --------------------------------------------------
__global__ void k_wide32_conv(const float* __restrict__ in, const float* __restrict__ w, const float* __restrict__ bias, const float* skip,
                              float* out, int H, int W, int ci_n, int co_n, int n_chunk, int chunk_base, const int* __restrict__ count_dev) {
  int cnt = n_chunk;
  if (count_dev) { const int live = *count_dev - chunk_base; cnt = live < cnt ? live : cnt; }
  const size_t total = (size_t)(cnt > 0 ? cnt : 0) * H * W * co_n;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int co = (int)(i % co_n);
    const size_t e = i / co_n;
    const int cell = (int)(e % (H * W)), y = cell / W, x = cell % W;
    float acc = bias[co];
    for (int tap = 0; tap < 9; ++tap) {
      const int yy = y + tap / 3 - 1, xx = x + tap % 3 - 1;
      if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
      const float* ip = in + (e + (tap / 3 - 1) * W + (tap % 3 - 1)) * ci_n;
      const float* wp = w + (size_t)tap * ci_n * co_n + co;
      for (int ci = 0; ci < ci_n; ++ci) acc = fmaf(ip[ci], wp[(size_t)ci * co_n], acc);
    }
    if (skip) acc += skip[i];
    out[i] = fmaxf(acc, 0.f);
  }
}

// ---- heads: P positions per CTA ----------------------------------------------------------------------------------------
struct WideHeadW {
  const float* conv;       // value conv w[F], b | policy conv w[F][2], b[2]   (BN folded): 3 F + 3 floats
};
constexpr int WIDE_HP = 4;  // positions per CTA
template <class G, bool BF16>
__global__ void __launch_bounds__(256) k_wide_heads(NetParams np, WideHeadW hw, int F, const void* __restrict__ act, const typename G::State* __restrict__ states,
                                                    int n_chunk, int chunk_base, const int* __restrict__ count_dev, float* __restrict__ policy,
                                                    float* __restrict__ value, int legal_only) {
  using Ge = TcGeom<G>;
  constexpr int HW = G::H * G::W, P = WIDE_HP, ST = (int)((sizeof(typename G::State) + 15) & ~15u);
  extern __shared__ __align__(16) float hsm[];
  float* sConv = hsm;                         // 3 F + 3
  float* hcell = sConv + 3 * F + 4;           // [3][P * HW]
  float* hsum = hcell + 3 * P * HW;           // [P * 4]
  float* red = hsum + P * 4;                  // [32]
  float* zbuf = red + 32;                     // wide heads: A floats (legal-only: 2 x 512 + rows); narrow: unused
  uint8_t* sState = reinterpret_cast<uint8_t*>(zbuf + (G::A > 32 ? G::A : 4));
  int cnt = n_chunk;
  if (count_dev) { const int live = *count_dev - chunk_base; cnt = live < cnt ? live : cnt; }
  const int base = blockIdx.x * P;            // first position of this CTA inside the chunk
  if (base >= cnt) return;
  const int tid = threadIdx.x;
  for (int i = tid; i < 3 * F + 3; i += blockDim.x) sConv[i] = hw.conv[i];
  if (tid < P && base + tid < cnt) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(states + chunk_base + base + tid);
    uint32_t* dst = reinterpret_cast<uint32_t*>(sState + tid * ST);
    for (int i = 0; i < (int)(sizeof(typename G::State) / 4); ++i) dst[i] = src[i];
  }
  __syncthreads();
  for (int item = tid; item < P * HW; item += blockDim.x) {
    const int p = item / HW, cell = item % HW, gp = base + p;
    float v = sConv[F], p0 = sConv[F + 1 + 2 * F], p1 = sConv[F + 1 + 2 * F + 1];
    if (gp < cnt) {
      if (BF16) {
        const int tile = gp / Ge::PPT, rt = (gp % Ge::PPT) * Ge::RPP + (cell / G::W + 1) * Ge::LW + cell % G::W;
        const uint8_t* row = reinterpret_cast<const uint8_t*>(act) + (size_t)tile * (F / 8) * WIDE_PLANE + (size_t)(rt + TC_MARGIN) * 16;
        for (int pl = 0; pl < F / 8; ++pl) {
          const uint4 u = *reinterpret_cast<const uint4*>(row + (size_t)pl * WIDE_PLANE);
          const uint32_t uw[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const float2 f = unpack_bf16(uw[k]);
            const int c = pl * 8 + 2 * k;
            v = fmaf(f.x, sConv[c], v); v = fmaf(f.y, sConv[c + 1], v);
            p0 = fmaf(f.x, sConv[F + 1 + c * 2], p0); p0 = fmaf(f.y, sConv[F + 1 + (c + 1) * 2], p0);
            p1 = fmaf(f.x, sConv[F + 1 + c * 2 + 1], p1); p1 = fmaf(f.y, sConv[F + 1 + (c + 1) * 2 + 1], p1);
          }
        }
      } else {
        const float* row = reinterpret_cast<const float*>(act) + ((size_t)gp * HW + cell) * F;
        for (int c = 0; c < F; ++c) {
          const float a = row[c];
          v = fmaf(a, sConv[c], v);
          p0 = fmaf(a, sConv[F + 1 + c * 2], p0);
          p1 = fmaf(a, sConv[F + 1 + c * 2 + 1], p1);
        }
      }
    }
    hcell[item] = fmaxf(v, 0.f);
    hcell[P * HW + item] = fmaxf(p0, 0.f);
    hcell[2 * P * HW + item] = fmaxf(p1, 0.f);
  }
  __syncthreads();
  if (tid < P * 3) {                                         // sum over H,W in a fixed order
    const int p = tid / 3, k = tid % 3;
    float s = 0.f;
    for (int c = 0; c < HW; ++c) s += hcell[k * P * HW + p * HW + c];
    hsum[p * 4 + k] = s;
  }
  __syncthreads();
  unsigned long long call = 0;
  if (np.eps > 0.f) call = *np.call_counter;
  CacheOut co;
  co.states = sState;
  co.stride = ST;
  co.state_bytes = (int)sizeof(typename G::State);
  co.legal_only = legal_only;
  // hcell (consumed above) doubles as the narrow heads' scratch, as in the fp32 kernel
  heads_finish<G, BF16>(np, np.head, np.pol_w, np.pol_b, hsum, red, chunk_base + base, chunk_base + cnt, P, call, policy, value,
                        Team{tid, (int)blockDim.x, 0}, nullptr, G::A > 32 ? zbuf : hcell, -1, co);
}
template <class G>
inline size_t wide_heads_smem(int F) {
  const int HW = G::H * G::W, ST = (int)((sizeof(typename G::State) + 15) & ~15u);
  return (size_t)(3 * F + 4 + 3 * WIDE_HP * HW + WIDE_HP * 4 + 32 + (G::A > 32 ? G::A : 4)) * 4 + (size_t)WIDE_HP * ST + 16;
}

}  // namespace bb
