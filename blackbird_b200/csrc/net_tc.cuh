// Policy-value network forward on the 5th-generation tensor cores (tcgen05 + TMEM), bf16 operands,
// fp32 accumulation.  Included by net.cu (shares NetParams / heads_finish with the fp32 kernel).
//
// The tower (NetworkFactory.py:40-103: conv3x3+BN+ReLU stem, `blocks` residual blocks) is nine
// implicit GEMMs per batch of NP positions, all operands resident in shared memory:
//
//   rows    = board cells in a padded raster: a position is (H+1) lines of (W+1) cells, line 0 and
//             the last column are zero padding shared with the neighbouring line / position, so
//             the 3x3 neighbour (dy,dx) of a row is simply row + dy*(W+1) + dx and always lands on
//             a real cell or on a zero row (no bounds tests, no im2col copy).  Positions never straddle
//             a 128-row tile (PPT = 128 / rows-per-position whole positions per tile, the tile's tail rows
//             stay zero), so the tiles of a batch are INDEPENDENT: tile t of layer l+1 needs only tile t
//             of layer l, and the layers of different tiles overlap freely (round 1 packed positions
//             densely, which made every tile wait for its two neighbours: 3.5 k cycles per layer);
//   A tile  = 128 consecutive rows x 16 input channels, bf16, K-major, SWIZZLE_NONE: the buffer is
//             [k-half][row][8 channels] so a row is 16 bytes and every tap of the convolution is
//             the SAME buffer addressed through a descriptor whose start address is shifted by
//             (dy*(W+1)+dx)*16 bytes;
//   B tile  = one tap's 16x16 weight block (BN folded on the host), bf16, K-major;
//   D tile  = 128 rows x 16 output channels, fp32, in TMEM (16 columns per tile).
//
// Warp roles (32 * (4 * TILES + 1) threads): the LAST warp issues every tcgen05.mma (one elected
// lane) and owns the TMEM allocation; each tile has its own four epilogue warps (one per TMEM lane
// quarter), so all tiles of a layer drain in parallel.  There is no CTA-wide barrier inside the
// tower: tile t of layer l+1 is issued as soon as the epilogue of tile t of layer l has arrived on its
// `ready` mbarrier, and an epilogue starts when the tile's tcgen05.commit arrives on its `full`
// mbarrier -- while the epilogue warps drain one tile the MMA warp already feeds the next.  Epilogue: tcgen05.ld (32 lanes x 16 columns per warp), + bias,
// (+ skip), ReLU, pack to bf16, two 16-byte st.shared per row into the next layer's A buffer.  The
// weights arrive with one cp.async.bulk (TMA) while the first batch's planes are decoded.  The
// last layer's epilogue feeds the heads from the fp32 registers (1x1 convs -> per-cell ReLU ->
// fixed-order sums).
#pragma once
#include <cuda_bf16.h>

namespace bb {

constexpr int TC_HEAD_FLOATS = 384;      // head parameters staged in shared memory: 51 + 3*dense + 1 (+ 3*A when A <= 32)
constexpr int TC_MARGIN = 16;            // zero rows in front of / behind the raster (|tap offset| <= W+2)

template <class G> struct TcBatch;
// A kernel that allocates TMEM runs ONE CTA per SM (cudaOccupancyMaxActiveBlocksPerMultiprocessor reports 1 for
// any kernel with tcgen05.alloc, two different TMEM kernels do not share an SM either, and the measured two-wave
// step agrees).  A single batch is a chain of dependent steps per layer (MMA -> commit -> epilogue -> arrive)
// that leaves the tensor pipe idle most of the time, so the CTA runs TWO independent pipelines: each has its
// own MMA-issuing warp, eight epilogue warps, activation buffers, TMEM columns, mbarriers and named barrier,
// and works on its own batch; they share the resident weights and interleave on the tensor pipe.
// Largest batch of one pipeline (positions): up to seven 128-row tiles.  DragonChess (17 input planes: a separate
// two-block input buffer) has room for two pipelines only.
// Two pipelines is the default: a third one (BB_TC_PIPES=3, four-tile batches) was measured and changes nothing in
// streaming mode -- with three batches in flight the tensor pipe is the shared bottleneck (nine 128x16x16 MMAs per
// tile and layer at ~47 cycles each, paced by the 4 KB A-operand read from shared memory, not by the math).
#ifndef BB_TC_PIPES
#define BB_TC_PIPES 2
#endif
// TILES = 128-row tiles of one pipeline's batch; NP = PPT * TILES positions
#ifndef BB_TC_TILES
#define BB_TC_TILES 4                            // streaming batches hold 3 - 5 positions; four tiles leave room for the heads warp's double buffers
#endif
template <> struct TcBatch<TicTacToe> { static constexpr int TILES = BB_TC_TILES, PIPES = BB_TC_PIPES; };   // 8 positions of 16 rows per tile: 32
template <> struct TcBatch<Connect4> { static constexpr int TILES = BB_TC_TILES, PIPES = BB_TC_PIPES; };    // 2 positions of 56 rows per tile: 8
template <> struct TcBatch<DragonChess> { static constexpr int TILES = 3, PIPES = 2; };                                  // 1 position of 81 rows per tile: 3 (+ a 4032-float policy scratch row per pipeline)

template <class G>
struct TcGeom {
  static constexpr int LW = G::W + 1;                       // line width
  static constexpr int RPP = (G::H + 1) * LW;               // rows per position
  static constexpr int PPT = 128 / RPP;                     // whole positions per 128-row tile
  static constexpr int TILES = TcBatch<G>::TILES;
  static constexpr int NP = PPT * TILES;                    // positions per batch
  static_assert(PPT >= 1, "a position must fit one tile");
  // row of (position p of the batch, line, column): positions never straddle a tile
  __host__ __device__ static constexpr int row_of(int p, int line, int col) { return (p / PPT) * 128 + (p % PPT) * RPP + line * LW + col; }
  static constexpr int ROWS = TILES * 128 + 2 * TC_MARGIN;  // rows allocated per k-half plane
  static constexpr int PLANE = ROWS * 16;                   // bytes per k-half plane
  static constexpr int BUF = 2 * PLANE;                     // one 16-channel activation buffer
  static constexpr int KB_IN = (G::C + 15) / 16;            // 16-channel blocks of the input planes
  static_assert(LW + 1 <= TC_MARGIN, "tap offsets must stay inside the margin");
  static_assert(TILES <= 8, "one mbarrier pair and 16 TMEM columns per tile, parity bits in one word");
  static constexpr int PIPES = TcBatch<G>::PIPES;           // independent pipelines per CTA
  static_assert(PIPES >= 1 && PIPES <= 3, "named barriers 1..3");
  // per pipeline: groups of four warps (one per TMEM lane quarter); group g drains tiles g, g + GROUPS, ...
  // One group is the default: streaming batches are one to three tiles, and the four warps a second group would
  // cost are worth more as search warps (register file: DESIGN.md "co-residency").  Splitting a tile's 16 output
  // channels over two groups instead was tried and lost: the epilogue's cost is per-row overhead, not per channel.
#ifndef BB_TC_EPI_GROUPS
#define BB_TC_EPI_GROUPS 1
#endif
  static constexpr int EPI_GROUPS = BB_TC_EPI_GROUPS;
  static constexpr int EPI_THREADS = EPI_GROUPS * 128;      // epilogue threads of one pipeline
  static constexpr int TEAM_THREADS = EPI_THREADS + 32;     // + its MMA-issuing warp: the threads that meet at the pipeline's named barrier
  // Narrow heads (A <= 32) get a warp of their own per pipeline: it turns the last layer's head maps into value / policy,
  // writes the answers and publishes them WHILE the epilogue and MMA warps already run the next batch's tower (double-
  // buffered hand-off through two mbarrier pairs).  Heads + publish were 20 - 26 % of a batch's time with the MMA warp idle.
#ifndef BB_TC_HEADS_WARP
#define BB_TC_HEADS_WARP 0            // measured: 65.5 M vs 70.4 M sims/s -- the lone warp takes 13 k cycles per batch, and answer LATENCY, not pipeline occupancy, is what the slots wait for
#endif
  static constexpr bool HEADS_WARP = BB_TC_HEADS_WARP && G::A <= 32;
  static constexpr int PIPE_THREADS = TEAM_THREADS + (HEADS_WARP ? 32 : 0);
  static constexpr int THREADS = PIPES * PIPE_THREADS;
  static constexpr int TMEM_PIPE_COLS = TILES * 16;         // one 16-column accumulator per tile
  static constexpr int TMEM_COLS = PIPES * TMEM_PIPE_COLS <= 32 ? 32 : PIPES * TMEM_PIPE_COLS <= 64 ? 64 : PIPES * TMEM_PIPE_COLS <= 128 ? 128
                                   : PIPES * TMEM_PIPE_COLS <= 256 ? 256 : 512;   // allocations are powers of two
  static_assert(PIPES * TMEM_PIPE_COLS <= 512, "TMEM has 512 columns");
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
// warp-uniform variant: the whole warp executes it, one elected lane commits (keeps the operands in
// uniform registers: UTCHMMA / UTCBAR read URs, and a divergent `if (lane == 0)` costs a R2UR waterfall per MMA)
__device__ __forceinline__ void umma_commit_elect(uint64_t* bar) {
  asm volatile(
      "{\n"
      ".reg .pred pe;\n"
      "elect.sync _|pe, 0xFFFFFFFF;\n"
      "@pe tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n"
      "}\n" ::"r"(smem_u32(bar)) : "memory");
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
__device__ __forceinline__ uint32_t pack_relu_bf16(float a, float b) {
  __nv_bfloat162 h = __hmax2(__floats2bfloat162_rn(a, b), __floats2bfloat162_rn(0.f, 0.f));
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float2 unpack_bf16(uint32_t u) {
  __nv_bfloat162 h = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(h);
}

// bulk global -> shared copy (TMA, non-tensor form) completing on an mbarrier
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// barrier among the epilogue threads of one pipeline (ids 4..6; the MMA warp never joins; literal ids, see Team)
template <int NTHREADS>
__device__ __forceinline__ void epi_bar_sync(int pipe) {
  if (pipe == 0) asm volatile("bar.sync 4, %0;" ::"n"(NTHREADS) : "memory");
  else if (pipe == 1) asm volatile("bar.sync 5, %0;" ::"n"(NTHREADS) : "memory");
  else asm volatile("bar.sync 6, %0;" ::"n"(NTHREADS) : "memory");
}

// The nine taps of one 128-row tile in ONE asm block: per tap a 32-bit add on the low descriptor word
// (start-address field, 16-byte units) for A (shifted rows) and for B (next 512-byte weight block).
// A single thread issues every MMA, so its instruction count per MMA is what paces the tensor pipe at
// this tiny N (measured: ~125 cycles per MMA with descriptors rebuilt in C, 8-cycle floor).
template <int LW, bool FIRST>
__device__ __forceinline__ void umma_taps9(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi, uint32_t idesc) {
  asm volatile(
      "{\n"
      ".reg .pred pz, po, pe;\n"
      ".reg .b32 al, bl;\n"
      "elect.sync _|pe, 0xFFFFFFFF;\n"
      ".reg .b64 ad, bd;\n"
      "setp.ne.u32 pz, %15, 0;\n"
      "setp.eq.u32 po, 0, 0;\n"
      "add.s32 al, %1, %6;\n  mov.b64 ad, {al, %2};\n  mov.b64 bd, {%3, %4};\n"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%0], ad, bd, %5, pz;\n"
      "add.s32 al, %1, %7;\n  mov.b64 ad, {al, %2};\n  add.s32 bl, %3, 32;\n  mov.b64 bd, {bl, %4};\n"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%0], ad, bd, %5, po;\n"
      "add.s32 al, %1, %8;\n  mov.b64 ad, {al, %2};\n  add.s32 bl, %3, 64;\n  mov.b64 bd, {bl, %4};\n"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%0], ad, bd, %5, po;\n"
      "add.s32 al, %1, %9;\n  mov.b64 ad, {al, %2};\n  add.s32 bl, %3, 96;\n  mov.b64 bd, {bl, %4};\n"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%0], ad, bd, %5, po;\n"
      "add.s32 al, %1, %10;\n mov.b64 ad, {al, %2};\n  add.s32 bl, %3, 128;\n mov.b64 bd, {bl, %4};\n"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%0], ad, bd, %5, po;\n"
      "add.s32 al, %1, %11;\n mov.b64 ad, {al, %2};\n  add.s32 bl, %3, 160;\n mov.b64 bd, {bl, %4};\n"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%0], ad, bd, %5, po;\n"
      "add.s32 al, %1, %12;\n mov.b64 ad, {al, %2};\n  add.s32 bl, %3, 192;\n mov.b64 bd, {bl, %4};\n"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%0], ad, bd, %5, po;\n"
      "add.s32 al, %1, %13;\n mov.b64 ad, {al, %2};\n  add.s32 bl, %3, 224;\n mov.b64 bd, {bl, %4};\n"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%0], ad, bd, %5, po;\n"
      "add.s32 al, %1, %14;\n mov.b64 ad, {al, %2};\n  add.s32 bl, %3, 256;\n mov.b64 bd, {bl, %4};\n"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%0], ad, bd, %5, po;\n"
      "}\n" ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc),
      "n"(-LW - 1), "n"(-LW), "n"(-LW + 1), "n"(-1), "n"(0), "n"(1), "n"(LW - 1), "n"(LW), "n"(LW + 1), "n"(FIRST ? 0 : 1)
      : "memory");
}

template <class G>
struct TcSmem {                                  // byte offsets into dynamic shared memory
  using Ge = TcGeom<G>;
  int w, bias, head, wbar, tmem, pipe0, pipe_stride, total;          // shared by both pipelines, then one block per pipeline
  int x, y, in, hsum, red, idx, st, zb, ent, claim, mbar;            // offsets inside a pipeline's block
  int hmap, info, hbar, idx_stride, st_stride2, ent_stride;          // heads warp: double-buffered head maps / batch info / mbarriers; strides of the doubled buffers
  __host__ __device__ TcSmem(int w_blocks, int layers) {
    int o = 0;
    w = o; o += w_blocks * 512;
    bias = o; o += layers * 16 * 4;
    head = o; o += TC_HEAD_FLOATS * 4;
    o = (o + 7) & ~7;
    wbar = o; o += 8;
    tmem = o; o += 8;
    o = (o + 127) & ~127;
    pipe0 = o;
    int q = 0;
    x = q; q += Ge::BUF;
    y = q; q += Ge::BUF;
    if (Ge::KB_IN == 1) in = y; else { in = q; q += Ge::KB_IN * Ge::BUF; }
    hsum = q; q += Ge::NP * 4 * 4;
    red = q; q += 32 * 4;
    constexpr int NB = Ge::HEADS_WARP ? 2 : 1;     // buffers the heads warp reads are doubled: it works on batch i while batch i + 1 is staged
    idx_stride = Ge::NP * 4;
    idx = q; q += NB * idx_stride;                 // streaming mode: slot id of every position of the batch
    q = (q + 15) & ~15;
    st_stride2 = Ge::NP * (int)((sizeof(typename G::State) + 15) & ~15u);
    st = q; q += NB * st_stride2;                  // the batch's packed states, staged once
    zb = q; q += (G::A > 32 ? G::A : Ge::NP * G::A) * 4;   // wide policy heads: one row of logits / probabilities; narrow ones: the batch's noise draws (heads_finish)
    ent_stride = 2 * Ge::NP * 4;
    ent = q; q += NB * ent_stride;                 // streaming mode: evaluation-cache entry (+1) this pipeline locked per position, and its tag
    claim = q; q += 2 * 4;                         // streaming mode: first ticket and size of the claimed batch
    q = (q + 7) & ~7;
    mbar = q; q += 2 * Ge::TILES * 8;              // full[TILES], ready[TILES]
    hbar = q; q += 4 * 8;                          // hfull[2] (epilogue -> heads warp), hfree[2] (heads warp -> epilogue)
    info = q; q += 2 * 4 * 4;                      // per buffer: positions of the batch (0 = leave), first ticket / base
    q = (q + 15) & ~15;
    hmap = q; if (Ge::HEADS_WARP) q += 2 * Ge::TILES * 128 * 16;   // per raster row (value, policy0, policy1) of the last layer's 1x1 head convolutions
    pipe_stride = (q + 127) & ~127;
    total = pipe0 + Ge::PIPES * pipe_stride;
  }
};

// The heads warp's work on one batch (narrow heads): GP lanes per position (8 up to 8 actions, 16 up to 16, else 32), 32 / GP
// positions per pass.  Lane j of a group sums cells j, j + GP, ... of the three head maps, the group combines them with a
// fixed-order xor tree, then the value head (dense units j, j + GP, ...) and the policy head (action j: logit, soft-max, its own
// noise draw, renormalisation) as in heads_finish.  In serve mode every pass ends with the release fence and the groups'
// publications (answer sequence number, evaluation-cache tag).
template <class G>
__device__ void heads_warp_batch(const NetParams& np, const float* sHead, const uint8_t* hmap, int cnt, bool serve, unsigned first, unsigned long long call,
                                 float* __restrict__ policy, float* __restrict__ value, const int* sIdx, const uint32_t* sEnt, const uint8_t* sState,
                                 int st_stride, const NetServe& sv) {
  using Ge = TcGeom<G>;
  constexpr int A = G::A, HW = G::H * G::W, GP = A <= 8 ? 8 : A <= 16 ? 16 : 32, PPW = 32 / GP, NPm = Ge::NP;
  const int lane = threadIdx.x & 31, g = lane / GP, j = lane % GP;
  const int D = np.dense, nh = 51 + 3 * D + 1;
  const float* d1w = sHead + 17 + 34;
  const float* d1b = d1w + D;
  const float* d2w = d1b + D;
  const float d2b = d2w[D];
  const float* pol_w = sHead + nh;
  const float* pol_b = pol_w + 2 * A;
  const bool cache = A <= ECACHE_MAX_ACTIONS && serve && sv.ecache != nullptr;
  for (int p0 = 0; p0 < cnt; p0 += PPW) {
    const int p = p0 + g;
    const bool live = p < cnt;
    const int pc = live ? p : p0;
    float sv_ = 0.f, s0 = 0.f, s1 = 0.f;                    // sums over H,W of relu(bn(1x1 conv)): value map, two policy maps
    for (int c = j; c < HW; c += GP) {
      const float4 m = *reinterpret_cast<const float4*>(hmap + Ge::row_of(pc, c / G::W + 1, c % G::W) * 16);
      sv_ += m.x; s0 += m.y; s1 += m.z;
    }
#pragma unroll
    for (int o = GP / 2; o; o >>= 1) {
      sv_ += __shfl_xor_sync(0xFFFFFFFFu, sv_, o, GP);
      s0 += __shfl_xor_sync(0xFFFFFFFFu, s0, o, GP);
      s1 += __shfl_xor_sync(0xFFFFFFFFu, s1, o, GP);
    }
    const size_t orow = serve ? (size_t)sIdx[pc] : (size_t)(first + (unsigned)pc);
    uint32_t* cent = nullptr;
    if (cache && live && sEnt[p]) cent = sv.ecache + (size_t)(sEnt[p] - 1u) * ECACHE_WORDS;
    float part = 0.f;                                       // value: dense(1->D)+sum, ReLU, dense(D->1), tanh
    for (int d = j; d < D; d += GP) part = fmaf(fmaxf(fmaf(sv_, d1w[d], (float)HW * d1b[d]), 0.f), d2w[d], part);
#pragma unroll
    for (int o = GP / 2; o; o >>= 1) part += __shfl_xor_sync(0xFFFFFFFFu, part, o, GP);
    if (live && j == 0) {
      const float v = tanhf(part + d2b);
      value[orow] = v;
      if (cent) {
        cent[13] = __float_as_uint(v);
        const uint32_t* st = reinterpret_cast<const uint32_t*>(sState + p * st_stride);
#pragma unroll
        for (int i = 0; i < 4; ++i) cent[i] = i * 4 < (int)sizeof(typename G::State) ? st[i] : 0u;
      }
    }
    const bool act = j < A;                                 // policy: dense(2->A)+sum, softmax, noise mix, renormalise
    const int a = act ? j : 0;
    const float z = act ? fmaf(s0, pol_w[a], fmaf(s1, pol_w[A + a], (float)HW * pol_b[a])) : -INFINITY;
    float m = z;
#pragma unroll
    for (int o = GP / 2; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xFFFFFFFFu, m, o, GP));
    const float e = act ? expf(z - m) : 0.f;
    float sum = e;
#pragma unroll
    for (int o = GP / 2; o; o >>= 1) sum += __shfl_xor_sync(0xFFFFFFFFu, sum, o, GP);
    float pa = e / sum;
    if (np.eps > 0.f && act && live)
      pa = (1.f - np.eps) * pa + np.eps * beta_noise(np.alpha, np.seed ^ (call * 0x632BE59BD9B4E019ull), ((unsigned long long)first + (unsigned)p) * A + a);
    float tot = pa;
#pragma unroll
    for (int o = GP / 2; o; o >>= 1) tot += __shfl_xor_sync(0xFFFFFFFFu, tot, o, GP);
    if (live && act) {
      const float pf = pa / tot;
      policy[orow * A + a] = pf;
      if (cent) cent[4 + a] = __float_as_uint(pf);
    }
    if (serve) {                                            // publish this pass: answers first, then sequence numbers and cache tags
      fence_acq_rel();
      __syncwarp();
      if (live && j == 0) {
        fence_acq_rel();
        *reinterpret_cast<volatile unsigned*>(&sv.resp_seq[orow]) = first + (unsigned)p + 1u;
        if (cent) *reinterpret_cast<volatile uint32_t*>(cent + 14) = sEnt[NPm + p];
      }
    }
  }
}

template <class G>
__global__ void __launch_bounds__(576, 2)   // <= 56 registers whatever the thread count: the CTA shares its SM's register file with a search CTA
k_net_forward_tc(NetParams np, TcParams tp, const typename G::State* __restrict__ states, int n_max,
                 const int* __restrict__ count_dev, float* __restrict__ policy, float* __restrict__ value, NetServe sv) {
  using Ge = TcGeom<G>;
  using State = typename G::State;
  constexpr int TILES = Ge::TILES;
  extern __shared__ __align__(128) uint8_t tc_smem[];
  const bool serve = sv.q != nullptr;                       // streaming mode: batches are claimed from the request ring
  const int n = serve ? 0x7FFFFFFF : (count_dev ? min(n_max, *count_dev) : n_max);
  // launch mode: spread the requests evenly over all pipelines of the grid (a tick of the engine is latency-bound,
  // one wave of equally loaded pipelines is what matters), at most NP positions per batch
  const int n_pipes = (int)gridDim.x * Ge::PIPES;
  const int ppc = min(Ge::NP, max(1, (n + n_pipes - 1) / n_pipes));
  const int n_batches = (n + ppc - 1) / ppc;
  if (!serve && (int)blockIdx.x * Ge::PIPES >= n_batches) return;     // before any allocation
  const TcSmem<G> L(tp.w_blocks, np.layers);
  const int tid_cta = threadIdx.x, lane = tid_cta & 31, warp = tid_cta >> 5;
  const int pipe = tid_cta / Ge::PIPE_THREADS;              // which pipeline this thread belongs to
  const int tid = tid_cta - pipe * Ge::PIPE_THREADS;        // thread index inside the pipeline
  const int pwarp = tid >> 5;                               // warp index inside the pipeline: 0..7 epilogue, 8 MMA issuer
  const bool is_mma = pwarp == 4 * Ge::EPI_GROUPS;
  const bool is_heads = Ge::HEADS_WARP && pwarp == 4 * Ge::EPI_GROUPS + 1;   // never joins the team's barrier
  const Team team{tid, Ge::TEAM_THREADS, 1 + pipe};         // named barrier 1 + pipeline: the pipelines never wait for each other
  uint8_t* sW = tc_smem + L.w;
  float* sBias = reinterpret_cast<float*>(tc_smem + L.bias);
  float* sHead = reinterpret_cast<float*>(tc_smem + L.head);
  uint64_t* wbar = reinterpret_cast<uint64_t*>(tc_smem + L.wbar);     // weights landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tc_smem + L.tmem);
  uint8_t* pb = tc_smem + L.pipe0 + pipe * L.pipe_stride;             // this pipeline's block
  uint8_t* const pb_ = pb;
  uint8_t* bufX = pb + L.x;
  uint8_t* bufY = pb + L.y;
  uint8_t* bufIn = pb + L.in;
  float* hsum = reinterpret_cast<float*>(pb + L.hsum);
  float* red = reinterpret_cast<float*>(pb + L.red);
  int* sIdx = reinterpret_cast<int*>(pb + L.idx);                     // (buffer 0; the heads-warp build switches per batch)
  constexpr int ST_STRIDE = (int)((sizeof(State) + 15) & ~15u);
  uint8_t* sState = pb + L.st;
  volatile unsigned* sClaim = reinterpret_cast<volatile unsigned*>(pb + L.claim);
  uint32_t* sEnt = reinterpret_cast<uint32_t*>(pb + L.ent);          // [NP] entry + 1 (0: not ours), then [NP] tags
  uint64_t* hfull = reinterpret_cast<uint64_t*>(pb + L.hbar);         // [2] epilogue warps -> heads warp: head maps of buffer b written
  uint64_t* hfree = hfull + 2;                                        // [2] heads warp -> epilogue warps: buffer b may be rewritten
  volatile unsigned* sInfo = reinterpret_cast<volatile unsigned*>(pb + L.info);   // [2][4]: positions of the batch (0 = leave), first ticket / base
  uint8_t* hmap = pb + L.hmap;
  uint64_t* full = reinterpret_cast<uint64_t*>(pb + L.mbar);          // MMA -> epilogue, per tile
  uint64_t* ready = full + TILES;                                     // epilogue -> MMA, per tile (4 warp arrivals)
  const int layers = np.layers;

  // ---- one-time setup (whole CTA) --------------------------------------------------------------------
  if (is_mma) {
    if (lane == 0) {
      for (int t = 0; t < TILES; ++t) { mbar_init(&full[t], 1); mbar_init(&ready[t], 4); }
      if (Ge::HEADS_WARP) for (int b = 0; b < 2; ++b) { mbar_init(&hfull[b], 4 * Ge::EPI_GROUPS); mbar_init(&hfree[b], 1); }
      if (pipe == 0) mbar_init(wbar, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      if (pipe == 0) bulk_g2s(sW, tp.w, (uint32_t)tp.w_blocks * 512u, wbar);
    }
    __syncwarp();
    if (pipe == 0) tmem_alloc(tmem_slot, Ge::TMEM_COLS);
  } else if (!is_heads) {
    if (pipe == 0) {
      for (int i = tid; i < layers * 16; i += Ge::EPI_THREADS) sBias[i] = np.tower_b[i];
      const int nh = 51 + 3 * np.dense + 1;                 // value conv, policy conv, d1 w/b, d2 w/b
      for (int i = tid; i < nh; i += Ge::EPI_THREADS) sHead[i] = np.head[i];
      if (G::A <= 32) {
        for (int i = tid; i < 2 * G::A; i += Ge::EPI_THREADS) sHead[nh + i] = np.pol_w[i];
        for (int i = tid; i < G::A; i += Ge::EPI_THREADS) sHead[nh + 2 * G::A + i] = np.pol_b[i];
      }
    }
    const int zero_bytes = L.hsum - L.x;                    // X, Y (and the separate input buffer), padding included
    for (int i = tid; i < zero_bytes / 16; i += Ge::EPI_THREADS) reinterpret_cast<uint4*>(bufX)[i] = make_uint4(0, 0, 0, 0);
  }
  tc_fence_before();
  fence_async_smem();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot + (uint32_t)(pipe * Ge::TMEM_PIPE_COLS);
  constexpr uint32_t IDESC = umma_idesc_bf16(128, 16);
  unsigned long long call = 0;
  if (np.eps > 0.f) call = *np.call_counter;
  // mbarrier phase parities, one bit per tile: batches use different numbers of tiles, so every tile's barriers
  // advance at their own pace (full: epilogue side, ready: MMA side)
  uint32_t full_par = 0, ready_par = 0;
  bool weights_in = false;
  unsigned pend_lo = 0, pend_n = 0;                         // serve mode, thread 0 of the pipeline: tickets claimed but not yet served

  int batch = (int)blockIdx.x * Ge::PIPES + pipe;
  if (is_heads) {
    // ---- heads warp: batch k of this pipeline arrives in buffer k & 1 ------------------------------------------
    if (lane == 0) { mbar_arrive(&hfree[0]); mbar_arrive(&hfree[1]); }      // both buffers start free
    for (unsigned k = 0;; ++k) {
      const unsigned b = k & 1u;
      mbar_wait(&hfull[b], (k >> 1) & 1u);
      const int hcnt = (int)sInfo[b * 4];
      if (hcnt == 0) break;                                   // the pipeline has run out of work
      const long long th0 = clock64();
      heads_warp_batch<G>(np, sHead, hmap + b * (TILES * 128 * 16), hcnt, serve, sInfo[b * 4 + 1], call, policy, value,
                          reinterpret_cast<const int*>(pb + L.idx + b * L.idx_stride), reinterpret_cast<const uint32_t*>(pb + L.ent + b * L.ent_stride),
                          pb + L.st + b * L.st_stride2, ST_STRIDE, sv);
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&hfree[b]);
        if (serve) atomicAdd(&sv.q->cyc_heads, (unsigned long long)(clock64() - th0));
      }
    }
  } else {
  unsigned bno = 0;                                           // batches this pipeline has started (heads-warp build: buffer = bno & 1)
  for (;; ++bno) {
    int base, cnt;
    if (Ge::HEADS_WARP) {                                     // this batch's copies of the buffers the heads warp reads
      const unsigned b = bno & 1u;
      sIdx = reinterpret_cast<int*>(pb + L.idx + b * L.idx_stride);
      sState = pb + L.st + b * L.st_stride2;
      sEnt = reinterpret_cast<uint32_t*>(pb + L.ent + b * L.ent_stride);
    }
    unsigned ticket0 = 0;
    long long tm0 = 0, tm1 = 0, tm2 = 0;
    if (serve && tid == 0) tm0 = clock64();
    if (serve) {
      if (tid == 0) {
        // Claim tickets with ONE atomicAdd (a compare-and-swap loop serialises the claimers: ~55 us of waiting per
        // batch was measured).  The size is open / divisor (at least 1, at most NP) from a possibly stale view, so a
        // claim can run past the tail.  Claimed tickets belong to this pipeline for good (pend_lo, pend_n); a batch
        // is whatever part of them has been issued by now -- never wait for the rest: the slots that would issue
        // them may all be waiting for answers of this very batch.
        unsigned c = 0, hh = 0, idle = 0;
        volatile unsigned* vt = reinterpret_cast<volatile unsigned*>(&sv.q->tail);
        volatile unsigned* vh = reinterpret_cast<volatile unsigned*>(&sv.q->head);
        volatile unsigned* vdone = reinterpret_cast<volatile unsigned*>(&sv.q->search_done);
        const unsigned srv = sv.q->servers, ctas = sv.q->search_ctas;
        for (;;) {
          if (pend_n == 0) {
            const int open = (int)(*vt - *vh);
            if (open > 0) {
              unsigned want = min((unsigned)Ge::NP, max(1u, ((unsigned)open + srv - 1u) / srv));
              // the tower's cost is per 128-row tile, not per position: fill the last tile while tickets are open
              const unsigned fill = (want + Ge::PPT - 1u) / Ge::PPT * Ge::PPT;
              want = max(want, min(min(fill, (unsigned)Ge::NP), (unsigned)open));
              pend_lo = atomicAdd(&sv.q->head, want);
              pend_n = want;
            }
          }
          if (pend_n) {
            const int have = (int)(*vt - pend_lo);
            if (have > 0) {
              c = min(pend_n, (unsigned)have);
              hh = pend_lo;
              pend_lo += c;
              pend_n -= c;
              break;
            }
          }
          if (*vdone >= ctas) {                                 // the search has left: nothing more will be issued
            fence_acq_rel();
            if (pend_n) { if ((int)(*vt - pend_lo) > 0) continue; pend_n = 0; }
            if ((int)(*vt - *vh) <= 0) { c = 0; break; }        // nobody left to ask
            continue;
          }
          __nanosleep(pend_n ? 100 : 200);
          if (++idle > (16u << 20)) {                           // seconds without a request or a finished search: give up, never hang
            if (blockIdx.x == 0) printf("[bbengine] streaming watchdog: network kernel idle (head %u tail %u, search CTAs done %u of %u)\n", *vh, *vt,
                                        *vdone, ctas);
            *reinterpret_cast<volatile unsigned*>(&sv.q->stop) = 2u;
            c = 0;
            break;
          }
        }
        sClaim[0] = hh;
        sClaim[1] = c;
      }
      team.sync();
      ticket0 = sClaim[0];
      cnt = (int)sClaim[1];
      base = 0;
      if (tid == 0) tm1 = clock64();
    } else {
      base = batch < n_batches ? batch * ppc : 0;
      cnt = batch < n_batches ? min(ppc, n - base) : 0;     // positions of this batch
      batch += n_pipes;
    }
    if (Ge::HEADS_WARP && !is_mma) {
      // buffer b is free once the heads warp has finished the batch before last; then hand it this batch's size and
      // first ticket (or, with no work left, the word to leave)
      const unsigned b = bno & 1u;
      mbar_wait(&hfree[b], (bno >> 1) & 1u);
      if (tid == 0) { sInfo[b * 4] = (unsigned)cnt; sInfo[b * 4 + 1] = serve ? ticket0 : (unsigned)base; }
      if (cnt == 0) {
        epi_bar_sync<Ge::EPI_THREADS>(pipe);
        if (lane == 0) mbar_arrive(&hfull[b]);
      }
    }
    if (cnt == 0) break;
    const int tiles = (cnt + Ge::PPT - 1) / Ge::PPT;        // 128-row tiles in use
    if (is_mma) {
      // ---- MMA issuer: layers x tiles as a wavefront ---------------------------------------------
      {                                                       // warp-uniform: every lane waits, one elected lane issues
        if (!weights_in) { mbar_wait(wbar, 0); weights_in = true; }
        const uint32_t w_addr = smem_u32(sW);
#ifdef BB_NET_TIMING_TOWER
        long long nt_wait = 0, nt_issue = 0;
#endif
        for (int layer = 0; layer < layers; ++layer) {
          const bool odd = layer & 1;
          const uint32_t in_addr = smem_u32(layer == 0 ? bufIn : (odd ? bufX : bufY));
          const int kblocks = layer == 0 ? Ge::KB_IN : 1;
          const int wblk0 = layer == 0 ? 0 : Ge::KB_IN * 9 + (layer - 1) * 9;
          for (int t = 0; t < tiles; ++t) {
            // the inputs of tile t are tile t of the previous layer's output, nothing else (positions do not straddle tiles)
#ifdef BB_NET_TIMING_TOWER
            const long long nt0 = clock64();
#endif
            mbar_wait(&ready[t], (ready_par >> t) & 1u);
            ready_par ^= 1u << t;
            tc_fence_after();
#ifdef BB_NET_TIMING_TOWER
            const long long nt1 = clock64();
            nt_wait += nt1 - nt0;
#endif
            {
              const uint64_t adesc = umma_desc(in_addr + (t * 128 + TC_MARGIN) * 16, Ge::PLANE, 128);
              const uint64_t bdesc = umma_desc(w_addr + wblk0 * 512, 256, 128);
              const uint32_t a_lo = (uint32_t)adesc, a_hi = (uint32_t)(adesc >> 32), b_lo = (uint32_t)bdesc, b_hi = (uint32_t)(bdesc >> 32);
              umma_taps9<Ge::LW, true>(tmem_base + t * 16, a_lo, a_hi, b_lo, b_hi, IDESC);
              for (int kb = 1; kb < kblocks; ++kb)          // further 16-channel blocks of the input planes (layer 0 only)
                umma_taps9<Ge::LW, false>(tmem_base + t * 16, a_lo + kb * (Ge::BUF >> 4), a_hi, b_lo + kb * 9 * 32, b_hi, IDESC);
            }
            umma_commit_elect(&full[t]);
#ifdef BB_NET_TIMING_TOWER
            nt_issue += clock64() - nt1;
#endif
          }
        }
#ifdef BB_NET_TIMING_TOWER
        if (serve && lane == 0) {
          atomicAdd(&sv.q->ndbg[0], (unsigned long long)nt_wait);
          atomicAdd(&sv.q->ndbg[1], (unsigned long long)nt_issue);
          atomicAdd(&sv.q->ndbg[5], (unsigned long long)(layers * tiles));
        }
#endif
      }
      __syncwarp();
    } else {
      // ---- epilogue warps: group wg drains tiles wg, wg + 2, ...; TMEM lane quarter q = warp % 4 -----------
      const int wg = pwarp >> 2, q = warp & 3;               // q follows the CTA-wide warp id: a warp reaches TMEM lanes 32 * (warp % 4) ...
      // stage the batch's packed states once (one thread per position: in streaming mode it waits for the ticket's
      // entry to be published and reads it through L2), so the global-memory latency is paid once per batch and
      // not once per tile
      if (tid < cnt) {
        const State* src;
        if (serve) {
          const unsigned tk = ticket0 + (unsigned)tid;
          volatile unsigned long long* ent = reinterpret_cast<volatile unsigned long long*>(&sv.q_ent[tk & sv.q_mask]);
          unsigned long long v = *ent;
          unsigned spin = 0;
          for (; !((unsigned)v == tk && ((unsigned)(v >> 32) & RING_PUBLISHED)) && spin < (64u << 20); ++spin) v = *ent;
          int slot = (int)((unsigned)(v >> 32) & ~RING_PUBLISHED);
          if (spin >= (64u << 20)) {                         // the ticket was issued (tail moved past it) but its entry never appeared: report, never serve garbage
            printf("[bbengine] streaming watchdog: ring entry of ticket %u never published (entry holds ticket %u, word %08x)\n", tk, (unsigned)v, (unsigned)(v >> 32));
            *reinterpret_cast<volatile unsigned*>(&sv.q->stop) = 2u;
            slot = 0;                                        // any valid slot: the call returns BB_ESTATE and the results are discarded
          } else {
            *ent = v & 0x7FFFFFFFFFFFFFFFull;                // entry read: ticket tk + ring may publish here
          }
          fence_acq_rel();                                   // acquire: the slot's state was written before the entry
          src = reinterpret_cast<const State*>(sv.q_state) + slot;
          sIdx[tid] = slot;                                  // where the answer goes
        } else {
          src = states + base + tid;
        }
        uint32_t* dst = reinterpret_cast<uint32_t*>(sState + tid * ST_STRIDE);
#pragma unroll
        for (int i = 0; i < (int)(sizeof(State) / 4); ++i) dst[i] = __ldcg(reinterpret_cast<const uint32_t*>(src) + i);
        if (G::A <= ECACHE_MAX_ACTIONS && serve) {           // evaluation cache: take the entry this answer will go to
          uint32_t own = 0u, tag = 0u;
          if (sv.ecache) {
            uint32_t k[4] = {0u, 0u, 0u, 0u};
#pragma unroll
            for (int i = 0; i < (int)(sizeof(State) / 4) && i < 4; ++i) k[i] = dst[i];
            uint32_t idx;
            ecache_hash(k, sv.ecache_mask, &idx, &tag);
            if (atomicExch(sv.ecache + (size_t)idx * ECACHE_WORDS + 14, ECACHE_LOCKED) != ECACHE_LOCKED) own = idx + 1u;
            fence_acq_rel();                                 // the entry reads as invalid before any of its payload changes
          }
          sEnt[tid] = own;
          sEnt[Ge::NP + tid] = tag;
        }
      }
      epi_bar_sync<Ge::EPI_THREADS>(pipe);
#ifdef BB_NET_TIMING_TOWER
      long long ne_wait = 0, ne_work = 0, ne_ld = 0, ne_math = 0;
      const long long ne_stage = clock64();
#endif
      // input planes (AsInputArray) decoded from the staged states into the raster rows; okmask: this thread's row
      // is a board cell of the batch, per tile
      uint32_t okmask = 0;
      for (int t = wg; t < tiles; t += Ge::EPI_GROUPS) {
        const int rt = q * 32 + lane, row = t * 128 + rt;     // row inside the tile / inside the batch raster
        const int pos = t * Ge::PPT + rt / Ge::RPP, l = rt % Ge::RPP, line = l / Ge::LW, col = l % Ge::LW;
        if (rt < Ge::PPT * Ge::RPP && pos < cnt && line >= 1 && col < G::W) {
          okmask |= 1u << t;
          const int cell = (line - 1) * G::W + col;
          const State s = *reinterpret_cast<const State*>(sState + pos * ST_STRIDE);
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
        fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&ready[t]);
      }

      for (int layer = 0; layer < layers; ++layer) {
        const bool odd = layer & 1;
        uint8_t* out = odd ? bufY : bufX;
        const bool skip = !odd && layer > 0;                // conv_2 + BN + block input (NetworkFactory.py:95-98)
        const bool last = layer == layers - 1;
        const float* bias = sBias + layer * 16;
        for (int t = wg; t < tiles; t += Ge::EPI_GROUPS) {
          const int row = t * 128 + q * 32 + lane;
          const bool cell_ok = (okmask >> t) & 1u;
#ifdef BB_NET_TIMING_TOWER
          const long long ne0 = clock64();
#endif
          mbar_wait(&full[t], (full_par >> t) & 1u);
          full_par ^= 1u << t;
          tc_fence_after();
#ifdef BB_NET_TIMING_TOWER
          const long long ne1 = clock64();
          ne_wait += ne1 - ne0;
#endif
          float acc[16];
          tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + t * 16, acc);
#ifdef BB_NET_TIMING_TOWER
          const long long ne2 = clock64();
          ne_ld += ne2 - ne1;
#endif
          if (cell_ok) {
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
            if (!last) {                                    // ReLU on the packed pairs: rounding is monotonic, so the order does not matter
              *reinterpret_cast<uint4*>(slot) = make_uint4(pack_relu_bf16(acc[0], acc[1]), pack_relu_bf16(acc[2], acc[3]),
                                                           pack_relu_bf16(acc[4], acc[5]), pack_relu_bf16(acc[6], acc[7]));
              *reinterpret_cast<uint4*>(slot + Ge::PLANE) = make_uint4(pack_relu_bf16(acc[8], acc[9]), pack_relu_bf16(acc[10], acc[11]),
                                                                       pack_relu_bf16(acc[12], acc[13]), pack_relu_bf16(acc[14], acc[15]));
            } else {
#pragma unroll
              for (int c = 0; c < 16; ++c) acc[c] = fmaxf(acc[c], 0.f);
              // heads from the fp32 registers: 1x1 convs (1 value + 2 policy filters) + BN + ReLU; the three
              // per-cell values replace this row's (now dead) slot in X
              float v = sHead[16], p0 = sHead[17 + 32], p1 = sHead[17 + 33];
#pragma unroll
              for (int c = 0; c < 16; ++c) {
                v = fmaf(acc[c], sHead[c], v);
                p0 = fmaf(acc[c], sHead[17 + c * 2], p0);
                p1 = fmaf(acc[c], sHead[17 + c * 2 + 1], p1);
              }
              float4* hdst = Ge::HEADS_WARP ? reinterpret_cast<float4*>(hmap + (bno & 1u) * (TILES * 128 * 16) + row * 16) : reinterpret_cast<float4*>(slot);
              *hdst = make_float4(fmaxf(v, 0.f), fmaxf(p0, 0.f), fmaxf(p1, 0.f), 0.f);
            }
          }
#ifdef BB_NET_TIMING_TOWER
          const long long ne3 = clock64();
          ne_math += ne3 - ne2;
#endif
          if (!last) {
            tc_fence_before();
            fence_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(&ready[t]);
          }
#ifdef BB_NET_TIMING_TOWER
          ne_work += clock64() - ne1;
#endif
        }
      }
#ifdef BB_NET_TIMING_TOWER
      if (serve && tid == 0) {
        atomicAdd(&sv.q->ndbg[2], (unsigned long long)ne_wait);
        atomicAdd(&sv.q->ndbg[3], (unsigned long long)ne_work);
        atomicAdd(&sv.q->ndbg[4], (unsigned long long)ne_ld);
        atomicAdd(&sv.q->ndbg[6], (unsigned long long)ne_math);
      }
      (void)ne_stage;
#endif
    }
    if constexpr (Ge::HEADS_WARP) {
      // ---- narrow heads: hand the batch's head maps to the heads warp and go on with the next batch ------------
      if (!is_mma) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&hfull[bno & 1u]);       // this warp's rows of the last layer are in hmap
      }
      if (serve && tid == 0) {
        tm2 = clock64();
        atomicAdd(&sv.q->batches, 1ull);
        atomicAdd(&sv.q->cyc_claim, (unsigned long long)(tm1 - tm0));
        atomicAdd(&sv.q->cyc_tower, (unsigned long long)(tm2 - tm1));
      }
    } else {
    // ---- wide heads: the whole pipeline (MMA warp included) meets here once per batch ------------------
    team.sync();
    if (serve && tid == 0) tm2 = clock64();
    // per-position sums over H,W of the three head maps: one warp per (position, map), fixed-order tree
    for (int pk = pwarp; pk < cnt * 3; pk += Ge::PIPE_THREADS / 32) {
      const int p = pk / 3, k = pk % 3;
      float s = 0.f;
      for (int c = lane; c < G::H * G::W; c += 32)
        s += reinterpret_cast<const float*>(bufX + (Ge::row_of(p, c / G::W + 1, c % G::W) + TC_MARGIN) * 16)[k];
#pragma unroll
      for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
      if (lane == 0) hsum[p * 4 + k] = s;
    }
    team.sync();
#ifdef BB_NET_TIMING_HEADS
    long long th1 = 0, th2 = 0, th3 = 0, th4 = 0;
    if (serve && tid == 0) th1 = clock64();
#endif
    {
      // (one warp per position instead of one thread -- lane a = action a, xor-tree softmax, each warp publishing its own
      // slot -- was measured in round 2 and LOST 4 %: the SM is issue-bound, five warps' worth of full-width instructions
      // cost the neighbouring epilogue and search warps more than the shorter dependent chain gains)
      const int nh = 51 + 3 * np.dense + 1;
      const float* pw = G::A <= 32 ? sHead + nh : np.pol_w;
      const float* pb = G::A <= 32 ? sHead + nh + 2 * G::A : np.pol_b;
      CacheOut co;
      co.states = sState; co.stride = ST_STRIDE; co.state_bytes = (int)sizeof(State);
      co.legal_only = sv.legal_only;
      if (G::A <= ECACHE_MAX_ACTIONS && serve && sv.ecache) { co.base = sv.ecache; co.ent = sEnt; }
#ifdef BB_NET_TIMING_HEADS
      if (serve) co.dbg = sv.q->hdbg;
#endif
      heads_finish<G, true>(np, sHead, pw, pb, hsum, red, serve ? (int)(ticket0 & 0x3FFFFFFFu) : base, n, cnt, call, policy, value, team, serve ? sIdx : nullptr,
                      reinterpret_cast<float*>(pb_ + L.zb), serve ? (long long)ticket0 : -1ll, co);
    }
#ifdef BB_NET_TIMING_HEADS
    if (serve && tid == 0) th2 = clock64();
#endif
    if (serve) {                                            // publish: answers first, then the per-slot sequence numbers (and the cache tags)
      fence_acq_rel();
#ifdef BB_NET_TIMING_HEADS
      if (tid == 0) th3 = clock64();
#endif
      team.sync();
#ifdef BB_NET_TIMING_HEADS
      if (tid == 0) th4 = clock64();
#endif
      if (tid < cnt) {
        fence_acq_rel();                                    // release by the publishing thread itself (the answers were written by its team mates before the barrier)
        *reinterpret_cast<volatile unsigned*>(&sv.resp_seq[sIdx[tid]]) = ticket0 + (unsigned)tid + 1u;
        if (G::A <= ECACHE_MAX_ACTIONS && sv.ecache && sEnt[tid])
          *reinterpret_cast<volatile uint32_t*>(sv.ecache + (size_t)(sEnt[tid] - 1u) * ECACHE_WORDS + 14) = sEnt[Ge::NP + tid];
      }
    }
    team.sync();                                          // X / hsum are rewritten by the next batch
#ifdef BB_NET_TIMING_HEADS
    if (serve && tid == 0) {
      const long long th5 = clock64();
      atomicAdd(&sv.q->ndbg[0], (unsigned long long)(th1 - tm2));   // hsum pass + sync
      atomicAdd(&sv.q->ndbg[1], (unsigned long long)(th2 - th1));   // heads_finish
      atomicAdd(&sv.q->ndbg[2], (unsigned long long)(th3 - th2));   // fence
      atomicAdd(&sv.q->ndbg[3], (unsigned long long)(th4 - th3));   // sync after fence
      atomicAdd(&sv.q->ndbg[7], (unsigned long long)(th5 - th4));   // publish + last sync
    }
#endif
    if (serve && tid == 0) {
      const long long tm3 = clock64();
      atomicAdd(&sv.q->batches, 1ull);
      atomicAdd(&sv.q->cyc_claim, (unsigned long long)(tm1 - tm0));
      atomicAdd(&sv.q->cyc_tower, (unsigned long long)(tm2 - tm1));
      atomicAdd(&sv.q->cyc_heads, (unsigned long long)(tm3 - tm2));
    }
    }
  }
  }
  tc_fence_before();
  __syncthreads();
  if (is_mma && pipe == 0) tmem_dealloc(*tmem_slot, Ge::TMEM_COLS);
}

}  // namespace bb
