// Trainer step on the GPU: Network.train (Network.py:66-84) = one sess.run(trainingOp) of the graph
// NetworkFactory.py builds -- forward with saved activations (:22-183), the loss AS WRITTEN (:185-204) and its
// gradients, and the Adam / momentum / SGD update of tf.compat.v1.train (:229-242).  fp32 throughout (the reference
// trains in fp32), NHWC activations, every reduction in a fixed order (run-to-run deterministic).
//
// What the loss is (restated, not fixed -- oracle/train_np.py has the derivation):
//   lossEvaluation = mean_n (value[n] - evalLabel[n])^2
//   policy fed to the loss = u / T,  u[n,a] = (1 - eps) softmax[n,a] + eps noise[a],  T = sum over the WHOLE [B,A] tensor
//   lossPolicy     = -(1/B^2) sum_a (sum_i log policy[i,a]) (sum_j label[j,a])        (mean over the B x B tensordot)
//   lossParam      = mean over {conv kernels, BN gamma, BN beta, dense kernels} of sum(v^2) / 2
// BN is the inference-mode affine map y = gamma (x - mean) / sqrt(var + 1e-3) + beta (training=False everywhere in
// the reference graph; mean / var never change), so each conv+BN pair is run as ONE folded convolution W' = W s,
// b' = (b - mean) s + beta with s = gamma / sqrt(var + 1e-3), and the gradients of W', b' are mapped back:
//   dW = dW' s,  db = db' s,  dbeta = db',  dgamma = (sum_{tap,ci} dW' W + db' (b - mean)) / sqrt(var + 1e-3).
//
// Kernels are plain SIMT: the template network is 16 filters wide and the reference's batch is 50 positions
// (parameters_template.yaml:24), so a step is launch-bound, not math-bound; the whole step (about 50 launches) is
// captured once per batch size into a CUDA graph and replayed.
#include <math.h>
#include <string.h>

#include <map>
#include <vector>

#include <tuple>

#include "common.cuh"
#include "noise.cuh"

namespace bb {

constexpr float TR_BN_EPS = 1e-3f;
constexpr int TR_MAX_DENSE = 64;

struct ConvBN {                    // offsets (floats) into the flat parameter vector / folded buffer
  size_t w, b, gamma, beta, mean, var;
  size_t wf, bf;                   // folded weights [k*k][ci][co], folded bias [co]
  int k, ci, co;
};

struct TrainDims {
  int H, W, HW, Cin, F, D, A, L;   // L = tower layers (1 + 2 blocks)
  int n_l2;                        // variables in lossParam
};

// ---- generic helpers ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float block_sum(float v, float* red) {     // fixed-order tree; red: 32 floats; result on every thread
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
  __syncthreads();
  if (lane == 0) red[wid] = v;
  __syncthreads();
  float r = 0.f;
  for (int i = 0; i < nw; ++i) r += red[i];
  return r;
}
__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xFFFFFFFFu, v, o));
  __syncthreads();
  if (lane == 0) red[wid] = v;
  __syncthreads();
  float r = red[0];
  for (int i = 1; i < nw; ++i) r = fmaxf(r, red[i]);
  return r;
}

// ---- BN folding and its transpose -----------------------------------------------------------------------------------
__global__ void k_tr_fold(const float* __restrict__ th, float* __restrict__ wf, ConvBN c) {
  const int n = c.k * c.k * c.ci * c.co;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n + c.co; i += gridDim.x * blockDim.x) {
    if (i < n) {
      const int co = i % c.co;
      const float s = th[c.gamma + co] / sqrtf(th[c.var + co] + TR_BN_EPS);
      wf[c.wf + i] = th[c.w + i] * s;
    } else {
      const int co = i - n;
      const float s = th[c.gamma + co] / sqrtf(th[c.var + co] + TR_BN_EPS);
      wf[c.bf + co] = (th[c.b + co] - th[c.mean + co]) * s + th[c.beta + co];
    }
  }
}
// gwf: gradients of the folded weights / biases (same offsets as wf) -> gradients of w, b, gamma, beta in g
__global__ void k_tr_unfold(const float* __restrict__ th, const float* __restrict__ gwf, float* __restrict__ g, ConvBN c) {
  __shared__ float red[32];
  const int co = blockIdx.x;                                    // one block per output channel
  const int rows = c.k * c.k * c.ci;
  const float inv = 1.0f / sqrtf(th[c.var + co] + TR_BN_EPS);
  const float s = th[c.gamma + co] * inv;
  float acc = 0.f;
  for (int r = threadIdx.x; r < rows; r += blockDim.x) {
    const size_t i = (size_t)r * c.co + co;
    const float d = gwf[c.wf + i];
    g[c.w + i] = d * s;
    acc += d * th[c.w + i];
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) {
    const float db = gwf[c.bf + co];
    g[c.b + co] = db * s;
    g[c.beta + co] = db;
    g[c.gamma + co] = (acc + db * (th[c.b + co] - th[c.mean + co])) * inv;
    g[c.mean + co] = 0.f;                                       // frozen statistics
    g[c.var + co] = 0.f;
  }
}

// ---- tower ------------------------------------------------------------------------------------------------------------
__global__ void k_tr_input(const int8_t* __restrict__ planes, float* __restrict__ x, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) x[i] = (float)planes[i];
}
// out[n,cell,co] = relu(b[co] + sum_tap sum_ci in[n,cell+off,ci] W[tap,ci,co] (+ skip[n,cell,co])); one thread per output element
__global__ void k_tr_conv_fwd(const float* __restrict__ in, const float* __restrict__ wgt, const float* __restrict__ bias,
                              const float* __restrict__ skip, float* __restrict__ out, int N, int H, int W, int ci_n, int co_n) {
  const size_t total = (size_t)N * H * W * co_n;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int co = (int)(i % co_n);
    const size_t pc = i / co_n;
    const int cell = (int)(pc % (H * W)), y = cell / W, x = cell % W;
    const size_t n = pc / (H * W);
    float acc = bias[co];
    for (int dy = -1; dy <= 1; ++dy) {
      const int yy = y + dy;
      if (yy < 0 || yy >= H) continue;
      for (int dx = -1; dx <= 1; ++dx) {
        const int xx = x + dx;
        if (xx < 0 || xx >= W) continue;
        const float* ip = in + ((n * H + yy) * W + xx) * ci_n;
        const float* wp = wgt + (size_t)((dy + 1) * 3 + (dx + 1)) * ci_n * co_n + co;
        for (int ci = 0; ci < ci_n; ++ci) acc = fmaf(ip[ci], wp[(size_t)ci * co_n], acc);
      }
    }
    if (skip) acc += skip[i];
    out[i] = fmaxf(acc, 0.f);
  }
}
// dpre = dout * (out > 0), in place; the block input of a residual block gets dpre added (skip connection)
__global__ void k_tr_relu_bwd(float* __restrict__ dout, const float* __restrict__ out, float* __restrict__ dskip, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float d = out[i] > 0.f ? dout[i] : 0.f;
    dout[i] = d;
    if (dskip) dskip[i] += d;
  }
}
// din[n,cell,ci] (+)= sum_tap sum_co dpre[n,cell-off,co] W[tap,ci,co]
__global__ void k_tr_conv_bwd_data(const float* __restrict__ dpre, const float* __restrict__ wgt, float* __restrict__ din, int accumulate,
                                   int N, int H, int W, int ci_n, int co_n) {
  const size_t total = (size_t)N * H * W * ci_n;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int ci = (int)(i % ci_n);
    const size_t pc = i / ci_n;
    const int cell = (int)(pc % (H * W)), y = cell / W, x = cell % W;
    const size_t n = pc / (H * W);
    float acc = 0.f;
    for (int dy = -1; dy <= 1; ++dy) {
      const int yy = y - dy;                                    // the output cell that read this input through tap (dy,dx)
      if (yy < 0 || yy >= H) continue;
      for (int dx = -1; dx <= 1; ++dx) {
        const int xx = x - dx;
        if (xx < 0 || xx >= W) continue;
        const float* dp = dpre + ((n * H + yy) * W + xx) * co_n;
        const float* wp = wgt + ((size_t)((dy + 1) * 3 + (dx + 1)) * ci_n + ci) * co_n;
        for (int co = 0; co < co_n; ++co) acc = fmaf(dp[co], wp[co], acc);
      }
    }
    din[i] = accumulate ? din[i] + acc : acc;
  }
}
// dW[tap,ci,co] = sum_{n,cell} in[n,cell+off,ci] dpre[n,cell,co]; db[co] = sum dpre.  One block per (tap,ci,co) / per co.
__global__ void k_tr_conv_bwd_w(const float* __restrict__ in, const float* __restrict__ dpre, float* __restrict__ dw, float* __restrict__ db,
                                int N, int H, int W, int ci_n, int co_n) {
  __shared__ float red[32];
  const int nW = 9 * ci_n * co_n;
  const int b = blockIdx.x;
  float acc = 0.f;
  if (b < nW) {
    const int co = b % co_n, ci = (b / co_n) % ci_n, tap = b / (co_n * ci_n);
    const int dy = tap / 3 - 1, dx = tap % 3 - 1;
    const int tot = N * H * W;
    for (int e = threadIdx.x; e < tot; e += blockDim.x) {
      const int cell = e % (H * W), y = cell / W, x = cell % W;
      const int yy = y + dy, xx = x + dx;
      if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
      const size_t n = e / (H * W);
      acc = fmaf(in[((n * H + yy) * W + xx) * ci_n + ci], dpre[(size_t)e * co_n + co], acc);
    }
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) dw[b] = acc;
  } else {
    const int co = b - nW;
    const int tot = N * H * W;
    for (int e = threadIdx.x; e < tot; e += blockDim.x) acc += dpre[(size_t)e * co_n + co];
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) db[co] = acc;
  }
}

// Row-per-thread convolutions for the template width (co = 16): a thread owns one raster cell and all of its output
// channels in registers; the layer's folded weights sit in shared memory and are read as warp-wide broadcasts, the
// activations are read once per tap as whole rows.  The element-per-thread kernels above issue two loads per FMA.
template <int CI, int CO>
__global__ void __launch_bounds__(128) k_tr_conv_fwd_row(const float* __restrict__ in, const float* __restrict__ wgt, const float* __restrict__ bias,
                                                          const float* __restrict__ skip, float* __restrict__ out, int N, int H, int W) {
  __shared__ float sw[9 * CI * CO];
  for (int i = threadIdx.x; i < 9 * CI * CO; i += blockDim.x) sw[i] = wgt[i];
  __syncthreads();
  const int total = N * H * W;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    const int cell = e % (H * W), y = cell / W, x = cell % W;
    float acc[CO];
#pragma unroll
    for (int co = 0; co < CO; ++co) acc[co] = bias[co];
#pragma unroll
    for (int tap = 0; tap < 9; ++tap) {
      const int yy = y + tap / 3 - 1, xx = x + tap % 3 - 1;
      if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
      const float* ip = in + ((size_t)e + (tap / 3 - 1) * W + (tap % 3 - 1)) * CI;
      float xv[CI];
#pragma unroll
      for (int ci = 0; ci < CI; ++ci) xv[ci] = ip[ci];
#pragma unroll
      for (int ci = 0; ci < CI; ++ci)
#pragma unroll
        for (int co = 0; co < CO; ++co) acc[co] = fmaf(xv[ci], sw[(tap * CI + ci) * CO + co], acc[co]);
    }
    float* op = out + (size_t)e * CO;
#pragma unroll
    for (int co = 0; co < CO; ++co) {
      float v = acc[co];
      if (skip) v += skip[(size_t)e * CO + co];
      op[co] = fmaxf(v, 0.f);
    }
  }
}
template <int CI, int CO>
__global__ void __launch_bounds__(128) k_tr_conv_bwd_data_row(const float* __restrict__ dpre, const float* __restrict__ wgt, float* __restrict__ din,
                                                               int accumulate, int N, int H, int W) {
  __shared__ float sw[9 * CI * CO];
  for (int i = threadIdx.x; i < 9 * CI * CO; i += blockDim.x) sw[i] = wgt[i];
  __syncthreads();
  const int total = N * H * W;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    const int cell = e % (H * W), y = cell / W, x = cell % W;
    float acc[CI];
#pragma unroll
    for (int ci = 0; ci < CI; ++ci) acc[ci] = 0.f;
#pragma unroll
    for (int tap = 0; tap < 9; ++tap) {
      const int yy = y - (tap / 3 - 1), xx = x - (tap % 3 - 1);       // the output cell that read this input through the tap
      if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
      const float* dp = dpre + ((size_t)e - (tap / 3 - 1) * W - (tap % 3 - 1)) * CO;
      float dv[CO];
#pragma unroll
      for (int co = 0; co < CO; ++co) dv[co] = dp[co];
#pragma unroll
      for (int ci = 0; ci < CI; ++ci)
#pragma unroll
        for (int co = 0; co < CO; ++co) acc[ci] = fmaf(dv[co], sw[(tap * CI + ci) * CO + co], acc[ci]);
    }
    float* ip = din + (size_t)e * CI;
#pragma unroll
    for (int ci = 0; ci < CI; ++ci) ip[ci] = accumulate ? ip[ci] + acc[ci] : acc[ci];
  }
}

// Tiled weight gradient for narrow layers (ci * co <= 1024): a CTA takes one tap and a chunk of TR_WCHUNK consecutive rows
// (cells of the batch raster), stages the chunk's shifted input rows and its dpre rows in shared memory (coalesced), and
// thread p owns the (ci, co) pairs p, p + 256, ...: one FMA per row and pair.  Partial sums go to part[chunk][tap][ci][co]
// (the bias gradient rides along as "tap 9"); k_tr_wgrad_reduce adds the chunks in a fixed order, so the result does not
// depend on scheduling.  The one-CTA-per-weight kernel above read the whole batch once per weight: 19 ms of a 20 ms step
// at batch 4096.
constexpr int TR_WCHUNK = 256;
template <int PPT>                                               // (ci, co) pairs per thread
__global__ void __launch_bounds__(256) k_tr_conv_bwd_w_tiled(const float* __restrict__ in, const float* __restrict__ dpre, float* __restrict__ part,
                                                              int N, int H, int W, int ci_n, int co_n) {
  extern __shared__ float tw_smem[];
  float* s_in = tw_smem;                                         // [TR_WCHUNK][ci_n]  (zero where the tap falls off the board)
  float* s_dp = tw_smem + TR_WCHUNK * ci_n;                      // [TR_WCHUNK][co_n]
  const int chunk = blockIdx.x, tap = blockIdx.y;                // tap 9: bias gradient (sum of dpre)
  const int total = N * H * W, row0 = chunk * TR_WCHUNK, rows = min(TR_WCHUNK, total - row0);
  const int dy = tap / 3 - 1, dx = tap % 3 - 1;
  for (int i = threadIdx.x; i < rows * co_n; i += 256) s_dp[i] = dpre[(size_t)row0 * co_n + i];
  if (tap < 9) {
    for (int i = threadIdx.x; i < rows * ci_n; i += 256) {
      const int r = i / ci_n, ci = i % ci_n, e = row0 + r;
      const int cell = e % (H * W), y = cell / W + dy, x = cell % W + dx;
      s_in[i] = (y >= 0 && y < H && x >= 0 && x < W) ? in[((size_t)e + dy * W + dx) * ci_n + ci] : 0.f;
    }
  }
  __syncthreads();
  const int pairs = ci_n * co_n;
  float* out = part + ((size_t)chunk * 10 + tap) * (size_t)(pairs > co_n ? pairs : co_n);
  if (tap == 9) {
    for (int co = threadIdx.x; co < co_n; co += 256) {
      float acc = 0.f;
      for (int r = 0; r < rows; ++r) acc += s_dp[r * co_n + co];
      out[co] = acc;
    }
    return;
  }
  float acc[PPT];
  int pci[PPT], pco[PPT];
#pragma unroll
  for (int q = 0; q < PPT; ++q) {
    const int p = threadIdx.x + q * 256;
    acc[q] = 0.f;
    pci[q] = p < pairs ? p / co_n : 0;
    pco[q] = p < pairs ? p % co_n : 0;
  }
  for (int r = 0; r < rows; ++r) {
#pragma unroll
    for (int q = 0; q < PPT; ++q) acc[q] = fmaf(s_in[r * ci_n + pci[q]], s_dp[r * co_n + pco[q]], acc[q]);
  }
#pragma unroll
  for (int q = 0; q < PPT; ++q) {
    const int p = threadIdx.x + q * 256;
    if (p < pairs) out[p] = acc[q];
  }
}
// The same for the template width (co = 16), all nine taps in one CTA: the chunk's dpre rows and its input rows with a halo
// of W + 1 rows either side are staged once; a thread owns one ci and four co for all taps (36 accumulators, ten shared
// loads per 36 FMAs), the CTA's NS row subsets are added in ascending order at the end.  Same partial layout as above.
template <int CI, int CH>
struct WRowCfg {
  static constexpr int T = CI * 4;                                  // threads per row subset
  static constexpr int NS = (256 / T) < 8 ? (256 / T) : 8;          // row subsets per CTA
  static constexpr int HALO = 9;                                    // W + 1 <= 9
  static constexpr int STAGE = CH * 16 + (CH + 2 * HALO) * CI + CH / 2;   // floats (masks: 256 u16)
  static constexpr int RED = NS * (9 * CI * 16 + 16);
  static constexpr int SMEM = (STAGE > RED ? STAGE : RED) * 4;
};
template <int CI, int CH>
__global__ void __launch_bounds__(256) k_tr_conv_bwd_w_row16(const float* __restrict__ in, const float* __restrict__ dpre, float* __restrict__ part,
                                                              int N, int H, int W) {
  typedef WRowCfg<CI, CH> C;
  extern __shared__ __align__(16) float tw_smem[];
  float* s_dp = tw_smem;                                            // [CH][16]
  float* s_in = tw_smem + CH * 16;                           // [CH + 2 HALO][CI]
  unsigned short* s_mask = (unsigned short*)(s_in + (CH + 2 * C::HALO) * CI);
  const int chunk = blockIdx.x, total = N * H * W, row0 = chunk * CH, tid = threadIdx.x;
  for (int i = tid; i < CH * 16; i += 256) s_dp[i] = (row0 + i / 16 < total) ? dpre[(size_t)row0 * 16 + i] : 0.f;
  for (int i = tid; i < (CH + 2 * C::HALO) * CI; i += 256) {
    const long long e = (long long)row0 - C::HALO + i / CI;
    s_in[i] = (e >= 0 && e < total) ? in[(size_t)e * CI + i % CI] : 0.f;
  }
  if (tid < CH) {
    const int e = row0 + tid, cell = e % (H * W), y = cell / W, x = cell % W;
    unsigned m = 0;
    if (e < total)
      for (int tap = 0; tap < 9; ++tap) {
        const int yy = y + tap / 3 - 1, xx = x + tap % 3 - 1;
        if (yy >= 0 && yy < H && xx >= 0 && xx < W) m |= 1u << tap;
      }
    s_mask[tid] = (unsigned short)m;
  }
  __syncthreads();
  const int sub = tid / C::T, lt = tid % C::T, ci = lt >> 2, cq = lt & 3;
  float acc[9][4], accb[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int t9 = 0; t9 < 9; ++t9) { acc[t9][0] = acc[t9][1] = acc[t9][2] = acc[t9][3] = 0.f; }
  if (sub < C::NS) {
    for (int r = sub; r < CH; r += C::NS) {
      const unsigned m = s_mask[r];
      const float4 dp = *reinterpret_cast<const float4*>(s_dp + r * 16 + cq * 4);
      accb[0] += dp.x; accb[1] += dp.y; accb[2] += dp.z; accb[3] += dp.w;
#pragma unroll
      for (int t9 = 0; t9 < 9; ++t9) {
        if (m >> t9 & 1) {
          const float xv = s_in[(r + C::HALO + (t9 / 3 - 1) * W + (t9 % 3 - 1)) * CI + ci];
          acc[t9][0] = fmaf(xv, dp.x, acc[t9][0]); acc[t9][1] = fmaf(xv, dp.y, acc[t9][1]);
          acc[t9][2] = fmaf(xv, dp.z, acc[t9][2]); acc[t9][3] = fmaf(xv, dp.w, acc[t9][3]);
        }
      }
    }
  }
  __syncthreads();                                                  // staging is dead: the same memory takes the subsets' partials
  constexpr int PER = 9 * CI * 16 + 16;
  if (sub < C::NS) {
    float* red = tw_smem + sub * PER;
#pragma unroll
    for (int t9 = 0; t9 < 9; ++t9)
#pragma unroll
      for (int k = 0; k < 4; ++k) red[(t9 * CI + ci) * 16 + cq * 4 + k] = acc[t9][k];
    if (ci == 0)
#pragma unroll
      for (int k = 0; k < 4; ++k) red[9 * CI * 16 + cq * 4 + k] = accb[k];
  }
  __syncthreads();
  constexpr int pairs = CI * 16;
  float* out = part + (size_t)chunk * 10 * pairs;
  for (int i = tid; i < PER; i += 256) {
    float a = 0.f;
#pragma unroll
    for (int q = 0; q < C::NS; ++q) a += tw_smem[q * PER + i];
    out[i] = a;                                                     // [tap][ci][co] for i < 9 pairs, then "tap 9" = bias at 9 pairs + co: the layout of the tiled kernel
  }
}
// dw[tap][ci][co] = sum over chunks (ascending) of part[chunk][tap][...]; db likewise from "tap 9"
// (a warp per output: lane l adds chunks l, l + 32, ... in ascending order, then the fixed xor tree)
__global__ void k_tr_wgrad_reduce(const float* __restrict__ part, int chunks, int pairs, int co_n, float* __restrict__ dw, float* __restrict__ db) {
  const int stride = pairs > co_n ? pairs : co_n;
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (i >= 9 * pairs + co_n) return;
  const int tap = i < 9 * pairs ? i / pairs : 9, p = i < 9 * pairs ? i % pairs : i - 9 * pairs;
  float acc = 0.f;
  for (int c = lane; c < chunks; c += 32) acc += part[((size_t)c * 10 + tap) * stride + p];
#pragma unroll
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
  if (lane == 0) { if (tap < 9) dw[i] = acc; else db[p] = acc; }
}

// ---- heads -------------------------------------------------------------------------------------------------------------
struct HeadPtrs {
  const float* wv; const float* bv;          // folded value conv  [F], [1]
  const float* wp; const float* bp;          // folded policy conv [F][2], [2]
  const float* d1w; const float* d1b;        // [D], [D]
  const float* d2w; const float* d2b;        // [D], [1]
  const float* pw; const float* pb;          // [2][A], [A]
};
struct HeadBufs {
  float* vmap;      // [N][HW]       relu(value conv)
  float* pmap;      // [N][HW][2]    relu(policy conv)
  float* s;         // [N]           sum_hw vmap
  float* t;         // [N][2]        sum_hw pmap
  float* h;         // [N][D]        relu(s w1 + HW b1)
  float* value;     // [N]
  float* soft;      // [N][A]
  float* dz;        // [N][A]
  float* dh;        // [N][D]
  float* d_o;       // [N]
  float* dvmap;     // [N][HW]
  float* dpmap;     // [N][HW][2]
  float* colsum;    // [A]  sum_j label[j,a]
  float* scal;      // [8]: 0 C = sum colsum, 1 T = sum u, 2 lossEval, 3 lossPolicy, 4 lossParam, 5 loss
};

// one block per position: value conv / policy conv + ReLU, sums over H,W, dense layers, tanh, logits, softmax
__global__ void k_tr_heads_fwd(const float* __restrict__ a, HeadPtrs P, HeadBufs B, TrainDims dm) {
  __shared__ float red[32];
  __shared__ float sh[TR_MAX_DENSE + 4];
  const int n = blockIdx.x, tid = threadIdx.x;
  float sv = 0.f, s0 = 0.f, s1 = 0.f;
  for (int cell = tid; cell < dm.HW; cell += blockDim.x) {
    const float* ap = a + ((size_t)n * dm.HW + cell) * dm.F;
    float v = P.bv[0], p0 = P.bp[0], p1 = P.bp[1];
    for (int c = 0; c < dm.F; ++c) {
      v = fmaf(ap[c], P.wv[c], v);
      p0 = fmaf(ap[c], P.wp[c * 2], p0);
      p1 = fmaf(ap[c], P.wp[c * 2 + 1], p1);
    }
    v = fmaxf(v, 0.f); p0 = fmaxf(p0, 0.f); p1 = fmaxf(p1, 0.f);
    B.vmap[(size_t)n * dm.HW + cell] = v;
    B.pmap[((size_t)n * dm.HW + cell) * 2] = p0;
    B.pmap[((size_t)n * dm.HW + cell) * 2 + 1] = p1;
    sv += v; s0 += p0; s1 += p1;
  }
  sv = block_sum(sv, red);
  s0 = block_sum(s0, red);
  s1 = block_sum(s1, red);
  if (tid < dm.D) {
    const float hv = fmaxf(fmaf(sv, P.d1w[tid], (float)dm.HW * P.d1b[tid]), 0.f);
    B.h[(size_t)n * dm.D + tid] = hv;
    sh[tid] = hv * P.d2w[tid];
  }
  __syncthreads();
  if (tid == 0) {
    float o = P.d2b[0];
    for (int d = 0; d < dm.D; ++d) o += sh[d];
    B.value[n] = tanhf(o);
    B.d_o[n] = o;                                               // pre-activation, for an accurate tanh gradient (overwritten by the backward pass)
    B.s[n] = sv;
    B.t[n * 2] = s0;
    B.t[n * 2 + 1] = s1;
  }
  float m = -INFINITY;
  float* z = B.soft + (size_t)n * dm.A;
  for (int k = tid; k < dm.A; k += blockDim.x) {
    const float zz = fmaf(s0, P.pw[k], fmaf(s1, P.pw[dm.A + k], (float)dm.HW * P.pb[k]));
    z[k] = zz;
    m = fmaxf(m, zz);
  }
  m = block_max(m, red);
  float sum = 0.f;
  for (int k = tid; k < dm.A; k += blockDim.x) { const float e = expf(z[k] - m); z[k] = e; sum += e; }
  sum = block_sum(sum, red);
  for (int k = tid; k < dm.A; k += blockDim.x) z[k] = z[k] / sum;
}

// colsum[a] = sum_j label[j,a] (one block per action)
__global__ void k_tr_colsum(const float* __restrict__ label, float* __restrict__ colsum, int N, int A) {
  __shared__ float red[32];
  const int a = blockIdx.x;
  float acc = 0.f;
  for (int j = threadIdx.x; j < N; j += blockDim.x) acc += label[(size_t)j * A + a];
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) colsum[a] = acc;
}
// single block: C, T, lossEvaluation, lossPolicy
__global__ void k_tr_loss(const float* __restrict__ evalLabel, const float* __restrict__ noise, float eps, HeadBufs B, TrainDims dm, int N) {
  __shared__ float red[32];
  float c = 0.f;
  for (int a = threadIdx.x; a < dm.A; a += blockDim.x) c += B.colsum[a];
  c = block_sum(c, red);
  float t = 0.f;
  for (size_t i = threadIdx.x; i < (size_t)N * dm.A; i += blockDim.x) t += (1.f - eps) * B.soft[i] + eps * (noise ? noise[i % dm.A] : 0.f);
  t = block_sum(t, red);
  float le = 0.f;
  for (int n = threadIdx.x; n < N; n += blockDim.x) { const float d = B.value[n] - evalLabel[n]; le += d * d; }
  le = block_sum(le, red);
  const float logT = logf(t);
  float lp = 0.f;
  for (size_t i = threadIdx.x; i < (size_t)N * dm.A; i += blockDim.x) {
    const float ca = B.colsum[i % dm.A];
    if (ca != 0.f) lp += ca * (logf((1.f - eps) * B.soft[i] + eps * (noise ? noise[i % dm.A] : 0.f)) - logT);
  }
  lp = block_sum(lp, red);
  if (threadIdx.x == 0) {
    B.scal[0] = c;
    B.scal[1] = t;
    B.scal[2] = le / (float)N;
    B.scal[3] = -lp / ((float)N * (float)N);
  }
}
// lossParam partial: one block per variable in the l2 list; part[v] = sum(v^2)/2
struct L2Var { size_t off; int n; };
__global__ void k_tr_l2(const float* __restrict__ th, const L2Var* __restrict__ vars, float* __restrict__ part) {
  __shared__ float red[32];
  const L2Var v = vars[blockIdx.x];
  float acc = 0.f;
  for (int i = threadIdx.x; i < v.n; i += blockDim.x) { const float x = th[v.off + i]; acc = fmaf(x, x, acc); }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) part[blockIdx.x] = 0.5f * acc;
}
__global__ void k_tr_loss_total(const float* __restrict__ part, int n_l2, float* __restrict__ scal) {
  float s = 0.f;
  for (int i = 0; i < n_l2; ++i) s += part[i];
  scal[4] = s / (float)n_l2;
  scal[5] = scal[2] + scal[3] + scal[4];
}

// one block per position: gradients through tanh / dense / softmax down to the head feature maps and dA_L
__global__ void k_tr_heads_bwd(const float* __restrict__ a, const float* __restrict__ evalLabel, const float* __restrict__ noise, float eps,
                               HeadPtrs P, HeadBufs B, float* __restrict__ dA, TrainDims dm, int N) {
  __shared__ float red[32];
  const int n = blockIdx.x, tid = threadIdx.x;
  // d/do (tanh(o) - label)^2 / N = 2 (tanh(o) - label) sech^2(o) / N.  Saturated values (|tanh| -> 1, labels are +-1) lose
  // every digit of 1 - value^2 and of value - label in fp32, so both factors are formed from e = exp(-2|o|) instead:
  // sech^2 = 4 e / (1 + e)^2, 1 - tanh|o| = 2 e / (1 + e).
  const float o = B.d_o[n], label = evalLabel[n];
  __syncthreads();                                              // everyone has read the pre-activation before thread 0 overwrites it
  const float e = expf(-2.f * fabsf(o)), e1 = 1.f + e;
  const float sech2 = 4.f * e / (e1 * e1);
  float diff = B.value[n] - label;
  if (label == 1.f && o > 0.f) diff = -2.f * e / e1;
  else if (label == -1.f && o < 0.f) diff = 2.f * e / e1;
  const float d_o = 2.f * diff / (float)N * sech2;
  if (tid == 0) B.d_o[n] = d_o;
  float ds = 0.f;
  if (tid < dm.D) {
    const float dh = B.h[(size_t)n * dm.D + tid] > 0.f ? d_o * P.d2w[tid] : 0.f;
    B.dh[(size_t)n * dm.D + tid] = dh;
    ds = dh * P.d1w[tid];
  }
  ds = block_sum(ds, red);
  // policy: dL/du = (N C / T - colsum[a] / u) / N^2, dsoft = (1 - eps) du, dz = soft (dsoft - sum dsoft soft)
  const float C = B.scal[0], T = B.scal[1], invN2 = 1.f / ((float)N * (float)N);
  const float* soft = B.soft + (size_t)n * dm.A;
  float* dz = B.dz + (size_t)n * dm.A;
  float dot = 0.f;
  for (int k = tid; k < dm.A; k += blockDim.x) {
    const float u = (1.f - eps) * soft[k] + eps * (noise ? noise[k] : 0.f);
    const float ca = B.colsum[k];
    const float du = ((float)N * C / T - (ca != 0.f ? ca / u : 0.f)) * invN2;
    const float dsf = (1.f - eps) * du;
    dz[k] = dsf;
    dot = fmaf(dsf, soft[k], dot);
  }
  dot = block_sum(dot, red);
  float dt0 = 0.f, dt1 = 0.f;
  for (int k = tid; k < dm.A; k += blockDim.x) {
    const float g = soft[k] * (dz[k] - dot);
    dz[k] = g;
    dt0 = fmaf(g, P.pw[k], dt0);
    dt1 = fmaf(g, P.pw[dm.A + k], dt1);
  }
  dt0 = block_sum(dt0, red);
  dt1 = block_sum(dt1, red);
  for (int cell = tid; cell < dm.HW; cell += blockDim.x) {
    const size_t pc = (size_t)n * dm.HW + cell;
    const float dv = B.vmap[pc] > 0.f ? ds : 0.f;
    const float dp0 = B.pmap[pc * 2] > 0.f ? dt0 : 0.f, dp1 = B.pmap[pc * 2 + 1] > 0.f ? dt1 : 0.f;
    B.dvmap[pc] = dv;
    B.dpmap[pc * 2] = dp0;
    B.dpmap[pc * 2 + 1] = dp1;
    float* da = dA + pc * dm.F;
    for (int c = 0; c < dm.F; ++c) da[c] = fmaf(dv, P.wv[c], fmaf(dp0, P.wp[c * 2], dp1 * P.wp[c * 2 + 1]));
  }
}

// parameter gradients of the heads: one block per output element
//   [0,F) dwv' | F dbv' | [F+1, F+1+2F) dwp'[c][k] | +2 dbp'[k] | D dw1 | D db1 | D dw2 | 1 db2 | 2A dWp | A dbp
struct HeadGradOut { float* wv; float* bv; float* wp; float* bp; float* d1w; float* d1b; float* d2w; float* d2b; float* pw; float* pb; };
__global__ void k_tr_head_grads(const float* __restrict__ a, HeadBufs B, HeadGradOut G, TrainDims dm, int N, int first) {
  __shared__ float red[32];
  int b = blockIdx.x + first;                                       // first = 3F + 3 when k_tr_head_conv_grads has the convolutions
  const int F = dm.F, D = dm.D, A = dm.A, HW = dm.HW;
  const int tot = N * HW;
  float acc = 0.f;
  float* dst;
  if (b < F) { for (int e = threadIdx.x; e < tot; e += blockDim.x) acc = fmaf(a[(size_t)e * F + b], B.dvmap[e], acc); dst = G.wv + b; }
  else if ((b -= F) < 1) { for (int e = threadIdx.x; e < tot; e += blockDim.x) acc += B.dvmap[e]; dst = G.bv; }
  else if ((b -= 1) < 2 * F) { const int c = b / 2, k = b % 2; for (int e = threadIdx.x; e < tot; e += blockDim.x) acc = fmaf(a[(size_t)e * F + c], B.dpmap[(size_t)e * 2 + k], acc); dst = G.wp + b; }
  else if ((b -= 2 * F) < 2) { for (int e = threadIdx.x; e < tot; e += blockDim.x) acc += B.dpmap[(size_t)e * 2 + b]; dst = G.bp + b; }
  else if ((b -= 2) < D) { for (int n = threadIdx.x; n < N; n += blockDim.x) acc = fmaf(B.dh[(size_t)n * D + b], B.s[n], acc); dst = G.d1w + b; }
  else if ((b -= D) < D) { for (int n = threadIdx.x; n < N; n += blockDim.x) acc += B.dh[(size_t)n * D + b]; acc *= (float)HW; dst = G.d1b + b; }
  else if ((b -= D) < D) { for (int n = threadIdx.x; n < N; n += blockDim.x) acc = fmaf(B.d_o[n], B.h[(size_t)n * D + b], acc); dst = G.d2w + b; }
  else if ((b -= D) < 1) { for (int n = threadIdx.x; n < N; n += blockDim.x) acc += B.d_o[n]; dst = G.d2b; }
  else if ((b -= 1) < 2 * A) { const int k = b / A, aa = b % A; for (int n = threadIdx.x; n < N; n += blockDim.x) acc = fmaf(B.t[n * 2 + k], B.dz[(size_t)n * A + aa], acc); dst = G.pw + b; }
  else { b -= 2 * A; for (int n = threadIdx.x; n < N; n += blockDim.x) acc += B.dz[(size_t)n * A + b]; acc *= (float)HW; dst = G.pb + b; }
  // HW * sum for the per-cell dense biases: multiply BEFORE the block sum is wrong for partial sums only if done per thread -- it is linear, fine
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) *dst = acc;
}

// The 1x1 head convolutions' parameter gradients (the first 3F + 3 outputs above), row-chunked: a CTA takes TR_HCHUNK rows,
// thread (subset, c) accumulates a[e][c] * {dvmap, dpmap0, dpmap1}[e] over its subset's rows (coalesced rows of a), the
// subsets are added in ascending order; part[chunk][3F + 3] is reduced by k_tr_head_conv_reduce (a warp per output).
// One block per output read the whole activation map with stride F: 0.6 ms of a 3.4 ms step at batch 4096.
constexpr int TR_HCHUNK = 1024;
__global__ void __launch_bounds__(256) k_tr_head_conv_grads(const float* __restrict__ a, HeadBufs B, float* __restrict__ part, int F, int total) {
  __shared__ float red[16 * 3 * 17];                                 // [NS <= 16][3][F + 1] with F <= 16 ... sized by the host check F <= 16
  const int NS = 256 / F < 16 ? 256 / F : 16;
  const int sub = threadIdx.x / F, c = threadIdx.x % F;
  const int row0 = blockIdx.x * TR_HCHUNK, rows = min(TR_HCHUNK, total - row0);
  float av = 0.f, ap0 = 0.f, ap1 = 0.f, bv = 0.f, bp0 = 0.f, bp1 = 0.f;
  if (sub < NS)
    for (int r = sub; r < rows; r += NS) {
      const size_t e = (size_t)row0 + r;
      const float x = a[e * F + c], dv = B.dvmap[e], d0 = B.dpmap[e * 2], d1 = B.dpmap[e * 2 + 1];
      av = fmaf(x, dv, av); ap0 = fmaf(x, d0, ap0); ap1 = fmaf(x, d1, ap1);
      bv += dv; bp0 += d0; bp1 += d1;
    }
  if (sub < NS) {
    float* r = red + sub * 3 * (F + 1);
    r[c] = av; r[(F + 1) + c] = ap0; r[2 * (F + 1) + c] = ap1;
    if (c == 0) { r[F] = bv; r[(F + 1) + F] = bp0; r[2 * (F + 1) + F] = bp1; }
  }
  __syncthreads();
  // output order of k_tr_head_grads: [0,F) dwv | F dbv | [F+1, 3F+1) dwp[c][k] | 3F+1, 3F+2 dbp[k]
  for (int o = threadIdx.x; o < 3 * F + 3; o += 256) {
    int idx;
    if (o < F) idx = o;
    else if (o == F) idx = F;
    else if (o < 3 * F + 1) { const int cc = (o - F - 1) / 2, k = (o - F - 1) % 2; idx = (1 + k) * (F + 1) + cc; }
    else idx = (1 + (o - 3 * F - 1)) * (F + 1) + F;
    float s = 0.f;
    for (int q = 0; q < NS; ++q) s += red[q * 3 * (F + 1) + idx];
    part[(size_t)blockIdx.x * (3 * F + 3) + o] = s;
  }
}
__global__ void k_tr_head_conv_reduce(const float* __restrict__ part, int chunks, HeadGradOut G, int F) {
  const int o = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31, n_out = 3 * F + 3;
  if (o >= n_out) return;
  float acc = 0.f;
  for (int c = lane; c < chunks; c += 32) acc += part[(size_t)c * n_out + o];
#pragma unroll
  for (int k = 16; k; k >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, k);
  if (lane) return;
  if (o < F) G.wv[o] = acc;
  else if (o == F) G.bv[0] = acc;
  else if (o < 3 * F + 1) G.wp[o - F - 1] = acc;
  else G.bp[o - 3 * F - 1] = acc;
}

// ---- L2 term and optimizers --------------------------------------------------------------------------------------------
__global__ void k_tr_add_l2(const float* __restrict__ th, float* __restrict__ g, const L2Var* __restrict__ vars, int n_l2) {
  const L2Var v = vars[blockIdx.y];
  const float inv = 1.0f / (float)n_l2;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < v.n; i += gridDim.x * blockDim.x) g[v.off + i] += th[v.off + i] * inv;
}
// ---- batch assembly from the compact record stream (records.cuh: final protobuf bytes) ----------------------------------
// rec_off[i] = first byte of example i in `stream`.  State{mctsEval f32 (optional), mctsPolicy f64[A], boardEncoding int8[LB], ...}
__global__ void k_tr_batch_from_stream(const uint8_t* __restrict__ stream, const unsigned long long* __restrict__ rec_off, int A, int LB,
                                       int8_t* __restrict__ planes, float* __restrict__ evalLabel, float* __restrict__ policy) {
  const int n = blockIdx.x;
  const uint8_t* p = stream + rec_off[n];
  float ev = 0.f;
  if (p[0] == 0x0d) {                                          // field 1, fixed32
    const uint32_t u = (uint32_t)p[1] | ((uint32_t)p[2] << 8) | ((uint32_t)p[3] << 16) | ((uint32_t)p[4] << 24);
    ev = __uint_as_float(u);
    p += 5;
  }
  p += 1;                                                      // tag 0x12
  while (*p & 0x80) ++p;                                       // length varint (8 A)
  ++p;
  for (int a = threadIdx.x; a < A; a += blockDim.x) {
    unsigned long long bits = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) bits |= (unsigned long long)p[a * 8 + k] << (8 * k);
    policy[(size_t)n * A + a] = (float)__longlong_as_double((long long)bits);    // np.vstack of f64 rows fed to a float32 placeholder
  }
  p += (size_t)8 * A;
  p += 1;                                                      // tag 0x1a
  while (*p & 0x80) ++p;
  ++p;
  for (int i = threadIdx.x; i < LB; i += blockDim.x) planes[(size_t)n * LB + i] = (int8_t)p[i];
  if (threadIdx.x == 0) evalLabel[n] = ev;
}

// one Beta(alpha, 1 - alpha) draw per action, shared by the whole batch (dist.sample([1, A])[0][:, 0], NetworkFactory.py:176-180)
__global__ void k_tr_noise(float alpha, unsigned long long key, int A, float* out) {
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a < A) out[a] = beta_noise(alpha, key, (unsigned long long)a);
}

}  // namespace bb

using namespace bb;

struct bb_trainer {
  bb_train_desc desc{};
  TrainDims dm{};
  int device = 0;
  std::vector<ConvBN> convs;            // tower layers, then value.conv, policy.conv
  size_t n_params = 0, n_fold = 0;
  size_t o_d1w = 0, o_d1b = 0, o_d2w = 0, o_d2b = 0, o_pw = 0, o_pb = 0;
  float *th = nullptr, *g = nullptr, *m = nullptr, *v = nullptr, *wf = nullptr, *gwf = nullptr;
  uint8_t* mask = nullptr;
  L2Var* l2vars = nullptr;
  float* l2part = nullptr;
  int cap_B = 0;                        // batch capacity of the activation buffers
  std::vector<float*> act, dact;        // [L+1]: input, tower outputs
  HeadBufs hb{};
  float* hpart = nullptr;               // [chunks][3F + 3] partial gradients of the head convolutions
  float* wpart = nullptr;               // [chunks][10][pairs] partial weight gradients of the tiled kernel (per batch capacity)
  int8_t* planes_own = nullptr;         // batch assembled from a record stream
  float *eval_own = nullptr, *policy_own = nullptr, *noise_dev = nullptr;
  unsigned long long* recoff_dev = nullptr;
  std::vector<void*> allocs, batch_allocs;
  unsigned long long steps = 0;
  cudaStream_t stream = nullptr;
  // one captured graph per (batch size, input pointers, noise on/off)
  typedef std::tuple<int, const void*, const void*, const void*, int> Key;
  std::map<Key, cudaGraphExec_t> graphs;
  float* lr_dev = nullptr;              // lr_t lives in device memory so that a captured graph can be replayed with a new value
  ~bb_trainer() {
    cudaSetDevice(device);
    for (auto& kv : graphs) cudaGraphExecDestroy(kv.second);
    for (void* p : allocs) cudaFree(p);
    for (void* p : batch_allocs) cudaFree(p);
    if (stream) cudaStreamDestroy(stream);
  }
};

namespace {

template <class T>
int tr_alloc(std::vector<void*>& list, T** p, size_t count) {
  void* q = nullptr;
  cudaError_t e = cudaMalloc(&q, (count ? count : 1) * sizeof(T));
  if (e != cudaSuccess) { set_error("cudaMalloc of %zu bytes failed: %s", count * sizeof(T), cudaGetErrorString(e)); return BB_ENOMEM; }
  cudaMemset(q, 0, (count ? count : 1) * sizeof(T));
  list.push_back(q);
  *p = (T*)q;
  return BB_OK;
}

int ensure_batch(bb_trainer* t, int B) {
  if (B <= t->cap_B) return BB_OK;
  for (auto& kv : t->graphs) cudaGraphExecDestroy(kv.second);
  t->graphs.clear();
  for (void* p : t->batch_allocs) cudaFree(p);
  t->batch_allocs.clear();
  const TrainDims& d = t->dm;
  const size_t n = (size_t)B, cells = n * d.HW;
  int rc;
  t->act.assign(d.L + 1, nullptr);
  t->dact.assign(d.L + 1, nullptr);
#define A_(p, cnt) if ((rc = tr_alloc(t->batch_allocs, &p, cnt)) != BB_OK) return rc
  A_(t->act[0], cells * d.Cin);
  A_(t->dact[0], cells * d.Cin);
  for (int l = 1; l <= d.L; ++l) { A_(t->act[l], cells * d.F); A_(t->dact[l], cells * d.F); }
  A_(t->hb.vmap, cells); A_(t->hb.pmap, cells * 2); A_(t->hb.s, n); A_(t->hb.t, n * 2); A_(t->hb.h, n * d.D); A_(t->hb.value, n);
  A_(t->hb.soft, n * d.A); A_(t->hb.dz, n * d.A); A_(t->hb.dh, n * d.D); A_(t->hb.d_o, n); A_(t->hb.dvmap, cells); A_(t->hb.dpmap, cells * 2);
  A_(t->hb.colsum, (size_t)d.A); A_(t->hb.scal, 8);
  A_(t->planes_own, cells * d.Cin); A_(t->eval_own, n); A_(t->policy_own, n * d.A); A_(t->recoff_dev, n);
  t->wpart = nullptr;
  t->hpart = nullptr;
  if (d.F <= 16) A_(t->hpart, ((cells + TR_HCHUNK - 1) / TR_HCHUNK) * (size_t)(3 * d.F + 3));
  {
    const size_t Fm = (size_t)(d.F > d.Cin ? d.F : d.Cin);
    if (Fm * d.F <= 1024 && Fm + d.F <= 96) {                                     // the tiled weight-gradient path applies to every tower layer
      const size_t chunks = (cells + 63) / 64;                                     // (the quarter-chunk form of small batches)
      A_(t->wpart, (chunks < 4 * 296 ? chunks : (cells + TR_WCHUNK - 1) / TR_WCHUNK + 1) * 10 * Fm * d.F);
    }
  }
#undef A_
  t->cap_B = B;
  return BB_OK;
}

inline unsigned blocks_for(size_t n, int per = 256) { size_t b = (n + per - 1) / per; return (unsigned)(b < 1 ? 1 : (b > 65535 ? 65535 : b)); }

// the whole step as a sequence of launches on t->stream (captured into a graph by the caller)
int enqueue_step(bb_trainer* t, const int8_t* planes, const float* evalLabel, const float* policy, int B, bool noisy) {
  const TrainDims& d = t->dm;
  cudaStream_t s = t->stream;
  const float eps = noisy ? t->desc.epsilon : 0.f;
  const float* noise = noisy ? t->noise_dev : nullptr;
  const size_t cells = (size_t)B * d.HW;
  for (const ConvBN& c : t->convs) k_tr_fold<<<blocks_for((size_t)c.k * c.k * c.ci * c.co + c.co), 256, 0, s>>>(t->th, t->wf, c);
  k_tr_input<<<blocks_for(cells * d.Cin), 256, 0, s>>>(planes, t->act[0], cells * d.Cin);
  for (int l = 0; l < d.L; ++l) {
    const ConvBN& c = t->convs[l];
    const float* skip = (l > 0 && l % 2 == 0) ? t->act[l - 1] : nullptr;       // act index = layer + 1: block input of conv_2 (layer l) is act[l-1]
    const unsigned rb = blocks_for(cells, 128);
    if (c.co == 16 && c.ci == 16) k_tr_conv_fwd_row<16, 16><<<rb, 128, 0, s>>>(t->act[l], t->wf + c.wf, t->wf + c.bf, skip, t->act[l + 1], B, d.H, d.W);
    else if (c.co == 16 && c.ci == 3) k_tr_conv_fwd_row<3, 16><<<rb, 128, 0, s>>>(t->act[l], t->wf + c.wf, t->wf + c.bf, skip, t->act[l + 1], B, d.H, d.W);
    else if (c.co == 16 && c.ci == 17) k_tr_conv_fwd_row<17, 16><<<rb, 128, 0, s>>>(t->act[l], t->wf + c.wf, t->wf + c.bf, skip, t->act[l + 1], B, d.H, d.W);
    else k_tr_conv_fwd<<<blocks_for(cells * c.co), 256, 0, s>>>(t->act[l], t->wf + c.wf, t->wf + c.bf, skip, t->act[l + 1], B, d.H, d.W, c.ci, c.co);
  }
  const ConvBN& cv = t->convs[d.L];
  const ConvBN& cp = t->convs[d.L + 1];
  HeadPtrs P{t->wf + cv.wf, t->wf + cv.bf, t->wf + cp.wf, t->wf + cp.bf, t->th + t->o_d1w, t->th + t->o_d1b,
             t->th + t->o_d2w, t->th + t->o_d2b, t->th + t->o_pw, t->th + t->o_pb};
  k_tr_heads_fwd<<<B, 128, 0, s>>>(t->act[d.L], P, t->hb, d);
  k_tr_colsum<<<d.A, 64, 0, s>>>(policy, t->hb.colsum, B, d.A);
  k_tr_loss<<<1, 1024, 0, s>>>(evalLabel, noise, eps, t->hb, d, B);
  k_tr_l2<<<d.n_l2, 128, 0, s>>>(t->th, t->l2vars, t->l2part);
  k_tr_loss_total<<<1, 1, 0, s>>>(t->l2part, d.n_l2, t->hb.scal);
  // ---- backward ----
  k_tr_heads_bwd<<<B, 128, 0, s>>>(t->act[d.L], evalLabel, noise, eps, P, t->hb, t->dact[d.L], d, B);
  HeadGradOut G{t->gwf + cv.wf, t->gwf + cv.bf, t->gwf + cp.wf, t->gwf + cp.bf, t->g + t->o_d1w, t->g + t->o_d1b,
                t->g + t->o_d2w, t->g + t->o_d2b, t->g + t->o_pw, t->g + t->o_pb};
  if (d.F <= 16 && t->hpart) {
    const int hchunks = (int)((cells + TR_HCHUNK - 1) / TR_HCHUNK);
    k_tr_head_conv_grads<<<hchunks, 256, 0, s>>>(t->act[d.L], t->hb, t->hpart, d.F, (int)cells);
    k_tr_head_conv_reduce<<<(3 * d.F + 3 + 7) / 8, 256, 0, s>>>(t->hpart, hchunks, G, d.F);
    k_tr_head_grads<<<3 * d.D + 1 + 3 * d.A, 128, 0, s>>>(t->act[d.L], t->hb, G, d, B, 3 * d.F + 3);
  } else {
    k_tr_head_grads<<<3 * d.F + 3 + 3 * d.D + 1 + 3 * d.A, 128, 0, s>>>(t->act[d.L], t->hb, G, d, B, 0);
  }
  // dact[l+1] holds dL/d(output of layer l).  Skip gradients: output of layer l (even, > 0) = relu(conv + act[l-1]), so
  // dpre of layer l is added to dact[l-1]; that buffer is first WRITTEN by the data gradient of layer l-1... which runs
  // later.  So: the data gradient of layer l-1 accumulates into dact[l-1] when a skip has already been stored there.
  std::vector<char> has_skip(d.L + 1, 0);
  for (int l = d.L - 1; l >= 0; --l) {
    const ConvBN& c = t->convs[l];
    float* dskip = nullptr;
    if (l > 0 && l % 2 == 0) {
      dskip = t->dact[l - 1];
      cudaMemsetAsync(dskip, 0, cells * d.F * sizeof(float), s);
      has_skip[l - 1] = 1;
    }
    k_tr_relu_bwd<<<blocks_for(cells * c.co), 256, 0, s>>>(t->dact[l + 1], t->act[l + 1], dskip, cells * c.co);
    const int pairs = c.ci * c.co;
    if (pairs <= 1024 && c.ci + c.co <= 96 && t->wpart) {                             // narrow layers: tiled, two-stage (deterministic) reduction
      int chunks = (int)((cells + TR_WCHUNK - 1) / TR_WCHUNK);
      const size_t smem = (size_t)TR_WCHUNK * (c.ci + c.co) * sizeof(float);
      const bool row16 = c.co == 16 && d.W <= 8 && (c.ci == 16 || c.ci == 17 || c.ci == 3);
      const bool small = cells < (size_t)TR_WCHUNK * 296;                              // fewer than two CTAs per SM at the full chunk: quarter chunks
      if (row16 && small) chunks = (int)((cells + 63) / 64);
#define BB_WROW(CI_, CH_) k_tr_conv_bwd_w_row16<CI_, CH_><<<chunks, 256, WRowCfg<CI_, CH_>::SMEM, s>>>(t->act[l], t->dact[l + 1], t->wpart, B, d.H, d.W)
      if (row16 && c.ci == 16) { if (small) BB_WROW(16, 64); else BB_WROW(16, 256); }
      else if (row16 && c.ci == 17) { if (small) BB_WROW(17, 64); else BB_WROW(17, 256); }
      else if (row16 && c.ci == 3) { if (small) BB_WROW(3, 64); else BB_WROW(3, 256); }
#undef BB_WROW
      else if (pairs <= 256) k_tr_conv_bwd_w_tiled<1><<<dim3(chunks, 10), 256, smem, s>>>(t->act[l], t->dact[l + 1], t->wpart, B, d.H, d.W, c.ci, c.co);
      else k_tr_conv_bwd_w_tiled<4><<<dim3(chunks, 10), 256, smem, s>>>(t->act[l], t->dact[l + 1], t->wpart, B, d.H, d.W, c.ci, c.co);
      k_tr_wgrad_reduce<<<(9 * pairs + c.co + 7) / 8, 256, 0, s>>>(t->wpart, chunks, pairs, c.co, t->gwf + c.wf, t->gwf + c.bf);
    } else {
      k_tr_conv_bwd_w<<<9 * c.ci * c.co + c.co, 128, 0, s>>>(t->act[l], t->dact[l + 1], t->gwf + c.wf, t->gwf + c.bf, B, d.H, d.W, c.ci, c.co);
    }
    if (l > 0) {
      if (c.ci == 16 && c.co == 16) k_tr_conv_bwd_data_row<16, 16><<<blocks_for(cells, 128), 128, 0, s>>>(t->dact[l + 1], t->wf + c.wf, t->dact[l], has_skip[l] ? 1 : 0, B, d.H, d.W);
      else k_tr_conv_bwd_data<<<blocks_for(cells * c.ci), 256, 0, s>>>(t->dact[l + 1], t->wf + c.wf, t->dact[l], has_skip[l] ? 1 : 0, B, d.H, d.W, c.ci, c.co);
    }
  }
  for (const ConvBN& c : t->convs) k_tr_unfold<<<c.co, 128, 0, s>>>(t->th, t->gwf, t->g, c);
  k_tr_add_l2<<<dim3(8, d.n_l2), 256, 0, s>>>(t->th, t->g, t->l2vars, d.n_l2);
  if (cudaGetLastError() != cudaSuccess) { set_error("trainer launch failed"); return BB_ECUDA; }
  return BB_OK;
}

}  // namespace

namespace bb {
// optimizer update (NetworkFactory.py:229-242): kind 0 Adam (m, v slots; lr_t = lr sqrt(1 - beta2^t) / (1 - beta1^t) computed on the
// host), 1 momentum (accumulator in m), 2 plain gradient descent; mask[i] = 0 for the frozen BN statistics.  lr is read
// from device memory so that the captured graph can be replayed with a new value.
__global__ void k_tr_update_dev(float* th, const float* g, float* m, float* v, const uint8_t* mask, size_t n, int kind, const float* lr, float momentum) {
  const float l = *lr;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    if (!mask[i]) continue;
    const float gi = g[i];
    if (kind == 0) {
      const float mi = 0.9f * m[i] + 0.1f * gi;
      const float vi = 0.999f * v[i] + 0.001f * gi * gi;
      m[i] = mi;
      v[i] = vi;
      th[i] -= l * mi / (sqrtf(vi) + 1e-8f);
    } else if (kind == 1) {
      const float acc = momentum * m[i] + gi;
      m[i] = acc;
      th[i] -= l * acc;
    } else {
      th[i] -= l * gi;
    }
  }
}
}  // namespace bb

extern "C" {

int bb_trainer_create(const bb_train_desc* desc, const float* w, size_t n_floats, bb_trainer** out) {
  if (!desc || !w || !out) { set_error("bb_trainer_create: null argument"); return BB_EINVAL; }
  *out = nullptr;
  int h, wd, cin;
  if (bb_game_plane_shape(desc->game, &h, &wd, &cin) != BB_OK) return BB_EINVAL;
  const int A = bb_game_actions(desc->game), F = desc->filters, D = desc->dense, blocks = desc->blocks;
  if (F < 1 || F > 1024 || D < 1 || D > TR_MAX_DENSE || blocks < 0) { set_error("trainer: filters in [1,1024], dense in [1,%d], blocks >= 0", TR_MAX_DENSE); return BB_EINVAL; }
  if (desc->optimizer < 0 || desc->optimizer > 2) { set_error("trainer: optimizer 0 (adam), 1 (momentum) or 2 (sgd)"); return BB_EINVAL; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { set_error("no CUDA device: the trainer has no CPU fallback"); return BB_ECUDA; }
  if (desc->device < 0 || desc->device >= ndev) { set_error("device %d out of range", desc->device); return BB_EINVAL; }
  auto* t = new bb_trainer();
  t->desc = *desc;
  t->device = desc->device;
  TrainDims& d = t->dm;
  d.H = h; d.W = wd; d.HW = h * wd; d.Cin = cin; d.F = F; d.D = D; d.A = A; d.L = 1 + 2 * blocks;
  size_t off = 0, foff = 0;
  std::vector<L2Var> l2;
  auto add_conv = [&](int k, int ci, int co) {
    ConvBN c{};
    c.k = k; c.ci = ci; c.co = co;
    c.w = off; off += (size_t)k * k * ci * co;
    c.b = off; off += co; c.gamma = off; off += co; c.beta = off; off += co; c.mean = off; off += co; c.var = off; off += co;
    c.wf = foff; foff += (size_t)k * k * ci * co;
    c.bf = foff; foff += co;
    l2.push_back({c.w, k * k * ci * co});
    l2.push_back({c.gamma, co});
    l2.push_back({c.beta, co});
    t->convs.push_back(c);
  };
  add_conv(3, cin, F);
  for (int l = 1; l < d.L; ++l) add_conv(3, F, F);
  add_conv(1, F, 1);
  t->o_d1w = off; off += D; l2.push_back({t->o_d1w, D});
  t->o_d1b = off; off += D;
  t->o_d2w = off; off += D; l2.push_back({t->o_d2w, D});
  t->o_d2b = off; off += 1;
  add_conv(1, F, 2);
  t->o_pw = off; off += 2 * (size_t)A; l2.push_back({t->o_pw, 2 * A});
  t->o_pb = off; off += A;
  // the value conv was added before the dense kernels: keep convs = tower, value.conv, policy.conv (already in that order)
  if (off != n_floats) { set_error("bb_trainer_create: expected %zu floats for this shape, got %zu", off, n_floats); delete t; return BB_EINVAL; }
  t->n_params = off;
  t->n_fold = foff;
  d.n_l2 = (int)l2.size();
  if (cudaSetDevice(t->device) != cudaSuccess) { set_error("cudaSetDevice failed"); delete t; return BB_ECUDA; }
#define BB_WATTR(CI_, CH_) cudaFuncSetAttribute(k_tr_conv_bwd_w_row16<CI_, CH_>, cudaFuncAttributeMaxDynamicSharedMemorySize, WRowCfg<CI_, CH_>::SMEM)
  BB_WATTR(16, 256); BB_WATTR(17, 256); BB_WATTR(3, 256); BB_WATTR(16, 64); BB_WATTR(17, 64); BB_WATTR(3, 64);
#undef BB_WATTR
  cudaFuncSetAttribute(k_tr_conv_bwd_w_tiled<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, TR_WCHUNK * 96 * (int)sizeof(float));
  cudaFuncSetAttribute(k_tr_conv_bwd_w_tiled<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, TR_WCHUNK * 96 * (int)sizeof(float));
  int rc = BB_OK;
#define A_(p, cnt) if (rc == BB_OK) rc = tr_alloc(t->allocs, &p, cnt)
  A_(t->th, off); A_(t->g, off); A_(t->m, off); A_(t->v, off); A_(t->wf, foff); A_(t->gwf, foff); A_(t->mask, off);
  A_(t->l2vars, l2.size()); A_(t->l2part, l2.size()); A_(t->noise_dev, (size_t)A); A_(t->lr_dev, 1);
#undef A_
  if (rc != BB_OK) { delete t; return rc; }
  std::vector<uint8_t> mask(off, 1);
  for (const ConvBN& c : t->convs)
    for (int i = 0; i < c.co; ++i) mask[c.mean + i] = mask[c.var + i] = 0;
  bool ok = cudaMemcpy(t->th, w, off * sizeof(float), cudaMemcpyHostToDevice) == cudaSuccess &&
            cudaMemcpy(t->mask, mask.data(), off, cudaMemcpyHostToDevice) == cudaSuccess &&
            cudaMemcpy(t->l2vars, l2.data(), l2.size() * sizeof(L2Var), cudaMemcpyHostToDevice) == cudaSuccess &&
            cudaStreamCreateWithFlags(&t->stream, cudaStreamNonBlocking) == cudaSuccess;
  if (!ok) { set_error("trainer set-up failed: %s", cudaGetErrorString(cudaGetLastError())); delete t; return BB_ECUDA; }
  *out = t;
  return BB_OK;
}

int bb_trainer_destroy(bb_trainer* t) { delete t; return BB_OK; }

static int trainer_run(bb_trainer* t, const int8_t* planes, const float* evalLabel, const float* policy, int B, float lr, const float* noise_host) {
  const bool noisy = t->desc.epsilon > 0.f;
  if (noisy) {
    if (noise_host) BB_CUDA(cudaMemcpyAsync(t->noise_dev, noise_host, (size_t)t->dm.A * sizeof(float), cudaMemcpyHostToDevice, t->stream));
    else k_tr_noise<<<grid_for(t->dm.A, 128), 128, 0, t->stream>>>(t->desc.alpha, t->desc.noise_seed ^ (0x9E3779B97F4A7C15ull * (t->steps + 1)), t->dm.A, t->noise_dev);
  }
  t->steps += 1;
  float lr_t = lr;
  if (t->desc.optimizer == 0) lr_t = (float)((double)lr * sqrt(1.0 - pow(0.999, (double)t->steps)) / (1.0 - pow(0.9, (double)t->steps)));
  BB_CUDA(cudaMemcpyAsync(t->lr_dev, &lr_t, sizeof(float), cudaMemcpyHostToDevice, t->stream));
  bb_trainer::Key key(B, (const void*)planes, (const void*)evalLabel, (const void*)policy, noisy ? 1 : 0);
  auto it = t->graphs.find(key);
  if (it == t->graphs.end()) {
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    BB_CUDA(cudaStreamBeginCapture(t->stream, cudaStreamCaptureModeThreadLocal));
    int rc = enqueue_step(t, planes, evalLabel, policy, B, noisy);
    if (rc == BB_OK) {
      k_tr_update_dev<<<blocks_for(t->n_params), 256, 0, t->stream>>>(t->th, t->g, t->m, t->v, t->mask, t->n_params, t->desc.optimizer, t->lr_dev, t->desc.momentum);
    }
    cudaError_t e = cudaStreamEndCapture(t->stream, &graph);
    if (rc != BB_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
    if (e != cudaSuccess) { set_error("trainer graph capture failed: %s", cudaGetErrorString(e)); return BB_ECUDA; }
    e = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) { set_error("trainer graph instantiate failed: %s", cudaGetErrorString(e)); return BB_ECUDA; }
    if (t->graphs.size() > 64) { for (auto& kv : t->graphs) cudaGraphExecDestroy(kv.second); t->graphs.clear(); }
    it = t->graphs.emplace(key, exec).first;
  }
  BB_CUDA(cudaGraphLaunch(it->second, t->stream));
  return BB_OK;
}

int bb_trainer_step(bb_trainer* t, const int8_t* planes_dev, const float* eval_dev, const float* policy_dev, int B, float lr, const float* noise_host) {
  if (!t) { set_error("null trainer handle"); return BB_EINVAL; }
  if (!planes_dev || !eval_dev || !policy_dev || B <= 0) { set_error("bb_trainer_step: null batch or B <= 0"); return BB_EINVAL; }
  BB_CUDA(cudaSetDevice(t->device));
  int rc = ensure_batch(t, B);
  if (rc) return rc;
  rc = trainer_run(t, planes_dev, eval_dev, policy_dev, B, lr, noise_host);
  if (rc) return rc;
  BB_CUDA(cudaStreamSynchronize(t->stream));
  return BB_OK;
}

int bb_trainer_step_stream(bb_trainer* t, const void* stream_dev, const uint64_t* rec_off_host, int B, float lr, const float* noise_host) {
  if (!t) { set_error("null trainer handle"); return BB_EINVAL; }
  if (!stream_dev || !rec_off_host || B <= 0) { set_error("bb_trainer_step_stream: null argument or B <= 0"); return BB_EINVAL; }
  BB_CUDA(cudaSetDevice(t->device));
  int rc = ensure_batch(t, B);
  if (rc) return rc;
  BB_CUDA(cudaMemcpyAsync(t->recoff_dev, rec_off_host, (size_t)B * sizeof(unsigned long long), cudaMemcpyHostToDevice, t->stream));
  k_tr_batch_from_stream<<<B, 128, 0, t->stream>>>((const uint8_t*)stream_dev, t->recoff_dev, t->dm.A, t->dm.HW * t->dm.Cin, t->planes_own, t->eval_own, t->policy_own);
  BB_CUDA(cudaGetLastError());
  rc = trainer_run(t, t->planes_own, t->eval_own, t->policy_own, B, lr, noise_host);
  if (rc) return rc;
  BB_CUDA(cudaStreamSynchronize(t->stream));
  return BB_OK;
}

int bb_trainer_losses(bb_trainer* t, float out[4]) {
  if (!t || !out) { set_error("bb_trainer_losses: null argument"); return BB_EINVAL; }
  if (!t->hb.scal) { set_error("bb_trainer_losses: no step has run"); return BB_ESTATE; }
  BB_CUDA(cudaSetDevice(t->device));
  float s[8];
  BB_CUDA(cudaMemcpy(s, t->hb.scal, sizeof s, cudaMemcpyDeviceToHost));
  out[0] = s[5]; out[1] = s[2]; out[2] = s[3]; out[3] = s[4];
  return BB_OK;
}

static int tr_copy_out(bb_trainer* t, const float* src, float* dst, size_t n) {
  if (!t || !dst) { set_error("trainer: null argument"); return BB_EINVAL; }
  if (n != t->n_params) { set_error("trainer: expected a buffer of %zu floats, got %zu", t->n_params, n); return BB_EINVAL; }
  BB_CUDA(cudaSetDevice(t->device));
  BB_CUDA(cudaMemcpy(dst, src, n * sizeof(float), cudaMemcpyDeviceToHost));
  return BB_OK;
}
int bb_trainer_weights(bb_trainer* t, float* weights_host, size_t n) { return tr_copy_out(t, t ? t->th : nullptr, weights_host, n); }
int bb_trainer_grads(bb_trainer* t, float* grads_host, size_t n) { return tr_copy_out(t, t ? t->g : nullptr, grads_host, n); }
int bb_trainer_set_weights(bb_trainer* t, const float* w, size_t n) {
  if (!t || !w) { set_error("trainer: null argument"); return BB_EINVAL; }
  if (n != t->n_params) { set_error("trainer: expected %zu floats, got %zu", t->n_params, n); return BB_EINVAL; }
  BB_CUDA(cudaSetDevice(t->device));
  BB_CUDA(cudaMemcpy(t->th, w, n * sizeof(float), cudaMemcpyHostToDevice));
  return BB_OK;
}
uint64_t bb_trainer_steps(bb_trainer* t) { return t ? t->steps : 0; }

}  // extern "C"
