// Policy-noise sampler shared by the network kernels (net.cu, net_tc.cuh) and the trainer (train.cu).
#pragma once
#include <stdint.h>

namespace bb {

// counter-based uniform in (0,1): splitmix64 of (seed, call, element, draw)
__device__ __forceinline__ float u01(unsigned long long key, unsigned long long ctr) {
  unsigned long long z = key + 0x9E3779B97F4A7C15ull * (ctr + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return ((float)(z >> 40) + 0.5f) * (1.0f / 16777216.0f);
}
// dist.sample(...)[..., 0] of Dirichlet([alpha, 1-alpha]) == Beta(alpha, 1-alpha)  (A.17), by Joehnk's method: X = U^(1/a),
// Y = V^(1/b), accept when X + Y <= 1, return X / (X + Y).  For b = 1 - a the acceptance probability is
// Gamma(1+a) Gamma(2-a) >= 0.785, so a draw costs ~2.4 uniforms and four MUFU operations on average (the first version,
// two Marsaglia-Tsang gammas with Box-Muller normals, logf and powf, cost ~3 k cycles per action and sat on the network
// pipelines' critical path: 7.4 k -> 49 k cycles of heads per batch with the template's epsilon 0.3).
__device__ __forceinline__ float beta_noise(float alpha, unsigned long long key, unsigned long long elem) {
  unsigned long long ctr = elem << 6;
  const float ia = 1.0f / alpha, ib = 1.0f / (1.0f - alpha);
  for (int it = 0; it < 32; ++it) {
    const float u = u01(key, ctr++), v = u01(key, ctr++);
    const float x = exp2f(ia * __log2f(u)), y = exp2f(ib * __log2f(v));
    const float t = x + y;
    if (t <= 1.0f && t > 0.f) return x / t;
  }
  return alpha;                                            // unreachable in practice: 0.215^32
}

}  // namespace bb
