// Counter-based dropout masks (Philox4x32-7).
//
// The reference draws its masks from torch's global CUDA generator inside four nn.Dropout modules
// (`src/llmtrain/models/gpt.py:28,30,97,129`, applied at :67, :71, :104, :179) and autograd keeps
// every mask alive until backward.  Here a mask is a pure function of
//     (seed, offset, site, element index)
// so forward and backward regenerate the same bits and nothing is stored: `seed` identifies the
// run (and rank), `offset` the micro-batch, `site` the dropout module (see DropSite), and the
// element index is `row * padded_cols + col` of the tensor the module acts on.
//
// One Philox call yields 4 x 32 bits = eight 16-bit lanes for eight CONSECUTIVE elements (a group
// is `index >> 3`); element k of the group is kept iff its lane >= threshold, threshold =
// round(p * 65536).  The keep probability is exact to 2^-16, and every kernel that holds 8, 32 or
// 64 consecutive columns of a row in one thread pays one call per 8 of them.
#pragma once

#include <cstdint>

namespace b200 {

struct DropoutParams {
  uint32_t threshold;  // drop iff lane < threshold; 0 disables dropout
  float scale;         // 1 / (1 - p)
  uint32_t seed_lo, seed_hi;
  uint32_t offset_lo, offset_hi;
  uint32_t site;
};

// site ids: 4 * layer + {0 attention probabilities, 1 attention residual, 2 MLP residual};
// the embedding dropout uses kDropSiteEmbedding.
constexpr uint32_t kDropSiteEmbedding = 0xFFFF0000u;

__host__ __device__ inline DropoutParams make_dropout(float p, uint64_t seed, uint64_t offset,
                                                      uint32_t site) {
  DropoutParams d;
  if (!(p > 0.f)) {
    d.threshold = 0;
    d.scale = 1.f;
  } else {
    uint32_t t = static_cast<uint32_t>(p * 65536.0f + 0.5f);
    if (t < 1) t = 1;
    if (t > 65535) t = 65535;
    d.threshold = t;
    d.scale = 65536.0f / static_cast<float>(65536u - t);  // exact inverse of the keep probability
  }
  d.seed_lo = static_cast<uint32_t>(seed);
  d.seed_hi = static_cast<uint32_t>(seed >> 32);
  d.offset_lo = static_cast<uint32_t>(offset);
  d.offset_hi = static_cast<uint32_t>(offset >> 32);
  d.site = site;
  return d;
}

#if defined(__CUDACC__)
struct Rand8 { uint32_t w[4]; };  // eight 16-bit lanes: element k -> (w[k >> 1] >> (16 * (k & 1))) & 0xFFFF

__device__ __forceinline__ Rand8 philox_group(const DropoutParams& d, uint64_t group) {
  constexpr uint32_t kM0 = 0xD2511F53u, kM1 = 0xCD9E8D57u;
  constexpr uint32_t kW0 = 0x9E3779B9u, kW1 = 0xBB67AE85u;
  uint32_t c0 = static_cast<uint32_t>(group), c1 = static_cast<uint32_t>(group >> 32);
  uint32_t c2 = d.site, c3 = d.offset_lo;
  uint32_t k0 = d.seed_lo, k1 = d.seed_hi ^ d.offset_hi;
#pragma unroll
  for (int r = 0; r < 7; ++r) {
    const uint32_t hi0 = __umulhi(kM0, c0), lo0 = kM0 * c0;
    const uint32_t hi1 = __umulhi(kM1, c2), lo1 = kM1 * c2;
    c0 = hi1 ^ c1 ^ k0;
    c1 = lo1;
    c2 = hi0 ^ c3 ^ k1;
    c3 = lo0;
    k0 += kW0;
    k1 += kW1;
  }
  return Rand8{{c0, c1, c2, c3}};
}

// keep factor (0 or scale) of element k (0..7) of a group
__device__ __forceinline__ float keep_factor(const DropoutParams& d, const Rand8& r, int k) {
  const uint32_t lane = (r.w[k >> 1] >> (16 * (k & 1))) & 0xFFFFu;
  return lane >= d.threshold ? d.scale : 0.f;
}
#endif

}  // namespace b200
