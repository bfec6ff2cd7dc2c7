// Global grad-norm + clip + AdamW + bf16 shadow refresh + gradient zeroing, as two passes over
// flat fp32 arenas.
//
// Replaces `torch.nn.utils.clip_grad_norm_` + `torch.optim.AdamW.step` + `zero_grad`
// (`src/llmtrain/training/trainer.py:362,390-394`; SURVEY §2.5 K13/K14, work item W10).  The
// reference sweeps parameter-sized memory three times per step plus a host sync for the norm;
// here pass 1 reads the gradients once (deterministic two-stage sum of squares, result stays on
// the device) and pass 2 reads p/m/v/g and writes p/m/v, the bf16 weight shadow the GEMMs
// consume, and zeros g for the next step — 16 B/param read, 14 B/param + 4 B/param written.
//
// Semantics matched to torch 2.11: clip coefficient = min(1, max_norm / (||g|| + 1e-6));
// decoupled weight decay p *= 1 - lr*wd applied to every parameter; bias-corrected moments with
// denom = sqrt(v)/sqrt(1 - beta2^t) + eps and step size lr / (1 - beta1^t).

#include <cmath>

#include "../common/host.cuh"
#include "../common/ptx.cuh"
#include "../../include/llmtrain_b200.h"

namespace b200 {
namespace optim {

constexpr int THREADS = 256;

__device__ __forceinline__ float block_sum(float v, float* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
  if (threadIdx.x < 32) {
    t = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  }
  return t;  // valid in thread 0
}

// partial[b] = sum over this CTA's grid-stride slice of g^2  (fixed assignment => deterministic)
__global__ void __launch_bounds__(THREADS)
sumsq_partial_kernel(const float* __restrict__ g, size_t n, float* __restrict__ partial) {
  pdl_enter();  // previous kernel finished, its writes are visible
  __shared__ float red[32];
  float acc = 0.f;
  const size_t n4 = n >> 2;
  const float4* g4 = reinterpret_cast<const float4*>(g);
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
       i += size_t(gridDim.x) * blockDim.x) {
    const float4 v = __ldg(g4 + i);
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const float v = g[(n4 << 2) + threadIdx.x];
    acc += v * v;
  }
  const float t = block_sum(acc, red);
  if (threadIdx.x == 0) partial[blockIdx.x] = t;
}

// out[0] (+)= sum(partial[0..nparts))   — one CTA, fixed order.
__global__ void __launch_bounds__(THREADS)
sumsq_finish_kernel(const float* __restrict__ partial, int nparts, float* __restrict__ out,
                    int accumulate) {
  pdl_enter();  // previous kernel finished, its writes are visible
  __shared__ float red[32];
  float acc = 0.f;
  for (int i = threadIdx.x; i < nparts; i += blockDim.x) acc += partial[i];
  const float t = block_sum(acc, red);
  if (threadIdx.x == 0) out[0] = accumulate ? out[0] + t : t;
}

struct AdamArgs {
  float lr, beta1, beta2, eps, weight_decay;
  float bias_corr1, bias_corr2_sqrt;  // 1 - beta1^t, sqrt(1 - beta2^t)
  float grad_scale;                   // applied to g before everything else (e.g. 1/world)
  float max_norm;                     // <= 0 disables clipping
};

__global__ void __launch_bounds__(THREADS)
adamw_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v,
             float* __restrict__ g, __nv_bfloat16* __restrict__ p16, size_t n,
             const float* __restrict__ sumsq, float* __restrict__ norm_out, AdamArgs a,
             int zero_grad) {
  pdl_enter();  // previous kernel finished, its writes are visible
  float coef = a.grad_scale;
  if (sumsq != nullptr) {
    const float norm = sqrtf(sumsq[0]) * a.grad_scale;
    if (a.max_norm > 0.f) coef *= fminf(1.0f, a.max_norm / (norm + 1e-6f));
    if (norm_out != nullptr && blockIdx.x == 0 && threadIdx.x == 0) norm_out[0] = norm;
  }
  const float decay = 1.0f - a.lr * a.weight_decay;
  const float step_size = a.lr / a.bias_corr1;
  const float one_m_b1 = 1.0f - a.beta1, one_m_b2 = 1.0f - a.beta2;
  const size_t n4 = n >> 2;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
       i += size_t(gridDim.x) * blockDim.x) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    float4 mv = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    const float4 gv = reinterpret_cast<const float4*>(g)[i];
    float* pp = reinterpret_cast<float*>(&pv);
    float* mp = reinterpret_cast<float*>(&mv);
    float* vp = reinterpret_cast<float*>(&vv);
    const float* gp = reinterpret_cast<const float*>(&gv);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float gj = gp[j] * coef;
      mp[j] = a.beta1 * mp[j] + one_m_b1 * gj;
      vp[j] = a.beta2 * vp[j] + one_m_b2 * gj * gj;
      const float denom = sqrtf(vp[j]) / a.bias_corr2_sqrt + a.eps;
      pp[j] = pp[j] * decay - step_size * (mp[j] / denom);
    }
    reinterpret_cast<float4*>(p)[i] = pv;
    reinterpret_cast<float4*>(m)[i] = mv;
    reinterpret_cast<float4*>(v)[i] = vv;
    if (p16 != nullptr) {
      reinterpret_cast<uint2*>(p16)[i] =
          make_uint2(pack_bf16x2(pp[0], pp[1]), pack_bf16x2(pp[2], pp[3]));
    }
    if (zero_grad) reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {  // scalar tail
    const size_t i = (n4 << 2) + threadIdx.x;
    const float gj = g[i] * coef;
    const float mj = a.beta1 * m[i] + one_m_b1 * gj;
    const float vj = a.beta2 * v[i] + one_m_b2 * gj * gj;
    const float pj = p[i] * decay - step_size * (mj / (sqrtf(vj) / a.bias_corr2_sqrt + a.eps));
    m[i] = mj; v[i] = vj; p[i] = pj;
    if (p16 != nullptr) p16[i] = __float2bfloat16(pj);
    if (zero_grad) g[i] = 0.f;
  }
}

inline int sumsq_grid(size_t n) {
  const size_t want = (n / 4 + THREADS - 1) / THREADS;
  const size_t cap = size_t(4 * sm_count());
  const size_t g = want < cap ? want : cap;
  return g < 1 ? 1 : static_cast<int>(g);
}

}  // namespace optim
}  // namespace b200

using namespace b200;

extern "C" size_t llmtrain_b200_grad_sumsq_scratch_bytes(size_t n) {
  return size_t(optim::sumsq_grid(n)) * sizeof(float);
}

// out[0] = (accumulate ? out[0] : 0) + sum(g[i]^2).  `scratch` holds the per-CTA partial sums.
extern "C" int llmtrain_b200_grad_sumsq(const float* g, size_t n, float* out, int accumulate,
                                        void* scratch, size_t scratch_bytes,
                                        cudaStream_t stream) {
  B200_REQUIRE((reinterpret_cast<uintptr_t>(g) & 15) == 0, "grad_sumsq: g must be 16-byte aligned");
  const int grid = optim::sumsq_grid(n);
  B200_REQUIRE(scratch_bytes >= size_t(grid) * sizeof(float), "grad_sumsq: scratch too small");
  B200_LAUNCH((optim::sumsq_partial_kernel), grid, optim::THREADS, 0, stream, g, n,
              static_cast<float*>(scratch));
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  B200_LAUNCH((optim::sumsq_finish_kernel), 1, optim::THREADS, 0, stream,
              static_cast<float*>(scratch), grid, out, accumulate);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}

extern "C" int llmtrain_b200_adamw(float* p, float* m, float* v, float* g, void* p_bf16, size_t n,
                                   const float* sumsq, float* norm_out, float lr, float beta1,
                                   float beta2, float eps, float weight_decay, int step,
                                   float grad_scale, float max_norm, int zero_grad,
                                   cudaStream_t stream) {
  B200_REQUIRE(step >= 1, "adamw: step counts from 1");
  B200_REQUIRE(((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(m) |
                 reinterpret_cast<uintptr_t>(v) | reinterpret_cast<uintptr_t>(g)) & 15) == 0,
               "adamw: arenas must be 16-byte aligned");
  if (n == 0) return 0;
  optim::AdamArgs a;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay;
  a.bias_corr1 = static_cast<float>(1.0 - std::pow(double(beta1), double(step)));
  a.bias_corr2_sqrt = static_cast<float>(std::sqrt(1.0 - std::pow(double(beta2), double(step))));
  a.grad_scale = grad_scale;
  a.max_norm = max_norm;
  const size_t want = (n / 4 + optim::THREADS - 1) / optim::THREADS;
  const size_t cap = size_t(8 * sm_count());
  const int grid = static_cast<int>(want < cap ? (want < 1 ? 1 : want) : cap);
  B200_LAUNCH((optim::adamw_kernel), grid, optim::THREADS, 0, stream, p, m, v, g,
              static_cast<__nv_bfloat16*>(p_bf16), n, sumsq, norm_out, a, zero_grad);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}
