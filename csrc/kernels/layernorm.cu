// LayerNorm forward / backward over bf16 activations (fp32 statistics and parameters).
//
// Replaces nn.LayerNorm at `src/llmtrain/models/gpt.py:86,93,142` (used at :100,102,183) and
// its autograd backward (work item W4).  Memory-bound: one warp owns one row,
// the row is held in registers between the statistics pass and the normalise pass, all global
// traffic is 128-bit.  Backward also carries the residual-stream gradient: the block structure
// x_{l+1} = x_l + f(LN(x_l)) means d x_l = d x_{l+1} + LN'(…), so the kernel adds its result
// into the running residual gradient in place.  dgamma / dbeta use a deterministic two-stage
// reduction (per-CTA partial rows in scratch, then a fixed-order column sum) — no float atomics.

#include "../common/host.cuh"
#include "../common/ptx.cuh"
#include "../../include/llmtrain_b200.h"

namespace b200 {
namespace ln {

constexpr int MAX_VEC = 8;  // uint4 (8 bf16) chunks per lane -> d <= 2048
constexpr int WARPS = 8;
constexpr float EPS = 1e-5f;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z),
         d = unpack_bf16x2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]),
                    pack_bf16x2(f[6], f[7]));
}

// Forward.  A warp owns one row at a time and walks rows with a grid stride; gamma/beta for the
// columns a lane owns are loaded ONCE into registers (they were 4x the load instructions of the
// activations when re-read per row).  NV = uint4 chunks per lane.
template <int NV>
__global__ void __launch_bounds__(WARPS * 32)
fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma,
           const float* __restrict__ beta, __nv_bfloat16* __restrict__ y,
           float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows, int d) {
  pdl_enter();  // previous kernel finished, its writes are visible
  const int lane = threadIdx.x & 31;
  const int nvec = d >> 3;
  float g[NV][8], bt[NV][8];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = lane + i * 32;
    if (c < nvec) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma) + 2 * c);
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma) + 2 * c + 1);
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta) + 2 * c);
      const float4 b1 = __ldg(reinterpret_cast<const float4*>(beta) + 2 * c + 1);
      g[i][0] = g0.x; g[i][1] = g0.y; g[i][2] = g0.z; g[i][3] = g0.w;
      g[i][4] = g1.x; g[i][5] = g1.y; g[i][6] = g1.z; g[i][7] = g1.w;
      bt[i][0] = b0.x; bt[i][1] = b0.y; bt[i][2] = b0.z; bt[i][3] = b0.w;
      bt[i][4] = b1.x; bt[i][5] = b1.y; bt[i][6] = b1.z; bt[i][7] = b1.w;
    }
  }
  const float inv_d = 1.0f / d;
  // Two rows are prefetched ahead of the one being normalised (three 1.5 KB rows in flight per
  // warp at d = 768): the grid is sized to ONE resident wave so the gamma/beta prologue is paid
  // once per warp, and the prefetch depth supplies the bytes in flight that extra waves did not.
  const int stride = gridDim.x * WARPS;
  int row = blockIdx.x * WARPS + (threadIdx.x >> 5);
  uint4 n0[NV], n1[NV];
  auto fetch = [&](uint4 (&dst)[NV], int r) {
    if (r < rows) {
      const uint4* xr = reinterpret_cast<const uint4*>(x + size_t(r) * d);
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int c = lane + i * 32;
        if (c < nvec) dst[i] = __ldg(xr + c);
      }
    }
  };
  fetch(n0, row);
  fetch(n1, row + stride);
  for (; row < rows; row += stride) {
    float v[NV][8];
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + i * 32;
      if (c < nvec) {
        unpack8(n0[i], v[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) sum += v[i][j];
      }
      n0[i] = n1[i];
    }
    fetch(n1, row + 2 * stride);
    const float mean = warp_sum(sum) * inv_d;
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      if (lane + i * 32 < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { const float t = v[i][j] - mean; sq += t * t; }
      }
    }
    const float rstd = rsqrtf(warp_sum(sq) * inv_d + EPS);
    uint4* yr = reinterpret_cast<uint4*>(y + size_t(row) * d);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + i * 32;
      if (c < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * g[i][j] + bt[i][j];
        yr[c] = pack8(o);
      }
    }
    if (lane == 0) {
      if (mean_out) mean_out[row] = mean;
      if (rstd_out) rstd_out[row] = rstd;
    }
  }
}

// Backward.  Each CTA walks rows blockIdx.x, blockIdx.x + gridDim.x, ... ; every lane keeps the
// dgamma/dbeta partial sums of the columns it owns in registers, warps are combined through
// shared memory in a fixed order, and the CTA writes ONE partial row to scratch.
// NV = uint4 chunks per lane (ceil(d / 256)); inputs stay packed (bf16) between the two passes.
// COLSUM: additionally produce column sums of the UPDATED residual gradient — that tensor is the
// output gradient of the preceding out_proj / mlp_proj, so its column sum is that layer's bias
// gradient and the separate reduction pass over it disappears.
template <int NV, bool COLSUM>
__global__ void __launch_bounds__(WARPS * 32)
bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
           const float* __restrict__ gamma, const float* __restrict__ mean,
           const float* __restrict__ rstd, __nv_bfloat16* __restrict__ dresid,
           float* __restrict__ partial /* [gridDim.x][NS][d], NS = COLSUM ? 3 : 2 */, int rows,
           int d) {
  pdl_enter();  // previous kernel finished, its writes are visible
  constexpr int NS = COLSUM ? 3 : 2;
  extern __shared__ float sh[];  // [WARPS][NS][d]
  float dc[COLSUM ? NV : 1][8];
#pragma unroll
  for (int i = 0; i < (COLSUM ? NV : 1); ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) dc[i][j] = 0.f;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int nvec = d >> 3;
  float dg[NV][8], db[NV][8], g[NV][8];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) { dg[i][j] = 0.f; db[i][j] = 0.f; g[i][j] = 0.f; }
    const int c = lane + i * 32;
    if (c < nvec) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma) + 2 * c);
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma) + 2 * c + 1);
      g[i][0] = g0.x; g[i][1] = g0.y; g[i][2] = g0.z; g[i][3] = g0.w;
      g[i][4] = g1.x; g[i][5] = g1.y; g[i][6] = g1.z; g[i][7] = g1.w;
    }
  }

  // Software-pipelined over rows: the three 16-byte streams of row r+stride (dy, x, dresid) are
  // requested before row r is reduced, so ~2x the bytes are in flight per warp.  With ~200
  // registers per thread only 8 warps fit on an SM, and one row per warp in flight left the
  // kernel latency-bound at a third of HBM bandwidth.
  const int stride = gridDim.x * WARPS;
  int row = blockIdx.x * WARPS + warp;
  uint4 na[NV], nb[NV], nr[NV];
  float nmu = 0.f, nrs = 0.f;
  auto fetch = [&](int r) {
    const uint4* dyr = reinterpret_cast<const uint4*>(dy + size_t(r) * d);
    const uint4* xr = reinterpret_cast<const uint4*>(x + size_t(r) * d);
    const uint4* drr = reinterpret_cast<const uint4*>(dresid + size_t(r) * d);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + i * 32;
      if (c < nvec) { na[i] = __ldg(dyr + c); nb[i] = __ldg(xr + c); nr[i] = drr[c]; }
    }
    nmu = mean[r];
    nrs = rstd[r];
  };
  constexpr bool PIPE = NV <= 4;  // wider rows would spill with two rows held in registers
  if (PIPE && row < rows) fetch(row);
  while (row < rows) {
    if (!PIPE) fetch(row);
    uint4 pa[NV], pb[NV], pr[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) { pa[i] = na[i]; pb[i] = nb[i]; pr[i] = nr[i]; }
    const float mu = nmu, rs = nrs;
    const int next = row + stride;
    if (PIPE && next < rows) fetch(next);
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + i * 32;
      if (c < nvec) {
        float a[8], b[8];
        unpack8(pa[i], a);
        unpack8(pb[i], b);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float h = (b[j] - mu) * rs;
          dg[i][j] += a[j] * h;
          db[i][j] += a[j];
          const float w = a[j] * g[i][j];
          s1 += w;
          s2 += w * h;
        }
      }
    }
    s1 = warp_sum(s1) / d;
    s2 = warp_sum(s2) / d;
    uint4* drr = reinterpret_cast<uint4*>(dresid + size_t(row) * d);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + i * 32;
      if (c < nvec) {
        float a[8], b[8], r[8];
        unpack8(pa[i], a);
        unpack8(pb[i], b);
        unpack8(pr[i], r);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float h = (b[j] - mu) * rs;
          r[j] += rs * (a[j] * g[i][j] - s1 - h * s2);
          if (COLSUM) dc[i][j] += r[j];
        }
        drr[c] = pack8(r);
      }
    }
    row = next;
  }

  float* my = sh + size_t(warp) * NS * d;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = lane + i * 32;
    if (c < nvec) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        my[c * 8 + j] = dg[i][j];
        my[d + c * 8 + j] = db[i][j];
        if (COLSUM) my[2 * d + c * 8 + j] = dc[i][j];
      }
    }
  }
  __syncthreads();
  float* out = partial + size_t(blockIdx.x) * NS * d;
  for (int idx = threadIdx.x; idx < NS * d; idx += blockDim.x) {
    float acc = 0.f;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) acc += sh[size_t(w) * NS * d + idx];
    out[idx] = acc;
  }
}

// Fixed-order column sum of `nparts` partial rows, accumulated into the fp32 gradient arena.
// Block = 32 columns x 8 part-lanes: lane py sums parts py, py+8, ... (coalesced 128 B reads per
// part row), then the 8 partials of a column are combined in a fixed order through smem.
constexpr int FIN_COLS = 32, FIN_LANES = 8;
__global__ void __launch_bounds__(FIN_COLS * FIN_LANES)
finish_kernel(const float* __restrict__ partial, int nparts, int d, int ns,
              float* __restrict__ dgamma, float* __restrict__ dbeta,
              float* __restrict__ dcolsum) {
  pdl_enter();  // previous kernel finished, its writes are visible
  __shared__ float sh[FIN_LANES][FIN_COLS];
  const int cx = threadIdx.x & (FIN_COLS - 1), py = threadIdx.x / FIN_COLS;
  const int idx = blockIdx.x * FIN_COLS + cx;
  float acc = 0.f;
  if (idx < ns * d) {
    for (int p = py; p < nparts; p += FIN_LANES) acc += partial[size_t(p) * ns * d + idx];
  }
  sh[py][cx] = acc;
  __syncthreads();
  if (py == 0 && idx < ns * d) {
    float t = sh[0][cx];
#pragma unroll
    for (int k = 1; k < FIN_LANES; ++k) t += sh[k][cx];
    if (idx < d) dgamma[idx] += t;
    else if (idx < 2 * d) dbeta[idx - d] += t;
    else dcolsum[idx - 2 * d] += t;
  }
}

inline int bwd_grid(int rows) {
  int g = 2 * sm_count();
  const int need = ceil_div(rows, WARPS);
  return g < need ? g : need;
}

}  // namespace ln
}  // namespace b200

using namespace b200;

extern "C" int llmtrain_b200_layernorm_fwd(const void* x, const float* gamma, const float* beta,
                                           void* y, float* mean, float* rstd, int rows, int d,
                                           cudaStream_t stream) {
  B200_REQUIRE(d % 8 == 0 && d <= ln::MAX_VEC * 256, "layernorm: d must be a multiple of 8, <= 2048");
  if (rows == 0) return 0;
  const auto* x_ = static_cast<const __nv_bfloat16*>(x);
  auto* y_ = static_cast<__nv_bfloat16*>(y);
  // One resident wave: as many CTAs per SM as the instantiation's registers allow, at most 2.
  const int need = ceil_div(rows, ln::WARPS);
#define B200_LN_FWD(NV)                                                                         \
  case NV: {                                                                                    \
    static int per_sm = 0;                                                                      \
    if (per_sm == 0) {                                                                          \
      B200_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(                            \
          &per_sm, ln::fwd_kernel<NV>, ln::WARPS * 32, 0));                                     \
      per_sm = per_sm < 1 ? 1 : (per_sm > 2 ? 2 : per_sm);                                      \
    }                                                                                           \
    int grid = per_sm * sm_count();                                                             \
    if (grid > need) grid = need;                                                               \
    B200_LAUNCH((ln::fwd_kernel<NV>), grid, ln::WARPS * 32, 0, stream, x_, gamma, beta, y_,     \
                mean, rstd, rows, d);                                                           \
    break;                                                                                      \
  }
  switch (ceil_div(d, 256)) {
    B200_LN_FWD(1) B200_LN_FWD(2) B200_LN_FWD(3) B200_LN_FWD(4)
    B200_LN_FWD(5) B200_LN_FWD(6) B200_LN_FWD(7) B200_LN_FWD(8)
    default: return fail("layernorm_fwd: unsupported width");
  }
#undef B200_LN_FWD
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}

extern "C" size_t llmtrain_b200_layernorm_bwd_scratch_bytes(int rows, int d) {
  return size_t(ln::bwd_grid(rows)) * 3 * d * sizeof(float);
}

extern "C" int llmtrain_b200_layernorm_bwd(const void* dy, const void* x, const float* gamma,
                                           const float* mean, const float* rstd, void* dresid,
                                           float* dgamma, float* dbeta, float* dcolsum,
                                           void* scratch, size_t scratch_bytes, int rows, int d,
                                           cudaStream_t stream) {
  B200_REQUIRE(d % 8 == 0 && d <= ln::MAX_VEC * 256, "layernorm: d must be a multiple of 8, <= 2048");
  if (rows == 0) return 0;
  const int grid = ln::bwd_grid(rows);
  const int ns = dcolsum != nullptr ? 3 : 2;
  B200_REQUIRE(scratch_bytes >= size_t(grid) * ns * d * sizeof(float),
               "layernorm_bwd: scratch too small (see llmtrain_b200_layernorm_bwd_scratch_bytes)");
  const size_t smem = size_t(ln::WARPS) * ns * d * sizeof(float);
  const auto* dy_ = static_cast<const __nv_bfloat16*>(dy);
  const auto* x_ = static_cast<const __nv_bfloat16*>(x);
  auto* dr_ = static_cast<__nv_bfloat16*>(dresid);
  auto* part_ = static_cast<float*>(scratch);
  const int nv = ceil_div(d, 256);
#define B200_LN_BWD(NV)                                                                        \
  case NV: {                                                                                   \
    static bool configured = false;                                                            \
    if (!configured) {                                                                         \
      B200_CHECK_CUDA(cudaFuncSetAttribute(ln::bwd_kernel<NV, false>,                          \
                                           cudaFuncAttributeMaxDynamicSharedMemorySize,        \
                                           200 * 1024));                                       \
      B200_CHECK_CUDA(cudaFuncSetAttribute(ln::bwd_kernel<NV, true>,                           \
                                           cudaFuncAttributeMaxDynamicSharedMemorySize,        \
                                           200 * 1024));                                       \
      configured = true;                                                                       \
    }                                                                                          \
    if (ns == 3) {                                                                             \
      B200_LAUNCH((ln::bwd_kernel<NV, true>), grid, ln::WARPS * 32, smem, stream, dy_, x_,     \
                  gamma, mean, rstd, dr_, part_, rows, d);                                     \
    } else {                                                                                   \
      B200_LAUNCH((ln::bwd_kernel<NV, false>), grid, ln::WARPS * 32, smem, stream, dy_, x_,    \
                  gamma, mean, rstd, dr_, part_, rows, d);                                     \
    }                                                                                          \
    break;                                                                                     \
  }
  switch (nv) {
    B200_LN_BWD(1) B200_LN_BWD(2) B200_LN_BWD(3) B200_LN_BWD(4)
    B200_LN_BWD(5) B200_LN_BWD(6) B200_LN_BWD(7) B200_LN_BWD(8)
    default: return fail("layernorm_bwd: unsupported width");
  }
#undef B200_LN_BWD
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  B200_LAUNCH((ln::finish_kernel), ceil_div(ns * d, ln::FIN_COLS), ln::FIN_COLS * ln::FIN_LANES, 0,
              stream, static_cast<float*>(scratch), grid, d, ns, dgamma, dbeta, dcolsum);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}
