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
           int d, int stages, int warp_bytes) {
  pdl_enter();  // previous kernel finished, its writes are visible
  constexpr int NS = COLSUM ? 3 : 2;
  // Per-warp window of `warp_bytes`: first a ring of `stages` row slots (dy | x | dresid rows,
  // filled by bulk copies), at the end the warp's NS x d partial column sums.
  extern __shared__ __align__(128) uint8_t sh_raw[];
  __shared__ uint64_t full_bar[WARPS][4];
  // All per-column state is held as packed fp32 pairs (f32x2, ptx.cuh): the kernel is paced by
  // instruction issue and dependency latency at 8 warps per SM (~200 registers per thread), and
  // FFMA2 halves the instruction count of both sweeps.
  f32x2 dc[COLSUM ? NV : 1][4];
#pragma unroll
  for (int i = 0; i < (COLSUM ? NV : 1); ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) dc[i][j] = splat2(0.f);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int nvec = d >> 3;
  f32x2 dg[NV][4], db[NV][4], g[NV][4];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
#pragma unroll
    for (int j = 0; j < 4; ++j) { dg[i][j] = splat2(0.f); db[i][j] = splat2(0.f); g[i][j] = splat2(0.f); }
    const int c = lane + i * 32;
    if (c < nvec) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma) + 2 * c);
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma) + 2 * c + 1);
      g[i][0] = pack2(g0.x, g0.y); g[i][1] = pack2(g0.z, g0.w);
      g[i][2] = pack2(g1.x, g1.y); g[i][3] = pack2(g1.z, g1.w);
    }
  }

  // Rows reach the warp through a private ring in shared memory: lane 0 issues three untiled bulk
  // copies (cp.async.bulk; dy, x and dresid rows are contiguous) per slot, `stages` rows ahead,
  // and refills a slot as soon as its row has been copied to registers.  Holding the prefetched
  // rows in registers instead (one row ahead was all that fitted) left the loads of the next row
  // exposed: ncu showed the warps idle on the long scoreboard with the issue slots 40% used.
  const int stride = gridDim.x * WARPS;
  const uint32_t row_bytes = uint32_t(d) * 2u;
  const uint32_t slot_bytes = 3u * row_bytes;
  uint8_t* ring = sh_raw + size_t(warp) * warp_bytes;
  uint64_t* full = full_bar[warp];
  if (lane == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(&full[s], 1);
    fence_barrier_init();
  }
  __syncwarp();
  auto request = [&](int r, int slot) {  // lane 0 only
    uint8_t* dst = ring + size_t(slot) * slot_bytes;
    mbar_expect_tx(&full[slot], slot_bytes);
    bulk_load_1d(dst, dy + size_t(r) * d, row_bytes, &full[slot]);
    bulk_load_1d(dst + row_bytes, x + size_t(r) * d, row_bytes, &full[slot]);
    bulk_load_1d(dst + 2 * row_bytes, dresid + size_t(r) * d, row_bytes, &full[slot]);
  };
  int row = blockIdx.x * WARPS + warp;
  if (lane == 0) {
    for (int s = 0; s < stages; ++s)
      if (row + s * stride < rows) request(row + s * stride, s);
  }
  int slot = 0;
  uint32_t parity = 0;
  while (row < rows) {
    const float mu = mean[row], rs = rstd[row];
    mbar_wait(&full[slot], parity);
    uint4 pa[NV], pb[NV], pr[NV];
    {
      const uint4* sa = reinterpret_cast<const uint4*>(ring + size_t(slot) * slot_bytes);
      const uint4* sb = reinterpret_cast<const uint4*>(ring + size_t(slot) * slot_bytes + row_bytes);
      const uint4* sr =
          reinterpret_cast<const uint4*>(ring + size_t(slot) * slot_bytes + 2 * row_bytes);
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int c = lane + i * 32;
        if (c < nvec) { pa[i] = sa[c]; pb[i] = sb[c]; pr[i] = sr[c]; }
      }
    }
    __syncwarp();  // every lane has its copy: the slot may be overwritten
    if (lane == 0 && row + stages * stride < rows) request(row + stages * stride, slot);
    if (++slot == stages) { slot = 0; parity ^= 1; }
    const int next = row + stride;
    // h = (x - mu) * rs = x * rs + (-mu * rs)
    const f32x2 rs2 = splat2(rs), nmurs2 = splat2(-mu * rs);
    f32x2 s1p = splat2(0.f), s2p = splat2(0.f);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + i * 32;
      if (c < nvec) {
        const uint32_t aw[4] = {pa[i].x, pa[i].y, pa[i].z, pa[i].w};
        const uint32_t bw[4] = {pb[i].x, pb[i].y, pb[i].z, pb[i].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 af = unpack_bf16x2(aw[j]), bf = unpack_bf16x2(bw[j]);
          const f32x2 a = pack2(af.x, af.y);
          const f32x2 h = fma2(pack2(bf.x, bf.y), rs2, nmurs2);
          dg[i][j] = fma2(a, h, dg[i][j]);
          db[i][j] = add2(db[i][j], a);
          const f32x2 w = mul2(a, g[i][j]);
          s1p = add2(s1p, w);
          s2p = fma2(w, h, s2p);
        }
      }
    }
    float s1, s2, t0, t1;
    unpack2(s1p, t0, t1);
    s1 = t0 + t1;
    unpack2(s2p, t0, t1);
    s2 = t0 + t1;
    s1 = warp_sum(s1) / d;
    s2 = warp_sum(s2) / d;
    const f32x2 ns1 = splat2(-s1), ns2 = splat2(-s2);
    uint4* drr = reinterpret_cast<uint4*>(dresid + size_t(row) * d);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + i * 32;
      if (c < nvec) {
        const uint32_t aw[4] = {pa[i].x, pa[i].y, pa[i].z, pa[i].w};
        const uint32_t bw[4] = {pb[i].x, pb[i].y, pb[i].z, pb[i].w};
        const uint32_t rw[4] = {pr[i].x, pr[i].y, pr[i].z, pr[i].w};
        uint32_t ow[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 af = unpack_bf16x2(aw[j]), bf = unpack_bf16x2(bw[j]),
                       rf = unpack_bf16x2(rw[j]);
          const f32x2 h = fma2(pack2(bf.x, bf.y), rs2, nmurs2);
          // r += rs * (a * g - s1 - h * s2)
          f32x2 t = fma2(pack2(af.x, af.y), g[i][j], ns1);
          t = fma2(h, ns2, t);
          const f32x2 r = fma2(t, rs2, pack2(rf.x, rf.y));
          if (COLSUM) dc[i][j] = add2(dc[i][j], r);
          float r0, r1;
          unpack2(r, r0, r1);
          ow[j] = pack_bf16x2(r0, r1);
        }
        drr[c] = make_uint4(ow[0], ow[1], ow[2], ow[3]);
      }
    }
    row = next;
  }

  float* my = reinterpret_cast<float*>(ring);  // the ring is idle now: reuse it for the sums
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = lane + i * 32;
    if (c < nvec) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float lo, hi;
        unpack2(dg[i][j], lo, hi);
        my[c * 8 + 2 * j] = lo; my[c * 8 + 2 * j + 1] = hi;
        unpack2(db[i][j], lo, hi);
        my[d + c * 8 + 2 * j] = lo; my[d + c * 8 + 2 * j + 1] = hi;
        if (COLSUM) {
          unpack2(dc[i][j], lo, hi);
          my[2 * d + c * 8 + 2 * j] = lo; my[2 * d + c * 8 + 2 * j + 1] = hi;
        }
      }
    }
  }
  __syncthreads();
  float* out = partial + size_t(blockIdx.x) * NS * d;
  for (int idx = threadIdx.x; idx < NS * d; idx += blockDim.x) {
    float acc = 0.f;
#pragma unroll
    for (int w = 0; w < WARPS; ++w)
      acc += reinterpret_cast<const float*>(sh_raw + size_t(w) * warp_bytes)[idx];
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
  int g = sm_count();  // one resident wave (the register footprint allows one CTA per SM)
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
  // per-warp window: ring of row slots (3 rows of d bf16 each), reused for NS x d partial sums
  const size_t slot_bytes = size_t(3) * d * 2;
  int stages = int((size_t(200) * 1024 / ln::WARPS) / slot_bytes);
  stages = stages > 3 ? 3 : (stages < 1 ? 1 : stages);
  size_t warp_bytes = stages * slot_bytes;
  if (warp_bytes < size_t(ns) * d * sizeof(float)) warp_bytes = size_t(ns) * d * sizeof(float);
  warp_bytes = (warp_bytes + 127) & ~size_t(127);
  const size_t smem = size_t(ln::WARPS) * warp_bytes;
  B200_REQUIRE(smem <= size_t(200) * 1024, "layernorm_bwd: shared-memory budget exceeded");
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
                  gamma, mean, rstd, dr_, part_, rows, d, stages, int(warp_bytes));            \
    } else {                                                                                   \
      B200_LAUNCH((ln::bwd_kernel<NV, false>), grid, ln::WARPS * 32, smem, stream, dy_, x_,    \
                  gamma, mean, rstd, dr_, part_, rows, d, stages, int(warp_bytes));            \
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
