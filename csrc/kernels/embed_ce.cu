// Embedding forward/backward, fused softmax-cross-entropy (forward + dlogits in one pass), bias
// gradient (column sums) and dtype casts.
//
// Reference sites: token+position embedding `src/llmtrain/models/gpt.py:176-179` (K1);
// cross-entropy + (masked) mean `src/llmtrain/models/gpt.py:256-269` (K11); the bias gradients
// are the autograd of the four nn.Linear biases (K12).  All of these are HBM-bound, so every
// kernel makes exactly one pass over its large operand with 128-bit accesses.

#include <cstdlib>

#include "../common/host.cuh"
#include "../common/ptx.cuh"
#include "../../include/llmtrain_b200.h"

namespace b200 {
namespace ew {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ------------------------------------------------------------------------------------------------
// x[m,:] = wte[ids[m],:] + wpe[m % T,:]        (bf16 tables, bf16 out; one warp per row)
// ------------------------------------------------------------------------------------------------
__global__ void embed_fwd_kernel(const int64_t* __restrict__ ids,
                                 const __nv_bfloat16* __restrict__ wte,
                                 const __nv_bfloat16* __restrict__ wpe,
                                 __nv_bfloat16* __restrict__ x, int rows, int T, int d, int vocab) {
  pdl_enter();  // previous kernel finished, its writes are visible
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  int64_t tok = ids[row];
  tok = tok < 0 ? 0 : (tok >= vocab ? vocab - 1 : tok);  // never read out of bounds
  const uint4* te = reinterpret_cast<const uint4*>(wte + size_t(tok) * d);
  const uint4* pe = reinterpret_cast<const uint4*>(wpe + size_t(row % T) * d);
  uint4* out = reinterpret_cast<uint4*>(x + size_t(row) * d);
  for (int c = lane; c < (d >> 3); c += 32) {
    const uint4 a = __ldg(te + c), b = __ldg(pe + c);
    const float2 a0 = unpack_bf16x2(a.x), a1 = unpack_bf16x2(a.y), a2 = unpack_bf16x2(a.z),
                 a3 = unpack_bf16x2(a.w);
    const float2 b0 = unpack_bf16x2(b.x), b1 = unpack_bf16x2(b.y), b2 = unpack_bf16x2(b.z),
                 b3 = unpack_bf16x2(b.w);
    out[c] = make_uint4(pack_bf16x2(a0.x + b0.x, a0.y + b0.y), pack_bf16x2(a1.x + b1.x, a1.y + b1.y),
                        pack_bf16x2(a2.x + b2.x, a2.y + b2.y), pack_bf16x2(a3.x + b3.x, a3.y + b3.y));
  }
}

// dwte[ids[m],:] += dx[m,:]   (fp32 arena; vector red.add — rows of one token may collide)
__global__ void embed_bwd_wte_kernel(const int64_t* __restrict__ ids,
                                     const __nv_bfloat16* __restrict__ dx,
                                     float* __restrict__ dwte, int rows, int d, int vocab) {
  pdl_enter();  // previous kernel finished, its writes are visible
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  int64_t tok = ids[row];
  if (tok < 0 || tok >= vocab) return;
  const uint2* src = reinterpret_cast<const uint2*>(dx + size_t(row) * d);
  float* dst = dwte + size_t(tok) * d;
  for (int c = lane; c < (d >> 2); c += 32) {
    const uint2 u = __ldg(src + c);
    const float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y);
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + 4 * c), "f"(f0.x),
                 "f"(f0.y), "f"(f1.x), "f"(f1.y)
                 : "memory");
  }
}

// Deterministic variant of the scatter (llmtrain_b200_set_deterministic): every vocabulary row gets
// ONE writer — the warp of the first position that holds the token — which walks the later
// positions in ascending order and adds their rows in registers before a single plain update of
// dwte[token].  first[] is found with atomicMin (an order-independent result).
__global__ void embed_first_kernel(const int64_t* __restrict__ ids, int* __restrict__ first,
                                   int rows, int vocab) {
  pdl_enter();
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= rows) return;
  const int64_t tok = ids[row];
  if (tok >= 0 && tok < vocab) atomicMin(first + tok, row);
}

constexpr int EMB_MAX_V4 = 16;  // float4 chunks per lane: d <= 2048
__global__ void embed_bwd_wte_ordered_kernel(const int64_t* __restrict__ ids,
                                             const __nv_bfloat16* __restrict__ dx,
                                             const int* __restrict__ first,
                                             float* __restrict__ dwte, int rows, int d, int vocab) {
  pdl_enter();
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const int64_t tok = ids[row];
  if (tok < 0 || tok >= vocab || first[tok] != row) return;  // warp-uniform
  const int nv = d >> 2;
  float4 acc[EMB_MAX_V4];
  auto add_row = [&](int r, bool init) {
    const uint2* src = reinterpret_cast<const uint2*>(dx + size_t(r) * d);
#pragma unroll
    for (int i = 0; i < EMB_MAX_V4; ++i) {
      const int c = lane + i * 32;
      if (c < nv) {
        const uint2 u = __ldg(src + c);
        const float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y);
        if (init) acc[i] = make_float4(f0.x, f0.y, f1.x, f1.y);
        else { acc[i].x += f0.x; acc[i].y += f0.y; acc[i].z += f1.x; acc[i].w += f1.y; }
      }
    }
  };
  add_row(row, true);
  for (int base = row + 1; base < rows; base += 32) {
    const int r = base + lane;
    const bool hit = r < rows && ids[r] == tok;
    unsigned m = __ballot_sync(0xffffffffu, hit);
    while (m != 0) {  // ascending positions
      const int bit = __ffs(m) - 1;
      m &= m - 1;
      add_row(base + bit, false);
    }
  }
  float4* dst = reinterpret_cast<float4*>(dwte + size_t(tok) * d);
#pragma unroll
  for (int i = 0; i < EMB_MAX_V4; ++i) {
    const int c = lane + i * 32;
    if (c < nv) {
      float4 cur = dst[c];
      cur.x += acc[i].x; cur.y += acc[i].y; cur.z += acc[i].z; cur.w += acc[i].w;
      dst[c] = cur;
    }
  }
}

// dwpe[t,:] += sum_b dx[b*T + t,:]   (one thread per (t, 4 columns); fixed order over b)
__global__ void embed_bwd_wpe_kernel(const __nv_bfloat16* __restrict__ dx,
                                     float* __restrict__ dwpe, int B, int T, int d) {
  pdl_enter();  // previous kernel finished, its writes are visible
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int per_row = d >> 2;
  if (idx >= T * per_row) return;
  const int t = idx / per_row, c = idx - t * per_row;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int b = 0; b < B; ++b) {
    const uint2 u = __ldg(reinterpret_cast<const uint2*>(dx + (size_t(b) * T + t) * d) + c);
    const float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y);
    acc.x += f0.x; acc.y += f0.y; acc.z += f1.x; acc.w += f1.y;
  }
  float4* out = reinterpret_cast<float4*>(dwpe + size_t(t) * d) + c;
  float4 cur = *out;
  cur.x += acc.x; cur.y += acc.y; cur.z += acc.z; cur.w += acc.w;
  *out = cur;
}

// ------------------------------------------------------------------------------------------------
// Fused cross-entropy: one CTA per row.  Pass 1 streams the (padded) vocabulary row once from HBM
// keeping an online (max, sum-exp) pair per thread; loss_row = lse - logit[label].  Pass 2
// (training only) re-reads the row — 100 KB, still L2-resident — and overwrites it in place with
// dlogits = (softmax - onehot) * grad_scale.  Padded columns [V, ldv) get zero.  Few registers and
// 256-thread CTAs keep 8 rows in flight per SM so the loads of one row hide the math of another.
// ------------------------------------------------------------------------------------------------
constexpr int CE_THREADS = 256;

__device__ __forceinline__ void online_merge(float& m, float& s, float m2, float s2) {
  const float mn = fmaxf(m, m2);
  const float a = (m == -INFINITY) ? 0.f : __expf(m - mn);
  const float b = (m2 == -INFINITY) ? 0.f : __expf(m2 - mn);
  s = s * a + s2 * b;
  m = mn;
}

template <bool WRITE_GRAD>
__global__ void __launch_bounds__(CE_THREADS)
ce_kernel(__nv_bfloat16* __restrict__ logits, const int64_t* __restrict__ labels,
          const uint8_t* __restrict__ row_valid, float* __restrict__ loss_rows, int V, int ldv,
          float grad_scale) {
  pdl_enter();  // previous kernel finished, its writes are visible
  __shared__ float red_m[CE_THREADS / 32], red_s[CE_THREADS / 32];
  __shared__ float bc_m, bc_s;
  const int row = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint4* rp = reinterpret_cast<uint4*>(logits + size_t(row) * ldv);
  const int nvec = ldv >> 3;
  const int label = static_cast<int>(labels[row]);
  // label == -100 is torch's ignore_index (`F.cross_entropy` in the reference's loss,
  // `src/llmtrain/models/gpt.py:256-260`): zero loss, zero gradient, still counted by the caller's
  // denominator.
  const bool valid = (row_valid == nullptr || row_valid[row] != 0) && label != -100;

  float m = -INFINITY, ssum = 0.f;
  for (int c = tid; c < nvec; c += CE_THREADS) {
    const uint4 u = rp[c];
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
    float f[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 t = unpack_bf16x2(w[j]);
      const int col = c * 8 + 2 * j;
      f[2 * j] = col < V ? t.x : -INFINITY;
      f[2 * j + 1] = col + 1 < V ? t.y : -INFINITY;
    }
    float cm = f[0];
#pragma unroll
    for (int j = 1; j < 8; ++j) cm = fmaxf(cm, f[j]);
    if (cm > -INFINITY) {
      float cs = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) cs += __expf(f[j] - cm);  // exp(-inf) = 0 for padding
      online_merge(m, ssum, cm, cs);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, m, o);
    const float s2 = __shfl_xor_sync(0xffffffffu, ssum, o);
    online_merge(m, ssum, m2, s2);
  }
  if (lane == 0) { red_m[warp] = m; red_s[warp] = ssum; }
  __syncthreads();
  if (warp == 0) {
    float mm = lane < CE_THREADS / 32 ? red_m[lane] : -INFINITY;
    float ss = lane < CE_THREADS / 32 ? red_s[lane] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float m2 = __shfl_xor_sync(0xffffffffu, mm, o);
      const float s2 = __shfl_xor_sync(0xffffffffu, ss, o);
      online_merge(mm, ss, m2, s2);
    }
    if (lane == 0) {
      bc_m = mm; bc_s = ss;
      const float picked = (label >= 0 && label < V)
                               ? __bfloat162float(logits[size_t(row) * ldv + label]) : 0.f;
      loss_rows[row] = valid ? (__logf(ss) + mm - picked) : 0.f;
    }
  }
  if (!WRITE_GRAD) return;
  __syncthreads();
  const float mx = bc_m;
  const float inv = valid ? grad_scale / bc_s : 0.f;
  const float hot = valid ? grad_scale : 0.f;
  for (int c = tid; c < nvec; c += CE_THREADS) {
    const uint4 u = rp[c];
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_bf16x2(w[j]);
      const int col = c * 8 + 2 * j;
      float g0 = 0.f, g1 = 0.f;
      if (col < V) g0 = __expf(f.x - mx) * inv - (col == label ? hot : 0.f);
      if (col + 1 < V) g1 = __expf(f.y - mx) * inv - (col + 1 == label ? hot : 0.f);
      o[j] = pack_bf16x2(g0, g1);
    }
    rp[c] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// Row-resident variant (the one the trainer runs when two rows fit in shared memory, i.e.
// ldv <= CE_ROW_MAX_LDV): a persistent CTA per SM owns rows blockIdx.x, +gridDim.x, ... and keeps
// each row in shared memory across three sweeps, so HBM sees ONE read and ONE write per logit
// instead of two reads and a write, and the next row streams in (untiled bulk copy,
// cp.async.bulk) underneath the math of the current one.
//   sweep A  exact row max on PACKED bf16 (4 max instructions per 8 logits)
//   sweep B  sum of exp2((x - max) * log2 e): one FFMA + one MUFU per logit
//   sweep C  dlogits = exp2(..) / sum * scale - onehot, stored straight to global (in place)
// ce_kernel's online (max, sum) merge spends ~10 instructions and 1.25 MUFU per logit in its
// first pass; with 100 K exponentials per row the MUFU pipe, not HBM, was the limiter.
// Chunk c of a row is touched by thread c % CE_ROW_THREADS in every sweep, so the owner can patch
// its padding columns to -inf in shared memory without a barrier.
constexpr int CE_ROW_THREADS = 512;
constexpr int CE_ROW_CHUNK = 16384;                             // bytes per bulk copy
constexpr int CE_ROW_MAX_LDV = (220 * 1024 / 2) / 2 / 64 * 64;  // two bf16 rows in <= 220 KB

__device__ __forceinline__ uint32_t bf16x2_max(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("max.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}

template <bool WRITE_GRAD>
__global__ void __launch_bounds__(CE_ROW_THREADS, 1)
ce_row_kernel(__nv_bfloat16* __restrict__ logits, const int64_t* __restrict__ labels,
              const uint8_t* __restrict__ row_valid, float* __restrict__ loss_rows, int rows,
              int V, int ldv, float grad_scale) {
  pdl_enter();  // previous kernel finished, its writes are visible
  extern __shared__ __align__(128) uint8_t ce_row_smem[];
  __shared__ uint64_t full[2];
  __shared__ float red_m[CE_ROW_THREADS / 32], red_s[CE_ROW_THREADS / 32];
  constexpr int NW = CE_ROW_THREADS / 32;
  constexpr float LOG2E = 1.4426950408889634f, LN2 = 0.6931471805599453f;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t row_bytes = uint32_t(ldv) * 2u;
  const uint32_t pitch = (row_bytes + 127u) & ~127u;
  const int nvec = ldv >> 3;
  const int first_pad_chunk = V >> 3;  // chunks >= this one contain columns >= V

  if (tid == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    fence_barrier_init();
  }
  __syncthreads();

  auto load_row = [&](int r, int b) {
    mbar_expect_tx(&full[b], row_bytes);
    const uint8_t* src = reinterpret_cast<const uint8_t*>(logits + size_t(r) * ldv);
    uint8_t* dst = ce_row_smem + size_t(b) * pitch;
    for (uint32_t off = 0; off < row_bytes; off += CE_ROW_CHUNK) {
      const uint32_t n = row_bytes - off < uint32_t(CE_ROW_CHUNK) ? row_bytes - off : CE_ROW_CHUNK;
      bulk_load_1d(dst + off, src + off, n, &full[b]);
    }
  };

  int row = blockIdx.x;
  if (tid == 0 && row < rows) load_row(row, 0);
  for (int it = 0; row < rows; row += gridDim.x, ++it) {
    const int b = it & 1;
    // Every thread finished reading buffer b^1 before the barrier that ended the last iteration.
    if (tid == 0 && row + int(gridDim.x) < rows) load_row(row + gridDim.x, b ^ 1);
    const int label = static_cast<int>(labels[row]);
    const bool valid = (row_valid == nullptr || row_valid[row] != 0) && label != -100;  // ignore_index
    uint4* rp = reinterpret_cast<uint4*>(ce_row_smem + size_t(b) * pitch);
    mbar_wait(&full[b], (it >> 1) & 1);

    // Padding columns [V, ldv) -> -inf, by the thread that owns the chunk in every sweep.
    for (int c = tid; c < nvec; c += CE_ROW_THREADS) {
      if (c >= first_pad_chunk) {
        __nv_bfloat16* e = reinterpret_cast<__nv_bfloat16*>(rp + c);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (c * 8 + j >= V) e[j] = __ushort_as_bfloat16(0xFF80u);
        fence_proxy_async_smem();  // ordered before the bulk load that later reuses this buffer
      }
    }

    // Sweep A: packed max.
    uint32_t pm = 0xFF80FF80u;
#pragma unroll 4
    for (int c = tid; c < nvec; c += CE_ROW_THREADS) {
      const uint4 u = rp[c];
      pm = bf16x2_max(pm, bf16x2_max(bf16x2_max(u.x, u.y), bf16x2_max(u.z, u.w)));
    }
    const float2 pmf = unpack_bf16x2(pm);
    float m = fmaxf(pmf.x, pmf.y);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) red_m[warp] = m;
    __syncthreads();
    m = red_m[0];
#pragma unroll
    for (int w = 1; w < NW; ++w) m = fmaxf(m, red_m[w]);
    const float mxl = m * LOG2E;

    // Sweep B: sum of exponentials.
    float s0 = 0.f, s1 = 0.f;
#pragma unroll 2
    for (int c = tid; c < nvec; c += CE_ROW_THREADS) {
      const uint4 u = rp[c];
      const float2 a = unpack_bf16x2(u.x), bb = unpack_bf16x2(u.y), cc = unpack_bf16x2(u.z),
                   dd = unpack_bf16x2(u.w);
      s0 += fast_exp2(fmaf(a.x, LOG2E, -mxl)) + fast_exp2(fmaf(bb.x, LOG2E, -mxl)) +
            fast_exp2(fmaf(cc.x, LOG2E, -mxl)) + fast_exp2(fmaf(dd.x, LOG2E, -mxl));
      s1 += fast_exp2(fmaf(a.y, LOG2E, -mxl)) + fast_exp2(fmaf(bb.y, LOG2E, -mxl)) +
            fast_exp2(fmaf(cc.y, LOG2E, -mxl)) + fast_exp2(fmaf(dd.y, LOG2E, -mxl));
    }
    float ssum = s0 + s1;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ssum += __shfl_xor_sync(0xffffffffu, ssum, o);
    if (lane == 0) red_s[warp] = ssum;
    __syncthreads();
    ssum = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) ssum += red_s[w];  // fixed order: every thread gets the same bits

    if (tid == 0) {
      const float picked = (label >= 0 && label < V)
                               ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(rp)[label])
                               : 0.f;
      loss_rows[row] = valid ? (__log2f(ssum) * LN2 + m - picked) : 0.f;
    }

    if (WRITE_GRAD) {
      // Sweep C: gradients straight to global memory (the row's own storage).
      const float inv = valid ? grad_scale / ssum : 0.f;
      const float hot = valid ? grad_scale : 0.f;
      const int label_chunk = (label >= 0 && label < V) ? (label >> 3) : -1;
      uint4* out = reinterpret_cast<uint4*>(logits + size_t(row) * ldv);
#pragma unroll 2
      for (int c = tid; c < nvec; c += CE_ROW_THREADS) {
        const uint4 u = rp[c];
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 f = unpack_bf16x2(w[j]);
          o[j] = pack_bf16x2(fast_exp2(fmaf(f.x, LOG2E, -mxl)) * inv,
                             fast_exp2(fmaf(f.y, LOG2E, -mxl)) * inv);
        }
        out[c] = make_uint4(o[0], o[1], o[2], o[3]);
      }
      // One-hot term: a scalar fix-up by the thread that stored the label's chunk (same thread,
      // same address, program order => it lands after the vector store).
      if (label_chunk >= 0 && (label_chunk % CE_ROW_THREADS) == tid) {
        const float x = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(rp)[label]);
        logits[size_t(row) * ldv + label] =
            __float2bfloat16(fast_exp2(fmaf(x, LOG2E, -mxl)) * inv - hot);
      }
    }
    __syncthreads();  // buffer b and red_* may be reused
  }
}

// loss_acc[0] += scale * sum(loss_rows)  — single CTA, fixed order => deterministic.
__global__ void loss_reduce_kernel(const float* __restrict__ loss_rows, int rows, float scale,
                                   float* __restrict__ loss_acc) {
  pdl_enter();  // previous kernel finished, its writes are visible
  __shared__ float red[32];
  float acc = 0.f;
  for (int i = threadIdx.x; i < rows; i += blockDim.x) acc += loss_rows[i];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
    t = warp_sum(t);
    if (threadIdx.x == 0) loss_acc[0] += t * scale;
  }
}

// ------------------------------------------------------------------------------------------------
// Column sums: db[n] += sum_m dy[m,n].  Grid (ceil(N/256), splits); a warp streams whole rows with
// 16-byte loads (a lane owns 8 adjacent columns, a warp 256); warps combine through smem; splits
// combine through scratch in a fixed order (second kernel).
// ------------------------------------------------------------------------------------------------
constexpr int CS_WARPS = 8;
constexpr int CS_COLS = 256;
__global__ void __launch_bounds__(CS_WARPS * 32)
colsum_partial_kernel(const __nv_bfloat16* __restrict__ dy, int ld, int M, int N,
                      float* __restrict__ partial /* [splits][N] */) {
  pdl_enter();  // previous kernel finished, its writes are visible
  __shared__ float sh[CS_WARPS][CS_COLS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int col = blockIdx.x * CS_COLS + lane * 8;
  const int rows_per = (M + gridDim.y - 1) / gridDim.y;
  const int r0 = blockIdx.y * rows_per;
  const int r1 = min(M, r0 + rows_per);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  const bool vec_ok = (col + 8 <= N) && ((ld & 7) == 0) &&
                      ((reinterpret_cast<uintptr_t>(dy) & 15) == 0);
  if (vec_ok) {
#pragma unroll 8
    for (int r = r0 + warp; r < r1; r += CS_WARPS) {
      const uint4 u = __ldg(reinterpret_cast<const uint4*>(dy + size_t(r) * ld + col));
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = unpack_bf16x2(w[j]);
        acc[2 * j] += f.x;
        acc[2 * j + 1] += f.y;
      }
    }
  } else if (col < N) {
    for (int r = r0 + warp; r < r1; r += CS_WARPS) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (col + j < N) acc[j] += __bfloat162float(dy[size_t(r) * ld + col + j]);
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) sh[warp][lane * 8 + j] = acc[j];
  __syncthreads();
  const int c = threadIdx.x;  // 256 threads <-> 256 columns
  if (blockIdx.x * CS_COLS + c < N) {
    float t = sh[0][c];
#pragma unroll
    for (int w = 1; w < CS_WARPS; ++w) t += sh[w][c];
    partial[size_t(blockIdx.y) * N + blockIdx.x * CS_COLS + c] = t;
  }
}
// out[n] += sum_s partial[s][n]: block = 32 columns x 8 split-lanes, fixed-order smem combine.
__global__ void __launch_bounds__(256)
colsum_finish_kernel(const float* __restrict__ partial, int splits, int N,
                     float* __restrict__ out) {
  pdl_enter();  // previous kernel finished, its writes are visible
  __shared__ float sh[8][32];
  const int cx = threadIdx.x & 31, sy = threadIdx.x >> 5;
  const int n = blockIdx.x * 32 + cx;
  float acc = 0.f;
  if (n < N) {
    for (int s = sy; s < splits; s += 8) acc += partial[size_t(s) * N + n];
  }
  sh[sy][cx] = acc;
  __syncthreads();
  if (sy == 0 && n < N) {
    float t = sh[0][cx];
#pragma unroll
    for (int k = 1; k < 8; ++k) t += sh[k][cx];
    out[n] += t;
  }
}
inline int colsum_splits(int M, int N) {
  const int col_blocks = ceil_div(N, CS_COLS);
  int s = ceil_div(6 * sm_count(), col_blocks);
  const int max_s = ceil_div(M, 64);
  if (s > max_s) s = max_s;
  if (s > 128) s = 128;
  return s < 1 ? 1 : s;
}

// ------------------------------------------------------------------------------------------------
// fp32 -> bf16 cast (weight shadow refresh), 8 elements per thread.
// ------------------------------------------------------------------------------------------------
__global__ void cast_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                            size_t n) {
  pdl_enter();  // previous kernel finished, its writes are visible
  const size_t i = (size_t(blockIdx.x) * blockDim.x + threadIdx.x) * 8;
  if (i + 8 <= n) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(src + i));
    const float4 b = __ldg(reinterpret_cast<const float4*>(src + i) + 1);
    *reinterpret_cast<uint4*>(dst + i) = make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w),
                                                    pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
  } else {
    for (size_t j = i; j < n; ++j) dst[j] = __float2bfloat16(src[j]);
  }
}

// ------------------------------------------------------------------------------------------------
// Device-resident synthetic token source (work item W13): counter-based generator, one thread per
// row, ids[b,t] = h(seed, row, t) mod vocab and labels[b,t] = ids[b,t+1] (next-token targets),
// so benchmark inputs never have to cross PCIe.  splitmix64 finaliser as the hash.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t mix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__global__ void synth_tokens_kernel(int64_t* __restrict__ ids, int64_t* __restrict__ labels,
                                    int rows, int T, int vocab, uint64_t seed, uint64_t row0) {
  pdl_enter();  // previous kernel finished, its writes are visible
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)rows * T) return;
  const int r = int(idx / T), t = int(idx - (long long)r * T);
  const uint64_t base = mix64(seed ^ mix64(row0 + uint64_t(r)));
  const int64_t cur = int64_t(mix64(base + uint64_t(t)) % uint64_t(vocab));
  const int64_t nxt = int64_t(mix64(base + uint64_t(t) + 1) % uint64_t(vocab));
  ids[idx] = cur;
  if (labels != nullptr) labels[idx] = nxt;
}

}  // namespace ew
}  // namespace b200

using namespace b200;

extern "C" int llmtrain_b200_synthetic_tokens(int64_t* ids, int64_t* labels, int rows, int T,
                                              int vocab, uint64_t seed, uint64_t first_row,
                                              cudaStream_t stream) {
  B200_REQUIRE(vocab > 0 && rows >= 0 && T > 0, "synthetic_tokens: bad shape");
  if (rows == 0) return 0;
  const long long n = (long long)rows * T;
  B200_LAUNCH((ew::synth_tokens_kernel), static_cast<unsigned>((n + 255) / 256), 256, 0, stream,
              ids, labels, rows, T, vocab, seed, first_row);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}

extern "C" int llmtrain_b200_embed_fwd(const int64_t* ids, const void* wte, const void* wpe,
                                       void* x, int B, int T, int d, int vocab,
                                       cudaStream_t stream) {
  B200_REQUIRE(d % 8 == 0, "embed: d must be a multiple of 8");
  const int rows = B * T;
  if (rows == 0) return 0;
  B200_LAUNCH((ew::embed_fwd_kernel), ceil_div(rows, 8), 256, 0, stream, ids,
              static_cast<const __nv_bfloat16*>(wte), static_cast<const __nv_bfloat16*>(wpe),
              static_cast<__nv_bfloat16*>(x), rows, T, d, vocab);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}

extern "C" int llmtrain_b200_embed_bwd(const int64_t* ids, const void* dx, float* dwte,
                                       float* dwpe, int B, int T, int d, int vocab,
                                       cudaStream_t stream) {
  B200_REQUIRE(d % 8 == 0, "embed: d must be a multiple of 8");
  const int rows = B * T;
  if (rows == 0) return 0;
  B200_LAUNCH((ew::embed_bwd_wte_kernel), ceil_div(rows, 8), 256, 0, stream, ids,
              static_cast<const __nv_bfloat16*>(dx), dwte, rows, d, vocab);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  B200_LAUNCH((ew::embed_bwd_wpe_kernel), ceil_div(T * (d / 4), 256), 256, 0, stream,
              static_cast<const __nv_bfloat16*>(dx), dwpe, B, T, d);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}

// Deterministic-order variant (see embed_bwd_wte_ordered_kernel): scratch holds vocab ints.
extern "C" size_t llmtrain_b200_embed_bwd_ordered_scratch_bytes(int vocab) {
  return size_t(vocab) * sizeof(int);
}
extern "C" int llmtrain_b200_embed_bwd_ordered(const int64_t* ids, const void* dx, float* dwte,
                                               float* dwpe, int B, int T, int d, int vocab,
                                               void* scratch, size_t scratch_bytes,
                                               cudaStream_t stream) {
  B200_REQUIRE(d % 8 == 0 && d <= ew::EMB_MAX_V4 * 128, "embed: d must be a multiple of 8, <= 2048");
  B200_REQUIRE(scratch_bytes >= size_t(vocab) * sizeof(int), "embed_bwd_ordered: scratch too small");
  const int rows = B * T;
  if (rows == 0) return 0;
  int* first = static_cast<int*>(scratch);
  B200_CHECK_CUDA(cudaMemsetAsync(first, 0x7f, size_t(vocab) * sizeof(int), stream));
  B200_LAUNCH((ew::embed_first_kernel), ceil_div(rows, 256), 256, 0, stream, ids, first, rows, vocab);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_LAUNCH((ew::embed_bwd_wte_ordered_kernel), ceil_div(rows, 8), 256, 0, stream, ids,
              static_cast<const __nv_bfloat16*>(dx), static_cast<const int*>(first), dwte, rows, d,
              vocab);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_LAUNCH((ew::embed_bwd_wpe_kernel), ceil_div(T * (d / 4), 256), 256, 0, stream,
              static_cast<const __nv_bfloat16*>(dx), dwpe, B, T, d);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(3);
  return 0;
}

extern "C" int llmtrain_b200_cross_entropy(void* logits, const int64_t* labels,
                                           const uint8_t* row_valid, float* loss_rows,
                                           float* loss_acc, int rows, int V, int ldv,
                                           float loss_scale, float grad_scale, int write_grad,
                                           cudaStream_t stream) {
  B200_REQUIRE(ldv % 8 == 0 && ldv >= V, "cross_entropy: row pitch must be >= V and a multiple of 8");
  if (rows == 0) return 0;
  auto* lg = static_cast<__nv_bfloat16*>(logits);
  static const bool row_path_off = std::getenv("LLMTRAIN_CE_TWO_PASS") != nullptr;
  if (!row_path_off && ldv <= ew::CE_ROW_MAX_LDV && (reinterpret_cast<uintptr_t>(logits) & 15) == 0) {
    const size_t smem = 2 * ((size_t(ldv) * 2 + 127) & ~size_t(127));
    static bool configured = false;
    if (!configured) {
      B200_CHECK_CUDA(cudaFuncSetAttribute(ew::ce_row_kernel<true>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, 225280));
      B200_CHECK_CUDA(cudaFuncSetAttribute(ew::ce_row_kernel<false>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, 225280));
      configured = true;
    }
    const int grid = rows < sm_count() ? rows : sm_count();
    if (write_grad)
      B200_LAUNCH((ew::ce_row_kernel<true>), grid, ew::CE_ROW_THREADS, smem, stream, lg, labels,
                  row_valid, loss_rows, rows, V, ldv, grad_scale);
    else
      B200_LAUNCH((ew::ce_row_kernel<false>), grid, ew::CE_ROW_THREADS, smem, stream, lg, labels,
                  row_valid, loss_rows, rows, V, ldv, grad_scale);
  } else if (write_grad) {
    B200_LAUNCH((ew::ce_kernel<true>), rows, ew::CE_THREADS, 0, stream, lg, labels, row_valid,
                loss_rows, V, ldv, grad_scale);
  } else {
    B200_LAUNCH((ew::ce_kernel<false>), rows, ew::CE_THREADS, 0, stream, lg, labels, row_valid,
                loss_rows, V, ldv, grad_scale);
  }
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  if (loss_acc != nullptr) {
    B200_LAUNCH((ew::loss_reduce_kernel), 1, 1024, 0, stream, loss_rows, rows, loss_scale,
                loss_acc);
    B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  }
  return 0;
}

extern "C" size_t llmtrain_b200_bias_grad_scratch_bytes(int M, int N) {
  return size_t(ew::colsum_splits(M, N)) * N * sizeof(float);
}

extern "C" int llmtrain_b200_bias_grad(const void* dy, int ld, int M, int N, float* dbias,
                                       void* scratch, size_t scratch_bytes, cudaStream_t stream) {
  B200_REQUIRE(ld % 2 == 0, "bias_grad: row pitch must be even");
  if (M == 0 || N == 0) return 0;
  const int splits = ew::colsum_splits(M, N);
  B200_REQUIRE(scratch_bytes >= size_t(splits) * N * sizeof(float), "bias_grad: scratch too small");
  dim3 grid(ceil_div(N, ew::CS_COLS), splits);
  B200_LAUNCH((ew::colsum_partial_kernel), grid, ew::CS_WARPS * 32, 0, stream,
              static_cast<const __nv_bfloat16*>(dy), ld, M, N, static_cast<float*>(scratch));
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  B200_LAUNCH((ew::colsum_finish_kernel), ceil_div(N, 32), 256, 0, stream,
              static_cast<float*>(scratch), splits, N, dbias);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}

extern "C" int llmtrain_b200_cast_f32_to_bf16(const float* src, void* dst, size_t n,
                                              cudaStream_t stream) {
  if (n == 0) return 0;
  const size_t threads = (n + 7) / 8;
  B200_LAUNCH((ew::cast_kernel), static_cast<unsigned>((threads + 255) / 256), 256, 0, stream, src,
              static_cast<__nv_bfloat16*>(dst), n);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}

// ------------------------------------------------------------------------------------------------
// Bias gradient for hosts that must not allocate.
//
// Some programs that link this library are not allowed to obtain device memory on their own once
// they are running: a serving or training process whose memory pool already owns every byte of
// HBM (a later cudaMalloc would fail or fragment the pool), a latency-critical loop that cannot
// afford the device-wide stall a cudaMalloc/cudaFree implies, or a thin language binding that can
// pass pointers it was given but has no allocator to ask.  llmtrain_b200_bias_grad needs a few
// hundred KB of temporary global memory for its deterministic two-stage column sum, so those
// callers could not use it.  This variant carries that temporary storage inside the library image
// instead: a fixed array in device global memory that belongs to the library's module on every GPU
// that has loaded it, until process exit.  With CUDA's default lazy module loading the module —
// and with it this 1 MiB — is only materialised on a GPU when something in it is first used
// there, which can be this function's first call: the driver may allocate at that moment.
//
// Consequences the caller has to know about:
//   * The array has a fixed capacity (kNoAllocFloats floats = 1 MiB per GPU).  A problem whose
//     partial sums do not fit is refused with an error; nothing is launched.
//   * All calls on one GPU use the same array and it is never cleared between calls.  That is
//     safe here because the first kernel overwrites every partial sum the second one reads, so no
//     element has to start at zero — but two calls whose kernels are in flight at the same time
//     (different streams, or different host threads on different streams) would overwrite each
//     other's partial sums.  Issue calls for one GPU on one stream, or order them with events.
//   * The device address of the array is looked up once per GPU, on the first call made while
//     that GPU is current.  That first call does host-side set-up work and may be the moment the
//     driver loads the module (see above): make it during initialisation — BEFORE the host's own
//     pool claims the rest of the device and before starting any stream capture — or run the
//     process with CUDA_MODULE_LOADING=EAGER so the memory is taken at library load.
// ------------------------------------------------------------------------------------------------
namespace b200 {
namespace ew {

constexpr size_t kNoAllocFloats = size_t(1) << 18;  // 262144 floats = 1 MiB
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) float g_bias_grad_partials[kNoAllocFloats];

// Host-side cache of the array's device address, one slot per GPU.  (In host code the name
// g_bias_grad_partials is only a handle for the runtime, never an address to hand to a kernel.)
static float* g_noalloc_addr[kNoAllocMaxDevices] = {};
static std::mutex g_noalloc_mutex;

static int noalloc_scratch(float** out) {
  int dev = -1;
  B200_CHECK_CUDA(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, "bias_grad_noalloc: device index out of range");
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (g_noalloc_addr[dev] == nullptr) {
    float* addr = nullptr;
    B200_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&addr), g_bias_grad_partials));
    B200_REQUIRE(addr != nullptr, "bias_grad_noalloc: symbol address lookup returned null");
    g_noalloc_addr[dev] = addr;
  }
  *out = g_noalloc_addr[dev];
  return 0;
}

}  // namespace ew
}  // namespace b200

extern "C" size_t llmtrain_b200_bias_grad_noalloc_capacity_bytes(void) {
  return ew::kNoAllocFloats * sizeof(float);
}

// dbias[n] (fp32) += sum_m dy[m,n] (bf16, pitch ld) — same result as llmtrain_b200_bias_grad, but
// the partial sums live in the library's built-in array (see the block comment above).
extern "C" int llmtrain_b200_bias_grad_noalloc(const void* dy, int ld, int M, int N, float* dbias,
                                               cudaStream_t stream) {
  B200_REQUIRE(ld % 2 == 0, "bias_grad_noalloc: row pitch must be even");
  B200_REQUIRE(M >= 0 && N >= 0, "bias_grad_noalloc: negative extent");
  if (M == 0 || N == 0) return 0;
  const int splits = ew::colsum_splits(M, N);
  if (size_t(splits) * size_t(N) > ew::kNoAllocFloats) {
    return fail("bias_grad_noalloc: " + std::to_string(size_t(splits) * size_t(N)) +
                " partial sums needed for M=" + std::to_string(M) + " N=" + std::to_string(N) +
                " but the built-in scratch holds " + std::to_string(ew::kNoAllocFloats) +
                " floats (1 MiB); use llmtrain_b200_bias_grad with caller-provided scratch");
  }
  float* partials = nullptr;
  if (ew::noalloc_scratch(&partials)) return 1;
  dim3 grid(ceil_div(N, ew::CS_COLS), splits);
  B200_LAUNCH((ew::colsum_partial_kernel), grid, ew::CS_WARPS * 32, 0, stream,
              static_cast<const __nv_bfloat16*>(dy), ld, M, N, partials);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  B200_LAUNCH((ew::colsum_finish_kernel), ceil_div(N, 32), 256, 0, stream, partials, splits, N,
              dbias);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}
