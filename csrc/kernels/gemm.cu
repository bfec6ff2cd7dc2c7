// Persistent, warp-specialised bf16 GEMM for sm_100a:  TMA -> smem ring -> tcgen05.mma -> TMEM
// (double-buffered accumulator) -> epilogue warps -> swizzled smem staging -> TMA store.
//
// Replaces every ATen GEMM the reference makes from `src/llmtrain/models/gpt.py` (qkv_proj :27,
// out_proj :29, mlp_fc :94, mlp_proj :96, lm_head :143) and their autograd dgrad / wgrad
// counterparts (SURVEY §2.5 K3,K5,K7,K8,K10,K12).  One kernel family, three operand layouts:
//
//   NT  C[M,N]  = A[M,K]   * B[N,K]^T      forward      (A, B both K-major)
//   NN  C[M,N]  = A[M,K]   * B[K,N]        dgrad        (A K-major, B MN-major)
//   TN  C[M,N] += A[K,M]^T * B[K,N]        wgrad        (A, B both MN-major; split-K; fp32
//                                                        accumulate into the gradient arena via
//                                                        TMA reduce-add)
//
// Fused epilogues: +bias, +bias & exact-erf GELU (writes pre- and post-activation), +bias &
// residual add, x GELU'(pre) for the MLP dgrad, fp32 accumulate.
//
// Tile: 128 x 256 x 64 per CTA, 2 x 256 TMEM columns.  18 warps: 0 = TMA producer, 1 = MMA issuer
// + TMEM owner, 2..17 = epilogue (warp w reads TMEM lanes 32*(w%4)..+31; four warps per lane
// quarter split the 256 columns into 64-column groups and work in 32-column pieces, which keeps a
// thread under 112 registers).  Sixteen epilogue warps because the fused epilogues (GELU, dGELU,
// residual) were latency-bound with eight: two warps per scheduler used under half of its issue
// slots and the GELU GEMM ran 35% slower than the plain one.  Static round-robin tile schedule with M fastest so concurrently running
// CTAs share the weight tile through L2.  Two flavours of the same kernel (template CTAS):
//   CTAS = 1  UMMA 128x256x16, cta_group::1, 4-stage ring of 48 KB, grid = #SMs.
//   CTAS = 2  CTA pair (2-CTA cluster on one TPC): UMMA 256x256x16, cta_group::2.  The pair owns a
//             256 x 256 tile; each CTA stages its own 128 rows of A and HALF of B (128 of the 256
//             columns), so a stage is 32 KB, the ring is 6 deep, and L2 -> smem traffic per FLOP
//             drops by a third.  The leader issues the MMAs for both tensor cores; each CTA's
//             epilogue drains its own TMEM.  This is the default for shapes with >= 2 row tiles.

#include <cstdlib>

#include "../common/host.cuh"
#include "../common/philox.cuh"
#include "../common/ptx.cuh"
#include "../../include/llmtrain_b200.h"

namespace b200 {
namespace gemm {

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 64;  // 64 bf16 = 128 B = one swizzle row
constexpr int UK = 16;
// smem ring depth: 4 x 48 KB everywhere except the GELU epilogue, which trades one stage for a
// second 4 KB staging buffer per epilogue warp (its two outputs then leave through two TMA stores
// in flight instead of one store plus scattered 16-byte global writes).
template <int EPI, int CTAS> struct Cfg {
  static constexpr int kStages = CTAS == 2 ? 6 : 4;
  static constexpr int kEpiBufs = 1;
};
template <int CTAS> struct Cfg<1 /*kBiasGelu*/, CTAS> {
  static constexpr int kStages = CTAS == 2 ? 5 : 3;
  static constexpr int kEpiBufs = 2;
};
constexpr int A_BYTES = BM * BK * 2;  // 16 KB
template <int CTAS> __host__ __device__ constexpr int b_bytes() { return BN / CTAS * BK * 2; }
template <int CTAS> __host__ __device__ constexpr int stage_bytes() { return A_BYTES + b_bytes<CTAS>(); }
constexpr int EPI_WARPS = 16;            // four warps per TMEM lane quarter, each owns 64 columns
constexpr int EPI_BUF_BYTES = 32 * 64;   // 32 rows x 64 B per staging buffer (SWIZZLE_64B boxes)
template <int EPI, int CTAS> __host__ __device__ constexpr int epi_bytes() {
  return EPI_WARPS * Cfg<EPI, CTAS>::kEpiBufs * EPI_BUF_BYTES;
}
template <int EPI, int CTAS> __host__ __device__ constexpr int smem_bytes() {
  return Cfg<EPI, CTAS>::kStages * stage_bytes<CTAS>() + epi_bytes<EPI, CTAS>() + 256 /*barriers*/;
}
constexpr int NUM_THREADS = 32 * (2 + EPI_WARPS);
constexpr int TMEM_COLS = 512;
// No alignment slack: the kernel has no static shared memory, so its dynamic window starts at the
// 1024 B-aligned base of the CTA's allocation.  Staying under 225 KB leaves room for one of the
// communicator's small, smem-free CTAs to share the SM during the overlapped bucket reduction.
// (shared memory per variant: smem_bytes<EPI, CTAS>())

enum Epi : int {
  kBias = 0,      // C = acc (+ bias)
  kBiasGelu = 1,  // x = acc + bias ; C = gelu(x) ; C2 = gelu'(x)   (C2 is what backward needs)
  kBiasResid = 2, // C = acc (+ bias) + R
  kDGelu = 3,     // C = acc * R                 (R holds gelu'(pre-activation) saved by forward)
  kAccumF32 = 4,  // C(fp32) += acc              (TMA reduce-add)
  kBiasDropResid = 5,  // C = dropout(acc + bias) + R   (mask regenerated in backward, philox.cuh)
  kRowDot = 6,    // C = acc; stats[b, h, t] = (lse * log2 e, sum over head h's 64 columns of C o R):
                  // the attention backward's per-row statistics out of out_proj's dgrad
};

struct Params {
  int M, N, K;
  int k_splits;          // >1 only with kAccumF32
  const float* bias;     // [N] fp32 or nullptr
  const __nv_bfloat16* R;  // residual / saved pre-activation, row pitch ldr
  int ldr;
  __nv_bfloat16* C2;       // second output of the GELU epilogue, row pitch ldc2
  int ldc2;
  int group_m;             // rasterisation: row-tiles per band (>= #row-tiles => M fastest)
  DropoutParams drop;      // kBiasDropResid only
  float* colsum;           // optional, kDGelu / kBias: colsum[n] += sum_m C[m, n]  (bias gradient
                           // of the layer that produced this GEMM's output gradient)
  const uint8_t* row_keep; // optional, kBiasResid / kBiasDropResid: rows with row_keep[m] == 0 get
                           // C = R (the branch is dropped, `gpt.py:73-74`: padded query rows)
  // kRowDot: rows are (b, t) with t < rd_T, columns are (head, 64); R is the attention output O,
  // C the gradient dO.  stats[(b * rd_H + h) * rd_tstat + t] = (lse2, D) with
  // D = sum_c bf16(dO)[b,t,h,c] * O[b,t,h,c] and lse2 = lse[b,h,t] * log2(e) (+inf for an empty row).
  const float* lse;
  float2* stats;
  int rd_T, rd_H, rd_tstat;
};

// Tile order.  Consecutive tiles walk `group_m` row-tiles for one column-tile, then the next
// column-tile, ...; after all column-tiles the next band of row-tiles starts.  A wave of resident
// CTAs therefore touches ~group_m A panels and ~wave/group_m B panels.  With group_m >= num_m this
// is plain M-fastest order (every CTA of a wave shares one weight tile), which is best while the
// whole A operand fits in L2; for larger A the band keeps HBM traffic from scaling with #tiles.
__device__ __forceinline__ void tile_coords(int mn, int num_m, int num_n, int group_m, int& mb,
                                            int& nb) {
  const int band = mn / (group_m * num_n);
  const int first = band * group_m;
  const int rows = min(group_m, num_m - first);
  const int idx = mn - band * group_m * num_n;
  mb = first + idx % rows;
  nb = idx / rows;
}

// (An epilogue that stored 16-byte vectors straight to global memory instead of staging + TMA
// store was measured slower — fc forward 70 -> 87 us — and removed; see profiles/README.md.)

// gelu(x) = x Phi(x) and gelu'(x) = Phi(x) + x phi(x) from ONE evaluation: forward stores the
// derivative (bf16) instead of the pre-activation, so the MLP dgrad epilogue is a single multiply.
// The GELU epilogue must keep up with a 128x256x768 mainloop (~6K cycles per tile) and is
// issue-bound, so everything is evaluated on packed fp32 pairs (f32x2, ptx.cuh).  libm's erff, and
// then an Abramowitz & Stegun 7.1.26 form (reciprocal + exponential + select per element), were
// the pacing stage of these GEMMs before.
// Phi(x) - 1/2 here is an odd degree-15 polynomial in x on [-4, 4] (least-squares fit on Chebyshev
// nodes, max abs error 5e-5 including the clamp — 40x below bf16 resolution), evaluated as
// x * Q(x^2) with packed FMAs: against the A&S form above it needs no reciprocal (one MUFU per
// element instead of two — the exponential for phi(x) remains) and no sign select.
__device__ __forceinline__ void gelu_erf_both2(float x0, float x1, f32x2& y, f32x2& dy) {
  constexpr float kC = -0.5f * 1.4426950408889634f;  // exp(-x^2/2) = exp2(x^2 * kC)
  const f32x2 x = pack2(x0, x1);
  const f32x2 xc = pack2(fminf(fmaxf(x0, -4.0f), 4.0f), fminf(fmaxf(x1, -4.0f), 4.0f));
  const f32x2 u = mul2(xc, xc);
  f32x2 q = fma2(splat2(-1.5807526398958771e-09f), u, splat2(1.2182773900804023e-07f));
  q = fma2(q, u, splat2(-4.10726079280721e-06f));
  q = fma2(q, u, splat2(8.080291445367038e-05f));
  q = fma2(q, u, splat2(-0.0010496044997125864f));
  q = fma2(q, u, splat2(0.009672095067799091f));
  q = fma2(q, u, splat2(-0.066192127764225f));
  q = fma2(q, u, splat2(0.3988598585128784f));
  const f32x2 cdf = fma2(q, xc, splat2(0.5f));
  float a0, a1;
  unpack2(mul2(mul2(x, x), splat2(kC)), a0, a1);
  const f32x2 e = pack2(fast_exp2(a0), fast_exp2(a1));  // exp(-x^2/2)
  y = mul2(x, cdf);
  dy = fma2(mul2(x, splat2(0.3989422804014327f)), e, cdf);  // Phi + x phi,  phi = e / sqrt(2 pi)
}

template <bool A_MN, bool B_MN, int EPI, int CTAS>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmC2,
            const Params p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  constexpr int STAGES = Cfg<EPI, CTAS>::kStages;
  constexpr int EPI_BUFS = Cfg<EPI, CTAS>::kEpiBufs;
  constexpr int STAGE_BYTES = stage_bytes<CTAS>();
  constexpr int BN_LOCAL = BN / CTAS;  // B columns staged by this CTA
  uint8_t* epi_smem = smem + STAGES * STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_smem + epi_bytes<EPI, CTAS>());
  uint64_t* full_bar = bars;                  // [STAGES]
  uint64_t* empty_bar = bars + STAGES;        // [STAGES]
  uint64_t* tfull_bar = bars + 2 * STAGES;    // [2]
  uint64_t* tempty_bar = bars + 2 * STAGES + 2;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // With CTAS = 2 a "tile" is the pair's 256 x 256 block; this CTA owns rows [rank*128, +128).
  const uint32_t rank = CTAS == 2 ? cluster_ctarank() : 0u;
  const bool leader = rank == 0;
  const int first_tile = blockIdx.x / CTAS;
  const int tile_stride = gridDim.x / CTAS;
  const int num_m = (p.M + BM * CTAS - 1) / (BM * CTAS);
  const int num_n = (p.N + BN - 1) / BN;
  const int tiles_mn = num_m * num_n;
  const int num_tiles = tiles_mn * p.k_splits;
  const int kb_total = (p.K + BK - 1) / BK;
  const int kb_per_split = (kb_total + p.k_splits - 1) / p.k_splits;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], EPI_WARPS * CTAS);  // the leader hears from both CTAs' epilogues
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    if (CTAS == 2) { tmem_alloc_pair(tmem_slot, TMEM_COLS); tmem_relinquish_pair(); }
    else { tmem_alloc(tmem_slot, TMEM_COLS); tmem_relinquish(); }
  }
  tc_fence_before_sync();
  if (CTAS == 2) cluster_sync_all();  // the peer's barriers exist before anything signals them
  else __syncthreads();
  tc_fence_after_sync();
  pdl_enter();  // barrier init + TMEM allocation above overlapped the previous kernel's tail
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================================== TMA producer =====================================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = first_tile; tile < num_tiles; tile += tile_stride) {
        const int split = tile / tiles_mn;
        const int mn = tile - split * tiles_mn;
        int mb, nb;
        tile_coords(mn, num_m, num_n, p.group_m, mb, nb);
        const int m0 = mb * (BM * CTAS) + int(rank) * BM;
        const int n0 = nb * BN + int(rank) * BN_LOCAL;  // this CTA's share of B
        const int kb_begin = split * kb_per_split;
        const int kb_end = min(kb_total, kb_begin + kb_per_split);
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * STAGE_BYTES;
          uint8_t* sb = sa + A_BYTES;
          // Pair: both CTAs' loads complete on the LEADER's barrier, which expects both shares.
          if (leader) mbar_expect_tx(&full_bar[stage], STAGE_BYTES * CTAS);
          const int k0 = kb * BK;
          auto load = [&](void* dst, const CUtensorMap* tm, int c0, int c1) {
            if (CTAS == 2) tma_load_2d_pair(dst, tm, &full_bar[stage], c0, c1);
            else tma_load_2d(dst, tm, &full_bar[stage], c0, c1);
          };
          if (!A_MN) {
            load(sa, &tmA, k0, m0);
          } else {
#pragma unroll
            for (int i = 0; i < BM / 64; ++i) load(sa + i * (BK * 128), &tmA, m0 + i * 64, k0);
          }
          if (!B_MN) {
            load(sb, &tmB, k0, n0);
          } else {
#pragma unroll
            for (int i = 0; i < BN_LOCAL / 64; ++i)
              load(sb + i * (BK * 128), &tmB, n0 + i * 64, k0);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================== MMA issuer =======================================
    // The whole warp of the leader CTA walks the loop and waits on the barriers; one elected
    // lane issues.  A stage's descriptors are built once, outside the elected region, and
    // advanced by one 64-bit add per instruction: inside `if (lane == 0)` (a divergent region,
    // so no uniform registers) every tcgen05.mma was preceded by ~20 dependent instructions,
    // ~100 cycles of issue against the 128 cycles the instruction takes to execute — no slack
    // for the barrier wait and the commit of a 4-instruction stage.
    if (leader) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM * CTAS, BN, A_MN, B_MN);
      // K-major: 32 B per UMMA_K inside the swizzled 128 B row (+2 in the address field);
      // SBO = 8 rows * 128 B.  MN-major: 16 K-rows * 128 B per UMMA_K (+128); LBO = one 64-wide
      // box (BK rows * 128 B).
      constexpr uint32_t a_step = A_MN ? (UK * 128) >> 4 : (UK * 2) >> 4;
      constexpr uint32_t b_step = B_MN ? (UK * 128) >> 4 : (UK * 2) >> 4;
      const bool issuer = elect_one();
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = first_tile; tile < num_tiles; tile += tile_stride) {
        const int split = tile / tiles_mn;
        const int kb_begin = split * kb_per_split;
        const int kb_end = min(kb_total, kb_begin + kb_per_split);
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after_sync();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after_sync();
          const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
          const uint32_t sb = sa + A_BYTES;
          const uint64_t da = A_MN ? umma_smem_desc(sa, BK * 128, 1024) : umma_smem_desc(sa, 0, 1024);
          const uint64_t db = B_MN ? umma_smem_desc(sb, BK * 128, 1024) : umma_smem_desc(sb, 0, 1024);
          if (issuer) {
#pragma unroll
            for (int k = 0; k < BK / UK; ++k) {
              if (CTAS == 2)
                umma_ss_pair(d_tmem, da + a_step * k, db + b_step * k, idesc, (kb > kb_begin) || (k > 0));
              else
                umma_ss(d_tmem, da + a_step * k, db + b_step * k, idesc, (kb > kb_begin) || (k > 0));
            }
            // frees this smem stage (in both CTAs of a pair) once the MMAs have read it
            if (CTAS == 2) umma_commit_pair(&empty_bar[stage], 0b11);
            else umma_commit(&empty_bar[stage]);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        // accumulator complete -> epilogue (of both CTAs)
        if (issuer) {
          if (CTAS == 2) umma_commit_pair(&tfull_bar[acc], 0b11);
          else umma_commit(&tfull_bar[acc]);
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ===================================== epilogue =========================================
    // Warp w reads TMEM lanes 32*(w%4)..+31 (hardware restriction) and the 64 accumulator columns
    // [64*colgrp, +64), in two pieces of 32 columns; one 2 KB staging buffer per warp and output.
    const int ew = warp - 2;        // 0..15
    const int quarter = warp & 3;   // TMEM lane quarter this warp may access
    const int colgrp = ew >> 2;     // which 64-column group of the tile
    uint8_t* stg = epi_smem + ew * EPI_BUFS * EPI_BUF_BYTES;
    uint8_t* stg2 = stg + (EPI_BUFS - 1) * EPI_BUF_BYTES;  // == stg unless the variant has two
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = first_tile; tile < num_tiles; tile += tile_stride) {
      const int split = tile / tiles_mn;
      const int mn = tile - split * tiles_mn;
      int mb, nb;
      tile_coords(mn, num_m, num_n, p.group_m, mb, nb);
      const int m0 = mb * (BM * CTAS) + int(rank) * BM;
      const int n0 = nb * BN + colgrp * 64;
      const int row = m0 + quarter * 32 + lane;  // this thread's output row
      const bool row_ok = row < p.M;
      // Residual / saved gelu' for this thread's row (64 bf16): issued BEFORE waiting for the
      // accumulator so the HBM latency hides behind this tile's mainloop.
      uint4 rv[8];
      if (EPI == kBiasResid || EPI == kDGelu || EPI == kBiasDropResid || EPI == kRowDot) {
#pragma unroll
        for (int q = 0; q < 8; ++q) rv[q] = make_uint4(0u, 0u, 0u, 0u);
        if (row_ok && n0 < p.N) {
          const __nv_bfloat16* rp = p.R + size_t(row) * p.ldr + n0;
          if (n0 + 64 <= p.N) {
#pragma unroll
            for (int q = 0; q < 8; ++q) rv[q] = __ldg(reinterpret_cast<const uint4*>(rp) + q);
          } else {
            uint32_t* rw = reinterpret_cast<uint32_t*>(rv);
#pragma unroll
            for (int j = 0; j < 64; j += 2) {  // static indices keep rv in registers
              const float lo = (n0 + j < p.N) ? __bfloat162float(rp[j]) : 0.f;
              const float hi = (n0 + j + 1 < p.N) ? __bfloat162float(rp[j + 1]) : 0.f;
              rw[j >> 1] = pack_bf16x2(lo, hi);
            }
          }
        }
      }
      // Bias for this warp's 64 columns: lane l keeps columns 2l, 2l+1 (one coalesced 8-byte load,
      // issued before the accumulator wait); values reach their consumers by shuffle, so no
      // memory latency sits inside the math loop.
      float2 bias2 = make_float2(0.f, 0.f);
      const bool add_bias =
          (EPI == kBias || EPI == kBiasGelu || EPI == kBiasResid || EPI == kBiasDropResid) &&
          p.bias != nullptr;
      if (add_bias) {
        const int bc = n0 + lane * 2;
        if (bc + 2 <= p.N) {
          bias2 = __ldg(reinterpret_cast<const float2*>(p.bias + bc));
        } else if (bc < p.N) {
          bias2.x = __ldg(p.bias + bc);
        }
      }
      bool keep_row = true;
      if ((EPI == kBiasResid || EPI == kBiasDropResid) && p.row_keep != nullptr && row_ok)
        keep_row = p.row_keep[row] != 0;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after_sync();
      const uint32_t t_row = tmem_base + (uint32_t(quarter * 32) << 16) + acc * BN + colgrp * 64;

      if (EPI == kAccumF32) {
        // 4 pieces of 16 fp32 columns (64 B rows), reduce-added into global fp32.
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
          if (n0 + c * 16 >= p.N) break;
          uint32_t v[16];
          tmem_ld_32x32b_x16(t_row + c * 16, v);
          tmem_ld_wait();
          tma_store_wait_read<0>();  // previous piece's TMA has finished reading the buffer
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 4; ++j)
            *reinterpret_cast<uint4*>(stg + swizzle64(lane * 64 + j * 16)) =
                make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_reduce_add_2d(&tmC, stg, n0 + c * 16, m0 + quarter * 32);
            tma_store_commit();
          }
        }
      } else {
        // 2 pieces of 32 bf16 columns (64 B rows).
        const uint32_t* rw = reinterpret_cast<const uint32_t*>(rv);
        float row_dot = 0.f;  // kRowDot
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int col0 = n0 + h * 32;
          if (col0 >= p.N) break;
          uint32_t v[32];
          tmem_ld_32x32b_x32(t_row + h * 32, v);
          tmem_ld_wait();

          uint32_t out[16];   // 32 bf16 packed
          uint32_t out2[16];  // gelu'(x) (GELU variant only)
          Rand8 rnd[4];
          if (EPI == kBiasDropResid) {
            // 32 consecutive columns of this thread's row = four 8-element Philox groups
            const uint64_t g0 = (uint64_t(row) * uint64_t(p.N) + uint64_t(col0)) >> 3;
#pragma unroll
            for (int g = 0; g < 4; ++g) rnd[g] = philox_group(p.drop, g0 + g);
          }
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            float x0 = __uint_as_float(v[j]);
            float x1 = __uint_as_float(v[j + 1]);
            if (add_bias) {
              // columns h*32 + j, +1 of this warp's 64 live in lane (h*32 + j) / 2
              x0 += __shfl_sync(0xffffffffu, bias2.x, (h * 32 + j) >> 1);
              x1 += __shfl_sync(0xffffffffu, bias2.y, (h * 32 + j) >> 1);
            }
            if (EPI == kBiasDropResid) {
              x0 *= keep_factor(p.drop, rnd[j >> 3], j & 7);
              x1 *= keep_factor(p.drop, rnd[j >> 3], (j & 7) + 1);
            }
            if (EPI == kBiasResid || EPI == kDGelu || EPI == kBiasDropResid) {
              const float2 f = unpack_bf16x2(rw[h * 16 + (j >> 1)]);
              if (EPI == kDGelu) { x0 *= f.x; x1 *= f.y; }
              else if (keep_row) { x0 += f.x; x1 += f.y; }
              else { x0 = f.x; x1 = f.y; }
            }
            if (EPI == kBiasGelu) {
              f32x2 y2, dy2;
              gelu_erf_both2(x0, x1, y2, dy2);
              float d0, d1;
              unpack2(dy2, d0, d1);
              out2[j / 2] = pack_bf16x2(d0, d1);
              unpack2(y2, x0, x1);
            }
            out[j / 2] = pack_bf16x2(x0, x1);
            if (EPI == kRowDot) {  // on the ROUNDED gradient: the same values the attention reads
              const float2 c = unpack_bf16x2(out[j / 2]);
              const float2 f = unpack_bf16x2(rw[h * 16 + (j >> 1)]);
              row_dot = fmaf(c.x, f.x, fmaf(c.y, f.y, row_dot));
            }
          }

          tma_store_wait_read<0>();  // previous piece's TMA stores have finished reading the buffers
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 4; ++j)
            *reinterpret_cast<uint4*>(stg + swizzle64(lane * 64 + j * 16)) =
                make_uint4(out[4 * j], out[4 * j + 1], out[4 * j + 2], out[4 * j + 3]);
          if (EPI == kBiasGelu) {
            // Second output, gelu'(x), through its own staging buffer: both TMA stores of this
            // piece are in flight together.
#pragma unroll
            for (int j = 0; j < 4; ++j)
              *reinterpret_cast<uint4*>(stg2 + swizzle64(lane * 64 + j * 16)) =
                  make_uint4(out2[4 * j], out2[4 * j + 1], out2[4 * j + 2], out2[4 * j + 3]);
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmC, stg, col0, m0 + quarter * 32);
            if (EPI == kBiasGelu) tma_store_2d(&tmC2, stg2, col0, m0 + quarter * 32);
            tma_store_commit();
          }
          if ((EPI == kDGelu || EPI == kBias) && p.colsum != nullptr) {
            // Column sums of the bf16 piece just staged (= the bias gradient of the layer whose
            // output gradient this GEMM produces), read back from shared memory and added to
            // global memory.  Rows
            // past M hold exact zeros (TMA zero-fills A), so they contribute nothing.  This
            // replaces a separate pass that re-read the whole [M, N] tensor from HBM.
            // lanes 0-15 take rows 0-15, lanes 16-31 rows 16-31 of column pair (lane & 15); one
            // 4-byte load per row, the two halves meet through a shuffle.
            float s0 = 0.f, s1 = 0.f;
            const uint32_t pr = uint32_t(lane & 15);        // column pair inside the piece
            const uint32_t r0 = uint32_t(lane >> 4) * 16u;  // first row of this lane's half
#pragma unroll
            for (int rr = 0; rr < 16; ++rr) {
              const uint32_t rowi = r0 + uint32_t(rr);
              const uint32_t off = rowi * 64u + (((pr >> 2) ^ ((rowi >> 1) & 3u)) << 4) +
                                   ((pr & 3u) << 2);
              const float2 f = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(stg + off));
              s0 += f.x;
              s1 += f.y;
            }
            s0 += __shfl_xor_sync(0xffffffffu, s0, 16);
            s1 += __shfl_xor_sync(0xffffffffu, s1, 16);
            // (A 128-byte bulk reduction, cp.reduce.async.bulk.add.f32, per piece was measured
            // slower than these atomics: 84.5 vs 80.6 us on the dGELU GEMM.)
            if (lane < 16) {
              const int cc = col0 + 2 * lane;
              if (cc < p.N) atomicAdd(p.colsum + cc, s0);
              if (cc + 1 < p.N) atomicAdd(p.colsum + cc + 1, s1);
            }
          }
        }
        if (EPI == kRowDot && row_ok && n0 < p.N) {
          // this warp's 64 columns are exactly one head: the thread owns D of (row, head)
          const int b = row / p.rd_T, t = row - b * p.rd_T, head = n0 >> 6;
          const size_t bh = size_t(b) * p.rd_H + head;
          const float l = __ldg(p.lse + bh * p.rd_T + t);
          p.stats[bh * p.rd_tstat + t] =
              make_float2(l == -INFINITY ? INFINITY : l * 1.4426950408889634f, row_dot);
        }
      }
      // TMEM reads for this accumulator are complete (wait::ld above): hand it back.
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) {
        if (CTAS == 2) mbar_arrive_cluster(&tempty_bar[acc], 0);  // the leader's MMA thread waits
        else mbar_arrive(&tempty_bar[acc]);
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    tma_store_wait<0>();  // global writes landed before the kernel exits
  }

  tc_fence_before_sync();
  if (CTAS == 2) {
    // Neither CTA may leave (or free TMEM) while the other can still signal its barriers or the
    // leader's MMAs can still read its shared memory.
    cluster_sync_all();
    if (warp == 1) tmem_dealloc_pair(tmem_base, TMEM_COLS);
  } else {
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------
// Host launcher
// ------------------------------------------------------------------------------------------------
template <bool A_MN, bool B_MN, int EPI, int CTAS>
int launch(const CUtensorMap& tA, const CUtensorMap& tB, const CUtensorMap& tC,
           const CUtensorMap& tC2, const Params& p, cudaStream_t stream) {
  auto kern = gemm_kernel<A_MN, B_MN, EPI, CTAS>;
  constexpr int kSmem = smem_bytes<EPI, CTAS>();
  static_assert(kSmem <= 227 * 1024, "gemm: shared memory budget exceeded");
  static bool configured = false;
  if (!configured) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    configured = true;
  }
  const int tiles = ceil_div(p.M, BM * CTAS) * ceil_div(p.N, BN) * p.k_splits;
  const int slots = sm_count() / CTAS;  // CTAs, or CTA pairs, that can be resident
  const int grid = (tiles < slots ? tiles : slots) * CTAS;
  B200_CHECK_CUDA(launch_kernel(kern, dim3(grid), dim3(NUM_THREADS), kSmem, stream, CTAS, tA, tB, tC,
                                tC2, p));
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}

// CTA pairs are the default; LLMTRAIN_GEMM_PAIR=0 selects the single-CTA kernel (A/B runs).
bool use_pair(int M) {
  static const int env = [] {
    const char* e = std::getenv("LLMTRAIN_GEMM_PAIR");
    return e == nullptr ? 1 : std::atoi(e);
  }();
  return env != 0 && M > BM;
}

int pick_splits(int M, int N, int K, int ctas) {
  const int tiles = ceil_div(M, BM * ctas) * ceil_div(N, BN);
  const int kb = ceil_div(K, BK);
  int s = (sm_count() / ctas) / (tiles > 0 ? tiles : 1);
  if (s < 1) s = 1;
  if (s > 16) s = 16;
  while (s > 1 && kb / s < 8) --s;  // keep >= 8 k-blocks per split
  // avoid empty trailing splits
  while (s > 1 && (s - 1) * ceil_div(kb, s) >= kb) --s;
  return s;
}

}  // namespace gemm
}  // namespace b200

using namespace b200;
using namespace b200::gemm;

// layout: 0 = NT, 1 = NN, 2 = TN (see header).  All pointers are device pointers.
static int gemm_dispatch(int layout, int epilogue, int M, int N, int K, const void* A, int lda,
                         const void* B, int ldb, void* C, int ldc, void* C2, int ldc2,
                         const float* bias, const void* R, int ldr, const DropoutParams& drop,
                         float* colsum, cudaStream_t stream, const uint8_t* row_keep = nullptr,
                         const Params* row_dot = nullptr) {
  B200_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem");
  Params p{};
  if (row_dot != nullptr) {
    p.lse = row_dot->lse;
    p.stats = row_dot->stats;
    p.rd_T = row_dot->rd_T;
    p.rd_H = row_dot->rd_H;
    p.rd_tstat = row_dot->rd_tstat;
  }
  B200_REQUIRE((epilogue == kRowDot) == (row_dot != nullptr), "gemm: row statistics go with kRowDot");
  p.drop = drop;
  p.colsum = colsum;
  p.row_keep = row_keep;
  B200_REQUIRE(row_keep == nullptr || epilogue == kBiasResid || epilogue == kBiasDropResid,
               "gemm: a row mask is available with the residual epilogues only");
  B200_REQUIRE(colsum == nullptr || epilogue == kDGelu || epilogue == kBias,
               "gemm: column sums are available with the plain and dGELU epilogues only");
  p.M = M; p.N = N; p.K = K; p.k_splits = 1;
  p.bias = bias;
  p.R = static_cast<const __nv_bfloat16*>(R);
  p.ldr = ldr;
  p.C2 = static_cast<__nv_bfloat16*>(C2);
  p.ldc2 = ldc2;
  const auto* a = static_cast<const __nv_bfloat16*>(A);
  const auto* b = static_cast<const __nv_bfloat16*>(B);
  CUtensorMap tA, tB, tC, tC2;
  std::memset(&tC2, 0, sizeof(tC2));
  const bool a_mn = (layout == 2);
  const bool b_mn = (layout == 1 || layout == 2);
  const bool pair = gemm::use_pair(M);
  {
    // Band the row-tiles once the A operand (M x K bf16) no longer sits comfortably in the 126 MB
    // L2 next to the weights and the output stream.
    const int num_m = ceil_div(M, BM * (pair ? 2 : 1));
    static const int forced = [] {
      const char* e = std::getenv("LLMTRAIN_GEMM_GROUP_M");
      return e == nullptr ? 0 : std::atoi(e);
    }();
    const bool big_a = size_t(M) * size_t(K) * 2 > (size_t(40) << 20);
    int g = forced > 0 ? forced : (big_a ? 8 : num_m);
    // Fused column sums add into ONE row of N floats: with M-fastest order every resident CTA
    // would hit the same few cache lines at once (measured: +22 us on the 72 us dGELU GEMM).
    // Narrow bands put the CTAs of a wave on different column tiles.
    static const int colsum_band = [] {
      const char* e = std::getenv("LLMTRAIN_GEMM_COLSUM_BAND");
      return e == nullptr ? 4 : std::atoi(e);
    }();
    if (colsum != nullptr && forced <= 0 && g > colsum_band) g = colsum_band;
    p.group_m = g < num_m ? g : num_m;
  }
  // K-major operand: memory [rows, K]; box = [tile rows, 64].  MN-major: memory [K, rows];
  // box = [64 k-rows, 64 cols].
  if (!a_mn) { if (make_tmap_2d(&tA, a, M, K, lda, BM, BK, Swz::B128)) return 1; }
  else       { if (make_tmap_2d(&tA, a, K, M, lda, BK, 64, Swz::B128)) return 1; }
  if (!b_mn) { if (make_tmap_2d(&tB, b, N, K, ldb, pair ? BN / 2 : BN, BK, Swz::B128)) return 1; }
  else       { if (make_tmap_2d(&tB, b, K, N, ldb, BK, 64, Swz::B128)) return 1; }

  if (epilogue == kAccumF32) {
    if (make_tmap_2d(&tC, static_cast<float*>(C), M, N, ldc, 32, 16, Swz::B64)) return 1;
    // split-K partials meet in L2 in arrival order; deterministic mode keeps one writer per element
    p.k_splits = deterministic() ? 1 : pick_splits(M, N, K, pair ? 2 : 1);
  } else {
    if (make_tmap_2d(&tC, static_cast<__nv_bfloat16*>(C), M, N, ldc, 32, 32, Swz::B64)) return 1;
    if (epilogue == kBiasGelu) {
      B200_REQUIRE(C2 != nullptr, "gemm: GELU epilogue needs the second output C2");
      if (make_tmap_2d(&tC2, static_cast<__nv_bfloat16*>(C2), M, N, ldc2, 32, 32, Swz::B64))
        return 1;
    }
    if (epilogue == kBiasResid || epilogue == kDGelu || epilogue == kBiasDropResid ||
        epilogue == kRowDot)
      B200_REQUIRE(R != nullptr, "gemm: this epilogue needs R");
    if (epilogue == kBiasDropResid)
      B200_REQUIRE(N % 8 == 0, "gemm: the dropout epilogue needs N to be a multiple of 8");
  }

#define B200_GEMM_CASE(L, AMN, BMN, E)                                                  \
  if (layout == L && epilogue == E)                                                     \
    return pair ? launch<AMN, BMN, E, 2>(tA, tB, tC, tC2, p, stream)                    \
                : launch<AMN, BMN, E, 1>(tA, tB, tC, tC2, p, stream);
  B200_GEMM_CASE(0, false, false, kBias)
  B200_GEMM_CASE(0, false, false, kBiasGelu)
  B200_GEMM_CASE(0, false, false, kBiasResid)
  B200_GEMM_CASE(0, false, false, kBiasDropResid)
  B200_GEMM_CASE(1, false, true, kBias)
  B200_GEMM_CASE(1, false, true, kDGelu)
  B200_GEMM_CASE(1, false, true, kRowDot)
  B200_GEMM_CASE(2, true, true, kAccumF32)
#undef B200_GEMM_CASE
  return fail("gemm: unsupported layout/epilogue combination");
}

extern "C" int llmtrain_b200_gemm(int layout, int epilogue, int M, int N, int K, const void* A,
                                  int lda, const void* B, int ldb, void* C, int ldc, void* C2,
                                  int ldc2, const float* bias, const void* R, int ldr,
                                  cudaStream_t stream) {
  B200_REQUIRE(epilogue != kBiasDropResid, "gemm: use llmtrain_b200_gemm_dropout_resid");
  return gemm_dispatch(layout, epilogue, M, N, K, A, lda, B, ldb, C, ldc, C2, ldc2, bias, R, ldr,
                       make_dropout(0.f, 0, 0, 0), nullptr, stream);
}

// As llmtrain_b200_gemm (epilogue 0 or 3 only) and additionally colsum[n] += sum_m C[m, n] over
// the bf16-rounded output — the bias gradient of the layer this GEMM back-propagates INTO, fused
// so that no separate reduction re-reads C.  fp32 atomics: the summation order is not fixed.
extern "C" int llmtrain_b200_gemm_colsum(int layout, int epilogue, int M, int N, int K,
                                         const void* A, int lda, const void* B, int ldb, void* C,
                                         int ldc, const float* bias, const void* R, int ldr,
                                         float* colsum, cudaStream_t stream) {
  B200_REQUIRE(colsum != nullptr, "gemm_colsum: colsum must not be null");
  return gemm_dispatch(layout, epilogue, M, N, K, A, lda, B, ldb, C, ldc, nullptr, 0, bias, R, ldr,
                       make_dropout(0.f, 0, 0, 0), colsum, stream);
}

// dO[M, H*64] = A[M,K] * B[K, H*64]  (NN layout: out_proj's dgrad) and, from the same epilogue, the
// attention backward's per-row statistics: stats[(b*H + h) * t_stat + t] = (lse[b,h,t] * log2 e,
// sum_c bf16(dO)[b,t,h,c] * O[b,t,h,c]) — rows are (b, t), M = B*T.  Replaces the statistics
// pre-pass of llmtrain_b200_attention_bwd (which re-read dO and O): call
// llmtrain_b200_attention_bwd_prestats afterwards with the same scratch.
extern "C" int llmtrain_b200_gemm_rowdot(int M, int N, int K, const void* A, int lda, const void* B,
                                         int ldb, void* dO, int ldc, const void* O, int ldo,
                                         const float* lse, void* stats, int T, int H, int t_stat,
                                         cudaStream_t stream) {
  B200_REQUIRE(T > 0 && H > 0 && N == H * 64 && M % T == 0, "gemm_rowdot: rows are (b, t), columns (head, 64)");
  B200_REQUIRE(lse != nullptr && stats != nullptr && t_stat >= T, "gemm_rowdot: statistics buffers");
  Params rd{};
  rd.lse = lse;
  rd.stats = static_cast<float2*>(stats);
  rd.rd_T = T;
  rd.rd_H = H;
  rd.rd_tstat = t_stat;
  return gemm_dispatch(1, kRowDot, M, N, K, A, lda, B, ldb, dO, ldc, nullptr, 0, nullptr, O, ldo,
                       make_dropout(0.f, 0, 0, 0), nullptr, stream, nullptr, &rd);
}

// C[M,N] = dropout(A[M,K] * B[N,K]^T + bias) + R   (NT layout; mask element index row * N + col,
// regenerated by llmtrain_b200_dropout with the same (p, seed, offset, site) in backward).
extern "C" int llmtrain_b200_gemm_dropout_resid(int M, int N, int K, const void* A, int lda,
                                                const void* B, int ldb, void* C, int ldc,
                                                const float* bias, const void* R, int ldr,
                                                float p, uint64_t seed, uint64_t offset,
                                                uint32_t site, cudaStream_t stream) {
  B200_REQUIRE(p >= 0.f && p < 1.f, "gemm_dropout_resid: p must be in [0, 1)");
  return gemm_dispatch(0, kBiasDropResid, M, N, K, A, lda, B, ldb, C, ldc, nullptr, 0, bias, R, ldr,
                       make_dropout(p, seed, offset, site), nullptr, stream);
}

// C = (row_keep[m] ? dropout(A B^T + bias) : 0) + R   (NT layout; p = 0: no dropout).  The attention
// branch of the reference multiplies its output by the padding mask before the residual add
// (`src/llmtrain/models/gpt.py:71-74`); this is that branch's out_proj with the multiply fused.
extern "C" int llmtrain_b200_gemm_resid_rowmask(int M, int N, int K, const void* A, int lda,
                                                const void* B, int ldb, void* C, int ldc,
                                                const float* bias, const void* R, int ldr,
                                                const uint8_t* row_keep, float p, uint64_t seed,
                                                uint64_t offset, uint32_t site,
                                                cudaStream_t stream) {
  B200_REQUIRE(p >= 0.f && p < 1.f, "gemm_resid_rowmask: p must be in [0, 1)");
  const bool drop = p > 0.f;
  return gemm_dispatch(0, drop ? kBiasDropResid : kBiasResid, M, N, K, A, lda, B, ldb, C, ldc,
                       nullptr, 0, bias, R, ldr, make_dropout(p, seed, offset, site), nullptr,
                       stream, row_keep);
}
