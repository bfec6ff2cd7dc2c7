// Causal flash attention forward on tcgen05 (head_dim 64, bf16 in/out, fp32 softmax state).
//
// Replaces the reference's materialised attention (`src/llmtrain/models/gpt.py:56-70`: scores
// (B,H,T,T) fp32, masked_fill, softmax, P@V — SURVEY §2.5 K4, work item W2).  Nothing of size T^2
// ever reaches HBM: per (batch, head, 128-query tile) one CTA walks the key/value tiles up to the
// diagonal.
//
//   TMA warp   : Q once, then K_j / V_j (3-D tensor maps over the fused [B, T, 3d] QKV buffer, so
//                no permute/copy is ever made and rows past T read 0); K is reloaded as soon as
//                S_j has consumed it, V as soon as PV_j has.
//   MMA warp   : S = Q K_j^T   (SS, 128x128x64, accumulator in TMEM cols [0,128))
//                PV = P_j V_j  (SS, 128x64x128, two TMEM accumulators of 64 cols), P_j read from a
//                double-buffered bf16 smem tile, V_j consumed as an MN-major operand straight from
//                its [kv, hd] layout.  Issue order: p_full(j) -> S_{j+1} -> PV_j.
//   4 softmax warps: thread i owns query row i (= TMEM lane i): two passes over S (max, then
//                exp2 + row-sum + bf16 pack into the swizzled P tile); PV_{j-1} is folded into the
//                register-resident output AFTER P_j has been handed over, so the tensor-core
//                latency of PV hides behind the next tile's softmax.
//
// Masking follows the reference's semantics: causal (col > row) and optional key padding are
// excluded from the softmax; a query whose keys are ALL excluded yields zeros (the reference
// zeroes such rows after out_proj, `gpt.py:73-74`).  LSE (natural log) is saved for backward.
//
// ~113 KB smem and 256 TMEM columns per CTA => two CTAs per SM, so one CTA's softmax overlaps
// the other's tensor-core work.

#include "../common/host.cuh"
#include "../common/philox.cuh"
#include "../common/ptx.cuh"
#include "../../include/llmtrain_b200.h"

namespace b200 {
namespace attn {

constexpr int BQ = 128;
constexpr int BKV = 128;
constexpr int HD = 64;
constexpr int TILE_BYTES = 128 * HD * 2;  // 16 KB: 128 rows x 128 B
constexpr int FWD_THREADS = 192;
constexpr int FWD_TMEM_COLS = 256;
constexpr int OFF_Q = 0;
constexpr int OFF_K = OFF_Q + TILE_BYTES;          // single stage (reloaded once S_j has read it)
constexpr int OFF_V = OFF_K + TILE_BYTES;          // single stage (reloaded once PV_j has read it)
constexpr int OFF_P = OFF_V + TILE_BYTES;          // 2 x 32 KB: P_j is written while PV_{j-1} runs
constexpr int P_BYTES = 2 * TILE_BYTES;
constexpr int OFF_BAR = OFF_P + 2 * P_BYTES;
constexpr int FWD_SMEM = OFF_BAR + 256;

struct FwdParams {
  int B, T, H;
  float scale_log2;              // head_dim^-0.5 * log2(e)
  const uint8_t* key_mask;       // [B, T] (1 = attend) or nullptr
  float* lse;                    // [B, H, T]
  DropoutParams drop;            // attention-probability dropout (`gpt.py:67`); threshold 0 = off
  int t_pad;                     // T rounded up to 8: row pitch of the mask's element index
};

// Mask modes of one 128x128 score tile:
//   kNoMask   tile strictly below the diagonal, no key padding: no per-element work at all
//   kDiag     the diagonal tile, no key padding.  Rows of warp w are 32w..32w+31 and kv0 == q0, so
//             32-column chunk c is entirely kept for c < w, entirely dropped for c > w, and only
//             chunk c == w needs the per-element test (column i of the chunk kept iff i <= lane).
//             The decision is warp-uniform, dropped chunks cost four zero stores and no MUFU.
//   kGeneral  key padding present: per-element causal + padding test.
// (A first version ran the general test on every diagonal tile: 22% of the tiles took 40% of the
// softmax warps' time.)
enum MaskMode : int { kNoMask = 0, kDiag = 1, kGeneral = 2 };

__device__ __forceinline__ bool keep_general(int col, int qrow, int T, const uint8_t* kmask) {
  bool ok = col <= qrow;
  if (kmask != nullptr && col < T) ok = ok && (kmask[col] != 0);
  return ok;
}

// max over this thread's row of the 128-column S tile
template <int MODE>
__device__ __forceinline__ float row_max(uint32_t t_row, int kv0, int qrow, int T,
                                         const uint8_t* kmask, int wchunk, int lane) {
  float mx = -INFINITY;
#pragma unroll 1
  for (int c = 0; c < BKV / 32; ++c) {
    if (MODE == kDiag && c > wchunk) break;
    uint32_t v[32];
    tmem_ld_32x32b_x32(t_row + c * 32, v);
    tmem_ld_wait();
    if (MODE == kNoMask || (MODE == kDiag && c < wchunk)) {
#pragma unroll
      for (int i = 0; i < 32; ++i) mx = fmaxf(mx, __uint_as_float(v[i]));
    } else if (MODE == kDiag) {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (i <= lane) mx = fmaxf(mx, __uint_as_float(v[i]));
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (keep_general(kv0 + c * 32 + i, qrow, T, kmask)) mx = fmaxf(mx, __uint_as_float(v[i]));
    }
  }
  return mx;
}

// p = exp2(s*scale - m) -> bf16 into the K-major 128B-swizzled P tile; returns the row sum
// DROP: the probabilities handed to the PV product are dropped and rescaled; the row sum (the
// softmax normaliser) is taken BEFORE dropout, as in softmax -> nn.Dropout.  `mask_base` is the
// mask element index of (this row, column kv0).
template <int MODE, bool DROP>
__device__ __forceinline__ float exp_pack(uint32_t t_row, uint8_t* sP, int r, int kv0, int qrow,
                                          int T, const uint8_t* kmask, float scale_log2,
                                          float m_use, int wchunk, int lane,
                                          const DropoutParams& dp, uint64_t mask_base) {
  float sum = 0.f;
#pragma unroll 1
  for (int c = 0; c < BKV / 32; ++c) {
    uint8_t* region = sP + (c >> 1) * TILE_BYTES;
    if (MODE == kDiag && c > wchunk) {  // whole chunk above the diagonal for every row of the warp
#pragma unroll
      for (int q = 0; q < 4; ++q)
        *reinterpret_cast<uint4*>(region + swizzle128(r * 128 + ((c & 1) * 4 + q) * 16)) =
            make_uint4(0u, 0u, 0u, 0u);
      continue;
    }
    uint32_t v[32];
    tmem_ld_32x32b_x32(t_row + c * 32, v);
    tmem_ld_wait();
    uint32_t packed[16];
    const bool plain = MODE == kNoMask || (MODE == kDiag && c < wchunk);
    Rand8 rnd{};
#pragma unroll
    for (int i = 0; i < 32; i += 2) {
      float e0 = fast_exp2(fmaf(__uint_as_float(v[i]), scale_log2, -m_use));
      float e1 = fast_exp2(fmaf(__uint_as_float(v[i + 1]), scale_log2, -m_use));
      if (!plain) {
        if (MODE == kDiag) {
          if (i > lane) e0 = 0.f;
          if (i + 1 > lane) e1 = 0.f;
        } else {
          if (!keep_general(kv0 + c * 32 + i, qrow, T, kmask)) e0 = 0.f;
          if (!keep_general(kv0 + c * 32 + i + 1, qrow, T, kmask)) e1 = 0.f;
        }
      }
      sum += e0 + e1;
      if (DROP) {
        if ((i & 7) == 0) rnd = philox_group(dp, (mask_base + uint64_t(c * 32 + i)) >> 3);
        e0 *= keep_factor(dp, rnd, i & 7);
        e1 *= keep_factor(dp, rnd, (i & 7) + 1);
      }
      packed[i >> 1] = pack_bf16x2(e0, e1);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t off = swizzle128(r * 128 + ((c & 1) * 4 + q) * 16);
      *reinterpret_cast<uint4*>(region + off) =
          make_uint4(packed[4 * q], packed[4 * q + 1], packed[4 * q + 2], packed[4 * q + 3]);
    }
  }
  return sum;
}

template <bool DROP>
__global__ void __launch_bounds__(FWD_THREADS, 2)
fwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmO,
           const FwdParams p) {
  // No static smem in this kernel, so the dynamic window starts at the (1024 B aligned) base of
  // the CTA's shared memory; relying on that instead of padding keeps two CTAs resident per SM.
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* q_full = bars + 0;
  uint64_t* k_full = bars + 1;
  uint64_t* k_empty = bars + 2;
  uint64_t* v_full = bars + 3;
  uint64_t* v_empty = bars + 4;
  uint64_t* s_full = bars + 5;
  uint64_t* p_full = bars + 6;
  uint64_t* pv_done = bars + 7;   // [2], one per PV accumulator
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = (p.T + BQ - 1) / BQ;
  const int qt = nq - 1 - int(blockIdx.x);  // heaviest (most kv tiles) first
  const int h = blockIdx.y, b = blockIdx.z;
  const int q0 = qt * BQ;
  const int d = p.H * HD;
  const int n_kv = qt + 1;

  // The TMA thread initialises the barriers itself and requests Q, K_0, V_0 BEFORE the CTA-wide
  // sync, so TMEM allocation and the barrier round trip overlap the load latency.
  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmO);
    mbar_init(q_full, 1);
    mbar_init(k_full, 1);
    mbar_init(k_empty, 1);
    mbar_init(v_full, 1);
    mbar_init(v_empty, 1);
    mbar_init(s_full, 1);
    mbar_init(p_full, 4);
    mbar_init(&pv_done[0], 1);
    mbar_init(&pv_done[1], 1);
    fence_barrier_init();
    pdl_wait();  // qkv comes from the previous kernel of the stream
    mbar_expect_tx(q_full, TILE_BYTES);
    tma_load_3d(smem + OFF_Q, &tmQKV, q_full, h * HD, q0, b);
    mbar_expect_tx(k_full, TILE_BYTES);
    tma_load_3d(smem + OFF_K, &tmQKV, k_full, d + h * HD, 0, b);
    mbar_expect_tx(v_full, TILE_BYTES);
    tma_load_3d(smem + OFF_V, &tmQKV, v_full, 2 * d + h * HD, 0, b);
  }
  if (warp == 5) {
    tmem_alloc(tmem_slot, FWD_TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  pdl_enter();  // everything above overlapped the previous kernel's tail
  const uint32_t tmem = *tmem_slot;
  const uint32_t tmem_S = tmem;           // 128 columns
  const uint32_t tmem_PV = tmem + 128;    // 2 x 64 columns

  if (warp == 4) {
    // ------------------------------- TMA producer -------------------------------------------
    if (lane == 0) {
      for (int j = 1; j < n_kv; ++j) {  // tile 0 was requested in the prologue
        mbar_wait(k_empty, (j - 1) & 1);   // S_{j-1} has consumed K
        mbar_expect_tx(k_full, TILE_BYTES);
        tma_load_3d(smem + OFF_K, &tmQKV, k_full, d + h * HD, j * BKV, b);
        mbar_wait(v_empty, (j - 1) & 1);   // PV_{j-1} has consumed V
        mbar_expect_tx(v_full, TILE_BYTES);
        tma_load_3d(smem + OFF_V, &tmQKV, v_full, 2 * d + h * HD, j * BKV, b);
      }
    }
  } else if (warp == 5) {
    // ------------------------------- MMA issuer ---------------------------------------------
    // Issue order per tile: ... p_full(j) -> S_{j+1} -> PV_j.  S_{j+1} goes first because the
    // softmax warps need it next; PV_j is only folded one iteration later.
    if (lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(BQ, BKV, false, false);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(BQ, HD, false, true);
      const uint32_t sQ = smem_u32(smem + OFF_Q);
      const uint32_t sK = smem_u32(smem + OFF_K);
      const uint32_t sV = smem_u32(smem + OFF_V);
      auto issue_s = [&](int j) {
        mbar_wait(k_full, j & 1);
        tc_fence_after_sync();
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_ss(tmem_S, umma_smem_desc(sQ + k * 32, 0, 1024), umma_smem_desc(sK + k * 32, 0, 1024),
                  idesc_s, k > 0);
        umma_commit(k_empty);
        umma_commit(s_full);
      };
      mbar_wait(q_full, 0);
      issue_s(0);
      for (int j = 0; j < n_kv; ++j) {
        mbar_wait(p_full, j & 1);  // P_j is in smem and S_j has been read out of TMEM
        if (j + 1 < n_kv) issue_s(j + 1);
        mbar_wait(v_full, j & 1);
        tc_fence_after_sync();
        const uint32_t sP = smem_u32(smem + OFF_P + (j & 1) * P_BYTES);
        const uint32_t acc = tmem_PV + (j & 1) * HD;
#pragma unroll
        for (int k = 0; k < BKV / 16; ++k)
          umma_ss(acc, umma_smem_desc(sP + (k >> 2) * TILE_BYTES + (k & 3) * 32, 0, 1024),
                  umma_smem_desc(sV + k * 2048, TILE_BYTES, 1024), idesc_pv, k > 0);
        umma_commit(v_empty);
        umma_commit(&pv_done[j & 1]);
      }
    }
  } else {
    // ------------------------------- softmax / output ---------------------------------------
    const int r = threadIdx.x;  // 0..127: query row inside the tile == TMEM lane
    const int qrow = q0 + r;
    const uint32_t lane_base = uint32_t(warp * 32) << 16;
    float m_run = -INFINITY, l_run = 0.f, alpha_prev = 0.f;
    float o[HD];
#pragma unroll
    for (int i = 0; i < HD; ++i) o[i] = 0.f;
    const uint8_t* kmask = p.key_mask ? p.key_mask + size_t(b) * p.T : nullptr;

    // o = o * alpha + PV_j, PV_j read from its TMEM accumulator
    auto fold = [&](int j, float alpha) {
      mbar_wait(&pv_done[j & 1], (j >> 1) & 1);
      tc_fence_after_sync();
#pragma unroll
      for (int c = 0; c < HD / 32; ++c) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_PV + (j & 1) * HD + lane_base + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[c * 32 + i] = o[c * 32 + i] * alpha + __uint_as_float(v[i]);
      }
      tc_fence_before_sync();
    };

    for (int j = 0; j < n_kv; ++j) {
      const int kv0 = j * BKV;
      uint8_t* sP = smem + OFF_P + (j & 1) * P_BYTES;
      mbar_wait(s_full, j & 1);
      tc_fence_after_sync();

      // Two passes over S in TMEM (max, then exp2 + pack).  The masked variant is only
      // instantiated for the diagonal tile / key-padding case, so off-diagonal tiles execute no
      // per-element integer or predicate instructions at all.
      // warp-uniform mode: general when key padding is present, diagonal for the last tile
      const int mode = kmask != nullptr ? kGeneral : (j == qt ? kDiag : kNoMask);
      const uint32_t t_row = tmem_S + lane_base;
      float mx;
      if (mode == kNoMask) mx = row_max<kNoMask>(t_row, kv0, qrow, p.T, kmask, warp, lane);
      else if (mode == kDiag) mx = row_max<kDiag>(t_row, kv0, qrow, p.T, kmask, warp, lane);
      else mx = row_max<kGeneral>(t_row, kv0, qrow, p.T, kmask, warp, lane);
      const float m_new = fmaxf(m_run, mx * p.scale_log2);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      const float alpha = fast_exp2(m_run - m_use);  // m_run = -inf -> 0
      float sum;
      const uint64_t mask_base =
          DROP ? (uint64_t((size_t(b) * p.H + h) * p.T + qrow) * uint64_t(p.t_pad) + uint64_t(kv0))
               : 0;
      if (mode == kNoMask)
        sum = exp_pack<kNoMask, DROP>(t_row, sP, r, kv0, qrow, p.T, kmask, p.scale_log2, m_use,
                                      warp, lane, p.drop, mask_base);
      else if (mode == kDiag)
        sum = exp_pack<kDiag, DROP>(t_row, sP, r, kv0, qrow, p.T, kmask, p.scale_log2, m_use, warp,
                                    lane, p.drop, mask_base);
      else
        sum = exp_pack<kGeneral, DROP>(t_row, sP, r, kv0, qrow, p.T, kmask, p.scale_log2, m_use,
                                       warp, lane, p.drop, mask_base);
      l_run = l_run * alpha + sum;
      m_run = m_new;

      fence_proxy_async_smem();
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);

      // Fold the PREVIOUS tile's PV while the tensor core works on S_{j+1} / PV_j: the PV latency
      // never sits between two softmax passes.
      if (j > 0) fold(j - 1, alpha_prev);
      alpha_prev = alpha;
    }
    fold(n_kv - 1, alpha_prev);

    // normalise, stage as bf16 (swizzled 128 B rows) in a P buffer, one TMA store per CTA
    const float inv_l = l_run > 0.f ? 1.0f / l_run : 0.f;
    uint8_t* stage = smem + OFF_P + (n_kv & 1) * P_BYTES;  // the buffer the last PV did NOT use
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      uint32_t w[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        w[u] = pack_bf16x2(o[q * 8 + 2 * u] * inv_l, o[q * 8 + 2 * u + 1] * inv_l);
      *reinterpret_cast<uint4*>(stage + swizzle128(r * 128 + q * 16)) =
          make_uint4(w[0], w[1], w[2], w[3]);
    }
    if (qrow < p.T) {
      p.lse[(size_t(b) * p.H + h) * p.T + qrow] =
          l_run > 0.f ? (m_run + log2f(l_run)) * 0.6931471805599453f : -INFINITY;
    }
    fence_proxy_async_smem();
    named_bar_sync(1, 128);
    if (threadIdx.x == 0) {
      tma_store_3d(&tmO, stage, h * HD, q0, b);
      tma_store_commit();
      tma_store_wait_read<0>();  // smem may go once it has been read; the writes land on their own
    }
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 5) tmem_dealloc(tmem, FWD_TMEM_COLS);
}

}  // namespace attn
}  // namespace b200

using namespace b200;

// qkv: [B, T, 3*H*64] bf16 (q | k | v, head-major inside each third).  out: [B, T, H*64] bf16.
// lse: [B, H, T] fp32.  key_mask: [B, T] uint8 or NULL.
static int attention_fwd_impl(const void* qkv, void* out, float* lse, const uint8_t* key_mask,
                              int B, int T, int H, int head_dim, const DropoutParams& drop,
                              float softmax_scale, cudaStream_t stream) {
  B200_REQUIRE(head_dim == attn::HD, "attention: the tcgen05 kernel is specialised for head_dim 64");
  B200_REQUIRE(B > 0 && T > 0 && H > 0, "attention: empty problem");
  const int d = H * attn::HD;
  CUtensorMap tQKV, tO;
  if (make_tmap_3d(&tQKV, static_cast<const __nv_bfloat16*>(qkv), B, T, 3 * d, 3 * d,
                   uint64_t(T) * 3 * d, 128, attn::HD, Swz::B128))
    return 1;
  if (make_tmap_3d(&tO, static_cast<__nv_bfloat16*>(out), B, T, d, d, uint64_t(T) * d, 128,
                   attn::HD, Swz::B128))
    return 1;
  attn::FwdParams p;
  p.B = B; p.T = T; p.H = H;
  // softmax_scale > 0 overrides head_dim^-0.5 (heads narrower than 64 run zero-padded to 64 lanes
  // and keep the scale of their true width)
  p.scale_log2 = 1.4426950408889634f *
                 (softmax_scale > 0.f ? softmax_scale : 1.0f / std::sqrt(float(head_dim)));
  p.key_mask = key_mask;
  p.lse = lse;
  p.drop = drop;
  p.t_pad = (T + 7) & ~7;
  static bool configured = false;
  if (!configured) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn::fwd_kernel<false>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         attn::FWD_SMEM));
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn::fwd_kernel<true>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         attn::FWD_SMEM));
    configured = true;
  }
  dim3 grid(ceil_div(T, attn::BQ), H, B);
  if (drop.threshold != 0)
    B200_LAUNCH((attn::fwd_kernel<true>), grid, attn::FWD_THREADS, attn::FWD_SMEM, stream, tQKV, tO, p);
  else
    B200_LAUNCH((attn::fwd_kernel<false>), grid, attn::FWD_THREADS, attn::FWD_SMEM, stream, tQKV, tO,
                p);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}

extern "C" int llmtrain_b200_attention_fwd(const void* qkv, void* out, float* lse,
                                           const uint8_t* key_mask, int B, int T, int H,
                                           int head_dim, cudaStream_t stream) {
  return attention_fwd_impl(qkv, out, lse, key_mask, B, T, H, head_dim,
                            make_dropout(0.f, 0, 0, 0), 0.f, stream);
}

// As above with dropout on the attention probabilities.  Mask element index:
// ((b*H + h)*T + q) * roundup(T, 8) + k  (llmtrain_b200_dropout_mask with rows = B*H*T, cols = T).
extern "C" int llmtrain_b200_attention_fwd_dropout(const void* qkv, void* out, float* lse,
                                                   const uint8_t* key_mask, int B, int T, int H,
                                                   int head_dim, float p, uint64_t seed,
                                                   uint64_t offset, uint32_t site,
                                                   cudaStream_t stream) {
  B200_REQUIRE(p >= 0.f && p < 1.f, "attention_fwd_dropout: p must be in [0, 1)");
  return attention_fwd_impl(qkv, out, lse, key_mask, B, T, H, head_dim,
                            make_dropout(p, seed, offset, site), 0.f, stream);
}

// Everything at once (what the training engine calls): optional dropout (p = 0: none) and an
// optional softmax scale (0 = head_dim^-0.5).
extern "C" int llmtrain_b200_attention_fwd_fused(const void* qkv, void* out, float* lse,
                                                 const uint8_t* key_mask, int B, int T, int H,
                                                 int head_dim, float softmax_scale, float p,
                                                 uint64_t seed, uint64_t offset, uint32_t site,
                                                 cudaStream_t stream) {
  B200_REQUIRE(p >= 0.f && p < 1.f, "attention_fwd_fused: p must be in [0, 1)");
  return attention_fwd_impl(qkv, out, lse, key_mask, B, T, H, head_dim,
                            make_dropout(p, seed, offset, site), softmax_scale, stream);
}
