// Causal flash attention backward on tcgen05 (head_dim 64): launcher, pre / post passes and the
// ROW-ORIENTED main kernel, which serves probability dropout (its keep mask is regenerated along
// key positions, the direction this orientation walks).  Everything else runs the key-row-oriented
// kernel of attention_bwd_t.cuh, which is faster; both are persistent and share the launcher.
//
// Autograd of `src/llmtrain/models/gpt.py:56-70` without ever storing P (work
// item W3).  One CTA owns a 128-row key/value tile j of one (batch, head) and walks the query
// tiles i >= j.  Five tensor-core products per (i, j) pair, all accumulating in TMEM:
//
//   S   = Q_i K_j^T                (128x128x64)      dP  = dO_i V_j^T              (128x128x64)
//   dV += P^T  dO_i                (128x64x128)      dK += dS^T Q_i               (128x64x128)
//   dQ_i = dS  K_j                 (128x64x128)   -> fp32 TMA reduce-add into the dQ buffer
//
// with P = exp2(S*scale*log2e - LSE_i*log2e) and dS = P o (dP - D_i) * scale computed by four
// warps that own one query row each (TMEM lane == row).  P and dS are written once to shared
// memory as bf16 in the 128-byte-swizzled K-major tile layout and are then consumed twice by the
// tensor core through two different descriptors: K-major (dQ = dS K) and MN-major (P^T, dS^T) —
// no transposes are materialised.  V_j, dO_i, Q_i, K_j are likewise used both K-major and
// MN-major straight from their TMA-loaded tiles.
//
// D_i = rowsum(dO o O) comes from a small pre-kernel.  dK/dV leave through one TMA store each into
// the fused [B, T, 3d] gradient buffer; dQ contributions from different kv tiles meet in an fp32
// buffer (TMA reduce-add of 32x32 slices) that a post-kernel converts into the q third of that
// same gradient buffer.
//
// Pipeline.  The tensor core executes its queue in order, so the issue order IS the schedule.  After
// the softmax-grad warps hand over P_i / dS_i the MMA warp issues
//     dV += P_i^T dO_i   ->   S_{i+1}, dP_{i+1}   ->   dK += dS_i^T Q_i   ->   dQ_i = dS_i K
// P has ONE shared-memory buffer (its only reader, the dV product, is first in the queue and has
// retired by the time S_{i+1}/dP_{i+1} complete, which is what the warps wait for before they write
// P_{i+1}); dS has two, so dK/dQ of tile i run underneath the math of tile i+1; dQ has two TMEM
// accumulators and is drained one iteration late.  The 32 KB saved on P hold the fp32 dQ tile on
// its way out: each warp stages its 32x32 slice and hands it to the TMA unit as one reduce-add
// into the fp32 dQ buffer (a post-kernel converts that buffer into the q third of dqkv).
// What limited this kernel when it had one CTA per item and issued from `if (lane == 0)`
// (B = 32, T = 1024, H = 12, 425 us per layer before this change; each
// variant measured with that work removed): per-thread dQ atomics 76 us, the softmax-grad math
// 55 us, all five MMAs together ~45 us; the remaining skeleton (barrier hops MMA <-> warps, TMA
// latency) is ~230 us, and the dsum / memset / convert passes around the kernel ~60 us.

#include "../common/host.cuh"
#include "../common/philox.cuh"
#include "../common/ptx.cuh"
#include "../../include/llmtrain_b200.h"
#include "attention_bwd_t.cuh"

namespace b200 {
namespace attn_bwd {

constexpr int BQ = 128;
constexpr int BKV = 128;
constexpr int HD = 64;
constexpr int TILE = 128 * HD * 2;  // 16 KB
constexpr int THREADS = 320;  // 8 softmax-grad warps + TMA warp + MMA warp
constexpr int TMEM_COLS = 512;
constexpr int COL_S = 0, COL_DP = 128, COL_DV = 256, COL_DK = 320, COL_DQ = 384;  // dQ: 2 x 64 cols
constexpr int OFF_K = 0;
constexpr int OFF_V = OFF_K + TILE;
constexpr int NQ = 2;                        // Q / dO ring depth
constexpr int OFF_Q = OFF_V + TILE;          // [NQ]
constexpr int OFF_DO = OFF_Q + NQ * TILE;    // [NQ]
constexpr int PDS_BYTES = 2 * TILE;          // one 128x128 bf16 tile = two 64-column regions
constexpr int OFF_P = OFF_DO + NQ * TILE;    // 32 KB, single buffer
constexpr int OFF_DS = OFF_P + PDS_BYTES;    // [2] x 32 KB
constexpr int OFF_DQS = OFF_DS + 2 * PDS_BYTES;  // 32 KB: fp32 dQ tile on its way out (8 warp slices)
constexpr int OFF_BAR = OFF_DQS + 2 * TILE;
constexpr int SMEM = OFF_BAR + 256 + 1024;

struct Params {
  int B, T, H;
  float scale;       // head_dim^-0.5
  float scale_log2;  // scale * log2(e)
  const float* lse;  // [B, H, T]
  const float* dsum; // [B, H, T]   D = rowsum(dO o O)
  const uint8_t* key_mask;
  float* dq32;       // [B, T, H*64] fp32 accumulator (zeroed by the launcher)
  float* dbias;      // optional [3*H*64]: += column sums of dqkv (bias gradient of qkv_proj)
  DropoutParams drop;  // attention-probability dropout; threshold 0 = off
  int t_pad;           // T rounded up to 8: row pitch of the mask's element index
  long long* trace;    // debug (LLMTRAIN_ATTN_TRACE): clock64 stamps of CTA 0, else nullptr
  int* turn;           // deterministic mode: [B, H, nq, 8] contribution counters (zeroed), else nullptr
};

// trace layout: [role 0 = softmax-grad warp 0, 1 = MMA, 2 = TMA][iteration < 16][4 stamps]
__device__ __forceinline__ void stamp(const Params& p, int role, int it, int k) {
  if (p.trace != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && it < 16)
    p.trace[(role * 16 + it) * 4 + k] = clock64();
}

// Mask modes of one 128x128 (query x key) tile, chosen warp-uniformly (the loops contain
// .sync.aligned TMEM loads):
//   kNoMask   strictly below the diagonal, no key padding, no dead rows
//   kDiag     the diagonal tile (q0 == kv0), no key padding, no dead rows.  A warp owns rows
//             32*quarter..+31, so 32-column chunk cc of the tile is kept whole for cc < quarter,
//             dropped whole for cc > quarter (zeros, no TMEM load, no MUFU), and tested per
//             element (i <= lane) only for cc == quarter.
//   kGeneral  key padding / rows past T / fully masked rows: per-element test.
enum MaskMode : int { kNoMask = 0, kDiag = 1, kGeneral = 2 };

// One thread: 64 columns of its row.  P = exp2(S*scale*log2e - lse*log2e), dS = P (dP - D) scale,
// both written as bf16 into one 64-column swizzled region (row r, 8 x 16 B chunks).
// DROP (attention-probability dropout, keep factor m in {0, 1/(1-p)} regenerated from the
// forward's counters): the tile handed to dV is P o m, and dS = P o (dP o m - D) * scale; D is
// still rowsum(dO o O) because O was computed from the dropped probabilities.
template <int MODE, bool DROP>
__device__ __forceinline__ void softmax_grad(uint32_t t_row, uint8_t* rp, uint8_t* rd, int r,
                                             int col0, int qrow, int T, const uint8_t* kmask,
                                             bool dead_row, float scale_log2, float scale,
                                             float lse2, float dsum, int quarter, int half,
                                             int lane, const DropoutParams& dp,
                                             uint64_t mask_base) {
#pragma unroll 1
  for (int c = 0; c < 2; ++c) {
    const int cc = half * 2 + c;  // 32-column chunk index inside the 128-column tile
    if (MODE == kDiag && cc > quarter) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t off = swizzle128(r * 128 + (c * 4 + q) * 16);
        *reinterpret_cast<uint4*>(rp + off) = make_uint4(0u, 0u, 0u, 0u);
        *reinterpret_cast<uint4*>(rd + off) = make_uint4(0u, 0u, 0u, 0u);
      }
      continue;
    }
    uint32_t sv[32], dv[32];
    tmem_ld_32x32b_x32(t_row + COL_S + c * 32, sv);
    tmem_ld_32x32b_x32(t_row + COL_DP + c * 32, dv);
    tmem_ld_wait();
    uint32_t pk[16], dk[16];
    const bool plain = MODE == kNoMask || (MODE == kDiag && cc < quarter);
    Rand8 rnd{};
#pragma unroll
    for (int i = 0; i < 32; i += 2) {
      float pe[2], de[2];
      if (DROP && (i & 7) == 0) rnd = philox_group(dp, (mask_base + uint64_t(c * 32 + i)) >> 3);
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        float pv = fast_exp2(fmaf(__uint_as_float(sv[i + u]), scale_log2, -lse2));
        if (!plain) {
          if (MODE == kDiag) {
            if (i + u > lane) pv = 0.f;
          } else {
            const int col = col0 + c * 32 + i + u;
            bool ok = !dead_row && (col <= qrow);
            if (kmask != nullptr && col < T) ok = ok && (kmask[col] != 0);
            if (!ok) pv = 0.f;
          }
        }
        if (DROP) {
          const float m = keep_factor(dp, rnd, (i & 7) + u);
          pe[u] = pv * m;
          de[u] = pv * (__uint_as_float(dv[i + u]) * m - dsum) * scale;
        } else {
          pe[u] = pv;
          de[u] = pv * (__uint_as_float(dv[i + u]) - dsum) * scale;
        }
      }
      pk[i >> 1] = pack_bf16x2(pe[0], pe[1]);
      dk[i >> 1] = pack_bf16x2(de[0], de[1]);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t off = swizzle128(r * 128 + (c * 4 + q) * 16);
      *reinterpret_cast<uint4*>(rp + off) =
          make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
      *reinterpret_cast<uint4*>(rd + off) =
          make_uint4(dk[4 * q], dk[4 * q + 1], dk[4 * q + 2], dk[4 * q + 3]);
    }
  }
}

// One work item = one (key tile jt, batch b, head h): the CTA keeps K_j / V_j in shared memory,
// sweeps the query tiles jt..nq-1 and leaves dK_j / dV_j.  Items are numbered heaviest first
// (jt = 0 has the most query tiles) and CTA c takes items c, c + grid, c + 2 grid, ...
struct Item {
  int jt, b, h, n_iter;
};
__device__ __forceinline__ Item make_item(const Params& p, int idx, int nq) {
  const int bh = p.B * p.H;
  Item w;
  w.jt = idx / bh;
  const int r = idx - w.jt * bh;
  w.b = r / p.H;
  w.h = r - w.b * p.H;
  w.n_iter = nq - w.jt;
  return w;
}

// Persistent: grid = min(items, SMs), one CTA per SM.  Every barrier runs on a counter that keeps
// counting across items (g = query-tile iterations, k = items), so an item boundary costs no
// barrier re-initialisation, no TMEM allocation and no CTA launch: the next item's first Q / dO
// tile is requested while the current item's last MMAs run, K / V follow as soon as those MMAs
// have retired (dkv_full), and the first score tile is issued while the softmax-grad warps still
// convert dK / dV.  The one-CTA-per-item version spent 8 300 of 41 000 cycles per heaviest item
// (and of 12 000 per lightest) outside its main loop.
template <bool DROP>
__global__ void __launch_bounds__(THREADS, 1)
bwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
           const __grid_constant__ CUtensorMap tmDKV, const __grid_constant__ CUtensorMap tmDQ,
           const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* kv_full = bars + 0;
  uint64_t* q_full = bars + 1;         // [NQ]
  uint64_t* q_empty = q_full + NQ;     // [NQ]
  uint64_t* s_full = q_empty + NQ;
  uint64_t* pds_full = s_full + 1;
  uint64_t* dq_full = pds_full + 1;    // [2]: dQ_g accumulates in TMEM buffer g&1 and is drained
                                       // one iteration later, behind the next tile's softmax-grad
  uint64_t* dkv_full = dq_full + 2;    // the item's last MMA has retired: dK / dV are complete and
                                       // K / V / P / dS in shared memory are no longer read
  uint64_t* dkv_drained = dkv_full + 1;  // 8 warps: dK / dV have left TMEM
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(dkv_drained + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = (p.T + BQ - 1) / BQ;
  const int d = p.H * HD;
  const int total = nq * p.B * p.H;
  const int first = blockIdx.x, stride = gridDim.x;

  // The TMA thread initialises the barriers itself and issues the first item's loads (K_j, V_j,
  // Q / dO of the first query tile) BEFORE the CTA-wide sync: TMEM allocation and the barrier
  // round trip then overlap the ~1 us load latency instead of preceding it.
  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmDO);
    tma_prefetch_desc(&tmDKV);
    tma_prefetch_desc(&tmDQ);
    mbar_init(kv_full, 1);
    for (int s = 0; s < NQ; ++s) {
      mbar_init(&q_full[s], 1);
      mbar_init(&q_empty[s], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(pds_full, 8);
    mbar_init(&dq_full[0], 1);
    mbar_init(&dq_full[1], 1);
    mbar_init(dkv_full, 1);
    mbar_init(dkv_drained, 8);
    fence_barrier_init();
    pdl_wait();  // qkv / dO come from the previous kernels of the stream
    const Item w = make_item(p, first, nq);
    mbar_expect_tx(kv_full, 2 * TILE);
    tma_load_3d(smem + OFF_K, &tmQKV, kv_full, d + w.h * HD, w.jt * BKV, w.b);
    tma_load_3d(smem + OFF_V, &tmQKV, kv_full, 2 * d + w.h * HD, w.jt * BKV, w.b);
    mbar_expect_tx(&q_full[0], 2 * TILE);
    tma_load_3d(smem + OFF_Q, &tmQKV, &q_full[0], w.h * HD, w.jt * BQ, w.b);
    tma_load_3d(smem + OFF_DO, &tmDO, &q_full[0], w.h * HD, w.jt * BQ, w.b);
  }
  if (warp == 9) {
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  pdl_enter();  // everything above overlapped the previous kernel's tail
  const uint32_t tmem = *tmem_slot;

  if (warp == 8) {
    // ------------------------------- TMA producer -------------------------------------------
    if (lane == 0) {
      int g = 0;
      for (int idx = first, k = 0; idx < total; idx += stride, ++k) {
        const Item w = make_item(p, idx, nq);
        for (int it = 0; it < w.n_iter; ++it, ++g) {
          if (k == 0 && it == 0) continue;  // requested in the prologue
          const int st = g % NQ;
          const uint32_t ph = uint32_t(g / NQ) & 1u;
          const int q0 = (w.jt + it) * BQ;
          stamp(p, 2, g, 0);
          mbar_wait(&q_empty[st], ph ^ 1u);
          stamp(p, 2, g, 1);
          mbar_expect_tx(&q_full[st], 2 * TILE);
          tma_load_3d(smem + OFF_Q + st * TILE, &tmQKV, &q_full[st], w.h * HD, q0, w.b);
          tma_load_3d(smem + OFF_DO + st * TILE, &tmDO, &q_full[st], w.h * HD, q0, w.b);
          if (it == 0) {
            // The item's first Q / dO tile is on its way; K / V have a single buffer and wait for
            // the previous item's last MMA.
            mbar_wait(dkv_full, uint32_t(k - 1) & 1u);
            mbar_expect_tx(kv_full, 2 * TILE);
            tma_load_3d(smem + OFF_K, &tmQKV, kv_full, d + w.h * HD, w.jt * BKV, w.b);
            tma_load_3d(smem + OFF_V, &tmQKV, kv_full, 2 * d + w.h * HD, w.jt * BKV, w.b);
          }
        }
      }
    }
  } else if (warp == 9) {
    // ------------------------------- MMA issuer ---------------------------------------------
    // The whole warp walks the loop, one elected lane issues, descriptors are built outside the
    // elected region and advanced by one add per instruction (see attention_bwd_t.cuh: inside
    // `if (lane == 0)` every tcgen05.mma cost ~85 cycles of issue).
    constexpr uint32_t id_s = umma_idesc_bf16(128, 128, false, false);   // Q K^T, dO V^T
    constexpr uint32_t id_t = umma_idesc_bf16(128, 64, true, true);     // P^T dO, dS^T Q
    constexpr uint32_t id_dq = umma_idesc_bf16(128, 64, false, true);   // dS K
    const bool leader = elect_one();
    const uint32_t sQ0 = smem_u32(smem + OFF_Q), sDO0 = smem_u32(smem + OFF_DO);
    // K-major tiles: +32 B per K step = +2 in the address field; MN-major: +2048 B = +128
    const uint64_t kK = umma_smem_desc(smem_u32(smem + OFF_K), 0, 1024);
    const uint64_t kV = umma_smem_desc(smem_u32(smem + OFF_V), 0, 1024);
    const uint64_t mK = umma_smem_desc(smem_u32(smem + OFF_K), TILE, 1024);
    const uint64_t mP = umma_smem_desc(smem_u32(smem + OFF_P), TILE, 1024);
    auto issue_scores = [&](int g) {  // S = Q K^T and dP = dO V^T of iteration g
      const int st = g % NQ;
      mbar_wait(&q_full[st], uint32_t(g / NQ) & 1u);
      tc_fence_after_sync();
      const uint64_t kQ = umma_smem_desc(sQ0 + st * TILE, 0, 1024);
      const uint64_t kDO = umma_smem_desc(sDO0 + st * TILE, 0, 1024);
      if (leader) {
        stamp(p, 1, g, 0);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) umma_ss(tmem + COL_S, kQ + 2 * k, kK + 2 * k, id_s, k > 0);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_ss(tmem + COL_DP, kDO + 2 * k, kV + 2 * k, id_s, k > 0);
        umma_commit(s_full);
      }
    };
    int g = 0;
    for (int idx = first, item = 0; idx < total; idx += stride, ++item) {
      const int n_iter = nq - idx / (p.B * p.H);
      mbar_wait(kv_full, uint32_t(item) & 1u);
      // S / dP of the previous item were read before its last pds_full, so the first score tile
      // of this item may be issued while the other warps are still in the previous epilogue.
      issue_scores(g);
      for (int it = 0; it < n_iter; ++it, ++g) {
        const int st = g % NQ;
        const uint64_t mQ = umma_smem_desc(sQ0 + st * TILE, TILE, 1024);
        const uint64_t mDO = umma_smem_desc(sDO0 + st * TILE, TILE, 1024);
        const uint32_t sDS = smem_u32(smem + OFF_DS + (g & 1) * PDS_BYTES);
        const uint64_t mDS = umma_smem_desc(sDS, TILE, 1024);
        // dS as the K-major left operand of dQ: two 64-wide regions TILE apart, 32 B per K step
        const uint64_t kDS0 = umma_smem_desc(sDS, 0, 1024), kDS1 = umma_smem_desc(sDS + TILE, 0, 1024);
        if (leader) stamp(p, 1, g, 1);
        mbar_wait(pds_full, uint32_t(g) & 1u);  // P_g / dS_g are in smem, S_g / dP_g have left TMEM
        if (it == 0 && item > 0) mbar_wait(dkv_drained, uint32_t(item - 1) & 1u);
        tc_fence_after_sync();
        if (leader) {
          stamp(p, 1, g, 2);
          // dV += P^T dO   (A, B both MN-major: 16 q-rows = 2048 B per K step).  First in the
          // queue: it is the only reader of the single P buffer.
#pragma unroll
          for (int k = 0; k < BQ / 16; ++k)
            umma_ss(tmem + COL_DV, mP + 128 * k, mDO + 128 * k, id_t, (it > 0) || (k > 0));
        }
        if (it + 1 < n_iter) issue_scores(g + 1);  // the softmax-grad warps wait on these
        if (leader) {
          // dK += dS^T Q
#pragma unroll
          for (int k = 0; k < BQ / 16; ++k)
            umma_ss(tmem + COL_DK, mDS + 128 * k, mQ + 128 * k, id_t, (it > 0) || (k > 0));
          umma_commit(&q_empty[st]);
          // dQ_g = dS K  (A K-major over 128 kv = two 64-wide regions; B = K_j MN-major)
#pragma unroll
          for (int k = 0; k < BKV / 16; ++k)
            umma_ss(tmem + COL_DQ + (g & 1) * HD, ((k >> 2) ? kDS1 : kDS0) + 2 * (k & 3),
                    mK + 128 * k, id_dq, k > 0);
          umma_commit(&dq_full[g & 1]);
          stamp(p, 1, g, 3);
        }
      }
      if (leader) umma_commit(dkv_full);
    }
  } else {
    // ------------------------------- softmax-grad / epilogue warps --------------------------
    // 8 warps: warp w owns TMEM lanes 32*(w%4)..+31 (query / kv row r) and the 64-column half
    // (w/4) of every 128-column tile — which is exactly one swizzled smem region of P / dS.
    const int quarter = warp & 3, half = warp >> 2;
    const int r = quarter * 32 + lane;  // 0..127
    const uint32_t lane_base = uint32_t(quarter * 32) << 16;
    // This warp's 4 KB staging slice: 32 rows of 128 bytes, swizzled.  It carries the warp's
    // 32 x 32 fp32 slice of every dQ tile and, at the end of an item, its 32 x 64 bf16 slice of
    // dV (half 0) or dK (half 1).  Each warp issues and tracks its own bulk copies, so the slice
    // needs no CTA-wide synchronisation.
    uint8_t* stage = smem + OFF_DQS + half * TILE + quarter * 4096;

    int g = 0;
    for (int idx = first, item = 0; idx < total; idx += stride, ++item) {
      const Item w = make_item(p, idx, nq);
      const int jt = w.jt, b = w.b, h = w.h, n_iter = w.n_iter;
      const int kv0 = jt * BKV;
      const uint8_t* kmask = p.key_mask ? p.key_mask + size_t(b) * p.T : nullptr;
      const size_t stat_base = (size_t(b) * p.H + h) * p.T;

      // dQ of iteration gi (query tile qt): TMEM -> registers -> the staging slice -> ONE TMA
      // reduce-add per warp into dq32.  (Per-thread red.global.add.v4 of the same data was the
      // single largest cost of this kernel: 453 MB of 16-byte L2 reductions per layer, 76 us of
      // 425 at B = 32; the TMA unit hands L2 whole 128-byte lines.)  Called one iteration late, so
      // the MMAs that precede dq_full(gi) are never waited for.  Rows past T are clipped by the map.
      auto drain_dq = [&](int gi, int qt) {
        mbar_wait(&dq_full[gi & 1], uint32_t(gi >> 1) & 1u);
        tc_fence_after_sync();
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem + COL_DQ + (gi & 1) * HD + lane_base + half * 32, v);
        tmem_ld_wait();
        tc_fence_before_sync();
        if (lane == 0) tma_store_wait_read<0>();  // the previous slice has left shared memory
        __syncwarp();
#pragma unroll
        for (int q = 0; q < 8; ++q)
          *reinterpret_cast<uint4*>(stage + swizzle128(lane * 128 + q * 16)) =
              make_uint4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          // Deterministic mode: the key tiles add into a query tile in ascending order.  This item
          // (key tile jt) waits until jt earlier contributions to this 32x32 slice have landed and
          // publishes its own once the reduction has completed.  An item with a smaller key tile
          // has a smaller index; every CTA is resident and walks its items in ascending index
          // order, so the unfinished item with the smallest index is always running and never
          // waits: the chain cannot deadlock.
          int* turn = nullptr;
          if (p.turn != nullptr) {
            turn = p.turn + ((size_t(b) * p.H + h) * nq + qt) * 8 + (warp & 7);
            const unsigned long long t0 = global_timer_ns();
            int seen;
            do {
              asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(turn) : "memory");
              if (seen != jt && global_timer_ns() - t0 > B200_MBAR_TIMEOUT_NS) mbar_timeout_trap();
            } while (seen != jt);
          }
          tma_reduce_add_3d(&tmDQ, stage, h * HD + half * 32, qt * BQ + quarter * 32, b);
          tma_store_commit();
          if (turn != nullptr) {
            tma_store_wait<0>();  // the reduction has been performed, not merely read from smem
            __threadfence();
            asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(turn), "r"(jt + 1) : "memory");
          }
        }
      };

      // Row statistics of the NEXT query tile are requested before this tile's math so their
      // global-memory latency is never exposed.
      float lse_next = 0.f, dsum_next = 0.f;
      {
        const int qrow0 = jt * BQ + r;
        if (qrow0 < p.T) {
          lse_next = p.lse[stat_base + qrow0];
          dsum_next = p.dsum[stat_base + qrow0];
        }
      }
      for (int it = 0; it < n_iter; ++it, ++g) {
        const int qt = jt + it;
        const int q0 = qt * BQ;
        const int qrow = q0 + r;
        const bool row_ok = qrow < p.T;
        const float lse2 = row_ok ? lse_next * 1.4426950408889634f : 0.f;
        const float dsum = row_ok ? dsum_next : 0.f;
        if (it + 1 < n_iter && qrow + BQ < p.T) {
          lse_next = p.lse[stat_base + qrow + BQ];
          dsum_next = p.dsum[stat_base + qrow + BQ];
        }
        const bool dead_row = !row_ok || (lse2 == -INFINITY);
        // warp-uniform on purpose: the variants contain .sync.aligned TMEM loads
        const int mode = (kmask != nullptr || __any_sync(0xffffffffu, dead_row) != 0)
                             ? kGeneral
                             : (it == 0 ? kDiag : kNoMask);

        if (threadIdx.x == 0) stamp(p, 0, g, 0);
        mbar_wait(s_full, uint32_t(g) & 1u);
        tc_fence_after_sync();
        if (threadIdx.x == 0) stamp(p, 0, g, 1);
        uint8_t* rp = smem + OFF_P + half * TILE;
        uint8_t* rd = smem + OFF_DS + (g & 1) * PDS_BYTES + half * TILE;
        const uint32_t t_row = tmem + lane_base + half * 64;
        const uint64_t mask_base =
            DROP ? (uint64_t((size_t(b) * p.H + h) * p.T + qrow) * uint64_t(p.t_pad) +
                    uint64_t(kv0 + half * 64))
                 : 0;
        if (mode == kNoMask)
          softmax_grad<kNoMask, DROP>(t_row, rp, rd, r, kv0 + half * 64, qrow, p.T, kmask, dead_row,
                                      p.scale_log2, p.scale, lse2, dsum, quarter, half, lane,
                                      p.drop, mask_base);
        else if (mode == kDiag)
          softmax_grad<kDiag, DROP>(t_row, rp, rd, r, kv0 + half * 64, qrow, p.T, kmask, dead_row,
                                    p.scale_log2, p.scale, lse2, dsum, quarter, half, lane, p.drop,
                                    mask_base);
        else
          softmax_grad<kGeneral, DROP>(t_row, rp, rd, r, kv0 + half * 64, qrow, p.T, kmask,
                                       dead_row, p.scale_log2, p.scale, lse2, dsum, quarter, half,
                                       lane, p.drop, mask_base);
        fence_proxy_async_smem();
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(pds_full);
        if (threadIdx.x == 0) stamp(p, 0, g, 2);

        if (it > 0) drain_dq(g - 1, qt - 1);
        if (threadIdx.x == 0) stamp(p, 0, g, 3);
      }
      drain_dq(g - 1, jt + n_iter - 1);

      // dV (warps 0-3) / dK (warps 4-7): 32 kv rows x 64 columns per warp, TMEM -> bf16 -> the
      // staging slice -> one TMA store per warp.
      mbar_wait(dkv_full, uint32_t(item) & 1u);
      tc_fence_after_sync();
      uint32_t packed[32];
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem + (half == 0 ? COL_DV : COL_DK) + lane_base + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int u = 0; u < 16; ++u)
          packed[c * 16 + u] =
              pack_bf16x2(__uint_as_float(v[2 * u]), __uint_as_float(v[2 * u + 1]));
      }
      tc_fence_before_sync();
      if (lane == 0) tma_store_wait_read<0>();  // the last dQ slice has left the staging slice
      __syncwarp();
      if (lane == 0) mbar_arrive(dkv_drained);  // the next item may overwrite dK / dV in TMEM
#pragma unroll
      for (int q = 0; q < 8; ++q)
        *reinterpret_cast<uint4*>(stage + swizzle128(lane * 128 + q * 16)) =
            make_uint4(packed[4 * q], packed[4 * q + 1], packed[4 * q + 2], packed[4 * q + 3]);
      fence_proxy_async_smem();
      __syncwarp();
      const int col0 = (half == 0 ? 2 * d : d) + h * HD;
      if (lane == 0) {
        tma_store_3d(&tmDKV, stage, col0, kv0 + quarter * 32, b);
        tma_store_commit();
      }
      if (p.dbias != nullptr) {
        // qkv_proj's bias gradient is the column sum of dqkv: the k and v thirds come from the
        // bf16 tile just staged (lane l: columns 2l, 2l+1 over the warp's 32 rows; conflict-free
        // in the swizzled layout).  Rows past T hold exact zeros.  The q third comes from
        // dq_convert_kernel.
        float s0 = 0.f, s1 = 0.f;
#pragma unroll 1
        for (int r8 = 0; r8 < 32; r8 += 8) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t off = uint32_t(r8 + j) * 128u +
                                 ((uint32_t(lane >> 2) ^ uint32_t(j)) << 4) +
                                 (uint32_t(lane & 3) << 2);
            const float2 f = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(stage + off));
            s0 += f.x;
            s1 += f.y;
          }
        }
        atomicAdd(p.dbias + col0 + 2 * lane, s0);
        atomicAdd(p.dbias + col0 + 2 * lane + 1, s1);
      }
    }
    if (lane == 0) tma_store_wait_read<0>();  // shared memory must outlive the last store's read
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem, TMEM_COLS);
}

// D[b,h,t] = sum_c dO[b,t,h*64+c] * O[b,t,h*64+c]; 8 lanes x 16 B cover one head (64 bf16), so a
// warp handles 4 (row, head) pairs per load instruction.
__global__ void dsum_kernel(const __nv_bfloat16* __restrict__ dO,
                            const __nv_bfloat16* __restrict__ O, float* __restrict__ dsum, int B,
                            int T, int H) {
  pdl_enter();  // previous kernel finished, its writes are visible
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long pair = gid >> 3;  // (row, head) index
  const int sub = int(gid & 7);
  const long long total = (long long)B * T * H;
  float v = 0.f;
  if (pair < total) {
    const size_t off = size_t(pair) * HD + sub * 8;  // [row][h][64] is contiguous in (row, h)
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(dO + off));
    const uint4 o = __ldg(reinterpret_cast<const uint4*>(O + off));
    const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, ow[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 x = unpack_bf16x2(aw[j]), y = unpack_bf16x2(ow[j]);
      v += x.x * y.x + x.y * y.y;
    }
  }
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  if (sub == 0 && pair < total) {
    const long long row = pair / H;
    const int h = int(pair - row * H);
    const int b = int(row / T), t = int(row - (long long)b * T);
    dsum[(size_t(b) * H + h) * T + t] = v;
  }
}

// dqkv[row, 0:d] = bf16(dq32[row, :])   (q third of the fused gradient buffer), and optionally
// dbias[0:d] += column sums of the bf16 result (q third of qkv_proj's bias gradient).
// blockDim = (d/8) * nsub: thread (g, sub) owns 8 columns and every nsub-th row of the block's slab.
constexpr int DQ_SLAB_ROWS = 64;
__global__ void dq_convert_kernel(const float* __restrict__ dq32, __nv_bfloat16* __restrict__ dqkv,
                                  float* __restrict__ dbias, int rows, int d, int nsub) {
  pdl_enter();  // previous kernel finished, its writes are visible
  extern __shared__ float dq_red[];  // [nsub][d] partial column sums (only with dbias)
  const int per_row = d >> 3;
  const int g = threadIdx.x % per_row, sub = threadIdx.x / per_row;
  const int row0 = blockIdx.x * DQ_SLAB_ROWS;
  const int row_end = min(rows, row0 + DQ_SLAB_ROWS);
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll 4
  for (int row = row0 + sub; row < row_end; row += nsub) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(dq32 + size_t(row) * d) + 2 * g);
    const float4 b = __ldg(reinterpret_cast<const float4*>(dq32 + size_t(row) * d) + 2 * g + 1);
    const uint4 o = make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y),
                               pack_bf16x2(b.z, b.w));
    *reinterpret_cast<uint4*>(dqkv + size_t(row) * 3 * d + g * 8) = o;
    if (dbias != nullptr) {
      const uint32_t w[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = unpack_bf16x2(w[j]);
        acc[2 * j] += f.x;
        acc[2 * j + 1] += f.y;
      }
    }
  }
  if (dbias == nullptr) return;
#pragma unroll
  for (int j = 0; j < 8; ++j) dq_red[sub * d + g * 8 + j] = acc[j];
  __syncthreads();
  // blocks finish together: start each one at a different column so they do not all queue on the
  // same cache lines
  const int rot = int((blockIdx.x * 67u) % unsigned(d));
  for (int i = threadIdx.x; i < d; i += blockDim.x) {
    const int c = (i + rot) % d;
    float t = 0.f;
    for (int k = 0; k < nsub; ++k) t += dq_red[k * d + c];
    atomicAdd(dbias + c, t);
  }
}

}  // namespace attn_bwd
}  // namespace b200

using namespace b200;

namespace {
// Scratch layout (every region 256-byte aligned): fp32 dQ accumulator [B,T,d] | D [B,H,T] (dropout
// path) | the deterministic mode's turn counters [B,H,nq,8] | statistics [B,H,nq*128] float2.
struct BwdScratch {
  size_t dq, dsum, turn, stats, total;
};
inline size_t up256(size_t x) { return (x + 255) & ~size_t(255); }
inline BwdScratch bwd_scratch(int B, int T, int H, int head_dim) {
  const size_t nq = size_t((T + 127) / 128);
  BwdScratch s;
  s.dq = 0;
  s.dsum = up256(size_t(B) * T * H * head_dim * sizeof(float));
  s.turn = s.dsum + up256(size_t(B) * H * T * sizeof(float));
  s.stats = s.turn + up256(size_t(B) * H * nq * 8 * sizeof(int));
  s.total = s.stats + up256(size_t(B) * H * nq * 128 * sizeof(float2));
  return s;
}
}  // namespace

extern "C" size_t llmtrain_b200_attention_bwd_scratch_bytes(int B, int T, int H, int head_dim) {
  return bwd_scratch(B, T, H, head_dim).total;
}

// qkv [B,T,3d], out/dout [B,T,d] bf16, lse [B,H,T]  ->  dqkv [B,T,3d] bf16.
static int attention_bwd_impl(const void* qkv, const void* out, const void* dout, const float* lse,
                              const uint8_t* key_mask, void* dqkv, void* scratch,
                              size_t scratch_bytes, int B, int T, int H, int head_dim,
                              const DropoutParams& drop, float* dbias, float softmax_scale,
                              cudaStream_t stream, bool stats_ready = false) {
  namespace ab = attn_bwd;
  namespace at = attn_bwd_t;
  B200_REQUIRE(head_dim == ab::HD, "attention_bwd: specialised for head_dim 64");
  B200_REQUIRE(B > 0 && T > 0 && H > 0, "attention_bwd: empty problem");
  B200_REQUIRE(!stats_ready || drop.threshold == 0,
               "attention_bwd: precomputed statistics are not used with probability dropout");
  const int d = H * ab::HD;
  const BwdScratch lay = bwd_scratch(B, T, H, head_dim);
  B200_REQUIRE(scratch_bytes >= lay.total, "attention_bwd: scratch too small");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(scratch) & 255) == 0, "attention_bwd: scratch alignment");
  uint8_t* base = static_cast<uint8_t*>(scratch);
  float* dq32 = reinterpret_cast<float*>(base + lay.dq);
  float* dsum = reinterpret_cast<float*>(base + lay.dsum);
  float2* stats = reinterpret_cast<float2*>(base + lay.stats);
  const size_t dq_bytes = size_t(B) * T * d * sizeof(float);
  B200_CHECK_CUDA(cudaMemsetAsync(dq32, 0, dq_bytes, stream));

  // Probability dropout regenerates its keep mask along key positions, which only the
  // row-oriented kernel walks; everything else takes the key-row-oriented kernel.
  const bool with_dropout = drop.threshold != 0;
  const int rows = B * T;
  const int nq = ceil_div(T, ab::BQ);
  const int t_stat = nq * ab::BQ;
  if (with_dropout) {
    B200_LAUNCH((ab::dsum_kernel), ceil_div(rows * H * 8, 256), 256, 0, stream,
                static_cast<const __nv_bfloat16*>(dout), static_cast<const __nv_bfloat16*>(out),
                dsum, B, T, H);
  } else if (!stats_ready) {
    B200_LAUNCH((at::stats_kernel), dim3(unsigned(ceil_div(t_stat * H * 8, 256)), unsigned(B)), 256, 0,
                stream,
                static_cast<const __nv_bfloat16*>(dout), static_cast<const __nv_bfloat16*>(out), lse,
                stats, B, T, H, t_stat);
  } else if (t_stat != T) {
    B200_LAUNCH((at::stats_pad_kernel), ceil_div(B * H * (t_stat - T), 256), 256, 0, stream, stats,
                B * H, T, t_stat);
  }
  B200_CHECK_CUDA(cudaGetLastError());
  if (with_dropout || !stats_ready || t_stat != T) B200_COUNT_LAUNCH(1);

  CUtensorMap tQKV, tDO, tDKV, tDQ;
  if (make_tmap_3d(&tQKV, static_cast<const __nv_bfloat16*>(qkv), B, T, 3 * d, 3 * d,
                   uint64_t(T) * 3 * d, 128, ab::HD, Swz::B128)) return 1;
  if (make_tmap_3d(&tDO, static_cast<const __nv_bfloat16*>(dout), B, T, d, d, uint64_t(T) * d, 128,
                   ab::HD, Swz::B128)) return 1;
  // dK / dV leave in per-warp boxes of 32 kv rows x 64 columns
  if (make_tmap_3d(&tDKV, static_cast<__nv_bfloat16*>(dqkv), B, T, 3 * d, 3 * d,
                   uint64_t(T) * 3 * d, 32, ab::HD, Swz::B128)) return 1;
  // fp32 accumulator [B, T, d], boxes of 32 rows x 32 columns (128-byte rows)
  if (make_tmap_3d(&tDQ, dq32, B, T, d, d, uint64_t(T) * d, 32, 32, Swz::B128)) return 1;

  const float scale = softmax_scale > 0.f ? softmax_scale : 1.0f / std::sqrt(float(head_dim));
  int* turn = nullptr;
  if (deterministic()) {
    turn = reinterpret_cast<int*>(base + lay.turn);
    B200_CHECK_CUDA(cudaMemsetAsync(turn, 0, size_t(B) * H * nq * 8 * sizeof(int), stream));
  }
  constexpr size_t kTraceWords = 3 * 16 * 4;
  long long* trace = nullptr;
  static const bool want_trace = std::getenv("LLMTRAIN_ATTN_TRACE") != nullptr;
  if (want_trace) {
    B200_CHECK_CUDA(cudaMalloc(&trace, kTraceWords * sizeof(long long)));
    B200_CHECK_CUDA(cudaMemsetAsync(trace, 0, kTraceWords * sizeof(long long), stream));
  }
  static bool configured = false;
  if (!configured) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(ab::bwd_kernel<true>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, ab::SMEM));
    B200_CHECK_CUDA(cudaFuncSetAttribute(at::bwd_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, at::SMEM));
    configured = true;
  }
  const int items = nq * H * B;
  const int grid = items < sm_count() ? items : sm_count();  // persistent: one CTA per SM
  if (with_dropout) {
    ab::Params p;
    p.B = B; p.T = T; p.H = H;
    p.scale = scale;
    p.scale_log2 = scale * 1.4426950408889634f;
    p.lse = lse;
    p.dsum = dsum;
    p.key_mask = key_mask;
    p.dq32 = dq32;
    p.dbias = dbias;
    p.drop = drop;
    p.t_pad = (T + 7) & ~7;
    p.turn = turn;
    p.trace = trace;
    B200_LAUNCH((ab::bwd_kernel<true>), grid, ab::THREADS, ab::SMEM, stream, tQKV, tDO, tDKV, tDQ, p);
  } else {
    at::Params p;
    p.B = B; p.T = T; p.H = H;
    p.t_stat = t_stat;
    {
      // Q / dO / fp32 dQ of one (b, h) pair: nq x 64 KB.  When all pairs together fit in L2 the
      // list is one group (best balance); otherwise a group is as many pairs as there are CTAs.
      static const int forced = [] {
        const char* e = std::getenv("LLMTRAIN_ATTN_BWD_GROUP");
        return e != nullptr ? std::atoi(e) : 0;
      }();
      const size_t pair_bytes = size_t(nq) * 64 * 1024;
      p.group = forced > 0 ? forced : (size_t(B) * H * pair_bytes <= (size_t(64) << 20) ? B * H : grid);
    }
    p.scale = scale;
    p.scale_log2 = scale * 1.4426950408889634f;
    p.stats = stats;
    p.key_mask = key_mask;
    p.dbias = dbias;
    p.trace = trace;
    p.turn = turn;
    B200_LAUNCH((at::bwd_kernel), grid, at::THREADS, at::SMEM, stream, tQKV, tDO, tDKV, tDQ, p);
  }
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  if (trace != nullptr) {  // debug: timeline of CTA 0's first 16 (query tile, key tile) pairs
    static long long host[kTraceWords];
    B200_CHECK_CUDA(cudaMemcpyAsync(host, trace, sizeof(host), cudaMemcpyDeviceToHost, stream));
    B200_CHECK_CUDA(cudaStreamSynchronize(stream));
    const long long t0 = host[1];
    for (int it = 0; it < 16 && host[it * 4] != 0; ++it) {
      std::fprintf(stderr, "attn bwd trace it=%2d", it);
      for (int role = 0; role < 3; ++role)
        for (int k = 0; k < 4; ++k)
          std::fprintf(stderr, "%s%7lld", k == 0 ? " |" : "",
                       host[(role * 16 + it) * 4 + k] == 0 ? -1LL : host[(role * 16 + it) * 4 + k] - t0);
      std::fprintf(stderr, "\n");
    }
    cudaFree(trace);
  }

  {
    const int per_row = d / 8;
    const int nsub = per_row >= 256 ? 1 : 256 / per_row;
    const size_t red_bytes = dbias != nullptr ? size_t(nsub) * d * sizeof(float) : 0;
    B200_LAUNCH((ab::dq_convert_kernel), ceil_div(rows, ab::DQ_SLAB_ROWS), per_row * nsub, red_bytes,
                stream, dq32, static_cast<__nv_bfloat16*>(dqkv), dbias, rows, d, nsub);
  }
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}

// qkv [B,T,3d], out/dout [B,T,d] bf16, lse [B,H,T]  ->  dqkv [B,T,3d] bf16.
extern "C" int llmtrain_b200_attention_bwd(const void* qkv, const void* out, const void* dout,
                                           const float* lse, const uint8_t* key_mask, void* dqkv,
                                           void* scratch, size_t scratch_bytes, int B, int T,
                                           int H, int head_dim, cudaStream_t stream) {
  return attention_bwd_impl(qkv, out, dout, lse, key_mask, dqkv, scratch, scratch_bytes, B, T, H,
                            head_dim, make_dropout(0.f, 0, 0, 0), nullptr, 0.f, stream);
}

// Backward of llmtrain_b200_attention_fwd_dropout: pass the same (p, seed, offset, site).
extern "C" int llmtrain_b200_attention_bwd_dropout(const void* qkv, const void* out,
                                                   const void* dout, const float* lse,
                                                   const uint8_t* key_mask, void* dqkv,
                                                   void* scratch, size_t scratch_bytes, int B,
                                                   int T, int H, int head_dim, float p,
                                                   uint64_t seed, uint64_t offset, uint32_t site,
                                                   cudaStream_t stream) {
  B200_REQUIRE(p >= 0.f && p < 1.f, "attention_bwd_dropout: p must be in [0, 1)");
  return attention_bwd_impl(qkv, out, dout, lse, key_mask, dqkv, scratch, scratch_bytes, B, T, H,
                            head_dim, make_dropout(p, seed, offset, site), nullptr, 0.f, stream);
}

// Everything at once (what the training engine calls): optional softmax scale (0 = head_dim^-0.5),
// optional dropout (p = 0: none) and, when dbias_qkv is not NULL, dbias_qkv[3d] += column sums of dqkv — qkv_proj's bias gradient
// (`gpt.py:27`) without a separate pass over dqkv.  fp32 atomics: summation order is not fixed.
extern "C" int llmtrain_b200_attention_bwd_fused(const void* qkv, const void* out, const void* dout,
                                                 const float* lse, const uint8_t* key_mask,
                                                 void* dqkv, float* dbias_qkv, void* scratch,
                                                 size_t scratch_bytes, int B, int T, int H,
                                                 int head_dim, float softmax_scale, float p,
                                                 uint64_t seed, uint64_t offset, uint32_t site,
                                                 cudaStream_t stream) {
  B200_REQUIRE(p >= 0.f && p < 1.f, "attention_bwd_fused: p must be in [0, 1)");
  return attention_bwd_impl(qkv, out, dout, lse, key_mask, dqkv, scratch, scratch_bytes, B, T, H,
                            head_dim, make_dropout(p, seed, offset, site), dbias_qkv,
                            softmax_scale, stream);
}

// Where llmtrain_b200_gemm_rowdot has to write the statistics for a following
// llmtrain_b200_attention_bwd_prestats with the same scratch and shape: pointer and row pitch.
extern "C" int llmtrain_b200_attention_bwd_stats_buffer(void* scratch, size_t scratch_bytes, int B,
                                                        int T, int H, int head_dim, void** stats,
                                                        int* t_stat) {
  B200_REQUIRE(scratch != nullptr && stats != nullptr && t_stat != nullptr, "attention_bwd_stats_buffer: null");
  const BwdScratch lay = bwd_scratch(B, T, H, head_dim);
  B200_REQUIRE(scratch_bytes >= lay.total, "attention_bwd_stats_buffer: scratch too small");
  *stats = static_cast<uint8_t*>(scratch) + lay.stats;
  *t_stat = ceil_div(T, 128) * 128;
  return 0;
}

// As llmtrain_b200_attention_bwd_fused without probability dropout, for callers whose out_proj
// dgrad already left the per-row statistics in the scratch (llmtrain_b200_gemm_rowdot): the
// pre-pass over dout / out is skipped.
extern "C" int llmtrain_b200_attention_bwd_prestats(const void* qkv, const void* out,
                                                    const void* dout, const float* lse,
                                                    const uint8_t* key_mask, void* dqkv,
                                                    float* dbias_qkv, void* scratch,
                                                    size_t scratch_bytes, int B, int T, int H,
                                                    int head_dim, float softmax_scale,
                                                    cudaStream_t stream) {
  return attention_bwd_impl(qkv, out, dout, lse, key_mask, dqkv, scratch, scratch_bytes, B, T, H,
                            head_dim, make_dropout(0.f, 0, 0, 0), dbias_qkv, softmax_scale, stream,
                            true);
}
