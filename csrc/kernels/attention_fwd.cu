// Causal flash attention forward on tcgen05 (head_dim 64, bf16 in/out, fp32 softmax state).
//
// Replaces the reference's materialised attention (`src/llmtrain/models/gpt.py:56-70`: scores
// (B,H,T,T) fp32, masked_fill, softmax, P@V — SURVEY §2.5 K4, work item W2).  Nothing of size T^2
// ever reaches HBM.  One CTA per SM owns TWO
// 128-row query tiles of one (batch, head) and everything the tensor core touches besides Q/K/V
// lives in tensor memory:
//
//   TMEM  [  0,128) S_0   [128,256) S_1   [256,320) O_0   [320,384) O_1
//         [384,448) P_0   [448,512) P_1      (P_t: bf16, two probabilities per 32-bit column)
//
//   warp 8  (TMA)  : Q_0, Q_1 once; K_j / V_j through a 3-deep ring shared by both query tiles.
//   warp 9  (MMA)  : S_t = Q_t K_j^T (SS, 128x128x64) and O_t += P_t V_j (TS: the A operand is read
//                    straight from TMEM, V_j MN-major from its TMA tile).  S_t(j+1) is issued as
//                    soon as the softmax group has pulled S_t(j) into registers (~130 cycles after
//                    it appeared), PV_t(j) when P_t(j) has been written: the score product of the
//                    next tile runs underneath the exponentials of the current one and neither
//                    product is on a softmax group's critical path.  The whole warp walks the loop
//                    (uniform control flow keeps descriptors in uniform registers); one elected
//                    lane issues.
//   warps 0-3 / 4-7: softmax group of tile 0 / 1, one query row per thread.  ONE pass over S: the
//                    whole 128-column row is pulled into registers, max, exp2, bf16 pack,
//                    tcgen05.st back into TMEM.  O is never read in the steady state: the running
//                    maximum is only raised (and O_t / l rescaled in place) when it grew by more
//                    than 2^8, which keeps exp2(s - m_ref) <= 256 — exact in fp32 accumulation and
//                    harmless in bf16 — and happens on the first tiles of a row only.
//
// setmaxnreg moves registers from the TMA/MMA warpgroup to the two softmax warpgroups (one row of
// S is 128 live registers).  Work is scheduled heaviest first (most key tiles).
//
// Masking follows the reference's semantics: causal (col > row) and optional key padding are
// excluded from the softmax; a query whose keys are ALL excluded yields zeros and LSE = -inf (the
// reference zeroes such rows after out_proj, `gpt.py:73-74`).  LSE (natural log) is saved for
// backward.  Rows past T read zeros through the 3-D tensor maps over the fused [B, T, 3d] QKV
// buffer (no permute/copy is ever made) and are clipped on store.

#define B200_MBAR_TRAP_INLINE  // see ptx.cuh: required by the setmaxnreg register split
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
constexpr int THREADS = 384;              // 2 softmax warpgroups + {TMA, MMA, 2 idle} warpgroup
constexpr int NST = 3;                    // K/V ring depth
constexpr int TMEM_COLS = 512;
constexpr int COL_S = 0;    // + 128 * t
constexpr int COL_O = 256;  // + 64 * t
constexpr int COL_P = 384;  // + 64 * t
constexpr int OFF_Q = 0;                          // [2]
constexpr int OFF_K = OFF_Q + 2 * TILE_BYTES;     // [NST]
constexpr int OFF_V = OFF_K + NST * TILE_BYTES;   // [NST]
constexpr int OFF_O = OFF_V + NST * TILE_BYTES;   // [2] bf16 output tiles on their way out
constexpr int OFF_BAR = OFF_O + 2 * TILE_BYTES;
constexpr int SMEM = OFF_BAR + 512;
constexpr float kRescaleThreshold = 8.0f;  // log2 units
// Of every 4 consecutive column pairs, this many take their exponential on the FMA pipe (the rest
// use MUFU.EX2, 16 per clock per SM — the scarce unit of this kernel: one 128x128 tile needs 1024
// MUFU cycles against 512 tensor-core cycles).
#ifndef B200_ATTN_POLY_OF4
#define B200_ATTN_POLY_OF4 1
#endif
constexpr int kPolyOf4 = B200_ATTN_POLY_OF4;

enum MaskMode : int { kNoMask = 0, kDiag = 1, kGeneral = 2 };

template <int N>
__device__ __forceinline__ void reg_alloc() {
  asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N));
}
template <int N>
__device__ __forceinline__ void reg_dealloc() {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N));
}

struct Params {
  int B, T, H;
  int nqb;                  // query blocks (256 rows) per sequence
  float scale_log2;         // softmax scale * log2(e)
  const uint8_t* key_mask;  // [B, T] (1 = attend) or nullptr
  float* lse;               // [B, H, T]
  DropoutParams drop;
  int t_pad;
  long long* trace;  // debug (LLMTRAIN_ATTN_TRACE): clock64 stamps of CTA 0, else nullptr
};

// trace layout: [role 0/1 = softmax group, 2 = MMA, 3 = inside softmax group 0][tile j < 32][4]
__device__ __forceinline__ void stamp(const Params& p, int role, int j, int k) {
  if (p.trace != nullptr && blockIdx.x == 0 && j < 32) p.trace[(role * 32 + j) * 4 + k] = clock64();
}

// One softmax step of one query row: S row (TMEM) -> P row (bf16, TMEM), running (m_ref, l).
// ONE copy of this code serves every tile: the mask mode only selects a short fix-up of the score
// registers.  (Three specialised copies were 110 KB of SASS; the diagonal copy, entered once per
// item, ran ~4x slower than the common one — instruction-cache misses on a cold 16 KB body.)
template <bool DROP>
__device__ __forceinline__ void softmax_tile(int mode, uint32_t t_s, uint32_t t_p, uint32_t t_o,
                                             uint64_t* s_read, uint64_t* pv_done, bool first,
                                             uint32_t pv_parity, int kv0,
                                             int qrow, int T, const uint8_t* kmask,
                                             float scale_log2, float& m_ref, float& l_run, int wq,
                                             int lane, const DropoutParams& dp,
                                             uint64_t mask_base, const Params& prm, int tj,
                                             bool tracer) {
  uint32_t v[BKV / 32][32];
  // All four chunks are loaded unconditionally, also on the diagonal tile: a tcgen05.ld under a
  // (warp-uniform) branch made ptxas keep the row in local memory, and the diagonal tiles — 22% of
  // all tiles — then cost 2.2x a full tile.
#pragma unroll
  for (int c = 0; c < BKV / 32; ++c) {
    tmem_ld_32x32b_x32(t_s + c * 32, v[c]);
    tmem_ld_wait();  // one load in flight: four back to back were measured 3x slower in total
  }
  // S_t is in registers: the tensor core may overwrite it with the next tile's scores
  tc_fence_before_sync();
  if (lane == 0) mbar_arrive(s_read);
  if (tracer) stamp(prm, 3, tj, 0);

  // masked scores become -inf: they drop out of the max and exp2 turns them into exact zeros
  if (mode == kDiag) {
    // diagonal tile (kv0 == q0): column 32c + i is kept iff 32c + i <= 32 wq + lane.  Chunk c is
    // kept whole for c < wq and dropped whole for c > wq (warp-uniform), per element for c == wq.
#pragma unroll
    for (int c = 0; c < BKV / 32; ++c) {
      const bool dead = c > wq, edge = c == wq;
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (dead || (edge && i > lane)) v[c][i] = 0xff800000u;
    }
  } else if (mode == kGeneral) {
    // key padding: one keep-bit per column, gathered with a rolled loop (no indexing of v, which
    // must stay in registers), then applied with static indices
    uint32_t keep[BKV / 32];
#pragma unroll
    for (int c = 0; c < BKV / 32; ++c) {
      uint32_t bits = 0;
#pragma unroll 1
      for (int i = 0; i < 32; ++i) {
        const int col = kv0 + c * 32 + i;
        bool ok = col <= qrow;
        if (kmask != nullptr && col < T) ok = ok && (kmask[col] != 0);
        bits |= uint32_t(ok) << i;
      }
      keep[c] = bits;
    }
#pragma unroll
    for (int c = 0; c < BKV / 32; ++c) {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (((keep[c] >> i) & 1u) == 0) v[c][i] = 0xff800000u;
    }
  }

  float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
  for (int c = 0; c < BKV / 32; ++c) {
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
#pragma unroll
      for (int u = 0; u < 4; ++u) mx[u] = fmaxf(mx[u], __uint_as_float(v[c][i + u]));
    }
  }
  const float m_tile = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) * scale_log2;

  // Raise the reference maximum only when it is far behind; the decision is warp-uniform because
  // the rescale of O contains .sync.aligned TMEM accesses.
  const bool grow = m_tile > m_ref + kRescaleThreshold;  // also true for the first finite maximum
  // PV_t(j-1) must have retired before O_t is rescaled or P_t is overwritten.  It has, long ago;
  // the check is deferred to the first P store so that its latency hides under the exponentials.
  bool pv_checked = first;
  if (__any_sync(0xffffffffu, grow)) {
    const float m_new = fmaxf(m_ref, m_tile);
    const float alpha = (m_new == -INFINITY) ? 1.f : fast_exp2(m_ref - m_new);  // -inf -> 0
    l_run *= alpha;
    if (!first) {
      mbar_wait(pv_done, pv_parity);
      tc_fence_after_sync();
      pv_checked = true;
#pragma unroll
      for (int c = 0; c < HD / 32; ++c) {
        uint32_t o[32];
        tmem_ld_32x32b_x32(t_o + c * 32, o);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
        tmem_st_32x32b_x32(t_o + c * 32, o);
      }
    }
    m_ref = m_new;
  }
  const float m_use = (m_ref == -INFINITY) ? 0.f : m_ref;
  if (tracer) stamp(prm, 3, tj, 1);

  const f32x2 sc2 = splat2(scale_log2), nm2 = splat2(-m_use);
  f32x2 sum2 = splat2(0.f);
#pragma unroll
  for (int c = 0; c < BKV / 32; ++c) {
    uint32_t pk[16];
    {
      Rand8 rnd{};
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        const f32x2 x2 =
            fma2(pack2(__uint_as_float(v[c][i]), __uint_as_float(v[c][i + 1])), sc2, nm2);
        float e0, e1;
        if (((i >> 1) & 3) < kPolyOf4) {
          exp2_poly_pair(x2, e0, e1);
        } else {
          float x0, x1;
          unpack2(x2, x0, x1);
          e0 = fast_exp2(x0);
          e1 = fast_exp2(x1);
        }
        sum2 = add2(sum2, pack2(e0, e1));
        if (DROP) {
          if ((i & 7) == 0) rnd = philox_group(dp, (mask_base + uint64_t(c * 32 + i)) >> 3);
          e0 *= keep_factor(dp, rnd, i & 7);
          e1 *= keep_factor(dp, rnd, (i & 7) + 1);
        }
        pk[i >> 1] = pack_bf16x2(e0, e1);
      }
    }
    if (c == 0 && !pv_checked) {
      mbar_wait(pv_done, pv_parity);
      tc_fence_after_sync();
    }
    tmem_st_32x32b_x16(t_p + c * 16, pk);
  }
  float s0, s1;
  unpack2(sum2, s0, s1);
  l_run += s0 + s1;
  if (tracer) stamp(prm, 3, tj, 2);
  tmem_st_wait();
  if (tracer) stamp(prm, 3, tj, 3);
}

// One unit of work: a 256-row query block of one (batch, head).
struct Item {
  int b, h, qb, q0, n0, n1, n_max;
};
// (Cutting the list into groups of (b, h) pairs whose K / V stay in L2 between query blocks, as
// the backward kernel does for Q / dO / dQ, was measured: the kernel alone 126 against 119 us
// (the rounds lose their balance) and no change in the training step.)
__device__ __forceinline__ Item make_item(const Params& p, int idx) {
  Item it;
  const int bh_count = p.B * p.H;
  it.qb = p.nqb - 1 - idx / bh_count;  // heaviest query blocks first
  const int bh = idx % bh_count;
  it.b = bh / p.H;
  it.h = bh % p.H;
  it.q0 = it.qb * 2 * BQ;
  // key tiles of query tile t: 0 .. (2 qb + t); a tile that starts past T is inactive
  it.n0 = 2 * it.qb + 1;
  it.n1 = (it.q0 + BQ < p.T) ? 2 * it.qb + 2 : 0;
  it.n_max = it.n1 > it.n0 ? it.n1 : it.n0;
  return it;
}

// Persistent: one CTA per SM walks items blockIdx.x, + gridDim.x, ... of the heaviest-first list.
// Every role keeps RUNNING counters for its barriers (a barrier completes once per use, whatever
// item the use belongs to), so nothing is re-initialised between items and the roles drift apart
// freely: while the softmax groups normalise and store item i, the TMA warp is already loading
// Q/K/V of item i+1 and the tensor core computes its first score tiles.  (With one CTA per item
// the prologue, the epilogue and the launch gap were ~40% of the kernel: 2430 cycles per tile
// averaged over the kernel against 1500 in steady state.)
template <bool DROP>
__global__ void __launch_bounds__(THREADS, 1)
fwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmO,
           const Params p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* q_full = bars + 0;                 // [2]
  uint64_t* q_free = bars + 2;                 // [2]  the item's last S_t product has read Q_t
  uint64_t* k_full = bars + 4;                 // [NST]
  uint64_t* k_empty = k_full + NST;            // [NST]
  uint64_t* v_full = k_empty + NST;            // [NST]
  uint64_t* v_empty = v_full + NST;            // [NST]
  uint64_t* s_full = v_empty + NST;            // [2]
  uint64_t* p_full = s_full + 2;               // [2]
  uint64_t* pv_done = p_full + 2;              // [2]
  uint64_t* s_read = pv_done + 2;              // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_read + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int d = p.H * HD;
  const int n_items = p.nqb * p.B * p.H;

  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmO);
    for (int t = 0; t < 2; ++t) {
      mbar_init(&q_full[t], 1);
      mbar_init(&q_free[t], 1);
      mbar_init(&s_full[t], 1);
      mbar_init(&p_full[t], 4);
      mbar_init(&pv_done[t], 1);
      mbar_init(&s_read[t], 4);
    }
    for (int s = 0; s < NST; ++s) {
      mbar_init(&k_full[s], 1);
      mbar_init(&k_empty[s], 1);
      mbar_init(&v_full[s], 1);
      mbar_init(&v_empty[s], 1);
    }
    fence_barrier_init();
  }
  if (warp == 9) {
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  pdl_enter();  // qkv comes from the previous kernel of the stream
  const uint32_t tmem = *tmem_slot;

  if (warp >= 8) {
    reg_dealloc<64>();
    if (warp == 8 && lane == 0) {
      // ------------------------------- TMA producer -------------------------------------------
      uint32_t kv = 0;            // K/V tiles requested so far (ring position)
      uint32_t q_uses[2] = {0, 0};  // items in which tile t was active so far
      for (int idx = blockIdx.x; idx < n_items; idx += gridDim.x) {
        const Item it = make_item(p, idx);
        for (int t = 0; t < 2; ++t) {
          if ((t == 0 ? it.n0 : it.n1) == 0) continue;
          // the previous item's last score product must have read this tile
          if (q_uses[t] > 0) mbar_wait(&q_free[t], (q_uses[t] - 1) & 1u);
          ++q_uses[t];
          mbar_expect_tx(&q_full[t], TILE_BYTES);
          tma_load_3d(smem + OFF_Q + t * TILE_BYTES, &tmQKV, &q_full[t], it.h * HD,
                      it.q0 + t * BQ, it.b);
        }
        for (int j = 0; j < it.n_max; ++j, ++kv) {
          const int s = int(kv % NST);
          const uint32_t ph = (kv / NST) & 1u;
          mbar_wait(&k_empty[s], ph ^ 1u);
          mbar_expect_tx(&k_full[s], TILE_BYTES);
          tma_load_3d(smem + OFF_K + s * TILE_BYTES, &tmQKV, &k_full[s], d + it.h * HD, j * BKV,
                      it.b, kEvictLast);
          mbar_wait(&v_empty[s], ph ^ 1u);
          mbar_expect_tx(&v_full[s], TILE_BYTES);
          tma_load_3d(smem + OFF_V + s * TILE_BYTES, &tmQKV, &v_full[s], 2 * d + it.h * HD,
                      j * BKV, it.b, kEvictLast);
        }
      }
    } else if (warp == 9) {
      // ------------------------------- MMA issuer ---------------------------------------------
      constexpr uint32_t idesc_s = umma_idesc_bf16(BQ, BKV, false, false);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(BQ, HD, false, true);
      const bool leader = elect_one();
      const uint32_t sQ0 = smem_u32(smem + OFF_Q), sK0 = smem_u32(smem + OFF_K),
                     sV0 = smem_u32(smem + OFF_V);
      uint32_t kv = 0;              // ring position of the current item's key tile 0
      uint32_t n_s[2] = {0, 0};     // S_t products issued so far
      uint32_t n_pv[2] = {0, 0};    // PV_t products issued so far
      uint32_t q_uses[2] = {0, 0};
      auto issue_s = [&](int t, uint32_t ring) {  // S_t = Q_t K^T, K in ring slot `ring`
        // S_t may only be overwritten once the softmax group has pulled the previous one out
        if (n_s[t] > 0) mbar_wait(&s_read[t], (n_s[t] - 1) & 1u);
        ++n_s[t];
        tc_fence_after_sync();
        const uint64_t dq = umma_smem_desc(sQ0 + t * TILE_BYTES, 0, 1024);
        const uint64_t dk = umma_smem_desc(sK0 + int(ring % NST) * TILE_BYTES, 0, 1024);
        if (leader) {
#pragma unroll
          for (int k = 0; k < HD / 16; ++k)  // +32 B per K step = +2 in the address field
            umma_ss(tmem + COL_S + t * BKV, dq + 2 * k, dk + 2 * k, idesc_s, k > 0);
          umma_commit(&s_full[t]);
        }
      };
      auto issue_pv = [&](int t, uint32_t ring, bool accumulate) {  // O_t (+)= P_t V
        mbar_wait(&p_full[t], n_pv[t] & 1u);  // P_t has been written
        ++n_pv[t];
        tc_fence_after_sync();
        const uint64_t dv = umma_smem_desc(sV0 + int(ring % NST) * TILE_BYTES, TILE_BYTES, 1024);
        if (leader) {
#pragma unroll
          for (int k = 0; k < BKV / 16; ++k)  // +2048 B per K step = +128 in the address field
            umma_ts(tmem + COL_O + t * HD, tmem + COL_P + t * HD + k * 8, dv + 128 * k, idesc_pv,
                    accumulate || (k > 0));
          umma_commit(&pv_done[t]);
        }
      };
      for (int idx = blockIdx.x; idx < n_items; idx += gridDim.x) {
        const Item it = make_item(p, idx);
        const int nt[2] = {it.n0, it.n1};
        for (int t = 0; t < 2; ++t)
          if (nt[t] > 0) mbar_wait(&q_full[t], q_uses[t]++ & 1u);
        // S_t(j+1) is issued as soon as S_t(j) has left TMEM (inside issue_s), PV_t(j) once P_t(j)
        // is written: the next tile's scores are computed under the current tile's exponentials.
        mbar_wait(&k_full[kv % NST], (kv / NST) & 1u);
        for (int t = 0; t < 2; ++t) {
          if (nt[t] == 0) continue;
          issue_s(t, kv);
          if (nt[t] == 1 && leader) umma_commit(&q_free[t]);  // Q_t is not needed again
        }
        if (leader) umma_commit(&k_empty[kv % NST]);
        for (int j = 0; j < it.n_max; ++j) {
          const uint32_t rv = kv + j, rn = kv + j + 1;
          bool k_ready = false, v_ready = false;
#pragma unroll
          for (int t = 0; t < 2; ++t) {
            if (j + 1 >= nt[t]) continue;
            if (!k_ready) {
              mbar_wait(&k_full[rn % NST], (rn / NST) & 1u);
              k_ready = true;
            }
            issue_s(t, rn);
            if (j + 2 == nt[t] && leader) umma_commit(&q_free[t]);  // that was the last S_t
            if ((t == 1 || j + 1 >= nt[1]) && leader) umma_commit(&k_empty[rn % NST]);
          }
#pragma unroll
          for (int t = 0; t < 2; ++t) {
            if (j >= nt[t]) continue;
            if (!v_ready) {
              mbar_wait(&v_full[rv % NST], (rv / NST) & 1u);
              v_ready = true;
            }
            if (idx == int(blockIdx.x)) stamp(p, 2, j, 2 * t);
            issue_pv(t, rv, j > 0);
            if (idx == int(blockIdx.x)) stamp(p, 2, j, 2 * t + 1);
            if ((t == 1 || j >= nt[1]) && leader) umma_commit(&v_empty[rv % NST]);
          }
        }
        kv += uint32_t(it.n_max);
      }
    }
  } else {
    // ------------------------------- softmax groups ------------------------------------------
    reg_alloc<216>();
    const int t = warp >> 2;         // query tile of this group
    const int wq = warp & 3;         // warp inside the group = 32-row quarter of the tile
    const int r = wq * 32 + lane;    // row inside the tile == TMEM lane
    const uint32_t lane_base = uint32_t(wq * 32) << 16;
    const uint32_t t_s = tmem + lane_base + COL_S + t * BKV;
    const uint32_t t_o = tmem + lane_base + COL_O + t * HD;
    const uint32_t t_p = tmem + lane_base + COL_P + t * HD;
    uint32_t tiles = 0;  // score tiles of this group so far (all items)

    for (int idx = blockIdx.x; idx < n_items; idx += gridDim.x) {
      const Item it = make_item(p, idx);
      const int n_t = t == 0 ? it.n0 : it.n1;
      if (n_t == 0) continue;
      const int b = it.b, h = it.h;
      const int qi = 2 * it.qb + t;  // query tile index == index of its diagonal key tile
      const int qrow = it.q0 + t * BQ + r;
      const uint8_t* kmask = p.key_mask ? p.key_mask + size_t(b) * p.T : nullptr;
      const bool traced = idx == int(blockIdx.x);
      float m_ref = -INFINITY, l_run = 0.f;
      const int item_no = (idx - int(blockIdx.x)) / int(gridDim.x);
      if (p.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0 && item_no < 32)
        p.trace[4 * 32 * 4 + 2 * item_no] = clock64();

      for (int j = 0; j < n_t; ++j, ++tiles) {
        const int kv0 = j * BKV;
        if (r == 0 && traced) stamp(p, t, j, 0);
        mbar_wait(&s_full[t], tiles & 1u);
        tc_fence_after_sync();
        if (r == 0 && traced) stamp(p, t, j, 1);
        const int mode = kmask != nullptr ? kGeneral : (j == qi ? kDiag : kNoMask);
        const uint64_t mask_base =
            DROP ? (uint64_t((size_t(b) * p.H + h) * p.T + qrow) * uint64_t(p.t_pad) + uint64_t(kv0))
                 : 0;
        const uint32_t pv_parity = (tiles - 1) & 1u;  // PV of the previous tile of this group
        const bool tr = threadIdx.x == 0 && traced;
        softmax_tile<DROP>(mode, t_s, t_p, t_o, &s_read[t], &pv_done[t], j == 0, pv_parity, kv0, qrow,
                           p.T, kmask, p.scale_log2, m_ref, l_run, wq, lane, p.drop, mask_base, p, j,
                           tr);
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[t]);
        if (r == 0 && traced) stamp(p, t, j, 2);
      }

      // O_t / l -> bf16 -> this group's staging tile -> one TMA store per query tile.  (The wait
      // also guarantees that P_t is free for the next item's first tile.)  The staging tile is
      // separate from Q_t so that the next item's Q can be loaded — and its first score tile
      // computed — while this one is still being normalised and stored: with O staged in the Q
      // tile every item paid ~5600 cycles of exposed load + MMA latency after its epilogue.
      mbar_wait(&pv_done[t], (tiles - 1) & 1u);
      tc_fence_after_sync();
      if (r == 0) tma_store_wait_read<0>();  // the previous item's store has read the staging tile
      named_bar_sync(1 + t, 128);
      // a row whose keys were all excluded never raised m_ref: zeros and LSE = -inf (its l holds
      // only the clamped 2^-126 terms of the polynomial exponential)
      const bool dead_row = m_ref == -INFINITY;
      const float inv_l = (!dead_row && l_run > 0.f) ? 1.0f / l_run : 0.f;
      uint8_t* stage = smem + OFF_O + t * TILE_BYTES;
#pragma unroll
      for (int c = 0; c < HD / 32; ++c) {
        uint32_t o[32];
        tmem_ld_32x32b_x32(t_o + c * 32, o);
        tmem_ld_wait();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint32_t w[4];
#pragma unroll
          for (int u = 0; u < 4; ++u)
            w[u] = pack_bf16x2(__uint_as_float(o[q * 8 + 2 * u]) * inv_l,
                               __uint_as_float(o[q * 8 + 2 * u + 1]) * inv_l);
          *reinterpret_cast<uint4*>(stage + swizzle128(r * 128 + (c * 4 + q) * 16)) =
              make_uint4(w[0], w[1], w[2], w[3]);
        }
      }
      if (qrow < p.T) {
        p.lse[(size_t(b) * p.H + h) * p.T + qrow] =
            (!dead_row && l_run > 0.f) ? (m_ref + log2f(l_run)) * 0.6931471805599453f : -INFINITY;
      }
      tc_fence_before_sync();
      fence_proxy_async_smem();
      named_bar_sync(1 + t, 128);
      if (r == 0) {
        tma_store_3d(&tmO, stage, h * HD, it.q0 + t * BQ, b);
        tma_store_commit();
      }
      if (p.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0 && item_no < 32)
        p.trace[4 * 32 * 4 + 2 * item_no + 1] = clock64();
    }
  }

  if (warp < 8 && (threadIdx.x & 127) == 0) tma_store_wait_read<0>();  // staging outlives its reader
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem, TMEM_COLS);
}

}  // namespace attn

static int attention_fwd_launch(const void* qkv, void* out, float* lse, const uint8_t* key_mask, int B, int T,
                     int H, const DropoutParams& drop, float scale_log2, cudaStream_t stream) {
  namespace a2 = attn;
  const int d = H * a2::HD;
  CUtensorMap tQKV, tO;
  if (make_tmap_3d(&tQKV, static_cast<const __nv_bfloat16*>(qkv), B, T, 3 * d, 3 * d,
                   uint64_t(T) * 3 * d, 128, a2::HD, Swz::B128))
    return 1;
  if (make_tmap_3d(&tO, static_cast<__nv_bfloat16*>(out), B, T, d, d, uint64_t(T) * d, 128, a2::HD,
                   Swz::B128))
    return 1;
  a2::Params p;
  p.B = B; p.T = T; p.H = H;
  p.nqb = ceil_div(T, 2 * a2::BQ);
  p.scale_log2 = scale_log2;
  p.key_mask = key_mask;
  p.lse = lse;
  p.drop = drop;
  p.t_pad = (T + 7) & ~7;
  p.trace = nullptr;
  static const bool want_trace = std::getenv("LLMTRAIN_ATTN_TRACE") != nullptr;
  if (want_trace) {
    B200_CHECK_CUDA(cudaMalloc(&p.trace, (4 * 32 * 4 + 64) * sizeof(long long)));
    B200_CHECK_CUDA(cudaMemsetAsync(p.trace, 0, (4 * 32 * 4 + 64) * sizeof(long long), stream));
  }
  static bool configured = false;
  if (!configured) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(a2::fwd_kernel<false>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, a2::SMEM + 1024));
    B200_CHECK_CUDA(cudaFuncSetAttribute(a2::fwd_kernel<true>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, a2::SMEM + 1024));
    configured = true;
  }
  const int items = p.nqb * B * H;
  const int grid = items < sm_count() ? items : sm_count();  // persistent: one CTA per SM
  if (drop.threshold != 0)
    B200_LAUNCH((a2::fwd_kernel<true>), grid, a2::THREADS, a2::SMEM + 1024, stream, tQKV, tO, p);
  else
    B200_LAUNCH((a2::fwd_kernel<false>), grid, a2::THREADS, a2::SMEM + 1024, stream, tQKV, tO, p);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  if (p.trace != nullptr) {  // debug: per-tile timeline of the heaviest CTA, cycles since its first stamp
    static long long host[4 * 32 * 4 + 64];
    B200_CHECK_CUDA(cudaMemcpyAsync(host, p.trace, sizeof(host), cudaMemcpyDeviceToHost, stream));
    B200_CHECK_CUDA(cudaStreamSynchronize(stream));
    const long long t0 = host[1];
    for (int j = 0; j < 32 && host[(1 * 32 + j) * 4] != 0; ++j) {
      std::fprintf(stderr, "attn trace j=%2d", j);
      for (int role = 0; role < 4; ++role)
        for (int k = 0; k < 4; ++k)
          std::fprintf(stderr, "%s%7lld", k == 0 ? " |" : "",
                       host[(role * 32 + j) * 4 + k] == 0 ? -1LL : host[(role * 32 + j) * 4 + k] - t0);
      std::fprintf(stderr, "\n");
    }
    std::fprintf(stderr, "attn trace items (tiles, start, end) of CTA 0:");
    for (int k = 0; k < 32 && host[4 * 32 * 4 + 2 * k] != 0; ++k)
      std::fprintf(stderr, " [%lld %lld]", host[4 * 32 * 4 + 2 * k] - t0,
                   host[4 * 32 * 4 + 2 * k + 1] - t0);
    std::fprintf(stderr, "\n");
    cudaFree(p.trace);
  }
  return 0;
}

}  // namespace b200

using namespace b200;

// qkv: [B, T, 3*H*64] bf16 (q | k | v, head-major inside each third).  out: [B, T, H*64] bf16.
// lse: [B, H, T] fp32.  key_mask: [B, T] uint8 or NULL.
static int attention_fwd_impl(const void* qkv, void* out, float* lse, const uint8_t* key_mask,
                              int B, int T, int H, int head_dim, const DropoutParams& drop,
                              float softmax_scale, cudaStream_t stream) {
  B200_REQUIRE(head_dim == attn::HD, "attention: the tcgen05 kernel is specialised for head_dim 64");
  B200_REQUIRE(B > 0 && T > 0 && H > 0, "attention: empty problem");
  // softmax_scale > 0 overrides head_dim^-0.5 (heads narrower than 64 run zero-padded to 64 lanes
  // and keep the scale of their true width)
  const float scale_log2 = 1.4426950408889634f *
                           (softmax_scale > 0.f ? softmax_scale : 1.0f / std::sqrt(float(head_dim)));
  return attention_fwd_launch(qkv, out, lse, key_mask, B, T, H, drop, scale_log2, stream);
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
