// Causal flash attention backward, key-row orientation (head_dim 64, no probability dropout).
//
// Same mathematics as attention_bwd.cu's row-oriented kernel (autograd of
// `src/llmtrain/models/gpt.py:56-70`), but the score tile is computed TRANSPOSED so that the two
// products that accumulate over query tiles take their left operand from tensor memory:
//
//   S^T  = K_j Q_i^T   (128 kv x 128 q, K = 64)        dP^T = V_j dO_i^T   (128 kv x 128 q, K = 64)
//   dV  += P^T  dO_i   (A = P^T  in TMEM, B = dO_i)    dK  += dS^T Q_i    (A = dS^T in TMEM, B = Q_i)
//   dQ_i = dS   K_j    (A = dS^T tile in shared memory read MN-major, B = K_j)
//
// TMEM lane == key row, so the softmax-grad warps hold one key row each and walk the query
// columns; the per-query statistics (LSE, D) are staged in shared memory by their own small ring
// and read as broadcasts.  P^T is written back as bf16 into its own 64 TMEM columns, dS^T over the
// fp32 dP^T columns it came from; only dS^T also goes to shared memory (single buffer).
//
// Why.  (1) Operand traffic: the row-oriented kernel reads 224 KB of MMA operands from shared
// memory per (i, j) pair (P and dS twice each as 128-row left operands of N = 64 products) and
// writes 64 KB of P / dS; here a pair reads 144 KB and writes 32 KB, and a TS product of N = 64
// runs at the tensor-core floor (32 cycles) where the SS form is shared-memory bound (48;
// bench/mma_probe.cu).  (2) The shared memory that held P and the second dS buffer holds a second
// K / V stage and a third Q / dO stage: the next item's K / V arrive while the current item still
// runs and its first score tile is issued in the slot where the next query tile's would be, so an
// item boundary costs the epilogue only (it was 6 000 - 8 000 cycles of a 33 000-cycle item).
//
// Tensor-core queue per pair (issue order == execution order):
//     [P^T_g, dS^T_g handed over]  S^T_{g+1}                    -> s_full: the exponentials start
//                                  dK += dS^T_g Q_g, dP^T_{g+1} -> dp_full
//                                  dQ_g = dS_g K    (after dQ_{g-1} has been read out of TMEM;
//                                                    first in the queue for an item's last pair)
//                                  dV += P^T_g dO_g             -> releases P^T and the Q / dO stage
// dP^T_{g+1} overwrites the columns dS^T_g lives in; it is queued behind the product that reads
// them.  What sets the pace is the softmax-grad warps (~2 400 of ~3 900 cycles per pair) plus the
// hand-over -> S^T hop; the dQ drain of the previous pair runs underneath that hop.

#pragma once

#include "../common/host.cuh"
#include "../common/ptx.cuh"

namespace b200 {
namespace attn_bwd_t {

constexpr int BQ = 128;
constexpr int BKV = 128;
constexpr int HD = 64;
constexpr int TILE = 128 * HD * 2;  // 16 KB
constexpr int THREADS = 320;        // 8 softmax-grad warps + TMA warp + MMA warp
constexpr int TMEM_COLS = 512;
// S^T and dP^T fp32 (128 columns each); dS^T bf16 is written back over dP^T (at the start of each
// warp half's 64 fp32 columns), P^T bf16 has its own 64 columns so that the next S^T can be
// queued first; dV, dK, dQ fp32 accumulators, 64 columns each.
constexpr int COL_ST = 0, COL_DPT = 128, COL_PT = 256, COL_DV = 320, COL_DK = 384, COL_DQ = 448;
constexpr int NKV = 2;  // K / V stages (one per item in flight)
constexpr int NQ = 3;   // Q / dO ring depth
constexpr int NS = 2;   // statistics ring depth
constexpr int OFF_K = 0;                      // [NKV]
constexpr int OFF_V = OFF_K + NKV * TILE;     // [NKV]
constexpr int OFF_Q = OFF_V + NKV * TILE;     // [NQ]
constexpr int OFF_DO = OFF_Q + NQ * TILE;     // [NQ]
constexpr int DS_BYTES = 2 * TILE;            // one 128x128 bf16 tile = two 64-column regions
constexpr int OFF_DS = OFF_DO + NQ * TILE;    // single buffer
constexpr int OFF_DQS = OFF_DS + DS_BYTES;    // 32 KB: eight 4 KB per-warp staging slices
constexpr int STAT_BYTES = BQ * 8;            // (LSE * log2 e, D) per query row
constexpr int OFF_STAT = OFF_DQS + 2 * TILE;  // [NS]
constexpr int OFF_DB = OFF_STAT + NS * STAT_BYTES;  // [2][64] fp32: dV / dK column sums of an item
constexpr int OFF_BAR = OFF_DB + 512;
constexpr int SMEM = OFF_BAR + 256;
static_assert(SMEM <= 227 * 1024, "shared memory budget");

struct Params {
  int B, T, H;
  int t_stat;          // T rounded up to a multiple of 128: row pitch of `stats`
  int group;           // (b, h) pairs per scheduling group (see Schedule)
  float scale;         // head_dim^-0.5
  float scale_log2;    // scale * log2(e)
  const float2* stats; // [B, H, t_stat] (LSE * log2 e or +inf for an empty / padded row, D)
  const uint8_t* key_mask;
  float* dbias;        // optional [3*H*64]: += column sums of the k and v thirds of dqkv
  long long* trace;    // debug (LLMTRAIN_ATTN_TRACE): clock64 stamps of CTA 0, else nullptr
  int* turn;           // deterministic mode: [B, H, nq, 8] contribution counters (zeroed), else nullptr
};

// trace layout: [role 0 = softmax-grad warp 0, 1 = MMA, 2 = TMA][iteration < 16][4 stamps]
__device__ __forceinline__ void stamp(const Params& p, int role, int it, int k) {
  if (p.trace != nullptr && blockIdx.x == 0 && it < 16)
    p.trace[(role * 16 + it) * 4 + k] = clock64();
}

// One work item = one (key tile jt, batch b, head h).  Order of the item list: the (b, h) pairs
// are cut into groups of `group` consecutive pairs; inside a group the items go key tile by key
// tile (jt = 0, the one with the most query tiles, first).  A group's Q / dO tiles and fp32 dQ
// slices (nq x 64 KB per pair) add up to ~32 MB, so they are still in L2 when the group's next key
// tile comes round; with one global heaviest-first list (all pairs of jt = 0, then all of jt = 1,
// ...) 100 MB pass between two uses and every re-read and every dQ reduction went to DRAM (841 MB
// read, 426 MB written per call at B = 32, H = 12 against ~250 MB of inputs).  CTA c takes
// positions c, c + grid, c + 2 grid, ...: the weights form a sawtooth much longer than one round,
// every CTA samples it evenly, and the list still ends with the lightest items.
struct Schedule {
  int nq, bh, group, total, grid, cta;
  __device__ __forceinline__ Schedule(const Params& p, int nq_)
      : nq(nq_), bh(p.B * p.H), total(nq_ * p.B * p.H), grid(int(gridDim.x)), cta(int(blockIdx.x)) {
    group = max(1, min(bh, p.group));
  }
  // position of this CTA's k-th item, or -1 when it has no more
  __device__ __forceinline__ int position(int k) const {
    const int pos = cta + k * grid;
    return pos < total ? pos : -1;
  }
};
struct Item {
  int jt, b, h, n_iter;
};
__device__ __forceinline__ Item make_item(const Params& p, const Schedule& s, int pos) {
  const int per_group = s.group * s.nq;
  const int grp = pos / per_group;
  const int within = pos - grp * per_group;
  const int first_pair = grp * s.group;
  const int pairs = min(s.group, s.bh - first_pair);  // the last group may be short
  Item w;
  w.jt = within / pairs;
  const int pair = first_pair + (within - w.jt * pairs);
  w.b = pair / p.H;
  w.h = pair - w.b * p.H;
  w.n_iter = s.nq - w.jt;
  return w;
}

// Of every 4 consecutive column pairs, this many take their exponential on the FMA pipe (a cubic
// on the split exponent, see exp2_poly_pair); the rest use MUFU.EX2, 16 per clock per SM — one
// 128x128 tile needs 1024 MUFU cycles.
#ifndef B200_ATTN_BWD_POLY_OF4
#define B200_ATTN_BWD_POLY_OF4 0
#endif
constexpr int kPolyOf4 = B200_ATTN_BWD_POLY_OF4;

// p = 2^(s * scale_log2 - lse2) for two columns
__device__ __forceinline__ void exp2_pair(float s0, float s1, float scale_log2, float l0, float l1,
                                          bool poly, float& p0, float& p1) {
  if (poly) {
    exp2_poly_pair(fma2(pack2(s0, s1), splat2(scale_log2), pack2(-l0, -l1)), p0, p1);
  } else {
    p0 = fast_exp2(fmaf(s0, scale_log2, -l0));
    p1 = fast_exp2(fmaf(s1, scale_log2, -l1));
  }
}

// Exponentials of one 32-column chunk, in place: v[i] = S^T -> P^T (fp32 bits).
template <bool MASKED>
__device__ __forceinline__ void exp_chunk(uint32_t (&v)[32], const float4* sc, float scale_log2,
                                          int lead) {
#pragma unroll
  for (int i = 0; i < 32; i += 2) {
    const float4 a = sc[i >> 1];  // (lse2, D) of columns i, i+1: one broadcast load
    float p0, p1;
    exp2_pair(__uint_as_float(v[i]), __uint_as_float(v[i + 1]), scale_log2, a.x, a.z,
              ((i >> 1) & 3) < kPolyOf4, p0, p1);
    if (MASKED) {
      if (i < lead) p0 = 0.f;
      if (i + 1 < lead) p1 = 0.f;
    }
    v[i] = __float_as_uint(p0);
    v[i + 1] = __float_as_uint(p1);
  }
}
// pk = bf16(P^T), dk = bf16(P^T o (dP^T - D)) of one chunk.
__device__ __forceinline__ void ds_chunk(const uint32_t (&pv)[32], const uint32_t (&dv)[32],
                                         const float4* sc, uint32_t (&pk)[16], uint32_t (&dk)[16]) {
#pragma unroll
  for (int i = 0; i < 32; i += 2) {
    const float4 a = sc[i >> 1];
    const float p0 = __uint_as_float(pv[i]), p1 = __uint_as_float(pv[i + 1]);
    pk[i >> 1] = pack_bf16x2(p0, p1);
    dk[i >> 1] = pack_bf16x2(p0 * (__uint_as_float(dv[i]) - a.y),
                             p1 * (__uint_as_float(dv[i + 1]) - a.w));
  }
}

// One pair for one thread, as a hand-scheduled sequence: the exponentials (the MUFU-bound part)
// need only S^T, so they start when S^T is complete and run while the tensor core still computes
// dK and dP^T; every TMEM load after the first is in flight under arithmetic.  `lead_base` =
// number of leading columns of the 64 that are masked for this thread (causal edge on the diagonal
// tile; 64 for a padded / masked key row); MASKED is chosen warp-uniformly.
template <bool MASKED>
__device__ __forceinline__ void softmax_grad_t(uint32_t t_row, uint32_t t_pt, uint8_t* rd,
                                               const float4* st, int r, float scale_log2,
                                               int lead_base, uint64_t* dp_full, uint32_t parity,
                                               uint64_t* pt_free, uint64_t* ds_free,
                                               const Params& p, int g) {
  const int lead0 = min(lead_base, 32), lead1 = min(max(lead_base - 32, 0), 32);
  uint32_t s0[32], s1[32], dv[32], pk[16], dk[16];
  tmem_ld_32x32b_x32(t_row + COL_ST, s0);
  tmem_ld_32x32b_x32(t_row + COL_ST + 32, s1);
  tmem_ld_wait();
  if (threadIdx.x == 0) stamp(p, 2, g, 2);
  exp_chunk<MASKED>(s0, st, scale_log2, lead0);
  mbar_wait(dp_full, parity);  // dP^T (queued behind S^T and dK) is complete
  tc_fence_after_sync();
  tmem_ld_32x32b_x32(t_row + COL_DPT, dv);
  exp_chunk<MASKED>(s1, st + 16, scale_log2, lead1);
  tmem_ld_wait();
  ds_chunk(s0, dv, st, pk, dk);
  tmem_ld_32x32b_x32(t_row + COL_DPT + 32, dv);
  if (threadIdx.x == 0) stamp(p, 2, g, 3);
  // The previous pair's results may still be in use: P^T by dV += P^T dO (pt_free), the dS^T tile
  // in shared memory by dQ = dS K (ds_free).  Both retired long ago; the waits are formalities.
  mbar_wait(pt_free, parity ^ 1u);
  mbar_wait(ds_free, parity ^ 1u);
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    if (c == 1) {
      tmem_ld_wait();
      ds_chunk(s1, dv, st + 16, pk, dk);
    }
    // dS^T: bf16 columns c*16..+15 of this half overwrite fp32 dP^T columns this thread has
    // already read (columns 32.. are still being loaded when chunk 0 is stored: disjoint)
    tmem_st_32x32b_x16(t_pt + c * 16, pk);
    tmem_st_32x32b_x16(t_row + COL_DPT + c * 16, dk);
#pragma unroll
    for (int q = 0; q < 4; ++q)
      *reinterpret_cast<uint4*>(rd + swizzle128(r * 128 + (c * 4 + q) * 16)) =
          make_uint4(dk[4 * q], dk[4 * q + 1], dk[4 * q + 2], dk[4 * q + 3]);
  }
}

// Persistent: grid = min(items, SMs), one CTA per SM; every barrier runs on a counter that keeps
// counting across items (g = (i, j) pairs, k = items).
__global__ void __launch_bounds__(THREADS, 1)
bwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
           const __grid_constant__ CUtensorMap tmDKV, const __grid_constant__ CUtensorMap tmDQ,
           const Params p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* kv_full = bars;             // [NKV]
  uint64_t* kv_empty = kv_full + NKV;   // [NKV] the item's last MMA has retired AND the 8 warps'
                                        //       dK / dV stores have left the stage (see epilogue)
  uint64_t* q_full = kv_empty + NKV;    // [NQ]  Q_i, dO_i have landed
  uint64_t* q_empty = q_full + NQ;      // [NQ]
  uint64_t* st_full = q_empty + NQ;     // [NS]  statistics of the pair have landed
  uint64_t* st_empty = st_full + NS;    // [NS]  8 warps: statistics have been read
  uint64_t* s_full = st_empty + NS;     // S^T of the next pair is in TMEM
  uint64_t* dp_full = s_full + 1;       // ... and dP^T (and dK of the current pair has retired)
  uint64_t* pds_full = dp_full + 1;     // 8 warps: P^T / dS^T are in TMEM, dS^T in shared memory
  uint64_t* dq_full = pds_full + 1;     // dQ_g is complete (and the dS^T tile no longer read)
  uint64_t* dq_read = dq_full + 1;      // 8 warps: dQ has left TMEM
  uint64_t* pt_free = dq_read + 1;      // dV += P^T dO of the pair has retired
  uint64_t* dkv_full = pt_free + 1;     // dK / dV of the item are complete
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(dkv_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = (p.T + BQ - 1) / BQ;
  const int d = p.H * HD;
  const Schedule sched(p, nq);

  if (warp == 8 && lane == 0) {
    if ((smem_u32(smem) & 1023u) != 0u) __trap();  // the swizzled tiles need 1 KB alignment
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmDO);
    tma_prefetch_desc(&tmDKV);
    tma_prefetch_desc(&tmDQ);
    for (int s = 0; s < NKV; ++s) {
      mbar_init(&kv_full[s], 1);
      mbar_init(&kv_empty[s], 1 + 8);
    }
    for (int s = 0; s < NQ; ++s) {
      mbar_init(&q_full[s], 1);
      mbar_init(&q_empty[s], 1);
    }
    for (int s = 0; s < NS; ++s) {
      mbar_init(&st_full[s], 1);
      mbar_init(&st_empty[s], 8);
    }
    mbar_init(s_full, 1);
    mbar_init(dp_full, 1);
    mbar_init(pds_full, 8);
    mbar_init(dq_full, 1);
    mbar_init(dq_read, 8);
    mbar_init(pt_free, 1);
    mbar_init(dkv_full, 1);
    fence_barrier_init();
  }
  if (threadIdx.x < 128) reinterpret_cast<float*>(smem + OFF_DB)[threadIdx.x] = 0.f;
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
      for (int k = 0, pos; (pos = sched.position(k)) >= 0; ++k) {
        const Item w = make_item(p, sched, pos);
        const int kvs = k % NKV;
        mbar_wait(&kv_empty[kvs], (uint32_t(k / NKV) & 1u) ^ 1u);
        mbar_expect_tx(&kv_full[kvs], 2 * TILE);
        tma_load_3d(smem + OFF_K + kvs * TILE, &tmQKV, &kv_full[kvs], d + w.h * HD, w.jt * BKV, w.b);
        tma_load_3d(smem + OFF_V + kvs * TILE, &tmQKV, &kv_full[kvs], 2 * d + w.h * HD, w.jt * BKV,
                    w.b);
        const float2* stat_row = p.stats + (size_t(w.b) * p.H + w.h) * p.t_stat;
        for (int it = 0; it < w.n_iter; ++it, ++g) {
          const int st = g % NQ, ss = g % NS;
          const int q0 = (w.jt + it) * BQ;
          stamp(p, 2, g, 0);
          mbar_wait(&q_empty[st], (uint32_t(g / NQ) & 1u) ^ 1u);
          stamp(p, 2, g, 1);
          mbar_expect_tx(&q_full[st], 2 * TILE);
          tma_load_3d(smem + OFF_Q + st * TILE, &tmQKV, &q_full[st], w.h * HD, q0, w.b);
          tma_load_3d(smem + OFF_DO + st * TILE, &tmDO, &q_full[st], w.h * HD, q0, w.b);
          mbar_wait(&st_empty[ss], (uint32_t(g / NS) & 1u) ^ 1u);
          mbar_expect_tx(&st_full[ss], STAT_BYTES);
          bulk_load_1d(smem + OFF_STAT + ss * STAT_BYTES, stat_row + q0, STAT_BYTES, &st_full[ss]);
        }
      }
    }
  } else if (warp == 9) {
    // ------------------------------- MMA issuer ---------------------------------------------
    // The whole warp walks the loop and waits on the barriers; one elected lane issues.  The
    // descriptors are built outside the elected region so that they live in uniform registers
    // and a product's instructions go out back to back (one 64-bit add each): with the loop
    // inside `if (lane == 0)` every tcgen05.mma was preceded by ~15 dependent instructions
    // (84 cycles measured) against a tensor-core floor of 32 cycles for a 128x64x16 product.
    constexpr uint32_t id_s = umma_idesc_bf16(128, 128, false, false);  // K Q^T, V dO^T
    constexpr uint32_t id_ts = umma_idesc_bf16(128, 64, false, true);   // P^T dO, dS^T Q (A in TMEM)
    constexpr uint32_t id_dq = umma_idesc_bf16(128, 64, true, true);    // dS K (A = dS^T, MN-major)
    const bool leader = elect_one();
    const uint32_t sK0 = smem_u32(smem + OFF_K), sV0 = smem_u32(smem + OFF_V),
                   sQ0 = smem_u32(smem + OFF_Q), sDO0 = smem_u32(smem + OFF_DO);
    // MN-major tiles: 16 rows = 2048 B per K step = +128 in the address field; the two
    // 64-element regions of a 128-wide operand are TILE apart.
    const uint64_t aDS = umma_smem_desc(smem_u32(smem + OFF_DS), TILE, 1024);
    int g = 0;
    {  // the very first score tile
      mbar_wait(&kv_full[0], 0);
      mbar_wait(&q_full[0], 0);
      tc_fence_after_sync();
      const uint64_t dK = umma_smem_desc(sK0, 0, 1024), dV = umma_smem_desc(sV0, 0, 1024);
      const uint64_t dQ = umma_smem_desc(sQ0, 0, 1024), dDO = umma_smem_desc(sDO0, 0, 1024);
      if (leader) {
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_ss(tmem + COL_ST, dK + 2 * k, dQ + 2 * k, id_s, k > 0);
        umma_commit(s_full);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_ss(tmem + COL_DPT, dV + 2 * k, dDO + 2 * k, id_s, k > 0);
        umma_commit(dp_full);
      }
    }
    for (int item = 0, pos; (pos = sched.position(item)) >= 0; ++item) {
      const int n_iter = make_item(p, sched, pos).n_iter;
      const int kvs = item % NKV;
      const bool more_items = sched.position(item + 1) >= 0;
      const uint64_t bK = umma_smem_desc(sK0 + kvs * TILE, TILE, 1024);
      for (int it = 0; it < n_iter; ++it, ++g) {
        const int st = g % NQ;
        const bool last = it + 1 == n_iter;
        const bool has_next = !last || more_items;
        const int nkvs = last ? (item + 1) % NKV : kvs;  // K / V stage of the next pair
        const int nst = (g + 1) % NQ;
        const uint64_t bDO = umma_smem_desc(sDO0 + st * TILE, TILE, 1024);
        const uint64_t bQ = umma_smem_desc(sQ0 + st * TILE, TILE, 1024);
        // K-major tiles of the next pair: +32 B per K step = +2 in the address field
        const uint64_t dK = umma_smem_desc(sK0 + nkvs * TILE, 0, 1024);
        const uint64_t dV = umma_smem_desc(sV0 + nkvs * TILE, 0, 1024);
        const uint64_t dQ = umma_smem_desc(sQ0 + nst * TILE, 0, 1024);
        const uint64_t dDO = umma_smem_desc(sDO0 + nst * TILE, 0, 1024);
        if (leader) stamp(p, 1, g, 1);
        // (landed long ago; waiting here keeps it out of the hand-over -> S^T path)
        if (!last) mbar_wait(&q_full[nst], uint32_t((g + 1) / NQ) & 1u);
        // P^T_g / dS^T_g are in place.  For it == 0 this also says that every warp has taken
        // the previous item's dK / dV out of TMEM (they do that before they arrive here).
        mbar_wait(pds_full, uint32_t(g) & 1u);
        if (leader) stamp(p, 1, g, 2);
        if (last) {
          // At the end of an item the warps need dQ_g before the next score tile (they drain it
          // and dK / dV before they start on the next item), so it goes first — and before this
          // warp waits for the next item's K / V, whose stage is only released by those warps.
          mbar_wait(dq_read, uint32_t(g - 1) & 1u);  // dQ_{g-1} has left TMEM
          tc_fence_after_sync();
          if (leader) {
#pragma unroll
            for (int k = 0; k < BKV / 16; ++k)
              umma_ss(tmem + COL_DQ, aDS + 128 * k, bK + 128 * k, id_dq, k > 0);
            umma_commit(dq_full);
          }
        }
        if (last && has_next) {
          mbar_wait(&kv_full[nkvs], uint32_t((item + 1) / NKV) & 1u);
          mbar_wait(&q_full[nst], uint32_t((g + 1) / NQ) & 1u);
        }
        tc_fence_after_sync();
        if (leader) {
          stamp(p, 1, g, 0);
          if (has_next) {
#pragma unroll
            for (int k = 0; k < HD / 16; ++k)
              umma_ss(tmem + COL_ST, dK + 2 * k, dQ + 2 * k, id_s, k > 0);
            umma_commit(s_full);
          }
          // dK += dS^T Q.  A: 128 kv lanes x 16 q per step = 8 TMEM columns; the bf16 halves sit
          // at the start of their warps' fp32 halves (columns 0 and 64 of dP^T).
#pragma unroll
          for (int k = 0; k < BQ / 16; ++k)
            umma_ts(tmem + COL_DK, tmem + COL_DPT + (k >> 2) * 64 + (k & 3) * 8, bQ + 128 * k, id_ts,
                    (it > 0) || (k > 0));
          if (has_next) {
#pragma unroll
            for (int k = 0; k < HD / 16; ++k)
              umma_ss(tmem + COL_DPT, dV + 2 * k, dDO + 2 * k, id_s, k > 0);
            umma_commit(dp_full);
          }
        }
        if (!last) {
          mbar_wait(dq_read, uint32_t(g - 1) & 1u);  // dQ_{g-1} has left TMEM
          tc_fence_after_sync();
          if (leader) {
            // dQ_g = dS K: A = the dS^T tile [kv][q] read MN-major, B = K_j MN-major.
#pragma unroll
            for (int k = 0; k < BKV / 16; ++k)
              umma_ss(tmem + COL_DQ, aDS + 128 * k, bK + 128 * k, id_dq, k > 0);
            umma_commit(dq_full);
          }
        }
        if (leader) {
          // dV += P^T dO
#pragma unroll
          for (int k = 0; k < BQ / 16; ++k)
            umma_ts(tmem + COL_DV, tmem + COL_PT + k * 8, bDO + 128 * k, id_ts, (it > 0) || (k > 0));
          umma_commit(pt_free);
          umma_commit(&q_empty[st]);
          if (last) {
            umma_commit(dkv_full);
            umma_commit(&kv_empty[kvs]);
          }
          stamp(p, 1, g, 3);
        }
      }
    }
  } else {
    // ------------------------------- softmax-grad / epilogue warps --------------------------
    // 8 warps: warp w owns TMEM lanes 32*(w%4)..+31 (key row r) and the 64-column half (w/4) of
    // every 128-column tile — which is exactly one swizzled smem region of dS^T.
    const int quarter = warp & 3, half = warp >> 2;
    const int r = quarter * 32 + lane;  // 0..127
    const uint32_t lane_base = uint32_t(quarter * 32) << 16;
    // Three per-warp 4 KB staging slices (32 rows of 128 bytes, swizzled), each handed to the TMA
    // unit by this warp's lane 0 and tracked in its bulk groups, so none needs a CTA-wide sync:
    //   stage     the warp's 32 x 32 fp32 slice of a dQ tile (TMA reduce-add), one pair late
    //   ds_slice  this warp's rows of its half of the dS^T tile: idle between an item's last
    //             dQ product and the next item's first softmax-grad, it carries the item's last
    //             dQ slice, so that slice does not wait for `stage` to be read out
    //   kv_slice  a piece of the finished item's K (warps 4-7) / V (warps 0-3) stage, which no
    //             product reads any more: carries dK / dV (32 rows x 64 bf16)
    uint8_t* stage = smem + OFF_DQS + half * TILE + quarter * 4096;
    uint8_t* ds_slice = smem + OFF_DS + half * TILE + quarter * 4096;
    int kv_pending = -1;  // K / V stage whose dK / dV store may still be reading shared memory

    int g = 0;
    for (int item = 0, pos; (pos = sched.position(item)) >= 0; ++item) {
      const Item w = make_item(p, sched, pos);
      const int jt = w.jt, b = w.b, h = w.h, n_iter = w.n_iter;
      const int kv0 = jt * BKV;
      const int kvs = item % NKV;
      bool row_dead = kv0 + r >= p.T;
      if (p.key_mask != nullptr && !row_dead) row_dead = p.key_mask[size_t(b) * p.T + kv0 + r] == 0;

      // dQ of pair gi (query tile qt): TMEM -> registers (x scale) -> a staging slice -> ONE TMA
      // reduce-add per warp into the fp32 dQ buffer.  Called one pair late, so the MMAs that
      // precede dq_full(gi) are never waited for.  Rows past T are clipped by the tensor map.
      // The TMEM load is issued by the caller (load_dq) so that, inside the loop, it is in flight
      // across the hand-over of P^T / dS^T.
      auto load_dq = [&](uint32_t (&v)[32]) {
        tmem_ld_32x32b_x32(tmem + COL_DQ + lane_base + half * 32, v);
      };
      auto drain_dq = [&](uint32_t (&v)[32], int qt, bool item_end) {
        tmem_ld_wait();
        tc_fence_before_sync();
        __syncwarp();
        uint8_t* dst = item_end ? ds_slice : stage;
        if (lane == 0) {
          mbar_arrive(dq_read);
          // `stage` must have been read out before it is rewritten; the same wait covers the
          // previous item's dK / dV store, whose K / V stage is then handed back to the TMA
          // warp.  (An item's last slice goes to ds_slice and waits only if a one-pair item has
          // left that store unreleased.)
          if (!item_end || kv_pending >= 0) {
            tma_store_wait_read<0>();
            if (kv_pending >= 0) mbar_arrive(&kv_empty[kv_pending]);
          }
        }
        kv_pending = -1;
        __syncwarp();
#pragma unroll
        for (int q = 0; q < 8; ++q)
          *reinterpret_cast<float4*>(dst + swizzle128(lane * 128 + q * 16)) =
              make_float4(__uint_as_float(v[4 * q]) * p.scale, __uint_as_float(v[4 * q + 1]) * p.scale,
                          __uint_as_float(v[4 * q + 2]) * p.scale,
                          __uint_as_float(v[4 * q + 3]) * p.scale);
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
          tma_reduce_add_3d(&tmDQ, dst, h * HD + half * 32, qt * BQ + quarter * 32, b);
          tma_store_commit();
          if (turn != nullptr) {
            tma_store_wait<0>();  // the reduction has been performed, not merely read from smem
            __threadfence();
            asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(turn), "r"(jt + 1) : "memory");
          }
        }
      };

      for (int it = 0; it < n_iter; ++it, ++g) {
        const int ss = g % NS;
        // Causal edge: on the diagonal tile query column c survives for key row r iff c >= r.
        int lead = 0;
        if (it == 0) lead = min(max(r - half * 64, 0), 64);
        if (row_dead) lead = 64;

        if (threadIdx.x == 0) stamp(p, 0, g, 0);
        if (it == 0 && item > 0) {
          // ds_slice carried the previous item's last dQ slice: all bulk groups but the newest
          // (the dK / dV store, which reads kv_slice) must have been read out.
          if (lane == 0) tma_store_wait_read<1>();
          __syncwarp();
        }
        mbar_wait(&st_full[ss], uint32_t(g / NS) & 1u);
        mbar_wait(s_full, uint32_t(g) & 1u);
        tc_fence_after_sync();
        if (threadIdx.x == 0) stamp(p, 0, g, 1);
        {
          const uint32_t t_row = tmem + lane_base + half * 64;
          const uint32_t t_pt = tmem + lane_base + COL_PT + half * 32;
          uint8_t* rd = smem + OFF_DS + half * TILE;
          const float4* st4 =
              reinterpret_cast<const float4*>(smem + OFF_STAT + ss * STAT_BYTES) + half * 32;
          // warp-uniform on purpose: the variants contain .sync.aligned TMEM instructions
          if (__any_sync(0xffffffffu, lead > 0))
            softmax_grad_t<true>(t_row, t_pt, rd, st4, r, p.scale_log2, lead, dp_full,
                                 uint32_t(g) & 1u, pt_free, dq_full, p, g);
          else
            softmax_grad_t<false>(t_row, t_pt, rd, st4, r, p.scale_log2, lead, dp_full,
                                  uint32_t(g) & 1u, pt_free, dq_full, p, g);
        }
        // dQ_{g-1} is complete (softmax_grad_t waited for it before it rewrote the dS^T tile).
        // Unconditional: for it == 0 the columns hold the previous item's drained dQ.
        uint32_t dqv[32];
        load_dq(dqv);
        tmem_st_wait();
        fence_proxy_async_smem();
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(pds_full);
          mbar_arrive(&st_empty[ss]);
        }
        if (threadIdx.x == 0) stamp(p, 0, g, 2);

        if (it > 0) {
          drain_dq(dqv, jt + it - 1, false);
        } else {
          tmem_ld_wait();
        }
        if (threadIdx.x == 0) stamp(p, 0, g, 3);
      }
      {
        mbar_wait(dq_full, uint32_t(g - 1) & 1u);
        tc_fence_after_sync();
        uint32_t dqv[32];
        load_dq(dqv);
        drain_dq(dqv, jt + n_iter - 1, true);
      }

      // dV (warps 0-3) / dK x scale (warps 4-7): 32 kv rows x 64 columns per warp, TMEM -> bf16
      // -> kv_slice -> one TMA store per warp.  The next item's first dV / dK products are issued
      // behind its first pds_full, at which this warp arrives only after this read.
      mbar_wait(dkv_full, uint32_t(item) & 1u);
      tc_fence_after_sync();
      uint8_t* kv_slice = smem + (half == 0 ? OFF_V : OFF_K) + kvs * TILE + quarter * 4096;
      const float out_scale = half == 0 ? 1.0f : p.scale;
      uint32_t packed[32];
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem + (half == 0 ? COL_DV : COL_DK) + lane_base + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int u = 0; u < 16; ++u)
          packed[c * 16 + u] = pack_bf16x2(__uint_as_float(v[2 * u]) * out_scale,
                                           __uint_as_float(v[2 * u + 1]) * out_scale);
      }
      tc_fence_before_sync();
#pragma unroll
      for (int q = 0; q < 8; ++q)
        *reinterpret_cast<uint4*>(kv_slice + swizzle128(lane * 128 + q * 16)) =
            make_uint4(packed[4 * q], packed[4 * q + 1], packed[4 * q + 2], packed[4 * q + 3]);
      fence_proxy_async_smem();
      __syncwarp();
      const int col0 = (half == 0 ? 2 * d : d) + h * HD;
      if (lane == 0) {
        tma_store_3d(&tmDKV, kv_slice, col0, kv0 + quarter * 32, b);
        tma_store_commit();
      }
      kv_pending = kvs;
      if (p.dbias != nullptr) {
        // qkv_proj's bias gradient is the column sum of dqkv: the k and v thirds come from the
        // bf16 tile just staged (lane l: columns 2l, 2l+1 over the warp's 32 rows; conflict-free
        // in the swizzled layout).  Rows past T hold exact zeros.  The four warps of a half meet
        // in a shared-memory accumulator so that an item costs 128 global atomics, not 512 (all
        // items of a layer hit the same 48 cache lines).  The q third comes from
        // dq_convert_kernel.
        float* acc = reinterpret_cast<float*>(smem + OFF_DB) + half * 64;
        float s0 = 0.f, s1 = 0.f;
#pragma unroll 1
        for (int r8 = 0; r8 < 32; r8 += 8) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t off = uint32_t(r8 + j) * 128u +
                                 ((uint32_t(lane >> 2) ^ uint32_t(j)) << 4) +
                                 (uint32_t(lane & 3) << 2);
            const float2 f = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(kv_slice + off));
            s0 += f.x;
            s1 += f.y;
          }
        }
        atomicAdd(acc + 2 * lane, s0);
        atomicAdd(acc + 2 * lane + 1, s1);
        if (half == 0)  // the four warps of this half (ids spelled out: a computed id makes the
          named_bar_sync(1, 128);  // compiler reserve all 16 hardware barriers)
        else
          named_bar_sync(2, 128);
        if (quarter == 0) {
          atomicAdd(p.dbias + col0 + 2 * lane, acc[2 * lane]);
          atomicAdd(p.dbias + col0 + 2 * lane + 1, acc[2 * lane + 1]);
          acc[2 * lane] = 0.f;  // for the next item (ordered by the barriers every pair passes)
          acc[2 * lane + 1] = 0.f;
        }
      }
    }
    if (lane == 0) tma_store_wait_read<0>();  // shared memory must outlive the last store's read
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem, TMEM_COLS);
}

// stats[b,h,t] = (LSE[b,h,t] * log2 e, D[b,h,t]) with D = sum_c dO[b,t,h*64+c] * O[b,t,h*64+c];
// an empty row (LSE = -inf) and the padding t in [T, t_stat) get (+inf, 0) so that their
// probabilities come out as exact zeros.  8 lanes x 16 B cover one head (64 bf16).
__global__ void stats_kernel(const __nv_bfloat16* __restrict__ dO, const __nv_bfloat16* __restrict__ O,
                             const float* __restrict__ lse, float2* __restrict__ stats, int B, int T,
                             int H, int t_stat) {
  pdl_enter();  // previous kernel finished, its writes are visible
  // grid = (ceil(t_stat * H * 8 / 256), B): 32-bit index arithmetic, one batch row per blockIdx.y
  const int gid = int(blockIdx.x * blockDim.x + threadIdx.x);
  const int pair = gid >> 3;  // (t < t_stat, head) index
  const int sub = gid & 7;
  const int total = t_stat * H;
  const int b = int(blockIdx.y);
  const int t = pair / H;
  const int h = pair - t * H;
  const bool real = pair < total && t < T;
  float v = 0.f;
  if (real) {
    const size_t off = ((size_t(b) * T + t) * H + h) * HD + sub * 8;
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
    float l2 = INFINITY;
    if (real) {
      const float l = lse[(size_t(b) * H + h) * T + t];
      if (l != -INFINITY) l2 = l * 1.4426950408889634f;
    }
    stats[(size_t(b) * H + h) * t_stat + t] = make_float2(l2, real ? v : 0.f);
  }
}

// The padding entries t in [T, t_stat) of every (b, h) row of `stats`: (+inf, 0).  Needed when the
// statistics come from out_proj's dgrad epilogue (llmtrain_b200_gemm_rowdot), which only writes
// real rows; launched only if T is not a multiple of 128.
__global__ void stats_pad_kernel(float2* __restrict__ stats, int BH, int T, int t_stat) {
  pdl_enter();
  const int pad = t_stat - T;
  const int i = int(blockIdx.x * blockDim.x + threadIdx.x);
  if (i < BH * pad) {
    const int bh = i / pad, t = T + (i - bh * pad);
    stats[size_t(bh) * t_stat + t] = make_float2(INFINITY, 0.f);
  }
}

}  // namespace attn_bwd_t
}  // namespace b200
