// GPT training runtime: owns the memory plan (flat parameter / gradient / activation arenas) and
// sequences the sm_100a kernels for a full forward, fused loss, and backward pass.
//
// This is the C++ replacement for what the reference delegates to torch autograd when
// `GPTAdapter.compute_loss` runs `GPT.forward` and `.backward()`
// (`src/llmtrain/models/gpt.py:167-184,214-271`, `src/llmtrain/training/trainer.py:385-387`):
// there is no graph, no per-op allocation and no Python in the loop — one call launches ~20
// kernels per layer on a caller-provided stream, every intermediate lives at a fixed offset of
// one activation arena sized for (max_batch x max_seq), and weight gradients accumulate in fp32
// straight into the flat gradient arena that the bucket engine reduces.
//
// Arena order (elements, every tensor 64-element aligned):
//   wpe | layer 0 .. L-1 {ln1.w ln1.b qkv.w qkv.b out.w out.b ln2.w ln2.b fc.w fc.b proj.w proj.b}
//   | ln_f.w ln_f.b | [token_embedding.w if untied] | <LM-head matrix>
// The LM-head matrix (token_embedding when tied) is LAST so that its bf16 shadow can carry zero
// rows up to the padded vocabulary (multiple of 128) without disturbing any other offset; backward
// finishes gradient regions from the end of the arena towards the front, which is the bucket
// order the comm engine wants.

#include <nvtx3/nvToolsExt.h>  // header-only NVTX3: no link dependency, no cost without a tool

#include <cmath>
#include <vector>

#include "../common/host.cuh"
#include "../common/philox.cuh"
#include "../../include/llmtrain_b200.h"

namespace b200 {
namespace engine {

constexpr size_t ALIGN_ELEMS = 64;
inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct LayerOffsets {
  size_t ln1_w, ln1_b, qkv_w, qkv_b, out_w, out_b, ln2_w, ln2_b, fc_w, fc_b, proj_w, proj_b;
};

struct LayerActs {
  size_t ln1, ln1_mean, ln1_rstd, qkv, atty, lse, resid1, ln2, ln2_mean, ln2_rstd, fch_pre,
      fch_act, resid2;
  // heads narrower than 64 only: zero-padded per-step copies of qkv_proj.weight (bf16 [3*dq, d]),
  // qkv_proj.bias (fp32 [3*dq]) and out_proj.weight (bf16 [d, dq]); see csrc/kernels/headpad.cu
  size_t wq_pad, bq_pad, wo_pad;
};

struct Gpt {
  int V = 0, Vp = 0, block = 0, d = 0, L = 0, H = 0, F = 0, tied = 1;
  int hd = 64;        // true head width (d / H)
  int dq = 0;         // width of the q / k / v / attention-output activations: H * 64
  bool padded = false;  // hd < 64: heads run zero-padded to 64 lanes
  int maxB = 0, maxT = 0;
  // parameter plan
  size_t wpe = 0, lnf_w = 0, lnf_b = 0, wte = 0, head = 0;
  std::vector<LayerOffsets> layer;
  size_t n_params = 0;   // elements in p32 / m / v / g32 (incl. alignment gaps)
  size_t n_shadow = 0;   // elements in the bf16 shadow (n_params + vocabulary padding rows)
  // activation plan (byte offsets)
  std::vector<LayerActs> acts;
  size_t a_x0 = 0, a_lnf = 0, a_lnf_mean = 0, a_lnf_rstd = 0, a_logits = 0, a_loss_rows = 0;
  size_t a_dresid = 0, a_dtmp = 0, a_dqkv = 0, a_dfc = 0, a_attn_scratch = 0, a_red_scratch = 0;
  size_t a_dmask = 0;  // residual-branch gradient with its dropout mask applied
  size_t a_datty = 0;  // gradient of the attention output, [M, dq]
  size_t a_t_dwq = 0, a_t_dbq = 0, a_t_dwo = 0;  // padded fp32 gradient temporaries (hd < 64)
  size_t attn_scratch_bytes = 0, red_scratch_bytes = 0, act_bytes = 0;
  // bound memory
  float *p32 = nullptr, *g32 = nullptr;
  __nv_bfloat16* p16 = nullptr;
  uint8_t* act = nullptr;
  // state of the last forward (needed by backward)
  int B = 0, T = 0;
  const int64_t* ids = nullptr;
  const uint8_t* key_mask = nullptr;
  bool have_forward = false;
  // dropout (`gpt.py:28,30,97,129`): probability and counters for the NEXT mode-2 forward and
  // its backward; drop_on records whether the last forward actually applied it.
  float drop_p = 0.f;
  uint64_t drop_seed = 0, drop_offset = 0;
  bool drop_on = false;
};

inline size_t take(size_t& cursor, size_t n) {
  const size_t at = cursor;
  cursor = align_up(cursor + n, ALIGN_ELEMS);
  return at;
}
inline size_t take_bytes(size_t& cursor, size_t n) {
  const size_t at = cursor;
  cursor = align_up(cursor + n, 256);
  return at;
}

void plan_params(Gpt& g) {
  size_t c = 0;
  const size_t d = g.d, F = g.F;
  g.wpe = take(c, size_t(g.block) * d);
  g.layer.resize(g.L);
  for (int l = 0; l < g.L; ++l) {
    LayerOffsets& o = g.layer[l];
    o.ln1_w = take(c, d);          o.ln1_b = take(c, d);
    o.qkv_w = take(c, 3 * d * d);  o.qkv_b = take(c, 3 * d);
    o.out_w = take(c, d * d);      o.out_b = take(c, d);
    o.ln2_w = take(c, d);          o.ln2_b = take(c, d);
    o.fc_w = take(c, F * d);       o.fc_b = take(c, F);
    o.proj_w = take(c, d * F);     o.proj_b = take(c, d);
  }
  g.lnf_w = take(c, d);
  g.lnf_b = take(c, d);
  if (g.tied) {
    g.wte = take(c, size_t(g.V) * d);
    g.head = g.wte;
  } else {
    g.wte = take(c, size_t(g.V) * d);
    g.head = take(c, size_t(g.V) * d);
  }
  g.n_params = c;
  g.n_shadow = align_up(g.head + size_t(g.Vp) * d, ALIGN_ELEMS);
  if (g.n_shadow < g.n_params) g.n_shadow = g.n_params;
}

void plan_acts(Gpt& g) {
  const size_t M = size_t(g.maxB) * g.maxT, d = g.d, F = g.F, dq = g.dq;
  size_t c = 0;
  g.a_x0 = take_bytes(c, M * d * 2);
  g.acts.resize(g.L);
  for (int l = 0; l < g.L; ++l) {
    LayerActs& a = g.acts[l];
    a.ln1 = take_bytes(c, M * d * 2);
    a.ln1_mean = take_bytes(c, M * 4);
    a.ln1_rstd = take_bytes(c, M * 4);
    a.qkv = take_bytes(c, M * 3 * dq * 2);
    a.atty = take_bytes(c, M * dq * 2);
    a.lse = take_bytes(c, size_t(g.maxB) * g.H * g.maxT * 4);
    a.resid1 = take_bytes(c, M * d * 2);
    a.ln2 = take_bytes(c, M * d * 2);
    a.ln2_mean = take_bytes(c, M * 4);
    a.ln2_rstd = take_bytes(c, M * 4);
    a.fch_pre = take_bytes(c, M * F * 2);
    a.fch_act = take_bytes(c, M * F * 2);
    a.resid2 = take_bytes(c, M * d * 2);
    a.wq_pad = a.bq_pad = a.wo_pad = 0;
    if (g.padded) {
      a.wq_pad = take_bytes(c, 3 * dq * d * 2);
      a.bq_pad = take_bytes(c, 3 * dq * 4);
      a.wo_pad = take_bytes(c, d * dq * 2);
    }
  }
  g.a_lnf = take_bytes(c, M * d * 2);
  g.a_lnf_mean = take_bytes(c, M * 4);
  g.a_lnf_rstd = take_bytes(c, M * 4);
  g.a_logits = take_bytes(c, M * size_t(g.Vp) * 2);
  g.a_loss_rows = take_bytes(c, M * 4);
  g.a_dresid = take_bytes(c, M * d * 2);
  g.a_dtmp = take_bytes(c, M * d * 2);
  g.a_dqkv = take_bytes(c, M * 3 * dq * 2);
  g.a_datty = take_bytes(c, M * dq * 2);
  if (g.padded) {
    g.a_t_dwq = take_bytes(c, 3 * dq * d * 4);
    g.a_t_dbq = take_bytes(c, 3 * dq * 4);
    g.a_t_dwo = take_bytes(c, d * dq * 4);
  }
  g.a_dfc = take_bytes(c, M * F * 2);
  g.a_dmask = take_bytes(c, M * d * 2);
  g.attn_scratch_bytes =
      llmtrain_b200_attention_bwd_scratch_bytes(g.maxB, g.maxT, g.H, 64);
  g.a_attn_scratch = take_bytes(c, g.attn_scratch_bytes);
  size_t red = llmtrain_b200_layernorm_bwd_scratch_bytes(int(M), g.d);
  const int widths[4] = {g.d, 3 * g.dq, g.F, g.d};
  for (int w : widths) {
    const size_t need = llmtrain_b200_bias_grad_scratch_bytes(int(M), w);
    if (need > red) red = need;
  }
  {
    const size_t need = llmtrain_b200_embed_bwd_ordered_scratch_bytes(g.V);
    if (need > red) red = need;
  }
  g.red_scratch_bytes = red;
  g.a_red_scratch = take_bytes(c, red);
  g.act_bytes = c;
}

// ---- thin typed views -------------------------------------------------------------------------
struct View {
  Gpt& g;
  explicit View(Gpt& g_) : g(g_) {}
  __nv_bfloat16* bf(size_t off) const { return reinterpret_cast<__nv_bfloat16*>(g.act + off); }
  float* f32(size_t off) const { return reinterpret_cast<float*>(g.act + off); }
  const __nv_bfloat16* w(size_t off) const { return g.p16 + off; }
  const float* p(size_t off) const { return g.p32 + off; }
  float* grad(size_t off) const { return g.g32 + off; }
};

#define B200_TRY(call)      \
  do {                      \
    if ((call) != 0) return 1; \
  } while (0)

// Phase / layer ranges for Nsight timelines (the reference has no tracing at all).
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

inline int linear_fwd(int epi, int M, int N, int K, const void* x, const void* w, void* y, void* y2,
                      const float* bias, const void* r, cudaStream_t s) {
  return llmtrain_b200_gemm(0, epi, M, N, K, x, K, w, K, y, N, y2, N, bias, r, N,
                            reinterpret_cast<llm_stream_t>(s));
}
// dX[M,K] = dY[M,N] @ W[N,K]
inline int linear_dgrad(int epi, int M, int N, int K, const void* dy, const void* w, void* dx,
                        const void* r, cudaStream_t s) {
  return llmtrain_b200_gemm(1, epi, M, K, N, dy, N, w, K, dx, K, nullptr, 0, nullptr, r, K,
                            reinterpret_cast<llm_stream_t>(s));
}
// dW[N,K] += dY[M,N]^T @ X[M,K]
inline int linear_wgrad(int M, int N, int K, const void* dy, int ld_dy, const void* x, float* dw,
                        cudaStream_t s) {
  return llmtrain_b200_gemm(2, 4, N, K, M, dy, ld_dy, x, K, dw, K, nullptr, 0, nullptr, nullptr, 0,
                            reinterpret_cast<llm_stream_t>(s));
}

// dropout site ids (philox.cuh): 4 * layer + {0 attention probabilities, 1 attention residual
// branch, 2 MLP residual branch}; the embedding uses kDropSiteEmbedding.
inline uint32_t drop_site(int layer, int which) { return uint32_t(4 * layer + which); }

int forward(Gpt& g, const int64_t* ids, const uint8_t* key_mask, int B, int T, bool train,
            cudaStream_t s, bool chunked_head = false) {
  View v(g);
  NvtxRange phase("gpt.forward");
  const int M = B * T, d = g.d, F = g.F, dq = g.dq;
  const float sm_scale = 1.0f / std::sqrt(float(g.hd));  // of the TRUE head width
  auto st = reinterpret_cast<llm_stream_t>(s);
  const bool drop = train && g.drop_p > 0.f;
  const float dp = g.drop_p;
  const uint64_t seed = g.drop_seed, off = g.drop_offset;
  g.drop_on = drop;
  B200_TRY(llmtrain_b200_embed_fwd(ids, v.w(g.wte), v.w(g.wpe), v.bf(g.a_x0), B, T, d, g.V, st));
  if (drop)
    B200_TRY(llmtrain_b200_dropout(v.bf(g.a_x0), v.bf(g.a_x0), M, d, d, dp, seed, off,
                                   kDropSiteEmbedding, st));
  const __nv_bfloat16* x = v.bf(g.a_x0);
  for (int l = 0; l < g.L; ++l) {
    NvtxRange layer_range("layer.fwd");
    const LayerOffsets& o = g.layer[l];
    const LayerActs& a = g.acts[l];
    B200_TRY(llmtrain_b200_layernorm_fwd(x, v.p(o.ln1_w), v.p(o.ln1_b), v.bf(a.ln1),
                                         v.f32(a.ln1_mean), v.f32(a.ln1_rstd), M, d, st));
    // q | k | v with 64-wide heads ([M, 3*dq]); narrower heads use this step's zero-padded weights
    const __nv_bfloat16* wq = v.w(o.qkv_w);
    const float* bq = v.p(o.qkv_b);
    const __nv_bfloat16* wo = v.w(o.out_w);
    if (g.padded) {
      B200_TRY(llmtrain_b200_head_pad(v.w(o.qkv_w), v.bf(a.wq_pad), 1, 3 * g.H, g.hd, d, 2, st));
      B200_TRY(llmtrain_b200_head_pad(v.p(o.qkv_b), v.f32(a.bq_pad), 1, 3 * g.H, g.hd, 1, 4, st));
      B200_TRY(llmtrain_b200_head_pad(v.w(o.out_w), v.bf(a.wo_pad), d, g.H, g.hd, 1, 2, st));
      wq = v.bf(a.wq_pad);
      bq = v.f32(a.bq_pad);
      wo = v.bf(a.wo_pad);
    }
    B200_TRY(linear_fwd(0, M, 3 * dq, d, v.bf(a.ln1), wq, v.bf(a.qkv), nullptr, bq, nullptr, s));
    B200_TRY(llmtrain_b200_attention_fwd_fused(v.bf(a.qkv), v.bf(a.atty), v.f32(a.lse), key_mask, B,
                                               T, g.H, 64, sm_scale, drop ? dp : 0.f, seed, off,
                                               drop_site(l, 0), st));
    if (key_mask != nullptr) {
      // padded query rows take no attention update (`gpt.py:73-74`); dropout (if any) fused too
      B200_TRY(llmtrain_b200_gemm_resid_rowmask(M, d, dq, v.bf(a.atty), dq, wo, dq, v.bf(a.resid1),
                                                d, v.p(o.out_b), x, d, key_mask, drop ? dp : 0.f,
                                                seed, off, drop_site(l, 1), st));
    } else if (drop) {
      B200_TRY(llmtrain_b200_gemm_dropout_resid(M, d, dq, v.bf(a.atty), dq, wo, dq, v.bf(a.resid1),
                                                d, v.p(o.out_b), x, d, dp, seed, off,
                                                drop_site(l, 1), st));
    } else {
      B200_TRY(llmtrain_b200_gemm(0, 2, M, d, dq, v.bf(a.atty), dq, wo, dq, v.bf(a.resid1), d,
                                  nullptr, 0, v.p(o.out_b), x, d, st));
    }
    B200_TRY(llmtrain_b200_layernorm_fwd(v.bf(a.resid1), v.p(o.ln2_w), v.p(o.ln2_b), v.bf(a.ln2),
                                         v.f32(a.ln2_mean), v.f32(a.ln2_rstd), M, d, st));
    // fch_pre receives gelu'(pre-activation): all that backward needs from this GEMM
    B200_TRY(linear_fwd(1, M, F, d, v.bf(a.ln2), v.w(o.fc_w), v.bf(a.fch_act), v.bf(a.fch_pre),
                        v.p(o.fc_b), nullptr, s));
    if (drop) {
      B200_TRY(llmtrain_b200_gemm_dropout_resid(M, d, F, v.bf(a.fch_act), F, v.w(o.proj_w), F,
                                                v.bf(a.resid2), d, v.p(o.proj_b), v.bf(a.resid1),
                                                d, dp, seed, off, drop_site(l, 2), st));
    } else {
      B200_TRY(linear_fwd(2, M, d, F, v.bf(a.fch_act), v.w(o.proj_w), v.bf(a.resid2), nullptr,
                          v.p(o.proj_b), v.bf(a.resid1), s));
    }
    x = v.bf(a.resid2);
  }
  B200_TRY(llmtrain_b200_layernorm_fwd(x, v.p(g.lnf_w), v.p(g.lnf_b), v.bf(g.a_lnf),
                                       v.f32(g.a_lnf_mean), v.f32(g.a_lnf_rstd), M, d, st));
  // logits[M, Vp] = lnf @ head^T   (pad rows of the shadow are zero -> pad logits are 0, ignored)
  // With a loss to follow, the head runs in row chunks interleaved with the loss kernel instead
  // (head_and_loss below).
  if (!chunked_head)
    B200_TRY(llmtrain_b200_gemm(0, 0, M, g.Vp, d, v.bf(g.a_lnf), d, v.w(g.head), d,
                                v.bf(g.a_logits), g.Vp, nullptr, 0, nullptr, nullptr, 0, st));
  g.B = B; g.T = T; g.ids = ids; g.key_mask = key_mask;
  g.have_forward = true;
  return 0;
}

// LM head + loss, optionally in row chunks (reference
// `src/llmtrain/models/gpt.py:143,184,256-269`).  The logits of all M rows are 3.3 GB at the
// bench shape, so "GEMM writes logits, loss kernel reads them back and rewrites them as dlogits"
// pays an HBM write, a read and another write per logit; per chunk of rows small enough to stay in
// the 126 MB L2 the loss kernel could find the GEMM's output still on chip.  MEASURED on B200
// (GPT-2 124M, mb 32, whole step): one shot 33.40 ms, chunks of 2048 / 1024 / 768 / 512 rows
// 33.47 / 33.51 / 33.84 / 34.12 ms — the wave quantisation of the small GEMMs (a 512-row chunk is
// 394 pair tiles on 74 pairs) and of the 148-CTA loss kernel costs more than the saved traffic, so
// one shot stays the default and LLMTRAIN_HEAD_CHUNK_ROWS=<rows> is the experiment knob.
// (dlogits must exist in full for the weight-gradient GEMM, whose reduction runs over ALL rows;
// recomputing them instead costs a fourth vocabulary-sized GEMM — 2.6 TFLOP, ~2 ms — against
// ~1.3 ms of loss-kernel traffic: profiles/README.md, "LM head".)
inline int head_chunk_rows(const Gpt& g, int M) {
  static const int forced = [] {
    const char* e = std::getenv("LLMTRAIN_HEAD_CHUNK_ROWS");
    return e != nullptr ? std::atoi(e) : 0;
  }();
  (void)g;
  if (forced <= 0) return M;
  int rows = forced / 256 * 256;  // whole CTA-pair row tiles
  if (rows < 256) rows = 256;
  return rows < M ? rows : M;
}

int head_and_loss(Gpt& g, const int64_t* labels, const uint8_t* key_mask, int M, float* loss_acc,
                  float loss_scale, float grad_scale, int write_grad, llm_stream_t st) {
  View v(g);
  const int d = g.d;
  const int chunk = head_chunk_rows(g, M);
  for (int r0 = 0; r0 < M; r0 += chunk) {
    const int rows = M - r0 < chunk ? M - r0 : chunk;
    __nv_bfloat16* logits = v.bf(g.a_logits) + size_t(r0) * g.Vp;
    B200_TRY(llmtrain_b200_gemm(0, 0, rows, g.Vp, d, v.bf(g.a_lnf) + size_t(r0) * d, d,
                                v.w(g.head), d, logits, g.Vp, nullptr, 0, nullptr, nullptr, 0, st));
    B200_TRY(llmtrain_b200_cross_entropy(logits, labels + r0,
                                         key_mask != nullptr ? key_mask + r0 : nullptr,
                                         v.f32(g.a_loss_rows) + r0, loss_acc, rows, g.V, g.Vp,
                                         loss_scale, grad_scale, write_grad, st));
  }
  return 0;
}

// Called right after the kernels that FINALISE the gradient range [lo, lo+len) of the arena have
// been enqueued on `s` (ranges arrive back-to-front, the embedding tables last).
using BucketHook = int (*)(void* ctx, size_t lo, size_t len);

int backward(Gpt& g, cudaStream_t s, BucketHook hook = nullptr, void* hook_ctx = nullptr) {
  B200_REQUIRE(g.have_forward, "gpt_backward: no forward pass to differentiate");
  View v(g);
  NvtxRange phase("gpt.backward");
  const int B = g.B, T = g.T, M = B * T, d = g.d, F = g.F, dq = g.dq;
  const float sm_scale = 1.0f / std::sqrt(float(g.hd));
  auto st = reinterpret_cast<llm_stream_t>(s);
  void* red = g.act + g.a_red_scratch;
  const size_t red_bytes = g.red_scratch_bytes;
  __nv_bfloat16* dlogits = v.bf(g.a_logits);
  __nv_bfloat16* dtmp = v.bf(g.a_dtmp);
  __nv_bfloat16* dresid = v.bf(g.a_dresid);
  __nv_bfloat16* dqkv = v.bf(g.a_dqkv);
  __nv_bfloat16* dfc = v.bf(g.a_dfc);
  __nv_bfloat16* dmask = v.bf(g.a_dmask);
  __nv_bfloat16* datty = v.bf(g.a_datty);
  // With dropout the gradient entering a residual-branch projection is dresid o mask: it is
  // materialised once (the GEMMs read it through TMA) and the projection's bias gradient is its
  // column sum, so the LayerNorm backwards stop emitting those sums.
  const bool drop = g.drop_on;
  const bool det = deterministic();  // fixed summation orders (llmtrain_b200_set_deterministic)
  const float dp = g.drop_p;
  const uint64_t seed = g.drop_seed, off = g.drop_offset;

  // LM head: dlnf = dlogits @ head ; dhead += dlogits^T @ lnf   (only the V real rows)
  B200_TRY(llmtrain_b200_gemm(1, 0, M, d, g.Vp, dlogits, g.Vp, v.w(g.head), d, dtmp, d, nullptr, 0,
                              nullptr, nullptr, 0, st));
  B200_TRY(llmtrain_b200_gemm(2, 4, g.V, d, M, dlogits, g.Vp, v.bf(g.a_lnf), d, v.grad(g.head), d,
                              nullptr, 0, nullptr, nullptr, 0, st));
  B200_CHECK_CUDA(cudaMemsetAsync(dresid, 0, size_t(M) * d * 2, s));
  const __nv_bfloat16* x_last = g.L > 0 ? v.bf(g.acts[g.L - 1].resid2) : v.bf(g.a_x0);
  // Every LayerNorm backward also emits the column sums of the residual gradient it has just
  // produced: that is the bias gradient of the projection whose output this residual value is.
  B200_TRY(llmtrain_b200_layernorm_bwd(dtmp, x_last, v.p(g.lnf_w), v.f32(g.a_lnf_mean),
                                       v.f32(g.a_lnf_rstd), dresid, v.grad(g.lnf_w),
                                       v.grad(g.lnf_b),
                                       (g.L > 0 && !drop) ? v.grad(g.layer[g.L - 1].proj_b)
                                                          : nullptr,
                                       red, red_bytes, M, d, st));

  for (int l = g.L - 1; l >= 0; --l) {
    NvtxRange layer_range("layer.bwd");
    const LayerOffsets& o = g.layer[l];
    const LayerActs& a = g.acts[l];
    const __nv_bfloat16* x_in = l > 0 ? v.bf(g.acts[l - 1].resid2) : v.bf(g.a_x0);
    // ---- MLP: resid2 = gelu(ln2 @ Wfc^T + bfc) @ Wproj^T + bproj + resid1
    const __nv_bfloat16* dproj = dresid;
    if (drop) {
      B200_TRY(llmtrain_b200_dropout(dresid, dmask, M, d, d, dp, seed, off, drop_site(l, 2), st));
      B200_TRY(llmtrain_b200_bias_grad(dmask, d, M, d, v.grad(o.proj_b), red, red_bytes, st));
      dproj = dmask;
    }
    B200_TRY(linear_wgrad(M, d, F, dproj, d, v.bf(a.fch_act), v.grad(o.proj_w), s));
    // dfc = (dproj @ Wproj) o gelu'(pre); its column sums (= fc's bias gradient) come out of the
    // same epilogue.
    if (det) {  // fixed-order bias gradient instead of the epilogue's atomics
      B200_TRY(llmtrain_b200_gemm(1, 3, M, F, d, dproj, d, v.w(o.proj_w), F, dfc, F, nullptr, 0,
                                  nullptr, v.bf(a.fch_pre), F, st));
      B200_TRY(llmtrain_b200_bias_grad(dfc, F, M, F, v.grad(o.fc_b), red, red_bytes, st));
    } else {
      B200_TRY(llmtrain_b200_gemm_colsum(1, 3, M, F, d, dproj, d, v.w(o.proj_w), F, dfc, F,
                                         nullptr, v.bf(a.fch_pre), F, v.grad(o.fc_b), st));
    }
    B200_TRY(linear_wgrad(M, F, d, dfc, F, v.bf(a.ln2), v.grad(o.fc_w), s));
    B200_TRY(linear_dgrad(0, M, F, d, dfc, v.w(o.fc_w), dtmp, nullptr, s));
    B200_TRY(llmtrain_b200_layernorm_bwd(dtmp, v.bf(a.resid1), v.p(o.ln2_w), v.f32(a.ln2_mean),
                                         v.f32(a.ln2_rstd), dresid, v.grad(o.ln2_w),
                                         v.grad(o.ln2_b),
                                         (drop || g.key_mask != nullptr) ? nullptr : v.grad(o.out_b), red,
                                         red_bytes, M, d, st));
    // ---- attention: resid1 = attn(ln1 @ Wqkv^T + bqkv) @ Wo^T + bo + x_in
    const __nv_bfloat16* dout = dresid;
    if (g.key_mask != nullptr) {  // no gradient enters the branch from padded query rows
      B200_TRY(llmtrain_b200_dropout_rowmask(dresid, dmask, M, d, d, g.key_mask, drop ? dp : 0.f,
                                             seed, off, drop_site(l, 1), st));
      B200_TRY(llmtrain_b200_bias_grad(dmask, d, M, d, v.grad(o.out_b), red, red_bytes, st));
      dout = dmask;
    } else if (drop) {
      B200_TRY(llmtrain_b200_dropout(dresid, dmask, M, d, d, dp, seed, off, drop_site(l, 1), st));
      B200_TRY(llmtrain_b200_bias_grad(dmask, d, M, d, v.grad(o.out_b), red, red_bytes, st));
      dout = dmask;
    }
    // With heads narrower than 64 the three attention-side parameter gradients are produced in
    // their padded shapes and folded back into the true-shape gradient arena.
    float* dwo = g.padded ? v.f32(g.a_t_dwo) : v.grad(o.out_w);
    float* dwq = g.padded ? v.f32(g.a_t_dwq) : v.grad(o.qkv_w);
    float* dbq = g.padded ? v.f32(g.a_t_dbq) : v.grad(o.qkv_b);
    const __nv_bfloat16* wq = g.padded ? v.bf(a.wq_pad) : v.w(o.qkv_w);
    const __nv_bfloat16* wo = g.padded ? v.bf(a.wo_pad) : v.w(o.out_w);
    if (g.padded) {
      B200_CHECK_CUDA(cudaMemsetAsync(dwo, 0, size_t(d) * dq * 4, s));
      B200_CHECK_CUDA(cudaMemsetAsync(dwq, 0, size_t(3) * dq * d * 4, s));
      B200_CHECK_CUDA(cudaMemsetAsync(dbq, 0, size_t(3) * dq * 4, s));
    }
    B200_TRY(linear_wgrad(M, d, dq, dout, d, v.bf(a.atty), dwo, s));
    // dqkv and, fused, qkv_proj's bias gradient (column sums of dqkv).  Without probability
    // dropout the attention kernel's per-row statistics (LSE, rowsum(dO o O)) come out of
    // out_proj's dgrad epilogue, which has dO in registers, instead of a pass over dO and O.
    static const bool fuse_stats = std::getenv("LLMTRAIN_ATTN_STATS_PASS") == nullptr;
    if (!drop && fuse_stats) {
      void* stats = nullptr;
      int t_stat = 0;
      B200_TRY(llmtrain_b200_attention_bwd_stats_buffer(g.act + g.a_attn_scratch,
                                                        g.attn_scratch_bytes, B, T, g.H, 64, &stats,
                                                        &t_stat));
      B200_TRY(llmtrain_b200_gemm_rowdot(M, dq, d, dout, d, wo, dq, datty, dq, v.bf(a.atty), dq,
                                         v.f32(a.lse), stats, T, g.H, t_stat, st));
      B200_TRY(llmtrain_b200_attention_bwd_prestats(
          v.bf(a.qkv), v.bf(a.atty), datty, v.f32(a.lse), g.key_mask, dqkv, det ? nullptr : dbq,
          g.act + g.a_attn_scratch, g.attn_scratch_bytes, B, T, g.H, 64, sm_scale, st));
    } else {
      B200_TRY(linear_dgrad(0, M, d, dq, dout, wo, datty, nullptr, s));
      B200_TRY(llmtrain_b200_attention_bwd_fused(
          v.bf(a.qkv), v.bf(a.atty), datty, v.f32(a.lse), g.key_mask, dqkv, det ? nullptr : dbq,
          g.act + g.a_attn_scratch, g.attn_scratch_bytes, B, T, g.H, 64, sm_scale,
          drop ? dp : 0.f, seed, off, drop_site(l, 0), st));
    }
    if (det) B200_TRY(llmtrain_b200_bias_grad(dqkv, 3 * dq, M, 3 * dq, dbq, red, red_bytes, st));
    B200_TRY(linear_wgrad(M, 3 * dq, d, dqkv, 3 * dq, v.bf(a.ln1), dwq, s));
    B200_TRY(linear_dgrad(0, M, 3 * dq, d, dqkv, wq, dtmp, nullptr, s));
    if (g.padded) {
      B200_TRY(llmtrain_b200_head_unpad_add(dwo, v.grad(o.out_w), d, g.H, g.hd, 1, st));
      B200_TRY(llmtrain_b200_head_unpad_add(dwq, v.grad(o.qkv_w), 1, 3 * g.H, g.hd, d, st));
      B200_TRY(llmtrain_b200_head_unpad_add(dbq, v.grad(o.qkv_b), 1, 3 * g.H, g.hd, 1, st));
    }
    B200_TRY(llmtrain_b200_layernorm_bwd(dtmp, x_in, v.p(o.ln1_w), v.f32(a.ln1_mean),
                                         v.f32(a.ln1_rstd), dresid, v.grad(o.ln1_w),
                                         v.grad(o.ln1_b),
                                         (l > 0 && !drop) ? v.grad(g.layer[l - 1].proj_b) : nullptr,
                                         red, red_bytes, M, d, st));
    if (hook != nullptr) {
      // layer l's parameters (+ ln_f for the last layer) are contiguous in the arena
      const size_t lo = o.ln1_w;
      const size_t hi = (l + 1 < g.L) ? g.layer[l + 1].ln1_w : (g.tied ? g.wte : g.wte);
      B200_TRY(hook(hook_ctx, lo, hi - lo));
    }
  }
  if (drop)
    B200_TRY(llmtrain_b200_dropout(dresid, dresid, M, d, d, dp, seed, off, kDropSiteEmbedding, st));
  if (det)
    B200_TRY(llmtrain_b200_embed_bwd_ordered(g.ids, dresid, v.grad(g.wte), v.grad(g.wpe), B, T, d,
                                             g.V, red, red_bytes, st));
  else
    B200_TRY(llmtrain_b200_embed_bwd(g.ids, dresid, v.grad(g.wte), v.grad(g.wpe), B, T, d, g.V, st));
  if (hook != nullptr) {
    const size_t first_layer = g.L > 0 ? g.layer[0].ln1_w : g.lnf_w;
    B200_TRY(hook(hook_ctx, 0, first_layer));              // position embedding (small, first)
    B200_TRY(hook(hook_ctx, g.wte, g.n_params - g.wte));  // token embedding (+ untied head)
  }
  g.have_forward = false;
  return 0;
}

}  // namespace engine
}  // namespace b200

// Backward with a host callback after each gradient bucket has been enqueued: hook(ctx, lo, len)
// is called once the kernels that finalise arena range [lo, lo+len) are on `stream` (ranges arrive
// back to front, the embedding tables last).  The NCCL fallback uses it to start one all-reduce
// per bucket on a side stream while backward continues (what the c10d Reducer does for the
// reference, `src/llmtrain/training/trainer.py:88-91,377-387`).
extern "C" int llmtrain_b200_gpt_backward_hooked(void* h, int (*hook)(void*, size_t, size_t),
                                                 void* ctx, llm_stream_t stream) {
  return b200::engine::backward(*static_cast<b200::engine::Gpt*>(h),
                                reinterpret_cast<cudaStream_t>(stream), hook, ctx);
}

// Internal (non-ABI) entry used by the communicator to interleave bucket reductions.
int b200_gpt_backward_with_hook(void* handle, cudaStream_t stream, b200::engine::BucketHook hook,
                                void* ctx) {
  return b200::engine::backward(*static_cast<b200::engine::Gpt*>(handle), stream, hook, ctx);
}
size_t b200_gpt_param_elems(void* handle) {
  return static_cast<b200::engine::Gpt*>(handle)->n_params;
}

namespace b200 {
namespace engine {

}  // namespace engine
}  // namespace b200

using namespace b200;
using b200::engine::Gpt;

extern "C" int llmtrain_b200_gpt_create(int vocab, int block_size, int d_model, int n_layers,
                                        int n_heads, int d_ff, int tied, int max_batch,
                                        int max_seq, void** handle) {
  B200_REQUIRE(handle != nullptr, "gpt_create: null handle");
  B200_REQUIRE(d_model % n_heads == 0 && d_model / n_heads <= 64 && (d_model / n_heads) % 8 == 0,
               "gpt_create: head_dim must be a multiple of 8 and at most 64 (narrower heads run "
               "zero-padded on the head_dim-64 attention kernels)");
  B200_REQUIRE(d_model % 8 == 0 && d_ff % 8 == 0 && d_model <= 2048,
               "gpt_create: d_model/d_ff must be multiples of 8 and d_model <= 2048");
  B200_REQUIRE(max_seq <= block_size && max_batch > 0 && max_seq > 0, "gpt_create: bad batch plan");
  auto* g = new Gpt();
  g->V = vocab;
  g->Vp = int(engine::align_up(size_t(vocab), 128));
  g->block = block_size; g->d = d_model; g->L = n_layers; g->H = n_heads; g->F = d_ff;
  g->tied = tied ? 1 : 0;
  g->hd = d_model / n_heads;
  g->dq = n_heads * 64;
  g->padded = g->hd < 64;
  g->maxB = max_batch; g->maxT = max_seq;
  engine::plan_params(*g);
  engine::plan_acts(*g);
  *handle = g;
  return 0;
}

extern "C" void llmtrain_b200_gpt_destroy(void* handle) { delete static_cast<Gpt*>(handle); }

extern "C" size_t llmtrain_b200_gpt_param_elems(void* h) { return static_cast<Gpt*>(h)->n_params; }
extern "C" size_t llmtrain_b200_gpt_shadow_elems(void* h) { return static_cast<Gpt*>(h)->n_shadow; }
extern "C" size_t llmtrain_b200_gpt_activation_bytes(void* h) { return static_cast<Gpt*>(h)->act_bytes; }
extern "C" int llmtrain_b200_gpt_padded_vocab(void* h) { return static_cast<Gpt*>(h)->Vp; }
extern "C" size_t llmtrain_b200_gpt_logits_offset(void* h) { return static_cast<Gpt*>(h)->a_logits; }
extern "C" size_t llmtrain_b200_gpt_loss_rows_offset(void* h) { return static_cast<Gpt*>(h)->a_loss_rows; }

// Offset (elements) of a named tensor inside the fp32 arenas.  layer < 0 for the globals
// (0 wpe, 1 ln_f.w, 2 ln_f.b, 3 token_embedding, 4 lm_head); otherwise which in [0,12) follows
// the LayerOffsets field order.
extern "C" long long llmtrain_b200_gpt_param_offset(void* h, int layer, int which) {
  Gpt& g = *static_cast<Gpt*>(h);
  if (layer < 0) {
    switch (which) {
      case 0: return (long long)g.wpe;
      case 1: return (long long)g.lnf_w;
      case 2: return (long long)g.lnf_b;
      case 3: return (long long)g.wte;
      case 4: return (long long)g.head;
      default: return -1;
    }
  }
  if (layer >= g.L || which < 0 || which >= 12) return -1;
  const size_t* fields = reinterpret_cast<const size_t*>(&g.layer[layer]);
  return (long long)fields[which];
}

extern "C" int llmtrain_b200_gpt_bind(void* h, float* p32, float* g32, void* p16, void* act,
                                      size_t act_bytes) {
  Gpt& g = *static_cast<Gpt*>(h);
  B200_REQUIRE(act_bytes >= g.act_bytes, "gpt_bind: activation arena too small");
  B200_REQUIRE(((reinterpret_cast<uintptr_t>(p32) | reinterpret_cast<uintptr_t>(g32) |
                 reinterpret_cast<uintptr_t>(p16) | reinterpret_cast<uintptr_t>(act)) & 255) == 0,
               "gpt_bind: arenas must be 256-byte aligned");
  g.p32 = p32; g.g32 = g32;
  g.p16 = static_cast<__nv_bfloat16*>(p16);
  g.act = static_cast<uint8_t*>(act);
  g.have_forward = false;
  return 0;
}

// bf16 shadow <- fp32 master (after init / checkpoint load); zero the vocabulary padding rows.
extern "C" int llmtrain_b200_gpt_refresh_shadow(void* h, llm_stream_t stream) {
  Gpt& g = *static_cast<Gpt*>(h);
  B200_REQUIRE(g.p32 != nullptr && g.p16 != nullptr, "gpt_refresh_shadow: arenas not bound");
  auto s = reinterpret_cast<cudaStream_t>(stream);
  if (llmtrain_b200_cast_f32_to_bf16(g.p32, g.p16, g.n_params, stream)) return 1;
  if (g.n_shadow > g.n_params) {
    B200_CHECK_CUDA(cudaMemsetAsync(g.p16 + g.n_params, 0, (g.n_shadow - g.n_params) * 2, s));
  }
  const size_t head_end = g.head + size_t(g.V) * g.d;
  if (g.n_params > head_end) {  // alignment gap after the head inside n_params
    B200_CHECK_CUDA(cudaMemsetAsync(g.p16 + head_end, 0, (g.n_params - head_end) * 2, s));
  }
  return 0;
}

// Forward pass.  mode 0: logits only (left in the arena as bf16 [B*T, Vp]).
// mode 1: + per-row loss, loss_acc[0] += loss_scale * sum(rows).   mode 2: as 1, and the logits
// are replaced by dlogits * grad_scale so that llmtrain_b200_gpt_backward can follow.
extern "C" int llmtrain_b200_gpt_forward(void* h, const int64_t* ids, const int64_t* labels,
                                         const uint8_t* key_mask, int B, int T, int mode,
                                         float* loss_acc, float loss_scale, float grad_scale,
                                         llm_stream_t stream) {
  Gpt& g = *static_cast<Gpt*>(h);
  B200_REQUIRE(g.act != nullptr, "gpt_forward: arenas not bound");
  B200_REQUIRE(B >= 1 && B <= g.maxB && T >= 1 && T <= g.maxT, "gpt_forward: batch exceeds plan");
  auto s = reinterpret_cast<cudaStream_t>(stream);
  B200_REQUIRE(mode == 0 || labels != nullptr, "gpt_forward: labels required for the loss");
  if (engine::forward(g, ids, key_mask, B, T, mode == 2, s, /*chunked_head=*/mode >= 1)) return 1;
  if (mode >= 1 && engine::head_and_loss(g, labels, key_mask, B * T, loss_acc, loss_scale,
                                         grad_scale, mode == 2 ? 1 : 0, stream))
    return 1;
  if (mode != 2) g.have_forward = false;
  return 0;
}

extern "C" int llmtrain_b200_gpt_backward(void* h, llm_stream_t stream) {
  return engine::backward(*static_cast<Gpt*>(h), reinterpret_cast<cudaStream_t>(stream));
}

// Dropout for the NEXT mode-2 forward and its backward (p = 0 turns it off; modes 0/1 never drop).
// `seed` identifies the run and rank, `offset` must change with every micro-batch.
extern "C" int llmtrain_b200_gpt_set_dropout(void* h, float p, uint64_t seed, uint64_t offset) {
  B200_REQUIRE(p >= 0.f && p < 1.f, "gpt_set_dropout: p must be in [0, 1)");
  Gpt& g = *static_cast<Gpt*>(h);
  g.drop_p = p;
  g.drop_seed = seed;
  g.drop_offset = offset;
  return 0;
}
