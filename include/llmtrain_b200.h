/*
 * llmtrain_b200 — C ABI of the sm_100a native library (libllmtrain_b200.so).
 *
 * Conventions
 *   - every function returns 0 on success and non-zero on failure; the failure text is
 *     returned by llmtrain_b200_last_error() (thread-local).
 *   - all data pointers are DEVICE pointers unless a parameter name starts with `h_`.
 *   - `stream` is a cudaStream_t (pass torch.cuda.current_stream().cuda_stream from Python).
 *   - activations / weights shadow are bf16, master weights / optimizer state / gradients fp32.
 *   - kernels are asynchronous on `stream`; nothing here synchronises the device unless noted.
 *
 * Each entry point cites the reference site it replaces (paths relative to /root/reference).
 */
#ifndef LLMTRAIN_B200_H_
#define LLMTRAIN_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* llm_stream_t;

const char* llmtrain_b200_last_error(void);
/* 1000*major + 10*minor of the library ABI. */
int llmtrain_b200_abi_version(void);
/* Kernels launched by this library so far in this process (monotonic). */
long long llmtrain_b200_launch_count(void);

/* Deterministic mode (process-wide, off by default).  The reference gets run-to-run equality of a
 * resumed and an uninterrupted run from torch.use_deterministic_algorithms on the CPU
 * (tests/test_checkpoint.py:58-62,301-320).  Here the fast kernels add floats in a
 * hardware-chosen order in five places; with the mode on they use a fixed order instead:
 * weight-gradient GEMMs run without split-K, attention backward adds the dQ contributions of the
 * key tiles in ascending order (turn counters), the fused bias-gradient column sums are replaced
 * by the two-stage reduction, and the embedding gradient is accumulated in position order.
 * Gradients are then bit-reproducible for a fixed (weights, batch, world size). */
void llmtrain_b200_set_deterministic(int on);
int llmtrain_b200_get_deterministic(void);

/* ---------------------------------------------------------------------------------------------
 * GEMM (tcgen05 + TMEM + TMA).  Replaces nn.Linear forward and its autograd dgrad/wgrad:
 * src/llmtrain/models/gpt.py:27,29,94,96,143.
 *   layout 0 (NT): C[M,N]  = A[M,K]   * B[N,K]^T
 *   layout 1 (NN): C[M,N]  = A[M,K]   * B[K,N]
 *   layout 2 (TN): C[M,N] += A[K,M]^T * B[K,N]      (C is fp32; split-K; TMA reduce-add)
 *   epilogue 0: C = acc (+bias)            1: x = acc+bias, C = gelu_erf(x), C2 = gelu_erf'(x)
 *            2: C = acc (+bias) + R        3: C = acc * R  (R = the C2 saved by 1)
 *            4: fp32 accumulate
 *   A, B, C(bf16), C2, R are bf16 with row pitches lda/ldb/ldc/ldc2/ldr in elements (each pitch
 *   a multiple of 8); bias is fp32 [N] or NULL.
 * ------------------------------------------------------------------------------------------- */
int llmtrain_b200_gemm(int layout, int epilogue, int M, int N, int K, const void* A, int lda,
                       const void* B, int ldb, void* C, int ldc, void* C2, int ldc2,
                       const float* bias, const void* R, int ldr, llm_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * LayerNorm (eps 1e-5).  Replaces nn.LayerNorm at src/llmtrain/models/gpt.py:86,93,142.
 *   fwd: y = (x-mean)*rstd*gamma+beta, x/y bf16 [rows,d], gamma/beta fp32; mean/rstd fp32 [rows]
 *        (may be NULL when no backward follows).
 *   bwd: dresid[rows,d] (bf16, in place) += LN'(dy);  dgamma/dbeta (fp32) += column sums;
 *        if dcolsum != NULL: dcolsum[d] (fp32) += column sums of the UPDATED dresid (= the bias
 *        gradient of the projection that produced this residual-stream value).
 *        Deterministic two-stage reduction through `scratch` (size from *_scratch_bytes).
 * ------------------------------------------------------------------------------------------- */
int llmtrain_b200_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y,
                                float* mean, float* rstd, int rows, int d, llm_stream_t stream);
size_t llmtrain_b200_layernorm_bwd_scratch_bytes(int rows, int d);
int llmtrain_b200_layernorm_bwd(const void* dy, const void* x, const float* gamma,
                                const float* mean, const float* rstd, void* dresid, float* dgamma,
                                float* dbeta, float* dcolsum, void* scratch, size_t scratch_bytes,
                                int rows, int d, llm_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Embedding.  Replaces src/llmtrain/models/gpt.py:176-179 and its backward.
 *   fwd: x[b*T+t,:] = wte[ids[b,t],:] + wpe[t,:]   (bf16 tables, bf16 out)
 *   bwd: dwte[ids,:] += dx (vector red.add), dwpe[t,:] += sum_b dx[b,t,:]   (fp32 arenas)
 * ------------------------------------------------------------------------------------------- */
int llmtrain_b200_embed_fwd(const int64_t* ids, const void* wte, const void* wpe, void* x, int B,
                            int T, int d, int vocab, llm_stream_t stream);
int llmtrain_b200_embed_bwd(const int64_t* ids, const void* dx, float* dwte, float* dwpe, int B,
                            int T, int d, int vocab, llm_stream_t stream);
/* Same gradients with a fixed summation order (positions ascending, one writer per vocabulary row);
 * used when llmtrain_b200_set_deterministic is on.  scratch: vocab ints. */
size_t llmtrain_b200_embed_bwd_ordered_scratch_bytes(int vocab);
int llmtrain_b200_embed_bwd_ordered(const int64_t* ids, const void* dx, float* dwte, float* dwpe,
                                    int B, int T, int d, int vocab, void* scratch,
                                    size_t scratch_bytes, llm_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Fused softmax cross-entropy.  Replaces F.cross_entropy(reduction="none") + (masked) mean at
 * src/llmtrain/models/gpt.py:256-269.  One read of the bf16 logits row [V valid of ldv columns];
 * loss_rows[r] = lse - logit[label] (0 for rows with row_valid==0); when write_grad != 0 the row
 * is overwritten in place with (softmax - onehot) * grad_scale (0 for invalid rows / pad columns).
 * If loss_acc != NULL: loss_acc[0] += loss_scale * sum(loss_rows)   (deterministic).
 * ------------------------------------------------------------------------------------------- */
int llmtrain_b200_cross_entropy(void* logits, const int64_t* labels, const uint8_t* row_valid,
                                float* loss_rows, float* loss_acc, int rows, int V, int ldv,
                                float loss_scale, float grad_scale, int write_grad,
                                llm_stream_t stream);

/* dbias[n] (fp32) += sum_m dy[m,n] (bf16, pitch ld).  Autograd of the nn.Linear biases. */
size_t llmtrain_b200_bias_grad_scratch_bytes(int M, int N);
int llmtrain_b200_bias_grad(const void* dy, int ld, int M, int N, float* dbias, void* scratch,
                            size_t scratch_bytes, llm_stream_t stream);

/* Same column sum for callers that cannot supply scratch because they cannot allocate device
 * memory themselves (a process whose pool already owns all of HBM, a latency-critical loop that
 * cannot take cudaMalloc's device-wide stall, a binding with no allocator).  This entry point has
 * no counterpart in the reference; it exists only for such hosts.  The partial sums go to a 1 MiB
 * array built into the library's device module (held by every GPU that has loaded the module,
 * until exit; with CUDA's lazy module loading that may only happen at the first call on a GPU).
 *   - refused with an error when colsum splits * N exceeds capacity_bytes()/4 partial sums;
 *   - every call on a GPU shares that array and nothing is cleared between calls (none of it needs
 *     to start at zero); calls whose kernels overlap on the GPU corrupt each other: keep them on
 *     one stream or order them with events;
 *   - the first call made on each GPU looks the array's address up and may make the driver load
 *     the module (taking the 1 MiB then): make that call during start-up, before the host's pool
 *     claims the device and before beginning any stream capture, or set
 *     CUDA_MODULE_LOADING=EAGER.
 * Performs no cudaMalloc/cudaFree and calls no allocating library routine. */
size_t llmtrain_b200_bias_grad_noalloc_capacity_bytes(void);
int llmtrain_b200_bias_grad_noalloc(const void* dy, int ld, int M, int N, float* dbias,
                                    llm_stream_t stream);

/* Device-resident synthetic next-token data: ids[r,t] = hash(seed, first_row + r, t) mod vocab,
 * labels[r,t] = ids[r,t+1] (labels may be NULL).  Capability of the reference's dummy_text
 * (src/llmtrain/data/dummy_text.py:33-51) at full sequence length, without host copies. */
int llmtrain_b200_synthetic_tokens(int64_t* ids, int64_t* labels, int rows, int T, int vocab,
                                   uint64_t seed, uint64_t first_row, llm_stream_t stream);

int llmtrain_b200_cast_f32_to_bf16(const float* src, void* dst, size_t n, llm_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Optimizer.  Replaces clip_grad_norm_ + AdamW.step + zero_grad
 * (src/llmtrain/training/trainer.py:362,390-394).
 *   grad_sumsq: out[0] = (accumulate ? out[0] : 0) + sum g^2  (two-stage, deterministic).
 *   adamw: g_eff = g * grad_scale * min(1, max_norm / (sqrt(sumsq)*grad_scale + 1e-6)); decoupled
 *          weight decay; bias-corrected moments (step counts from 1); optionally refreshes the
 *          bf16 shadow, writes the pre-clip norm to norm_out, and zeroes g.
 * ------------------------------------------------------------------------------------------- */
size_t llmtrain_b200_grad_sumsq_scratch_bytes(size_t n);
int llmtrain_b200_grad_sumsq(const float* g, size_t n, float* out, int accumulate, void* scratch,
                             size_t scratch_bytes, llm_stream_t stream);
int llmtrain_b200_adamw(float* p, float* m, float* v, float* g, void* p_bf16, size_t n,
                        const float* sumsq, float* norm_out, float lr, float beta1, float beta2,
                        float eps, float weight_decay, int step, float grad_scale, float max_norm,
                        int zero_grad, llm_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Causal flash attention (tcgen05, head_dim 64).  Replaces src/llmtrain/models/gpt.py:56-70.
 *   qkv [B,T,3*H*64] bf16 (q|k|v thirds, head-major), out [B,T,H*64] bf16, lse [B,H,T] fp32,
 *   key_mask [B,T] uint8 (1 = attend) or NULL.
 * ------------------------------------------------------------------------------------------- */
int llmtrain_b200_attention_fwd(const void* qkv, void* out, float* lse, const uint8_t* key_mask,
                                int B, int T, int H, int head_dim, llm_stream_t stream);

/* Backward of the above (autograd of gpt.py:56-70).  dout [B,T,H*64] bf16 -> dqkv [B,T,3*H*64]
 * bf16 (all three thirds fully overwritten).  `scratch` (256-byte aligned, size from
 * *_scratch_bytes) holds the fp32 dQ accumulator and the row statistics D = rowsum(dO o O). */
size_t llmtrain_b200_attention_bwd_scratch_bytes(int B, int T, int H, int head_dim);
int llmtrain_b200_attention_bwd(const void* qkv, const void* out, const void* dout,
                                const float* lse, const uint8_t* key_mask, void* dqkv,
                                void* scratch, size_t scratch_bytes, int B, int T, int H,
                                int head_dim, llm_stream_t stream);

/* As llmtrain_b200_gemm with epilogue 0 (plain) or 3 (x gelu'), plus colsum[n] += sum_m C[m, n]
 * over the bf16-rounded output: the bias gradient of the layer this GEMM back-propagates into
 * (replaces the reduction behind nn.Linear's bias.grad, src/llmtrain/models/gpt.py:27,94), fused
 * so no separate pass re-reads C.  fp32 atomics: summation order is not fixed. */
int llmtrain_b200_gemm_colsum(int layout, int epilogue, int M, int N, int K, const void* A, int lda,
                              const void* B, int ldb, void* C, int ldc, const float* bias,
                              const void* R, int ldr, float* colsum, llm_stream_t stream);

/* dO[M, H*64] = A[M,K] * B[K, H*64] (NN: out_proj's dgrad, src/llmtrain/models/gpt.py:29) and, from
 * the same epilogue, the attention backward's per-row statistics (rows are (b, t), M = B*T):
 * stats[(b*H + h) * t_stat + t] = (lse[b,h,t] * log2 e, sum_c bf16(dO)[b,t,h,c] * O[b,t,h,c]) as
 * float pairs.  `stats` / `t_stat` come from llmtrain_b200_attention_bwd_stats_buffer; follow with
 * llmtrain_b200_attention_bwd_prestats (same scratch), which then skips its pass over dO and O. */
int llmtrain_b200_gemm_rowdot(int M, int N, int K, const void* A, int lda, const void* B, int ldb,
                              void* dO, int ldc, const void* O, int ldo, const float* lse,
                              void* stats, int T, int H, int t_stat, llm_stream_t stream);
int llmtrain_b200_attention_bwd_stats_buffer(void* scratch, size_t scratch_bytes, int B, int T,
                                             int H, int head_dim, void** stats, int* t_stat);
int llmtrain_b200_attention_bwd_prestats(const void* qkv, const void* out, const void* dout,
                                         const float* lse, const uint8_t* key_mask, void* dqkv,
                                         float* dbias_qkv, void* scratch, size_t scratch_bytes,
                                         int B, int T, int H, int head_dim, float softmax_scale,
                                         llm_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Dropout (csrc/common/philox.cuh, csrc/kernels/dropout.cu).  Replaces the four nn.Dropout modules
 * of src/llmtrain/models/gpt.py:28,30,97,129.  A mask is a pure function of (seed, offset, site,
 * element index = row * roundup(cols, 8) + col): forward and backward regenerate it, nothing is
 * stored.  keep probability = 1 - round(p * 65536) / 65536; kept values are scaled by its inverse.
 * ------------------------------------------------------------------------------------------- */
/* y = x o mask / (1 - p), bf16 [rows, cols] with row pitch ld (x == y allowed) */
int llmtrain_b200_dropout(const void* x, void* y, long long rows, int cols, int ld, float p,
                          uint64_t seed, uint64_t offset, uint32_t site, llm_stream_t stream);
/* out[rows, cols] uint8 = 1 where kept: the mask the fused kernels apply (tests build the
 * PyTorch reference from it) */
int llmtrain_b200_dropout_mask(uint8_t* out, long long rows, int cols, float p, uint64_t seed,
                               uint64_t offset, uint32_t site, llm_stream_t stream);
/* C[M,N] = dropout(A[M,K] B[N,K]^T + bias) + R ; mask rows = M, cols = N */
int llmtrain_b200_gemm_dropout_resid(int M, int N, int K, const void* A, int lda, const void* B,
                                     int ldb, void* C, int ldc, const float* bias, const void* R,
                                     int ldr, float p, uint64_t seed, uint64_t offset,
                                     uint32_t site, llm_stream_t stream);
/* C = (row_keep[m] ? dropout(A B^T + bias) : 0) + R  (p = 0: no dropout): out_proj of the attention
 * branch with the reference's padding-mask multiply fused (gpt.py:71-74). */
int llmtrain_b200_gemm_resid_rowmask(int M, int N, int K, const void* A, int lda, const void* B,
                                     int ldb, void* C, int ldc, const float* bias, const void* R,
                                     int ldr, const uint8_t* row_keep, float p, uint64_t seed,
                                     uint64_t offset, uint32_t site, llm_stream_t stream);
/* y = x o dropout_mask o row_keep[:, None] (p = 0: row mask only); backward of that branch. */
int llmtrain_b200_dropout_rowmask(const void* x, void* y, long long rows, int cols, int ld,
                                  const uint8_t* row_keep, float p, uint64_t seed, uint64_t offset,
                                  uint32_t site, llm_stream_t stream);
/* attention with dropout on the probabilities; mask rows = B*H*T (index (b*H+h)*T+q), cols = T */
int llmtrain_b200_attention_fwd_dropout(const void* qkv, void* out, float* lse,
                                        const uint8_t* key_mask, int B, int T, int H, int head_dim,
                                        float p, uint64_t seed, uint64_t offset, uint32_t site,
                                        llm_stream_t stream);
int llmtrain_b200_attention_bwd_dropout(const void* qkv, const void* out, const void* dout,
                                        const float* lse, const uint8_t* key_mask, void* dqkv,
                                        void* scratch, size_t scratch_bytes, int B, int T, int H,
                                        int head_dim, float p, uint64_t seed, uint64_t offset,
                                        uint32_t site, llm_stream_t stream);

/* attention backward with every option (what the engine calls): dropout as above (p = 0: none)
 * , softmax scale (0 = default) and, when dbias_qkv != NULL, dbias_qkv[3d] += column sums of dqkv = qkv_proj's bias gradient
 * (src/llmtrain/models/gpt.py:27) without another pass over dqkv.  fp32 atomics. */
int llmtrain_b200_attention_bwd_fused(const void* qkv, const void* out, const void* dout,
                                      const float* lse, const uint8_t* key_mask, void* dqkv,
                                      float* dbias_qkv, void* scratch, size_t scratch_bytes, int B,
                                      int T, int H, int head_dim, float softmax_scale, float p,
                                      uint64_t seed, uint64_t offset, uint32_t site,
                                      llm_stream_t stream);
/* forward counterpart: softmax_scale = 0 means head_dim^-0.5; a positive value overrides it (heads
 * narrower than 64 are run zero-padded to 64 lanes by the engine and keep their true scale). */
int llmtrain_b200_attention_fwd_fused(const void* qkv, void* out, float* lse,
                                      const uint8_t* key_mask, int B, int T, int H, int head_dim,
                                      float softmax_scale, float p, uint64_t seed, uint64_t offset,
                                      uint32_t site, llm_stream_t stream);

/* Head padding (csrc/kernels/headpad.cu): heads narrower than 64 (any d_model % n_heads == 0 is
 * legal in the reference, src/llmtrain/models/gpt.py:21-25) run zero-padded to 64 lanes.  A tensor
 * is seen as [outer][groups][width][inner]; head_pad widens `width` from hd to 64 with zeros
 * (elem_bytes 2 = bf16, 4 = fp32), head_unpad_add folds an fp32 gradient back (dst += src). */
int llmtrain_b200_head_pad(const void* src, void* dst, long long outer, int groups, int hd,
                           long long inner, int elem_bytes, llm_stream_t stream);
int llmtrain_b200_head_unpad_add(const float* src, float* dst, long long outer, int groups, int hd,
                                 long long inner, llm_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * GPT runtime (csrc/engine/gpt_engine.cu).  Replaces autograd over GPT.forward + compute_loss
 * (src/llmtrain/models/gpt.py:167-184,214-271; src/llmtrain/training/trainer.py:385-387).
 *
 * The engine plans memory but owns none: the caller allocates (e.g. as torch tensors, so that
 * nn.Parameters can alias the master weights) and binds
 *   p32 / g32 : fp32 [gpt_param_elems]      master weights / accumulated gradients
 *   p16       : bf16 [gpt_shadow_elems]     weight shadow consumed by the GEMMs
 *   act       : bytes [gpt_activation_bytes] activations + backward temporaries + scratch
 * gpt_param_offset maps tensors to arena offsets: layer < 0 -> which = 0 position_embedding,
 * 1 ln_f.weight, 2 ln_f.bias, 3 token_embedding, 4 lm_head (== 3 when tied); layer >= 0 ->
 * which = 0..11 = ln_1.w ln_1.b qkv.w qkv.b out.w out.b ln_2.w ln_2.b fc.w fc.b proj.w proj.b.
 * ------------------------------------------------------------------------------------------- */
int llmtrain_b200_gpt_create(int vocab, int block_size, int d_model, int n_layers, int n_heads,
                             int d_ff, int tied, int max_batch, int max_seq, void** handle);
void llmtrain_b200_gpt_destroy(void* handle);
size_t llmtrain_b200_gpt_param_elems(void* handle);
size_t llmtrain_b200_gpt_shadow_elems(void* handle);
size_t llmtrain_b200_gpt_activation_bytes(void* handle);
int llmtrain_b200_gpt_padded_vocab(void* handle);
size_t llmtrain_b200_gpt_logits_offset(void* handle);
size_t llmtrain_b200_gpt_loss_rows_offset(void* handle);
long long llmtrain_b200_gpt_param_offset(void* handle, int layer, int which);
int llmtrain_b200_gpt_bind(void* handle, float* p32, float* g32, void* p16, void* act,
                           size_t act_bytes);
int llmtrain_b200_gpt_refresh_shadow(void* handle, llm_stream_t stream);
/* mode 0: logits (bf16 [B*T, padded_vocab] at logits_offset).  mode 1: + loss rows and
 * loss_acc[0] += loss_scale * sum.  mode 2: as 1 and logits := dlogits * grad_scale (then call
 * gpt_backward).  key_mask [B,T] uint8 (1 = real token) or NULL; ids/labels int64 [B,T]. */
int llmtrain_b200_gpt_forward(void* handle, const int64_t* ids, const int64_t* labels,
                              const uint8_t* key_mask, int B, int T, int mode, float* loss_acc,
                              float loss_scale, float grad_scale, llm_stream_t stream);
/* Dropout for the NEXT mode-2 forward and its backward (p = 0: off; modes 0/1 never drop).  seed
 * identifies run and rank, offset must change with every micro-batch. */
int llmtrain_b200_gpt_set_dropout(void* handle, float p, uint64_t seed, uint64_t offset);
/* g32 += d loss / d params for the last mode-2 forward. */
int llmtrain_b200_gpt_backward(void* handle, llm_stream_t stream);
/* As llmtrain_b200_gpt_backward; hook(ctx, lo, len) is called on the host each time the kernels
 * that finalise gradient-arena range [lo, lo+len) have been enqueued (back to front), so a
 * transport of the caller's choice can reduce that bucket while backward continues (c10d Reducer
 * behaviour, reference trainer.py:88-91).  A non-zero return aborts backward. */
int llmtrain_b200_gpt_backward_hooked(void* gpt, int (*hook)(void*, size_t, size_t), void* ctx,
                                      llm_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * NVLink peer-memory communicator (csrc/comm/nvl.cu).  Replaces DistributedDataParallel's bucketed
 * all-reduce + clip_grad_norm_ + AdamW.step (src/llmtrain/training/trainer.py:87-91,377-394).
 * One process per GPU, 2..8 ranks on one NVLink domain.  create() cudaMallocs symmetric buffers
 * for an arena of n fp32 elements (n % 8 == 0); export() writes 3 cudaIpcMemHandle_t (192 bytes)
 * to host memory; the caller all-gathers those rank-major and passes them to connect().
 *   nvl_step:      g (fp32 local gradient sum) -> bf16 wire -> rank-ordered reduce-scatter ->
 *                  gather fused with *grad_scale, global-norm clip, AdamW on p/m/v, bf16 shadow
 *                  refresh, g := 0.  Replicas stay bit-identical.
 *   nvl_allreduce: out = scale * sum_r bf16(in_r)   (tests, bandwidth sweep)
 * A peer that never arrives trips a wall-clock watchdog (timeout_sec) and traps with a message.
 * ------------------------------------------------------------------------------------------- */
int llmtrain_b200_nvl_create(int rank, int world, size_t n, double timeout_sec, void** handle);
int llmtrain_b200_nvl_export(void* handle, void* h_out_192_bytes);
int llmtrain_b200_nvl_connect(void* handle, const void* h_all_handles);
void llmtrain_b200_nvl_destroy(void* handle);
int llmtrain_b200_nvl_barrier(void* handle, llm_stream_t stream);
int llmtrain_b200_nvl_step(void* handle, float* g, float* p, float* m, float* v, void* p16,
                           float* norm_out, float lr, float beta1, float beta2, float eps,
                           float weight_decay, int step, float grad_scale, float max_norm,
                           llm_stream_t stream);
int llmtrain_b200_nvl_allreduce(void* handle, float* in, float* out, float scale,
                                llm_stream_t stream);
/* One launch, every rank reads all peers (latency-optimal, count <= 262144, arena >= 2*count). */
int llmtrain_b200_nvl_allreduce_oneshot(void* handle, const float* in, float* out, size_t count,
                                        float scale, llm_stream_t stream);
/* Copy-engine broadcast of fp32 values from `root` (initial parameter sync). */
int llmtrain_b200_nvl_broadcast(void* handle, float* buf, size_t count, int root,
                                llm_stream_t stream);
/* Overlapped path: backward of a bound GPT runtime with per-layer buckets reduced on the
 * communicator's stream (owners' reduced slices reach all ranks via copy engines), then the
 * local clip + AdamW tail. */
int llmtrain_b200_nvl_backward_overlapped(void* nvl_handle, void* gpt_handle, float* g,
                                          llm_stream_t stream);
int llmtrain_b200_nvl_finish_step(void* nvl_handle, float* g, float* p, float* m, float* v,
                                  void* p16, float* norm_out, float lr, float beta1, float beta2,
                                  float eps, float weight_decay, int step, float grad_scale,
                                  float max_norm, llm_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * NVLink-SHARP all-reduce (csrc/comm/nvls.cu): the NVSwitch reduces in flight
 * (multimem.ld_reduce / multimem.st on a multicast mapping).  Large-message alternative to the
 * peer-to-peer two-shot scheme for what DistributedDataParallel's NCCL all-reduce does in the
 * reference (src/llmtrain/training/trainer.py:87-91).  Setup, all ranks in lock step:
 *   create -> (rank 0) export 80 bytes -> caller broadcasts them -> connect -> host barrier ->
 *   bind -> host barrier.  Cross-GPU ordering inside collectives uses nvl_handle's flag barrier.
 * ------------------------------------------------------------------------------------------- */
int llmtrain_b200_nvls_supported(void);
int llmtrain_b200_nvls_create(void* nvl_handle, int rank, int world, size_t n, int use_fabric,
                              void** handle);
int llmtrain_b200_nvls_export(void* handle, void* out_80_bytes);
int llmtrain_b200_nvls_connect(void* handle, const void* wire_80_bytes);
int llmtrain_b200_nvls_bind(void* handle);
void* llmtrain_b200_nvls_buffer(void* handle);     /* this rank's view (ordinary memory) */
void* llmtrain_b200_nvls_mc_buffer(void* handle);  /* the team's view (multimem.* only) */
/* overlapped training path in NVLink-SHARP mode: the gradient arena passed to
 * nvl_backward_overlapped / nvl_finish_step must then be `uc` itself (NULL, NULL detaches) */
int llmtrain_b200_nvl_attach_nvls(void* nvl_handle, float* uc, float* mc);
/* Where llmtrain_b200_nvl_backward_overlapped leaves the cross-rank SUM of the gradient buckets
 * (fp32, arena layout, valid once the communicator's stream has drained and every rank has passed
 * a barrier): the symmetric `full` buffer on the peer-to-peer path, the arena `g` itself when the
 * buckets are reduced in the switch.  For verification (bench.py ddp_check, the DDP-equivalence
 * test; reference behaviour being checked: DDP's averaged gradients, trainer.py:377-387). */
const float* llmtrain_b200_nvl_reduced(void* nvl_handle, const float* g);
/* buffer[i] := scale * sum_r buffer_r[i] on every rank (count % (4*world) == 0).  in / out: optional
 * fp32 [count] copied in before / out after; NULL, NULL runs the collective alone, in place. */
int llmtrain_b200_nvls_allreduce(void* handle, const float* in, float* out, size_t count,
                                 float scale, llm_stream_t stream);
void llmtrain_b200_nvls_destroy(void* handle);

#ifdef __cplusplus
}
#endif
#endif /* LLMTRAIN_B200_H_ */
