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

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* llm_stream_t;

const char* llmtrain_b200_last_error(void);
/* 1000*major + 10*minor of the library ABI. */
int llmtrain_b200_abi_version(void);

/* ---------------------------------------------------------------------------------------------
 * GEMM (tcgen05 + TMEM + TMA).  Replaces nn.Linear forward and its autograd dgrad/wgrad:
 * src/llmtrain/models/gpt.py:27,29,94,96,143.
 *   layout 0 (NT): C[M,N]  = A[M,K]   * B[N,K]^T
 *   layout 1 (NN): C[M,N]  = A[M,K]   * B[K,N]
 *   layout 2 (TN): C[M,N] += A[K,M]^T * B[K,N]      (C is fp32; split-K; TMA reduce-add)
 *   epilogue 0: C = acc (+bias)            1: C2 = acc+bias, C = gelu_erf(C2)
 *            2: C = acc (+bias) + R        3: C = acc * gelu_erf'(R)      4: fp32 accumulate
 *   A, B, C(bf16), C2, R are bf16 with row pitches lda/ldb/ldc/ldc2/ldr in elements (each pitch
 *   a multiple of 8); bias is fp32 [N] or NULL.
 * ------------------------------------------------------------------------------------------- */
int llmtrain_b200_gemm(int layout, int epilogue, int M, int N, int K, const void* A, int lda,
                       const void* B, int ldb, void* C, int ldc, void* C2, int ldc2,
                       const float* bias, const void* R, int ldr, llm_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* LLMTRAIN_B200_H_ */
