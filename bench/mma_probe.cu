// Cycles per tcgen05.mma (kind::f16, bf16, cta_group::1, M = 128, K = 16) by operand source and
// layout, measured with clock64 around a queue of back-to-back instructions on one CTA per SM.
// Build and run on a B200:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Icsrc -o /tmp/mma_probe \
//        bench/mma_probe.cu -lcuda && /tmp/mma_probe
// The numbers decide which products of the attention kernels can reach the tensor-core floor
// (128 * N / 256 cycles per instruction) and which are paced by their operand fetch.

#include <cstdio>
#include <cstdlib>

#include "common/ptx.cuh"

using namespace b200;

constexpr int kTile = 16384;
constexpr int kIssues = 512;

// MODE: 0 SS  A K-major, B K-major      1 SS  A K-major, B MN-major
//       2 SS  A MN-major, B MN-major    3 TS  B MN-major              4 TS  B K-major
// WARP: 1 = the whole warp runs the loop and an elected lane issues (descriptors stay in uniform
//       registers, one 64-bit add per instruction); 0 = the loop sits inside `if (thread == 0)`.
template <int MODE, int N, int WARP>
__global__ void __launch_bounds__(128, 1) probe(long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  // the shuffle tells the compiler that the value is warp-uniform
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
  }
  for (int i = threadIdx.x; i < 6 * kTile / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;  // small bf16 values
  fence_proxy_async_smem();
  if (warp == 0) {
    tmem_alloc(&slot, 512);
    tmem_relinquish();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = slot;
  constexpr bool a_mn = MODE == 2, b_mn = MODE == 1 || MODE == 2 || MODE == 3;
  constexpr uint32_t id = umma_idesc_bf16(128, N, a_mn, b_mn);
  if (warp == 0 && (WARP || threadIdx.x == 0)) {
    const uint32_t sA = smem_u32(smem), sB = smem_u32(smem + 2 * kTile);
    const uint64_t da = a_mn ? umma_smem_desc(sA, kTile, 1024) : umma_smem_desc(sA, 0, 1024);
    const uint64_t db = b_mn ? umma_smem_desc(sB, kTile, 1024) : umma_smem_desc(sB, 0, 1024);
    const bool leader = WARP ? elect_one() : true;
    const long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < kIssues / 8; ++i) {
      if (leader) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint64_t ak = da + (a_mn ? 128 * k : 2 * (k & 3));
          const uint64_t bk = db + (b_mn ? 128 * k : 2 * (k & 3));
          if (MODE >= 3)
            umma_ts(tmem + 256, tmem + k * 8, bk, id, (i | k) > 0);
          else
            umma_ss(tmem + 256, ak, bk, id, (i | k) > 0);
        }
      }
    }
    const long long t1 = clock64();
    if (leader) umma_commit(&bar);
    mbar_wait(&bar, 0);
    const long long t2 = clock64();
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      out[0] = t1 - t0;
      out[1] = t2 - t0;
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

template <int MODE, int N, int WARP>
void run(long long* out, const char* name) {
  cudaFuncSetAttribute(probe<MODE, N, WARP>, cudaFuncAttributeMaxDynamicSharedMemorySize, 6 * kTile);
  long long h[2] = {0, 0};
  for (int rep = 0; rep < 2; ++rep) {
    probe<MODE, N, WARP><<<148, 128, 6 * kTile>>>(out);
    if (cudaDeviceSynchronize() != cudaSuccess) {
      std::printf("%s: %s\n", name, cudaGetErrorString(cudaGetLastError()));
      std::exit(1);
    }
  }
  cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
  std::printf("%s  N %3d  %s : issue %6.1f  complete %6.1f cycles per MMA (floor %d)\n",
              WARP ? "warp + elect " : "if (thread==0)", N, name, double(h[0]) / kIssues,
              double(h[1]) / kIssues, 128 * N / 256);
}

int main() {
  long long* out;
  cudaMalloc(&out, 16);
#define ROW(W, N)                                      \
  run<0, N, W>(out, "SS A K-major  B K-major ");       \
  run<1, N, W>(out, "SS A K-major  B MN-major");       \
  run<2, N, W>(out, "SS A MN-major B MN-major");       \
  run<3, N, W>(out, "TS            B MN-major");       \
  run<4, N, W>(out, "TS            B K-major ");
  ROW(0, 64) ROW(0, 128) ROW(1, 64) ROW(1, 128) ROW(1, 256)
  return 0;
}
