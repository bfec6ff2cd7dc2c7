// NVLink / NVSwitch peer-memory gradient reduction fused with the optimiser — the replacement for
// torch DDP's c10d Reducer + ProcessGroup all-reduce (`src/llmtrain/training/trainer.py:87-91,
// 377-394`; SURVEY §2.6 X3, work items W7/W8/W10/W12).  No NCCL call is made on this path.
//
// One process per GPU.  Each rank cudaMalloc's three *symmetric* buffers and publishes their CUDA
// IPC handles (exchanged by the caller over the bootstrap process group):
//     send  : bf16 [n]        this rank's gradient as it goes on the wire
//     red   : fp32 [chunk]    the reduced 1/W slice this rank owns
//     sync  : flags[W] + norm partial (one 4 KB page)
// and maps every peer's buffers, so kernels dereference peer memory directly (LDG/STG over NVLink).
//
// One optimiser step = five launches on the caller's stream, no host synchronisation:
//   1. pack      send[i] = bf16(g32[i])                                   (fp32 accumulate -> bf16 wire)
//   2. barrier   flag exchange with release/acquire at system scope, monotonically increasing epoch
//   3. scatter   owner r sums slice r over all ranks IN RANK ORDER (fp32), stores it to red and
//                produces sum(g^2) of the slice (deterministic two-stage reduction)
//   4. barrier
//   5. gather    every rank walks all slices, pulling the reduced values from their owners, and
//                applies average + global-norm clip + AdamW + bf16 shadow refresh + grad zeroing in
//                the same pass — the all-gather never lands in memory.
// Because each element is reduced by exactly one owner in a fixed order and every rank consumes
// the same fp32 value, replicas stay bit-identical without a parameter broadcast.
//
// A bounded spin (wall-clock, `timeout`) turns a lost peer into a trap with a diagnostic instead
// of a hang.  Buffers are not reused before all peers have consumed them: the next
// step's pack happens after this step's second barrier, and the next scatter after the next
// first barrier, which a peer only signals once its gather has finished.

#include <cmath>
#include <cstdlib>
#include <vector>

#include "../common/host.cuh"
#include "../common/ptx.cuh"
#include "../../include/llmtrain_b200.h"

namespace b200 {
namespace nvl {

constexpr int MAX_RANKS = 8;
constexpr int SYNC_BYTES = 4096;
constexpr int NORM_SLOT = 64;  // float index inside the sync page (flags occupy [0, MAX_RANKS))
constexpr int THREADS = 512;

struct Peers {
  const __nv_bfloat16* send[MAX_RANKS];
  const float* red[MAX_RANKS];
  uint32_t* sync[MAX_RANKS];
  float* full[MAX_RANKS];  // fp32 [n]: the reduced gradient, pushed by the slice owners
};

constexpr int MAX_BUCKETS = 128;
// Overlapped bucket kernels must CO-RESIDE with the backward GEMM CTAs (1 per SM, ~54K of the 64K
// registers, ~224 KB smem): small CTAs, no smem, <= 80 registers per thread.
constexpr int BUCKET_GRID = 148;
constexpr int OVL_THREADS = 128;
// Buckets that only become final when backward is over (the embedding tables) find the GPU idle:
// they get a full-size grid instead.
constexpr int TAIL_GRID = 1024;
constexpr int PARTIAL_STRIDE = TAIL_GRID * (OVL_THREADS / 32);  // one slot per warp per bucket

struct Comm {
  int rank = 0, world = 1, device = 0;
  size_t n = 0, chunk = 0;
  __nv_bfloat16* send = nullptr;
  float* red = nullptr;
  uint32_t* sync = nullptr;
  float* partials = nullptr;  // per-CTA sum-of-squares partials (local scratch)
  int scatter_grid = 0;
  float* full = nullptr;             // symmetric fp32 [n] (overlapped path)
  float* bucket_partials = nullptr;  // [MAX_BUCKETS][BUCKET_GRID]
  int buckets_this_step = 0;
  int bucket_grid = BUCKET_GRID;
  bool dma_gather = true;  // push reduced slices with cudaMemcpyAsync (copy engines) vs SM stores
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ready_events[MAX_BUCKETS] = {};
  cudaEvent_t done_event = nullptr;
  Peers peers{};
  void* opened[4 * MAX_RANKS] = {};
  uint32_t epoch = 0;
  unsigned long long timeout_ns = 1800ull * 1000000000ull;
  bool peers_open = false;
  uint32_t oneshot_calls = 0;
  // NVLink-SHARP mode of the overlapped path (llmtrain_b200_nvl_attach_nvls): the gradient arena
  // itself lives in multicast-bound memory; nvls_uc is this rank's view, nvls_mc the team's.
  float* nvls_uc = nullptr;
  float* nvls_mc = nullptr;
};

__device__ __forceinline__ unsigned long long now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Thread p tells peer p "rank `rank` reached `epoch`", then waits until peer p has told us.
__global__ void barrier_kernel(Peers peers, int rank, int world, uint32_t epoch,
                               unsigned long long timeout_ns) {
  const int p = threadIdx.x;
  if (p < world) {
    __threadfence_system();
    st_release_sys(peers.sync[p] + rank, epoch);
    const uint32_t* mine = peers.sync[rank] + p;
    const unsigned long long t0 = now_ns();
    // epochs only grow; compare with wrap-around in mind
    while (static_cast<int32_t>(ld_acquire_sys(mine) - epoch) < 0) {
      if (now_ns() - t0 > timeout_ns) {
        printf("llmtrain nvl: rank %d timed out waiting for rank %d at epoch %u\n", rank, p, epoch);
        __trap();
      }
    }
  }
  __syncthreads();
  __threadfence_system();
}

__global__ void __launch_bounds__(THREADS)
pack_kernel(const float* __restrict__ g, __nv_bfloat16* __restrict__ send, size_t n) {
  const size_t n8 = n >> 3;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n8;
       i += size_t(gridDim.x) * blockDim.x) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(g) + 2 * i);
    const float4 b = __ldg(reinterpret_cast<const float4*>(g) + 2 * i + 1);
    reinterpret_cast<uint4*>(send)[i] = make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w),
                                                   pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
  }
}

// red[j] = sum_r send_r[lo + j]  for j in my slice; partial[blockIdx.x] = sum red^2
template <int W>
__global__ void __launch_bounds__(THREADS)
scatter_kernel(Peers peers, int rank, size_t lo, size_t len, float* __restrict__ red_out,
               float* __restrict__ partial) {
  __shared__ float sh[THREADS / 32];
  const size_t n8 = len >> 3;  // len is a multiple of 8
  float sq = 0.f;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n8;
       i += size_t(gridDim.x) * blockDim.x) {
    uint4 v[W];
#pragma unroll
    for (int r = 0; r < W; ++r) v[r] = reinterpret_cast<const uint4*>(peers.send[r] + lo)[i];
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < W; ++r) {  // fixed rank order => deterministic
      const uint32_t w[4] = {v[r].x, v[r].y, v[r].z, v[r].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = unpack_bf16x2(w[j]);
        acc[2 * j] += f.x;
        acc[2 * j + 1] += f.y;
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) sq += acc[j] * acc[j];
    reinterpret_cast<float4*>(red_out)[2 * i] = make_float4(acc[0], acc[1], acc[2], acc[3]);
    reinterpret_cast<float4*>(red_out)[2 * i + 1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = sq;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < THREADS / 32 ? sh[threadIdx.x] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (threadIdx.x == 0) partial[blockIdx.x] = t;
  }
}

__global__ void norm_finish_kernel(const float* __restrict__ partial, int nparts,
                                   float* __restrict__ out) {
  __shared__ float sh[32];
  float acc = 0.f;
  for (int i = threadIdx.x; i < nparts; i += blockDim.x) acc += partial[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (threadIdx.x == 0) out[0] = t;
  }
}

struct Adam {
  float lr, beta1, beta2, eps, weight_decay, bias_corr1, bias_corr2_sqrt, grad_scale, max_norm;
};

// Walk the whole arena; element i's reduced gradient lives in the `red` buffer of rank i / chunk.
__global__ void __launch_bounds__(THREADS)
gather_adamw_kernel(Peers peers, int world, size_t n, size_t chunk, float* __restrict__ p,
                    float* __restrict__ m, float* __restrict__ v, float* __restrict__ g,
                    __nv_bfloat16* __restrict__ p16, float* __restrict__ out_reduced,
                    float* __restrict__ norm_out, Adam a, int apply_adam) {
  float total = 0.f;
  for (int r = 0; r < world; ++r)
    total += reinterpret_cast<const float*>(peers.sync[r])[NORM_SLOT];  // same order on all ranks
  const float norm = sqrtf(total) * a.grad_scale;
  float coef = a.grad_scale;
  if (a.max_norm > 0.f) coef *= fminf(1.0f, a.max_norm / (norm + 1e-6f));
  if (norm_out != nullptr && blockIdx.x == 0 && threadIdx.x == 0) norm_out[0] = norm;
  const float decay = 1.0f - a.lr * a.weight_decay;
  const float step_size = a.lr / a.bias_corr1;
  const float omb1 = 1.0f - a.beta1, omb2 = 1.0f - a.beta2;
  const size_t n4 = n >> 2;  // n and chunk are multiples of 8
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
       i += size_t(gridDim.x) * blockDim.x) {
    const size_t e = i << 2;
    const int owner = int(e / chunk);
    const float4 gr = *reinterpret_cast<const float4*>(peers.red[owner] + (e - size_t(owner) * chunk));
    if (!apply_adam) {
      reinterpret_cast<float4*>(out_reduced)[i] =
          make_float4(gr.x * a.grad_scale, gr.y * a.grad_scale, gr.z * a.grad_scale,
                      gr.w * a.grad_scale);
      continue;
    }
    float4 pv = reinterpret_cast<float4*>(p)[i];
    float4 mv = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float* pp = reinterpret_cast<float*>(&pv);
    float* mp = reinterpret_cast<float*>(&mv);
    float* vp = reinterpret_cast<float*>(&vv);
    const float* gp = reinterpret_cast<const float*>(&gr);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float gj = gp[j] * coef;
      mp[j] = a.beta1 * mp[j] + omb1 * gj;
      vp[j] = a.beta2 * vp[j] + omb2 * gj * gj;
      pp[j] = pp[j] * decay - step_size * (mp[j] / (sqrtf(vp[j]) / a.bias_corr2_sqrt + a.eps));
    }
    reinterpret_cast<float4*>(p)[i] = pv;
    reinterpret_cast<float4*>(m)[i] = mv;
    reinterpret_cast<float4*>(v)[i] = vv;
    reinterpret_cast<uint2*>(p16)[i] = make_uint2(pack_bf16x2(pp[0], pp[1]), pack_bf16x2(pp[2], pp[3]));
    reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// Overlapped path, one bucket: rank r owns sub-slice r of [lo, lo+len); it sums that sub-slice over
// all ranks (rank order, fp32) and PUSHES the result into every rank's `full` buffer, so that after
// the last bucket each rank holds the complete reduced gradient locally.
template <int W>
__global__ void __launch_bounds__(OVL_THREADS, 4)
scatter_push_kernel(Peers peers, size_t lo, size_t len, float* __restrict__ partial,
                    int only_rank /* >= 0: store to that rank's buffer only (DMA does the rest) */) {
  const size_t n8 = len >> 3;
  float sq = 0.f;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n8;
       i += size_t(gridDim.x) * blockDim.x) {
    uint4 v[W];
#pragma unroll
    for (int r = 0; r < W; ++r) v[r] = reinterpret_cast<const uint4*>(peers.send[r] + lo)[i];
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < W; ++r) {
      const uint32_t w[4] = {v[r].x, v[r].y, v[r].z, v[r].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = unpack_bf16x2(w[j]);
        acc[2 * j] += f.x;
        acc[2 * j + 1] += f.y;
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) sq += acc[j] * acc[j];
    const float4 a = make_float4(acc[0], acc[1], acc[2], acc[3]);
    const float4 b = make_float4(acc[4], acc[5], acc[6], acc[7]);
    if (only_rank >= 0) {
      float4* dst = reinterpret_cast<float4*>(peers.full[only_rank] + lo) + 2 * i;
      dst[0] = a;
      dst[1] = b;
    } else {
#pragma unroll
      for (int r = 0; r < W; ++r) {
        float4* dst = reinterpret_cast<float4*>(peers.full[r] + lo) + 2 * i;
        dst[0] = a;
        dst[1] = b;
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
  // No shared memory on purpose (see OVL_THREADS): one partial per warp, summed later in a fixed
  // order by norm_finish_kernel.
  if ((threadIdx.x & 31) == 0)
    partial[blockIdx.x * (OVL_THREADS / 32) + (threadIdx.x >> 5)] = sq;
}

// NVLink-SHARP bucket reduction, in place on the gradient arena: this rank's sub-slice is summed
// over all ranks INSIDE the switch (multimem.ld_reduce) and the sum is written back to every
// rank's arena (multimem.st).  No pack, no send buffer, no gather; 1/W of the bucket per rank and
// two instructions per 16 bytes, so it disturbs the backward GEMMs it runs next to far less than
// the peer-to-peer pull.  Also leaves the slice's sum of squares (global-norm clipping).
__global__ void __launch_bounds__(OVL_THREADS, 4)
nvls_bucket_kernel(float* __restrict__ mc, size_t lo, size_t len, float* __restrict__ partial) {
  const size_t n4 = len >> 2;
  float* base = mc + lo;
  float sq = 0.f;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
       i += size_t(gridDim.x) * blockDim.x) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(base + 4 * i)
                 : "memory");
    sq += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(base + 4 * i),
                 "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
  if ((threadIdx.x & 31) == 0)
    partial[blockIdx.x * (OVL_THREADS / 32) + (threadIdx.x >> 5)] = sq;
}

__global__ void __launch_bounds__(OVL_THREADS, 4)
pack_range_kernel(const float* __restrict__ g, __nv_bfloat16* __restrict__ send, size_t lo,
                  size_t len) {
  const size_t n8 = len >> 3;
  const float4* src = reinterpret_cast<const float4*>(g + lo);
  uint4* dst = reinterpret_cast<uint4*>(send + lo);
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n8;
       i += size_t(gridDim.x) * blockDim.x) {
    const float4 a = __ldg(src + 2 * i), b = __ldg(src + 2 * i + 1);
    dst[i] = make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y),
                        pack_bf16x2(b.z, b.w));
  }
}

// Local AdamW over the complete reduced gradient in `full` (overlapped path).
__global__ void __launch_bounds__(THREADS)
adamw_full_kernel(Peers peers, int world, size_t n, const float* full /* may alias g */,
                  float* __restrict__ p, float* __restrict__ m, float* __restrict__ v, float* g,
                  __nv_bfloat16* __restrict__ p16,
                  float* __restrict__ norm_out, Adam a) {
  float total = 0.f;
  for (int r = 0; r < world; ++r)
    total += reinterpret_cast<const float*>(peers.sync[r])[NORM_SLOT];
  const float norm = sqrtf(total) * a.grad_scale;
  float coef = a.grad_scale;
  if (a.max_norm > 0.f) coef *= fminf(1.0f, a.max_norm / (norm + 1e-6f));
  if (norm_out != nullptr && blockIdx.x == 0 && threadIdx.x == 0) norm_out[0] = norm;
  const float decay = 1.0f - a.lr * a.weight_decay;
  const float step_size = a.lr / a.bias_corr1;
  const float omb1 = 1.0f - a.beta1, omb2 = 1.0f - a.beta2;
  const size_t n4 = n >> 2;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
       i += size_t(gridDim.x) * blockDim.x) {
    const float4 gr = reinterpret_cast<const float4*>(full)[i];
    float4 pv = reinterpret_cast<float4*>(p)[i];
    float4 mv = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float* pp = reinterpret_cast<float*>(&pv);
    float* mp = reinterpret_cast<float*>(&mv);
    float* vp = reinterpret_cast<float*>(&vv);
    const float* gp = reinterpret_cast<const float*>(&gr);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float gj = gp[j] * coef;
      mp[j] = a.beta1 * mp[j] + omb1 * gj;
      vp[j] = a.beta2 * vp[j] + omb2 * gj * gj;
      pp[j] = pp[j] * decay - step_size * (mp[j] / (sqrtf(vp[j]) / a.bias_corr2_sqrt + a.eps));
    }
    reinterpret_cast<float4*>(p)[i] = pv;
    reinterpret_cast<float4*>(m)[i] = mv;
    reinterpret_cast<float4*>(v)[i] = vv;
    reinterpret_cast<uint2*>(p16)[i] = make_uint2(pack_bf16x2(pp[0], pp[1]), pack_bf16x2(pp[2], pp[3]));
    reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// ------------------------------------------------------------------------------------------------
// One-shot all-reduce for small messages (<= ONESHOT_MAX elements): a single CTA packs the local
// values to bf16 in its send buffer, exchanges flags with every peer, then every rank reads ALL
// peers' buffers and reduces locally in rank order.  One launch instead of five: latency-optimal
// where the two-shot pipeline is launch-bound.  The send buffer alternates between two halves by
// call parity, so a fast rank's next call cannot overwrite data a slow peer is still reading.
// ------------------------------------------------------------------------------------------------
constexpr size_t ONESHOT_MAX = 1u << 18;  // elements (1 MiB of fp32 payload)

template <int W>
__global__ void __launch_bounds__(1024)
oneshot_kernel(Peers peers, int rank, uint32_t epoch, size_t half_off, const float* __restrict__ in,
               float* __restrict__ out, size_t n, float scale, unsigned long long timeout_ns) {
  __nv_bfloat16* mine = const_cast<__nv_bfloat16*>(peers.send[rank]) + half_off;
  for (size_t i = threadIdx.x; i < n; i += blockDim.x) mine[i] = __float2bfloat16(in[i]);
  __syncthreads();
  if (threadIdx.x < W) {
    const int p = threadIdx.x;
    __threadfence_system();
    st_release_sys(peers.sync[p] + rank, epoch);
    const uint32_t* flag = peers.sync[rank] + p;
    const unsigned long long t0 = now_ns();
    while (static_cast<int32_t>(ld_acquire_sys(flag) - epoch) < 0) {
      if (now_ns() - t0 > timeout_ns) {
        printf("llmtrain nvl: rank %d timed out waiting for rank %d (one-shot, epoch %u)\n", rank,
               p, epoch);
        __trap();
      }
    }
  }
  __syncthreads();
  __threadfence_system();
  for (size_t i = threadIdx.x; i < n; i += blockDim.x) {
    float acc = 0.f;
#pragma unroll
    for (int r = 0; r < W; ++r) acc += __bfloat162float(peers.send[r][half_off + i]);
    out[i] = acc * scale;
  }
}

inline int launch_barrier(Comm& c, cudaStream_t s) {
  c.epoch += 1;
  barrier_kernel<<<1, 32, 0, s>>>(c.peers, c.rank, c.world, c.epoch, c.timeout_ns);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}

inline int launch_scatter(Comm& c, cudaStream_t s) {
  const size_t lo = size_t(c.rank) * c.chunk;
  const size_t hi = lo + c.chunk < c.n ? lo + c.chunk : c.n;
  const size_t len = hi > lo ? hi - lo : 0;
#define B200_SCATTER(W)                                                                      \
  case W:                                                                                    \
    scatter_kernel<W><<<c.scatter_grid, THREADS, 0, s>>>(c.peers, c.rank, lo, len, c.red,    \
                                                         c.partials);                        \
    break;
  switch (c.world) {
    B200_SCATTER(2) B200_SCATTER(3) B200_SCATTER(4) B200_SCATTER(5) B200_SCATTER(6)
    B200_SCATTER(7) B200_SCATTER(8)
    default: return fail("nvl: world size must be in [2, 8]");
  }
#undef B200_SCATTER
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  norm_finish_kernel<<<1, 256, 0, s>>>(c.partials, c.scatter_grid,
                                       reinterpret_cast<float*>(c.sync) + NORM_SLOT);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}

}  // namespace nvl
}  // namespace b200

using namespace b200;
using b200::nvl::Comm;

// Creates the communicator and its symmetric buffers for an arena of n fp32 elements (n must be a
// multiple of 8).  Checks that every peer device is reachable with P2P.
extern "C" int llmtrain_b200_nvl_create(int rank, int world, size_t n, double timeout_sec,
                                        void** handle) {
  B200_REQUIRE(handle != nullptr, "nvl_create: null handle");
  B200_REQUIRE(world >= 2 && world <= nvl::MAX_RANKS, "nvl_create: world size must be in [2, 8]");
  B200_REQUIRE(rank >= 0 && rank < world, "nvl_create: bad rank");
  B200_REQUIRE(n % 8 == 0 && n > 0, "nvl_create: arena length must be a positive multiple of 8");
  auto* c = new Comm();
  c->rank = rank; c->world = world; c->n = n;
  B200_CHECK_CUDA(cudaGetDevice(&c->device));
  size_t chunk = (n + world - 1) / world;
  chunk = (chunk + 1023) / 1024 * 1024;
  c->chunk = chunk;
  c->timeout_ns = static_cast<unsigned long long>(timeout_sec * 1e9);
  B200_CHECK_CUDA(cudaMalloc(&c->send, n * sizeof(__nv_bfloat16)));
  B200_CHECK_CUDA(cudaMalloc(&c->red, chunk * sizeof(float)));
  B200_CHECK_CUDA(cudaMalloc(&c->sync, nvl::SYNC_BYTES));
  B200_CHECK_CUDA(cudaMemset(c->sync, 0, nvl::SYNC_BYTES));
  B200_CHECK_CUDA(cudaMemset(c->red, 0, chunk * sizeof(float)));
  c->scatter_grid = 2 * sm_count();
  B200_CHECK_CUDA(cudaMalloc(&c->partials, size_t(c->scatter_grid) * sizeof(float)));
  B200_CHECK_CUDA(cudaMalloc(&c->full, n * sizeof(float)));
  B200_CHECK_CUDA(cudaMemset(c->full, 0, n * sizeof(float)));
  B200_CHECK_CUDA(cudaMalloc(&c->bucket_partials,
                             size_t(nvl::MAX_BUCKETS) * nvl::PARTIAL_STRIDE * sizeof(float)));
  B200_CHECK_CUDA(cudaMemset(c->bucket_partials, 0,
                             size_t(nvl::MAX_BUCKETS) * nvl::PARTIAL_STRIDE * sizeof(float)));
  {
    // Highest priority: the few small comm CTAs should win any CTA slot that frees up, otherwise
    // a bucket's reduction trails far behind the backward pass it is meant to overlap.
    int least = 0, greatest = 0;
    B200_CHECK_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    B200_CHECK_CUDA(cudaStreamCreateWithPriority(&c->comm_stream, cudaStreamNonBlocking, greatest));
  }
  if (const char* env = std::getenv("LLMTRAIN_NVL_DMA_GATHER")) c->dma_gather = std::atoi(env) != 0;
  if (const char* env = std::getenv("LLMTRAIN_NVL_BUCKET_GRID")) {
    const int v = std::atoi(env);
    if (v >= 1 && v <= nvl::TAIL_GRID) c->bucket_grid = v;
  }
  for (auto& ev : c->ready_events)
    B200_CHECK_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  B200_CHECK_CUDA(cudaEventCreateWithFlags(&c->done_event, cudaEventDisableTiming));
  B200_CHECK_CUDA(cudaDeviceSynchronize());
  *handle = c;
  return 0;
}

// out: 4 consecutive cudaIpcMemHandle_t (send, red, sync, full) = 4 * 64 bytes (host memory).
extern "C" int llmtrain_b200_nvl_export(void* handle, void* h_out) {
  Comm& c = *static_cast<Comm*>(handle);
  auto* out = static_cast<cudaIpcMemHandle_t*>(h_out);
  B200_CHECK_CUDA(cudaIpcGetMemHandle(&out[0], c.send));
  B200_CHECK_CUDA(cudaIpcGetMemHandle(&out[1], c.red));
  B200_CHECK_CUDA(cudaIpcGetMemHandle(&out[2], c.sync));
  B200_CHECK_CUDA(cudaIpcGetMemHandle(&out[3], c.full));
  return 0;
}

// h_all: world * 4 handles, rank-major (the all-gathered output of nvl_export).
extern "C" int llmtrain_b200_nvl_connect(void* handle, const void* h_all) {
  Comm& c = *static_cast<Comm*>(handle);
  const auto* all = static_cast<const cudaIpcMemHandle_t*>(h_all);
  for (int r = 0; r < c.world; ++r) {
    if (r == c.rank) {
      c.peers.send[r] = c.send;
      c.peers.red[r] = c.red;
      c.peers.sync[r] = c.sync;
      c.peers.full[r] = c.full;
      continue;
    }
    void* ptrs[4] = {nullptr, nullptr, nullptr, nullptr};
    for (int k = 0; k < 4; ++k) {
      B200_CHECK_CUDA(cudaIpcOpenMemHandle(&ptrs[k], all[r * 4 + k], cudaIpcMemLazyEnablePeerAccess));
      c.opened[r * 4 + k] = ptrs[k];
    }
    c.peers.send[r] = static_cast<const __nv_bfloat16*>(ptrs[0]);
    c.peers.red[r] = static_cast<const float*>(ptrs[1]);
    c.peers.sync[r] = static_cast<uint32_t*>(ptrs[2]);
    c.peers.full[r] = static_cast<float*>(ptrs[3]);
  }
  c.peers_open = true;
  return 0;
}

extern "C" void llmtrain_b200_nvl_destroy(void* handle) {
  auto* c = static_cast<Comm*>(handle);
  if (c == nullptr) return;
  cudaDeviceSynchronize();
  for (void* p : c->opened)
    if (p != nullptr) cudaIpcCloseMemHandle(p);
  cudaFree(c->send);
  cudaFree(c->red);
  cudaFree(c->sync);
  cudaFree(c->partials);
  cudaFree(c->full);
  cudaFree(c->bucket_partials);
  for (auto& ev : c->ready_events)
    if (ev != nullptr) cudaEventDestroy(ev);
  if (c->done_event != nullptr) cudaEventDestroy(c->done_event);
  if (c->comm_stream != nullptr) cudaStreamDestroy(c->comm_stream);
  delete c;
}

extern "C" int llmtrain_b200_nvl_barrier(void* handle, llm_stream_t stream) {
  Comm& c = *static_cast<Comm*>(handle);
  B200_REQUIRE(c.peers_open, "nvl: peers are not connected");
  return nvl::launch_barrier(c, reinterpret_cast<cudaStream_t>(stream));
}

static int nvl_run(Comm& c, float* g, float* p, float* m, float* v, void* p16, float* out_reduced,
                   float* norm_out, const nvl::Adam& a, int apply_adam, cudaStream_t s) {
  B200_REQUIRE(c.peers_open, "nvl: peers are not connected");
  const int grid = 4 * sm_count();
  nvl::pack_kernel<<<grid, nvl::THREADS, 0, s>>>(g, c.send, c.n);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  if (nvl::launch_barrier(c, s)) return 1;
  if (nvl::launch_scatter(c, s)) return 1;
  if (nvl::launch_barrier(c, s)) return 1;
  nvl::gather_adamw_kernel<<<grid, nvl::THREADS, 0, s>>>(
      c.peers, c.world, c.n, c.chunk, p, m, v, g, static_cast<__nv_bfloat16*>(p16), out_reduced,
      norm_out, a, apply_adam);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}

// Fused step: g (fp32, local sum over micro-batches) is reduced over ranks, averaged (grad_scale,
// typically 1/world), clipped to max_norm by the GLOBAL norm, and applied with AdamW to p/m/v; the
// bf16 shadow is refreshed and g zeroed.  norm_out (may be NULL) receives the pre-clip norm.
extern "C" int llmtrain_b200_nvl_step(void* handle, float* g, float* p, float* m, float* v,
                                      void* p16, float* norm_out, float lr, float beta1,
                                      float beta2, float eps, float weight_decay, int step,
                                      float grad_scale, float max_norm, llm_stream_t stream) {
  Comm& c = *static_cast<Comm*>(handle);
  B200_REQUIRE(step >= 1, "nvl_step: step counts from 1");
  nvl::Adam a;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay;
  a.bias_corr1 = static_cast<float>(1.0 - std::pow(double(beta1), double(step)));
  a.bias_corr2_sqrt = static_cast<float>(std::sqrt(1.0 - std::pow(double(beta2), double(step))));
  a.grad_scale = grad_scale; a.max_norm = max_norm;
  return nvl_run(c, g, p, m, v, p16, nullptr, norm_out, a, 1, reinterpret_cast<cudaStream_t>(stream));
}

// Plain all-reduce (sum * scale) of an fp32 buffer of the communicator's length, bf16 on the wire:
// out[i] = scale * sum_r bf16(in_r[i]).  Used by tests and the bandwidth sweep.
extern "C" int llmtrain_b200_nvl_allreduce(void* handle, float* in, float* out, float scale,
                                           llm_stream_t stream) {
  Comm& c = *static_cast<Comm*>(handle);
  nvl::Adam a{};
  a.grad_scale = scale;
  a.max_norm = 0.f;
  a.bias_corr1 = 1.f; a.bias_corr2_sqrt = 1.f;
  return nvl_run(c, in, nullptr, nullptr, nullptr, nullptr, out, nullptr, a, 0,
                 reinterpret_cast<cudaStream_t>(stream));
}

// ------------------------------------------------------------------------------------------------
// Overlapped path: bucket reductions run on the communicator's own stream while backward continues.
// ------------------------------------------------------------------------------------------------
int b200_gpt_backward_with_hook(void* handle, cudaStream_t stream,
                                int (*hook)(void*, size_t, size_t), void* ctx);

namespace {
struct HookCtx {
  Comm* comm;
  float* g;
  cudaStream_t compute;
};

// [lo, lo+len) of the arena is final on the compute stream: pack -> barrier -> reduce + push.
int reduce_bucket(void* raw, size_t lo, size_t len) {
  HookCtx& h = *static_cast<HookCtx*>(raw);
  Comm& c = *h.comm;
  if (len == 0) return 0;
  B200_REQUIRE(lo % 8 == 0 && len % 8 == 0, "nvl: bucket bounds must be multiples of 8 elements");
  B200_REQUIRE(c.buckets_this_step < nvl::MAX_BUCKETS, "nvl: too many buckets in one step");
  const int b = c.buckets_this_step++;
  cudaStream_t cs = c.comm_stream;
  B200_CHECK_CUDA(cudaEventRecord(c.ready_events[b], h.compute));
  B200_CHECK_CUDA(cudaStreamWaitEvent(cs, c.ready_events[b], 0));
  // Ranges touching either end of the arena are the embedding tables: backward is finished.
  const bool tail = (lo == 0) || (lo + len == c.n);
  const int grid = tail ? nvl::TAIL_GRID : c.bucket_grid;
  if (c.nvls_mc != nullptr && h.g == c.nvls_uc) {
    // NVLink-SHARP mode: the arena is multicast-bound, the switch does the sum.
    if (nvl::launch_barrier(c, cs)) return 1;  // every rank's bucket is final
    size_t nsub = (len + c.world - 1) / c.world;
    nsub = (nsub + 7) / 8 * 8;
    const size_t nlo = lo + size_t(c.rank) * nsub;
    const size_t nhi = nlo + nsub < lo + len ? nlo + nsub : lo + len;
    const size_t nlen = nhi > nlo ? nhi - nlo : 0;
    float* npart = c.bucket_partials + size_t(b) * nvl::PARTIAL_STRIDE;
    B200_CHECK_CUDA(cudaMemsetAsync(npart, 0, nvl::PARTIAL_STRIDE * sizeof(float), cs));
    if (nlen > 0) {
      nvl::nvls_bucket_kernel<<<grid, nvl::OVL_THREADS, 0, cs>>>(c.nvls_mc, nlo, nlen, npart);
      B200_CHECK_CUDA(cudaGetLastError());
      B200_COUNT_LAUNCH(1);
    }
    return 0;
  }
  nvl::pack_range_kernel<<<grid, nvl::OVL_THREADS, 0, cs>>>(h.g, c.send, lo, len);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  if (nvl::launch_barrier(c, cs)) return 1;
  size_t sub = (len + c.world - 1) / c.world;
  sub = (sub + 7) / 8 * 8;
  const size_t my_lo = lo + size_t(c.rank) * sub;
  const size_t my_hi = my_lo + sub < lo + len ? my_lo + sub : lo + len;
  const size_t my_len = my_hi > my_lo ? my_hi - my_lo : 0;
  float* part = c.bucket_partials + size_t(b) * nvl::PARTIAL_STRIDE;
  B200_CHECK_CUDA(cudaMemsetAsync(part, 0, nvl::PARTIAL_STRIDE * sizeof(float), cs));
#define B200_PUSH(W)                                                                          \
  case W:                                                                                     \
    nvl::scatter_push_kernel<W><<<grid, nvl::OVL_THREADS, 0, cs>>>(                            \
        c.peers, my_lo, my_len, part, c.dma_gather ? c.rank : -1);                            \
    break;
  switch (c.world) {
    B200_PUSH(2) B200_PUSH(3) B200_PUSH(4) B200_PUSH(5) B200_PUSH(6) B200_PUSH(7) B200_PUSH(8)
    default: return fail("nvl: world size must be in [2, 8]");
  }
#undef B200_PUSH
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  if (c.dma_gather && my_len > 0) {
    // All-gather by the copy engines: the owner's reduced sub-slice goes to every peer's `full`
    // buffer without occupying issue slots on SMs that are busy with backward GEMMs.
    for (int r = 0; r < c.world; ++r) {
      if (r == c.rank) continue;
      B200_CHECK_CUDA(cudaMemcpyAsync(c.peers.full[r] + my_lo, c.full + my_lo,
                                      my_len * sizeof(float), cudaMemcpyDeviceToDevice, cs));
    }
  }
  return 0;
}
}  // namespace

// Backward of the bound GPT runtime with every gradient bucket reduced across ranks as soon as it
// is final.  Follow with llmtrain_b200_nvl_finish_step on the same stream.
extern "C" int llmtrain_b200_nvl_backward_overlapped(void* nvl_handle, void* gpt_handle, float* g,
                                                     llm_stream_t stream) {
  Comm& c = *static_cast<Comm*>(nvl_handle);
  B200_REQUIRE(c.peers_open, "nvl: peers are not connected");
  c.buckets_this_step = 0;
  HookCtx ctx{&c, g, reinterpret_cast<cudaStream_t>(stream)};
  return b200_gpt_backward_with_hook(gpt_handle, ctx.compute, &reduce_bucket, &ctx);
}

// After the last bucket: global norm from all ranks' partials, then clip + AdamW + shadow refresh
// + gradient zeroing over the locally complete reduced gradient.
extern "C" int llmtrain_b200_nvl_finish_step(void* nvl_handle, float* g, float* p, float* m,
                                             float* v, void* p16, float* norm_out, float lr,
                                             float beta1, float beta2, float eps,
                                             float weight_decay, int step, float grad_scale,
                                             float max_norm, llm_stream_t stream) {
  Comm& c = *static_cast<Comm*>(nvl_handle);
  B200_REQUIRE(step >= 1, "nvl_finish_step: step counts from 1");
  B200_REQUIRE(c.buckets_this_step > 0, "nvl_finish_step: no buckets were reduced this step");
  auto s = reinterpret_cast<cudaStream_t>(stream);
  cudaStream_t cs = c.comm_stream;
  // this rank's sum of squares = all bucket partials, fixed order
  nvl::norm_finish_kernel<<<1, 1024, 0, cs>>>(c.bucket_partials,
                                              c.buckets_this_step * nvl::PARTIAL_STRIDE,
                                             reinterpret_cast<float*>(c.sync) + nvl::NORM_SLOT);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  if (nvl::launch_barrier(c, cs)) return 1;  // every rank's pushes and norm partial are visible
  B200_CHECK_CUDA(cudaEventRecord(c.done_event, cs));
  B200_CHECK_CUDA(cudaStreamWaitEvent(s, c.done_event, 0));
  nvl::Adam a;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay;
  a.bias_corr1 = static_cast<float>(1.0 - std::pow(double(beta1), double(step)));
  a.bias_corr2_sqrt = static_cast<float>(std::sqrt(1.0 - std::pow(double(beta2), double(step))));
  a.grad_scale = grad_scale; a.max_norm = max_norm;
  // NVLink-SHARP mode: the reduced gradient is the arena itself
  const float* reduced = (c.nvls_mc != nullptr && g == c.nvls_uc) ? g : c.full;
  nvl::adamw_full_kernel<<<8 * sm_count(), nvl::THREADS, 0, s>>>(
      c.peers, c.world, c.n, reduced, p, m, v, g, static_cast<__nv_bfloat16*>(p16), norm_out, a);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  c.buckets_this_step = 0;
  return 0;
}

// Switch the overlapped path to NVLink-SHARP: `uc` / `mc` are this rank's and the team's view of a
// multicast-bound fp32 buffer of at least n elements (llmtrain_b200_nvls_buffer / _mc_buffer).
// From now on llmtrain_b200_nvl_backward_overlapped(.., g == uc, ..) reduces each bucket in the
// switch, in place on g, and llmtrain_b200_nvl_finish_step reads the reduced gradient from g.
// NULL, NULL detaches.
extern "C" int llmtrain_b200_nvl_attach_nvls(void* handle, float* uc, float* mc) {
  Comm& c = *static_cast<Comm*>(handle);
  B200_REQUIRE((uc == nullptr) == (mc == nullptr), "nvl_attach_nvls: pass both views or neither");
  c.nvls_uc = uc;
  c.nvls_mc = mc;
  return 0;
}

extern "C" const float* llmtrain_b200_nvl_reduced(void* handle, const float* g) {
  Comm& c = *static_cast<Comm*>(handle);
  return (c.nvls_mc != nullptr && g == c.nvls_uc) ? g : c.full;
}

// Small-message all-reduce: out[i] = scale * sum_r bf16(in_r[i]) for count <= 262144 elements
// (count may be smaller than the communicator's arena; the arena must hold 2 * count elements).
extern "C" int llmtrain_b200_nvl_allreduce_oneshot(void* handle, const float* in, float* out,
                                                   size_t count, float scale,
                                                   llm_stream_t stream) {
  Comm& c = *static_cast<Comm*>(handle);
  B200_REQUIRE(c.peers_open, "nvl: peers are not connected");
  B200_REQUIRE(count > 0 && count <= nvl::ONESHOT_MAX, "nvl_allreduce_oneshot: count out of range");
  B200_REQUIRE(2 * count <= c.n, "nvl_allreduce_oneshot: communicator arena too small");
  auto s = reinterpret_cast<cudaStream_t>(stream);
  const size_t half_off = (c.oneshot_calls++ & 1u) ? (c.n / 2) : 0;
  c.epoch += 1;
#define B200_ONESHOT(W)                                                                        \
  case W:                                                                                      \
    nvl::oneshot_kernel<W><<<1, 1024, 0, s>>>(c.peers, c.rank, c.epoch, half_off, in, out,     \
                                              count, scale, c.timeout_ns);                     \
    break;
  switch (c.world) {
    B200_ONESHOT(2) B200_ONESHOT(3) B200_ONESHOT(4) B200_ONESHOT(5) B200_ONESHOT(6)
    B200_ONESHOT(7) B200_ONESHOT(8)
    default: return fail("nvl: world size must be in [2, 8]");
  }
#undef B200_ONESHOT
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}

// Broadcast `count` fp32 values from `root` to every rank through the symmetric `full` buffer:
// root stages locally, flag barrier, the others pull with a copy-engine transfer from root's
// mapped buffer, flag barrier (root may then reuse the buffer).  Replaces the parameter broadcast
// DistributedDataParallel performs in its constructor.
extern "C" int llmtrain_b200_nvl_broadcast(void* handle, float* buf, size_t count, int root,
                                           llm_stream_t stream) {
  Comm& c = *static_cast<Comm*>(handle);
  B200_REQUIRE(c.peers_open, "nvl: peers are not connected");
  B200_REQUIRE(count <= c.n && root >= 0 && root < c.world, "nvl_broadcast: bad arguments");
  auto s = reinterpret_cast<cudaStream_t>(stream);
  if (c.rank == root)
    B200_CHECK_CUDA(cudaMemcpyAsync(c.full, buf, count * sizeof(float), cudaMemcpyDeviceToDevice, s));
  if (nvl::launch_barrier(c, s)) return 1;
  if (c.rank != root)
    B200_CHECK_CUDA(cudaMemcpyAsync(buf, c.peers.full[root], count * sizeof(float),
                                    cudaMemcpyDeviceToDevice, s));
  if (nvl::launch_barrier(c, s)) return 1;
  return 0;
}
