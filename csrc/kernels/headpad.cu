// Head padding: run attention heads narrower than 64 on the head_dim-64 tcgen05 kernels.
//
// The reference accepts any d_model % n_heads == 0 (`src/llmtrain/models/gpt.py:21-25`) and all of
// its own presets use 16- or 32-wide heads.  Rather than a second family of attention kernels,
// the engine widens such heads to 64 lanes with zeros: zero q/k lanes add nothing to q.k, zero v
// lanes produce zero output lanes, and the matching weight rows / columns are zero so nothing
// leaks into the rest of the network.  Master weights, gradients and checkpoints keep their true
// shapes; only per-step bf16 copies of qkv_proj.weight (rows), qkv_proj.bias and out_proj.weight
// (columns) are padded, and their gradients are folded back.
//
// Both kernels see a tensor as [outer][groups][width][inner] and change `width` (hd <-> 64):
//   rows of a [G*hd, cols] matrix :  outer = 1,    inner = cols
//   cols of a [rows, G*hd] matrix :  outer = rows, inner = 1
//   a [G*hd] vector               :  outer = 1,    inner = 1

#include "../common/host.cuh"
#include "../common/ptx.cuh"
#include "../../include/llmtrain_b200.h"

namespace b200 {
namespace headpad {

constexpr int PADW = 64;

template <typename T>
__global__ void __launch_bounds__(256)
pad_kernel(const T* __restrict__ src, T* __restrict__ dst, long long outer, int groups, int hd,
           long long inner) {
  pdl_enter();  // previous kernel finished, its writes are visible
  const long long total = outer * groups * PADW * inner;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long in = i % inner;
    const long long t = i / inner;
    const int j = int(t % PADW);
    const long long og = t / PADW;  // outer * groups + group
    dst[i] = j < hd ? src[(og * hd + j) * inner + in] : T(0.f);
  }
}

// dst[true shape] += src[padded shape]   (fp32 gradients)
__global__ void __launch_bounds__(256)
unpad_add_kernel(const float* __restrict__ src, float* __restrict__ dst, long long outer,
                 int groups, int hd, long long inner) {
  pdl_enter();  // previous kernel finished, its writes are visible
  const long long total = outer * groups * hd * inner;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long in = i % inner;
    const long long t = i / inner;
    const int j = int(t % hd);
    const long long og = t / hd;
    dst[i] += src[(og * PADW + j) * inner + in];
  }
}

inline unsigned grid_for(long long total) {
  long long g = (total + 255) / 256;
  const long long cap = 8LL * sm_count();
  return unsigned(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace headpad
}  // namespace b200

using namespace b200;

// dst[outer][groups][64][inner] = src[outer][groups][hd][inner] with zeros in lanes hd..63.
// elem_bytes: 2 (bf16) or 4 (fp32).
extern "C" int llmtrain_b200_head_pad(const void* src, void* dst, long long outer, int groups,
                                      int hd, long long inner, int elem_bytes,
                                      cudaStream_t stream) {
  B200_REQUIRE(hd > 0 && hd <= headpad::PADW && groups > 0 && outer > 0 && inner > 0,
               "head_pad: bad shape");
  const long long total = outer * groups * headpad::PADW * inner;
  if (elem_bytes == 2) {
    B200_LAUNCH((headpad::pad_kernel<__nv_bfloat16>), headpad::grid_for(total), 256, 0, stream,
                static_cast<const __nv_bfloat16*>(src), static_cast<__nv_bfloat16*>(dst), outer,
                groups, hd, inner);
  } else if (elem_bytes == 4) {
    B200_LAUNCH((headpad::pad_kernel<float>), headpad::grid_for(total), 256, 0, stream,
                static_cast<const float*>(src), static_cast<float*>(dst), outer, groups, hd, inner);
  } else {
    return fail("head_pad: elem_bytes must be 2 or 4");
  }
  B200_COUNT_LAUNCH(1);
  return 0;
}

// dst[outer][groups][hd][inner] += src[outer][groups][64][inner]   (fp32)
extern "C" int llmtrain_b200_head_unpad_add(const float* src, float* dst, long long outer,
                                            int groups, int hd, long long inner,
                                            cudaStream_t stream) {
  B200_REQUIRE(hd > 0 && hd <= headpad::PADW && groups > 0 && outer > 0 && inner > 0,
               "head_unpad_add: bad shape");
  const long long total = outer * groups * hd * inner;
  B200_LAUNCH((headpad::unpad_add_kernel), headpad::grid_for(total), 256, 0, stream, src, dst, outer,
              groups, hd, inner);
  B200_COUNT_LAUNCH(1);
  return 0;
}
