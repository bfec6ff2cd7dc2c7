// Stand-alone dropout passes over bf16 [rows, cols] tensors, plus a mask dump for tests.
//
// Forward dropout is fused where the value is produced (GEMM epilogue for the two residual
// branches, softmax for the attention probabilities — `src/llmtrain/models/gpt.py:67,71,104`);
// this file covers the places where the masked tensor must exist in memory on its own:
//   * the embedding dropout (`gpt.py:179`), forward and backward, and
//   * backward through the two residual-branch dropouts: the masked gradient is the A operand of
//     the projection's dgrad / wgrad GEMMs, which only TMA ever touches.
// All of them regenerate the mask from (seed, offset, site, element index) — see philox.cuh.

#include "../common/host.cuh"
#include "../common/philox.cuh"
#include "../common/ptx.cuh"
#include "../../include/llmtrain_b200.h"

namespace b200 {
namespace drop {

// One thread per 8 consecutive columns (one Philox call, one 16-byte load, one 16-byte store).
__global__ void __launch_bounds__(256)
apply_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, long long rows,
             int cols, int ld, int padded_cols, const DropoutParams dp,
             const uint8_t* __restrict__ row_keep) {
  pdl_enter();  // previous kernel finished, its writes are visible
  const int groups_per_row = padded_cols >> 3;
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= rows * groups_per_row) return;
  const long long row = gid / groups_per_row;
  const int g = int(gid - row * groups_per_row);
  const int col = g * 8;
  const Rand8 r = philox_group(dp, uint64_t(row) * uint64_t(groups_per_row) + uint64_t(g));
  const __nv_bfloat16* src = x + size_t(row) * ld + col;
  __nv_bfloat16* dst = y + size_t(row) * ld + col;
  if (row_keep != nullptr && row_keep[row] == 0) {  // the whole row is dropped (padded position)
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (col + k < cols) dst[k] = __float2bfloat16(0.f);
    return;
  }
  if (col + 8 <= cols && (ld & 7) == 0) {
    const uint4 u = *reinterpret_cast<const uint4*>(src);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_bf16x2(w[j]);
      o[j] = pack_bf16x2(f.x * keep_factor(dp, r, 2 * j), f.y * keep_factor(dp, r, 2 * j + 1));
    }
    *reinterpret_cast<uint4*>(dst) = make_uint4(o[0], o[1], o[2], o[3]);
  } else {
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (col + k < cols)
        dst[k] = __float2bfloat16(__bfloat162float(src[k]) * keep_factor(dp, r, k));
  }
}

__global__ void __launch_bounds__(256)
mask_kernel(uint8_t* __restrict__ out, long long rows, int cols, int padded_cols,
            const DropoutParams dp) {
  pdl_enter();  // previous kernel finished, its writes are visible
  const int groups_per_row = padded_cols >> 3;
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= rows * groups_per_row) return;
  const long long row = gid / groups_per_row;
  const int g = int(gid - row * groups_per_row);
  const Rand8 r = philox_group(dp, uint64_t(row) * uint64_t(groups_per_row) + uint64_t(g));
#pragma unroll
  for (int k = 0; k < 8; ++k)
    if (g * 8 + k < cols) out[size_t(row) * cols + g * 8 + k] = keep_factor(dp, r, k) != 0.f;
}

}  // namespace drop
}  // namespace b200

using namespace b200;

// y = x o mask / (1 - p) over bf16 [rows, cols] (row pitch ld elements; x == y allowed).
extern "C" int llmtrain_b200_dropout(const void* x, void* y, long long rows, int cols, int ld,
                                     float p, uint64_t seed, uint64_t offset, uint32_t site,
                                     cudaStream_t stream) {
  B200_REQUIRE(rows >= 0 && cols > 0 && ld >= cols, "dropout: bad shape");
  B200_REQUIRE(p >= 0.f && p < 1.f, "dropout: p must be in [0, 1)");
  if (rows == 0) return 0;
  const int padded = (cols + 7) & ~7;
  const long long n = rows * (padded >> 3);
  const DropoutParams dp = make_dropout(p, seed, offset, site);
  B200_LAUNCH((drop::apply_kernel), unsigned((n + 255) / 256), 256, 0, stream,
              static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), rows, cols, ld,
              padded, dp, static_cast<const uint8_t*>(nullptr));
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}

// y = x o dropout_mask(p, ...) o row_keep[:, None]   (p = 0: only the row mask).  Backward of the
// attention branch when a padding mask is present: the reference multiplies that branch by the
// mask (`gpt.py:73-74`), so no gradient flows into it from padded positions.
extern "C" int llmtrain_b200_dropout_rowmask(const void* x, void* y, long long rows, int cols,
                                             int ld, const uint8_t* row_keep, float p,
                                             uint64_t seed, uint64_t offset, uint32_t site,
                                             cudaStream_t stream) {
  B200_REQUIRE(rows >= 0 && cols > 0 && ld >= cols, "dropout_rowmask: bad shape");
  B200_REQUIRE(p >= 0.f && p < 1.f, "dropout_rowmask: p must be in [0, 1)");
  B200_REQUIRE(row_keep != nullptr, "dropout_rowmask: row_keep must not be null");
  if (rows == 0) return 0;
  const int padded = (cols + 7) & ~7;
  const long long n = rows * (padded >> 3);
  const DropoutParams dp = make_dropout(p, seed, offset, site);
  B200_LAUNCH((drop::apply_kernel), unsigned((n + 255) / 256), 256, 0, stream,
              static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), rows, cols, ld,
              padded, dp, row_keep);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}

// out[rows, cols] (uint8) = 1 where the element is KEPT.  The mask every fused kernel applies for
// the same (p, seed, offset, site) on a tensor with `cols` columns — used by the tests to build
// the PyTorch reference of an op with dropout.
extern "C" int llmtrain_b200_dropout_mask(uint8_t* out, long long rows, int cols, float p,
                                          uint64_t seed, uint64_t offset, uint32_t site,
                                          cudaStream_t stream) {
  B200_REQUIRE(rows >= 0 && cols > 0, "dropout_mask: bad shape");
  if (rows == 0) return 0;
  const int padded = (cols + 7) & ~7;
  const long long n = rows * (padded >> 3);
  const DropoutParams dp = make_dropout(p, seed, offset, site);
  B200_LAUNCH((drop::mask_kernel), unsigned((n + 255) / 256), 256, 0, stream, out, rows, cols,
              padded, dp);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  return 0;
}
