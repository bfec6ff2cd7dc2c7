// Host-side helpers: error plumbing for the C ABI, TMA tensor-map construction.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>

namespace b200 {

// ---- error reporting through the C ABI: functions return 0 on success, non-zero on failure,
// ---- and the text is available from llmtrain_b200_last_error().
inline std::string& last_error_slot() {
  thread_local std::string slot;
  return slot;
}
inline int fail(const std::string& what) {
  last_error_slot() = what;
  return 1;
}
#define B200_CHECK_CUDA(expr)                                                              \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess) {                                                               \
      return ::b200::fail(std::string(#expr) + " failed: " + cudaGetErrorString(_e) +      \
                          " (" __FILE__ ":" + std::to_string(__LINE__) + ")");             \
    }                                                                                      \
  } while (0)
#define B200_REQUIRE(cond, msg)                                                            \
  do {                                                                                     \
    if (!(cond)) return ::b200::fail(std::string(msg) + " [" #cond "]");                   \
  } while (0)

// ---- cuTensorMapEncodeTiled, resolved at run time so the library never links libcuda
// ---- (it must import on a CPU-only build box).
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                   const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess) {
      fn = reinterpret_cast<EncodeTiledFn>(sym);
    }
  });
  return fn;
}

enum class Swz { None, B128, B64 };

// 2-D row-major tensor [rows, cols] with `ld` elements between rows; box = [box_rows, box_cols].
// Out-of-bounds elements are read as zero and never written.
template <typename T>
inline int make_tmap_2d(CUtensorMap* out, const T* base, uint64_t rows, uint64_t cols, uint64_t ld,
                        uint32_t box_rows, uint32_t box_cols, Swz swz) {
  EncodeTiledFn enc = encode_tiled_fn();
  B200_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is unavailable (no CUDA driver?)");
  CUtensorMapDataType dt;
  if (sizeof(T) == 2) {
    dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  } else {
    dt = CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * sizeof(T)};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  B200_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base must be 16-byte aligned");
  B200_REQUIRE((ld * sizeof(T)) % 16 == 0, "TMA row pitch must be a multiple of 16 bytes");
  B200_REQUIRE(box_cols * sizeof(T) <= 128 || swz == Swz::None,
               "swizzled TMA box inner extent must be <= 128 bytes");
  B200_REQUIRE(swz != Swz::B64 || box_cols * sizeof(T) <= 64,
               "64-byte-swizzled TMA box inner extent must be <= 64 bytes");
  CUresult r = enc(out, dt, 2, const_cast<T*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swz == Swz::B128  ? CU_TENSOR_MAP_SWIZZLE_128B
                   : swz == Swz::B64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                     : CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    return fail("cuTensorMapEncodeTiled failed with CUresult " + std::to_string(int(r)));
  }
  return 0;
}

// 3-D tensor [d2, d1, d0] (d0 contiguous) with byte-free element pitches ld1 (between d1 rows) and
// ld2 (between d2 slabs); box = [1, box1, box0].  Used for per-sequence views [B, T, C] so that
// rows past T are zero-filled on load and clipped on store instead of bleeding into the next
// sequence.
template <typename T>
inline int make_tmap_3d(CUtensorMap* out, const T* base, uint64_t d2, uint64_t d1, uint64_t d0,
                        uint64_t ld1, uint64_t ld2, uint32_t box1, uint32_t box0, Swz swz) {
  EncodeTiledFn enc = encode_tiled_fn();
  B200_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is unavailable (no CUDA driver?)");
  CUtensorMapDataType dt =
      sizeof(T) == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  cuuint64_t dims[3] = {d0, d1, d2};
  cuuint64_t strides[2] = {ld1 * sizeof(T), ld2 * sizeof(T)};
  cuuint32_t box[3] = {box0, box1, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  B200_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base must be 16-byte aligned");
  B200_REQUIRE((ld1 * sizeof(T)) % 16 == 0 && (ld2 * sizeof(T)) % 16 == 0,
               "TMA pitches must be multiples of 16 bytes");
  CUresult r = enc(out, dt, 3, const_cast<T*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swz == Swz::B128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    return fail("cuTensorMapEncodeTiled(3d) failed with CUresult " + std::to_string(int(r)));
  }
  return 0;
}

// Number of kernels this library has launched (all threads); bench.py reports the delta over its
// timed region as "gpu_launches".
inline std::atomic<long long>& launch_counter() {
  static std::atomic<long long> counter{0};
  return counter;
}
#define B200_COUNT_LAUNCH(n) ::b200::launch_counter().fetch_add((n), std::memory_order_relaxed)

// Deterministic mode (llmtrain_b200_set_deterministic): kernels whose fast path sums floats in a
// hardware-chosen order switch to a fixed order — split-K weight gradients run unsplit, attention
// backward takes turns on dQ, the embedding scatter and the bias gradients use ordered kernels.
inline std::atomic<int>& deterministic_flag() {
  static std::atomic<int> flag{0};
  return flag;
}
inline bool deterministic() { return deterministic_flag().load(std::memory_order_relaxed) != 0; }

// Kernel launch with programmatic dependent launch (see pdl_wait in ptx.cuh) and an optional
// thread-block cluster.  LLMTRAIN_PDL=0 falls back to plain stream order (A/B measurements).
// ONLY kernels that call pdl_enter()/pdl_wait() before their first global access may come here.
inline bool pdl_enabled() {
  static const bool on = [] {
    const char* e = std::getenv("LLMTRAIN_PDL");
    return e == nullptr || std::atoi(e) != 0;
  }();
  return on;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                 cudaStream_t stream, int cluster_x, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[2];
  unsigned n = 0;
  if (cluster_x > 1) {
    attrs[n].id = cudaLaunchAttributeClusterDimension;
    attrs[n].val.clusterDim.x = unsigned(cluster_x);
    attrs[n].val.clusterDim.y = 1;
    attrs[n].val.clusterDim.z = 1;
    ++n;
  }
  if (pdl_enabled()) {
    attrs[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attrs;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#define B200_LAUNCH(kernel, grid, block, smem, stream, ...)                                    \
  B200_CHECK_CUDA(::b200::launch_kernel((kernel), dim3(grid), dim3(block), (smem), (stream), 1, \
                                        __VA_ARGS__))

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

inline int sm_count() {
  static int cached = 0;
  if (cached == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&cached, cudaDevAttrMultiProcessorCount, dev);
    if (cached <= 0) cached = 148;
  }
  return cached;
}

}  // namespace b200
