// NVLink-SHARP (NVLS) all-reduce: the NVSwitch adds the ranks' values in flight.
//
// Replaces, for large gradient buckets, what the reference gets from NCCL's ring/tree all-reduce
// inside DistributedDataParallel (`src/llmtrain/training/trainer.py:87-91`; SURVEY work item W9).
// Every rank binds one physical allocation of its own into a MULTICAST object shared by the team;
// through the multicast address
//     multimem.ld_reduce.add.v4.f32  returns sum_r buffer_r[i]   (one load, reduced in the switch)
//     multimem.st.v4.f32             writes a value into buffer_r[i] of EVERY rank
// so an all-reduce is: each rank walks its 1/W slice with ld_reduce -> scale -> st.  Per GPU that
// is N/W floats pulled and N/W floats pushed over NVLink instead of 2 * N * (W-1)/W in the
// two-shot peer-to-peer scheme of nvl.cu.
//
// Setup needs the driver's virtual-memory API (resolved at run time, the library does not link
// libcuda): rank 0 creates the multicast object and exports it (fabric handle, 64 bytes, shipped
// by the caller through torch.distributed; or a POSIX fd picked up with pidfd_getfd when fabric
// handles are not available); every rank adds its device, binds its memory and maps both the
// unicast and the multicast view.  Cross-GPU ordering uses the flag barrier of the peer-to-peer
// communicator (llmtrain_b200_nvl_barrier).

#include <sys/syscall.h>
#include <unistd.h>

#include <cstring>

#include "../common/host.cuh"
#include "../../include/llmtrain_b200.h"

namespace b200 {
namespace nvls {

template <typename Fn>
inline Fn driver_fn(const char* name) {
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint(name, &sym, cudaEnableDefault, &qres) != cudaSuccess ||
      qres != cudaDriverEntryPointSuccess)
    return nullptr;
  return reinterpret_cast<Fn>(sym);
}

struct Driver {
  CUresult (*memCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*,
                        unsigned long long) = nullptr;
  CUresult (*memRelease)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*addrReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*addrFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*memMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle,
                     unsigned long long) = nullptr;
  CUresult (*memUnmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*setAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*exportHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType,
                           unsigned long long) = nullptr;
  CUresult (*importHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) =
      nullptr;
  CUresult (*mcCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
  CUresult (*mcAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
  CUresult (*mcBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t,
                        size_t, unsigned long long) = nullptr;
  CUresult (*mcUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
  CUresult (*mcGranularity)(size_t*, const CUmulticastObjectProp*,
                            CUmulticastGranularity_flags) = nullptr;
  CUresult (*deviceGet)(CUdevice*, int) = nullptr;
  CUresult (*deviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
  bool ok = false;
};

inline const Driver& driver() {
  static Driver d = [] {
    Driver x;
    x.memCreate = driver_fn<decltype(x.memCreate)>("cuMemCreate");
    x.memRelease = driver_fn<decltype(x.memRelease)>("cuMemRelease");
    x.addrReserve = driver_fn<decltype(x.addrReserve)>("cuMemAddressReserve");
    x.addrFree = driver_fn<decltype(x.addrFree)>("cuMemAddressFree");
    x.memMap = driver_fn<decltype(x.memMap)>("cuMemMap");
    x.memUnmap = driver_fn<decltype(x.memUnmap)>("cuMemUnmap");
    x.setAccess = driver_fn<decltype(x.setAccess)>("cuMemSetAccess");
    x.exportHandle = driver_fn<decltype(x.exportHandle)>("cuMemExportToShareableHandle");
    x.importHandle = driver_fn<decltype(x.importHandle)>("cuMemImportFromShareableHandle");
    x.mcCreate = driver_fn<decltype(x.mcCreate)>("cuMulticastCreate");
    x.mcAddDevice = driver_fn<decltype(x.mcAddDevice)>("cuMulticastAddDevice");
    x.mcBindMem = driver_fn<decltype(x.mcBindMem)>("cuMulticastBindMem");
    x.mcUnbind = driver_fn<decltype(x.mcUnbind)>("cuMulticastUnbind");
    x.mcGranularity = driver_fn<decltype(x.mcGranularity)>("cuMulticastGetGranularity");
    x.deviceGet = driver_fn<decltype(x.deviceGet)>("cuDeviceGet");
    x.deviceGetAttribute = driver_fn<decltype(x.deviceGetAttribute)>("cuDeviceGetAttribute");
    x.ok = x.memCreate && x.memRelease && x.addrReserve && x.addrFree && x.memMap && x.memUnmap &&
           x.setAccess && x.exportHandle && x.importHandle && x.mcCreate && x.mcAddDevice &&
           x.mcBindMem && x.mcUnbind && x.mcGranularity && x.deviceGet && x.deviceGetAttribute;
    return x;
  }();
  return d;
}

#define B200_CHECK_CU(call)                                                                   \
  do {                                                                                        \
    CUresult r_ = (call);                                                                     \
    if (r_ != CUDA_SUCCESS)                                                                   \
      return ::b200::fail(std::string(#call) + " failed with CUresult " + std::to_string(int(r_))); \
  } while (0)

struct Team {
  void* nvl = nullptr;  // peer-to-peer communicator that supplies the cross-GPU barrier
  int rank = 0, world = 1, device = 0;
  CUdevice cu_device = 0;
  size_t n = 0;       // fp32 elements usable by callers
  size_t bytes = 0;   // bound size (multiple of the multicast granularity)
  CUmemAllocationHandleType handle_type = CU_MEM_HANDLE_TYPE_FABRIC;
  CUmemGenericAllocationHandle mc = 0, mem = 0;
  bool have_mc = false, have_mem = false, bound = false;
  CUdeviceptr ucast = 0, mcast = 0;
  int export_fd = -1;
};

struct Wire {  // what rank 0 hands to the team (fits the 192-byte slots the caller already ships)
  int kind;            // 0 = fabric handle, 1 = (pid, fd) of a POSIX file descriptor
  int pid, fd, pad;
  unsigned char fabric[64];
};

inline CUmulticastObjectProp mc_prop(const Team& t) {
  CUmulticastObjectProp p;
  std::memset(&p, 0, sizeof(p));
  p.numDevices = unsigned(t.world);
  p.size = t.bytes;
  p.handleTypes = t.handle_type;
  return p;
}

// 16-byte multicast load-reduce + store over this rank's slice; four independent requests per
// thread and iteration keep enough bytes in flight to fill the NVLink pipes.
__device__ __forceinline__ float4 mc_ld_reduce(const float* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void mc_st(float* p, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__global__ void __launch_bounds__(512)
allreduce_kernel(float* __restrict__ mc, size_t lo, size_t len, float scale) {
  const size_t n4 = len >> 2;
  float* base = mc + lo;
  const size_t stride = size_t(gridDim.x) * blockDim.x;
  size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n4; i += 4 * stride) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = mc_ld_reduce(base + 4 * (i + u * stride));
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      v[u].x *= scale; v[u].y *= scale; v[u].z *= scale; v[u].w *= scale;
      mc_st(base + 4 * (i + u * stride), v[u]);
    }
  }
  for (; i < n4; i += stride) {
    float4 v = mc_ld_reduce(base + 4 * i);
    v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
    mc_st(base + 4 * i, v);
  }
}

}  // namespace nvls
}  // namespace b200

using namespace b200;
using nvls::Team;

// 1 if the current device reports multicast support and the driver entry points resolve.
extern "C" int llmtrain_b200_nvls_supported(void) {
  const nvls::Driver& d = nvls::driver();
  if (!d.ok) return 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  CUdevice cu;
  if (d.deviceGet(&cu, dev) != CUDA_SUCCESS) return 0;
  int v = 0;
  if (d.deviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cu) != CUDA_SUCCESS) return 0;
  return v != 0;
}

// Step 1 (all ranks).  n = fp32 elements; rank 0 also creates the multicast object.
// use_fabric: 1 = share it as a fabric handle, 0 = as a POSIX fd picked up through pidfd_getfd.
extern "C" int llmtrain_b200_nvls_create(void* nvl_handle, int rank, int world, size_t n,
                                         int use_fabric, void** handle) {
  B200_REQUIRE(handle != nullptr && nvl_handle != nullptr, "nvls_create: null handle");
  B200_REQUIRE(world >= 2 && world <= 8 && rank >= 0 && rank < world, "nvls_create: bad team");
  B200_REQUIRE(n > 0 && n % (size_t(4) * world) == 0,
               "nvls_create: n must be a positive multiple of 4 * world");
  const nvls::Driver& d = nvls::driver();
  B200_REQUIRE(d.ok, "nvls_create: the driver's multicast / virtual-memory API is unavailable");
  auto* t = new Team();
  t->nvl = nvl_handle;
  t->rank = rank; t->world = world; t->n = n;
  B200_CHECK_CUDA(cudaGetDevice(&t->device));
  B200_CHECK_CUDA(cudaFree(nullptr));  // make sure the primary context exists
  B200_CHECK_CU(d.deviceGet(&t->cu_device, t->device));
  t->handle_type = use_fabric ? CU_MEM_HANDLE_TYPE_FABRIC : CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  t->bytes = n * sizeof(float);
  CUmulticastObjectProp p = nvls::mc_prop(*t);
  size_t gran = 0;
  B200_CHECK_CU(d.mcGranularity(&gran, &p, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  t->bytes = (t->bytes + gran - 1) / gran * gran;
  if (rank == 0) {
    p = nvls::mc_prop(*t);
    B200_CHECK_CU(d.mcCreate(&t->mc, &p));
    t->have_mc = true;
  }
  *handle = t;
  return 0;
}

// Step 2 (rank 0): 80 bytes describing the multicast object, to be broadcast by the caller.
extern "C" int llmtrain_b200_nvls_export(void* handle, void* out_80_bytes) {
  Team& t = *static_cast<Team*>(handle);
  B200_REQUIRE(t.rank == 0 && t.have_mc, "nvls_export: only rank 0 owns the multicast object");
  const nvls::Driver& d = nvls::driver();
  nvls::Wire w;
  std::memset(&w, 0, sizeof(w));
  if (t.handle_type == CU_MEM_HANDLE_TYPE_FABRIC) {
    w.kind = 0;
    B200_CHECK_CU(d.exportHandle(w.fabric, t.mc, CU_MEM_HANDLE_TYPE_FABRIC, 0));
  } else {
    int fd = -1;
    B200_CHECK_CU(d.exportHandle(&fd, t.mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    t.export_fd = fd;
    w.kind = 1;
    w.pid = int(getpid());
    w.fd = fd;
  }
  std::memcpy(out_80_bytes, &w, sizeof(w));
  return 0;
}

// Step 3 (all ranks): import rank 0's object (no-op on rank 0) and add this device to it.
extern "C" int llmtrain_b200_nvls_connect(void* handle, const void* wire_80_bytes) {
  Team& t = *static_cast<Team*>(handle);
  const nvls::Driver& d = nvls::driver();
  if (t.rank != 0) {
    nvls::Wire w;
    std::memcpy(&w, wire_80_bytes, sizeof(w));
    if (w.kind == 0) {
      B200_CHECK_CU(d.importHandle(&t.mc, w.fabric, CU_MEM_HANDLE_TYPE_FABRIC));
    } else {
#if defined(SYS_pidfd_open) && defined(SYS_pidfd_getfd)
      const int pidfd = int(syscall(SYS_pidfd_open, w.pid, 0));
      B200_REQUIRE(pidfd >= 0, "nvls_connect: pidfd_open on rank 0's process failed");
      const int fd = int(syscall(SYS_pidfd_getfd, pidfd, w.fd, 0));
      close(pidfd);
      B200_REQUIRE(fd >= 0, "nvls_connect: pidfd_getfd failed (ptrace permission?)");
      const CUresult r = d.importHandle(&t.mc, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                        CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
      close(fd);
      if (r != CUDA_SUCCESS)
        return fail("nvls_connect: cuMemImportFromShareableHandle failed with CUresult " +
                    std::to_string(int(r)));
#else
      return fail("nvls_connect: pidfd syscalls are not available on this platform");
#endif
    }
    t.have_mc = true;
  }
  B200_CHECK_CU(d.mcAddDevice(t.mc, t.cu_device));
  return 0;
}

// Step 4 (all ranks, after a host-side barrier: every device must have been added): allocate this
// rank's memory, bind it, map the unicast and the multicast view.  Another host-side barrier must
// follow before the first collective.
extern "C" int llmtrain_b200_nvls_bind(void* handle) {
  Team& t = *static_cast<Team*>(handle);
  const nvls::Driver& d = nvls::driver();
  CUmemAllocationProp ap;
  std::memset(&ap, 0, sizeof(ap));
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = t.device;
  ap.requestedHandleTypes = t.handle_type;
  B200_CHECK_CU(d.memCreate(&t.mem, t.bytes, &ap, 0));
  t.have_mem = true;
  B200_CHECK_CU(d.mcBindMem(t.mc, 0, t.mem, 0, t.bytes, 0));
  t.bound = true;
  CUmemAccessDesc acc;
  std::memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = t.device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  B200_CHECK_CU(d.addrReserve(&t.ucast, t.bytes, 0, 0, 0));
  B200_CHECK_CU(d.memMap(t.ucast, t.bytes, 0, t.mem, 0));
  B200_CHECK_CU(d.setAccess(t.ucast, t.bytes, &acc, 1));
  B200_CHECK_CU(d.addrReserve(&t.mcast, t.bytes, 0, 0, 0));
  B200_CHECK_CU(d.memMap(t.mcast, t.bytes, 0, t.mc, 0));
  B200_CHECK_CU(d.setAccess(t.mcast, t.bytes, &acc, 1));
  B200_CHECK_CUDA(cudaMemset(reinterpret_cast<void*>(t.ucast), 0, t.bytes));
  B200_CHECK_CUDA(cudaDeviceSynchronize());
  return 0;
}

// This rank's view of the registered buffer (fp32 [n]); collectives operate on it in place.
extern "C" void* llmtrain_b200_nvls_buffer(void* handle) {
  return reinterpret_cast<void*>(static_cast<Team*>(handle)->ucast);
}

// The team's multicast view of the same buffer (only multimem.* instructions may touch it).
extern "C" void* llmtrain_b200_nvls_mc_buffer(void* handle) {
  return reinterpret_cast<void*>(static_cast<Team*>(handle)->mcast);
}

// buffer_r[i] := scale * sum_r buffer_r[i] on every rank, for i < count (count % (4*world) == 0).
// If `in` is not NULL it is first copied into the buffer, if `out` is not NULL the result is
// copied out (both fp32 [count]); pass NULL / NULL to time the collective alone.
extern "C" int llmtrain_b200_nvls_allreduce(void* handle, const float* in, float* out, size_t count,
                                            float scale, cudaStream_t stream) {
  Team& t = *static_cast<Team*>(handle);
  B200_REQUIRE(t.bound, "nvls_allreduce: llmtrain_b200_nvls_bind has not run");
  B200_REQUIRE(count > 0 && count <= t.n && count % (size_t(4) * t.world) == 0,
               "nvls_allreduce: count must be a multiple of 4 * world within the buffer");
  float* mine = reinterpret_cast<float*>(t.ucast);
  if (in != nullptr)
    B200_CHECK_CUDA(cudaMemcpyAsync(mine, in, count * sizeof(float), cudaMemcpyDeviceToDevice,
                                    stream));
  // every rank's input is in place before anyone reduces ...
  if (llmtrain_b200_nvl_barrier(t.nvl, stream)) return 1;
  const size_t chunk = count / t.world;
  const size_t n4 = chunk >> 2;
  long long blocks = (long long)((n4 + 511) / 512);
  const long long cap = 4LL * sm_count();
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  nvls::allreduce_kernel<<<unsigned(blocks), 512, 0, stream>>>(
      reinterpret_cast<float*>(t.mcast), size_t(t.rank) * chunk, chunk, scale);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_COUNT_LAUNCH(1);
  // ... and every slice has been broadcast before anyone reads the result
  if (llmtrain_b200_nvl_barrier(t.nvl, stream)) return 1;
  if (out != nullptr)
    B200_CHECK_CUDA(cudaMemcpyAsync(out, mine, count * sizeof(float), cudaMemcpyDeviceToDevice,
                                    stream));
  return 0;
}

extern "C" void llmtrain_b200_nvls_destroy(void* handle) {
  if (handle == nullptr) return;
  Team* t = static_cast<Team*>(handle);
  const nvls::Driver& d = nvls::driver();
  cudaDeviceSynchronize();
  if (t->mcast) { d.memUnmap(t->mcast, t->bytes); d.addrFree(t->mcast, t->bytes); }
  if (t->ucast) { d.memUnmap(t->ucast, t->bytes); d.addrFree(t->ucast, t->bytes); }
  if (t->bound) d.mcUnbind(t->mc, t->cu_device, 0, t->bytes);
  if (t->have_mem) d.memRelease(t->mem);
  if (t->have_mc) d.memRelease(t->mc);
  if (t->export_fd >= 0) close(t->export_fd);
  delete t;
}
