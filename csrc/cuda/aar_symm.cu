// Symmetric memory layer: ring layout, CUDA VMM allocations that can be mapped by peer processes
// (POSIX fd export/import), NVLS multicast objects, and host-mapped doorbell / completion words.
//
// Stands in for the part of Akka remoting that gives every worker an address its peers can send to
// (application.conf:5-16, AllreduceMaster.scala:70-79): here "an address" is a peer-mapped HBM region
// with the same layout on every rank, so `region[j] + offset` is how rank i names rank j's ring slot.
// The driver API is reached through cudaGetDriverEntryPoint, so the library loads on hosts without libcuda
// (the CPU-only build check).
#include <cuda.h>
#include <cuda_runtime.h>
#include <unistd.h>

#include <cstdint>
#include <cstring>

#include "aar_device.cuh"
#include "aar_host.h"
#include "akka_allreduce_b200.h"

using aar_host::fail;

namespace {

template <typename F>
bool driverFn(const char* name, F* out) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult st;
  if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &st) != cudaSuccess || st != cudaDriverEntryPointSuccess || !fn) {
    cudaGetLastError();
    return false;
  }
  *out = reinterpret_cast<F>(fn);
  return true;
}

const char* cuErr(CUresult r) {
  static thread_local char buf[64];
  snprintf(buf, sizeof buf, "CUresult %d", (int)r);
  return buf;
}

#define DRV(name, ...)                                                                           \
  using name##_fn = __VA_ARGS__;                                                                 \
  name##_fn name = nullptr;                                                                      \
  if (!driverFn(#name, &name)) return fail(AAR_EUNSUPPORTED, "driver entry point %s unavailable", #name)

uint64_t alignUp(uint64_t v, uint64_t a) { return (v + a - 1) / a * a; }

CUmemAllocationProp deviceProp(int device) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof prop);
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

}  // namespace

AAR_API const char* aar_last_error(void) { return aar_host::g_lastError.c_str(); }

AAR_API int aar_abi_sizes(int* layoutBytes, int* rankParamsBytes) {
  if (layoutBytes) *layoutBytes = (int)sizeof(AarLayout);
  if (rankParamsBytes) *rankParamsBytes = (int)sizeof(AarRankParams);
  return AAR_OK;
}

AAR_API int aar_layout_compute(int P, int dataSize, int maxChunkSize, int maxLag, int dtype, int oneshotMaxElems, int withStage,
                               AarLayout* out) {
  if (!out) return fail(AAR_EINVAL, "aar_layout_compute: null output");
  if (P <= 0 || P > AAR_MAXP) return fail(AAR_EINVAL, "aar_layout_compute: P=%d outside 1..%d", P, AAR_MAXP);
  if (dataSize <= 0 || maxChunkSize <= 0 || maxLag < 0) return fail(AAR_EINVAL, "aar_layout_compute: bad sizes");
  if (dtype != AAR_F32 && dtype != AAR_BF16) return fail(AAR_EINVAL, "aar_layout_compute: bad dtype %d", dtype);
  AarLayout l;
  memset(&l, 0, sizeof l);
  l.P = P;
  l.L = maxLag + 1;
  l.dataSize = dataSize;
  l.maxChunk = maxChunkSize;
  l.dtype = dtype;
  l.elemSize = dtype == AAR_F32 ? 4 : 2;
  const int V = 16 / l.elemSize;
  l.stepSize = (int)(((long long)dataSize + P - 1) / P);
  l.blkStride = (int)alignUp(l.stepSize, V);
  l.C = (l.stepSize + maxChunkSize - 1) / maxChunkSize;
  l.tileElems = aar_dev::kThreads * aar_dev::kUnroll * V;
  const int chunkSpan = maxChunkSize < l.stepSize ? maxChunkSize : l.stepSize;
  l.tilesPerChunk = (chunkSpan + l.tileElems - 1) / l.tileElems;
  for (int i = 0; i < P; ++i) {
    long long s = (long long)i * l.stepSize, e = (long long)(i + 1) * l.stepSize;
    if (s > dataSize) s = dataSize;
    if (e > dataSize) e = dataSize;
    l.totalChunks += (int)((e - s + maxChunkSize - 1) / maxChunkSize);
  }
  uint64_t off = 0;
  auto take = [&](uint64_t bytes) {
    uint64_t at = off;
    off = alignUp(off + bytes, 256);
    return at;
  };
  const uint64_t ring = (uint64_t)l.L * P * l.blkStride * l.elemSize;
  l.scatterOff = take(ring);
  l.reduceOff = take(ring);
  if (oneshotMaxElems > 0 && dataSize <= oneshotMaxElems) l.oneshotOff = take((uint64_t)l.L * P * alignUp(dataSize, V) * l.elemSize) + 0;
  if (withStage) l.stageOff = take((uint64_t)l.L * alignUp(dataSize, V) * l.elemSize);
  // offsets of 0 mean "absent" for the optional sections; the scatter ring legitimately sits at 0
  l.ctrlBytes = off;
  l.sflagOff = take((uint64_t)l.L * l.C * AAR_MAXP * 4);
  l.rflagOff = take((uint64_t)l.L * P * l.C * 8);
  l.peerBaseOff = take(AAR_MAXP * 4);
  l.k1DecOff = take((uint64_t)l.L * AAR_MAXP * 8);
  l.k1CtrOff = take((uint64_t)l.L * P * l.C * 4);
  l.k2DecOff = take((uint64_t)l.L * l.C * 8);
  l.k2CtrOff = take((uint64_t)l.L * l.C * 4);
  l.k3DecOff = take((uint64_t)l.L * 8);
  l.k3CtrOff = take((uint64_t)l.L * 4);
  l.countsOff = take((uint64_t)l.L * (uint64_t)P * l.C * 4);
  l.statusOff = take(16 * 4);
  l.oneshotFlagOff = take((uint64_t)l.L * AAR_MAXP * 4);
  l.barrierOff = take(8 * 4 * 16);
  l.traceOff = take((uint64_t)AAR_TRACE_CTAS * AAR_TRACE_MARKS * 8);
  l.egDecOff = take((uint64_t)l.L * P * l.C * 8);  // appended: every older offset stays where it was
  // pieces of the in-switch path: about 16 per block, 1..8 tiles each
  {
    const int tilesPerBlock = (l.stepSize + l.tileElems - 1) / l.tileElems;
    int tpp = tilesPerBlock / 16;
    l.tilesPerPiece = tpp < 1 ? 1 : (tpp > 8 ? 8 : tpp);
    l.piecesPerChunk = (l.tilesPerChunk + l.tilesPerPiece - 1) / l.tilesPerPiece;
  }
  l.pflagSOff = take((uint64_t)l.L * l.C * l.piecesPerChunk * AAR_MAXP * 4);
  l.pflagROff = take((uint64_t)l.L * P * l.C * l.piecesPerChunk * 8);
  l.slotBaseOff = take((uint64_t)l.L * AAR_MAXP * 4);
  l.regionBytes = alignUp(off, 2ull << 20);
  *out = l;
  return AAR_OK;
}

AAR_API int aar_region_alloc_local(size_t bytes, void** ptr) {
  if (!ptr) return fail(AAR_EINVAL, "aar_region_alloc_local: null");
  cudaError_t e = cudaMalloc(ptr, bytes);
  if (e != cudaSuccess) return fail(AAR_ECUDA, "cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e));
  e = cudaMemset(*ptr, 0, bytes);
  if (e != cudaSuccess) return fail(AAR_ECUDA, "cudaMemset: %s", cudaGetErrorString(e));
  return AAR_OK;
}

AAR_API int aar_region_free_local(void* ptr) {
  cudaError_t e = cudaFree(ptr);
  if (e != cudaSuccess) return fail(AAR_ECUDA, "cudaFree: %s", cudaGetErrorString(e));
  return AAR_OK;
}

// ------------------------------------------------------------------------------------------------ VMM
AAR_API int aar_vmm_granularity(int device, int multicast, int nDevices, size_t* out) {
  if (!out) return fail(AAR_EINVAL, "aar_vmm_granularity: null");
  cudaFree(0);
  DRV(cuMemGetAllocationGranularity, CUresult(*)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags));
  CUmemAllocationProp prop = deviceProp(device);
  size_t g = 0;
  CUresult r = cuMemGetAllocationGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
  if (r != CUDA_SUCCESS) return fail(AAR_ECUDA, "cuMemGetAllocationGranularity: %s", cuErr(r));
  if (multicast) {
    DRV(cuMulticastGetGranularity, CUresult(*)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags));
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof mp);
    mp.numDevices = nDevices;
    mp.size = g;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    r = cuMulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED);
    if (r != CUDA_SUCCESS) return fail(AAR_ECUDA, "cuMulticastGetGranularity: %s", cuErr(r));
    if (mg > g) g = mg;
  }
  *out = g;
  return AAR_OK;
}

static int mapAndAccess(CUmemGenericAllocationHandle h, size_t bytes, int device, void** ptr) {
  DRV(cuMemAddressReserve, CUresult(*)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long));
  DRV(cuMemMap, CUresult(*)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long));
  DRV(cuMemSetAccess, CUresult(*)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t));
  DRV(cuMemUnmap, CUresult(*)(CUdeviceptr, size_t));
  DRV(cuMemAddressFree, CUresult(*)(CUdeviceptr, size_t));
  CUdeviceptr va = 0;
  CUresult r = cuMemAddressReserve(&va, bytes, 0, 0, 0);
  if (r != CUDA_SUCCESS) return fail(AAR_ECUDA, "cuMemAddressReserve(%zu): %s", bytes, cuErr(r));
  r = cuMemMap(va, bytes, 0, h, 0);
  if (r != CUDA_SUCCESS) {
    cuMemAddressFree(va, bytes);  // failure paths give back what they took
    return fail(AAR_ECUDA, "cuMemMap: %s", cuErr(r));
  }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof acc);
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = cuMemSetAccess(va, bytes, &acc, 1);
  if (r != CUDA_SUCCESS) {
    cuMemUnmap(va, bytes);
    cuMemAddressFree(va, bytes);
    return fail(AAR_ECUDA, "cuMemSetAccess(device %d): %s", device, cuErr(r));
  }
  *ptr = reinterpret_cast<void*>(va);
  return AAR_OK;
}

AAR_API int aar_vmm_create(int device, size_t bytes, AarVmm* out) {
  if (!out || bytes == 0) return fail(AAR_EINVAL, "aar_vmm_create: bad arguments");
  cudaError_t ce = cudaSetDevice(device);
  if (ce != cudaSuccess) return fail(AAR_ECUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(ce));
  cudaFree(0);
  DRV(cuMemCreate, CUresult(*)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long));
  DRV(cuMemExportToShareableHandle, CUresult(*)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long));
  DRV(cuMemRelease, CUresult(*)(CUmemGenericAllocationHandle));
  CUmemAllocationProp prop = deviceProp(device);
  CUmemGenericAllocationHandle h = 0;
  CUresult r = cuMemCreate(&h, bytes, &prop, 0);
  if (r != CUDA_SUCCESS) return fail(AAR_ECUDA, "cuMemCreate(%zu bytes on device %d): %s (size must be a multiple of the granularity)", bytes, device, cuErr(r));
  void* ptr = nullptr;
  int rc = mapAndAccess(h, bytes, device, &ptr);
  if (rc) {
    cuMemRelease(h);
    return rc;
  }
  AarVmm partial;  // what aar_vmm_release needs to undo the steps above
  partial.handle = (uint64_t)h;
  partial.ptr = ptr;
  partial.bytes = bytes;
  partial.fd = -1;
  partial.device = device;
  ce = cudaMemset(ptr, 0, bytes);
  if (ce == cudaSuccess) ce = cudaDeviceSynchronize();
  if (ce != cudaSuccess) {
    aar_vmm_release(&partial);
    return fail(AAR_ECUDA, "zero-fill of the region failed: %s", cudaGetErrorString(ce));
  }
  int fd = -1;
  r = cuMemExportToShareableHandle(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) {
    aar_vmm_release(&partial);
    return fail(AAR_ECUDA, "cuMemExportToShareableHandle: %s", cuErr(r));
  }
  out->handle = (uint64_t)h;
  out->ptr = ptr;
  out->bytes = bytes;
  out->fd = fd;
  out->device = device;
  return AAR_OK;
}

AAR_API int aar_vmm_import(int device, int fd, size_t bytes, AarVmm* out) {
  if (!out || fd < 0) return fail(AAR_EINVAL, "aar_vmm_import: bad arguments");
  cudaFree(0);
  DRV(cuMemImportFromShareableHandle, CUresult(*)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType));
  DRV(cuMemRelease, CUresult(*)(CUmemGenericAllocationHandle));
  CUmemGenericAllocationHandle h = 0;
  CUresult r = cuMemImportFromShareableHandle(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) return fail(AAR_ECUDA, "cuMemImportFromShareableHandle(fd %d): %s", fd, cuErr(r));
  void* ptr = nullptr;
  int rc = mapAndAccess(h, bytes, device, &ptr);
  if (rc) {
    cuMemRelease(h);
    return rc;
  }
  out->handle = (uint64_t)h;
  out->ptr = ptr;
  out->bytes = bytes;
  out->fd = -1;
  out->device = device;
  return AAR_OK;
}

AAR_API int aar_vmm_release(AarVmm* v) {
  if (!v || !v->ptr) return AAR_OK;
  DRV(cuMemUnmap, CUresult(*)(CUdeviceptr, size_t));
  DRV(cuMemAddressFree, CUresult(*)(CUdeviceptr, size_t));
  DRV(cuMemRelease, CUresult(*)(CUmemGenericAllocationHandle));
  cuMemUnmap((CUdeviceptr)v->ptr, v->bytes);
  cuMemAddressFree((CUdeviceptr)v->ptr, v->bytes);
  cuMemRelease((CUmemGenericAllocationHandle)v->handle);
  if (v->fd >= 0) close(v->fd);
  v->ptr = nullptr;
  v->fd = -1;
  return AAR_OK;
}

// ------------------------------------------------------------------------------------------------ multicast (NVLS)
AAR_API int aar_mc_supported(int device) {
  int v = 0;
  using attr_fn = CUresult (*)(int*, CUdevice_attribute, CUdevice);
  attr_fn cuDeviceGetAttribute = nullptr;
  cudaFree(0);
  if (!driverFn("cuDeviceGetAttribute", &cuDeviceGetAttribute)) return 0;
  if (cuDeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, (CUdevice)device) != CUDA_SUCCESS) return 0;
  return v;
}

AAR_API int aar_mc_create(int nDevices, size_t bytes, uint64_t* mcHandle, int* fd) {
  if (!mcHandle || !fd) return fail(AAR_EINVAL, "aar_mc_create: null");
  cudaFree(0);
  DRV(cuMulticastCreate, CUresult(*)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*));
  DRV(cuMemExportToShareableHandle, CUresult(*)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long));
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof mp);
  mp.numDevices = nDevices;
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h = 0;
  CUresult r = cuMulticastCreate(&h, &mp);
  if (r != CUDA_SUCCESS) return fail(AAR_ECUDA, "cuMulticastCreate(%d devices, %zu bytes): %s", nDevices, bytes, cuErr(r));
  int f = -1;
  r = cuMemExportToShareableHandle(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) {
    using rel_fn = CUresult (*)(CUmemGenericAllocationHandle);
    rel_fn rel = nullptr;
    if (driverFn("cuMemRelease", &rel)) rel(h);
    return fail(AAR_ECUDA, "export of the multicast handle: %s", cuErr(r));
  }
  *mcHandle = (uint64_t)h;
  *fd = f;
  return AAR_OK;
}

AAR_API int aar_mc_import(int fd, uint64_t* mcHandle) {
  if (!mcHandle) return fail(AAR_EINVAL, "aar_mc_import: null");
  cudaFree(0);
  DRV(cuMemImportFromShareableHandle, CUresult(*)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType));
  CUmemGenericAllocationHandle h = 0;
  CUresult r = cuMemImportFromShareableHandle(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) return fail(AAR_ECUDA, "import of the multicast handle: %s", cuErr(r));
  *mcHandle = (uint64_t)h;
  return AAR_OK;
}

AAR_API int aar_mc_add_device(uint64_t mcHandle, int device) {
  DRV(cuMulticastAddDevice, CUresult(*)(CUmemGenericAllocationHandle, CUdevice));
  CUresult r = cuMulticastAddDevice((CUmemGenericAllocationHandle)mcHandle, (CUdevice)device);
  if (r != CUDA_SUCCESS) return fail(AAR_ECUDA, "cuMulticastAddDevice(%d): %s", device, cuErr(r));
  return AAR_OK;
}

AAR_API int aar_mc_bind_and_map(uint64_t mcHandle, int device, const AarVmm* mem, size_t bytes, void** mcPtr) {
  if (!mem || !mcPtr) return fail(AAR_EINVAL, "aar_mc_bind_and_map: null");
  DRV(cuMulticastBindMem, CUresult(*)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long));
  CUresult r = cuMulticastBindMem((CUmemGenericAllocationHandle)mcHandle, 0, (CUmemGenericAllocationHandle)mem->handle, 0, bytes, 0);
  if (r != CUDA_SUCCESS) return fail(AAR_ECUDA, "cuMulticastBindMem: %s", cuErr(r));
  return mapAndAccess((CUmemGenericAllocationHandle)mcHandle, bytes, device, mcPtr);
}

AAR_API int aar_mc_release(uint64_t mcHandle, void* mcPtr, size_t bytes) {
  DRV(cuMemUnmap, CUresult(*)(CUdeviceptr, size_t));
  DRV(cuMemAddressFree, CUresult(*)(CUdeviceptr, size_t));
  DRV(cuMemRelease, CUresult(*)(CUmemGenericAllocationHandle));
  if (mcPtr) {
    cuMemUnmap((CUdeviceptr)mcPtr, bytes);
    cuMemAddressFree((CUdeviceptr)mcPtr, bytes);
  }
  if (mcHandle) cuMemRelease((CUmemGenericAllocationHandle)mcHandle);
  return AAR_OK;
}

// ------------------------------------------------------------------------------------------------ misc
AAR_API int aar_host_words_alloc(size_t nWords, unsigned** hostPtr, unsigned** devPtr) {
  if (!hostPtr || !devPtr) return fail(AAR_EINVAL, "aar_host_words_alloc: null");
  void* h = nullptr;
  cudaError_t e = cudaHostAlloc(&h, nWords * sizeof(unsigned), cudaHostAllocMapped | cudaHostAllocPortable);
  if (e != cudaSuccess) return fail(AAR_ECUDA, "cudaHostAlloc: %s", cudaGetErrorString(e));
  memset(h, 0, nWords * sizeof(unsigned));
  void* d = nullptr;
  e = cudaHostGetDevicePointer(&d, h, 0);
  if (e != cudaSuccess) return fail(AAR_ECUDA, "cudaHostGetDevicePointer: %s", cudaGetErrorString(e));
  *hostPtr = static_cast<unsigned*>(h);
  *devPtr = static_cast<unsigned*>(d);
  return AAR_OK;
}

AAR_API int aar_host_words_free(unsigned* hostPtr) {
  cudaError_t e = cudaFreeHost(hostPtr);
  if (e != cudaSuccess) return fail(AAR_ECUDA, "cudaFreeHost: %s", cudaGetErrorString(e));
  return AAR_OK;
}

AAR_API int aar_device_sm_count(int device, int* out) {
  cudaError_t e = cudaDeviceGetAttribute(out, cudaDevAttrMultiProcessorCount, device);
  if (e != cudaSuccess) return fail(AAR_ECUDA, "cudaDeviceGetAttribute: %s", cudaGetErrorString(e));
  return AAR_OK;
}

AAR_API int aar_enable_peer_access(int device, int peer) {
  cudaError_t e = cudaSetDevice(device);
  if (e == cudaSuccess) e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    e = cudaSuccess;
  }
  if (e != cudaSuccess) return fail(AAR_ECUDA, "cudaDeviceEnablePeerAccess(%d -> %d): %s", device, peer, cudaGetErrorString(e));
  return AAR_OK;
}
