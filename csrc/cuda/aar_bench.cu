// Link micro-benchmarks: raw NVLink P2P load/store and NVLS multimem throughput as seen by one GPU while all ranks run
// the same loop.  These numbers are the measured roofline denominators for the fused kernels (profiles/).
#include <cuda_runtime.h>
#include <stdint.h>

#include "aar_device.cuh"
#include "aar_host.h"

using namespace aar_dev;

namespace {

__device__ __forceinline__ uint4 mm_ld_weak(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.weak.global.add.v4.f32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void mm_st_weak(void* mc, const uint4& v) {
  asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// mode 0: multimem.ld_reduce   1: multimem.st   2: ld_reduce + st   3: peer ld   4: peer st   5: local copy
// mode 6: ld_reduce.weak + st.weak
template <int U>
__global__ void __launch_bounds__(512, 1) linkBench(int mode, char* mc, char* peer, char* local, size_t bytes, unsigned* sink) {
  const size_t nv = bytes / 16;
  const size_t per = (nv + gridDim.x - 1) / gridDim.x;
  const size_t lo = (size_t)blockIdx.x * per, hi = (lo + per < nv) ? lo + per : nv;
  unsigned acc = 0;
  for (size_t base = lo + threadIdx.x; base < hi; base += (size_t)blockDim.x * U) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + (size_t)u * blockDim.x;
      if (i < hi) {
        if (mode == 0 || mode == 2) v[u] = multimem_ld_reduce_f32x4(mc + i * 16);
        else if (mode == 6) v[u] = mm_ld_weak(mc + i * 16);
        else if (mode == 3) v[u] = __ldcg(reinterpret_cast<const uint4*>(peer) + i);
        else if (mode == 5) v[u] = __ldcs(reinterpret_cast<const uint4*>(local) + i);
        else v[u] = make_uint4((unsigned)i, 1, 2, 3);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + (size_t)u * blockDim.x;
      if (i < hi) {
        if (mode == 1 || mode == 2) multimem_st_x4(mc + bytes + i * 16, v[u]);
        else if (mode == 6) mm_st_weak(mc + bytes + i * 16, v[u]);
        else if (mode == 4) st_peer(peer + bytes + i * 16, v[u]);
        else if (mode == 5) __stcs(reinterpret_cast<uint4*>(local + bytes) + i, v[u]);
        else acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
      }
    }
  }
  if (acc == 0x12345678u) *sink = acc;
}

struct PeerSet {
  char* ptr[16];
  int n;
};

// mode 0: every CTA stores to ONE peer (CTA b -> peer b % n), different CTAs different peers (K1-like)
// mode 1: every CTA stores the same vectors to ALL peers (K2-like broadcast); bytes = per-peer payload
// mode 2: every CTA loads from ONE peer (pull)
// mode 3: like 0 but 32-byte stores (st.global.v8.b32)
template <int U>
__global__ void __launch_bounds__(512, 1) all2allBench(int mode, PeerSet peers, char* local, size_t bytesPerPeer, unsigned* sink) {
  const int n = peers.n;
  unsigned acc = 0;
  if (mode == 0 || mode == 2 || mode == 3) {
    const int peer = blockIdx.x % n;
    const int ctasForPeer = (gridDim.x - peer + n - 1) / n;
    const int k = blockIdx.x / n;
    const size_t nv = bytesPerPeer / 16;
    const size_t per = (nv + ctasForPeer - 1) / ctasForPeer;
    const size_t lo = (size_t)k * per, hi = (lo + per < nv) ? lo + per : nv;
    char* dst = peers.ptr[peer];
    if (mode == 3) {
      for (size_t base = lo + 2 * threadIdx.x; base < hi; base += (size_t)blockDim.x * 2 * U) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const size_t i = base + (size_t)u * blockDim.x * 2;
          if (i + 1 < hi) {
            const unsigned x = (unsigned)i;
            asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst + i * 16), "r"(x), "r"(x), "r"(x), "r"(x), "r"(x),
                         "r"(x), "r"(x), "r"(x)
                         : "memory");
          }
        }
      }
      return;
    }
    for (size_t base = lo + threadIdx.x; base < hi; base += (size_t)blockDim.x * U) {
      uint4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const size_t i = base + (size_t)u * blockDim.x;
        if (i < hi) v[u] = (mode == 2) ? __ldcg(reinterpret_cast<const uint4*>(dst) + i) : __ldcs(reinterpret_cast<const uint4*>(local) + i);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const size_t i = base + (size_t)u * blockDim.x;
        if (i < hi) {
          if (mode == 0) st_peer(dst + bytesPerPeer + i * 16, v[u]);
          else acc ^= v[u].x ^ v[u].w;
        }
      }
    }
  } else {
    const size_t nv = bytesPerPeer / 16;
    const size_t per = (nv + gridDim.x - 1) / gridDim.x;
    const size_t lo = (size_t)blockIdx.x * per, hi = (lo + per < nv) ? lo + per : nv;
    for (size_t base = lo + threadIdx.x; base < hi; base += (size_t)blockDim.x * U) {
      uint4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const size_t i = base + (size_t)u * blockDim.x;
        if (i < hi) v[u] = __ldcs(reinterpret_cast<const uint4*>(local) + i);
      }
      for (int q = 0; q < n; ++q) {
        char* dst = peers.ptr[(q + blockIdx.x) % n];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const size_t i = base + (size_t)u * blockDim.x;
          if (i < hi) st_peer(dst + bytesPerPeer + i * 16, v[u]);
        }
      }
    }
  }
  if (acc == 0x12345678u) *sink = acc;
}

}  // namespace

AAR_API int aar_bench_all2all(int mode, int grid, int unroll, void** peers, int nPeers, void* local, size_t bytesPerPeer, int iters, void* stream) {
  if (mode < 0 || mode > 3 || grid <= 0 || nPeers <= 0 || nPeers > 16) return aar_host::fail(-1, "aar_bench_all2all: bad arguments");
  static unsigned* sink = nullptr;
  if (!sink) cudaMalloc(&sink, 4);
  PeerSet ps;
  ps.n = nPeers;
  for (int i = 0; i < nPeers; ++i) ps.ptr[i] = (char*)peers[i];
  for (int i = 0; i < iters; ++i) {
    if (unroll >= 8) all2allBench<8><<<grid, 512, 0, (cudaStream_t)stream>>>(mode, ps, (char*)local, bytesPerPeer, sink);
    else all2allBench<4><<<grid, 512, 0, (cudaStream_t)stream>>>(mode, ps, (char*)local, bytesPerPeer, sink);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return aar_host::fail(-2, "aar_bench_all2all: %s", cudaGetErrorString(e));
  return 0;
}

// Runs `iters` launches on `stream`; the caller times them with events.  The buffers need 2 * bytes (read half, write half).
AAR_API int aar_bench_link(int mode, int grid, int unroll, void* mc, void* peer, void* local, size_t bytes, int iters, void* stream) {
  if (mode < 0 || mode > 6 || grid <= 0) return aar_host::fail(-1, "aar_bench_link: bad arguments");
  if ((mode <= 2 || mode == 6) && !mc) return aar_host::fail(-3, "aar_bench_link: multimem modes need the multicast mapping");
  static unsigned* sink = nullptr;
  if (!sink) cudaMalloc(&sink, 4);
  for (int i = 0; i < iters; ++i) {
    if (unroll >= 16)
      linkBench<16><<<grid, 512, 0, (cudaStream_t)stream>>>(mode, (char*)mc, (char*)peer, (char*)local, bytes, sink);
    else if (unroll >= 8)
      linkBench<8><<<grid, 512, 0, (cudaStream_t)stream>>>(mode, (char*)mc, (char*)peer, (char*)local, bytes, sink);
    else if (unroll >= 4)
      linkBench<4><<<grid, 512, 0, (cudaStream_t)stream>>>(mode, (char*)mc, (char*)peer, (char*)local, bytes, sink);
    else
      linkBench<2><<<grid, 512, 0, (cudaStream_t)stream>>>(mode, (char*)mc, (char*)peer, (char*)local, bytes, sink);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return aar_host::fail(-2, "aar_bench_link: %s", cudaGetErrorString(e));
  return 0;
}
