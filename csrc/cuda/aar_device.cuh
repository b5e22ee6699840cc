// Device-side building blocks of the sm_100a data plane: memory-model helpers (system-scope
// release/acquire over NVLink), geometry, 16-byte vector tiles with fp32 accumulation.
//
// Why explicit release/acquire: the reference relies on Akka's per-pair FIFO ("reduce t would never
// come after reduce t+1", AllreduceSpec.scala:520,666).  NVLink stores from one GPU to another carry no
// such ordering, so every "message" here is data stores followed by a st.release.sys of an epoch-tagged
// flag, and every receive is a ld.acquire.sys poll of that flag.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "akka_allreduce_b200.h"

namespace aar_dev {

constexpr int kThreads = 512;
constexpr int kUnroll = 4;  // 16-byte vectors per thread per tile (tile = kThreads * kUnroll vectors = 32 KiB)
constexpr int kCopyUnroll = 8;  // vectors in flight per thread in the copy loops (64 KiB per CTA)

// ---------------------------------------------------------------- memory model helpers
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t ld_acquire_sys_u64(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t ld_relaxed_sys_u64(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t ld_acquire_gpu_u64(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_release_sys_u64(uint64_t* p, uint64_t v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_release_gpu_u64(uint64_t* p, uint64_t v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// 16-byte accesses.  Ring data was written by other GPUs: read it through L2 only (.cg) so a stale L1 line
// can never be observed; source data is streamed (.cs); remote stores are plain st.global (NVLink posted writes).
__device__ __forceinline__ uint4 ld_ring(const void* p) { return __ldcg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ uint4 ld_stream(const void* p) { return __ldcs(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ void st_peer(void* p, const uint4& v) {
  asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void st_stream(void* p, const uint4& v) { __stcs(reinterpret_cast<uint4*>(p), v); }

// NVLS: one load returns the sum over every GPU bound to the multicast object (the reduction happens in the
// NVSwitch), one store lands in every GPU's copy.  Accumulation of bf16 is in fp32 inside the switch.
__device__ __forceinline__ uint4 multimem_ld_reduce_f32x4(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_x4(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ---------------------------------------------------------------- element traits
template <typename T>
struct Elem;
template <>
struct Elem<float> {
  static constexpr int kPerVec = 4;
  __device__ static void unpack(const uint4& v, float* f) {
    f[0] = __uint_as_float(v.x);
    f[1] = __uint_as_float(v.y);
    f[2] = __uint_as_float(v.z);
    f[3] = __uint_as_float(v.w);
  }
  __device__ static uint4 pack(const float* f) {
    return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
  }
  __device__ static float toFloat(float v) { return v; }
  __device__ static float fromFloat(float v) { return v; }
};
template <>
struct Elem<__nv_bfloat16> {
  static constexpr int kPerVec = 8;
  __device__ static void unpack(const uint4& v, float* f) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      f[2 * i] = __uint_as_float(w[i] << 16);
      f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  __device__ static uint4 pack(const float* f) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
  __device__ static float toFloat(__nv_bfloat16 v) { return __bfloat162float(v); }
  __device__ static __nv_bfloat16 fromFloat(float v) { return __float2bfloat16_rn(v); }
};

// ---------------------------------------------------------------- geometry (mirrors csrc/core/geometry.h)
__device__ __forceinline__ int blockStart(const AarLayout& l, int i) {
  long long s = (long long)i * l.stepSize;
  return (int)(s < l.dataSize ? s : l.dataSize);
}
__device__ __forceinline__ int blockLen(const AarLayout& l, int i) { return blockStart(l, i + 1) - blockStart(l, i); }
__device__ __forceinline__ int numChunks(const AarLayout& l, int i) { return (blockLen(l, i) + l.maxChunk - 1) / l.maxChunk; }
__device__ __forceinline__ int chunkLen(const AarLayout& l, int block, int c) {
  int r = blockLen(l, block) - c * l.maxChunk;
  return r < 0 ? 0 : (r > l.maxChunk ? l.maxChunk : r);
}
__device__ __forceinline__ int numTiles(const AarLayout& l, int len) { return (len + l.tileElems - 1) / l.tileElems; }

// ---------------------------------------------------------------- waits
struct Deadline {
  unsigned long long start, budget;
  __device__ explicit Deadline(unsigned long long timeoutNs) : start(timeoutNs ? globaltimer_ns() : 0ull), budget(timeoutNs) {}
  __device__ bool expired() const { return budget != 0 && globaltimer_ns() - start > budget; }
};

// True when this round must complete with whatever has arrived (the catch-up of AllreduceWorker.scala:91-97):
//   * AAR_FLAG_FORCE, or
//   * the master's newest round (host doorbell) left this round outside the maxLag window, or
//   * a peer is known to have FINISHED a round more than maxLag ahead of this one.  That is the device-side form of the
//     reference's peer-ahead self-start (AllreduceWorker.scala:123-126): progress of the peers reaches a worker through
//     the data plane (their retire words land in my region), not only through the master.
__device__ __forceinline__ bool forcedByHost(const AarRankParams& p, const AarLayout& l) {
  if (p.flags & AAR_FLAG_FORCE) return true;
  const uint32_t* base = reinterpret_cast<const uint32_t*>(reinterpret_cast<const char*>(p.region[p.rank]) + l.peerBaseOff);
  int newestDone = -1;
  for (int j = 0; j < l.P; ++j) {
    const int done = (int)ld_relaxed_sys_u32(base + j) - 1;  // newest round peer j has retired
    newestDone = done > newestDone ? done : newestDone;
  }
  if (newestDone - p.round > p.maxLag) return true;
  if (p.doorbell) {
    int newest = (int)ld_relaxed_sys_u32(reinterpret_cast<const uint32_t*>(p.doorbell));
    if (newest - p.round > p.maxLag) return true;
  }
  return false;
}

// ---------------------------------------------------------------- tile movers
// Copy `n` elements; the 16-byte path is taken only when both pointers and the tile origin are 16-byte aligned.
// kRemoteDst: destination is a peer GPU's ring (plain st.global); kRingSrc: source was written by peers (L2-only loads).
template <typename T, bool kRemoteDst, bool kRingSrc>
__device__ __forceinline__ void copyTile(T* dst, const T* src, int n, bool vecOk) {
  constexpr int V = Elem<T>::kPerVec;
  const int nv = vecOk ? n / V : 0;
  for (int base = threadIdx.x; base < nv; base += kThreads * kCopyUnroll) {
    uint4 v[kCopyUnroll];
#pragma unroll
    for (int u = 0; u < kCopyUnroll; ++u) {
      const int i = base + u * kThreads;
      if (i < nv) v[u] = kRingSrc ? ld_ring(reinterpret_cast<const uint4*>(src) + i) : ld_stream(reinterpret_cast<const uint4*>(src) + i);
    }
#pragma unroll
    for (int u = 0; u < kCopyUnroll; ++u) {
      const int i = base + u * kThreads;
      if (i < nv) {
        if (kRemoteDst)
          st_peer(reinterpret_cast<uint4*>(dst) + i, v[u]);
        else
          st_stream(reinterpret_cast<uint4*>(dst) + i, v[u]);
      }
    }
  }
  for (int i = nv * V + threadIdx.x; i < n; i += kThreads) dst[i] = kRingSrc ? __ldcg(src + i) : src[i];
}

template <typename T>
__device__ __forceinline__ void zeroTile(T* dst, int n, bool vecOk) {
  constexpr int V = Elem<T>::kPerVec;
  if (vecOk) {
    const int nv = n / V;
    for (int i = threadIdx.x; i < nv; i += kThreads) st_stream(reinterpret_cast<uint4*>(dst) + i, make_uint4(0, 0, 0, 0));
    for (int i = nv * V + threadIdx.x; i < n; i += kThreads) dst[i] = Elem<T>::fromFloat(0.f);
  } else {
    for (int i = threadIdx.x; i < n; i += kThreads) dst[i] = Elem<T>::fromFloat(0.f);
  }
}

}  // namespace aar_dev
