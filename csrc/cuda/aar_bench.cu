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

// ------------------------------------------------------------------------------------------------ TMA bulk copy
// A copy role that holds no data in registers: ONE thread per CTA drives a ring of shared-memory tiles with the bulk
// async-copy engine -- cp.async.bulk global -> shared (completion counted in bytes on an mbarrier), then cp.async.bulk
// shared -> (peer) global as a bulk group.  The fused kernels' copy roles (stage, gather, push) spend 512 threads x 128
// registers per SM on the same job; this measures what a 32-thread CTA can move, to decide whether those roles should be
// rebuilt on it (it would free SMs for the caller's compute).  Written without GPU access: every wait is bounded and
// reports through `err` instead of hanging.
namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ bool mbar_wait(uint32_t bar, uint32_t parity) {
  for (unsigned tries = 0; tries < (1u << 20); ++tries) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) return true;
  }
  return false;
}

// blockDim.x = 32 * W: lane 0 of every warp drives its own ring of kStages tiles (its own slice of shared memory, its own
// barriers, every W-th tile of the CTA's share).  W > 1 answers whether the cost of issuing a bulk copy (a few hundred
// cycles of the issuing thread) is a per-thread latency that several issuers hide, or a limit of the SM's copy engine.
template <int kStages>
__global__ void __launch_bounds__(256, 1) bulkCopyBench(const char* src, char* dst, size_t bytes, int tileBytes, unsigned* err) {
  extern __shared__ __align__(128) unsigned char smemAll[];
  const int W = blockDim.x / 32, w = threadIdx.x / 32;
  unsigned char* smem = smemAll + (size_t)w * ((size_t)kStages * tileBytes + 64);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)kStages * tileBytes);
  if ((threadIdx.x & 31) != 0) return;  // one thread per warp issues everything; the copy engine does the work
  for (int s = 0; s < kStages; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bars + s)) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");  // make the initialised barriers visible to the async proxy
  const size_t nTiles = (bytes + tileBytes - 1) / tileBytes;
  const size_t me = (size_t)blockIdx.x * W + w, lanes = (size_t)gridDim.x * W;
  const size_t mine = nTiles > me ? (nTiles - me - 1) / lanes + 1 : 0;
  auto tileOff = [&](size_t k) { return (me + k * lanes) * (size_t)tileBytes; };
  auto tileLen = [&](size_t k) {
    const size_t off = tileOff(k);
    return (uint32_t)(bytes - off < (size_t)tileBytes ? bytes - off : (size_t)tileBytes);
  };
  auto load = [&](size_t k) {
    const int s = (int)(k % kStages);
    const uint32_t bar = smem_u32(bars + s), n = tileLen(k);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(n) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem + (size_t)s * tileBytes)),
                 "l"(src + tileOff(k)), "r"(n), "r"(bar)
                 : "memory");
  };
  for (size_t k = 0; k < mine && k < (size_t)kStages; ++k) load(k);
  for (size_t k = 0; k < mine; ++k) {
    const int s = (int)(k % kStages);
    if (!mbar_wait(smem_u32(bars + s), (uint32_t)((k / kStages) & 1))) {
      atomicExch(err, 1u);
      return;
    }
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + tileOff(k)), "r"(smem_u32(smem + (size_t)s * tileBytes)), "r"(tileLen(k))
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    // the store of tile k-1 has finished READING its stage once at most one newer group is pending: refill that stage
    if (k >= 1 && k - 1 + kStages < mine) {
      asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      load(k - 1 + kStages);
    }
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // writes complete (not only their reads) before the CTA exits
}

}  // namespace

AAR_API int aar_bench_bulk_copy_multi(int grid, int warps, int stages, int tileBytes, const void* src, void* dst, size_t bytes, int iters, void* stream,
                                      unsigned* errOut);

// `iters` launches of the bulk-copy kernel on `stream` (timed by the caller).  src / dst: `bytes` each, 16-byte aligned,
// bytes a multiple of 16; dst may be a peer mapping.  stages in {2, 4, 8}; tileBytes a multiple of 128 with
// stages * tileBytes + 64 <= 227 KiB.  *errOut (device word, may be NULL) becomes 1 if a wait ran out of patience.
AAR_API int aar_bench_bulk_copy(int grid, int stages, int tileBytes, const void* src, void* dst, size_t bytes, int iters, void* stream, unsigned* errOut) {
  return aar_bench_bulk_copy_multi(grid, 1, stages, tileBytes, src, dst, bytes, iters, stream, errOut);
}

AAR_API int aar_bench_bulk_copy_multi(int grid, int warps, int stages, int tileBytes, const void* src, void* dst, size_t bytes, int iters, void* stream,
                                      unsigned* errOut) {
  if (grid <= 0 || warps < 1 || warps > 8 || tileBytes <= 0 || tileBytes % 128 || bytes % 16 || !src || !dst)
    return aar_host::fail(-1, "aar_bench_bulk_copy: bad arguments");
  const size_t smem = (size_t)warps * ((size_t)stages * tileBytes + 64);
  if (smem > 227 * 1024) return aar_host::fail(-1, "aar_bench_bulk_copy: %d stages x %d bytes do not fit in shared memory", stages, tileBytes);
  static unsigned* fallbackErr = nullptr;
  if (!errOut) {
    if (!fallbackErr && cudaMalloc(&fallbackErr, 4) != cudaSuccess) return aar_host::fail(-2, "aar_bench_bulk_copy: cudaMalloc failed");
    cudaMemsetAsync(fallbackErr, 0, 4, (cudaStream_t)stream);
    errOut = fallbackErr;
  }
  void (*k)(const char*, char*, size_t, int, unsigned*) = nullptr;
  if (stages == 2) k = bulkCopyBench<2>;
  else if (stages == 4) k = bulkCopyBench<4>;
  else if (stages == 8) k = bulkCopyBench<8>;
  else return aar_host::fail(-1, "aar_bench_bulk_copy: stages must be 2, 4 or 8");
  cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return aar_host::fail(-2, "aar_bench_bulk_copy: shared memory opt-in: %s", cudaGetErrorString(e));
  for (int i = 0; i < iters; ++i) k<<<grid, 32 * warps, smem, (cudaStream_t)stream>>>((const char*)src, (char*)dst, bytes, tileBytes, errOut);
  e = cudaGetLastError();
  if (e != cudaSuccess) return aar_host::fail(-2, "aar_bench_bulk_copy: %s", cudaGetErrorString(e));
  return 0;
}

// ------------------------------------------------------------------------------------------------ litmus: bulk store + flag
// Does completion of a bulk async-group (cp.async.bulk.wait_group) mean that the bytes have LANDED in a peer GPU's memory,
// so that a flag written afterwards -- with no fence -- can never be seen before the data?  The kernels want to know: a
// system fence on an SM that streams remote stores waits for all of them, newer ones included (measured: tens of
// microseconds), while the bulk-group wait is precise.  Sender CTA b on one GPU, receiver CTA b on the other, `iters` rounds:
//   sender:   fill shared memory with the round number, bulk-store `bytes` of it into the peer's buffer,
//             mode 0: wait_group 0, then st.relaxed.sys flag            (the candidate)
//             mode 1: NO wait, flag at once                              (control: must show violations, or the test sees nothing)
//             mode 2: wait_group 0, fence.acq_rel.sys, st.relaxed flag   (the textbook form)
//   receiver: polls the flag (relaxed), then reads the whole buffer through L2 and counts words that are not the round number.
namespace {

__global__ void __launch_bounds__(256, 1) litmusBulkFlag(int sender, int mode, char* peerBuf, char* myBuf, uint32_t* peerWord, uint32_t* myWord, int bytes,
                                                         int iters, unsigned long long* violations, unsigned* gaveUp) {
  extern __shared__ __align__(128) unsigned char smem[];  // one 32 KiB tile
  const int b = blockIdx.x;
  char* pbuf = peerBuf + (size_t)b * bytes;
  char* mbuf = myBuf + (size_t)b * bytes;
  // word 2b: flag (sender -> receiver), word 2b+1: ack (receiver -> sender); each side polls its OWN copy
  uint32_t* peerFlag = peerWord + 32 * b;
  uint32_t* myFlag = myWord + 32 * b;
  uint32_t* peerAck = peerWord + 32 * b + 16;
  uint32_t* myAck = myWord + 32 * b + 16;
  unsigned long long bad = 0;
  for (int k = 1; k <= iters; ++k) {
    if (sender) {
      for (int i = threadIdx.x; i < 32768 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = (uint32_t)k;
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncthreads();
      if (threadIdx.x == 0) {
        unsigned spins = 0;
        while (ld_relaxed_sys_u32(myAck) != (uint32_t)(k - 1)) {
          if (++spins > (1u << 26)) {
            atomicExch(gaveUp, 1u);
            break;
          }
        }
        for (int o = 0; o < bytes; o += 32768) {
          asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(pbuf + o), "r"(smem_u32(smem)), "r"(32768) : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        if (mode != 1) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
        if (mode == 2) asm volatile("fence.acq_rel.sys;" ::: "memory");
        asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(peerFlag), "r"((uint32_t)k) : "memory");
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // shared memory is rewritten next round
      }
      __syncthreads();
    } else {
      if (threadIdx.x == 0) {
        unsigned spins = 0;
        while (ld_relaxed_sys_u32(myFlag) != (uint32_t)k) {
          if (++spins > (1u << 26)) {
            atomicExch(gaveUp, 1u);
            break;
          }
        }
      }
      __syncthreads();
      for (int i = threadIdx.x; i < bytes / 4; i += blockDim.x) {
        uint32_t v;
        asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mbuf + (size_t)i * 4) : "memory");
        if (v != (uint32_t)k) ++bad;
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        asm volatile("fence.acq_rel.sys;" ::: "memory");
        asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(peerAck), "r"((uint32_t)k) : "memory");
      }
    }
  }
  if (bad) atomicAdd(violations, bad);
}

}  // namespace

// One launch on the caller's GPU; the peer runs the other role at the same time.  Buffers: grid * bytes each; words: grid * 32
// u32 each (zeroed by the caller before both sides start).  bytes: multiple of 32768.
AAR_API int aar_bench_litmus_bulk_flag(int sender, int mode, int grid, void* peerBuf, void* myBuf, void* peerWords, void* myWords, int bytes, int iters,
                                       unsigned long long* violations, unsigned* gaveUp, void* stream) {
  if (grid <= 0 || bytes <= 0 || bytes % 32768 || iters <= 0) return aar_host::fail(-1, "aar_bench_litmus_bulk_flag: bad arguments");
  litmusBulkFlag<<<grid, 256, 32768, (cudaStream_t)stream>>>(sender, mode, (char*)peerBuf, (char*)myBuf, (uint32_t*)peerWords, (uint32_t*)myWords, bytes, iters,
                                                              violations, gaveUp);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return aar_host::fail(-2, "aar_bench_litmus_bulk_flag: %s", cudaGetErrorString(e));
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
