// sm_100a kernels of the threshold allreduce: K1 scatter_push, K2 reduce_bcast, K3 gather_flush
// (+ K6 ring_advance and K8 force_complete folded in), their single-launch fusion, K4 NVLS, K5 one-shot and
// the single-device cluster emulation.  Reference steps replaced:
//   S1/S2 fetch+scatter   AllreduceWorker.scala:171-209   -> phaseScatter
//   S3/S4 store+threshold AllreduceWorker.scala:106-127, DataBuffer.scala:31-43 -> arrival flags + chunkDecision
//   S5/S6 reduce+broadcast AllreduceWorker.scala:230-251  -> phaseReduceBcast
//   S7/S8 store+round threshold AllreduceWorker.scala:129-150, DataBuffer.scala:69-75 -> gatherDecision
//   S9/S10 complete+flush+up() AllreduceWorker.scala:180-192,253-268, DataBuffer.scala:63-67 -> phaseGather
//   S11 catch-up          AllreduceWorker.scala:90-97     -> forcedByHost() inside every wait
// No tensor-core work exists in this problem (one fp32 add per element per contributor); every phase is
// NVLink- or HBM-bound, so the design effort goes into bytes moved, flags and overlap.
//
// Work decomposition: every (destination, chunk) "message" of the reference is cut into segments of whole
// 32 KiB tiles so that a grid of one CTA per SM has about two segments per CTA; a CTA streams its segment
// without synchronising, then pays ONE system-scope fence + one atomic, and the CTA that finishes a message
// last publishes its arrival flag.
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>

#include "aar_device.cuh"
#include "aar_host.h"
#include "aar_protocol.cuh"

using namespace aar_dev;

namespace {


// kStage = false: K1 scatter_push -- block `dest` of my input goes into peer dest's scatter ring over NVLink.
// kStage = true : NVLS variant -- the input is staged in MY region (the switch will pull it), only the flag travels.
//
// kElide (AAR_FLAG_ELIDE_LATE, strict variants): a rank that starts a round after some peer has already RETIRED it is a
// whole round late.  With thresholds < 1 most owners have then decided their chunks without it, and a copy nobody will
// read only keeps the rank late.  Such a rank looks at the owner's decision word (one remote load per message segment)
// and skips the copy when the decision excludes it.  Skipped segments still take part in the last-finisher election --
// the counter must be back at zero when the slot is reused -- but taint the message (upper half of the counter), and a
// tainted message is never flagged.  Modelled in csrc/tests/protocol_litmus.cpp (`elide`).
template <typename T, bool kStage, bool kElide = false>
__device__ void phaseScatter(const AarRankParams& p, const AarLayout& l, int cta, int nctas, Smem& sm, int unitsPerCta = 2) {
  if (p.flags & AAR_FLAG_SKIP_SCATTER) return;
  const bool elide = kElide && !kStage && (p.flags & AAR_FLAG_ELIDE_LATE);
  // shared words of the elision live here, not in Smem: only the strict instantiations carry them, the default kernels'
  // shared layout and code stay exactly as measured
  __shared__ int k1Late;   // some peer has already retired this round
  __shared__ int k1Skip;   // current unit: its owner has decided the chunk without this rank
  if (elide) {
    if (threadIdx.x == 0) {
      const uint32_t* base = reinterpret_cast<const uint32_t*>(regionOf(p, p.rank) + l.peerBaseOff);
      int late = 0;
      for (int j = 0; j < l.P; ++j) late |= ((int)ld_relaxed_sys_u32(base + j) > p.round) ? 1 : 0;
      k1Late = late;
    }
    __syncthreads();
  }
  const int P = l.P, C = l.C;
  const int slot = p.round % l.L;
  const uint32_t epoch = (uint32_t)(p.round + 1);
  char* my = regionOf(p, p.rank);
  const bool vecOk = geomVecOk<T>(l) && (reinterpret_cast<uintptr_t>(p.input) % 16 == 0);
  const SegPlan plan(l, (long long)P * C, nctas, unitsPerCta);
  const int S = plan.segs, segElems = plan.segElems(l);
  T* stage = reinterpret_cast<T*>(my + l.stageOff) + (long long)slot * paddedData(l);
  const bool inPlace = kStage && (reinterpret_cast<const T*>(p.input) == stage);  // the source wrote straight into the stage
  const long long total = (long long)C * P * S;
  // chunk-major, then segment, then destination staggered from self (AllreduceWorker.scala:195-196).  Consecutive
  // units go to different peers so every CTA spreads its pushes over all NVLink destinations; when the grid size is a
  // multiple of P a CTA would otherwise always draw the same destination, so each pass rotates by one.
  const bool rotate = (nctas % P) == 0;
  for (long long idx = cta; idx < total; idx += nctas) {
    const int i = (int)((idx % P + (rotate ? idx / nctas : 0)) % P);
    const int s = (int)((idx / P) % S);
    const int c = (int)(idx / ((long long)S * P));
    const int dest = (i + p.rank) % P;
    const int clen = chunkLen(l, dest, c);
    const int off = s * segElems;
    if (off >= clen) continue;
    // The shipped NVLS path stages without flow control: the owner of a block is the only reader of that part of the
    // staging slot, and with thresholds = 1 and no wait that gives up it has finished reading round R - L before this
    // rank could retire R - L.  Once waits can give up that no longer holds (litmus model, `nvls` = 1 with a lagger), so
    // the strict variant keeps the check: an owner that has not retired the slot's previous tenant gets no staging write
    // and no flag from this rank for this round.
    if ((!kStage || kElide) && !scatterDecision(p, l, dest, sm)) continue;
    const int n = min(segElems, clen - off);
    const long long inOff = blockStart(l, dest) + (long long)c * l.maxChunk + off;
    const T* src = reinterpret_cast<const T*>(p.input) + inOff;
    bool skip = false;
    if (elide && k1Late) {  // CTA-uniform: k1Late was published before a barrier
      if (threadIdx.x == 0) {
        const uint64_t d = ld_relaxed_sys_u64(reinterpret_cast<const uint64_t*>(regionOf(p, dest) + l.k2DecOff) + (long long)slot * C + c);
        k1Skip = ((d >> 32) == (uint64_t)epoch && !((d >> p.rank) & 1u)) ? 1 : 0;
      }
      __syncthreads();
      skip = k1Skip != 0;
    }
    if (skip) {
      // nothing to move
    } else if (kStage) {
      if (!inPlace) copyTile<T, false, false>(stage + inOff, src, n, vecOk);
    } else {
      T* dst = reinterpret_cast<T*>(regionOf(p, dest) + l.scatterOff) + ((long long)slot * P + p.rank) * l.blkStride +
               (long long)c * l.maxChunk + off;
      copyTile<T, true, false>(dst, src, n, vecOk);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();
      unsigned* ctr = reinterpret_cast<unsigned*>(my + l.k1CtrOff) + ((long long)slot * P + dest) * C + c;
      if (kElide) {
        const unsigned old = atomicAdd(ctr, skip ? 0x10001u : 1u);
        if ((old & 0xffffu) + 1 == plan.liveSegs(l, clen)) {
          atomicExch(ctr, 0u);
          if (!skip && (old >> 16) == 0) {  // no segment of this message was skipped: publish its arrival
            __threadfence_system();
            uint32_t* flag = reinterpret_cast<uint32_t*>(regionOf(p, dest) + l.sflagOff) + ((long long)slot * C + c) * AAR_MAXP + p.rank;
            st_release_sys_u32(flag, epoch);
          }
        }
      } else {
        const unsigned old = atomicAdd(ctr, 1u);
        if (old + 1 == plan.liveSegs(l, clen)) {  // last segment of this "ScatterBlock": publish its arrival
          atomicExch(ctr, 0u);
          __threadfence_system();
          uint32_t* flag = reinterpret_cast<uint32_t*>(regionOf(p, dest) + l.sflagOff) + ((long long)slot * C + c) * AAR_MAXP + p.rank;
          st_release_sys_u32(flag, epoch);
        }
      }
    }
    // Nobody starts the next segment before the fence above has drained: the SM-wide write drain would otherwise
    // wait for the new stores too and the arrival flag of an early chunk would only go out at the end of the phase.
    __syncthreads();
  }
}


// Ordered fp32 sum of the masked sources (peer-index order from 0.0f, AllreduceWorker.scala:245-249) and store of
// the result while it is still in registers.
//   kBcast = true : into every peer's reduce ring at element offset outElemOff (K2)
//   kBcast = false: to the local pointer `localOut` (K5 one-shot)
// PMAX bounds the source loop at compile time so all PMAX x U loads are issued before the first add
// (memory-level parallelism: 16 outstanding 16-byte loads per thread).
// kFC = true (AAR_FLAG_BCAST_FC): destinations whose bit in destMask is clear are skipped (their reduce-ring slot is
// not ours to write yet, see bcastDestMask).  A separate instantiation so that the default path compiles unchanged.
template <typename T, bool kBcast, int PMAX, int U, bool kFC = false>
__device__ __forceinline__ void reduceSpanVec(const AarRankParams& p, const AarLayout& l, const T* rows, long long rowStride,
                                              long long outElemOff, T* localOut, int nv, unsigned mask, float divisor,
                                              unsigned destMask = 0xffffffffu) {
  constexpr int V = Elem<T>::kPerVec;
  const int P = l.P;
  for (int base = threadIdx.x; base < nv; base += kThreads * U) {
    float acc[U][V];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int k = 0; k < V; ++k) acc[u][k] = 0.f;
    // groups of PMAX sources: all loads of a group are in flight before its first add; order stays 0..P-1
    for (int g = 0; g < P; g += PMAX) {
      uint4 v[PMAX][U];
#pragma unroll
      for (int s = 0; s < PMAX; ++s) {
        const int src = g + s;
        if (src < P && ((mask >> src) & 1u)) {
          const uint4* sp = reinterpret_cast<const uint4*>(rows + (long long)src * rowStride);
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int i = base + u * kThreads;
            if (i < nv) v[s][u] = ld_ring(sp + i);
          }
        }
      }
#pragma unroll
      for (int s = 0; s < PMAX; ++s) {
        const int src = g + s;
        if (src < P && ((mask >> src) & 1u)) {
#pragma unroll
          for (int u = 0; u < U; ++u) {
            float f[V];
            Elem<T>::unpack(v[s][u], f);
#pragma unroll
            for (int k = 0; k < V; ++k) acc[u][k] += f[k];
          }
        }
      }
    }
    uint4 outv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (divisor != 1.f) {
#pragma unroll
        for (int k = 0; k < V; ++k) acc[u][k] = acc[u][k] / divisor;
      }
      outv[u] = Elem<T>::pack(acc[u]);
    }
    if (kBcast) {
      for (int i2 = 0; i2 < P; ++i2) {
        const int dest = (i2 + p.rank) % P;
        if (kFC && !((destMask >> dest) & 1u)) continue;
        uint4* op = reinterpret_cast<uint4*>(reinterpret_cast<T*>(regionOf(p, dest) + l.reduceOff) + outElemOff);
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int i = base + u * kThreads;
          if (i < nv) st_peer(op + i, outv[u]);
        }
      }
    } else {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int i = base + u * kThreads;
        if (i < nv) st_stream(reinterpret_cast<uint4*>(localOut) + i, outv[u]);
      }
    }
  }
}

template <typename T, bool kBcast, bool kFC = false>
__device__ __forceinline__ void reduceSpan(const AarRankParams& p, const AarLayout& l, const T* rows, long long rowStride,
                                           long long outElemOff, T* localOut, int n, unsigned mask, float divisor, bool vecOk,
                                           unsigned destMask = 0xffffffffu) {
  constexpr int V = Elem<T>::kPerVec;
  const int P = l.P;
  const int nv = vecOk ? n / V : 0;
  if (nv > 0) {
    // 8 x 16-byte loads in flight per thread in every variant (64 KiB per CTA)
    if (kFC) {
      if (P <= 2) reduceSpanVec<T, kBcast, 2, 2, true>(p, l, rows, rowStride, outElemOff, localOut, nv, mask, divisor, destMask);
      else if (P <= 4) reduceSpanVec<T, kBcast, 4, 2, true>(p, l, rows, rowStride, outElemOff, localOut, nv, mask, divisor, destMask);
      else reduceSpanVec<T, kBcast, 4, 1, true>(p, l, rows, rowStride, outElemOff, localOut, nv, mask, divisor, destMask);
    } else if (!kBcast) {
      // one-shot (small, latency-bound messages): four loads in flight per thread keep the loop free of spills
      if (P <= 2) reduceSpanVec<T, kBcast, 2, 2>(p, l, rows, rowStride, outElemOff, localOut, nv, mask, divisor);
      else if (P <= 4) reduceSpanVec<T, kBcast, 4, 2>(p, l, rows, rowStride, outElemOff, localOut, nv, mask, divisor);
      else reduceSpanVec<T, kBcast, 4, 1>(p, l, rows, rowStride, outElemOff, localOut, nv, mask, divisor);
    } else {
      if (P <= 2) reduceSpanVec<T, kBcast, 2, 2>(p, l, rows, rowStride, outElemOff, localOut, nv, mask, divisor);
      else if (P <= 4) reduceSpanVec<T, kBcast, 4, 2>(p, l, rows, rowStride, outElemOff, localOut, nv, mask, divisor);
      else reduceSpanVec<T, kBcast, 4, 1>(p, l, rows, rowStride, outElemOff, localOut, nv, mask, divisor);
    }
  }
  for (int i = nv * V + threadIdx.x; i < n; i += kThreads) {
    float a = 0.f;
    for (int src = 0; src < P; ++src)
      if ((mask >> src) & 1u) a += Elem<T>::toFloat(__ldcg(rows + (long long)src * rowStride + i));
    if (divisor != 1.f) a = a / divisor;
    const T r = Elem<T>::fromFloat(a);
    if (kBcast) {
      for (int i2 = 0; i2 < P; ++i2) {
        const int dest = (i2 + p.rank) % P;
        if (kFC && !((destMask >> dest) & 1u)) continue;
        (reinterpret_cast<T*>(regionOf(p, dest) + l.reduceOff) + outElemOff)[i] = r;
      }
    } else {
      localOut[i] = r;
    }
  }
}

// NVLS (K4): the NVSwitch sums the staged inputs of all ranks (multimem.ld_reduce) and replicates the result into
// every rank's reduce ring (multimem.st).  Per direction each GPU moves about N(1 + 1/P) bytes instead of 2N(P-1)/P.
template <typename T>
__device__ __forceinline__ void nvlsSpan(const char* mcStage, char* mcOut, int nv) {
  constexpr int U = 8;
  for (int base = threadIdx.x; base < nv; base += kThreads * U) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int i = base + u * kThreads;
      if (i < nv) {
        if (sizeof(T) == 4) v[u] = multimem_ld_reduce_f32x4(mcStage + (size_t)i * 16);
        else v[u] = multimem_ld_reduce_bf16x8(mcStage + (size_t)i * 16);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int i = base + u * kThreads;
      if (i < nv) multimem_st_x4(mcOut + (size_t)i * 16, v[u]);
    }
  }
}


template <typename T, bool kNvls, bool kFC = false>
__device__ void phaseReduceBcast(const AarRankParams& p, const AarLayout& l, int cta, int nctas, Smem& sm, int unitsPerCta = 2) {
  const int P = l.P, C = l.C;
  const int myChunks = numChunks(l, p.rank);
  if (myChunks == 0) return;
  const int slot = p.round % l.L;
  const uint64_t epoch = (uint64_t)(p.round + 1);
  char* my = regionOf(p, p.rank);
  const bool vecOk = geomVecOk<T>(l);
  const SegPlan plan(l, myChunks, nctas, kNvls ? 1 : unitsPerCta);
  const int S = plan.segs, segElems = plan.segElems(l);
  const int total = myChunks * S;   // <= chunks x tiles of one block: far below 2^31
  unsigned destMask = 0xffffffffu;
  if (kFC && !kNvls && cta < total) destMask = bcastDestMask(p, l, sm);
  for (int idx = cta; idx < total; idx += nctas) {
    const int s = idx % S;
    const int c = idx / S;
    const int clen = chunkLen(l, p.rank, c);
    const int off = s * segElems;
    if (off >= clen) continue;
    if (sm.k2Chunk != c) {
      __syncthreads();  // everybody has read the previous mask
      if (threadIdx.x < 32) {
        unsigned m = chunkDecision(p, l, reinterpret_cast<uint64_t*>(my + l.k2DecOff) + (long long)slot * C + c,
                                   reinterpret_cast<const uint32_t*>(my + l.sflagOff) + ((long long)slot * C + c) * AAR_MAXP,
                                   kNvls ? P : p.minRequired);
        if (threadIdx.x == 0) {
          if (sm.k2Chunk < 0) traceMark(p, l, 2);
          sm.k2Mask = m;
          sm.k2Chunk = c;
        }
      }
      __syncthreads();
    }
    const unsigned mask = sm.k2Mask;
    const int count = __popc(mask);
    const int n = min(segElems, clen - off);
    const long long outElemOff = ((long long)slot * P + p.rank) * l.blkStride + (long long)c * l.maxChunk + off;
    // Strict NVLS: a decision that was forced over an incomplete set cannot be honoured by the switch (it would add the
    // absent ranks' stale staging slots to the sum while `count` says they are not in it).  Such a chunk is not published
    // at all: every rank's gather gives up on it and flushes zeros with count 0 -- "never arrived" -- instead of a sum
    // that does not match its count.
    const bool nvlsIncomplete = kNvls && kFC && count != P;
    if (nvlsIncomplete) {
      // nothing to move
    } else if (kNvls) {
      // the switch sums every bound GPU: only valid when all P contributed (the launcher enforces thresholds = 1)
      const char* mcStage = reinterpret_cast<const char*>(p.mcRegion) + l.stageOff +
                            ((long long)slot * paddedData(l) + blockStart(l, p.rank) + (long long)c * l.maxChunk + off) * sizeof(T);
      char* mcOut = reinterpret_cast<char*>(p.mcRegion) + l.reduceOff + outElemOff * sizeof(T);
      nvlsSpan<T>(mcStage, mcOut, n / Elem<T>::kPerVec);
    } else {
      const float divisor = ((p.flags & AAR_FLAG_NORMALIZE) && count > 1) ? (float)count : 1.f;
      const T* rows = reinterpret_cast<const T*>(my + l.scatterOff) + (long long)slot * P * l.blkStride + (long long)c * l.maxChunk + off;
      if (kFC) reduceSpan<T, true, true>(p, l, rows, l.blkStride, outElemOff, nullptr, n, mask, divisor, vecOk, destMask);
      else reduceSpan<T, true>(p, l, rows, l.blkStride, outElemOff, nullptr, n, mask, divisor, vecOk);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();
      unsigned* ctr = reinterpret_cast<unsigned*>(my + l.k2CtrOff) + (long long)slot * C + c;
      const unsigned old = atomicAdd(ctr, 1u);
      if (old + 1 == (unsigned)((clen + segElems - 1) / segElems)) {  // last live segment of this "ReduceBlock": tell every peer, count included
        atomicExch(ctr, 0u);
        __threadfence_system();
        const uint64_t word = (epoch << 32) | (uint64_t)count;
        for (int i2 = 0; i2 < (nvlsIncomplete ? 0 : P); ++i2) {
          const int dest = (i2 + p.rank) % P;
          if (kFC && !((destMask >> dest) & 1u)) continue;  // no data was written there: no flag either
          uint64_t* flag = reinterpret_cast<uint64_t*>(regionOf(p, dest) + l.rflagOff) + ((long long)slot * P + p.rank) * C + c;
          st_release_sys_u64(flag, word);
        }
      }
    }
    __syncthreads();  // see phaseScatter: let the fence drain before new stores are issued
  }
}



template <typename T>
__device__ void phaseGatherBody(const AarRankParams& p, const AarLayout& l, int cta, int nctas, Smem& sm) {
  const int P = l.P, C = l.C;
  const int slot = p.round % l.L;
  char* my = regionOf(p, p.rank);
  gatherDecision(p, l, slot, sm);
  traceMark(p, l, 4);
  const int* counts = reinterpret_cast<const int*>(my + l.countsOff) + (long long)slot * l.totalChunks;
  // AAR_FLAG_VIEW_OUTPUT: the caller consumes the ring slot itself ([P][blkStride] == the flushed vector when
  // blkStride == stepSize), stream-ordered; only chunks that never arrived are touched (zero-filled in place).
  const bool view = (p.flags & AAR_FLAG_VIEW_OUTPUT) != 0;
  const bool vecOk = geomVecOk<T>(l) && (view || reinterpret_cast<uintptr_t>(p.output) % 16 == 0);
  const SegPlan plan(l, (long long)P * C, nctas);
  const int S = plan.segs, segElems = plan.segElems(l);
  const long long total = (long long)P * C * S;
  for (long long idx = cta; idx < total; idx += nctas) {
    const int s = (int)(idx % S);
    const int c = (int)((idx / S) % C);
    const int owner = (int)(idx / ((long long)S * C));
    const int clen = chunkLen(l, owner, c);
    const int off = s * segElems;
    if (off >= clen) continue;
    const int n = min(segElems, clen - off);
    const int cnt = __ldcg(counts + owner * C + c);
    T* ring = reinterpret_cast<T*>(my + l.reduceOff) + ((long long)slot * P + owner) * l.blkStride + (long long)c * l.maxChunk + off;
    if (view) {
      if (cnt < 0) zeroTile<T>(ring, n, vecOk);
      continue;
    }
    T* dst = reinterpret_cast<T*>(p.output) + blockStart(l, owner) + (long long)c * l.maxChunk + off;
    if (cnt >= 0) {
      copyTile<T, false, true>(dst, ring, n, vecOk);
    } else {
      zeroTile<T>(dst, n, vecOk);  // chunks whose ReduceBlock never arrived stay zero (AllreduceWorker.scala:183)
    }
  }
}

template <typename T>
__device__ void phaseGather(const AarRankParams& p, const AarLayout& l, int cta, int nctas, Smem& sm) {
  phaseGatherBody<T>(p, l, cta, nctas, sm);
  retireRound(p, l, nctas);
}

// Eager gather (thComplete = 1.0 only): every chunk is required, so nothing has to be decided -- a CTA copies chunk
// (owner, c) as soon as its ReduceBlock flag lands, while other chunks are still in flight.  A chunk that never shows
// up (host-forced round, device timeout) is zero-filled like in phaseGather.
//
// kStrict (AAR_FLAG_EAGER_DEC): the segments of one chunk are copied by different CTAs, each with its own wait.  When a
// wait gives up, a CTA that looked earlier may zero-fill its segment while one that looks later sees the flag and copies:
// the chunk comes out half and half (found by csrc/tests/protocol_litmus.cpp, `eagerDecision`).  The strict variant takes
// ONE decision per (round, chunk) in a CAS word -- "arrived with this count" or "abandoned" -- made by the first CTA that
// stops waiting and adopted by everybody else; a ReduceBlock that lands after an "abandoned" is ignored, like a late
// ScatterBlock after the contributor decision.  A separate instantiation: the default path compiles unchanged.
template <typename T, bool kStrict = false>
__device__ void phaseGatherEager(const AarRankParams& p, const AarLayout& l, int cta, int nctas, Smem& sm) {
  const int P = l.P, C = l.C;
  const int slot = p.round % l.L;
  const uint64_t epoch = (uint64_t)(p.round + 1);
  char* my = regionOf(p, p.rank);
  const bool view = (p.flags & AAR_FLAG_VIEW_OUTPUT) != 0;
  const bool vecOk = geomVecOk<T>(l) && (view || reinterpret_cast<uintptr_t>(p.output) % 16 == 0);
  const SegPlan plan(l, (long long)P * C, nctas);
  const int S = plan.segs, segElems = plan.segElems(l);
  const long long total = (long long)P * C * S;
  const uint64_t* flags = reinterpret_cast<const uint64_t*>(my + l.rflagOff) + (long long)slot * P * C;
  // chunk-major (chunk c of every owner is reduced at about the same time), owners staggered from self
  for (long long idx = cta; idx < total; idx += nctas) {
    const int s = (int)(idx % S);
    const int owner = ((int)((idx / S) % P) + p.rank) % P;
    const int c = (int)(idx / ((long long)S * P));
    const int clen = chunkLen(l, owner, c);
    const int off = s * segElems;
    if (off >= clen) continue;
    const int key = owner * C + c;
    if (sm.gKey != key) {
      __syncthreads();
      if (threadIdx.x == 0) {
        Deadline dl(p.timeoutNs);
        unsigned spins = 0, backoff = 32;
        int cnt = -1;
        uint64_t* dec = reinterpret_cast<uint64_t*>(my + l.egDecOff) + (long long)slot * P * C + key;
        uint64_t d = 0;
        bool decided = false;
        if (kStrict) {
          d = ld_acquire_gpu_u64(dec);
          if ((d >> 32) == epoch) {  // somebody already stopped waiting for this chunk: adopt
            cnt = (int)(d & 0xffffffffu) - 1;
            decided = true;
          }
        }
        while (!decided) {
          const uint64_t f = ld_relaxed_sys_u64(flags + key);
          if ((f >> 32) == epoch) {
            cnt = (int)(f & 0xffffffffu);
            break;
          }
          if ((p.flags & AAR_FLAG_FORCE) || ((++spins & 255u) == 0)) {
            if (forcedByHost(p, l)) break;
            if (dl.expired()) {
              atomicOr(statusWord(p, l, 0), AAR_ST_TIMEOUT_GATHER);
              break;
            }
          }
          __nanosleep(backoff);
          if (backoff < 512) backoff *= 2;
        }
        __threadfence_system();
        if (kStrict && !decided) {
          // the fence above is the acquire for the flag and, being cumulative, the release for the word below: a CTA
          // that adopts "arrived" through the word also sees the owner's data
          const uint64_t want = (epoch << 32) | (uint64_t)(uint32_t)(cnt + 1);
          const uint64_t old = atomicCAS(reinterpret_cast<unsigned long long*>(dec), (unsigned long long)d, (unsigned long long)want);
          if (old != d && (old >> 32) == epoch) {
            cnt = (int)(old & 0xffffffffu) - 1;
            __threadfence();
          }
        }
        sm.gKey = key;
        sm.gCnt = cnt;
        if (s == 0 && p.outCounts) p.outCounts[key] = cnt < 0 ? 0 : cnt;
      }
      __syncthreads();
    }
    const int cnt = sm.gCnt;
    const int n = min(segElems, clen - off);
    T* ring = reinterpret_cast<T*>(my + l.reduceOff) + ((long long)slot * P + owner) * l.blkStride + (long long)c * l.maxChunk + off;
    if (view) {
      if (cnt < 0) zeroTile<T>(ring, n, vecOk);
      continue;
    }
    T* dst = reinterpret_cast<T*>(p.output) + blockStart(l, owner) + (long long)c * l.maxChunk + off;
    if (cnt >= 0) copyTile<T, false, true>(dst, ring, n, vecOk);
    else zeroTile<T>(dst, n, vecOk);
  }
}

// ------------------------------------------------------------------------------------------------ K5 one-shot
// Small messages are latency-bound: one hop instead of two.  Every rank pushes its whole vector into slot
// [round % L][src] of every peer's one-shot buffer (+ one arrival flag per peer), then sums the arrived set locally.
// thReduce decides how many sources to wait for; thComplete has nothing left to wait for.
template <typename T>
__device__ void phaseOneshotPush(const AarRankParams& p, const AarLayout& l, int cta, int nctas, Smem& sm) {
  if (p.flags & AAR_FLAG_SKIP_SCATTER) return;
  const int P = l.P;
  const int slot = p.round % l.L;
  const int Tn = numTiles(l, l.dataSize);
  const long long stride = paddedData(l);
  char* my = regionOf(p, p.rank);
  const bool vecOk = reinterpret_cast<uintptr_t>(p.input) % 16 == 0;
  for (long long idx = cta; idx < (long long)P * Tn; idx += nctas) {
    const int t = (int)(idx % Tn);
    const int dest = ((int)(idx / Tn) + p.rank) % P;
    if (!scatterDecision(p, l, dest, sm)) continue;
    const int off = t * l.tileElems;
    const int n = min(l.tileElems, l.dataSize - off);
    T* dst = reinterpret_cast<T*>(regionOf(p, dest) + l.oneshotOff) + ((long long)slot * P + p.rank) * stride + off;
    copyTile<T, true, false>(dst, reinterpret_cast<const T*>(p.input) + off, n, vecOk);
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();
      unsigned* ctr = reinterpret_cast<unsigned*>(my + l.k1CtrOff) + ((long long)slot * P + dest) * l.C;
      const unsigned old = atomicAdd(ctr, 1u);
      if (old + 1 == (unsigned)Tn) {
        atomicExch(ctr, 0u);
        __threadfence_system();
        st_release_sys_u32(reinterpret_cast<uint32_t*>(regionOf(p, dest) + l.oneshotFlagOff) + slot * AAR_MAXP + p.rank, (uint32_t)(p.round + 1));
      }
    }
  }
}

template <typename T>
__device__ void phaseOneshotReduce(const AarRankParams& p, const AarLayout& l, int cta, int nctas, Smem& sm) {
  const int P = l.P;
  const int slot = p.round % l.L;
  const int Tn = numTiles(l, l.dataSize);
  const long long stride = paddedData(l);
  char* my = regionOf(p, p.rank);
  if (threadIdx.x < 32) {
    unsigned m = chunkDecision(p, l, reinterpret_cast<uint64_t*>(my + l.k2DecOff) + (long long)slot * l.C,
                               reinterpret_cast<const uint32_t*>(my + l.oneshotFlagOff) + slot * AAR_MAXP, p.minRequired);
    if (threadIdx.x == 0) sm.k2Mask = m;
  }
  __syncthreads();
  const unsigned mask = sm.k2Mask;
  const int count = __popc(mask);
  const float divisor = ((p.flags & AAR_FLAG_NORMALIZE) && count > 1) ? (float)count : 1.f;
  const bool vecOk = reinterpret_cast<uintptr_t>(p.output) % 16 == 0;
  if (cta == 0 && p.outCounts)
    for (int i = threadIdx.x; i < l.totalChunks; i += kThreads) p.outCounts[i] = count;
  for (int t = cta; t < Tn; t += nctas) {
    const int off = t * l.tileElems;
    const int n = min(l.tileElems, l.dataSize - off);
    const T* rows = reinterpret_cast<const T*>(my + l.oneshotOff) + (long long)slot * P * stride + off;
    reduceSpan<T, false>(p, l, rows, stride, 0, reinterpret_cast<T*>(p.output) + off, n, mask, divisor, vecOk);
  }
  retireRound(p, l, nctas);
}


// phases: bit 0 = scatter, bit 1 = reduce+broadcast, bit 2 = gather.  7 = one launch per round per rank: the three
// phases back to back; chunk flags let a CTA start reducing chunk c as soon as its contributors landed while other
// CTAs are still pushing later chunks.
template <typename T, bool kNvls, bool kFC = false>
__global__ void __launch_bounds__(kThreads, 1) roundKernel(const AarRankParams p, const AarLayout l, int phases) {
  __shared__ Smem sm;
  initSmem(sm);
  traceMark(p, l, 0);
  if (phases & 1) phaseScatter<T, kNvls, kFC>(p, l, blockIdx.x, gridDim.x, sm);
  traceMark(p, l, 1);
  if (phases & 2) phaseReduceBcast<T, kNvls, kFC>(p, l, blockIdx.x, gridDim.x, sm);
  traceMark(p, l, 3);
  if (phases & 4) phaseGather<T>(p, l, blockIdx.x, gridDim.x, sm);
  traceMark(p, l, 5);
}

// Two-shot P2P, one launch per round, CTAs specialised by role so K1, K2 and K3 overlap chunk by chunk:
//   CTAs [0, nPush)               K1 scatter_push
//   CTAs [nPush, nPush+nReduce)   K2 reduce_bcast
//   the rest                      K3 gather (eager per chunk when every chunk is required, else the round decision)
template <typename T, bool kFC = false>
__global__ void __launch_bounds__(kThreads, 1) p2pPipelinedKernel(const AarRankParams p, const AarLayout l, int split) {
  __shared__ Smem sm;
  initSmem(sm);
  traceMark(p, l, 0);
  const int nctas = gridDim.x;
  const int nPush = split & 0xffff, nReduce = split >> 16;
  const int b = blockIdx.x;
  if (b < nPush) {
    phaseScatter<T, false, kFC>(p, l, b, nPush, sm, 1);
    traceMark(p, l, 1);
  } else if (b < nPush + nReduce) {
    phaseReduceBcast<T, false, kFC>(p, l, b - nPush, nReduce, sm, 1);
    traceMark(p, l, 3);
  } else {
    const int cta = b - nPush - nReduce, n = nctas - nPush - nReduce;
    if (p.minChunksRequired >= l.totalChunks && !(p.flags & AAR_FLAG_FORCE)) {
      phaseGatherEager<T, kFC>(p, l, cta, n, sm);
      traceMark(p, l, 4);
    } else {
      phaseGatherBody<T>(p, l, cta, n, sm);
    }
  }
  retireRound(p, l, nctas);
  traceMark(p, l, 5);
}

// K4, one launch per round, CTAs specialised by role so the stages overlap chunk by chunk -- and, because consecutive
// rounds alternate between two streams, round by round: the stage CTAs sit first in block order and leave early, so
// the next round's stage CTAs take their SMs while this round is still reducing and gathering.
//   CTAs [0, nStage)                  copy my input into the multicast-bound staging area, arrival flags per chunk
//   CTAs [nStage, nStage + nReduce)   in-switch reduce + multicast store of my block (few CTAs saturate NVLS: measured
//                                     multimem throughput does not grow beyond ~8-16 CTAs, profiles/r1_link_bench_8gpu.json)
//   the rest                          copy every reduced chunk to the caller's output as soon as its flag lands
template <typename T, bool kStrict = false>
__global__ void __launch_bounds__(kThreads, 1) nvlsPipelinedKernel(const AarRankParams p, const AarLayout l, int split) {
  __shared__ Smem sm;
  initSmem(sm);
  traceMark(p, l, 0);
  const int nctas = gridDim.x;
  const int nStage = split & 0xffff, nReduce = split >> 16;
  const int b = blockIdx.x;
  if (b < nStage) {
    phaseScatter<T, true, kStrict>(p, l, b, nStage, sm, 1);
    traceMark(p, l, 1);
  } else if (b < nStage + nReduce) {
    phaseReduceBcast<T, true, kStrict>(p, l, b - nStage, nReduce, sm);
    traceMark(p, l, 3);
  } else {
    phaseGatherEager<T, kStrict>(p, l, b - nStage - nReduce, nctas - nStage - nReduce, sm);
    traceMark(p, l, 4);
  }
  retireRound(p, l, nctas);
  traceMark(p, l, 5);
}

template <typename T>
__global__ void __launch_bounds__(kThreads, 1) oneshotKernel(const AarRankParams p, const AarLayout l, int phase) {
  __shared__ Smem sm;
  initSmem(sm);
  if (phase == 0 || phase == 2) phaseOneshotPush<T>(p, l, blockIdx.x, gridDim.x, sm);
  if (phase == 1 || phase == 2) phaseOneshotReduce<T>(p, l, blockIdx.x, gridDim.x, sm);
}

__global__ void delayKernel(unsigned long long ns) {
  const unsigned long long t0 = globaltimer_ns();
  while (globaltimer_ns() - t0 < ns) __nanosleep(256);
}

// ------------------------------------------------------------------------------------------------ host side
template <typename K>
int residentGrid(K kernel, int want) {
  int dev = 0, sms = 0, perSm = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kernel, kThreads, 0) != cudaSuccess) return -1;
  int cap = sms * (perSm > 0 ? perSm : 1);
  if (want <= 0) want = sms;
  return want < cap ? want : cap;
}

template <typename K>
int launch(K kernel, const AarRankParams* p, const AarLayout* l, int grid, void* stream, const char* what, int phases) {
  if (!p || !l) return aar_host::fail(AAR_EINVAL, "%s: null params", what);
  if (p->rank < 0 || p->rank >= l->P || l->P > AAR_MAXP) return aar_host::fail(AAR_EINVAL, "%s: bad rank %d for P=%d", what, p->rank, l->P);
  // every CTA may wait on work that another CTA of the same grid still has to do: the grid must be co-resident
  int g = residentGrid(kernel, grid);
  if (g <= 0) return aar_host::fail(AAR_ECUDA, "%s: occupancy query failed: %s", what, cudaGetErrorString(cudaGetLastError()));
  kernel<<<g, kThreads, 0, (cudaStream_t)stream>>>(*p, *l, phases);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return aar_host::fail(AAR_ECUDA, "%s: launch failed: %s", what, cudaGetErrorString(e));
  return AAR_OK;
}

int launchRound(const AarRankParams* p, const AarLayout* lay, int grid, void* stream, const char* what, int phases) {
  if (!lay) return aar_host::fail(AAR_EINVAL, "%s: null layout", what);
  if (p && (p->flags & AAR_FLAG_BCAST_FC) && (phases & 3)) {  // opt-in: strict scatter / broadcast variants
    if (lay->dtype == AAR_F32) return launch(roundKernel<float, false, true>, p, lay, grid, stream, what, phases);
    if (lay->dtype == AAR_BF16) return launch(roundKernel<__nv_bfloat16, false, true>, p, lay, grid, stream, what, phases);
  }
  if (lay->dtype == AAR_F32) return launch(roundKernel<float, false>, p, lay, grid, stream, what, phases);
  if (lay->dtype == AAR_BF16) return launch(roundKernel<__nv_bfloat16, false>, p, lay, grid, stream, what, phases);
  return aar_host::fail(AAR_EINVAL, "%s: unknown dtype %d", what, lay->dtype);
}

}  // namespace

AAR_API int aar_launch_scatter(const AarRankParams* p, const AarLayout* lay, int grid, void* stream) {
  return launchRound(p, lay, grid, stream, "aar_launch_scatter", 1);
}
AAR_API int aar_launch_reduce_bcast(const AarRankParams* p, const AarLayout* lay, int grid, void* stream) {
  return launchRound(p, lay, grid, stream, "aar_launch_reduce_bcast", 2);
}
AAR_API int aar_launch_gather(const AarRankParams* p, const AarLayout* lay, int grid, void* stream) {
  return launchRound(p, lay, grid, stream, "aar_launch_gather", 4);
}
AAR_API int aar_launch_round_fused(const AarRankParams* p, const AarLayout* lay, int grid, void* stream) {
  return launchRound(p, lay, grid, stream, "aar_launch_round_fused", 7);
}

// Role-specialised two-shot round.  split = nPush | nReduce << 16 (0 = heuristic); the remaining CTAs gather.
AAR_API int aar_launch_round_pipelined(const AarRankParams* p, const AarLayout* lay, int grid, int split, void* stream) {
  const char* what = "aar_launch_round_pipelined";
  if (!p || !lay) return aar_host::fail(AAR_EINVAL, "%s: null params", what);
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int g = grid > 0 ? grid : sms;
  if (g < 3) g = 3;
  int nPush = split & 0xffff, nReduce = split >> 16;
  if (nPush <= 0 || nReduce <= 0 || nPush + nReduce >= g) {
    nPush = g * 3 / 8;
    nReduce = g * 3 / 8;
  }
  if (nPush < 1) nPush = 1;
  if (nReduce < 1) nReduce = 1;
  const int sp = nPush | (nReduce << 16);
  if (p->flags & (AAR_FLAG_BCAST_FC | AAR_FLAG_EAGER_DEC)) {  // strict variants do both (each only matters in its own regime)
    if (lay->dtype == AAR_F32) return launch(p2pPipelinedKernel<float, true>, p, lay, g, stream, what, sp);
    if (lay->dtype == AAR_BF16) return launch(p2pPipelinedKernel<__nv_bfloat16, true>, p, lay, g, stream, what, sp);
  }
  if (lay->dtype == AAR_F32) return launch(p2pPipelinedKernel<float>, p, lay, g, stream, what, sp);
  if (lay->dtype == AAR_BF16) return launch(p2pPipelinedKernel<__nv_bfloat16>, p, lay, g, stream, what, sp);
  return aar_host::fail(AAR_EINVAL, "%s: unknown dtype %d", what, lay->dtype);
}

AAR_API int aar_launch_oneshot(const AarRankParams* p, const AarLayout* lay, int phase, int grid, void* stream) {
  if (!lay || lay->oneshotOff == 0) return aar_host::fail(AAR_EINVAL, "aar_launch_oneshot: layout has no one-shot buffers");
  if (phase < 0 || phase > 2) return aar_host::fail(AAR_EINVAL, "aar_launch_oneshot: phase must be 0, 1 or 2");
  if (lay->dtype == AAR_F32) return launch(oneshotKernel<float>, p, lay, grid, stream, "aar_launch_oneshot", phase);
  if (lay->dtype == AAR_BF16) return launch(oneshotKernel<__nv_bfloat16>, p, lay, grid, stream, "aar_launch_oneshot", phase);
  return aar_host::fail(AAR_EINVAL, "aar_launch_oneshot: unknown dtype %d", lay->dtype);
}

AAR_API int aar_launch_nvls(const AarRankParams* p, const AarLayout* lay, int phases, int grid, void* stream) {
  const char* what = "aar_launch_nvls";
  if (!p || !lay) return aar_host::fail(AAR_EINVAL, "%s: null params", what);
  if (!p->mcRegion || lay->stageOff == 0) return aar_host::fail(AAR_EUNSUPPORTED, "%s: needs a multicast mapping and the staging area", what);
  if ((phases & 7) == 0) return aar_host::fail(AAR_EINVAL, "%s: phases is a bit mask 1..7 (| nReduceCtas << 8)", what);
  const int V = 16 / lay->elemSize;
  if (lay->stepSize % V || lay->maxChunk % V || lay->dataSize % V)
    return aar_host::fail(AAR_EUNSUPPORTED, "%s: multimem needs 16-byte aligned blocks and chunks", what);
  const int total = lay->totalChunks;
  if (p->minRequired != lay->P || p->minChunksRequired != total || (p->flags & (AAR_FLAG_SKIP_SCATTER | AAR_FLAG_EXACT_QUORUM | AAR_FLAG_NORMALIZE)))
    return aar_host::fail(AAR_EUNSUPPORTED, "%s: the switch sums every bound GPU, so thresholds must be 1.0 (use the P2P paths otherwise)", what);
  const int mask = phases & 7, nReduceArg = phases >> 8;
  if (mask == 7) {
    // one launch: role-specialised pipeline; needs at least one CTA of each role
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int g = grid > 0 ? grid : sms;
    if (g < 3) g = 3;
    int nReduce = nReduceArg & 0xff, nStage = (nReduceArg >> 8) & 0xff;
    if (nReduce <= 0) nReduce = 16;
    if (nStage <= 0) nStage = g / 3;
    if (nReduce > g / 3) nReduce = g / 3;
    if (nStage + nReduce >= g) nStage = g - nReduce - 1;
    const int sp = nStage | (nReduce << 16);
    if (p->flags & AAR_FLAG_EAGER_DEC) {
      if (lay->dtype == AAR_F32) return launch(nvlsPipelinedKernel<float, true>, p, lay, g, stream, what, sp);
      if (lay->dtype == AAR_BF16) return launch(nvlsPipelinedKernel<__nv_bfloat16, true>, p, lay, g, stream, what, sp);
    }
    if (lay->dtype == AAR_F32) return launch(nvlsPipelinedKernel<float>, p, lay, g, stream, what, sp);
    if (lay->dtype == AAR_BF16) return launch(nvlsPipelinedKernel<__nv_bfloat16>, p, lay, g, stream, what, sp);
    return aar_host::fail(AAR_EINVAL, "%s: unknown dtype %d", what, lay->dtype);
  }
  if (lay->dtype == AAR_F32) return launch(roundKernel<float, true>, p, lay, grid, stream, what, mask);
  if (lay->dtype == AAR_BF16) return launch(roundKernel<__nv_bfloat16, true>, p, lay, grid, stream, what, mask);
  return aar_host::fail(AAR_EINVAL, "%s: unknown dtype %d", what, lay->dtype);
}

AAR_API int aar_local_round(const AarRankParams* params, int nranks, const AarLayout* lay, unsigned skipMask, int grid, void* stream) {
  if (!params || !lay || nranks != lay->P) return aar_host::fail(AAR_EINVAL, "aar_local_round: need one params entry per rank");
  // Phase order matters on a single device: no kernel may wait for a later launch.
  for (int r = 0; r < nranks; ++r) {
    if ((skipMask >> r) & 1u) continue;
    int rc = aar_launch_scatter(&params[r], lay, grid, stream);
    if (rc) return rc;
  }
  for (int r = 0; r < nranks; ++r) {
    int rc = aar_launch_reduce_bcast(&params[r], lay, grid, stream);
    if (rc) return rc;
  }
  for (int r = 0; r < nranks; ++r) {
    int rc = aar_launch_gather(&params[r], lay, grid, stream);
    if (rc) return rc;
  }
  return AAR_OK;
}

AAR_API int aar_launch_delay(unsigned long long ns, void* stream) {
  delayKernel<<<1, 1, 0, (cudaStream_t)stream>>>(ns);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return aar_host::fail(AAR_ECUDA, "aar_launch_delay: %s", cudaGetErrorString(e));
  return AAR_OK;
}

// ------------------------------------------------------------------------------------------------ fixed-scratch variant
// One full allreduce round of `nranks` emulated ranks on the current device whose ring buffers, flag words and
// counters -- everything aar_local_round() expects the caller to have allocated as "regions" -- live in the
// file-scope array below instead.  For callers that are not allowed to allocate device memory themselves: code
// that runs inside someone else's memory plan (an inference server with a frozen arena, a plugin called from a
// host application that owns the allocator, a signal-safe or real-time path where cudaMalloc's implicit
// synchronisation and possible failure are not acceptable).  Nothing here calls cudaMalloc/cudaFree or any
// routine that allocates; inputs and outputs still come from the caller.
//
// The array is part of the library image: every GPU that loads libaar_cuda.so carries AAR_STATIC_SCRATCH_BYTES of
// it until the process exits.  All calls on a device share it and it keeps what the previous call left behind, so
// each call first clears the words that must start at zero (flags, election counters, decision words); the ring
// payload needs no clearing because a chunk is only read after its flag of the current call is seen.  Two calls that
// overlap on the same GPU (different streams, different threads) race on the array: serialise them.
//
// `g_staticScratch` names a device symbol; in host code the identifier is only a handle, not an address a kernel
// can use.  The device address is looked up once per device with cudaGetSymbolAddress(), with that device current,
// on the first call made on that device.  That first call should therefore come before any stream capture.
namespace {

__device__ __align__(256) unsigned char g_staticScratch[AAR_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_staticMutex;
void* g_staticBase[kMaxDevices] = {};

int staticScratchBase(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return aar_host::fail(AAR_ECUDA, "aar_local_allreduce_static: cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices) return aar_host::fail(AAR_EUNSUPPORTED, "aar_local_allreduce_static: device index %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_staticMutex);
  if (!g_staticBase[dev]) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_staticScratch);
    if (e != cudaSuccess || !ptr)
      return aar_host::fail(AAR_ECUDA, "aar_local_allreduce_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
    g_staticBase[dev] = ptr;
  }
  *out = g_staticBase[dev];
  return AAR_OK;
}

}  // namespace

AAR_API int aar_local_allreduce_static(const void* const* inputs, void* const* outputs, int nranks, int dataSize, int maxChunkSize, int dtype,
                                       void* stream) {
  const char* what = "aar_local_allreduce_static";
  if (!inputs || !outputs) return aar_host::fail(AAR_EINVAL, "%s: null input/output table", what);
  if (nranks < 1 || nranks > AAR_MAXP) return aar_host::fail(AAR_EINVAL, "%s: nranks=%d outside 1..%d", what, nranks, AAR_MAXP);
  for (int r = 0; r < nranks; ++r)
    if (!inputs[r] || !outputs[r]) return aar_host::fail(AAR_EINVAL, "%s: rank %d has a null input or output", what, r);
  AarLayout lay;
  int rc = aar_layout_compute(nranks, dataSize, maxChunkSize, /*maxLag=*/0, dtype, /*oneshot=*/0, /*stage=*/0, &lay);
  if (rc) return rc;
  // bytes one emulated rank really uses (the layout's own regionBytes is padded to the 2 MiB VMM granularity)
  // = end of the last layout section (the eager-gather decision words, appended after the trace marks)
  const uint64_t perRank = (lay.egDecOff + (uint64_t)lay.L * lay.P * lay.C * 8 + 255) / 256 * 256;
  const uint64_t need = perRank * (uint64_t)nranks;
  if (need > (uint64_t)AAR_STATIC_SCRATCH_BYTES)
    return aar_host::fail(AAR_EINVAL,
                          "%s: %d ranks x %d elements need %llu bytes of scratch but the built-in array holds %llu; use aar_local_round() with "
                          "caller-allocated regions for this size",
                          what, nranks, dataSize, (unsigned long long)need, (unsigned long long)AAR_STATIC_SCRATCH_BYTES);
  void* base = nullptr;
  rc = staticScratchBase(&base);
  if (rc) return rc;
  char* b = static_cast<char*>(base);
  // flags, counters and decision words of the previous call must not be mistaken for this call's
  for (int r = 0; r < nranks; ++r) {
    cudaError_t e = cudaMemsetAsync(b + (uint64_t)r * perRank + lay.ctrlBytes, 0, perRank - lay.ctrlBytes, (cudaStream_t)stream);
    if (e != cudaSuccess) return aar_host::fail(AAR_ECUDA, "%s: clearing the control words: %s", what, cudaGetErrorString(e));
  }
  AarRankParams params[AAR_MAXP];
  for (int r = 0; r < nranks; ++r) {
    AarRankParams& p = params[r];
    memset(&p, 0, sizeof p);
    p.rank = r;
    p.round = 0;
    p.minRequired = nranks;               // plain allreduce: every rank contributes, every chunk is required
    p.minChunksRequired = lay.totalChunks;
    p.maxLag = 0;
    p.timeoutNs = 2000000000ull;          // liveness guard only; every flag is already set when a phase looks at it
    p.input = inputs[r];
    p.output = outputs[r];
    for (int j = 0; j < nranks; ++j) p.region[j] = b + (uint64_t)j * perRank;
  }
  return aar_local_round(params, nranks, &lay, 0u, 0, stream);
}
