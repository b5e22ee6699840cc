// Device-side protocol pieces shared by every kernel of the data plane: the per-CTA scratch words, the segment plan,
// and the decisions that several CTAs (and several kernels of one round) must agree on -- slot flow control per
// destination, the contributor set of a chunk, the round's gather snapshot -- plus the retirement of a ring slot.
// Reference steps: DataBuffer.scala:31-33,69-75 (thresholds), AllreduceWorker.scala:91-97 (catch-up), DataBuffer.scala:63-67 (up()).
#pragma once
#include "aar_device.cuh"

namespace aar_dev {
struct Smem {
  unsigned k1Known, k1Go;       // per-destination flow-control decisions cached by this CTA
  int k2Chunk;                  // chunk whose contributor mask is cached
  unsigned k2Mask;
  int k3Leader;
  int k3Count;
  int scratch;
  int gKey, gCnt;               // eager gather: (owner, chunk) whose flag this CTA has already seen, and its count
};

static __device__ __forceinline__ char* regionOf(const AarRankParams& p, int r) { return reinterpret_cast<char*>(p.region[r]); }
static __device__ __forceinline__ uint32_t* statusWord(const AarRankParams& p, const AarLayout& l, int i) {
  return reinterpret_cast<uint32_t*>(regionOf(p, p.rank) + l.statusOff) + i;
}
// Tracing (the reference has none): per-CTA %globaltimer stamps at phase boundaries.
static __device__ __forceinline__ void traceMark(const AarRankParams& p, const AarLayout& l, int mark) {
  if ((p.flags & AAR_FLAG_TRACE) && threadIdx.x == 0 && blockIdx.x < AAR_TRACE_CTAS)
    reinterpret_cast<unsigned long long*>(regionOf(p, p.rank) + l.traceOff)[blockIdx.x * AAR_TRACE_MARKS + mark] = globaltimer_ns();
}
// the same from whichever thread runs the role
static __device__ __forceinline__ void traceMarkAny(const AarRankParams& p, const AarLayout& l, int mark) {
  if ((p.flags & AAR_FLAG_TRACE) && blockIdx.x < AAR_TRACE_CTAS)
    reinterpret_cast<unsigned long long*>(regionOf(p, p.rank) + l.traceOff)[blockIdx.x * AAR_TRACE_MARKS + mark] = globaltimer_ns();
}
template <typename T>
static __device__ __forceinline__ bool geomVecOk(const AarLayout& l) {
  constexpr int V = Elem<T>::kPerVec;
  return (l.stepSize % V == 0) && (l.maxChunk % V == 0) && (l.blkStride % V == 0);
}
static __device__ __forceinline__ long long paddedData(const AarLayout& l) {
  const int V = 16 / l.elemSize;
  return ((long long)l.dataSize + V - 1) / V * V;
}

// Segments per message so that `pairs` messages give about two units per CTA; whole tiles per segment.
struct SegPlan {
  int tilesPerSeg, segs;
  // unitsPerCta = 2 balances well when segments are cheap to finish; = 1 gives every CTA exactly one segment and therefore
  // one fence, so its memory pipeline never drains in the middle (used where the link latency is long: NVLS reduce).
  __device__ SegPlan(const AarLayout& l, long long pairs, int nctas, int unitsPerCta = 2) {
    const int Tm = l.tilesPerChunk;
    long long want = ((long long)unitsPerCta * nctas + pairs - 1) / (pairs > 0 ? pairs : 1);
    int S = (int)(want < 1 ? 1 : (want > Tm ? Tm : want));
    tilesPerSeg = (Tm + S - 1) / S;
    segs = (Tm + tilesPerSeg - 1) / tilesPerSeg;
  }
  __device__ int segElems(const AarLayout& l) const { return tilesPerSeg * l.tileElems; }
  // segments of a message of `len` elements that are not empty
  __device__ unsigned liveSegs(const AarLayout& l, int len) const { return (unsigned)((len + segElems(l) - 1) / segElems(l)); }
};

// ------------------------------------------------------------------------------------------------ K1
// Flow control (replaces the unbounded actor mailbox): round R may overwrite slot R % L of peer `dest` only
// after that peer retired round R - L.  One decision per (round, dest), shared by all CTAs through a CAS word.
static __device__ bool scatterDecision(const AarRankParams& p, const AarLayout& l, int dest, Smem& sm) {
  if (!((sm.k1Known >> dest) & 1u)) {
    // Every thread must have evaluated the condition above before thread 0 changes k1Known: a warp that read the
    // updated bit would skip this block and pair its next barrier with the one below.
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned go = 1;
      if (p.round >= l.L) {
        char* my = regionOf(p, p.rank);
        uint64_t* word = reinterpret_cast<uint64_t*>(my + l.k1DecOff) + (p.round % l.L) * AAR_MAXP + dest;
        const uint64_t epoch = (uint64_t)(p.round + 1);
        uint64_t v = ld_acquire_gpu_u64(word);
        if ((v >> 32) == epoch) {
          go = (unsigned)(v & 1u);
        } else {
          const uint32_t need = (uint32_t)(p.round - l.L + 1);
          const uint32_t* base = reinterpret_cast<const uint32_t*>(my + l.slotBaseOff) + (p.round % l.L) * AAR_MAXP + dest;
          Deadline dl(p.timeoutNs);
          unsigned spins = 0;
          for (;;) {
            if (ld_acquire_sys_u32(base) >= need) break;
            if (((++spins & 63u) == 0) && (forcedByHost(p, l) || dl.expired())) {
              go = 0;  // the peer fell out of the window: its copy of this contribution is dropped
              break;
            }
            __nanosleep(64);
          }
          __threadfence();
          uint64_t old = atomicCAS(reinterpret_cast<unsigned long long*>(word), (unsigned long long)v,
                                   (unsigned long long)((epoch << 32) | go));
          if (old != v && (old >> 32) == epoch) go = (unsigned)(old & 1u);
          if (!go) atomicOr(statusWord(p, l, 0), AAR_ST_PUSH_DROPPED);
        }
      }
      sm.k1Known |= (1u << dest);
      if (go) sm.k1Go |= (1u << dest);
    }
    __syncthreads();
  }
  return (sm.k1Go >> dest) & 1u;
}

// ------------------------------------------------------------------------------------------------ K2
// Contributor set of one of my chunks: wait until at least minRequired sources have flagged it (or the host forces
// the round), snapshot the arrived mask, and make that snapshot the single decision for all CTAs (exactly-once
// firing, DataBuffer.scala:31-33 -- without the == overshoot bug B4).  Runs in warp 0.
static __device__ unsigned chunkDecision(const AarRankParams& p, const AarLayout& l, uint64_t* word, const uint32_t* flags, int minRequired) {
  const int lane = threadIdx.x & 31;
  const uint64_t epoch = (uint64_t)(p.round + 1);
  uint64_t v = 0;
  if (lane == 0) v = ld_acquire_gpu_u64(word);
  v = __shfl_sync(0xffffffffu, v, 0);
  if ((v >> 32) == epoch) return (unsigned)(v & 0xffffffffu);

  Deadline dl(p.timeoutNs);
  unsigned arrived = 0, spins = 0, backoff = 32;
  bool forced = false;
  // Poll with relaxed loads (one 64-byte line holds all P flags) and pay for the acquire once, after the quorum is
  // seen.  The host doorbell lives across PCIe: look at it rarely (the AAR_FLAG_FORCE bit is checked up front).
  for (;;) {
    uint32_t f = (lane < l.P) ? ld_relaxed_sys_u32(flags + lane) : 0u;
    arrived = __ballot_sync(0xffffffffu, lane < l.P && f == (uint32_t)epoch);
    if (__popc(arrived) >= minRequired) break;
    const bool look = (p.flags & AAR_FLAG_FORCE) || ((++spins & 31u) == 0);
    if (look) {
      int stop = 0;
      if (lane == 0) {
        if (forcedByHost(p, l)) stop = 1;
        else if (dl.expired()) stop = 2;
      }
      stop = __shfl_sync(0xffffffffu, stop, 0);
      if (stop) {
        if (stop == 2 && lane == 0) atomicOr(statusWord(p, l, 0), AAR_ST_TIMEOUT_REDUCE);
        forced = true;
        // one more look so the snapshot is as fresh as the decision
        f = (lane < l.P) ? ld_relaxed_sys_u32(flags + lane) : 0u;
        arrived = __ballot_sync(0xffffffffu, lane < l.P && f == (uint32_t)epoch);
        break;
      }
    }
    __nanosleep(backoff);
    if (backoff < 512) backoff *= 2;
  }
  __threadfence_system();  // acquire: the flagged sources' data is visible to everything after this point
  if ((p.flags & AAR_FLAG_EXACT_QUORUM) && !forced) {
    while (__popc(arrived) > minRequired) arrived &= ~(1u << (31 - __clz(arrived)));  // keep the lowest ids
  }
  unsigned mask = arrived;
  if (lane == 0) {
    __threadfence();
    uint64_t old = atomicCAS(reinterpret_cast<unsigned long long*>(word), (unsigned long long)v,
                             (unsigned long long)((epoch << 32) | mask));
    if (old != v && (old >> 32) == epoch) mask = (unsigned)(old & 0xffffffffu);
  }
  return __shfl_sync(0xffffffffu, mask, 0);
}

// Slot flow control for the broadcast (thresholds < 1 only; found by csrc/tests/protocol_litmus.cpp).  With a partial
// quorum the owner can finish reducing round R while a peer that was not in the quorum is still gathering round R - L
// from the very reduce-ring slot the broadcast of round R writes.  The scatter already knows the answer: its per-
// destination decision "peer retired round R - L, go" (or "peer fell out of the window, drop").  The broadcast adopts
// that decision -- it waits for it like a push CTA would and skips the destinations that were dropped.  peerBase is
// monotone, so a "go" seen by any CTA stays true.  Called by all threads of the CTA.
static __device__ unsigned bcastDestMask(const AarRankParams& p, const AarLayout& l, Smem& sm) {
  unsigned m = 0;
  for (int i2 = 0; i2 < l.P; ++i2) {
    const int dest = (i2 + p.rank) % l.P;
    if (dest == p.rank || scatterDecision(p, l, dest, sm)) m |= 1u << dest;
  }
  if (p.flags & AAR_FLAG_ELIDE_LATE) {
    // late-work elision: a destination that has already retired this round will never gather this rank's chunks
    __shared__ unsigned k2Dest;
    __syncthreads();
    if (threadIdx.x == 0) {
      // "has retired THIS round": the slot's word (rounds retire out of order, the newest-round word would not do)
      const uint32_t* base = reinterpret_cast<const uint32_t*>(regionOf(p, p.rank) + l.slotBaseOff) + (p.round % l.L) * AAR_MAXP;
      unsigned e = m;
      for (int d = 0; d < l.P; ++d)
        if (d != p.rank && (int)ld_relaxed_sys_u32(base + d) > p.round) e &= ~(1u << d);
      k2Dest = e;
    }
    __syncthreads();
    m = k2Dest;
  }
  return m;
}

// K6 ring_advance: the last CTA retires the slot -- publish "round+1 rounds retired" to every peer (their flow
// control) and to the host (CompleteAllreduce is posted from there).  No bulk zeroing: epochs make stale data
// unreachable (the reference re-allocates and zero-fills the slot, DataBuffer.scala:63-67).
struct NoCleanup {
  __device__ void operator()(const AarRankParams&, const AarLayout&) const {}
};
// `cleanup` runs on thread 0 of the round's last CTA after every other CTA has finished and before the slot is published
// as retired (used by paths that repair a partially flushed chunk).
template <class Cleanup>
static __device__ void retireRoundWith(const AarRankParams& p, const AarLayout& l, int nctas, Cleanup cleanup) {
  char* my = regionOf(p, p.rank);
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    unsigned* ctr = reinterpret_cast<unsigned*>(my + l.k3CtrOff) + (p.round % l.L);
    const unsigned old = atomicAdd(ctr, 1u);
    if (old + 1 == (unsigned)nctas) {
      atomicExch(ctr, 0u);
      __threadfence();
      cleanup(p, l);
      // Retirement is per ring slot: peers read slotBase[slot][me] before they overwrite MY slot (round R needs "retired R - L"
      // in slot R % L), so rounds of one rank may retire in any order -- round r+1 before round r when its quorum lands first
      // (AllreduceWorker.scala:253-268).  peerBase[me] is only raised: the newest round this rank has retired (catch-up predicate).
      __threadfence_system();
      const int slot = p.round % l.L;
      for (int i2 = 0; i2 < l.P; ++i2) {
        const int dest = (i2 + p.rank) % l.P;
        st_release_sys_u32(reinterpret_cast<uint32_t*>(regionOf(p, dest) + l.slotBaseOff) + slot * AAR_MAXP + p.rank, (uint32_t)(p.round + 1));
        asm volatile("red.relaxed.sys.global.max.u32 [%0], %1;" ::"l"(reinterpret_cast<uint32_t*>(regionOf(p, dest) + l.peerBaseOff) + p.rank),
                     "r"((uint32_t)(p.round + 1))
                     : "memory");
      }
      atomicMax(statusWord(p, l, 1), (uint32_t)(p.round + 1));  // rounds flushed
      if (p.hostDone) st_release_sys_u32(p.hostDone + slot, (uint32_t)(p.round + 1));
    }
  }
}

static __device__ void retireRound(const AarRankParams& p, const AarLayout& l, int nctas) { retireRoundWith(p, l, nctas, NoCleanup{}); }

// ------------------------------------------------------------------------------------------------ K3
// One CTA becomes the leader of the round: it counts reduced-chunk arrivals against minChunksRequired
// (DataBuffer.scala:69-75 with the true chunk total, B3), snapshots which chunks made it together with their
// contributor counts (B5), and publishes the snapshot; every CTA then gathers from that one snapshot.
static __device__ void gatherDecision(const AarRankParams& p, const AarLayout& l, int slot, Smem& sm) {
  char* my = regionOf(p, p.rank);
  uint64_t* word = reinterpret_cast<uint64_t*>(my + l.k3DecOff) + slot;
  const uint64_t epoch = (uint64_t)(p.round + 1);
  const uint64_t busy = (epoch << 32) | 1u, done = (epoch << 32) | 2u;
  if (threadIdx.x == 0) {
    int leader = 0;
    uint64_t v = ld_acquire_gpu_u64(word);
    if (v != done && v != busy) {
      uint64_t old = atomicCAS(reinterpret_cast<unsigned long long*>(word), (unsigned long long)v, (unsigned long long)busy);
      leader = (old == v) ? 1 : 0;
    }
    sm.k3Leader = leader;
  }
  __syncthreads();
  const int P = l.P, C = l.C;
  int* counts = reinterpret_cast<int*>(my + l.countsOff) + (long long)slot * l.totalChunks;
  if (sm.k3Leader) {
    const uint64_t* flags = reinterpret_cast<const uint64_t*>(my + l.rflagOff) + (long long)slot * P * C;
    Deadline dl(p.timeoutNs);
    unsigned spins = 0, backoff = 32;
    for (;;) {
      if (threadIdx.x == 0) sm.k3Count = 0;
      __syncthreads();
      int local = 0;
      for (int i = threadIdx.x; i < P * C; i += kThreads) {
        const int owner = i / C, c = i % C;
        if (c < numChunks(l, owner) && (ld_relaxed_sys_u64(flags + i) >> 32) == epoch) ++local;
      }
      if (local) atomicAdd(&sm.k3Count, local);
      __syncthreads();
      int stop = (sm.k3Count >= p.minChunksRequired) ? 1 : 0;
      if (!stop && threadIdx.x == 0 && ((p.flags & AAR_FLAG_FORCE) || (++spins & 127u) == 0)) {
        if (forcedByHost(p, l)) sm.scratch = 1;
        else if (dl.expired()) {
          sm.scratch = 1;
          atomicOr(statusWord(p, l, 0), AAR_ST_TIMEOUT_GATHER);
        }
      }
      __syncthreads();
      if (stop || sm.scratch) break;
      __nanosleep(backoff);
      if (backoff < 512) backoff *= 2;
    }
    __threadfence_system();  // acquire side of the rflag releases
    // snapshot (a superset of what was counted: anything that landed in between is welcome)
    for (int i = threadIdx.x; i < P * C; i += kThreads) {
      const int owner = i / C, c = i % C;
      if (c >= numChunks(l, owner)) continue;
      const uint64_t f = ld_relaxed_sys_u64(flags + i);
      const int cnt = ((f >> 32) == epoch) ? (int)(f & 0xffffffffu) : -1;
      counts[i] = cnt;  // valid (owner, c) pairs are exactly the indices [0, totalChunks)
      if (p.outCounts) p.outCounts[i] = cnt < 0 ? 0 : cnt;
    }
    __threadfence_system();  // chunks that landed between the count and the snapshot
    __syncthreads();
    if (threadIdx.x == 0) {
      sm.scratch = 0;
      __threadfence();
      st_release_gpu_u64(word, done);
    }
  } else {
    if (threadIdx.x == 0) {
      unsigned backoff = 64;
      while (ld_acquire_gpu_u64(word) != done) {
        __nanosleep(backoff);
        if (backoff < 1024) backoff *= 2;
      }
      __threadfence_system();
    }
  }
  __syncthreads();
}

static __device__ __forceinline__ void initSmem(Smem& sm) {
  if (threadIdx.x == 0) {
    sm.k1Known = 0;
    sm.k1Go = 0;
    sm.k2Chunk = -1;
    sm.k2Mask = 0;
    sm.k3Leader = 0;
    sm.k3Count = 0;
    sm.scratch = 0;
    sm.gKey = -1;
    sm.gCnt = -1;
  }
  __syncthreads();
}

}  // namespace aar_dev
