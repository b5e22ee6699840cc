// Bulk async-copy engine primitives (cp.async.bulk + mbarrier) and the tile pipeline the role-specialised kernels of
// aar_bulk.cu are built on.
//
// Why: the reference's scatter / broadcast / flush are array copies plus a message send (AllreduceWorker.scala:194-238,
// 180-192).  On the GPU the first generation of these roles moved every byte through registers (LDG.128 -> STG.128, 512
// threads x 128 registers per SM) and published a chunk with a system fence that drains the whole SM's store queue.  Here
// ONE thread per CTA drives a ring of shared-memory tiles: global -> shared by cp.async.bulk (completion counted in bytes on
// an mbarrier), shared -> (peer) global by cp.async.bulk as a bulk group.  A unit of work is complete when its bulk groups
// are, which the issuing thread learns from cp.async.bulk.wait_group without stalling the tiles behind it -- so arrival
// flags go out per segment while the copy keeps streaming.  Measured on B200 (bench/bulk_bench.py): 45-48 GB/s per CTA
// against 28-35 GB/s for the register loop.
//
// All bulk loads here are the plain (non-multicast) form; a peer GPU's memory is only ever the DESTINATION of a bulk store
// or the source of a plain bulk load.
#pragma once
#include "aar_protocol.cuh"

namespace aar_dev {

constexpr int kCopyTile = 32768;    // bytes per tile of the copy roles (== one layout tile: kThreads * kUnroll * 16)
constexpr int kCopyStages = 6;      // 192 KiB of loads + stores in flight per CTA
constexpr int kStageBytes = 221184; // 216 KiB: 6 x 32 KiB copy tiles, or 3 x (8 + 1) x 8 KiB reduce tiles at P = 8
constexpr int kZeroTile = 8192;     // zero-filled tile behind the stages: source of the "chunk never arrived" stores
constexpr int kBulkDynSmem = kStageBytes + kZeroTile;
constexpr int kMaxStages = 8;
constexpr int kDoneLag = 8;         // a unit is checked for completion once this many newer bulk groups have been committed

static __device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

static __device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
static __device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
static __device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
static __device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(done)
      : "r"(bar), "r"(parity)
      : "memory");
  return done != 0;
}
static __device__ __forceinline__ void bulk_g2s(uint32_t dstSmem, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dstSmem), "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
static __device__ __forceinline__ void bulk_s2g(void* dst, uint32_t srcSmem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(srcSmem), "r"(bytes) : "memory");
}
static __device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
static __device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
static __device__ __forceinline__ void bulk_wait() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// generic-proxy accesses before / async-proxy accesses after (and the other way round) are ordered across this fence
static __device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }
static __device__ __forceinline__ void fence_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
static __device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Units whose last byte is written, handed from the thread that issues the copies to the thread that publishes them.
// Why two threads: the flag of a unit needs a system fence after the unit's bulk stores have completed (a fence-less flag
// written after cp.async.bulk.wait_group WAS seen before the data by a peer: bench/litmus_bulk_flag.py, 64 stale words in
// 1.3 M rounds), and a fence issued by the thread that keeps the copies going stalls that thread for as long as the stores
// in flight take to be acknowledged.  The publishing thread fences (once per batch of finished units), the issuing thread
// never stops.  (Tried and measured worse: pausing the stores at the end of every unit so that the issuing thread's own
// fence is cheap -- the pause costs one link round trip per 256 KiB, 18 instead of 42 GB/s per pushing CTA.)
struct DoneRing {
  uint32_t a[16], b[16], c[16];
  uint32_t tail, head, closed;
};

struct BulkShared {
  unsigned long long full[kMaxStages];      // mbarriers: "the loads of the tile in this stage have landed"
  unsigned long long computed[kMaxStages];  // reduce role: "the result tile of this stage is written (and its inputs are read)"
  unsigned long long outFree[kMaxStages];   // reduce role: "the stores of this stage's result tile have read it"
  unsigned long long jobDst[kMaxStages];    // per stage: destination of the tile (copy roles) / byte offset inside the block (reduce)
  uint32_t jobBytes[kMaxStages];
  uint32_t jobInfo[kMaxStages];             // copy roles: bit 0 last tile of its unit, bit 1 zero-fill (no load)
  uint32_t unitA[kMaxStages], unitB[kMaxStages], unitC[kMaxStages];  // the unit's words for complete(), parked with its last tile
  uint32_t fifoGroup[16], fifoA[16], fifoB[16], fifoC[16];  // units whose stores are committed but not yet known complete
  DoneRing done;
  int bcast[4];
};

static __device__ __forceinline__ uint32_t ld_acquire_cta_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
  return v;
}
static __device__ __forceinline__ void st_release_cta_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}
static __device__ __forceinline__ void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
static __device__ __forceinline__ uint64_t atom_cas_acq_rel_gpu_u64(uint64_t* p, uint64_t cmp, uint64_t val) {
  uint64_t old;
  asm volatile("atom.acq_rel.gpu.global.cas.b64 %0, [%1], %2, %3;" : "=l"(old) : "l"(p), "l"(cmp), "l"(val) : "memory");
  return old;
}

// Wait for an mbarrier phase.  A wait that never ends here means a bad address or a broken pipeline, not a slow peer (the
// data is local or already flagged): report and trap rather than hang the GPU.
static __device__ __forceinline__ void mbar_wait(const AarRankParams& p, const AarLayout& l, uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const unsigned long long t0 = globaltimer_ns();
  unsigned spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 1023u) == 0 && globaltimer_ns() - t0 > 8000000000ull) {
      atomicOr(statusWord(p, l, 0), AAR_ST_BULK_FAULT);
      __threadfence_system();
      __trap();
    }
  }
}

// (Re)arm the pipeline of this CTA: called by every thread at the start of a role.  computedCount = number of arrivals
// that complete a `computed` phase (the reduce role's compute warps).
static __device__ __forceinline__ void bulkInit(BulkShared& bs, unsigned char* dsm, uint32_t computedCount = 1) {
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int s = 0; s < kMaxStages; ++s) {
      mbar_init(smem_u32(&bs.full[s]), 1);
      mbar_init(smem_u32(&bs.computed[s]), computedCount);
      mbar_init(smem_u32(&bs.outFree[s]), 1);
    }
    bs.done.tail = bs.done.head = bs.done.closed = 0;
    mbar_init_fence();
  }
  uint4* z = reinterpret_cast<uint4*>(dsm + kStageBytes);
  for (int i = threadIdx.x; i < kZeroTile / 16; i += kThreads) z[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();  // the zero tile is read by bulk stores
  __syncthreads();
}

// issuing side of the hand-over
static __device__ __forceinline__ void donePush(DoneRing& d, uint32_t a, uint32_t b, uint32_t c) {
  const uint32_t t = d.tail;  // only this thread writes it
  while (t - ld_acquire_cta_u32(&d.head) >= 16u) __nanosleep(32);
  d.a[t & 15] = a;
  d.b[t & 15] = b;
  d.c[t & 15] = c;
  st_release_cta_u32(&d.tail, t + 1);
}
static __device__ __forceinline__ void doneClose(DoneRing& d) { st_release_cta_u32(&d.closed, 1u); }
// publishing side: runs until the ring is closed and empty
template <class Role>
static __device__ void doneDrain(DoneRing& d, Role& role) {
  uint32_t h = 0;
  for (;;) {
    uint32_t t = ld_acquire_cta_u32(&d.tail);
    if (h == t) {
      if (ld_acquire_cta_u32(&d.closed)) {
        t = ld_acquire_cta_u32(&d.tail);
        if (h == t) return;
        continue;
      }
      __nanosleep(64);
      continue;
    }
    // ONE fence for everything that is waiting, then one relaxed store per flag
    fence_proxy_async();
    fence_sys();
    for (; h != t; ++h) role.complete(d.a[h & 15], d.b[h & 15], d.c[h & 15]);
    st_release_cta_u32(&d.head, h);
  }
}

// ------------------------------------------------------------------------------------------------ single-thread decisions
// The decisions of aar_protocol.cuh for a role whose only active thread is thread 0 (no barriers, caches in registers).
// Same words, same CAS protocol: CTAs running either form agree.

// Slot flow control for destination `dest` (see scatterDecision): 1 go, 0 drop, -1 not known yet (only when !blocking).
static __device__ int destDecision1(const AarRankParams& p, const AarLayout& l, int dest, unsigned& known, unsigned& go, bool blocking) {
  if ((known >> dest) & 1u) return (int)((go >> dest) & 1u);
  unsigned g = 1;
  if (p.round >= l.L) {
    char* my = regionOf(p, p.rank);
    uint64_t* word = reinterpret_cast<uint64_t*>(my + l.k1DecOff) + (p.round % l.L) * AAR_MAXP + dest;
    const uint64_t epoch = (uint64_t)(p.round + 1);
    const uint64_t v = ld_acquire_gpu_u64(word);
    if ((v >> 32) == epoch) {
      g = (unsigned)(v & 1u);
    } else {
      const uint32_t need = (uint32_t)(p.round - l.L + 1);
      const uint32_t* base = reinterpret_cast<const uint32_t*>(my + l.slotBaseOff) + (p.round % l.L) * AAR_MAXP + dest;
      if (ld_acquire_sys_u32(base) < need) {
        if (!blocking) return -1;
        Deadline dl(p.timeoutNs);
        unsigned spins = 0;
        for (;;) {
          if (ld_acquire_sys_u32(base) >= need) break;
          if (((++spins & 63u) == 0) && (forcedByHost(p, l) || dl.expired())) {
            g = 0;  // the peer fell out of the window: nothing is written into its slot
            break;
          }
          __nanosleep(64);
        }
      }
      const uint64_t old = atom_cas_acq_rel_gpu_u64(word, v, (epoch << 32) | g);
      if (old != v && (old >> 32) == epoch) g = (unsigned)(old & 1u);
      if (!g) atomicOr(statusWord(p, l, 0), AAR_ST_PUSH_DROPPED);
    }
  }
  known |= 1u << dest;
  if (g) go |= 1u << dest;
  return (int)g;
}

// ---- piece flags
// The bulk kernels publish arrival per PIECE (a fixed sub-chunk granule of the layout, at most 8 tiles = 256 KiB), not per
// chunk: exactly one CTA moves a piece, so its flag needs no election among CTAs -- one fence of the publishing thread, one
// relaxed store -- and the consumer of a reduced chunk can start on its first piece.  What the reference calls a ScatterBlock
// has arrived from a source when ALL pieces of the chunk carry the round's epoch; the contributor decision stays per chunk.
static __device__ __forceinline__ int piecesOf(const AarLayout& l, int clen) {
  const int pe = l.tilesPerPiece * l.tileElems;
  return (clen + pe - 1) / pe;
}
static __device__ __forceinline__ uint32_t* pflagS(const AarRankParams& p, const AarLayout& l, int rankOfRegion, int slot, int c, int s) {
  return reinterpret_cast<uint32_t*>(regionOf(p, rankOfRegion) + l.pflagSOff) + (((long long)slot * l.C + c) * l.piecesPerChunk + s) * AAR_MAXP;
}
static __device__ __forceinline__ uint64_t* pflagR(const AarRankParams& p, const AarLayout& l, int rankOfRegion, int slot, int owner, int c, int s) {
  return reinterpret_cast<uint64_t*>(regionOf(p, rankOfRegion) + l.pflagROff) + (((long long)slot * l.P + owner) * l.C + c) * l.piecesPerChunk + s;
}

// Contributor set of one of my chunks (see chunkDecision): wait until at least minRequired sources have delivered every piece
// of the chunk (or the round is forced), snapshot the arrived mask, make it the single decision for all CTAs.  Returns false
// only when !blocking and the quorum is not there yet.
static __device__ bool chunkDecision1(const AarRankParams& p, const AarLayout& l, int c, int clen, int minRequired, bool blocking, unsigned& maskOut) {
  const int slot = p.round % l.L;
  const uint64_t epoch = (uint64_t)(p.round + 1);
  uint64_t* word = reinterpret_cast<uint64_t*>(regionOf(p, p.rank) + l.k2DecOff) + (long long)slot * l.C + c;
  const uint64_t v = ld_acquire_gpu_u64(word);
  if ((v >> 32) == epoch) {
    maskOut = (unsigned)(v & 0xffffffffu);
    return true;
  }
  const int np = piecesOf(l, clen);
  const unsigned all = l.P >= 32 ? 0xffffffffu : ((1u << l.P) - 1u);
  auto look = [&](bool acquire) {
    unsigned a = all;
    for (int s = np - 1; s >= 0 && a; --s) {  // pieces land roughly in order: the last one decides most polls
      const uint32_t* f = pflagS(p, l, p.rank, slot, c, s);
      unsigned here = 0;
      for (int j = 0; j < l.P; ++j)
        if (((a >> j) & 1u) && (acquire ? ld_acquire_sys_u32(f + j) : ld_relaxed_sys_u32(f + j)) == (uint32_t)epoch) here |= 1u << j;
      a &= here;
    }
    return a;
  };
  unsigned arrived = look(false);
  bool forced = false;
  if (__popc(arrived) < minRequired) {
    if (!blocking) return false;
    Deadline dl(p.timeoutNs);
    unsigned spins = 0, backoff = 32;
    for (;;) {
      arrived = look(false);
      if (__popc(arrived) >= minRequired) break;
      if ((p.flags & AAR_FLAG_FORCE) || ((++spins & 31u) == 0)) {
        int stop = forcedByHost(p, l) ? 1 : (dl.expired() ? 2 : 0);
        if (stop) {
          if (stop == 2) atomicOr(statusWord(p, l, 0), AAR_ST_TIMEOUT_REDUCE);
          forced = true;
          break;
        }
      }
      __nanosleep(backoff);
      if (backoff < 512) backoff *= 2;
    }
  }
  // acquire: one more pass with acquire loads (a fence here would wait for the stores this SM has in flight); as fresh as the
  // decision, and the flagged sources' data is visible to everything after this point
  arrived = look(true);
  if ((p.flags & AAR_FLAG_EXACT_QUORUM) && !forced) {
    while (__popc(arrived) > minRequired) arrived &= ~(1u << (31 - __clz(arrived)));  // keep the lowest ids
  }
  unsigned mask = arrived;
  const uint64_t old = atom_cas_acq_rel_gpu_u64(word, v, (epoch << 32) | mask);  // release: whoever adopts the word also sees the data
  if (old != v && (old >> 32) == epoch) mask = (unsigned)(old & 0xffffffffu);
  maskOut = mask;
  return true;
}

// ------------------------------------------------------------------------------------------------ the copy pipeline
// A copy role describes its work as a sequence of units {src, dst, bytes}; src == nullptr means "fill with zeros",
// bytes == 0 means "nothing to move, only the completion action".  Role::nextUnit(u, blocking) returns 1 (unit), 0 (no
// more units) or -1 (the next unit's precondition -- a flag, a flow-control word -- is not there yet; only when !blocking).
//
// Role::complete(a, b, c) runs on the CTA's publishing thread (doneDrain) once the unit's last byte is written.
struct CopyUnit {
  const char* src;
  char* dst;
  uint32_t bytes;
  uint32_t a, b, c;  // opaque to the pipeline: handed back to complete()
};

template <class Role>
static __device__ void runCopyRole(const AarRankParams& p, const AarLayout& l, Role& role, BulkShared& bs, unsigned char* dsm) {
  // thread 0 only
  const uint32_t stage0 = smem_u32(dsm), zero = smem_u32(dsm + kStageBytes);
  uint32_t L = 0, S = 0, head = 0, tail = 0, phaseBits = 0;
  bool exhausted = false, haveUnit = false;
  CopyUnit u;
  uint32_t off = 0;
  auto retireDone = [&](uint32_t lastCompleteGroup) {
    while (head != tail && bs.fifoGroup[head & 15] <= lastCompleteGroup) {
      if (Role::kPublishes) donePush(bs.done, bs.fifoA[head & 15], bs.fifoB[head & 15], bs.fifoC[head & 15]);
      ++head;
    }
  };
  for (;;) {
    while (L - S <= (uint32_t)(kCopyStages - 2) && !exhausted) {
      if (!haveUnit) {
        if (L == S && head != tail) {  // about to wait with nothing in flight: do not sit on finished units' flags
          bulk_wait<0>();
          retireDone(S - 1);
        }
        const int r = role.nextUnit(u, L == S);
        if (r > 0 && L == 0) traceMarkAny(p, l, 6);  // first unit in hand (its flag / flow-control word was there)
        if (r == 0) {
          exhausted = true;
          break;
        }
        if (r < 0) break;
        if (u.bytes == 0) {
          if (Role::kPublishes) donePush(bs.done, u.a, u.b, u.c);
          continue;
        }
        haveUnit = true;
        off = 0;
      }
      const uint32_t n = (u.bytes - off < (uint32_t)kCopyTile) ? u.bytes - off : (uint32_t)kCopyTile;
      const int st = (int)(L % kCopyStages);
      const bool last = off + n == u.bytes;
      bs.jobDst[st] = (unsigned long long)(u.dst + off);
      bs.jobBytes[st] = n;
      bs.jobInfo[st] = (last ? 1u : 0u) | (u.src ? 0u : 2u);
      if (last) {
        bs.unitA[st] = u.a;  // parked until the store is issued
        bs.unitB[st] = u.b;
        bs.unitC[st] = u.c;
      }
      if (u.src) {
        const uint32_t bar = smem_u32(&bs.full[st]);
        mbar_expect_tx(bar, n);
        bulk_g2s(stage0 + (uint32_t)st * kCopyTile, u.src + off, n, bar);
      }
      ++L;
      off += n;
      if (last) haveUnit = false;
    }
    if (S == L) {
      if (exhausted) break;
      continue;  // the next nextUnit() call blocks
    }
    const int st = (int)(S % kCopyStages);
    const uint32_t info = bs.jobInfo[st], n = bs.jobBytes[st];
    char* dst = reinterpret_cast<char*>(bs.jobDst[st]);
    if (info & 2u) {
      for (uint32_t o = 0; o < n; o += kZeroTile) bulk_s2g(dst + o, zero, (n - o < (uint32_t)kZeroTile) ? n - o : (uint32_t)kZeroTile);
    } else {
      mbar_wait(p, l, smem_u32(&bs.full[st]), (phaseBits >> st) & 1u);
      phaseBits ^= 1u << st;
      bulk_s2g(dst, stage0 + (uint32_t)st * kCopyTile, n);
    }
    bulk_commit();
    if (Role::kPublishes && (info & 1u)) {
      bs.fifoGroup[tail & 15] = S;
      bs.fifoA[tail & 15] = bs.unitA[st];
      bs.fifoB[tail & 15] = bs.unitB[st];
      bs.fifoC[tail & 15] = bs.unitC[st];
      ++tail;
    }
    ++S;
    bulk_wait_read<1>();  // the store before this one has finished reading its stage: that stage may be refilled
    if (head != tail && S > (uint32_t)kDoneLag && bs.fifoGroup[head & 15] + kDoneLag < S) {
      bulk_wait<kDoneLag>();  // groups up to S - 1 - kDoneLag are complete (written, visible to this thread)
      retireDone(S - 1 - kDoneLag);
    }
  }
  bulk_wait<0>();
  if (S > 0) retireDone(S - 1);
  if (Role::kPublishes) doneClose(bs.done);
}

}  // namespace aar_dev
