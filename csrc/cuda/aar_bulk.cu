// The allreduce round on the bulk async-copy engine: K1 scatter_push, K2 reduce_bcast, K3 gather_flush (and their NVLS
// forms) rebuilt on cp.async.bulk + mbarrier, one launch per round with the CTAs specialised by role.
//
// Reference steps: scatter AllreduceWorker.scala:194-209 -> PushRole / StageRole; store + threshold +
// reduce + broadcast :106-127,230-251 -> roleReduceP2P / roleReduceNvls; store + round threshold + flush + up() :129-150,
// 180-192,253-268 -> the gather roles + retireRound.  Flags, decision words, thresholds and the ring are those of
// aar_kernels.cu (aar_protocol.cuh): a rank may run any round with either generation of kernels.
//
// What is different from the register kernels:
//   * every copy is bulk: global -> shared ring -> (peer) global, issued by ONE thread per CTA; the other threads of a
//     copy CTA are idle (the reduce role uses all 512 for the adds, reading and writing shared-memory tiles only);
//   * completion of a segment = completion of its bulk groups (cp.async.bulk.wait_group), so an arrival flag leaves while
//     the tiles behind it keep streaming -- no system fence that drains the SM's whole store queue per segment;
//   * a role's CTAs sweep the block chunk by chunk TOGETHER (segments sized to the role's CTA count), so chunk c is flagged
//     after 1/C of the role's time and the next role starts on it: push, reduce and gather of one round overlap;
//   * AAR_FLAG_SELF_DIRECT: a rank's own contribution is read from its input instead of being copied into its own ring,
//     and (when every chunk is required) its own reduced block is written straight to the output;
//   * the in-switch (NVLS) form publishes and consumes PIECES (sub-chunk granules): it has no contributor decision to
//     take, so nothing forces it to wait for a whole 4 MiB chunk;
//   * always strict: the broadcast obeys the slot flow control, the eager gather decides once per chunk.
#include <cuda_runtime.h>

#include <mutex>

#include "aar_bulk.cuh"
#include "aar_host.h"

using namespace aar_dev;

namespace {

__device__ __forceinline__ bool eagerMode(const AarRankParams& p, const AarLayout& l) {
  return p.minChunksRequired >= l.totalChunks && !(p.flags & AAR_FLAG_FORCE);
}
__device__ __forceinline__ bool selfDirect(const AarRankParams& p) { return (p.flags & AAR_FLAG_SELF_DIRECT) != 0; }
__device__ __forceinline__ bool selfOut(const AarRankParams& p, const AarLayout& l) {
  return selfDirect(p) && eagerMode(p, l) && !(p.flags & AAR_FLAG_VIEW_OUTPUT);
}

// Segments per message such that `lanes` messages (the ones a role works on at the same time: one chunk of every
// destination) give about one segment per CTA: the role's CTAs finish chunk c together and move on to chunk c + 1.
struct SweepPlan {
  int tilesPerSeg, segs;
  __device__ SweepPlan(const AarLayout& l, int lanes, int nctas) {
    const int Tm = l.tilesPerChunk;
    int want = (nctas + lanes - 1) / (lanes > 0 ? lanes : 1);
    const int S = want < 1 ? 1 : (want > Tm ? Tm : want);
    tilesPerSeg = (Tm + S - 1) / S;
    segs = (Tm + tilesPerSeg - 1) / tilesPerSeg;
  }
  __device__ int segElems(const AarLayout& l) const { return tilesPerSeg * l.tileElems; }
  __device__ unsigned liveSegs(const AarLayout& l, int len) const { return (unsigned)((len + segElems(l) - 1) / segElems(l)); }
};

// ------------------------------------------------------------------------------------------------ K1 on the copy engine
template <typename T>
struct PushRole {
  const AarRankParams& p;
  const AarLayout& l;
  int nctas, slot, R, first, S, segElems;
  uint32_t epoch;
  long long total, idx;
  bool rotate;
  unsigned known, go;

  __device__ PushRole(const AarRankParams& p_, const AarLayout& l_, int cta, int nctas_) : p(p_), l(l_), nctas(nctas_), known(0), go(0) {
    slot = p.round % l.L;
    epoch = (uint32_t)(p.round + 1);
    first = selfDirect(p) ? 1 : 0;
    R = l.P - first;
    const SweepPlan plan(l, R > 0 ? R : 1, nctas);
    S = plan.segs;
    segElems = plan.segElems(l);
    total = (long long)l.C * R * S;
    idx = cta;
    rotate = R > 0 && (nctas % R) == 0;
  }
  // chunk-major, then segment, then destination staggered from self (AllreduceWorker.scala:195-196); when the CTA count is
  // a multiple of the destination count each pass rotates by one so a CTA does not always draw the same peer
  __device__ int nextUnit(CopyUnit& u, bool blocking) {
    for (; idx < total; idx += nctas) {
      const int i = (int)((idx % R + (rotate ? idx / nctas : 0)) % R);
      const int s = (int)((idx / R) % S);
      const int c = (int)(idx / ((long long)S * R));
      const int dest = (p.rank + first + i) % l.P;
      const int clen = chunkLen(l, dest, c);
      const int off = s * segElems;
      if (off >= clen) continue;
      const int d = destDecision1(p, l, dest, known, go, blocking);
      if (d < 0) return -1;
      if (d == 0) continue;  // the peer fell out of the window: its copy of this contribution is dropped
      const int n = min(segElems, clen - off);
      const long long inOff = blockStart(l, dest) + (long long)c * l.maxChunk + off;
      u.src = reinterpret_cast<const char*>(reinterpret_cast<const T*>(p.input) + inOff);
      u.dst = regionOf(p, dest) + l.scatterOff + (((long long)slot * l.P + p.rank) * l.blkStride + (long long)c * l.maxChunk + off) * (long long)sizeof(T);
      u.bytes = (uint32_t)n * (uint32_t)sizeof(T);
      u.a = (uint32_t)dest;
      u.b = (uint32_t)c;
      u.c = (uint32_t)clen;
      idx += nctas;
      return 1;
    }
    return 0;
  }
  // last segment of this "ScatterBlock": publish its arrival
  __device__ void complete(uint32_t dest, uint32_t c, uint32_t clen) {
    fence_proxy_async();
    fence_sys();
    unsigned* ctr = reinterpret_cast<unsigned*>(regionOf(p, p.rank) + l.k1CtrOff) + ((long long)slot * l.P + dest) * l.C + c;
    const unsigned old = atomicAdd(ctr, 1u);
    if (old + 1 == (unsigned)((clen + segElems - 1) / segElems)) {
      atomicExch(ctr, 0u);
      fence_sys();
      uint32_t* flag = reinterpret_cast<uint32_t*>(regionOf(p, dest) + l.sflagOff) + ((long long)slot * l.C + c) * AAR_MAXP + p.rank;
      st_release_sys_u32(flag, epoch);
    }
  }
};

template <typename T>
__device__ void rolePush(const AarRankParams& p, const AarLayout& l, int cta, int nctas, BulkShared& bs, unsigned char* dsm) {
  bulkInit(bs, dsm);
  if (p.flags & AAR_FLAG_SKIP_SCATTER) return;
  if (threadIdx.x != 0) return;
  if (selfDirect(p) && cta == 0) {
    // my own contribution stays where it is (the reduce role reads it from the input): only its arrival is announced
    const int slot = p.round % l.L;
    uint32_t* flags = reinterpret_cast<uint32_t*>(regionOf(p, p.rank) + l.sflagOff) + (long long)slot * l.C * AAR_MAXP + p.rank;
    const int mine = numChunks(l, p.rank);
    for (int c = 0; c < mine; ++c) st_release_sys_u32(flags + (long long)c * AAR_MAXP, (uint32_t)(p.round + 1));
  }
  PushRole<T> role(p, l, cta, nctas);
  if (role.R > 0) runCopyRole(p, l, role, bs, dsm);
}

// NVLS: the input is staged in MY region (the switch pulls it); only the flag travels, one per piece.
template <typename T>
struct StageRole {
  const AarRankParams& p;
  const AarLayout& l;
  int nctas, slot, Sp, pieceElems;
  uint32_t epoch;
  long long total, idx;
  unsigned known, go;
  char* stage;
  bool inPlace;

  __device__ StageRole(const AarRankParams& p_, const AarLayout& l_, int cta, int nctas_) : p(p_), l(l_), nctas(nctas_), known(0), go(0) {
    slot = p.round % l.L;
    epoch = (uint32_t)(p.round + 1);
    Sp = l.piecesPerChunk;
    pieceElems = l.tilesPerPiece * l.tileElems;
    total = (long long)l.C * l.P * Sp;
    idx = cta;
    stage = regionOf(p, p.rank) + l.stageOff + (long long)slot * paddedData(l) * (long long)sizeof(T);
    inPlace = reinterpret_cast<const char*>(p.input) == stage;  // the source wrote straight into the stage
  }
  __device__ int nextUnit(CopyUnit& u, bool blocking) {
    for (; idx < total; idx += nctas) {
      const int b = ((int)(idx % l.P) + p.rank) % l.P;
      const int s = (int)((idx / l.P) % Sp);
      const int c = (int)(idx / ((long long)Sp * l.P));
      const int clen = chunkLen(l, b, c);
      const int off = s * pieceElems;
      if (off >= clen) continue;
      // Owner b is the only reader of this part of my staging slot; it must have retired the slot's previous tenant
      // (round - L) before the slot is rewritten.  An owner that fell out of the window gets no staging write and no flag.
      const int d = destDecision1(p, l, b, known, go, blocking);
      if (d < 0) return -1;
      if (d == 0) continue;
      const int n = min(pieceElems, clen - off);
      const long long inOff = blockStart(l, b) + (long long)c * l.maxChunk + off;
      u.src = reinterpret_cast<const char*>(reinterpret_cast<const T*>(p.input) + inOff);
      u.dst = stage + inOff * (long long)sizeof(T);
      u.bytes = inPlace ? 0u : (uint32_t)n * (uint32_t)sizeof(T);
      u.a = (uint32_t)b;
      u.b = (uint32_t)c;
      u.c = (uint32_t)s;
      idx += nctas;
      return 1;
    }
    return 0;
  }
  __device__ void complete(uint32_t b, uint32_t c, uint32_t s) {
    fence_proxy_async();
    fence_sys();
    uint32_t* flag = reinterpret_cast<uint32_t*>(regionOf(p, (int)b) + l.pflagSOff) + (((long long)slot * l.C + c) * Sp + s) * AAR_MAXP + p.rank;
    st_release_sys_u32(flag, epoch);
  }
};

template <typename T>
__device__ void roleStage(const AarRankParams& p, const AarLayout& l, int cta, int nctas, BulkShared& bs, unsigned char* dsm) {
  bulkInit(bs, dsm);
  if (p.flags & AAR_FLAG_SKIP_SCATTER) return;
  if (threadIdx.x != 0) return;
  StageRole<T> role(p, l, cta, nctas);
  runCopyRole(p, l, role, bs, dsm);
}

// ------------------------------------------------------------------------------------------------ K2 on the copy engine
// Ordered fp32 sum of the decided contributor set (peer-index order from 0.0f, AllreduceWorker.scala:245-249), tile by
// tile: thread 0 bulk-loads the tile of every contributor into shared memory, all 512 threads add (one 16-byte vector
// each) and write the result tile to shared memory, thread 0 bulk-stores it into every destination's reduce ring.  The
// loads run up to (stages - 2) tiles ahead of the adds, across segment and chunk boundaries whenever the next chunk's
// contributor set is already decided.
template <typename T>
__device__ void roleReduceP2P(const AarRankParams& p, const AarLayout& l, int cta, int nctas, BulkShared& bs, unsigned char* dsm) {
  bulkInit(bs, dsm);
  constexpr int V = Elem<T>::kPerVec;
  const int P = l.P, C = l.C;
  const int myChunks = numChunks(l, p.rank);
  if (myChunks == 0) return;
  const int slot = p.round % l.L;
  const uint64_t epoch = (uint64_t)(p.round + 1);
  char* my = regionOf(p, p.rank);
  const bool sd = selfDirect(p), so = selfOut(p, l);
  const int tileB = P <= 8 ? 8192 : 4096;
  const int stageB = (P + 1) * tileB;
  int nst = kStageBytes / stageB;
  if (nst > 6) nst = 6;
  const SweepPlan plan(l, 1, nctas);
  const int S = plan.segs, segElems = plan.segElems(l);
  const long long total = (long long)myChunks * S;
  const uint32_t stage0 = smem_u32(dsm);
  const long long blockIn = blockStart(l, p.rank);

  auto unitLen = [&](long long idx) {  // elements of unit idx (0: nothing there)
    const int c = (int)(idx / S), s = (int)(idx % S);
    const int r = chunkLen(l, p.rank, c) - s * segElems;
    return r <= 0 ? 0 : (r < segElems ? r : segElems);
  };
  auto tilesOf = [&](int n) { return (int)(((long long)n * (long long)sizeof(T) + tileB - 1) / tileB); };

  // load side (thread 0)
  long long lIdx = cta;
  int lTile = 0, lTiles = 0, lOrd = -1, lLen = 0;
  bool lOpen = false;
  unsigned known = 0, go = 0, cachedMask = 0, lDest = 0;
  int cachedChunk = -1;
  uint32_t GL = 0, head = 0, tail = 0;
  // compute side (all threads)
  long long cIdx = cta;
  int cTile = 0, cTiles = 0, cOrd = -1, cLen = 0;
  uint32_t G = 0, phaseBits = 0;

  auto complete = [&](uint32_t c, uint32_t destCount, uint32_t clen) {
    const unsigned dm = destCount & 0xffffu, count = destCount >> 16;
    fence_proxy_async();
    fence_sys();
    unsigned* ctr = reinterpret_cast<unsigned*>(my + l.k2CtrOff) + (long long)slot * C + c;
    const unsigned old = atomicAdd(ctr, 1u);
    if (old + 1 == (unsigned)(((int)clen + segElems - 1) / segElems)) {  // last segment of this "ReduceBlock": tell every peer, count included
      atomicExch(ctr, 0u);
      if (so && p.outCounts) p.outCounts[p.rank * C + (int)c] = (int)count;
      fence_sys();
      const uint64_t word = (epoch << 32) | (uint64_t)count;
      for (int i2 = 0; i2 < P; ++i2) {
        const int dest = (i2 + p.rank) % P;
        if (!((dm >> dest) & 1u)) continue;  // no data was written there: no flag either
        st_release_sys_u64(reinterpret_cast<uint64_t*>(regionOf(p, dest) + l.rflagOff) + ((long long)slot * P + p.rank) * C + c, word);
      }
    }
  };
  auto retireDone = [&](uint32_t lastCompleteGroup) {
    while (head != tail && bs.fifoGroup[head & 15] <= lastCompleteGroup) {
      complete(bs.fifoA[head & 15], bs.fifoB[head & 15], bs.fifoC[head & 15]);
      ++head;
    }
  };

  for (;;) {
    if (threadIdx.x == 0) {
      while (GL - G <= (uint32_t)(nst - 2)) {
        if (!lOpen) {
          while (lIdx < total && unitLen(lIdx) == 0) lIdx += nctas;
          if (lIdx >= total) break;
          const int c = (int)(lIdx / S);
          const bool block = GL == G;
          if (block && head != tail) {  // about to wait with nothing in flight: do not sit on finished segments' flags
            bulk_wait<0>();
            retireDone(G - 1);
          }
          if (c != cachedChunk) {
            unsigned m = 0;
            if (!chunkDecision1(p, l, reinterpret_cast<uint64_t*>(my + l.k2DecOff) + (long long)slot * C + c,
                                reinterpret_cast<const uint32_t*>(my + l.sflagOff) + ((long long)slot * C + c) * AAR_MAXP, p.minRequired, block, m))
              break;
            fence_proxy_async();  // the flagged rows are read by bulk loads
            if (cachedChunk < 0) traceMark(p, l, 2);
            cachedChunk = c;
            cachedMask = m;
          }
          // Slot flow control for the broadcast: round R may overwrite slot R % L of a peer's reduce ring only after that
          // peer retired round R - L; a peer that fell out of the window is skipped (no data, no flag).
          unsigned dm = 0;
          bool ready = true;
          for (int i2 = 0; i2 < P; ++i2) {
            const int dest = (i2 + p.rank) % P;
            const int d = destDecision1(p, l, dest, known, go, block);
            if (d < 0) {
              ready = false;
              break;
            }
            if (d) dm |= 1u << dest;
          }
          if (!ready) break;
          ++lOrd;
          bs.openMask[lOrd % kMaxStages] = cachedMask;
          lDest = dm;
          lLen = unitLen(lIdx);
          lTiles = tilesOf(lLen);
          lTile = 0;
          lOpen = true;
        }
        const int st = (int)(GL % (uint32_t)nst);
        const int c = (int)(lIdx / S), s = (int)(lIdx % S);
        const long long elemOff = (long long)c * l.maxChunk + (long long)s * segElems;  // inside my block
        const uint32_t bytes = (uint32_t)min((long long)tileB, (long long)lLen * (long long)sizeof(T) - (long long)lTile * tileB);
        const uint32_t bar = smem_u32(&bs.full[st]);
        const int nsrc = __popc(cachedMask);
        bs.jobBytes[st] = bytes;
        bs.jobInfo[st] = lDest;
        if (nsrc == 0) {
          asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");  // nothing to load: the phase completes at once
        } else {
          mbar_expect_tx(bar, bytes * (uint32_t)nsrc);
          int j = 0;
          for (int src = 0; src < P; ++src) {
            if (!((cachedMask >> src) & 1u)) continue;
            const char* row = (src == p.rank && sd)
                                  ? reinterpret_cast<const char*>(reinterpret_cast<const T*>(p.input) + blockIn + elemOff)
                                  : my + l.scatterOff + (((long long)slot * P + src) * l.blkStride + elemOff) * (long long)sizeof(T);
            bulk_g2s(stage0 + (uint32_t)(st * stageB + j * tileB), row + (long long)lTile * tileB, bytes, bar);
            ++j;
          }
        }
        ++GL;
        if (++lTile == lTiles) {
          lOpen = false;
          lIdx += nctas;
        }
      }
    }
    // ---- compute side: every thread, in lockstep
    if (cTile == cTiles) {
      while (cIdx < total && unitLen(cIdx) == 0) cIdx += nctas;
      if (cIdx >= total) break;
      ++cOrd;
      cLen = unitLen(cIdx);
      cTiles = tilesOf(cLen);
      cTile = 0;
    }
    const int st = (int)(G % (uint32_t)nst);
    mbar_wait(p, l, smem_u32(&bs.full[st]), (phaseBits >> st) & 1u);
    phaseBits ^= 1u << st;
    const unsigned mask = bs.openMask[cOrd % kMaxStages];
    const int count = __popc(mask);
    const float divisor = ((p.flags & AAR_FLAG_NORMALIZE) && count > 1) ? (float)count : 1.f;
    const uint32_t bytes = (uint32_t)min((long long)tileB, (long long)cLen * (long long)sizeof(T) - (long long)cTile * tileB);
    const unsigned char* in = dsm + (size_t)st * stageB;
    unsigned char* out = dsm + (size_t)st * stageB + (size_t)P * tileB;
    for (uint32_t i = threadIdx.x; i < bytes / 16; i += kThreads) {
      float acc[V];
#pragma unroll
      for (int k = 0; k < V; ++k) acc[k] = 0.f;
      int j = 0;
      for (int src = 0; src < P; ++src) {
        if (!((mask >> src) & 1u)) continue;
        const uint4 v = *reinterpret_cast<const uint4*>(in + (size_t)j * tileB + (size_t)i * 16);
        float f[V];
        Elem<T>::unpack(v, f);
#pragma unroll
        for (int k = 0; k < V; ++k) acc[k] += f[k];
        ++j;
      }
      if (divisor != 1.f) {
#pragma unroll
        for (int k = 0; k < V; ++k) acc[k] = acc[k] / divisor;
      }
      *reinterpret_cast<uint4*>(out + (size_t)i * 16) = Elem<T>::pack(acc);
    }
    fence_proxy_async_smem();  // the result tile is read by the bulk stores below
    __syncthreads();
    if (threadIdx.x == 0) {
      const int c = (int)(cIdx / S), s = (int)(cIdx % S);
      const long long elemOff = (long long)c * l.maxChunk + (long long)s * segElems;
      const long long byteOff = elemOff * (long long)sizeof(T) + (long long)cTile * tileB;
      const unsigned dm = bs.jobInfo[st];
      const uint32_t outS = stage0 + (uint32_t)(st * stageB + P * tileB);
      for (int i2 = 0; i2 < P; ++i2) {
        const int dest = (i2 + p.rank) % P;
        if (!((dm >> dest) & 1u)) continue;
        char* dst = (dest == p.rank && so)
                        ? reinterpret_cast<char*>(p.output) + blockIn * (long long)sizeof(T) + byteOff
                        : regionOf(p, dest) + l.reduceOff + ((long long)slot * P + p.rank) * l.blkStride * (long long)sizeof(T) + byteOff;
        bulk_s2g(dst, outS, bytes);
      }
      bulk_commit();
      if (cTile + 1 == cTiles) {
        bs.fifoGroup[tail & 15] = G;
        bs.fifoA[tail & 15] = (uint32_t)c;
        bs.fifoB[tail & 15] = dm | ((uint32_t)count << 16);
        bs.fifoC[tail & 15] = (uint32_t)chunkLen(l, p.rank, c);
        ++tail;
      }
      bulk_wait_read<1>();
      if (head != tail && G + 1 > (uint32_t)kDoneLag && bs.fifoGroup[head & 15] + kDoneLag < G + 1) {
        bulk_wait<kDoneLag>();
        retireDone(G - kDoneLag);
      }
    }
    ++G;
    if (++cTile == cTiles) cIdx += nctas;
  }
  if (threadIdx.x == 0) {
    bulk_wait<0>();
    if (G > 0) retireDone(G - 1);
    traceMark(p, l, 3);
  }
}

// NVLS (K4): the NVSwitch sums the staged inputs of all ranks (multimem.ld_reduce) and replicates the result into every
// rank's reduce ring (multimem.st); per direction each GPU moves about N(1 + 1/P) bytes instead of 2N(P-1)/P.
template <typename T>
__device__ __forceinline__ void nvlsSpanBulk(const char* mcStage, char* mcOut, int nv) {
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

// One piece of my block at a time: wait until every rank has staged it, let the switch reduce and replicate it, tell
// every rank.  A piece that some rank never stages (it fell out of the window, or the round was forced) is not reduced
// and not published: the switch cannot leave a rank out of the sum, so "count < P" does not exist on this path -- the
// gather of every rank gives up on it and the chunk is flushed as zeros with count 0.
template <typename T>
__device__ void roleReduceNvls(const AarRankParams& p, const AarLayout& l, int cta, int nctas, BulkShared& bs) {
  __syncthreads();
  const int P = l.P, C = l.C, Sp = l.piecesPerChunk;
  const int myChunks = numChunks(l, p.rank);
  if (myChunks == 0) return;
  const int slot = p.round % l.L;
  const uint32_t epoch = (uint32_t)(p.round + 1);
  char* my = regionOf(p, p.rank);
  const int pieceElems = l.tilesPerPiece * l.tileElems;
  const long long total = (long long)myChunks * Sp;
  bool first = true;
  for (long long idx = cta; idx < total; idx += nctas) {
    const int c = (int)(idx / Sp), s = (int)(idx % Sp);
    const int clen = chunkLen(l, p.rank, c);
    const int off = s * pieceElems;
    if (off >= clen) continue;
    const int n = min(pieceElems, clen - off);
    if (threadIdx.x == 0) {
      const uint32_t* flags = reinterpret_cast<const uint32_t*>(my + l.pflagSOff) + (((long long)slot * C + c) * Sp + s) * AAR_MAXP;
      Deadline dl(p.timeoutNs);
      unsigned spins = 0, backoff = 32;
      int ok = 0;
      for (;;) {
        int have = 0;
        for (int j = 0; j < P; ++j) have += ld_relaxed_sys_u32(flags + j) == epoch ? 1 : 0;
        if (have == P) {
          ok = 1;
          break;
        }
        if ((p.flags & AAR_FLAG_FORCE) || ((++spins & 31u) == 0)) {
          if (forcedByHost(p, l)) break;
          if (dl.expired()) {
            atomicOr(statusWord(p, l, 0), AAR_ST_TIMEOUT_REDUCE);
            break;
          }
        }
        __nanosleep(backoff);
        if (backoff < 512) backoff *= 2;
      }
      __threadfence_system();  // acquire side of the stage flags
      bs.bcast[0] = ok;
      if (first) traceMark(p, l, 2);
    }
    first = false;
    __syncthreads();
    const int ok = bs.bcast[0];
    const long long outElemOff = ((long long)slot * P + p.rank) * l.blkStride + (long long)c * l.maxChunk + off;
    if (ok) {
      const char* mcStage = reinterpret_cast<const char*>(p.mcRegion) + l.stageOff +
                            ((long long)slot * paddedData(l) + blockStart(l, p.rank) + (long long)c * l.maxChunk + off) * (long long)sizeof(T);
      char* mcOut = reinterpret_cast<char*>(p.mcRegion) + l.reduceOff + outElemOff * (long long)sizeof(T);
      nvlsSpanBulk<T>(mcStage, mcOut, n / Elem<T>::kPerVec);
    }
    __syncthreads();
    if (threadIdx.x == 0 && ok) {
      __threadfence_system();
      for (int i2 = 0; i2 < P; ++i2) {
        const int dest = (i2 + p.rank) % P;
        uint32_t* flag = reinterpret_cast<uint32_t*>(regionOf(p, dest) + l.pflagROff) + (((long long)slot * P + p.rank) * C + c) * Sp + s;
        st_release_sys_u32(flag, epoch);
      }
    }
  }
  if (threadIdx.x == 0) traceMark(p, l, 3);
}

// ------------------------------------------------------------------------------------------------ K3 on the copy engine
// Eager gather (every chunk required): a segment of chunk (owner, c) is copied as soon as the chunk's ReduceBlock flag
// lands, while other chunks are still in flight; a chunk that never shows up (forced round, timeout) is zero-filled --
// decided once per chunk (eagerChunkDecision1).
template <typename T>
struct GatherEagerRole {
  const AarRankParams& p;
  const AarLayout& l;
  int nctas, slot, R, first, S, segElems, lastKey, lastCnt;
  long long total, idx;
  bool view;

  __device__ GatherEagerRole(const AarRankParams& p_, const AarLayout& l_, int cta, int nctas_) : p(p_), l(l_), nctas(nctas_), lastKey(-1), lastCnt(-1) {
    slot = p.round % l.L;
    first = selfOut(p, l) ? 1 : 0;  // my own block was written by my reduce role
    R = l.P - first;
    const SweepPlan plan(l, R > 0 ? R : 1, nctas);
    S = plan.segs;
    segElems = plan.segElems(l);
    total = (long long)l.C * R * S;
    idx = cta;
    view = (p.flags & AAR_FLAG_VIEW_OUTPUT) != 0;
  }
  // chunk-major (chunk c of every owner is reduced at about the same time), owners staggered from self
  __device__ int nextUnit(CopyUnit& u, bool blocking) {
    for (; idx < total; idx += nctas) {
      const int s = (int)(idx % S);
      const int owner = (p.rank + first + (int)((idx / S) % R)) % l.P;
      const int c = (int)(idx / ((long long)S * R));
      const int clen = chunkLen(l, owner, c);
      const int off = s * segElems;
      if (off >= clen) continue;
      const int key = owner * l.C + c;
      if (key != lastKey) {
        const int cnt = eagerChunkDecision1(p, l, key, blocking);
        if (cnt == -2) return -1;
        fence_proxy_async();  // the chunk is read by bulk loads
        lastKey = key;
        lastCnt = cnt;
      }
      const int cnt = lastCnt;
      if (s == 0 && p.outCounts) p.outCounts[key] = cnt < 0 ? 0 : cnt;
      const int n = min(segElems, clen - off);
      char* ring = regionOf(p, p.rank) + l.reduceOff + (((long long)slot * l.P + owner) * l.blkStride + (long long)c * l.maxChunk + off) * (long long)sizeof(T);
      idx += nctas;
      if (view) {  // the caller reads the ring slot itself: only chunks that never arrived are touched
        if (cnt >= 0) {
          idx -= nctas;
          continue;
        }
        u.src = nullptr;
        u.dst = ring;
      } else {
        u.src = cnt >= 0 ? ring : nullptr;  // chunks whose ReduceBlock never arrived stay zero (AllreduceWorker.scala:183)
        u.dst = reinterpret_cast<char*>(p.output) + (blockStart(l, owner) + (long long)c * l.maxChunk + off) * (long long)sizeof(T);
      }
      u.bytes = (uint32_t)n * (uint32_t)sizeof(T);
      u.a = u.b = u.c = 0;
      return 1;
    }
    return 0;
  }
  __device__ void complete(uint32_t, uint32_t, uint32_t) {}
};

// Gather from the round's snapshot (thComplete < 1 or a forced round): gatherDecision() has counted the arrivals against
// minChunksRequired and recorded which chunks made it; nothing is waited for here.
template <typename T>
struct GatherSnapshotRole {
  const AarRankParams& p;
  const AarLayout& l;
  int nctas, slot, S, segElems;
  long long total, idx;
  bool view;
  const int* counts;

  __device__ GatherSnapshotRole(const AarRankParams& p_, const AarLayout& l_, int cta, int nctas_) : p(p_), l(l_), nctas(nctas_) {
    slot = p.round % l.L;
    const SweepPlan plan(l, l.P, nctas);
    S = plan.segs;
    segElems = plan.segElems(l);
    total = (long long)l.C * l.P * S;
    idx = cta;
    view = (p.flags & AAR_FLAG_VIEW_OUTPUT) != 0;
    counts = reinterpret_cast<const int*>(regionOf(p, p.rank) + l.countsOff) + (long long)slot * l.totalChunks;
  }
  __device__ int nextUnit(CopyUnit& u, bool) {
    for (; idx < total; idx += nctas) {
      const int s = (int)(idx % S);
      const int owner = (p.rank + (int)((idx / S) % l.P)) % l.P;
      const int c = (int)(idx / ((long long)S * l.P));
      const int clen = chunkLen(l, owner, c);
      const int off = s * segElems;
      if (off >= clen) continue;
      const int cnt = __ldcg(counts + owner * l.C + c);
      if (view && cnt >= 0) continue;
      const int n = min(segElems, clen - off);
      char* ring = regionOf(p, p.rank) + l.reduceOff + (((long long)slot * l.P + owner) * l.blkStride + (long long)c * l.maxChunk + off) * (long long)sizeof(T);
      if (view) {
        u.src = nullptr;
        u.dst = ring;
      } else {
        u.src = cnt >= 0 ? ring : nullptr;
        u.dst = reinterpret_cast<char*>(p.output) + (blockStart(l, owner) + (long long)c * l.maxChunk + off) * (long long)sizeof(T);
      }
      u.bytes = (uint32_t)n * (uint32_t)sizeof(T);
      u.a = u.b = u.c = 0;
      idx += nctas;
      return 1;
    }
    return 0;
  }
  __device__ void complete(uint32_t, uint32_t, uint32_t) {}
};

template <typename T>
__device__ void roleGatherP2P(const AarRankParams& p, const AarLayout& l, int cta, int nctas, BulkShared& bs, Smem& sm, unsigned char* dsm) {
  bulkInit(bs, dsm);
  if (eagerMode(p, l)) {
    if (threadIdx.x == 0) {
      GatherEagerRole<T> role(p, l, cta, nctas);
      if (role.R > 0) runCopyRole(p, l, role, bs, dsm);
      traceMark(p, l, 4);
    }
  } else {
    gatherDecision(p, l, p.round % l.L, sm);  // every thread
    if (threadIdx.x == 0) {
      traceMark(p, l, 4);
      fence_proxy_async();  // the ring was written by peers' generic stores, acquired through the snapshot
      GatherSnapshotRole<T> role(p, l, cta, nctas);
      runCopyRole(p, l, role, bs, dsm);
    }
  }
}

// NVLS gather: piece by piece.  Exactly one CTA handles a piece, so "arrived or not" needs no shared decision; a piece
// that never arrives marks its chunk abandoned (the eager-gather decision word), and the round's last CTA zero-fills
// abandoned chunks as a whole before the slot retires (NvlsCleanup), so a chunk is never left half copied.
template <typename T>
struct GatherPieceRole {
  const AarRankParams& p;
  const AarLayout& l;
  int nctas, slot, Sp, pieceElems;
  uint32_t epoch;
  long long total, idx;
  bool view;

  __device__ GatherPieceRole(const AarRankParams& p_, const AarLayout& l_, int cta, int nctas_) : p(p_), l(l_), nctas(nctas_) {
    slot = p.round % l.L;
    epoch = (uint32_t)(p.round + 1);
    Sp = l.piecesPerChunk;
    pieceElems = l.tilesPerPiece * l.tileElems;
    total = (long long)l.C * l.P * Sp;
    idx = cta;
    view = (p.flags & AAR_FLAG_VIEW_OUTPUT) != 0;
  }
  __device__ int nextUnit(CopyUnit& u, bool blocking) {
    for (; idx < total; idx += nctas) {
      const int owner = (p.rank + (int)(idx % l.P)) % l.P;
      const int s = (int)((idx / l.P) % Sp);
      const int c = (int)(idx / ((long long)Sp * l.P));
      const int clen = chunkLen(l, owner, c);
      const int off = s * pieceElems;
      if (off >= clen) continue;
      char* my = regionOf(p, p.rank);
      const uint32_t* flag = reinterpret_cast<const uint32_t*>(my + l.pflagROff) + (((long long)slot * l.P + owner) * l.C + c) * Sp + s;
      bool arrived = ld_relaxed_sys_u32(flag) == epoch;
      if (!arrived) {
        if (!blocking) return -1;
        Deadline dl(p.timeoutNs);
        unsigned spins = 0, backoff = 32;
        for (;;) {
          if (ld_relaxed_sys_u32(flag) == epoch) {
            arrived = true;
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
      }
      fence_sys();          // acquire side of the piece flag
      fence_proxy_async();  // ... for the bulk loads that follow
      const int key = owner * l.C + c;
      if (!arrived) {
        uint64_t* dec = reinterpret_cast<uint64_t*>(my + l.egDecOff) + (long long)slot * l.P * l.C + key;
        atomicExch(reinterpret_cast<unsigned long long*>(dec), (unsigned long long)((uint64_t)epoch << 32));  // abandoned
      }
      if (s == 0 && p.outCounts) p.outCounts[key] = arrived ? l.P : 0;
      const int n = min(pieceElems, clen - off);
      char* ring = my + l.reduceOff + (((long long)slot * l.P + owner) * l.blkStride + (long long)c * l.maxChunk + off) * (long long)sizeof(T);
      idx += nctas;
      if (view) {
        if (arrived) {
          idx -= nctas;
          continue;
        }
        u.src = nullptr;
        u.dst = ring;
      } else {
        u.src = arrived ? ring : nullptr;
        u.dst = reinterpret_cast<char*>(p.output) + (blockStart(l, owner) + (long long)c * l.maxChunk + off) * (long long)sizeof(T);
      }
      u.bytes = (uint32_t)n * (uint32_t)sizeof(T);
      u.a = u.b = u.c = 0;
      return 1;
    }
    return 0;
  }
  __device__ void complete(uint32_t, uint32_t, uint32_t) {}
};

template <typename T>
__device__ void roleGatherNvls(const AarRankParams& p, const AarLayout& l, int cta, int nctas, BulkShared& bs, unsigned char* dsm) {
  bulkInit(bs, dsm);
  if (threadIdx.x != 0) return;
  GatherPieceRole<T> role(p, l, cta, nctas);
  runCopyRole(p, l, role, bs, dsm);
  traceMark(p, l, 4);
}

// Runs on thread 0 of the round's last CTA, after every other CTA has finished: chunks that were abandoned piecewise are
// zero-filled as a whole and reported with count 0.
template <typename T>
struct NvlsCleanup {
  unsigned char* dsm;
  __device__ void operator()(const AarRankParams& p, const AarLayout& l) const {
    char* my = regionOf(p, p.rank);
    const int slot = p.round % l.L;
    const uint64_t abandoned = (uint64_t)(p.round + 1) << 32;
    const uint64_t* dec = reinterpret_cast<const uint64_t*>(my + l.egDecOff) + (long long)slot * l.P * l.C;
    const bool view = (p.flags & AAR_FLAG_VIEW_OUTPUT) != 0;
    const uint32_t zero = smem_u32(dsm + kStageBytes);
    bool any = false;
    for (int owner = 0; owner < l.P; ++owner) {
      for (int c = 0; c < numChunks(l, owner); ++c) {
        const int key = owner * l.C + c;
        if (ld_acquire_gpu_u64(dec + key) != abandoned) continue;
        any = true;
        if (p.outCounts) p.outCounts[key] = 0;
        const long long bytes = (long long)chunkLen(l, owner, c) * (long long)sizeof(T);
        char* dst = view ? my + l.reduceOff + (((long long)slot * l.P + owner) * l.blkStride + (long long)c * l.maxChunk) * (long long)sizeof(T)
                         : reinterpret_cast<char*>(p.output) + (blockStart(l, owner) + (long long)c * l.maxChunk) * (long long)sizeof(T);
        for (long long o = 0; o < bytes; o += kZeroTile) {
          bulk_s2g(dst + o, zero, (uint32_t)(bytes - o < kZeroTile ? bytes - o : kZeroTile));
          bulk_commit();
          bulk_wait_read<8>();
        }
      }
    }
    if (any) {
      bulk_wait<0>();
      fence_proxy_async();
    }
  }
};

// ------------------------------------------------------------------------------------------------ kernels
// phases == 7: one launch per round, CTAs specialised by role (split = nA | nB << 16):
//   CTAs [0, nA)          K1 scatter_push (NVLS: stage)
//   CTAs [nA, nA + nB)    K2 reduce_bcast (NVLS: in-switch reduce + multicast store)
//   the rest              K3 gather
// phases in {1, 2, 4}: the whole grid runs one phase (single-device emulation launches the phases of all ranks in order).
template <typename T, bool kNvls>
__global__ void __launch_bounds__(kThreads, 1) bulkRoundKernel(const AarRankParams p, const AarLayout l, int split, int phases) {
  extern __shared__ __align__(1024) unsigned char dsm[];
  __shared__ BulkShared bs;
  __shared__ Smem sm;
  initSmem(sm);
  traceMark(p, l, 0);
  const int nctas = gridDim.x, b = blockIdx.x;
  int role = 0, cta = b, n = nctas;  // role bit mask
  if (phases == 7) {
    const int nA = split & 0xffff, nB = split >> 16;
    if (b < nA) {
      role = 1;
      n = nA;
    } else if (b < nA + nB) {
      role = 2;
      cta = b - nA;
      n = nB;
    } else {
      role = 4;
      cta = b - nA - nB;
      n = nctas - nA - nB;
    }
  } else {
    role = phases;
  }
  if (role & 1) {
    if (kNvls) roleStage<T>(p, l, cta, n, bs, dsm);
    else rolePush<T>(p, l, cta, n, bs, dsm);
    if (threadIdx.x == 0) traceMark(p, l, 1);
  }
  if (role & 2) {
    if (kNvls) roleReduceNvls<T>(p, l, cta, n, bs);
    else roleReduceP2P<T>(p, l, cta, n, bs, dsm);
  }
  if (role & 4) {
    if (kNvls) roleGatherNvls<T>(p, l, cta, n, bs, dsm);
    else roleGatherP2P<T>(p, l, cta, n, bs, sm, dsm);
  }
  if (phases & 4) {
    if (kNvls) retireRoundWith(p, l, nctas, NvlsCleanup<T>{dsm});
    else retireRound(p, l, nctas);
  }
  traceMark(p, l, 5);
}

// ------------------------------------------------------------------------------------------------ host side
struct KernelInfo {
  bool ready = false;
  int maxGrid = 0;
};
std::mutex g_mutex;
KernelInfo g_info[64][4];  // [device][kernel]

template <typename K>
int prepare(K kernel, int which, int* maxGrid) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return aar_host::fail(AAR_ECUDA, "aar_launch_round_bulk: cudaGetDevice failed");
  std::lock_guard<std::mutex> lock(g_mutex);
  KernelInfo& ki = g_info[dev][which];
  if (!ki.ready) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBulkDynSmem);
    if (e != cudaSuccess) return aar_host::fail(AAR_ECUDA, "aar_launch_round_bulk: shared memory opt-in (%d bytes): %s", kBulkDynSmem, cudaGetErrorString(e));
    int sms = 0, perSm = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kernel, kThreads, kBulkDynSmem);
    if (e != cudaSuccess || perSm < 1 || sms < 1)
      return aar_host::fail(AAR_ECUDA, "aar_launch_round_bulk: occupancy query failed: %s", cudaGetErrorString(e));
    ki.maxGrid = sms * perSm;
    ki.ready = true;
  }
  *maxGrid = ki.maxGrid;
  return AAR_OK;
}

int supported(const AarRankParams* p, const AarLayout* lay, int algo, const char** why) {
  const int V = 16 / lay->elemSize;
  *why = nullptr;
  if (lay->stepSize % V || lay->maxChunk % V || lay->dataSize % V || lay->blkStride % V) *why = "blocks and chunks must be multiples of 16 bytes";
  else if (reinterpret_cast<uintptr_t>(p->input) % 16) *why = "the input must be 16-byte aligned";
  else if (!(p->flags & AAR_FLAG_VIEW_OUTPUT) && reinterpret_cast<uintptr_t>(p->output) % 16) *why = "the output must be 16-byte aligned";
  else if (lay->tileElems * lay->elemSize != kCopyTile) *why = "layout tile size differs from the copy tile";
  else if (algo == 1) {
    if (!p->mcRegion || lay->stageOff == 0) *why = "the in-switch path needs a multicast mapping and the staging area";
    else if (p->minRequired != lay->P || p->minChunksRequired != lay->totalChunks || (p->flags & (AAR_FLAG_SKIP_SCATTER | AAR_FLAG_EXACT_QUORUM | AAR_FLAG_NORMALIZE)))
      *why = "the switch sums every bound GPU, so thresholds must be 1.0 (use the P2P path otherwise)";
  }
  return *why == nullptr;
}

}  // namespace

AAR_API int aar_bulk_supported(const AarRankParams* p, const AarLayout* lay, int algo) {
  if (!p || !lay) return 0;
  const char* why = nullptr;
  return supported(p, lay, algo, &why);
}

AAR_API int aar_launch_round_bulk(const AarRankParams* p, const AarLayout* lay, int algo, int phases, int grid, int split, void* stream) {
  const char* what = "aar_launch_round_bulk";
  if (!p || !lay) return aar_host::fail(AAR_EINVAL, "%s: null params", what);
  if (p->rank < 0 || p->rank >= lay->P || lay->P > AAR_MAXP) return aar_host::fail(AAR_EINVAL, "%s: bad rank %d for P=%d", what, p->rank, lay->P);
  if (algo != 0 && algo != 1) return aar_host::fail(AAR_EINVAL, "%s: algo must be 0 (P2P) or 1 (NVLS)", what);
  if (phases != 7 && phases != 1 && phases != 2 && phases != 4) return aar_host::fail(AAR_EINVAL, "%s: phases must be 7, 1, 2 or 4", what);
  if (lay->dtype != AAR_F32 && lay->dtype != AAR_BF16) return aar_host::fail(AAR_EINVAL, "%s: unknown dtype %d", what, lay->dtype);
  const char* why = nullptr;
  if (!supported(p, lay, algo, &why)) return aar_host::fail(AAR_EUNSUPPORTED, "%s: %s", what, why);
  const int which = (algo ? 2 : 0) + (lay->dtype == AAR_BF16 ? 1 : 0);
  int maxGrid = 0, rc = 0;
  switch (which) {
    case 0: rc = prepare(bulkRoundKernel<float, false>, which, &maxGrid); break;
    case 1: rc = prepare(bulkRoundKernel<__nv_bfloat16, false>, which, &maxGrid); break;
    case 2: rc = prepare(bulkRoundKernel<float, true>, which, &maxGrid); break;
    default: rc = prepare(bulkRoundKernel<__nv_bfloat16, true>, which, &maxGrid); break;
  }
  if (rc) return rc;
  // CTAs of one grid wait on each other: the grid must be co-resident
  int g = grid > 0 ? grid : maxGrid;
  if (g > maxGrid) g = maxGrid;
  int sp = 0;
  if (phases == 7) {
    if (g < 3) return aar_host::fail(AAR_EINVAL, "%s: the role-specialised launch needs at least 3 CTAs", what);
    int nA = split & 0xffff, nB = split >> 16;
    if (nA <= 0 || nB <= 0 || nA + nB >= g) {
      if (algo == 1) {
        nB = g / 8 > 0 ? g / 8 : 1;   // few CTAs saturate the switch
        nA = (g - nB) * 2 / 5;
      } else {
        nA = g / 5;                   // remote copies only: the link, not the SMs, bounds this role
        nB = g * 2 / 5;
      }
    }
    if (nA < 1) nA = 1;
    if (nB < 1) nB = 1;
    if (nA + nB >= g) return aar_host::fail(AAR_EINVAL, "%s: split %d + %d leaves no gather CTA in a grid of %d", what, nA, nB, g);
    sp = nA | (nB << 16);
  }
  cudaStream_t s = (cudaStream_t)stream;
  switch (which) {
    case 0: bulkRoundKernel<float, false><<<g, kThreads, kBulkDynSmem, s>>>(*p, *lay, sp, phases); break;
    case 1: bulkRoundKernel<__nv_bfloat16, false><<<g, kThreads, kBulkDynSmem, s>>>(*p, *lay, sp, phases); break;
    case 2: bulkRoundKernel<float, true><<<g, kThreads, kBulkDynSmem, s>>>(*p, *lay, sp, phases); break;
    default: bulkRoundKernel<__nv_bfloat16, true><<<g, kThreads, kBulkDynSmem, s>>>(*p, *lay, sp, phases); break;
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return aar_host::fail(AAR_ECUDA, "%s: launch failed: %s", what, cudaGetErrorString(e));
  return AAR_OK;
}
