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
//   * arrival is published per PIECE (a fixed sub-chunk granule of the layout, <= 256 KiB) instead of per chunk: exactly
//     one CTA moves a piece, so its flag costs one fence of the CTA's publishing thread and one relaxed store -- no
//     election, no second fence; the contributor decision is still taken per chunk (a source has delivered a chunk when
//     all its pieces are flagged), the gather starts on a reduced chunk's first piece, and the in-switch (NVLS) form,
//     which has no decision to take, pipelines piece by piece throughout;
//   * always strict: the broadcast obeys the slot flow control, a chunk that is given up is zero-filled as a whole.
// Every rank of a job must run the bulk kernels for a given geometry (they do not read or write the per-chunk flags of
// aar_kernels.cu); RoundEngine selects them from the layout, which is the same on all ranks.
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

// ------------------------------------------------------------------------------------------------ K1 on the copy engine
template <typename T>
struct PushRole {
  static constexpr bool kPublishes = true;
  const AarRankParams& p;
  const AarLayout& l;
  int nctas, slot, R, first, Sp, pieceElems;
  uint32_t epoch;
  int total, idx;
  bool rotate;
  unsigned known, go;

  __device__ PushRole(const AarRankParams& p_, const AarLayout& l_, int cta, int nctas_) : p(p_), l(l_), nctas(nctas_), known(0), go(0) {
    slot = p.round % l.L;
    epoch = (uint32_t)(p.round + 1);
    first = selfDirect(p) ? 1 : 0;
    R = l.P - first;
    Sp = l.piecesPerChunk;
    pieceElems = l.tilesPerPiece * l.tileElems;
    total = l.C * R * Sp;
    idx = cta;
    rotate = R > 0 && (nctas % R) == 0;
  }
  // chunk-major, then piece, then destination staggered from self (AllreduceWorker.scala:195-196): the role's CTAs sweep
  // the block chunk by chunk together, so chunk c is complete after (c + 1) / C of the role's time; when the CTA count is a
  // multiple of the destination count each pass rotates by one so a CTA does not always draw the same peer
  __device__ int nextUnit(CopyUnit& u, bool blocking) {
    for (; idx < total; idx += nctas) {
      const int i = (idx % R + (rotate ? idx / nctas : 0)) % R;
      const int s = (idx / R) % Sp;
      const int c = idx / (Sp * R);
      const int dest = (p.rank + first + i) % l.P;
      const int clen = chunkLen(l, dest, c);
      const int off = s * pieceElems;
      if (off >= clen) continue;
      const int d = destDecision1(p, l, dest, known, go, blocking);
      if (d < 0) return -1;
      if (d == 0) continue;  // the peer fell out of the window: its copy of this contribution is dropped
      const int n = min(pieceElems, clen - off);
      const long long inOff = blockStart(l, dest) + (long long)c * l.maxChunk + off;
      u.src = reinterpret_cast<const char*>(reinterpret_cast<const T*>(p.input) + inOff);
      u.dst = regionOf(p, dest) + l.scatterOff + (((long long)slot * l.P + p.rank) * l.blkStride + (long long)c * l.maxChunk + off) * (long long)sizeof(T);
      u.bytes = (uint32_t)n * (uint32_t)sizeof(T);
      u.a = (uint32_t)dest;
      u.b = (uint32_t)c;
      u.c = (uint32_t)s;
      idx += nctas;
      return 1;
    }
    return 0;
  }
  // this piece of the "ScatterBlock" has landed (runs on the publishing thread, after its fence)
  __device__ void complete(uint32_t dest, uint32_t c, uint32_t s) {
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(pflagS(p, l, (int)dest, slot, (int)c, (int)s) + p.rank), "r"(epoch) : "memory");
  }
};

template <typename T>
__device__ void rolePush(const AarRankParams& p, const AarLayout& l, int cta, int nctas, BulkShared& bs, unsigned char* dsm) {
  bulkInit(bs, dsm);
  if (p.flags & AAR_FLAG_SKIP_SCATTER) return;
  if (threadIdx.x == 32) {  // publishing thread: arrival flags of the pieces thread 0 has finished
    PushRole<T> role(p, l, cta, nctas);
    doneDrain(bs.done, role);
    return;
  }
  if (threadIdx.x != 0) return;
  if (selfDirect(p) && cta == 0) {
    // my own contribution stays where it is (the reduce role reads it from the input): only its arrival is announced
    const int slot = p.round % l.L;
    const int mine = numChunks(l, p.rank);
    for (int c = 0; c < mine; ++c) {
      const int np = piecesOf(l, chunkLen(l, p.rank, c));
      for (int s = 0; s < np; ++s)  // relaxed: the input was complete before this kernel started, there is nothing to order
        asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(pflagS(p, l, p.rank, slot, c, s) + p.rank), "r"((uint32_t)(p.round + 1)) : "memory");
    }
  }
  PushRole<T> role(p, l, cta, nctas);
  if (role.R > 0) runCopyRole(p, l, role, bs, dsm);
  else doneClose(bs.done);
}

// NVLS: the input is staged in MY region (the switch pulls it); only the flag travels, one per piece.
template <typename T>
struct StageRole {
  static constexpr bool kPublishes = true;
  const AarRankParams& p;
  const AarLayout& l;
  int nctas, slot, Sp, pieceElems;
  uint32_t epoch;
  int total, idx;
  unsigned known, go;
  char* stage;
  bool inPlace;

  __device__ StageRole(const AarRankParams& p_, const AarLayout& l_, int cta, int nctas_) : p(p_), l(l_), nctas(nctas_), known(0), go(0) {
    slot = p.round % l.L;
    epoch = (uint32_t)(p.round + 1);
    Sp = l.piecesPerChunk;
    pieceElems = l.tilesPerPiece * l.tileElems;
    total = l.C * l.P * Sp;
    idx = cta;
    stage = regionOf(p, p.rank) + l.stageOff + (long long)slot * paddedData(l) * (long long)sizeof(T);
    inPlace = reinterpret_cast<const char*>(p.input) == stage;  // the source wrote straight into the stage
  }
  __device__ int nextUnit(CopyUnit& u, bool blocking) {
    for (; idx < total; idx += nctas) {
      const int b = (idx % l.P + p.rank) % l.P;
      const int s = (idx / l.P) % Sp;
      const int c = idx / (Sp * l.P);
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
  __device__ void complete(uint32_t b, uint32_t c, uint32_t s) {  // on the publishing thread, after its fence
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(pflagS(p, l, (int)b, slot, (int)c, (int)s) + p.rank), "r"(epoch) : "memory");
  }
};

template <typename T>
__device__ void roleStage(const AarRankParams& p, const AarLayout& l, int cta, int nctas, BulkShared& bs, unsigned char* dsm) {
  bulkInit(bs, dsm);
  if (p.flags & AAR_FLAG_SKIP_SCATTER) return;
  if (threadIdx.x != 0 && threadIdx.x != 32) return;
  StageRole<T> role(p, l, cta, nctas);
  if (threadIdx.x == 32) doneDrain(bs.done, role);
  else runCopyRole(p, l, role, bs, dsm);
}

// ------------------------------------------------------------------------------------------------ K2 on the copy engine
// Ordered fp32 sum of the decided contributor set (peer-index order from 0.0f, AllreduceWorker.scala:245-249), tile by
// tile, with the CTA's warps specialised:
//   warp 0 (one thread)   loader: takes the chunk's contributor decision and the destinations' flow-control decisions, then
//                         bulk-loads the tile of every contributor into the stage (own contribution straight from the input
//                         with AAR_FLAG_SELF_DIRECT); runs ahead of the adds by the number of stages, across segments and chunks
//   warps 3..15           adders: two (or one) 16-byte vectors per thread, shared memory in, shared memory out
//   warp 1 (one thread)   storer: bulk-stores the result tile into every destination's reduce ring (own block straight to the
//                         output when every chunk is required), one bulk group per tile
//   warp 2 (one thread)   publisher: the ReduceBlock flags (epoch | count) of the pieces whose stores have completed -- the
//                         only thread that fences, and it has no copies of its own in flight
// Stages hand over through mbarriers: full (loads landed) -> computed (result written, inputs consumed) -> outFree (stores
// have read the result).  Issuing one bulk copy costs the issuing thread about 300 cycles (measured), which is why loads,
// stores and adds each get their own thread(s) instead of taking turns on one.
constexpr int kAdderWarp0 = 3, kAdderWarps = kThreads / 32 - kAdderWarp0, kAdders = kAdderWarps * 32;  // 13 warps, 416 threads

struct ReducePublisher {
  const AarRankParams& p;
  const AarLayout& l;
  int slot;
  bool so;
  // this piece of the "ReduceBlock" has landed everywhere it was sent: tell those peers, count included (runs after the
  // publishing thread's fence)
  __device__ void complete(uint32_t c, uint32_t destCount, uint32_t s) {
    const int P = l.P;
    const unsigned dm = destCount & 0xffffu, count = destCount >> 16;
    if (so && s == 0 && p.outCounts) p.outCounts[p.rank * l.C + (int)c] = (int)count;
    const uint64_t word = ((uint64_t)(p.round + 1) << 32) | (uint64_t)count;
    for (int i2 = 0; i2 < P; ++i2) {
      const int dest = (i2 + p.rank) % P;
      if (!((dm >> dest) & 1u)) continue;  // no data was written there: no flag either
      asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(pflagR(p, l, dest, slot, p.rank, (int)c, (int)s)), "l"(word) : "memory");
    }
  }
};

template <typename T>
__device__ void roleReduceP2P(const AarRankParams& p, const AarLayout& l, int cta, int nctas, BulkShared& bs, unsigned char* dsm) {
  bulkInit(bs, dsm, kAdderWarps);
  constexpr int V = Elem<T>::kPerVec;
  const int P = l.P;
  const int myChunks = numChunks(l, p.rank);
  if (myChunks == 0) return;
  const int slot = p.round % l.L;
  char* my = regionOf(p, p.rank);
  const bool sd = selfDirect(p), so = selfOut(p, l);
  const int tileB = kAdders * 16 * (P <= 4 ? 2 : 1) / (P <= 8 ? 1 : 2);  // 13312 / 6656 / 3328 bytes per contributor per tile
  const int stageB = (P + 1) * tileB;
  int nst = kStageBytes / stageB;
  if (nst > kMaxStages) nst = kMaxStages;
  const int S = l.piecesPerChunk, segElems = l.tilesPerPiece * l.tileElems;  // units are pieces, swept chunk by chunk
  const int total = myChunks * S;
  const uint32_t stage0 = smem_u32(dsm);
  const long long blockIn = blockStart(l, p.rank);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr uint32_t kStop = 0xffffffffu;

  if (warp == 0) {
    if (lane != 0) return;
    // ------------------------------------------------------------ loader
    unsigned known = 0, go = 0, mask = 0;
    int cachedChunk = -1, st = 0;
    uint32_t use = 0;  // how often stage `st` has been used before
    auto arm = [&](uint32_t bytes, unsigned contributors, unsigned dm, long long byteOff, uint32_t a, uint32_t b, uint32_t c, bool last) -> uint32_t {
      // the stage's previous tenant must be gone altogether -- inputs added, result stored, its job words read by the storer:
      // phase (use - 1) of `outFree` (which the storer signals one tile after the stores were issued)
      mbar_wait(p, l, smem_u32(&bs.outFree[st]), (use + 1u) & 1u);
      bs.jobBytes[st] = bytes;
      bs.jobInfo[st] = contributors;
      bs.jobDst[st] = (unsigned long long)byteOff;
      bs.unitA[st] = a;
      bs.unitB[st] = dm | (b << 16);
      bs.unitC[st] = c | (last ? 0x80000000u : 0u);
      return smem_u32(&bs.full[st]);
    };
    auto advance = [&]() {
      if (++st == nst) {
        st = 0;
        ++use;
      }
    };
    for (int idx = cta; idx < total; idx += nctas) {
      const int c = idx / S, s = idx - c * S;
      const int clen = chunkLen(l, p.rank, c);
      const int rest = clen - s * segElems;
      if (rest <= 0) continue;
      const int len = rest < segElems ? rest : segElems;
      if (c != cachedChunk) {
        chunkDecision1(p, l, c, clen, p.minRequired, true, mask);
        fence_proxy_async();  // the flagged rows are read by bulk loads
        if (cachedChunk < 0) traceMark(p, l, 2);
        cachedChunk = c;
      }
      // Slot flow control for the broadcast: round R may overwrite slot R % L of a peer's reduce ring only after that peer
      // retired round R - L; a peer that fell out of the window is skipped (no data, no flag).
      unsigned dm = 0;
      for (int i2 = 0; i2 < P; ++i2) {
        const int dest = (i2 + p.rank) % P;
        if (destDecision1(p, l, dest, known, go, true)) dm |= 1u << dest;
      }
      const int nsrc = __popc(mask);
      const long long unitOff = ((long long)c * l.maxChunk + (long long)s * segElems) * (long long)sizeof(T);
      const long long lenB = (long long)len * (long long)sizeof(T);
      for (long long o = 0; o < lenB; o += tileB) {
        const uint32_t bytes = (uint32_t)(lenB - o < tileB ? lenB - o : tileB);
        const uint32_t bar = arm(bytes, mask, dm, unitOff + o, (uint32_t)c, (uint32_t)nsrc, (uint32_t)s, o + tileB >= lenB);
        if (nsrc == 0) {
          mbar_arrive(bar);  // nothing to load: the phase completes at once
        } else {
          mbar_expect_tx(bar, bytes * (uint32_t)nsrc);
          int j = 0;
          for (unsigned m = mask; m; m &= m - 1, ++j) {
            const int src = __ffs(m) - 1;
            const char* row = (src == p.rank && sd) ? reinterpret_cast<const char*>(p.input) + blockIn * (long long)sizeof(T)
                                                    : my + l.scatterOff + ((long long)slot * P + src) * l.blkStride * (long long)sizeof(T);
            bulk_g2s(stage0 + (uint32_t)(st * stageB + j * tileB), row + unitOff + o, bytes, bar);
          }
        }
        advance();
      }
    }
    mbar_arrive(arm(kStop, 0, 0, 0, 0, 0, 0, false));  // no more tiles
    traceMarkAny(p, l, 6);
    return;
  }

  if (warp == 1) {
    if (lane != 0) return;
    // ------------------------------------------------------------ storer
    int st = 0, prev = -1;
    uint32_t use = 0, G = 0, head = 0, tail = 0;
    auto retireDone = [&](uint32_t lastCompleteGroup) {
      while (head != tail && bs.fifoGroup[head & 15] <= lastCompleteGroup) {
        donePush(bs.done, bs.fifoA[head & 15], bs.fifoB[head & 15], bs.fifoC[head & 15]);
        ++head;
      }
    };
    for (;;) {
      mbar_wait(p, l, smem_u32(&bs.computed[st]), use & 1u);
      const uint32_t bytes = bs.jobBytes[st];
      if (bytes == kStop) break;
      const unsigned dm = bs.unitB[st] & 0xffffu;
      const long long byteOff = (long long)bs.jobDst[st];
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
      if (G == 0) traceMarkAny(p, l, 7);
      if (bs.unitC[st] & 0x80000000u) {  // last tile of its piece
        bs.fifoGroup[tail & 15] = G;
        bs.fifoA[tail & 15] = bs.unitA[st];
        bs.fifoB[tail & 15] = bs.unitB[st];
        bs.fifoC[tail & 15] = bs.unitC[st] & 0x7fffffffu;
        ++tail;
      }
      bulk_wait_read<1>();  // the stores of the tile before this one have read their result tile: its stage may be rewritten
      if (prev >= 0) mbar_arrive(smem_u32(&bs.outFree[prev]));
      prev = st;
      if (head != tail && G + 1 > (uint32_t)kDoneLag && bs.fifoGroup[head & 15] + kDoneLag < G + 1) {
        bulk_wait<kDoneLag>();
        retireDone(G - kDoneLag);
      }
      ++G;
      if (++st == nst) {
        st = 0;
        ++use;
      }
    }
    bulk_wait<0>();
    if (G > 0) retireDone(G - 1);
    doneClose(bs.done);
    traceMarkAny(p, l, 3);
    return;
  }

  if (warp == 2) {
    if (lane != 0) return;
    // ------------------------------------------------------------ publisher
    ReducePublisher pub{p, l, slot, so};
    doneDrain(bs.done, pub);
    return;
  }

  // ---------------------------------------------------------------- adders
  const int t = threadIdx.x - kAdderWarp0 * 32;
  int st = 0;
  uint32_t use = 0;
  for (;;) {
    mbar_wait(p, l, smem_u32(&bs.full[st]), use & 1u);
    const uint32_t bytes = bs.jobBytes[st];
    if (bytes == kStop) {
      if (lane == 0) mbar_arrive(smem_u32(&bs.computed[st]));
      break;
    }
    const int count = __popc(bs.jobInfo[st]);
    const float divisor = ((p.flags & AAR_FLAG_NORMALIZE) && count > 1) ? (float)count : 1.f;
    // the result tile of this stage is free once the stores of its previous tenant have read it
    mbar_wait(p, l, smem_u32(&bs.outFree[st]), (use + 1u) & 1u);
    const unsigned char* in = dsm + (size_t)st * stageB;
    unsigned char* out = dsm + (size_t)st * stageB + (size_t)P * tileB;
    for (uint32_t i = (uint32_t)t; i < bytes / 16; i += kAdders) {
      float acc[V];
#pragma unroll
      for (int k = 0; k < V; ++k) acc[k] = 0.f;
      const unsigned char* src = in + (size_t)i * 16;
      for (int j = 0; j < count; ++j, src += tileB) {  // contributors sit in the stage in peer-index order
        const uint4 v = *reinterpret_cast<const uint4*>(src);
        float f[V];
        Elem<T>::unpack(v, f);
#pragma unroll
        for (int k = 0; k < V; ++k) acc[k] += f[k];
      }
      if (divisor != 1.f) {
#pragma unroll
        for (int k = 0; k < V; ++k) acc[k] = acc[k] / divisor;
      }
      *reinterpret_cast<uint4*>(out + (size_t)i * 16) = Elem<T>::pack(acc);
    }
    fence_proxy_async_smem();  // the result tile is read by the storer's bulk stores
    __syncwarp();
    if (lane == 0) mbar_arrive(smem_u32(&bs.computed[st]));
    if (++st == nst) {
      st = 0;
      ++use;
    }
  }
}

// NVLS (K4): the NVSwitch sums the staged inputs of all ranks (multimem.ld_reduce) and replicates the result into every
// rank's reduce ring (multimem.st); per direction each GPU moves about N(1 + 1/P) bytes instead of 2N(P-1)/P.
template <typename T>
__device__ __forceinline__ void nvlsSpanBulk(const char* mcStage, char* mcOut, int nv, float divisor) {
  constexpr int U = 8;
  constexpr int V = Elem<T>::kPerVec;
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
    if (divisor != 1.f) {  // AAR_FLAG_NORMALIZE: the switch sums all P ranks, so the contributor count is P
#pragma unroll
      for (int u = 0; u < U; ++u) {
        float f[V];
        Elem<T>::unpack(v[u], f);
#pragma unroll
        for (int k = 0; k < V; ++k) f[k] = f[k] / divisor;
        v[u] = Elem<T>::pack(f);
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
  const int P = l.P, Sp = l.piecesPerChunk;
  const int myChunks = numChunks(l, p.rank);
  if (myChunks == 0) return;
  const int slot = p.round % l.L;
  const uint32_t epoch = (uint32_t)(p.round + 1);
  char* my = regionOf(p, p.rank);
  const int pieceElems = l.tilesPerPiece * l.tileElems;
  const int total = myChunks * Sp;
  bool first = true;
  for (int idx = cta; idx < total; idx += nctas) {
    const int c = idx / Sp, s = idx - c * Sp;
    const int clen = chunkLen(l, p.rank, c);
    const int off = s * pieceElems;
    if (off >= clen) continue;
    const int n = min(pieceElems, clen - off);
    if (threadIdx.x == 0) {
      const uint32_t* flags = pflagS(p, l, p.rank, slot, c, s);
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
      nvlsSpanBulk<T>(mcStage, mcOut, n / Elem<T>::kPerVec, ((p.flags & AAR_FLAG_NORMALIZE) && P > 1) ? (float)P : 1.f);
    }
    __syncthreads();
    if (threadIdx.x == 0 && ok) {
      __threadfence_system();
      const uint64_t word = ((uint64_t)epoch << 32) | (uint64_t)P;
      for (int i2 = 0; i2 < P; ++i2) {
        const int dest = (i2 + p.rank) % P;
        asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(pflagR(p, l, dest, slot, p.rank, c, s)), "l"(word) : "memory");
      }
    }
  }
  if (threadIdx.x == 0) traceMark(p, l, 3);
}

// ------------------------------------------------------------------------------------------------ K3 on the copy engine
// Eager gather (every chunk required): piece (owner, c, s) is copied to the output as soon as its flag lands, while the
// rest of the chunk is still being reduced.  Exactly one CTA handles a piece, so "arrived or not" needs no shared decision
// -- except for giving up: a piece that never shows up (forced round, timeout) marks its CHUNK abandoned in the eager
// decision word, later pieces of that chunk are not waited for, and the round's last CTA zero-fills abandoned chunks as a
// whole before the slot retires (ChunkCleanup), so a chunk is never flushed half copied (AllreduceWorker.scala:183: a
// ReduceBlock that never arrived stays zero).
template <typename T>
struct GatherPieceRole {
  static constexpr bool kPublishes = false;
  const AarRankParams& p;
  const AarLayout& l;
  int nctas, slot, R, first, Sp, pieceElems;
  uint64_t epoch;
  int total, idx;
  bool view;

  __device__ GatherPieceRole(const AarRankParams& p_, const AarLayout& l_, int cta, int nctas_, bool skipOwn) : p(p_), l(l_), nctas(nctas_) {
    slot = p.round % l.L;
    epoch = (uint64_t)(p.round + 1);
    first = skipOwn ? 1 : 0;  // my own block was written by my reduce role
    R = l.P - first;
    Sp = l.piecesPerChunk;
    pieceElems = l.tilesPerPiece * l.tileElems;
    total = l.C * R * Sp;
    idx = cta;
    view = (p.flags & AAR_FLAG_VIEW_OUTPUT) != 0;
  }
  // chunk-major (chunk c of every owner is reduced at about the same time), then piece, owners staggered from self
  __device__ int nextUnit(CopyUnit& u, bool blocking) {
    for (; idx < total; idx += nctas) {
      const int owner = (p.rank + first + idx % R) % l.P;
      const int s = (idx / R) % Sp;
      const int c = idx / (Sp * R);
      const int clen = chunkLen(l, owner, c);
      const int off = s * pieceElems;
      if (off >= clen) continue;
      char* my = regionOf(p, p.rank);
      const int key = owner * l.C + c;
      const uint64_t* flag = pflagR(p, l, p.rank, slot, owner, c, s);
      uint64_t* dec = reinterpret_cast<uint64_t*>(my + l.egDecOff) + (long long)slot * l.P * l.C + key;
      const uint64_t abandoned = epoch << 32;
      uint64_t f = ld_relaxed_sys_u64(flag);
      bool arrived = (f >> 32) == epoch;
      if (!arrived) {
        if (!blocking) return -1;
        Deadline dl(p.timeoutNs);
        unsigned spins = 0, backoff = 32;
        for (;;) {
          f = ld_relaxed_sys_u64(flag);
          if ((f >> 32) == epoch) {
            arrived = true;
            break;
          }
          if ((spins & 15u) == 0 && ld_acquire_gpu_u64(dec) == abandoned) break;  // somebody already gave this chunk up
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
      if (arrived) {
        (void)ld_acquire_sys_u64(flag);  // acquire (a fence would wait for the copies this thread has in flight)
        fence_proxy_async();             // ... for the bulk loads that follow
      } else {
        atomicExch(reinterpret_cast<unsigned long long*>(dec), (unsigned long long)abandoned);
      }
      if (s == 0 && p.outCounts) p.outCounts[key] = arrived ? (int)(f & 0xffffffffu) : 0;
      const int n = min(pieceElems, clen - off);
      char* ring = my + l.reduceOff + (((long long)slot * l.P + owner) * l.blkStride + (long long)c * l.maxChunk + off) * (long long)sizeof(T);
      if (view) {  // the caller reads the ring slot itself: only pieces that never arrived are touched
        if (arrived) continue;
        u.src = nullptr;
        u.dst = ring;
      } else {
        u.src = arrived ? ring : nullptr;
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

// Round decision for thComplete < 1 or a forced round (see gatherDecision in aar_protocol.cuh, which reads per-chunk flags):
// one CTA becomes the leader, counts the chunks whose EVERY piece has landed against minChunksRequired
// (DataBuffer.scala:69-75 with the true chunk total), snapshots which chunks made it with their contributor counts, and
// publishes the snapshot; every CTA then gathers from that one snapshot.
static __device__ void gatherDecisionPieces(const AarRankParams& p, const AarLayout& l, int slot, Smem& sm) {
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
  auto chunkCount = [&](int owner, int c) {  // contributor count of a chunk whose every piece has landed, else -1
    const int np = piecesOf(l, chunkLen(l, owner, c));
    int cnt = -1;
    for (int s = 0; s < np; ++s) {
      const uint64_t f = ld_relaxed_sys_u64(pflagR(p, l, p.rank, slot, owner, c, s));
      if ((f >> 32) != epoch) return -1;
      cnt = (int)(f & 0xffffffffu);
    }
    return cnt;
  };
  if (sm.k3Leader) {
    Deadline dl(p.timeoutNs);
    unsigned spins = 0, backoff = 32;
    for (;;) {
      if (threadIdx.x == 0) sm.k3Count = 0;
      __syncthreads();
      int local = 0;
      for (int i = threadIdx.x; i < P * C; i += kThreads) {
        const int owner = i / C, c = i % C;
        if (c < numChunks(l, owner) && chunkCount(owner, c) >= 0) ++local;
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
    __threadfence_system();  // acquire side of the piece flags
    for (int i = threadIdx.x; i < P * C; i += kThreads) {
      const int owner = i / C, c = i % C;
      if (c >= numChunks(l, owner)) continue;
      const int cnt = chunkCount(owner, c);
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

// Gather from the round's snapshot: nothing is waited for here.
template <typename T>
struct GatherSnapshotRole {
  static constexpr bool kPublishes = false;
  const AarRankParams& p;
  const AarLayout& l;
  int nctas, slot, Sp, pieceElems;
  int total, idx;
  bool view;
  const int* counts;

  __device__ GatherSnapshotRole(const AarRankParams& p_, const AarLayout& l_, int cta, int nctas_) : p(p_), l(l_), nctas(nctas_) {
    slot = p.round % l.L;
    Sp = l.piecesPerChunk;
    pieceElems = l.tilesPerPiece * l.tileElems;
    total = l.C * l.P * Sp;
    idx = cta;
    view = (p.flags & AAR_FLAG_VIEW_OUTPUT) != 0;
    counts = reinterpret_cast<const int*>(regionOf(p, p.rank) + l.countsOff) + (long long)slot * l.totalChunks;
  }
  __device__ int nextUnit(CopyUnit& u, bool) {
    for (; idx < total; idx += nctas) {
      const int s = idx % Sp;
      const int owner = (p.rank + (idx / Sp) % l.P) % l.P;
      const int c = idx / (Sp * l.P);
      const int clen = chunkLen(l, owner, c);
      const int off = s * pieceElems;
      if (off >= clen) continue;
      const int cnt = __ldcg(counts + owner * l.C + c);
      if (view && cnt >= 0) continue;
      const int n = min(pieceElems, clen - off);
      char* ring = regionOf(p, p.rank) + l.reduceOff + (((long long)slot * l.P + owner) * l.blkStride + (long long)c * l.maxChunk + off) * (long long)sizeof(T);
      if (view) {
        u.src = nullptr;
        u.dst = ring;
      } else {
        u.src = cnt >= 0 ? ring : nullptr;  // chunks whose ReduceBlock never arrived stay zero (AllreduceWorker.scala:183)
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
__device__ void roleGather(const AarRankParams& p, const AarLayout& l, int cta, int nctas, BulkShared& bs, Smem& sm, unsigned char* dsm, bool nvls) {
  bulkInit(bs, dsm);
  if (nvls || eagerMode(p, l)) {
    if (threadIdx.x == 0) {
      GatherPieceRole<T> role(p, l, cta, nctas, !nvls && selfOut(p, l));
      if (role.R > 0) runCopyRole(p, l, role, bs, dsm);
      traceMark(p, l, 4);
    }
  } else {
    gatherDecisionPieces(p, l, p.round % l.L, sm);  // every thread
    if (threadIdx.x == 0) {
      traceMark(p, l, 4);
      fence_proxy_async();  // the ring was written by peers, acquired through the snapshot; bulk loads read it
      GatherSnapshotRole<T> role(p, l, cta, nctas);
      runCopyRole(p, l, role, bs, dsm);
    }
  }
}

// Runs on thread 0 of the round's last CTA, after every other CTA has finished: chunks that were given up piecewise are
// zero-filled as a whole and reported with count 0.
template <typename T>
struct ChunkCleanup {
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
  if (role & 4) roleGather<T>(p, l, cta, n, bs, sm, dsm, kNvls);
  if (phases & 4) retireRoundWith(p, l, nctas, ChunkCleanup<T>{dsm});
  traceMark(p, l, 5);
}

// ------------------------------------------------------------------------------------------------ in-switch, one hop
// For small and medium messages the two-hop in-switch round (reduce my block, multicast it, everybody gathers) is three
// dependent steps of a few microseconds each.  One hop less: every rank stages its vector, tells EVERY rank when all of it is
// staged (one flag per source), and then reduces the WHOLE vector through the switch (multimem.ld_reduce) straight into its own
// output -- no multimem.st, no reduce ring, no gather.  The switch pulls each GPU's vector P times instead of once, so this only
// pays while P x bytes over the link is cheaper than a hop (the selector table decides; measured crossover in profiles/).
//   CTAs [0, nA)   stage my input (bulk copies), the last piece's publisher tells every rank
//   the rest       wait for all P flags, in-switch reduce of my share of the pieces into the output
template <typename T>
struct StageAllRole {
  static constexpr bool kPublishes = true;
  const AarRankParams& p;
  const AarLayout& l;
  int nctas, slot, pieceElems, total, idx;
  uint32_t epoch;
  char* stage;
  bool inPlace, go;

  __device__ StageAllRole(const AarRankParams& p_, const AarLayout& l_, int cta, int nctas_, bool go_) : p(p_), l(l_), nctas(nctas_), go(go_) {
    slot = p.round % l.L;
    epoch = (uint32_t)(p.round + 1);
    pieceElems = l.tilesPerPiece * l.tileElems;
    total = (l.dataSize + pieceElems - 1) / pieceElems;
    idx = cta;
    stage = regionOf(p, p.rank) + l.stageOff + (long long)slot * paddedData(l) * (long long)sizeof(T);
    inPlace = reinterpret_cast<const char*>(p.input) == stage;
  }
  __device__ int nextUnit(CopyUnit& u, bool) {
    if (!go || idx >= total) return 0;
    const long long off = (long long)idx * pieceElems;
    const int n = (int)min((long long)pieceElems, (long long)l.dataSize - off);
    u.src = reinterpret_cast<const char*>(reinterpret_cast<const T*>(p.input) + off);
    u.dst = stage + off * (long long)sizeof(T);
    u.bytes = inPlace ? 0u : (uint32_t)n * (uint32_t)sizeof(T);
    u.a = u.b = u.c = 0;
    idx += nctas;
    return 1;
  }
  // after the publishing thread's fence: count the piece; whoever counts the last one tells every rank
  __device__ void complete(uint32_t, uint32_t, uint32_t) {
    unsigned* ctr = reinterpret_cast<unsigned*>(regionOf(p, p.rank) + l.k1CtrOff) + (long long)slot * l.P * l.C;
    unsigned old;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(old) : "l"(ctr) : "memory");
    if (old + 1 == (unsigned)total) {
      atomicExch(ctr, 0u);
      for (int i2 = 0; i2 < l.P; ++i2) {
        const int dest = (i2 + p.rank) % l.P;
        st_release_sys_u32(reinterpret_cast<uint32_t*>(regionOf(p, dest) + l.oneshotFlagOff) + slot * AAR_MAXP + p.rank, epoch);
      }
    }
  }
};

template <typename T>
__global__ void __launch_bounds__(kThreads, 1) nvlsOneHopKernel(const AarRankParams p, const AarLayout l, int nA) {
  extern __shared__ __align__(1024) unsigned char dsm[];
  __shared__ BulkShared bs;
  const int nctas = gridDim.x, b = blockIdx.x;
  const int slot = p.round % l.L;
  const uint32_t epoch = (uint32_t)(p.round + 1);
  const int P = l.P;
  traceMark(p, l, 0);
  if (b < nA) {
    bulkInit(bs, dsm);
    if (threadIdx.x == 0 || threadIdx.x == 32) {
      // Every rank reads my whole staging slot: it may be rewritten once ALL of them have retired the slot's previous tenant
      // (both threads take the same decide-once decisions).
      bool go = true;
      unsigned known = 0, goMask = 0;
      for (int d = 0; d < P; ++d) go = destDecision1(p, l, d, known, goMask, true) && go;
      StageAllRole<T> role(p, l, b, nA, go);
      if (threadIdx.x == 32) doneDrain(bs.done, role);
      else runCopyRole(p, l, role, bs, dsm);
    }
    if (threadIdx.x == 0) traceMark(p, l, 1);
  } else {
    const int cta = b - nA, n = nctas - nA;
    __syncthreads();
    if (threadIdx.x == 0) {
      const uint32_t* flags = reinterpret_cast<const uint32_t*>(regionOf(p, p.rank) + l.oneshotFlagOff) + slot * AAR_MAXP;
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
        if (backoff < 256) backoff *= 2;
      }
      for (int j = 0; j < P && ok; ++j) (void)ld_acquire_sys_u32(flags + j);
      bs.bcast[0] = ok;
      traceMark(p, l, 2);
    }
    __syncthreads();
    const int ok = bs.bcast[0];
    constexpr int V = Elem<T>::kPerVec;
    const float divisor = ((p.flags & AAR_FLAG_NORMALIZE) && P > 1) ? (float)P : 1.f;
    const char* mcStage = reinterpret_cast<const char*>(p.mcRegion) + l.stageOff + (long long)slot * paddedData(l) * (long long)sizeof(T);
    char* out = reinterpret_cast<char*>(p.output);
    const int nvTotal = l.dataSize / V;
    if (cta == 0 && p.outCounts)
      for (int i = threadIdx.x; i < l.totalChunks; i += kThreads) p.outCounts[i] = ok ? P : 0;
    // 8 x 16-byte switch loads in flight per thread; a rank that gave up (forced round, timeout) flushes zeros with count 0
    constexpr int U = 8;
    for (long long base = (long long)cta * kThreads * U + threadIdx.x; base < nvTotal; base += (long long)n * kThreads * U) {
      uint4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = base + (long long)u * kThreads;
        if (i < nvTotal) {
          if (!ok) v[u] = make_uint4(0, 0, 0, 0);
          else if (sizeof(T) == 4) v[u] = multimem_ld_reduce_f32x4(mcStage + (size_t)i * 16);
          else v[u] = multimem_ld_reduce_bf16x8(mcStage + (size_t)i * 16);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = base + (long long)u * kThreads;
        if (i < nvTotal) {
          if (ok && divisor != 1.f) {
            float f[V];
            Elem<T>::unpack(v[u], f);
#pragma unroll
            for (int k = 0; k < V; ++k) f[k] = f[k] / divisor;
            v[u] = Elem<T>::pack(f);
          }
          st_stream(out + (size_t)i * 16, v[u]);
        }
      }
    }
    if (threadIdx.x == 0) traceMark(p, l, 3);
  }
  retireRound(p, l, nctas);
  traceMark(p, l, 5);
}

// ------------------------------------------------------------------------------------------------ host side
struct KernelInfo {
  bool ready = false;
  int maxGrid = 0;
};
std::mutex g_mutex;
KernelInfo g_info[64][6];  // [device][kernel]

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
  else if (algo == 1 || algo == 2) {
    if (algo == 2 && lay->oneshotFlagOff == 0) *why = "layout without the one-shot flag words";
    else if (!p->mcRegion || lay->stageOff == 0) *why = "the in-switch path needs a multicast mapping and the staging area";
    else if (p->minRequired != lay->P || p->minChunksRequired != lay->totalChunks || (p->flags & (AAR_FLAG_SKIP_SCATTER | AAR_FLAG_EXACT_QUORUM)))
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
  if (algo < 0 || algo > 2) return aar_host::fail(AAR_EINVAL, "%s: algo must be 0 (P2P), 1 (NVLS) or 2 (NVLS, one hop)", what);
  if (algo == 2) {
    if (phases != 7) return aar_host::fail(AAR_EINVAL, "%s: the one-hop in-switch round is one launch (phases = 7)", what);
    if (lay->dtype != AAR_F32 && lay->dtype != AAR_BF16) return aar_host::fail(AAR_EINVAL, "%s: unknown dtype %d", what, lay->dtype);
    const char* why2 = nullptr;
    if (!supported(p, lay, algo, &why2)) return aar_host::fail(AAR_EUNSUPPORTED, "%s: %s", what, why2);
    const bool bf = lay->dtype == AAR_BF16;
    int maxGrid2 = 0;
    int rc2 = bf ? prepare(nvlsOneHopKernel<__nv_bfloat16>, 5, &maxGrid2) : prepare(nvlsOneHopKernel<float>, 4, &maxGrid2);
    if (rc2) return rc2;
    int g2 = grid > 0 ? grid : maxGrid2;
    if (g2 > maxGrid2) g2 = maxGrid2;
    const int pieceElems = lay->tilesPerPiece * lay->tileElems;
    const int pieces = (lay->dataSize + pieceElems - 1) / pieceElems;
    int nA2 = split & 0xffff;
    if (nA2 <= 0) nA2 = g2 / 4;
    if (nA2 > pieces) nA2 = pieces;
    if (nA2 < 1) nA2 = 1;
    const long long vecs = (long long)lay->dataSize * lay->elemSize / 16;
    long long nB2 = (vecs + (long long)kThreads * 8 - 1) / ((long long)kThreads * 8);   // one pass of 8 vectors per thread
    if (nB2 < 1) nB2 = 1;
    if (nB2 > g2 - nA2) nB2 = g2 - nA2;
    if (nB2 < 1) return aar_host::fail(AAR_EINVAL, "%s: no CTA left for the reduce", what);
    const int grid2 = nA2 + (int)nB2;
    if (bf) nvlsOneHopKernel<__nv_bfloat16><<<grid2, kThreads, kBulkDynSmem, (cudaStream_t)stream>>>(*p, *lay, nA2);
    else nvlsOneHopKernel<float><<<grid2, kThreads, kBulkDynSmem, (cudaStream_t)stream>>>(*p, *lay, nA2);
    cudaError_t e2 = cudaGetLastError();
    if (e2 != cudaSuccess) return aar_host::fail(AAR_ECUDA, "%s: launch failed: %s", what, cudaGetErrorString(e2));
    return AAR_OK;
  }
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
        nA = g / 9;                   // 16 of 148: remote copies only, and 16 CTAs already saturate the link (measured, 2 GPUs)
        nB = g / 2;
      }
    }
    if (nA < 1) nA = 1;
    if (nB < 1) nB = 1;
    if (nA + nB >= g) return aar_host::fail(AAR_EINVAL, "%s: split %d + %d leaves no gather CTA in a grid of %d", what, nA, nB, g);
    // A role never needs more CTAs than it has pieces to move: small messages get a small grid (less to launch, to initialise
    // and to wait for at retirement).
    const int Sp = lay->piecesPerChunk, P = lay->P;
    const bool sd = (p->flags & AAR_FLAG_SELF_DIRECT) && algo == 0;
    const int remote = sd ? P - 1 : P;
    long long unitsA = (long long)lay->C * (remote > 0 ? remote : 1) * Sp, unitsB = (long long)lay->C * Sp, unitsC = (long long)lay->C * P * Sp;
    int nC = g - nA - nB;
    if (unitsA < nA) nA = (int)(unitsA < 1 ? 1 : unitsA);
    if (unitsB < nB) nB = (int)(unitsB < 1 ? 1 : unitsB);
    if (unitsC < nC) nC = (int)(unitsC < 1 ? 1 : unitsC);
    g = nA + nB + nC;
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
