// Executable model of the DEVICE protocol (docs/PROTOCOL.md) on host threads, meant to be run under ThreadSanitizer.
//
// SURVEY.md section 5 asks for litmus tests of the flag protocol.  compute-sanitizer is closed on the build pool, so the
// protocol -- not the GPU memory model -- is checked here: every "CTA" is a host thread, every rank's symmetric region is
// a struct of plain float rings plus std::atomic flag / decision / election words, and the phases below follow
// csrc/cuda/aar_kernels.cu step for step (same index math, same publish / observe / decide-once / last-finisher
// structure, same flow control, two rounds of one rank in flight, in-order retirement).  The ring payload is deliberately
// NON-atomic: if any path reads or overwrites it without a happens-before edge through the flags, TSAN reports a race.
// Device fences + relaxed atomics are modelled by acquire/release on the atomic operations themselves (TSAN does not
// model std::atomic_thread_fence); that yields the same happens-before edges.
//
// Checks per run: no TSAN report; every flushed vector equals the peer-index-ordered fp32 sum over exactly the
// contributor set its owner decided (bit-exact); with thresholds = 1 that set is everybody; counts match; no deadlock.
//
// Degraded mode (`lagger` = 1): one rank is slow in EVERY round, so it falls out of the maxLag window.  That exercises what the
// straggler-free runs never touch: waits that give up (the device-side form of the catch-up loop: a peer is known to have
// retired a round more than maxLag ahead, or a spin budget standing in for the %globaltimer deadline), pushes that are
// dropped by the per-destination decide-once word, the broadcast adopting those decisions, forced chunk decisions over
// an incomplete contributor set, and gathers that zero-fill what never arrived.  Extra check there: a chunk is flushed
// either whole or as zeros, never half and half (`eagerDecision` = 0 reproduces that hole in the eager gather).
//
//   protocol_litmus [seed] [P] [rounds] [thresholdPercent] [maxLag] [overlap 0|1] [ctasPerKernel] [bcastFlowControl 0|1]
//                   [lagger 0|1] [eagerDecision 0|1] [spinBudget] [elide 0|1] [nvls 0|1|2] [oneshot 0|1]
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <random>
#include <thread>
#include <vector>

namespace {

struct Barrier {  // kernel boundary / __syncthreads stand-in
  explicit Barrier(int n) : n_(n) {}
  void wait() {
    std::unique_lock<std::mutex> g(m_);
    int gen = gen_;
    if (++count_ == n_) {
      count_ = 0;
      ++gen_;
      cv_.notify_all();
    } else {
      cv_.wait(g, [&] { return gen != gen_; });
    }
  }
  std::mutex m_;
  std::condition_variable cv_;
  int n_, count_ = 0, gen_ = 0;
};

struct Config {
  int P = 4, L = 2, C = 3, S = 2;   // ranks, ring slots (maxLag+1), chunks per block, segments per message
  int chunkLen = 8;                // elements per chunk (segment = chunkLen / S)
  int G = 4;                       // CTAs per kernel
  int rounds = 12;
  int minRequired = 4, minChunks = 12;
  bool overlap = true;
  bool bcastFlowControl = true;    // K2 adopts K1's per-destination decision (wait for the slot, or skip a dropped destination)
  bool lagger = false;             // rank 0 is slow in every round: degraded-mode paths (give-ups, drops, forced decisions)
  bool eagerDecision = true;       // eager gather: one decision per (round, chunk) shared by all CTAs (proposed fix)
  int nvls = 0;                    // 1 = NVLS path as shipped (staging slots without flow control, the switch sums every rank),
                                   // 2 = strict NVLS (staging obeys the slot flow control, incomplete decisions are not published)
                                   // 3 = in-switch round in one hop (nvlsOneHopKernel), normal mode only
  bool oneshot = false;            // K5: every rank pushes its whole vector to every peer and sums the arrived set locally
  bool elide = false;              // a rank that is a whole round late skips work nobody will read (AAR_FLAG_ELIDE_LATE)
  int spinBudget = 1 << 30;        // stand-in for the %globaltimer deadline of a wait (default: timeoutNs = 0, unbounded)
  unsigned seed = 1;
  int blockLen() const { return C * chunkLen; }
  int dataSize() const { return P * blockLen(); }
};

struct Region {
  std::vector<float> scatter, reduce;                 // [L][P][blockLen]   plain memory on purpose
  std::vector<float> stage;                           // [L][dataSize]      NVLS staging area (read by the owners through the switch)
  std::vector<float> oneshot;                         // [L][P][dataSize]   K5 one-shot buffers, written by peers
  std::vector<std::atomic<uint32_t>> osflag;          // [L][P]
  std::vector<std::atomic<uint32_t>> sflag;           // [L][C][P]
  std::vector<std::atomic<uint64_t>> rflag;           // [L][P][C]
  std::vector<std::atomic<uint32_t>> peerBase;        // [P]     newest round each peer has retired (+1): catch-up predicate only
  std::vector<std::atomic<uint32_t>> slotBase;        // [L][P]  per ring slot: round + 1 of the last round each peer retired there (flow control)
  std::vector<std::atomic<uint32_t>> slotFlushed;     // [L]     this rank's completion word per slot (what the host polls)
  std::vector<std::atomic<uint32_t>> k1Ctr;           // [L][P][C]
  std::vector<std::atomic<uint64_t>> k1Dec;           // [L][P]   epoch << 32 | go
  std::vector<std::atomic<uint64_t>> egDec;           // [L][P*C] epoch << 32 | (count + 1), 0 = abandoned   (eager gather)
  std::vector<std::atomic<uint64_t>> k2Dec;           // [L][C]
  std::vector<std::atomic<uint32_t>> k2Ctr;           // [L][C]
  std::vector<std::atomic<uint64_t>> k3Dec;           // [L]
  std::vector<std::atomic<uint32_t>> k3Ctr;           // [L]
  std::vector<int> counts;                            // [L][P*C]           plain, published through k3Dec
  std::atomic<uint32_t> flushed{0};
  std::vector<std::atomic<int>> consumed;             // [L] newest round whose output slot the host has consumed (-1 = none)
  std::vector<std::vector<float>> output;             // [L][dataSize]      what the sink would see
  std::vector<std::vector<int>> outCounts;            // [L][P*C]
  explicit Region(const Config& c)
      : scatter((size_t)c.L * c.P * c.blockLen()), reduce((size_t)c.L * c.P * c.blockLen()), stage((size_t)c.L * c.dataSize()),
        oneshot(c.oneshot ? (size_t)c.L * c.P * c.dataSize() : 0), osflag((size_t)c.L * c.P), sflag((size_t)c.L * c.C * c.P),
        rflag((size_t)c.L * c.P * c.C), peerBase(c.P), slotBase((size_t)c.L * c.P), slotFlushed(c.L), k1Ctr((size_t)c.L * c.P * c.C), k1Dec((size_t)c.L * c.P),
        egDec((size_t)c.L * c.P * c.C), k2Dec((size_t)c.L * c.C), k2Ctr((size_t)c.L * c.C),
        k3Dec(c.L), k3Ctr(c.L), counts((size_t)c.L * c.P * c.C), consumed(c.L), output(c.L, std::vector<float>(c.dataSize())),
        outCounts(c.L, std::vector<int>(c.P * c.C)) {
    for (auto& x : consumed) x.store(-1);
  }
};

struct World {
  Config cfg;
  std::vector<Region*> region;
  std::vector<std::vector<std::vector<float>>> input;  // [round][rank][dataSize]
  std::atomic<long> failures{0};
  std::atomic<long> drops{0}, forcedChunks{0}, forcedRounds{0}, abandoned{0};  // how often the degraded paths ran
  std::atomic<long> elidedPushes{0}, elidedBcasts{0};
};

float inputValue(int round, int rank, int i) { return (float)((round * 131 + rank * 17 + i * 7) % 1000) * 0.125f + (float)rank; }

void jitter(std::mt19937& rng) {
  unsigned v = rng() % 64;
  if (v == 0) std::this_thread::sleep_for(std::chrono::microseconds(rng() % 200));
  else if (v < 8) std::this_thread::yield();
}

// forcedByHost without a host: a peer is known to have retired a round more than maxLag ahead of this one
bool forced(World& w, int rank, int round) {
  const Config& c = w.cfg;
  int newestDone = -1;
  for (int j = 0; j < c.P; ++j) newestDone = std::max(newestDone, (int)w.region[rank]->peerBase[j].load(std::memory_order_relaxed) - 1);
  return newestDone - round > c.L - 1;
}

// One go / drop decision per (round, dest), shared by every CTA of the rank (scatterDecision in aar_kernels.cu)
bool scatterDecision(World& w, int rank, int round, int dest) {
  const Config& c = w.cfg;
  if (round < c.L) return true;
  Region& my = *w.region[rank];
  std::atomic<uint64_t>& word = my.k1Dec[(size_t)(round % c.L) * c.P + dest];
  const uint64_t epoch = (uint64_t)round + 1;
  uint64_t v = word.load(std::memory_order_acquire);
  if ((v >> 32) == epoch) return v & 1;
  uint64_t go = 1;
  int spins = 0;
  while ((int)my.slotBase[(size_t)(round % c.L) * c.P + dest].load(std::memory_order_acquire) < round - c.L + 1) {
    if ((++spins & 63) == 0 && (forced(w, rank, round) || spins > c.spinBudget)) {
      go = 0;  // the peer fell out of the window: its copy of this contribution is dropped
      w.drops++;
      break;
    }
    std::this_thread::yield();
  }
  if (!word.compare_exchange_strong(v, (epoch << 32) | go, std::memory_order_acq_rel, std::memory_order_acquire)) {
    if ((v >> 32) == epoch) return v & 1;
    w.failures++;
  }
  return go != 0;
}

// ---------------------------------------------------------------------------------------------- K1 scatter_push
void phaseScatter(World& w, int rank, int round, int cta, int nctas, std::mt19937& rng) {
  const Config& c = w.cfg;
  const int slot = round % c.L;
  const uint32_t epoch = (uint32_t)round + 1;
  Region& my = *w.region[rank];
  const int segLen = c.chunkLen / c.S;
  const long total = (long)c.C * c.P * c.S;
  // Late-work elision: some peer has already RETIRED this round, so this rank is at least a round behind.  Owners that
  // have decided a chunk's contributor set without this rank will never read its copy: do not send it.
  bool late = false;
  if (c.elide)
    for (int j = 0; j < c.P; ++j) late = late || (int)my.peerBase[j].load(std::memory_order_relaxed) > round;
  for (long idx = cta; idx < total; idx += nctas) {
    const int i = (int)(idx % c.P), s = (int)((idx / c.P) % c.S), ch = (int)(idx / ((long)c.S * c.P));
    const int dest = (i + rank) % c.P;
    Region& peer = *w.region[dest];
    // flow control: the previous tenant of this slot at `dest` must be retired -- or the push is dropped.  The shipped NVLS
    // path stages without it (the owner of a block is the only reader of that part of the staging slot, and at thresholds
    // = 1 it has finished reading round R - L before this rank could retire R - L); the strict variant keeps the check.
    if (c.nvls != 1 && !scatterDecision(w, rank, round, dest)) continue;
    jitter(rng);
    if (c.nvls) {
      const float* src = w.input[round][rank].data() + dest * c.blockLen() + ch * c.chunkLen + s * segLen;
      float* dst = my.stage.data() + (size_t)slot * c.dataSize() + dest * c.blockLen() + ch * c.chunkLen + s * segLen;
      for (int k = 0; k < segLen; ++k) dst[k] = src[k];
      std::atomic<uint32_t>& sctr = my.k1Ctr[(size_t)(slot * c.P + dest) * c.C + ch];
      if (sctr.fetch_add(1, std::memory_order_acq_rel) + 1 == (uint32_t)c.S) {  // only the flag travels
        sctr.store(0, std::memory_order_relaxed);
        peer.sflag[(size_t)(slot * c.C + ch) * c.P + rank].store(epoch, std::memory_order_release);
      }
      continue;
    }
    bool skip = false;
    if (late) {  // one remote load of the owner's decision word
      const uint64_t d = peer.k2Dec[(size_t)slot * c.C + ch].load(std::memory_order_relaxed);
      skip = (d >> 32) == (uint64_t)epoch && !((d >> rank) & 1);
    }
    if (!skip) {
      const float* src = w.input[round][rank].data() + dest * c.blockLen() + ch * c.chunkLen + s * segLen;
      float* dst = peer.scatter.data() + ((size_t)(slot * c.P + rank) * c.blockLen()) + ch * c.chunkLen + s * segLen;
      for (int k = 0; k < segLen; ++k) dst[k] = src[k];
    } else {
      w.elidedPushes++;
    }
    // Segments still count (the election word must be back at zero when the slot is reused); a skipped one taints the
    // message in the upper half of the word, and a tainted message is never flagged.
    std::atomic<uint32_t>& ctr = my.k1Ctr[(size_t)(slot * c.P + dest) * c.C + ch];
    const uint32_t old = ctr.fetch_add(skip ? 0x10001u : 1u, std::memory_order_acq_rel);
    if ((old & 0xffffu) + 1 == (uint32_t)c.S) {  // last segment: publish the ScatterBlock
      ctr.store(0, std::memory_order_relaxed);
      if (!skip && (old >> 16) == 0) peer.sflag[(size_t)(slot * c.C + ch) * c.P + rank].store(epoch, std::memory_order_release);
    }
  }
}

// ---------------------------------------------------------------------------------------------- K2 reduce_bcast
unsigned chunkDecision(World& w, int rank, int round, int ch, std::mt19937& rng) {
  const Config& c = w.cfg;
  const int slot = round % c.L;
  const uint64_t epoch = (uint64_t)round + 1;
  Region& my = *w.region[rank];
  std::atomic<uint64_t>& word = my.k2Dec[(size_t)slot * c.C + ch];
  uint64_t v = word.load(std::memory_order_acquire);
  if ((v >> 32) == epoch) return (unsigned)v;
  unsigned arrived = 0;
  int spins = 0;
  for (;;) {
    arrived = 0;
    for (int src = 0; src < c.P; ++src)
      if (my.sflag[(size_t)(slot * c.C + ch) * c.P + src].load(std::memory_order_acquire) == (uint32_t)epoch) arrived |= 1u << src;
    if (__builtin_popcount(arrived) >= c.minRequired) break;
    if ((++spins & 31) == 0 && (forced(w, rank, round) || spins > c.spinBudget)) {
      w.forcedChunks++;
      arrived = 0;  // forced: one more look so the snapshot is as fresh as the decision, then go with what is there
      for (int src = 0; src < c.P; ++src)
        if (my.sflag[(size_t)(slot * c.C + ch) * c.P + src].load(std::memory_order_acquire) == (uint32_t)epoch) arrived |= 1u << src;
      break;
    }
    jitter(rng);
    std::this_thread::yield();
  }
  uint64_t want = (epoch << 32) | arrived;
  if (!word.compare_exchange_strong(v, want, std::memory_order_acq_rel, std::memory_order_acquire)) {
    if ((v >> 32) == epoch) return (unsigned)v;  // somebody else decided first: adopt
    // word still belongs to an older round (cannot happen: only this round's CTAs write it now)
    w.failures++;
  }
  return arrived;
}

void phaseReduce(World& w, int rank, int round, int cta, int nctas, std::mt19937& rng) {
  const Config& c = w.cfg;
  const int slot = round % c.L;
  const uint64_t epoch = (uint64_t)round + 1;
  Region& my = *w.region[rank];
  const int segLen = c.chunkLen / c.S;
  // Flow control for the broadcast.  With thresholds < 1 an owner can reduce round R over a quorum that does not contain
  // a lagging peer and would then overwrite that peer's reduce-ring slot while it still gathers round R - L (found by
  // this model: data race between the store below and copyChunk).  So the broadcast adopts the scatter's per-destination
  // decision: it waits for it like a push CTA would, and leaves a dropped destination alone (bcastDestMask in the kernels).
  unsigned destMask = ~0u;
  if (c.bcastFlowControl && !c.nvls && cta < (long)c.C * c.S) {
    destMask = 0;
    for (int i = 0; i < c.P; ++i) {
      const int d = (i + rank) % c.P;
      if (d == rank || scatterDecision(w, rank, round, d)) destMask |= 1u << d;
    }
  }
  // Late-work elision: a destination that has already retired this round will never gather this chunk
  if (c.elide && cta < (long)c.C * c.S)
    for (int d = 0; d < c.P; ++d)
      if (d != rank && (int)my.slotBase[(size_t)(round % c.L) * c.P + d].load(std::memory_order_acquire) > round && (destMask >> d & 1)) {
        destMask &= ~(1u << d);
        w.elidedBcasts++;
      }
  for (long idx = cta; idx < (long)c.C * c.S; idx += nctas) {
    const int s = (int)(idx % c.S), ch = (int)(idx / c.S);
    const unsigned mask = chunkDecision(w, rank, round, ch, rng);
    const int count = __builtin_popcount(mask);
    jitter(rng);
    if (c.nvls) {
      // The switch cannot leave a rank out.  Shipped kernel: sums every staging slot whatever the decision says.  Strict: a
      // decision forced over an incomplete set is not carried out and not published (every gather gives up on the chunk).
      const bool incomplete = count != c.P;
      if (!(c.nvls == 2 && incomplete)) {
        for (int k = 0; k < segLen; ++k) {
          const int e = ch * c.chunkLen + s * segLen + k;
          float acc = 0.f;
          for (int src = 0; src < c.P; ++src) acc += w.region[src]->stage[(size_t)slot * c.dataSize() + rank * c.blockLen() + e];  // multimem.ld_reduce
          for (int d = 0; d < c.P; ++d) w.region[d]->reduce[(size_t)(slot * c.P + rank) * c.blockLen() + e] = acc;               // multimem.st
        }
      }
      std::atomic<uint32_t>& nctr = my.k2Ctr[(size_t)slot * c.C + ch];
      if (nctr.fetch_add(1, std::memory_order_acq_rel) + 1 == (uint32_t)c.S) {
        nctr.store(0, std::memory_order_relaxed);
        if (!(c.nvls == 2 && incomplete))
          for (int d = 0; d < c.P; ++d)
            w.region[d]->rflag[(size_t)(slot * c.P + rank) * c.C + ch].store((epoch << 32) | (uint64_t)count, std::memory_order_release);
      }
      continue;
    }
    for (int k = 0; k < segLen; ++k) {
      const int e = ch * c.chunkLen + s * segLen + k;
      float acc = 0.f;
      for (int src = 0; src < c.P; ++src)
        if (mask >> src & 1) acc += my.scatter[(size_t)(slot * c.P + src) * c.blockLen() + e];
      for (int i = 0; i < c.P; ++i) {
        const int d = (i + rank) % c.P;
        if (!(destMask >> d & 1)) continue;
        w.region[d]->reduce[(size_t)(slot * c.P + rank) * c.blockLen() + e] = acc;
      }
    }
    std::atomic<uint32_t>& ctr = my.k2Ctr[(size_t)slot * c.C + ch];
    if (ctr.fetch_add(1, std::memory_order_acq_rel) + 1 == (uint32_t)c.S) {  // last segment: publish the ReduceBlock + count
      ctr.store(0, std::memory_order_relaxed);
      for (int i = 0; i < c.P; ++i) {
        const int d = (i + rank) % c.P;
        if (!(destMask >> d & 1)) continue;  // no data was written there: no flag either
        w.region[d]->rflag[(size_t)(slot * c.P + rank) * c.C + ch].store((epoch << 32) | (uint64_t)count, std::memory_order_release);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------- K3 gather_flush
void copyChunk(World& w, int rank, int round, int owner, int ch, int s, int cnt) {
  const Config& c = w.cfg;
  const int slot = round % c.L, segLen = c.chunkLen / c.S;
  Region& my = *w.region[rank];
  for (int k = 0; k < segLen; ++k) {
    const int e = ch * c.chunkLen + s * segLen + k;
    my.output[slot][owner * c.blockLen() + e] = cnt >= 0 ? my.reduce[(size_t)(slot * c.P + owner) * c.blockLen() + e] : 0.f;
  }
}

void phaseGather(World& w, int rank, int round, int cta, int nctas, std::mt19937& rng) {
  const Config& c = w.cfg;
  const int slot = round % c.L;
  const uint64_t epoch = (uint64_t)round + 1;
  Region& my = *w.region[rank];
  const bool eager = c.minChunks >= c.P * c.C;
  if (!eager) {  // one leader counts arrivals against the round threshold and publishes a snapshot
    const uint64_t busy = (epoch << 32) | 1, done = (epoch << 32) | 2;
    std::atomic<uint64_t>& word = my.k3Dec[slot];
    uint64_t v = word.load(std::memory_order_acquire);
    bool leader = false;
    if (v != busy && v != done) leader = word.compare_exchange_strong(v, busy, std::memory_order_acq_rel, std::memory_order_acquire);
    if (leader) {
      int spins = 0;
      for (;;) {
        int n = 0;
        for (int i = 0; i < c.P * c.C; ++i)
          if ((my.rflag[(size_t)slot * c.P * c.C + i].load(std::memory_order_acquire) >> 32) == epoch) ++n;
        if (n >= c.minChunks) break;
        if ((++spins & 31) == 0 && (forced(w, rank, round) || spins > c.spinBudget)) {
          w.forcedRounds++;
          break;
        }
        jitter(rng);
        std::this_thread::yield();
      }
      for (int i = 0; i < c.P * c.C; ++i) {
        const uint64_t f = my.rflag[(size_t)slot * c.P * c.C + i].load(std::memory_order_acquire);
        const int cnt = (f >> 32) == epoch ? (int)(uint32_t)f : -1;
        my.counts[(size_t)slot * c.P * c.C + i] = cnt;
        my.outCounts[slot][i] = cnt < 0 ? 0 : cnt;
      }
      word.store(done, std::memory_order_release);
    } else {
      while (word.load(std::memory_order_acquire) != done) std::this_thread::yield();
    }
  }
  for (long idx = cta; idx < (long)c.P * c.C * c.S; idx += nctas) {
    const int s = (int)(idx % c.S), owner = ((int)((idx / c.S) % c.P) + rank) % c.P, ch = (int)(idx / ((long)c.S * c.P));
    int cnt;
    if (eager) {
      const int key = owner * c.C + ch;
      std::atomic<uint64_t>& dec = my.egDec[(size_t)slot * c.P * c.C + key];
      uint64_t d = c.eagerDecision ? dec.load(std::memory_order_acquire) : 0;
      if (c.eagerDecision && (d >> 32) == epoch) {
        cnt = (int)(uint32_t)d - 1;
        if (cnt >= 0) (void)my.rflag[(size_t)(slot * c.P + owner) * c.C + ch].load(std::memory_order_acquire);  // acquire the data
      } else {
        uint64_t f;
        int spins = 0;
        cnt = -1;
        for (;;) {
          f = my.rflag[(size_t)(slot * c.P + owner) * c.C + ch].load(std::memory_order_acquire);
          if ((f >> 32) == epoch) {
            cnt = (int)(uint32_t)f;
            break;
          }
          if ((++spins & 31) == 0 && (forced(w, rank, round) || spins > c.spinBudget)) {  // give up: zero-fill
            w.abandoned++;
            break;
          }
          std::this_thread::yield();
        }
        if (c.eagerDecision) {  // first CTA to stop waiting decides for everybody: arrived with this count, or abandoned
          if (!dec.compare_exchange_strong(d, (epoch << 32) | (uint64_t)(cnt + 1), std::memory_order_acq_rel, std::memory_order_acquire)) {
            if ((d >> 32) != epoch) w.failures++;
            cnt = (int)(uint32_t)d - 1;
            if (cnt >= 0) (void)my.rflag[(size_t)(slot * c.P + owner) * c.C + ch].load(std::memory_order_acquire);
          }
        }
      }
      if (s == 0) my.outCounts[slot][key] = cnt < 0 ? 0 : cnt;
    } else {
      cnt = my.counts[(size_t)slot * c.P * c.C + owner * c.C + ch];
    }
    jitter(rng);
    copyChunk(w, rank, round, owner, ch, s, cnt);
  }
}

// ---------------------------------------------------------------------------------------------- K5 one-shot
// Tiles of C * chunkLen / S elements; one arrival flag per (round, source).  The contributor decision reuses chunkDecision's
// word for chunk 0 of the slot, as the kernel does.
void phaseOneshot(World& w, int rank, int round, int cta, int nctas, std::mt19937& rng) {
  const Config& c = w.cfg;
  const int slot = round % c.L;
  const uint64_t epoch = (uint64_t)round + 1;
  Region& my = *w.region[rank];
  const int tile = c.chunkLen / c.S, Tn = c.dataSize() / tile;
  for (long idx = cta; idx < (long)c.P * Tn; idx += nctas) {  // push
    const int t = (int)(idx % Tn), dest = ((int)(idx / Tn) + rank) % c.P;
    if (!scatterDecision(w, rank, round, dest)) continue;
    jitter(rng);
    Region& peer = *w.region[dest];
    float* dst = peer.oneshot.data() + (size_t)(slot * c.P + rank) * c.dataSize() + (size_t)t * tile;
    for (int k = 0; k < tile; ++k) dst[k] = w.input[round][rank][(size_t)t * tile + k];
    std::atomic<uint32_t>& ctr = my.k1Ctr[(size_t)(slot * c.P + dest) * c.C];
    if (ctr.fetch_add(1, std::memory_order_acq_rel) + 1 == (uint32_t)Tn) {
      ctr.store(0, std::memory_order_relaxed);
      peer.osflag[(size_t)slot * c.P + rank].store((uint32_t)epoch, std::memory_order_release);
    }
  }
  // reduce: one contributor decision per round
  std::atomic<uint64_t>& word = my.k2Dec[(size_t)slot * c.C];
  uint64_t v = word.load(std::memory_order_acquire);
  unsigned mask;
  if ((v >> 32) == epoch) {
    mask = (unsigned)v;
  } else {
    int spins = 0;
    for (;;) {
      mask = 0;
      for (int src = 0; src < c.P; ++src)
        if (my.osflag[(size_t)slot * c.P + src].load(std::memory_order_acquire) == (uint32_t)epoch) mask |= 1u << src;
      if (__builtin_popcount(mask) >= c.minRequired) break;
      if ((++spins & 31) == 0 && (forced(w, rank, round) || spins > c.spinBudget)) {
        w.forcedChunks++;
        mask = 0;
        for (int src = 0; src < c.P; ++src)
          if (my.osflag[(size_t)slot * c.P + src].load(std::memory_order_acquire) == (uint32_t)epoch) mask |= 1u << src;
        break;
      }
      std::this_thread::yield();
    }
    if (!word.compare_exchange_strong(v, (epoch << 32) | mask, std::memory_order_acq_rel, std::memory_order_acquire)) {
      if ((v >> 32) == epoch) mask = (unsigned)v;
      else w.failures++;
    }
  }
  const int count = __builtin_popcount(mask);
  if (cta == 0)
    for (int i = 0; i < c.P * c.C; ++i) my.outCounts[slot][i] = count;
  for (int t = cta; t < Tn; t += nctas) {
    jitter(rng);
    for (int k = 0; k < tile; ++k) {
      float acc = 0.f;
      for (int src = 0; src < c.P; ++src)
        if (mask >> src & 1) acc += my.oneshot[(size_t)(slot * c.P + src) * c.dataSize() + (size_t)t * tile + k];
      my.output[slot][(size_t)t * tile + k] = acc;
    }
  }
}

void verifyOneshot(World& w, int rank, int round) {
  const Config& c = w.cfg;
  const int slot = round % c.L;
  Region& my = *w.region[rank];
  const uint64_t dec = my.k2Dec[(size_t)slot * c.C].load(std::memory_order_acquire);
  if ((dec >> 32) != (uint64_t)round + 1) {
    w.failures++;
    return;
  }
  if (__builtin_popcount((unsigned)dec) != my.outCounts[slot][0]) w.failures++;
  for (int e = 0; e < c.dataSize(); ++e) {
    float want = 0.f;
    for (int src = 0; src < c.P; ++src)
      if ((unsigned)dec >> src & 1) want += w.input[round][src][e];
    if (my.output[slot][e] != want) w.failures++;
  }
}

// ---------------------------------------------------------------------------------------------- in-switch round in one hop
// nvlsOneHopKernel (csrc/cuda/aar_bulk.cu): stage CTAs copy the input into this rank's staging slot once EVERY rank has retired the
// slot's previous tenant (all of them read it), the CTA that finishes the last tile flags every rank; the other CTAs wait for all
// P flags and sum the P staging slots (the switch's reads of peer memory) straight into the output.  Normal mode only: the
// kernel's give-up decision is per CTA, so a forced round is outside what this model checks (docs/PROTOCOL.md).
void phaseOneHop(World& w, int rank, int round, int cta, int nctas, std::mt19937& rng) {
  const Config& c = w.cfg;
  const int slot = round % c.L;
  const uint32_t epoch = (uint32_t)round + 1;
  Region& my = *w.region[rank];
  const int tile = c.chunkLen / c.S, Tn = c.dataSize() / tile;
  const int nA = std::max(1, nctas / 2);
  if (cta < nA) {
    bool go = true;
    for (int d = 0; d < c.P; ++d) go = scatterDecision(w, rank, round, d) && go;
    if (!go) return;
    for (int t = cta; t < Tn; t += nA) {
      jitter(rng);
      float* dst = my.stage.data() + (size_t)slot * c.dataSize() + (size_t)t * tile;
      for (int k = 0; k < tile; ++k) dst[k] = w.input[round][rank][(size_t)t * tile + k];
      std::atomic<uint32_t>& ctr = my.k1Ctr[(size_t)slot * c.P * c.C];
      if (ctr.fetch_add(1, std::memory_order_acq_rel) + 1 == (uint32_t)Tn) {
        ctr.store(0, std::memory_order_relaxed);
        for (int i = 0; i < c.P; ++i) w.region[(i + rank) % c.P]->osflag[(size_t)slot * c.P + rank].store(epoch, std::memory_order_release);
      }
    }
    return;
  }
  int spins = 0;
  for (;;) {
    int have = 0;
    for (int src = 0; src < c.P; ++src) have += my.osflag[(size_t)slot * c.P + src].load(std::memory_order_acquire) == epoch;
    if (have == c.P) break;
    if ((++spins & 31) == 0 && spins > c.spinBudget) {
      w.failures++;   // normal mode: nobody may give up
      return;
    }
    std::this_thread::yield();
  }
  if (cta == nA)
    for (int i = 0; i < c.P * c.C; ++i) my.outCounts[slot][i] = c.P;
  for (int t = cta - nA; t < Tn; t += nctas - nA) {
    jitter(rng);
    for (int k = 0; k < tile; ++k) {
      const size_t e = (size_t)t * tile + k;
      float acc = 0.f;
      for (int src = 0; src < c.P; ++src) acc += w.region[src]->stage[(size_t)slot * c.dataSize() + e];   // multimem.ld_reduce
      my.output[slot][e] = acc;
    }
  }
}

void verifyOneHop(World& w, int rank, int round) {
  const Config& c = w.cfg;
  const int slot = round % c.L;
  Region& my = *w.region[rank];
  if (my.outCounts[slot][0] != c.P) w.failures++;
  for (int e = 0; e < c.dataSize(); ++e) {
    float want = 0.f;
    for (int src = 0; src < c.P; ++src) want += w.input[round][src][e];
    if (my.output[slot][e] != want) w.failures++;
  }
}

// K6: last CTA of the kernel retires the slot.  Retirement is per ring slot (peers' flow control reads slotBase[slot][me]), so
// with two rounds in flight round r+1 may retire before round r; peerBase[me] is only raised
void retire(World& w, int rank, int round, int nctas) {
  const Config& c = w.cfg;
  Region& my = *w.region[rank];
  std::atomic<uint32_t>& ctr = my.k3Ctr[round % c.L];
  if (ctr.fetch_add(1, std::memory_order_acq_rel) + 1 == (uint32_t)nctas) {
    ctr.store(0, std::memory_order_relaxed);
    for (int i = 0; i < c.P; ++i) {
      Region& peer = *w.region[(i + rank) % c.P];
      peer.slotBase[(size_t)(round % c.L) * c.P + rank].store((uint32_t)round + 1, std::memory_order_release);
      uint32_t cur = peer.peerBase[rank].load(std::memory_order_relaxed);
      while (cur < (uint32_t)round + 1 && !peer.peerBase[rank].compare_exchange_weak(cur, (uint32_t)round + 1, std::memory_order_relaxed)) {
      }
    }
    my.slotFlushed[round % c.L].store((uint32_t)round + 1, std::memory_order_release);
    uint32_t f = my.flushed.load(std::memory_order_relaxed);
    while (f < (uint32_t)round + 1 && !my.flushed.compare_exchange_weak(f, (uint32_t)round + 1, std::memory_order_release)) {
    }
  }
}

// What the host does when the completion word moves: check the flushed vector against the oracle over the DECIDED masks.
void verifyRound(World& w, int rank, int round) {
  const Config& c = w.cfg;
  const int slot = round % c.L;
  Region& my = *w.region[rank];
  for (int owner = 0; owner < c.P; ++owner)
    for (int ch = 0; ch < c.C; ++ch) {
      const int cnt = my.outCounts[slot][owner * c.C + ch];
      const uint64_t dec = w.region[owner]->k2Dec[(size_t)slot * c.C + ch].load(std::memory_order_acquire);
      for (int k = 0; k < c.chunkLen; ++k) {
        const int e = owner * c.blockLen() + ch * c.chunkLen + k;
        float want = 0.f;
        if (cnt > 0) {
          if ((dec >> 32) != (uint64_t)round + 1) {  // the owner has moved on and reused the word: cannot re-derive, skip
            continue;
          }
          for (int src = 0; src < c.P; ++src)
            if ((unsigned)dec >> src & 1) want += w.input[round][src][e];
          if (__builtin_popcount((unsigned)dec) != cnt) w.failures++;
        }
        if (my.output[slot][e] != want) w.failures++;
      }
    }
}

}  // namespace

int main(int argc, char** argv) {
  World w;
  Config& c = w.cfg;
  c.seed = argc > 1 ? (unsigned)atoi(argv[1]) : 1;
  c.P = argc > 2 ? atoi(argv[2]) : 4;
  c.rounds = argc > 3 ? atoi(argv[3]) : 12;
  const int thPct = argc > 4 ? atoi(argv[4]) : 100;
  c.L = (argc > 5 ? atoi(argv[5]) : 1) + 1;
  c.overlap = argc > 6 ? atoi(argv[6]) != 0 : true;
  c.G = argc > 7 ? atoi(argv[7]) : 4;
  c.bcastFlowControl = argc > 8 ? atoi(argv[8]) != 0 : true;
  c.lagger = argc > 9 ? atoi(argv[9]) != 0 : false;
  c.eagerDecision = argc > 10 ? atoi(argv[10]) != 0 : true;
  c.spinBudget = argc > 11 ? atoi(argv[11]) : (1 << 30);
  c.elide = argc > 12 ? atoi(argv[12]) != 0 : false;
  c.nvls = argc > 13 ? atoi(argv[13]) : 0;
  c.oneshot = argc > 14 ? atoi(argv[14]) != 0 : false;
  c.minRequired = std::max(1, (int)((float)thPct / 100.f * (float)c.P));
  c.minChunks = std::max(1, (int)((float)thPct / 100.f * (float)(c.P * c.C)));
  if (c.L < 2) c.overlap = false;
  for (int r = 0; r < c.P; ++r) w.region.push_back(new Region(c));
  w.input.assign(c.rounds, std::vector<std::vector<float>>(c.P, std::vector<float>(c.dataSize())));
  for (int r = 0; r < c.rounds; ++r)
    for (int k = 0; k < c.P; ++k)
      for (int i = 0; i < c.dataSize(); ++i) w.input[r][k][i] = inputValue(r, k, i);

  const int streams = c.overlap ? 2 : 1;
  std::vector<std::thread> threads;
  std::vector<Barrier*> barriers;
  for (int rank = 0; rank < c.P; ++rank)
    for (int st = 0; st < streams; ++st) {
      Barrier* b = new Barrier(c.G);
      barriers.push_back(b);
      for (int cta = 0; cta < c.G; ++cta)
        threads.emplace_back([&w, rank, st, cta, streams, b] {
          const Config& cf = w.cfg;
          std::mt19937 rng(cf.seed * 7919u + rank * 131u + st * 17u + cta);
          for (int round = st; round < cf.rounds; round += streams) {  // kernels of one stream run one after another
            // the worker launches round R only after the sink has consumed round R - L (same output / count buffers)
            if (round >= cf.L)
              while (w.region[rank]->consumed[round % cf.L].load(std::memory_order_acquire) != round - cf.L) std::this_thread::yield();
            // ... and only after EVERY round up to R - L of this rank has been flushed (the engine makes round R's stream wait for
            // them): with per-slot retirement nothing else keeps one stream from running arbitrarily far ahead of the other, and
            // a peer that sees round R retired must be able to conclude that nobody is still in a round <= R - L
            for (int q = round - cf.L; q > round - 2 * cf.L && q >= 0; --q)
              while ((int)w.region[rank]->slotFlushed[q % cf.L].load(std::memory_order_acquire) < q + 1) std::this_thread::yield();
            if (rank == (round % cf.P) && (rng() % 3 == 0)) std::this_thread::sleep_for(std::chrono::microseconds(300));  // straggler
            if (cf.lagger && rank == 0 && cta == 0) std::this_thread::sleep_for(std::chrono::microseconds(1500 + rng() % 2000));
            if (cf.nvls == 3) {
              phaseOneHop(w, rank, round, cta, cf.G, rng);
            } else if (cf.oneshot) {
              phaseOneshot(w, rank, round, cta, cf.G, rng);
            } else {
              phaseScatter(w, rank, round, cta, cf.G, rng);
              phaseReduce(w, rank, round, cta, cf.G, rng);
              phaseGather(w, rank, round, cta, cf.G, rng);
            }
            retire(w, rank, round, cf.G);
            b->wait();  // kernel boundary: the stream's next kernel starts when every CTA of this one is done
            if (cta == 0) {
              // host side: wait for the completion word, then consume this round's output slot before it is reused
              while ((int)w.region[rank]->slotFlushed[round % cf.L].load(std::memory_order_acquire) < round + 1) std::this_thread::yield();
              if (cf.nvls == 3) verifyOneHop(w, rank, round);
              else if (cf.oneshot) verifyOneshot(w, rank, round);
              else verifyRound(w, rank, round);
              w.region[rank]->consumed[round % cf.L].store(round, std::memory_order_release);
            }
            b->wait();
          }
        });
    }
  for (auto& t : threads) t.join();
  long bad = w.failures.load();
  if (c.nvls) printf("nvls=%d ", c.nvls);
  if (c.oneshot) printf("oneshot ");
  printf("protocol_litmus seed=%u P=%d rounds=%d th=%d%% L=%d overlap=%d G=%d fc=%d lagger=%d eagerDecision=%d: failures=%ld"
         " (drops=%ld forcedChunks=%ld forcedRounds=%ld abandoned=%ld elidedPushes=%ld elidedBcasts=%ld)\n", c.seed, c.P,
         c.rounds, thPct, c.L, (int)c.overlap, c.G, (int)c.bcastFlowControl, (int)c.lagger, (int)c.eagerDecision, bad, w.drops.load(),
         w.forcedChunks.load(), w.forcedRounds.load(), w.abandoned.load(), w.elidedPushes.load(), w.elidedBcasts.load());
  return bad == 0 ? 0 : 1;
}
