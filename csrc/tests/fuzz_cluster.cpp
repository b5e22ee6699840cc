// Whole in-process clusters under AddressSanitizer / UBSan: random geometry (including uneven and empty blocks), thresholds, ring
// depth, pipeline depth, message loss, parked (straggling) workers and late joiners.
//
// Invariants checked on every flushed vector (the ones the reference's tests pin for one worker, here for all of them):
//   * no worker flushes a round twice; rounds below startRound are never flushed;
//   * every element is a sum of k inputs with 0 <= k <= P: since all workers use the source data[i] = i + round, the value
//     must be k * (i + round) for the k the count array reports for that element's chunk;
//   * with thresholds = 1, no loss and nobody parked, every round 0..maxRound is flushed by every worker with k = P.
//
//   fuzz_cluster [seed] [cases]
#include <cstdio>
#include <cstdlib>
#include <map>
#include <random>
#include <set>
#include <vector>

#include "../core/local_cluster.h"

using namespace aar;

int main(int argc, char** argv) {
  const unsigned seed = argc > 1 ? (unsigned)atoi(argv[1]) : 1;
  const int cases = argc > 2 ? atoi(argv[2]) : 300;
  std::mt19937 rng(seed);
  auto pick = [&](int lo, int hi) { return lo + (int)(rng() % (unsigned)(hi - lo + 1)); };
  const float ths[] = {1.f, 0.9f, 0.75f, 0.6f, 0.5f};
  long bad = 0, flushedTotal = 0, cleanCases = 0;
  long kind[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // twice, below start, size, count range, clean count, value, missing round, not finished
  for (int k = 0; k < cases; ++k) {
    MasterConfig cfg;
    const int P = pick(1, 6);
    cfg.totalWorkers = P;
    cfg.dataSize = pick(0, 40);
    cfg.maxChunkSize = pick(1, 9);
    cfg.maxLag = pick(0, 4);
    cfg.maxRound = pick(0, 12);
    cfg.pipelineDepth = pick(0, 3);  // the master caps it at maxLag (running further ahead would push workers out of their window)
    cfg.startRound = pick(0, 4) == 0 ? pick(1, 6) : 0;
    const bool clean = pick(0, 2) == 0;  // full thresholds, no faults: the strong invariant applies
    cfg.thAllreduce = clean ? 1.f : ths[pick(0, 4)];
    cfg.thReduce = clean ? 1.f : ths[pick(0, 4)];
    cfg.thComplete = clean ? 1.f : ths[pick(0, 4)];
    std::vector<std::set<std::pair<int, int>>> seen((size_t)P);  // (generation, round) per worker
    LocalCluster* clp = nullptr;
    const int n = cfg.dataSize, start = cfg.startRound;
    auto makeSource = [n](int) {
      return [n](const AllReduceInputRequest& r) {
        AllReduceInput in;
        in.data.resize((size_t)n);
        for (int i = 0; i < n; ++i) in.data[(size_t)i] = (float)(i + r.iteration);
        return in;
      };
    };
    auto makeSink = [&](int w) {
      return [&, w](const AllReduceOutput& o) {
        ++flushedTotal;
        // a late joiner makes the master re-initialise everyone: fewer peers before, round numbering restarts after
        const WorkerState st = clp->worker(w).state();
        const int P = st.numPeers;
        Geometry geo(P, n, st.maxChunkSize);
        if (!seen[(size_t)w].insert({st.generation, o.iteration}).second) ++bad, ++kind[0];  // flushed twice
        if (o.iteration < start) ++bad, ++kind[1];
        if ((int)o.data.size() != n) {
          ++bad, ++kind[2];
          return;
        }
        // element -> (owner, chunk) -> count
        for (int owner = 0, ci = 0; owner < P; ++owner)
          for (int c = 0; c < geo.numChunks(owner); ++c, ++ci) {
            const int cnt = ci < (int)o.count.size() ? o.count[(size_t)ci] : -1;
            if (cnt < 0 || cnt > P) ++bad, ++kind[3];
            if (clean && cnt != P) ++bad, ++kind[4];
            const int s = geo.blockStart(owner) + c * geo.maxChunkSize, len = geo.chunkLen(owner, c);
            for (int i = s; i < s + len; ++i)
              if (o.data[(size_t)i] != (float)cnt * (float)(i + o.iteration)) ++bad, ++kind[5];
          }
      };
    };
    LocalCluster cl(cfg, P, makeSource, makeSink);
    clp = &cl;
    if (!clean) {
      const unsigned lossPermille = (unsigned)pick(0, 150);
      cl.setDrop([&rng, lossPermille](Address, Address to, const Message& m) {
        // control messages from / to the master are reliable here (the TCP transport does not lose them either)
        if (to == LocalCluster::kMasterAddr || std::holds_alternative<InitWorkers>(m) || std::holds_alternative<StartAllreduce>(m)) return false;
        return rng() % 1000 < lossPermille;
      });
    }
    const int late = (!clean && P > 1 && pick(0, 3) == 0) ? pick(1, P - 1) : -1;  // this worker joins after the others have started
    for (int i = 0; i < P; ++i)
      if (i != late) cl.join(i);
    int steps = 0;
    while (cl.pending() > 0 && steps++ < 4000) {
      cl.run(pick(1, 50));
      if (!clean && P > 1) {
        const int w = pick(0, P - 1);
        const int action = pick(0, 9);
        if (action == 0) cl.hold(w);
        else if (action <= 2) cl.release(w);
        else if (action == 3 && late >= 0 && steps > 3) cl.join(late);
      }
    }
    for (int i = 0; i < P; ++i) cl.release(i);
    cl.run();
    if (clean) {
      ++cleanCases;
      for (int w = 0; w < P; ++w)
        for (int r = start; r <= cfg.maxRound; ++r)
          if (!seen[(size_t)w].count({1, r})) ++bad, ++kind[6];
      if (start <= cfg.maxRound && !cl.master().finished()) {
        ++bad, ++kind[7];
        fprintf(stderr, "stuck: P=%d dataSize=%d chunk=%d maxLag=%d maxRound=%d depth=%d start=%d (master round %d, quorum %d)\n", P, cfg.dataSize,
                cfg.maxChunkSize, cfg.maxLag, cfg.maxRound, cfg.pipelineDepth, start, cl.master().round(), cl.master().quorumRound());
      }
    }
  }
  printf("fuzz_cluster seed=%u cases=%d (%ld clean): flushed=%ld violations=%ld [twice %ld, below start %ld, size %ld, count range %ld, clean count %ld, "
         "value %ld, missing round %ld, not finished %ld]\n",
         seed, cases, cleanCases, flushedTotal, bad, kind[0], kind[1], kind[2], kind[3], kind[4], kind[5], kind[6], kind[7]);
  return bad == 0 ? 0 : 1;
}
