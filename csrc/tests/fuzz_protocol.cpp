// Random message sequences against the C++ worker and master, meant to be built with -fsanitize=address,undefined.
//
// The Python-level tests only see crashes; this driver lets the sanitizers see out-of-bounds indexing, signed overflow and
// use-after-free in the protocol engine when messages carry hostile fields (ids, chunk numbers and rounds outside every
// range, empty or oversized payloads, re-initialisation in the middle of a round, completions for unknown rounds).
// Contract under test: a bad message is rejected with an exception or ignored -- never undefined behaviour.
//
//   fuzz_protocol [seed] [cases]
#include <cstdio>
#include <cstdlib>
#include <random>
#include <stdexcept>

#include "../core/master.h"
#include "../core/worker.h"

using namespace aar;

namespace {
struct Sink : WorkerPort {
  long sent = 0;
  void send(Address, Message) override { ++sent; }
};
}  // namespace

int main(int argc, char** argv) {
  const unsigned seed = argc > 1 ? (unsigned)atoi(argv[1]) : 1;
  const int cases = argc > 2 ? atoi(argv[2]) : 2000;
  std::mt19937 rng(seed);
  auto pick = [&](int lo, int hi) { return lo + (int)(rng() % (unsigned)(hi - lo + 1)); };
  const int rounds[] = {-2, -1, 0, 1, 2, 3, 7, 100, 3000, 2147483647};  // 3000: a long but finite catch-up
  long accepted = 0, rejected = 0, flushed = 0;
  for (int k = 0; k < cases; ++k) {
    const int P = pick(1, 5), n = pick(0, 24), chunk = pick(1, 6);
    Sink port;
    Worker w([n](const AllReduceInputRequest& r) {
               AllReduceInput in;
               in.data.resize((size_t)n);
               for (int i = 0; i < n; ++i) in.data[i] = (float)(i + r.iteration);
               return in;
             },
             [&](const AllReduceOutput&) { ++flushed; }, &port);
    MasterConfig mc;
    mc.totalWorkers = P;
    mc.dataSize = n;
    mc.maxChunkSize = chunk;
    mc.maxRound = pick(0, 20);
    mc.maxLag = pick(0, 3);
    mc.pipelineDepth = pick(0, 3);
    mc.thAllreduce = (float[]){1.f, 0.75f, 0.5f}[pick(0, 2)];
    Master master(mc, &port, 0);
    auto init = [&]() {
      InitWorkers i;
      for (int r = 0; r < P; ++r) i.workers[r] = 100;
      i.master = 100;
      i.destId = pick(0, P - 1);
      i.thReduce = (float[]){1.f, 0.75f, 0.5f}[pick(0, 2)];
      i.thComplete = (float[]){1.f, 0.75f, 0.5f}[pick(0, 2)];
      i.maxLag = pick(0, 3);
      i.dataSize = n;
      i.maxChunkSize = chunk;
      i.startRound = pick(0, 3) == 0 ? pick(0, 50) : 0;
      return i;
    };
    const int steps = pick(1, 40);
    for (int s = 0; s < steps; ++s) {
      try {
        switch (s == 0 ? 0 : pick(0, 7)) {
          case 0:
            w.receive(init());
            break;
          case 1:
            w.receive(StartAllreduce{rounds[pick(0, 9)], pick(0, 2)});
            break;
          case 2: {
            ScatterBlock b;
            b.value.assign((size_t)pick(0, 9), 1.f);
            b.srcId = pick(-1, P + 1);
            b.destId = pick(-1, P + 1);
            b.chunkId = pick(-1, 8);
            b.round = rounds[pick(0, 9)];
            b.generation = pick(0, 2);
            w.receive(b);
            break;
          }
          case 3: {
            ReduceBlock b;
            b.value.assign((size_t)pick(0, 9), 2.f);
            b.srcId = pick(-1, P + 1);
            b.destId = pick(-1, P + 1);
            b.chunkId = pick(-1, 8);
            b.round = rounds[pick(0, 9)];
            b.count = pick(-1, 70);
            b.generation = pick(0, 2);
            w.receive(b);
            break;
          }
          case 4:
            master.memberUp(pick(1, 8), pick(0, 3) ? "worker" : "master");
            break;
          case 5:
            master.terminated(pick(1, 8));
            break;
          case 6:
            master.receive(CompleteAllreduce{pick(-1, P + 1), rounds[pick(0, 9)], pick(0, 3)});
            break;
          default:
            master.receive(CompleteAllreduce{pick(0, P - 1), master.round(), master.generation()});
            break;
        }
        ++accepted;
      } catch (const std::exception&) {
        ++rejected;
      }
    }
  }
  printf("fuzz_protocol seed=%u cases=%d: accepted=%ld rejected=%ld flushed=%ld\n", seed, cases, accepted, rejected, flushed);
  return 0;
}
