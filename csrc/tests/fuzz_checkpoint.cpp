// Damaged checkpoint files against the loaders, meant to be built with -fsanitize=address,undefined: a valid worker, master and
// ring snapshot is written, then flipped, truncated and spliced at random before it is loaded again.  Contract: a damaged file is
// refused with an exception or -- when the damage only hits payload floats -- loaded; never a crash, an out-of-bounds access or
// an allocation sized by a corrupt length field.
//
//   fuzz_checkpoint [seed] [cases] [scratch dir]
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <random>
#include <string>
#include <vector>

#include "../core/checkpoint.h"

using namespace aar;

namespace {
struct Null : WorkerPort {
  void send(Address, Message) override {}
};
std::string slurp(const std::string& p) {
  std::ifstream f(p, std::ios::binary);
  return std::string((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
}
void spit(const std::string& p, const std::string& s) {
  std::ofstream f(p, std::ios::binary | std::ios::trunc);
  f.write(s.data(), (std::streamsize)s.size());
}
std::string damage(std::string s, std::mt19937& rng) {
  if (s.empty()) return s;
  switch (rng() % 3) {
    case 0:
      for (int k = 0, n = 1 + (int)(rng() % 8); k < n; ++k) s[rng() % s.size()] = (char)rng();
      break;
    case 1:
      s.resize(rng() % s.size());
      break;
    default: {
      size_t at = rng() % s.size();
      for (int k = 0; k < 8 && at + k < s.size(); ++k) s[at + k] = (char)rng();
    }
  }
  return s;
}
}  // namespace

int main(int argc, char** argv) {
  const unsigned seed = argc > 1 ? (unsigned)atoi(argv[1]) : 1;
  const int cases = argc > 2 ? atoi(argv[2]) : 500;
  const std::string dir = argc > 3 ? argv[3] : ".";
  std::mt19937 rng(seed);
  Null port;
  auto source = [](const AllReduceInputRequest& r) {
    AllReduceInput in;
    in.data.assign(8, (float)r.iteration);
    return in;
  };
  auto freshWorker = [&]() { return Worker(source, [](const AllReduceOutput&) {}, &port); };

  // valid files
  Worker w = freshWorker();
  InitWorkers init;
  for (int r = 0; r < 4; ++r) init.workers[r] = 100;
  init.master = 100;
  init.destId = 1;
  init.thReduce = init.thComplete = 0.75f;
  init.maxLag = 2;
  init.dataSize = 8;
  init.maxChunkSize = 2;
  w.receive(init);
  w.receive(StartAllreduce{0, 0});
  ScatterBlock sb;
  sb.value = {1.f, 2.f};
  sb.srcId = 0;
  sb.destId = 1;
  w.receive(sb);
  const std::string wp = dir + "/w.ckpt", mp = dir + "/m.ckpt", rp = dir + "/r.ckpt", bad = dir + "/bad.ckpt";
  saveWorkerCheckpoint(wp, w);
  MasterConfig mc;
  mc.totalWorkers = 2;
  Master m(mc, &port, 0);
  m.memberUp(5, "worker");
  m.memberUp(6, "worker");
  m.receive(CompleteAllreduce{0, 0, 1});
  saveMasterCheckpoint(mp, m);
  RingSnapshot rs;
  rs.rank = 1;
  rs.P = 2;
  rs.L = 2;
  rs.dataSize = 8;
  rs.maxChunkSize = 2;
  rs.round = 7;
  rs.scatterRing = {1.f, 2.f, 3.f};
  rs.reduceRing = {4.f};
  rs.scatterFlags = {1, 2, 3};
  rs.reduceFlags = {5, 6};
  saveRingSnapshot(rp, rs);
  const std::string good[3] = {slurp(wp), slurp(mp), slurp(rp)};

  long refused = 0, loaded = 0;
  for (int k = 0; k < cases; ++k) {
    const int which = (int)(rng() % 3);
    spit(bad, damage(good[which], rng));
    try {
      if (which == 0) {
        Worker fresh = freshWorker();
        loadWorkerCheckpoint(bad, fresh);
        fresh.receive(StartAllreduce{fresh.state().maxRound + 1, 0});  // the restored worker must be usable
      } else if (which == 1) {
        Master fresh(mc, &port, 0);
        loadMasterCheckpoint(bad, fresh);
        Master resumed(mc, &port, 0);
        resumeMasterFromCheckpoint(bad, resumed);
      } else {
        (void)loadRingSnapshot(bad);
      }
      ++loaded;
    } catch (const std::exception&) {
      ++refused;
    }
  }
  printf("fuzz_checkpoint seed=%u cases=%d: refused=%ld loaded=%ld\n", seed, cases, refused, loaded);
  return 0;
}
