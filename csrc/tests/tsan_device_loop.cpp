// The device-plane round loop (csrc/core/device_loop.cpp) end to end without Python and without a GPU, for the sanitizers:
// one master and three workers over real TCP sockets; each worker's data plane is a C function table over an in-process
// "fabric" whose completion words are written by a separate device thread per rank (as the GPU writes host-mapped words),
// in an order that lets later rounds flush before earlier ones.  One worker never hears StartAllreduce (drop-start fault
// injection) and must follow its peers through the board; an observer thread per worker reads the loop's state while it runs
// (what a metrics endpoint does).  Built with -fsanitize=thread / address,undefined by
// tests/test_tsan.py and tests/test_sanitizers.py.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <mutex>
#include <thread>
#include <vector>

#include "../core/device_loop.h"
#include "../core/tcp_transport.h"

using namespace aar;

namespace {

constexpr int P = 3, kRounds = 80, kSlots = 3, kMaxLag = 2, kData = 64;

struct Fabric {
  std::mutex mu;
  std::vector<int> enqueuedBy = std::vector<int>(kRounds, 0);   // ranks that have launched round r
  std::atomic<int> board[P];
  Fabric() {
    for (auto& b : board) b.store(-1);
  }
};

struct RankPlane {
  Fabric* fabric = nullptr;
  int rank = 0;
  std::atomic<unsigned> done[kSlots];       // host-mapped completion words: round + 1 of the newest round flushed in the slot
  std::atomic<int> doorbell{-1};
  std::atomic<int> launched{-1};            // newest round this rank has launched
  std::atomic<bool> stop{false};
  std::atomic<long> badArgs{0};
  RankPlane() {
    for (auto& d : done) d.store(0);
  }
};

int fabricEnqueue(void* e, int round, const void* in, void* out, int* counts, unsigned flags, void* stream, unsigned long long delay) {
  auto* p = static_cast<RankPlane*>(e);
  if (in != reinterpret_cast<const void*>(uintptr_t(0x1000 + round)) || out == nullptr || counts == nullptr || flags != 0 || delay != 0 ||
      stream != reinterpret_cast<void*>(uintptr_t(0x77)))
    ++p->badArgs;
  {
    std::lock_guard<std::mutex> g(p->fabric->mu);
    ++p->fabric->enqueuedBy[(size_t)round];
  }
  p->launched.store(round, std::memory_order_release);
  return 0;
}
int fabricDoorbell(void* e, int newest) {
  auto* p = static_cast<RankPlane*>(e);
  int cur = p->doorbell.load();
  while (newest > cur && !p->doorbell.compare_exchange_weak(cur, newest)) {
  }
  return 0;
}
int fabricFlushed(void* e, int round) {
  auto* p = static_cast<RankPlane*>(e);
  return p->done[round % kSlots].load(std::memory_order_acquire) >= (unsigned)(round + 1);
}
int fabricSync(void*) { return 0; }

// The "GPU" of one rank: a launched round whose quorum (all ranks) has launched it gets flushed -- the newest eligible round
// first, so rounds complete out of order whenever two are eligible at once.
void deviceThread(RankPlane* p) {
  std::vector<char> flushed((size_t)kRounds, 0);
  while (!p->stop.load()) {
    const int top = p->launched.load(std::memory_order_acquire);
    bool any = false;
    for (int r = top; r >= 0 && r > top - kSlots; --r) {
      if (flushed[(size_t)r]) continue;
      bool quorum;
      {
        std::lock_guard<std::mutex> g(p->fabric->mu);
        quorum = p->fabric->enqueuedBy[(size_t)r] == P;
      }
      const bool forced = p->doorbell.load() - r > kMaxLag;   // catch-up: complete with what has arrived
      if (!quorum && !forced) continue;
      p->done[r % kSlots].store((unsigned)(r + 1), std::memory_order_release);
      flushed[(size_t)r] = 1;
      any = true;
    }
    if (!any) std::this_thread::sleep_for(std::chrono::microseconds(50));
  }
}

}  // namespace

int main() {
  MasterConfig cfg;
  cfg.totalWorkers = P;
  cfg.dataSize = kData;
  cfg.maxChunkSize = 16;
  cfg.maxRound = kRounds - 1;
  cfg.maxLag = kMaxLag;
  cfg.pipelineDepth = kMaxLag;
  MasterNode master(cfg, "127.0.0.1", 0);
  const int port = master.port();
  Fabric fabric;
  std::atomic<long> sunk{0}, outOfOrder{0}, selfStarts{0}, bad{0}, duplicates{0};
  std::vector<std::thread> threads;
  for (int w = 0; w < P; ++w) {
    threads.emplace_back([&, w] {
      WorkerNode node("127.0.0.1", port);
      RankPlane plane;
      plane.fabric = &fabric;
      std::vector<char> seen((size_t)kRounds, 0);
      int fetchedNext = 0;
      DeviceLoop loop(
          [&](int round) {
            if (round != fetchedNext++) ++bad;   // every round fetched exactly once, in order
            RoundBuffers b;
            b.input = uintptr_t(0x1000 + round);
            b.output = 0x2000;
            b.counts = 0x3000;
            b.stream = 0x77;
            b.numel = kData;
            return b;
          },
          [&](int round, int slot) {
            if (slot != round % kSlots) ++bad;
            if (seen[(size_t)round]++) ++duplicates;
            ++sunk;
          },
          &node);
      std::thread device;
      loop.setInit([&](const InitWorkers& m) {
        plane.rank = m.destId;
        DevicePlane dp;
        dp.engine = &plane;
        dp.enqueue = fabricEnqueue;
        dp.doorbell = fabricDoorbell;
        dp.flushed = fabricFlushed;
        dp.synchronize = fabricSync;
        dp.slots = kSlots;
        dp.maxLag = kMaxLag;
        dp.dataSize = m.dataSize;
        loop.setPlane(dp, m.destId, m.generation, m.master, m.startRound);
        loop.setDropStart(m.destId == P - 1);   // the last rank never hears the master's clock
        const int me = m.destId;
        loop.setBoard([&fabric, me](int newest) {
                        int cur = fabric.board[me].load();
                        while (newest > cur && !fabric.board[me].compare_exchange_weak(cur, newest)) {
                        }
                      },
                      [&fabric] {
                        int best = -1;
                        for (auto& b : fabric.board) best = std::max(best, b.load());
                        return best;
                      });
        if (!device.joinable()) device = std::thread(deviceThread, &plane);
      });
      // a metrics endpoint reading the loop's state from its own thread while the loop runs
      std::atomic<bool> watching{true};
      std::atomic<long> observed{0};
      std::thread observer([&] {
        while (watching.load()) {
          const DeviceLoop::Snapshot s = loop.snapshot();
          if (s.round > s.maxScattered + 1 || (long long)s.completed.size() > kSlots) ++bad;
          observed += s.stats.delivered >= 0;
          std::this_thread::sleep_for(std::chrono::microseconds(20));
        }
      });
      loop.run(node, 60.0);
      watching.store(false);
      observer.join();
      plane.stop.store(true);
      if (device.joinable()) device.join();
      outOfOrder += loop.stats().outOfOrder;
      selfStarts += loop.stats().selfStarts;
      bad += plane.badArgs.load();
      if (loop.stats().delivered != kRounds) ++bad;
    });
  }
  const bool ok = master.run(60.0);
  for (auto& t : threads) t.join();
  printf("tsan_device_loop: finished=%d sunk=%ld duplicates=%ld bad=%ld out_of_order=%ld self_starts=%ld\n", ok ? 1 : 0, sunk.load(),
         duplicates.load(), bad.load(), outOfOrder.load(), selfStarts.load());
  return (ok && sunk.load() == (long)P * kRounds && duplicates.load() == 0 && bad.load() == 0 && selfStarts.load() >= kRounds) ? 0 : 1;   // the deaf rank started every round itself; a hearing rank may beat the master's message too
}
