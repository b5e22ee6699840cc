// Race-detection harness for the host runtime (the reference has none; its master even mutates actor state from a
// Future callback, AllreduceMaster.scala:40-47).  One master and three CPU-data-plane workers, each with its own event
// loop thread, talk over real TCP sockets for 60 rounds.  Built with -fsanitize=thread by tests/test_tsan.py; any data
// race in the coordinator, the transport or the logger fails the run.
#include <atomic>
#include <cstdio>
#include <thread>
#include <vector>

#include "../core/tcp_transport.h"

using namespace aar;

int main() {
  const int P = 3, n = 24, rounds = 60;
  MasterConfig cfg;
  cfg.totalWorkers = P;
  cfg.dataSize = n;
  cfg.maxChunkSize = 4;
  cfg.maxRound = rounds - 1;
  cfg.maxLag = 2;
  cfg.pipelineDepth = 1;
  MasterNode master(cfg, "127.0.0.1", 0);
  const int port = master.port();
  std::atomic<long> flushed{0}, bad{0};
  std::vector<std::thread> threads;
  for (int w = 0; w < P; ++w) {
    threads.emplace_back([&, w] {
      WorkerNode node("127.0.0.1", port);
      DataSource src = [&](const AllReduceInputRequest& r) {
        AllReduceInput in;
        in.data.resize(n);
        for (int i = 0; i < n; ++i) in.data[i] = (float)i + (float)r.iteration;
        return in;
      };
      DataSink sink = [&](const AllReduceOutput& out) {
        ++flushed;
        for (int i = 0; i < n; ++i)
          if (out.data[i] != (float)P * ((float)i + (float)out.iteration)) ++bad;
      };
      Worker worker(src, sink, &node);
      node.setHandler([&](const Message& m) { worker.receive(m); });
      node.run(60.0);
    });
  }
  bool ok = master.run(60.0);
  for (auto& t : threads) t.join();
  printf("tsan_cluster: finished=%d flushed=%ld bad=%ld\n", ok ? 1 : 0, flushed.load(), bad.load());
  return (ok && flushed.load() == (long)P * rounds && bad.load() == 0) ? 0 : 1;
}
