// CLI entry point of one CPU-data-plane worker (the GPU worker is `python -m akka_allreduce_b200.cli.worker`).
//
//   aar_worker [port=2553] [dataSize=10] [--conf application.conf] [--set k=v]... [--timeout seconds] [--check]
//
// Same positional arguments as `sbt "runMain sample.cluster.allreduce.AllreduceWorker ..."`
// (AllreduceWorker.scala:272-302, README.md:7): synthetic source data[i] = i + iteration, sink logs the
// iteration.  `--check` additionally verifies every flushed element whose contributor count is c against
// c * (i + iteration), which holds because all workers use the same source.
#include <cmath>
#include <cstdio>

#include "../core/tcp_transport.h"
#include "cli_common.h"

using namespace aar;

int main(int argc, char** argv) {
  try {
    bool check = false;
    std::vector<char*> rest;
    for (int i = 0; i < argc; ++i) {
      if (std::string(argv[i]) == "--check")
        check = true;
      else
        rest.push_back(argv[i]);
    }
    CliArgs args = parseCli((int)rest.size(), rest.data());
    if (args.help) {
      puts("usage: aar_worker [port=2553] [dataSize=10] [--conf application.conf] [--set k=v]... [--timeout seconds] [--check]\n"
           "CPU data plane, source data[i] = i + iteration.  The master is found at allreduce.transport.{hostname,port}.\n"
           "GPU workers: python -m akka_allreduce_b200.cli.worker");
      return 0;
    }
    Config conf = loadLayered(args, "worker");
    AllreduceSettings s = AllreduceSettings::fromConfig(conf);
    Logger::instance().setLevel(Logger::parse(s.loglevel));
    int port = args.positional.size() > 0 ? atoi(args.positional[0].c_str()) : 2553;
    int dataSize = args.positional.size() > 1 ? atoi(args.positional[1].c_str()) : 10;

    WorkerNode node(s.hostname, s.port, s.hostname, port, s.heartbeatMs);
    long flushed = 0, bad = 0;
    Worker* self = nullptr;
    DataSource source = [dataSize](const AllReduceInputRequest& req) {
      AllReduceInput in;
      in.data.resize(dataSize);
      for (int i = 0; i < dataSize; ++i) in.data[i] = (float)i + (float)req.iteration;
      AAR_DEBUG("Data source at #%d (%d floats)", req.iteration, dataSize);
      return in;
    };
    DataSink sink = [&](const AllReduceOutput& out) {
      ++flushed;
      AAR_INFO("----Data output at #%d (%zu floats, %zu chunk counts)", out.iteration, out.data.size(), out.count.size());
      if (check && self) {
        const Geometry& g = self->geometry();
        int ci = 0;
        for (int p = 0; p < g.P; ++p)
          for (int c = 0; c < g.numChunks(p); ++c, ++ci)
            for (int j = 0; j < g.chunkLen(p, c); ++j) {
              int i = g.blockStart(p) + c * g.maxChunkSize + j;
              float want = (float)out.count[ci] * ((float)i + (float)out.iteration);
              if (out.data[i] != want) ++bad;
            }
      }
    };
    Worker worker(source, sink, &node);
    self = &worker;
    node.setHandler([&](const Message& m) { worker.receive(m); });
    bool ok = node.run(args.timeoutSec);
    // shutdown=1: the master said so; master_lost=1: its link closed without a Shutdown (crash, network) -- a different exit code,
    // so that a supervisor can tell "job done" from "coordinator gone"
    const bool lost = ok && !node.shutdownByMaster();
    printf("worker %d: shutdown=%d master_lost=%d rounds_flushed=%ld mismatches=%ld\n", worker.id(), ok ? 1 : 0, lost ? 1 : 0, flushed, bad);
    if (lost) return 3;
    return (ok && bad == 0) ? 0 : 2;
  } catch (const std::exception& e) {
    fprintf(stderr, "aar_worker: %s\n", e.what());
    return 1;
  }
}
