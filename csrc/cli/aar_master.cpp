// CLI entry point of the coordinator.
//
//   aar_master [port=2551] [totalWorkers=2] [dataSize=totalWorkers*5] [maxChunkSize=2]
//              [--conf application.conf] [--set allreduce.th-reduce=1.0]... [--timeout seconds]
//              [--checkpoint file [--checkpoint-every rounds] [--resume]]
//
// Same positional arguments as `sbt "runMain sample.cluster.allreduce.AllreduceMaster ..."`
// (AllreduceMaster.scala:101-138, README.md:3).  The thresholds / maxLag / maxRound the reference
// hard-codes at :105-109 come from the `allreduce { }` block of the config file, with the same defaults.
#include <cstdio>

#include "../core/checkpoint.h"
#include "../core/tcp_transport.h"
#include "cli_common.h"

using namespace aar;

int main(int argc, char** argv) {
  try {
    CliArgs args = parseCli(argc, argv);
    if (args.help) {
      puts("usage: aar_master [port=2551] [totalWorkers=2] [dataSize=totalWorkers*5] [maxChunkSize=2]\n"
           "                  [--conf application.conf] [--set allreduce.th-reduce=1.0]... [--timeout seconds]\n"
           "                  [--checkpoint file [--checkpoint-every rounds=100] [--resume]]\n"
           "Thresholds, maxLag and maxRound come from the allreduce { } block of the config file.");
      return 0;
    }
    Config conf = loadLayered(args, "master");
    // reference defaults when there is no file at all
    Config defaults = Config::parseString(
        "allreduce { th-allreduce = 1.0\n th-reduce = 0.9\n th-complete = 0.8\n max-lag = 1\n max-round = 100\n total-workers = 2\n max-chunk-size = 2 }");
    conf = conf.withFallback(defaults);
    auto& p = args.positional;
    if (p.size() > 0) conf.set("allreduce.transport.port", p[0]);
    if (p.size() > 1) conf.set("allreduce.total-workers", p[1]);
    if (p.size() > 2) conf.set("allreduce.data-size", p[2]);
    if (p.size() > 3) conf.set("allreduce.max-chunk-size", p[3]);
    AllreduceSettings s = AllreduceSettings::fromConfig(conf);
    Logger::instance().setLevel(Logger::parse(s.loglevel));
    AAR_INFO("-------\n Port = %d \n Number of Workers = %d \n Message Size = %d \n Max Chunk Size = %d", s.port, s.master.totalWorkers,
             s.master.dataSize, s.master.maxChunkSize);
    MasterNode node(s.master, s.hostname, s.port, s.autoDownAfterMs);
    if (!args.checkpoint.empty()) {
      bool exists = false;
      if (FILE* f = fopen(args.checkpoint.c_str(), "rb")) {
        fclose(f);
        exists = true;
      }
      if (args.resume && exists) {
        int start = resumeMasterFromCheckpoint(args.checkpoint, node.master());
        printf("master: resuming at round %d from %s\n", start, args.checkpoint.c_str());
      }
      node.setCheckpoint(args.checkpoint, args.checkpointEvery);
    }
    bool ok = node.run(args.timeoutSec);
    auto m = node.master().metrics().asMap();
    printf("master: finished=%d rounds_started=%.0f rounds_per_sec=%.1f\n", ok ? 1 : 0, m["rounds_started"], m["rounds_per_sec"]);
    return ok ? 0 : 2;
  } catch (const std::exception& e) {
    fprintf(stderr, "aar_master: %s\n", e.what());
    return 1;
  }
}
