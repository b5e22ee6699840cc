// See local_cluster.h.
#include "local_cluster.h"

#include <chrono>

namespace aar {

double cpuBaselineSeconds(const MasterConfig& cfg, int numWorkers, double* checksum, int64_t* messages) {
  double sum = 0;
  auto makeSource = [&](int) -> DataSource {
    int n = cfg.dataSize;
    return [n](const AllReduceInputRequest& req) {
      AllReduceInput in;
      in.data.resize(n);
      for (int i = 0; i < n; ++i) in.data[i] = (float)i + (float)req.iteration;
      return in;
    };
  };
  auto makeSink = [&](int w) -> DataSink {
    if (w != 0) return [](const AllReduceOutput&) {};
    return [&sum](const AllReduceOutput& out) {
      double s = 0;
      for (float v : out.data) s += v;
      sum = s;
    };
  };
  LocalCluster cluster(cfg, numWorkers, makeSource, makeSink);
  auto t0 = std::chrono::steady_clock::now();
  cluster.joinAll();
  cluster.run();
  auto t1 = std::chrono::steady_clock::now();
  if (checksum) *checksum = sum;
  if (messages) *messages = cluster.delivered();
  return std::chrono::duration<double>(t1 - t0).count();
}

}  // namespace aar
