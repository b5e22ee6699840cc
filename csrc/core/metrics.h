// Structured counters exported by the coordinator and the workers.
// The reference enables akka-cluster-metrics + Sigar (application.conf:26-34, build.sbt:20,26)
// but never reads them; what an allreduce operator actually needs is per-round accounting:
// rounds/s, contributor counts, lag, forced completions, stale drops.
#pragma once
#include <array>
#include <cstdint>
#include <map>
#include <sstream>
#include <string>

namespace aar {

struct WorkerMetrics {
  int64_t roundsCompleted = 0;
  int64_t roundsForced = 0;        // completed by the catch-up path (AllreduceWorker.scala:91-97)
  int64_t roundsOutOfOrder = 0;    // completed while an older round was still open
  int64_t scatterSent = 0, scatterReceived = 0;
  int64_t reduceSent = 0, reduceReceived = 0;
  int64_t staleDrops = 0;          // messages for finished rounds
  int64_t duplicateDrops = 0;      // repeated (src, chunk, round)
  int64_t selfStarts = 0;          // peer-ahead StartAllreduce synthesised locally
  int64_t deferredBeforeInit = 0;  // messages queued until InitWorkers (B10)
  int64_t chunksReduced = 0;
  int64_t contributorsSum = 0;     // sum over reduced chunks of their contributor count
  int64_t maxLagSeen = 0;          // max(maxRound - round) observed
  std::array<int64_t, 65> contributorHist{};  // histogram of contributor counts per reduced chunk

  std::map<std::string, double> asMap() const {
    std::map<std::string, double> m;
    m["rounds_completed"] = (double)roundsCompleted;
    m["rounds_forced"] = (double)roundsForced;
    m["rounds_out_of_order"] = (double)roundsOutOfOrder;
    m["scatter_sent"] = (double)scatterSent;
    m["scatter_received"] = (double)scatterReceived;
    m["reduce_sent"] = (double)reduceSent;
    m["reduce_received"] = (double)reduceReceived;
    m["stale_drops"] = (double)staleDrops;
    m["duplicate_drops"] = (double)duplicateDrops;
    m["self_starts"] = (double)selfStarts;
    m["deferred_before_init"] = (double)deferredBeforeInit;
    m["chunks_reduced"] = (double)chunksReduced;
    m["mean_contributors"] = chunksReduced ? (double)contributorsSum / (double)chunksReduced : 0.0;
    m["max_lag_seen"] = (double)maxLagSeen;
    return m;
  }
};

struct MasterMetrics {
  int64_t membersUp = 0;
  int64_t workersAdmitted = 0;
  int64_t workersTerminated = 0;
  int64_t initBroadcasts = 0;   // number of (re)initialisations (late joiners re-init everyone)
  int64_t roundsStarted = 0;
  int64_t completionsCounted = 0;
  int64_t completionsIgnored = 0;  // wrong round or stale generation
  double firstStartSec = 0, lastStartSec = 0;

  std::map<std::string, double> asMap() const {
    std::map<std::string, double> m;
    m["members_up"] = (double)membersUp;
    m["workers_admitted"] = (double)workersAdmitted;
    m["workers_terminated"] = (double)workersTerminated;
    m["init_broadcasts"] = (double)initBroadcasts;
    m["rounds_started"] = (double)roundsStarted;
    m["completions_counted"] = (double)completionsCounted;
    m["completions_ignored"] = (double)completionsIgnored;
    double dt = lastStartSec - firstStartSec;
    m["rounds_per_sec"] = (roundsStarted > 1 && dt > 0) ? (double)(roundsStarted - 1) / dt : 0.0;
    return m;
  }
};

}  // namespace aar
