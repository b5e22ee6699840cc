// Coordinator: worker admission, dense id assignment, round clock, completion quorum.
//
// Behavioural parity with AllreduceMaster.scala:15-98:
//   * a member that comes up with role "worker" is registered in join order (`:70-79`);
//   * once workers >= floor(totalWorkers * thAllreduce) every worker gets InitWorkers and round 0
//     starts (`:40-47`); a later joiner re-initialises everyone and restarts at round 0;
//   * CompleteAllreduce is counted only for the current round; the round advances when
//     numComplete >= totalWorkers * thAllreduce (float compare, no truncation) and round < maxRound (`:58-67`);
//   * a terminated worker is removed from the map (`:50-56`).
// Differences by design: single-threaded event handling (the reference mutates actor state from a
// Future callback, B7); a generation number tags every (re)initialisation so completions from a
// previous incarnation are ignored; ids are re-densified on re-initialisation (the reference leaves
// holes that crash scatter()); an optional pipelined round clock (`pipelineDepth`) lets the master
// run ahead of the quorum by up to that many rounds so the control-plane latency hides behind the
// device-resident maxLag ring.  pipelineDepth = 0 is the reference's clock.
#pragma once
#include <map>
#include <string>
#include <vector>

#include "messages.h"
#include "metrics.h"
#include "worker.h"

namespace aar {

struct MasterConfig {
  int totalWorkers = 2;
  float thAllreduce = 1.f;
  float thReduce = 1.f;
  float thComplete = 1.f;
  int maxLag = 1;
  int dataSize = 10;
  int maxRound = 100;
  int maxChunkSize = 2;
  int pipelineDepth = 0;
  int startRound = 0;  // resume: rounds below this are taken as done; an admission (re)starts here instead of at 0
};

class Master {
 public:
  Master(MasterConfig cfg, WorkerPort* port, Address self = kNoAddress) : cfg_(cfg), port_(port), self_(self) {}

  // Cluster event: a member joined.  Non-worker roles are ignored for registration but, like the
  // reference, still re-evaluate the admission predicate.
  void memberUp(Address addr, const std::string& role);
  void terminated(Address addr);
  void receive(const Message& m);  // only CompleteAllreduce is meaningful
  void onComplete(const CompleteAllreduce& c);

  int round() const { return round_; }
  int numComplete() const { return numCompleteOf(round_); }
  int generation() const { return generation_; }
  bool started() const { return round_ >= 0; }
  bool finished() const { return started() && quorumRound_ >= cfg_.maxRound; }
  const std::map<int, Address>& workers() const { return workers_; }
  const MasterConfig& config() const { return cfg_; }
  const MasterMetrics& metrics() const { return metrics_; }
  int quorumRound() const { return quorumRound_; }

  // checkpoint
  void restore(int round, int quorumRound, int generation, const std::map<int, Address>& workers, const std::map<int, int>& completes);
  const std::map<int, int>& completes() const { return completes_; }
  // Continue another coordinator's numbering (checkpoint.h, resumeMasterFromCheckpoint): before the first admission only.
  void resumeAt(int startRound, int generation) {
    cfg_.startRound = startRound;
    generation_ = generation;
  }

 private:
  void initWorkers();
  void startAllreduce(int round);
  void maybeAdmit();
  void advance();
  int numCompleteOf(int r) const {
    auto it = completes_.find(r);
    return it == completes_.end() ? 0 : it->second;
  }

  MasterConfig cfg_;
  WorkerPort* port_;
  Address self_;
  std::map<int, Address> workers_;
  int round_ = -1;        // highest round started (reference: `round`)
  int quorumRound_ = -1;  // highest round r such that every round <= r reached its completion quorum
  int generation_ = 0;
  std::map<int, int> completes_;  // round -> completions counted
  std::map<int, std::vector<int>> completedBy_;
  MasterMetrics metrics_;
};

}  // namespace aar
