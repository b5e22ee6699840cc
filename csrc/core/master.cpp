// See master.h.
#include "master.h"

#include <algorithm>
#include <ctime>

#include "logging.h"

namespace aar {

static double nowSec() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec + ts.tv_nsec * 1e-9;
}

// AllreduceMaster.scala:38-47,70-82
void Master::memberUp(Address addr, const std::string& role) {
  metrics_.membersUp++;
  if (role == "worker") {
    for (auto& kv : workers_)
      if (kv.second == addr) return;  // already registered
    int idx = workers_.empty() ? 0 : workers_.rbegin()->first + 1;
    workers_[idx] = addr;
    metrics_.workersAdmitted++;
    AAR_INFO("master: worker %d registered (%zu of %d)", idx, workers_.size(), cfg_.totalWorkers);
  }
  maybeAdmit();
}

void Master::maybeAdmit() {
  // (totalWorkers * thAllreduce).toInt -- float multiply then truncation (AllreduceMaster.scala:42).
  int need = (int)((float)cfg_.totalWorkers * cfg_.thAllreduce);
  if ((int)workers_.size() >= std::max(1, need)) {
    initWorkers();
    round_ = cfg_.startRound - 1;
    quorumRound_ = cfg_.startRound - 1;
    completes_.clear();
    completedBy_.clear();
    advance();
  }
}

// AllreduceMaster.scala:84-89
void Master::initWorkers() {
  // Re-densify ids (join order preserved) so that peers.get((i + id) % P) is total.
  std::map<int, Address> dense;
  int k = 0;
  for (auto& kv : workers_) dense[k++] = kv.second;
  workers_.swap(dense);
  generation_++;
  metrics_.initBroadcasts++;
  for (auto& kv : workers_) {
    InitWorkers m;
    m.workers = workers_;
    m.master = self_;
    m.destId = kv.first;
    m.thReduce = cfg_.thReduce;
    m.thComplete = cfg_.thComplete;
    m.maxLag = cfg_.maxLag;
    m.dataSize = cfg_.dataSize;
    m.maxChunkSize = cfg_.maxChunkSize;
    m.generation = generation_;
    m.startRound = cfg_.startRound;
    port_->send(kv.second, std::move(m));
  }
}

// AllreduceMaster.scala:91-97
void Master::startAllreduce(int round) {
  double t = nowSec();
  if (metrics_.roundsStarted == 0) metrics_.firstStartSec = t;
  metrics_.lastStartSec = t;
  metrics_.roundsStarted++;
  AAR_DEBUG("master: start allreduce round %d", round);
  for (auto& kv : workers_) port_->send(kv.second, StartAllreduce{round, generation_});
}

void Master::advance() {
  const float quorum = (float)cfg_.totalWorkers * cfg_.thAllreduce;
  while (quorumRound_ < round_ && (float)numCompleteOf(quorumRound_ + 1) >= quorum) {
    completes_.erase(quorumRound_ + 1);
    completedBy_.erase(quorumRound_ + 1);
    quorumRound_++;
  }
  // Reference clock: start round r+1 when round r reached quorum and r < maxRound.  With a pipeline
  // depth d the master may be d rounds ahead of the newest quorum.
  // The depth is capped at maxLag: a master further ahead than the workers' window would make them force-complete rounds
  // that nothing is wrong with (found by csrc/tests/fuzz_cluster.cpp: thresholds 1.0, no faults, partial sums).
  const int depth = std::min(cfg_.pipelineDepth, cfg_.maxLag);
  while (round_ < cfg_.maxRound && round_ < quorumRound_ + 1 + depth) {
    round_++;
    startAllreduce(round_);
  }
}

void Master::receive(const Message& m) {
  if (std::holds_alternative<CompleteAllreduce>(m)) onComplete(std::get<CompleteAllreduce>(m));
}

// AllreduceMaster.scala:58-67
void Master::onComplete(const CompleteAllreduce& c) {
  bool counted = false;
  if (c.generation == generation_ && c.round > quorumRound_ && c.round <= round_) {
    auto& by = completedBy_[c.round];
    if (std::find(by.begin(), by.end(), c.srcId) == by.end()) {
      by.push_back(c.srcId);
      completes_[c.round]++;
      counted = true;
    }
  }
  if (counted) {
    metrics_.completionsCounted++;
    advance();
  } else {
    metrics_.completionsIgnored++;
  }
}

// AllreduceMaster.scala:50-56
void Master::terminated(Address addr) {
  for (auto it = workers_.begin(); it != workers_.end();) {
    if (it->second == addr) {
      AAR_INFO("master: worker %d terminated, removing it from the set", it->first);
      it = workers_.erase(it);
      metrics_.workersTerminated++;
    } else {
      ++it;
    }
  }
}

void Master::restore(int round, int quorumRound, int generation, const std::map<int, Address>& workers,
                     const std::map<int, int>& completes) {
  round_ = round;
  quorumRound_ = quorumRound;
  generation_ = generation;
  workers_ = workers;
  completes_ = completes;
  completedBy_.clear();
}

}  // namespace aar
