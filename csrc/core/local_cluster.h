// In-process cluster: one Master and N Workers wired through a deterministic FIFO router.
//
// This is the "CPU restatement" of the reference named in BASELINE.md section 2: it runs the same
// scatter / store / reduce / broadcast / flush message flow (AllreduceWorker.scala:194-268) with the
// same chunking and thresholds, single-threaded, and serves three purposes:
//   1. oracle for the CUDA data plane (bit-exact sums over the reported contributor set),
//   2. the timed CPU baseline (BASELINE.json config 1 and the CPU column of the sweep),
//   3. multi-rank protocol tests the reference does not have (it only tests one worker against a
//      fake cluster, AllreduceSpec.scala:757-763).
// Akka guarantees FIFO per sender/receiver pair; a single global FIFO is a legal schedule of that.
// Fault injection: a worker can be `hold`-ed (its mailbox is parked, i.e. a straggler) and released,
// and an arbitrary drop predicate can discard messages in flight.
#pragma once
#include <deque>
#include <functional>
#include <memory>

#include "master.h"
#include "worker.h"

namespace aar {

class LocalCluster {
 public:
  static constexpr Address kMasterAddr = 0;
  static Address workerAddr(int i) { return 1 + i; }

  using DropFn = std::function<bool(Address from, Address to, const Message&)>;

  LocalCluster(MasterConfig cfg, int numWorkers, std::function<DataSource(int)> makeSource, std::function<DataSink(int)> makeSink)
      : masterPort_(this, kMasterAddr), master_(cfg, &masterPort_, kMasterAddr) {
    for (int i = 0; i < numWorkers; ++i) {
      ports_.push_back(std::make_unique<Port>(this, workerAddr(i)));
      workers_.push_back(std::make_unique<Worker>(makeSource(i), makeSink(i), ports_.back().get()));
      held_.push_back(false);
      parked_.emplace_back();
    }
  }

  // Members come up in index order (join order decides the worker id, AllreduceMaster.scala:75).
  void joinAll() {
    for (size_t i = 0; i < workers_.size(); ++i) master_.memberUp(workerAddr((int)i), "worker");
  }
  void join(int i) { master_.memberUp(workerAddr(i), "worker"); }

  // Deliver queued messages until the queue drains (or the budget is spent). Returns messages delivered.
  int64_t run(int64_t maxMessages = -1) {
    int64_t n = 0;
    while (!queue_.empty() && (maxMessages < 0 || n < maxMessages)) {
      Envelope e = std::move(queue_.front());
      queue_.pop_front();
      if (drop_ && drop_(e.from, e.to, e.msg)) {
        ++dropped_;
        continue;
      }
      deliver(std::move(e));
      ++n;
    }
    delivered_ += n;
    return n;
  }

  void hold(int worker) { held_[worker] = true; }
  void release(int worker) {
    held_[worker] = false;
    // Parked messages keep their relative order and go to the back of the global queue.
    for (auto& e : parked_[worker]) queue_.push_back(std::move(e));
    parked_[worker].clear();
  }
  void setDrop(DropFn f) { drop_ = std::move(f); }
  void terminate(int worker) { master_.terminated(workerAddr(worker)); }

  Master& master() { return master_; }
  Worker& worker(int i) { return *workers_[i]; }
  int numWorkers() const { return (int)workers_.size(); }
  int64_t delivered() const { return delivered_; }
  int64_t dropped() const { return dropped_; }
  size_t pending() const { return queue_.size(); }

 private:
  struct Envelope {
    Address from, to;
    Message msg;
  };
  struct Port : WorkerPort {
    LocalCluster* c;
    Address self;
    Port(LocalCluster* cl, Address s) : c(cl), self(s) {}
    void send(Address to, Message msg) override { c->queue_.push_back(Envelope{self, to, std::move(msg)}); }
  };

  void deliver(Envelope e) {
    if (e.to == kMasterAddr) {
      master_.receive(e.msg);
      return;
    }
    int w = (int)(e.to - 1);
    if (w < 0 || w >= (int)workers_.size()) return;  // dead letter
    if (held_[w]) {
      parked_[w].push_back(std::move(e));
      return;
    }
    workers_[w]->receive(e.msg);
  }

  Port masterPort_;
  Master master_;
  std::vector<std::unique_ptr<Port>> ports_;
  std::vector<std::unique_ptr<Worker>> workers_;
  std::vector<bool> held_;
  std::vector<std::deque<Envelope>> parked_;
  std::deque<Envelope> queue_;
  DropFn drop_;
  int64_t delivered_ = 0, dropped_ = 0;
};

// Timed CPU baseline: P workers, synthetic source data[i] = i + iteration (AllreduceWorker.scala:285-291),
// `rounds` rounds through the in-process router. Returns wall seconds; `checksum` receives the sum of
// the last flushed vector of worker 0 so the work cannot be optimised away.
double cpuBaselineSeconds(const MasterConfig& cfg, int numWorkers, double* checksum, int64_t* messages);

}  // namespace aar
