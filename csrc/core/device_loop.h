// The device-plane worker's round loop in C++ (AllreduceWorker.scala:84-104, :253-268 on the host side of the GPU data plane):
//   StartAllreduce(r) -> doorbell, fetch + launch every not-yet-scattered round -> completion words -> sink + CompleteAllreduce,
// in whatever order the device flushes, with the peer-ahead self-start read from the shared board.  Python is called for the
// caller's source and sink only (and once per InitWorkers to build the data plane); the loop itself blocks in poll() on the
// control socket while nothing is in flight and runs without the interpreter lock otherwise.
//
// The host runtime stays free of CUDA: the data plane is a handle plus four C entry points (csrc/cuda/aar_engine.cu:
// aar_engine_enqueue_ex, aar_engine_set_doorbell, aar_engine_round_flushed, aar_engine_synchronize) handed over as addresses.
#pragma once

#include <cstdint>
#include <functional>
#include <map>
#include <mutex>
#include <set>

#include "messages.h"
#include "worker.h"

namespace aar {

class WorkerNode;

struct DevicePlane {
  using EnqueueFn = int (*)(void* engine, int round, const void* input, void* output, int* counts, unsigned extraFlags, void* callerStream,
                            unsigned long long delayNs);
  using DoorbellFn = int (*)(void* engine, int newestRound);
  using FlushedFn = int (*)(void* engine, int round);
  using SyncFn = int (*)(void* engine);
  void* engine = nullptr;
  EnqueueFn enqueue = nullptr;
  DoorbellFn doorbell = nullptr;
  FlushedFn flushed = nullptr;
  SyncFn synchronize = nullptr;
  int slots = 2;           // L = maxLag + 1 ring slots
  int maxLag = 1;
  long long dataSize = 0;  // elements per round, fixed at initialisation (AllreduceWorker.scala:174-176)
  unsigned extraFlags = 0;
  bool valid() const { return engine && enqueue && doorbell && flushed; }
};

// What the source hands back for one round: device pointers, the stream the input was produced on, and the element count.
struct RoundBuffers {
  uintptr_t input = 0, output = 0, counts = 0, stream = 0;
  long long numel = 0;
};

struct DeviceLoopStats {
  long long delivered = 0, outOfOrder = 0, selfStarts = 0, forced = 0, launched = 0;
  std::map<int, long long> lagHist;  // (newest round - round) at the moment a round reached the sink
};

class DeviceLoop {
 public:
  using Source = std::function<RoundBuffers(int round)>;
  using Sink = std::function<void(int round, int slot)>;
  using Init = std::function<void(const InitWorkers&)>;   // builds the data plane and calls setPlane()
  using BoardPost = std::function<void(int newest)>;
  using BoardNewest = std::function<int()>;

  DeviceLoop(Source source, Sink sink, WorkerPort* port) : source_(std::move(source)), sink_(std::move(sink)), port_(port) {}

  void setInit(Init fn) { init_ = std::move(fn); }
  void setBoard(BoardPost post, BoardNewest newest) {
    boardPost_ = std::move(post);
    boardNewest_ = std::move(newest);
  }
  // (Re)initialisation: wipes all round state (AllreduceWorker.scala:44 "clear round info to start over").
  void setPlane(const DevicePlane& plane, int id, int generation, Address master, int startRound);
  void clearPlane() { plane_ = DevicePlane(); }
  // Fault injection: a worker the master's StartAllreduce never reaches (it must then start its rounds because its peers have).
  void setDropStart(bool drop) { dropStart_ = drop; }

  void receive(const Message& msg);   // InitWorkers -> init callback; StartAllreduce -> onStart; everything else is not ours
  void onStart(int round);            // AllreduceWorker.scala:84-104
  bool peerAhead();                   // AllreduceWorker.scala:123-126 through the board; true if rounds were started
  int deliverFlushed(bool block);     // AllreduceWorker.scala:253-268; returns rounds delivered
  void drain();                       // job over: ring the doorbell far ahead, wait for the device, deliver what flushed
  bool busy() const { return plane_.valid() && stats_.delivered + startRound_ < (long long)maxScattered_ + 1; }   // loop thread only
  // Serve until the master says Shutdown / its link closes (true) or the timeout expires (false).
  bool run(WorkerNode& node, double timeoutSec);

  // Observers: safe from any thread while the loop runs (a metrics endpoint reads them from its own thread).
  struct Snapshot {
    int id, generation, round, maxRound, maxScattered, startRound;
    std::set<int> completed;
    DeviceLoopStats stats;
  };
  Snapshot snapshot() const {
    std::lock_guard<std::mutex> g(mu_);
    return Snapshot{id_, generation_, round_, maxRound_, maxScattered_, startRound_, completed_, stats_};
  }
  int id() const { return snapshot().id; }
  int generation() const { return snapshot().generation; }
  int round() const { return snapshot().round; }
  int maxRound() const { return snapshot().maxRound; }
  int maxScattered() const { return snapshot().maxScattered; }
  int startRound() const { return snapshot().startRound; }
  std::set<int> completed() const { return snapshot().completed; }
  DeviceLoopStats stats() const { return snapshot().stats; }

 private:
  Source source_;
  Sink sink_;
  Init init_;
  BoardPost boardPost_;
  BoardNewest boardNewest_;
  WorkerPort* port_;
  DevicePlane plane_;
  Address master_ = kNoAddress;
  bool dropStart_ = false;
  int id_ = -1, generation_ = 0, startRound_ = 0;
  int round_ = 0;           // oldest round that has not reached the sink (the reference's `round`)
  int maxRound_ = -1, maxScattered_ = -1;
  std::set<int> completed_;  // rounds >= round_ that have reached the sink (the reference's `completed`)
  DeviceLoopStats stats_;
  mutable std::mutex mu_;    // guards the round state and the counters against observers; never held across a callback
};

}  // namespace aar
