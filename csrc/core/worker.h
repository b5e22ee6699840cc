// Per-rank round engine of the host (CPU) data plane.
//
// Behavioural parity with AllreduceWorker.scala:35-270 (state machine described in SURVEY.md
// section 2.4): InitWorkers -> StartAllreduce (catch-up, fetch, scatter) -> ScatterBlock
// (store, threshold, reduce, broadcast) -> ReduceBlock (store, round threshold, complete, flush).
// The engine is transport-agnostic: it talks to a `WorkerPort`, so the same code runs under the
// in-process router (oracle / CPU baseline / tests) and under the TCP transport (separate processes).
//
// Design differences from the reference actor (see SURVEY.md 2.4 "latent bugs"):
//   B1 chunks are cut by the destination block's length, B2 empty tail blocks are legal,
//   B3 the round threshold counts real chunks, B4 duplicates cannot overshoot a firing count,
//   B5 contributor counts reach the sink, B6 only arrived sources are summed (in peer order,
//   from 0.0f, so results are bit-identical to the reference's sum over zero-filled rows),
//   B10 messages that arrive before InitWorkers are queued instead of being re-sent in a spin.
#pragma once
#include <deque>
#include <functional>
#include <set>

#include "data_buffer.h"
#include "geometry.h"
#include "messages.h"
#include "metrics.h"

namespace aar {

// Outbound side of a worker: deliver `msg` to the actor at `to`.
class WorkerPort {
 public:
  virtual ~WorkerPort() = default;
  virtual void send(Address to, Message msg) = 0;
};

using DataSource = std::function<AllReduceInput(const AllReduceInputRequest&)>;
using DataSink = std::function<void(const AllReduceOutput&)>;

// Snapshot of every field the reference actor keeps (AllreduceWorker.scala:12-33), used by checkpoint.
struct WorkerState {
  int id = -1;
  float thReduce = 1.f, thComplete = 1.f;
  int maxLag = 0, round = -1, maxRound = -1, maxScattered = -1;
  std::vector<int> completed;
  int dataSize = 0, maxChunkSize = 1024, numPeers = 0, generation = 0;
};

class Worker {
 public:
  Worker(DataSource source, DataSink sink, WorkerPort* port) : source_(std::move(source)), sink_(std::move(sink)), port_(port) {}

  void receive(const Message& m);

  // Introspection (tests, checkpoint, metrics).
  int id() const { return id_; }
  int round() const { return round_; }
  int maxRound() const { return maxRound_; }
  int maxScattered() const { return maxScattered_; }
  const std::set<int>& completed() const { return completed_; }
  const Geometry& geometry() const { return geo_; }
  const WorkerMetrics& metrics() const { return metrics_; }
  bool initialized() const { return id_ >= 0; }
  const std::map<int, Address>& peers() const { return peers_; }
  Address masterAddress() const { return master_; }
  WorkerState state() const;
  DataBuffer& scatterBuffer() { return scatterBuf_; }
  DataBuffer& reduceBuffer() { return reduceBuf_; }
  // Restore bookkeeping saved by state(); buffers are restored separately through their raw accessors.
  void restore(const WorkerState& s, const std::map<int, Address>& peers, Address master);

 private:
  void onInit(const InitWorkers& m);
  void onStart(const StartAllreduce& m);
  void onScatter(const ScatterBlock& m);
  void onReduce(const ReduceBlock& m);

  void allocate();
  void fetch(int round);
  void scatter(int round);
  // Sum chunk `c` of my block over the sources that arrived for `round`; returns contributor count.
  int reduce(int round, int c, std::vector<float>& out);
  void broadcast(const std::vector<float>& reduced, int c, int round, int count);
  void complete(int round, bool forced);
  void flush(int round);

  DataSource source_;
  DataSink sink_;
  WorkerPort* port_;

  int id_ = -1;
  Address master_ = kNoAddress;
  std::map<int, Address> peers_;
  float thReduce_ = 1.f, thComplete_ = 1.f;
  int maxLag_ = 0;
  int round_ = -1, maxRound_ = -1, maxScattered_ = -1;
  int generation_ = 0;
  std::set<int> completed_;

  Geometry geo_;
  std::vector<float> data_;
  DataBuffer scatterBuf_, reduceBuf_;
  int minRequired_ = 1, minChunksRequired_ = 1;

  std::deque<Message> deferred_;  // messages received before InitWorkers
  WorkerMetrics metrics_;
};

}  // namespace aar
