// Control- and data-plane records of the host protocol engine.
//
// Same five records as the reference wire protocol (AllreduceMessage.scala:7-20) plus the
// caller-facing trio (DataWrapper.scala:3-7).  On the CUDA data plane ScatterBlock and
// ReduceBlock never exist as messages (they are NVLink stores + release flags); they are
// materialised only by the CPU engine, which doubles as oracle and timed CPU baseline.
#pragma once
#include <stdexcept>
#include <cstdint>
#include <map>
#include <string>
#include <variant>
#include <vector>

namespace aar {

// A well-formed frame whose CONTENT violates the protocol and that came from a peer: ids or sizes outside the geometry, a round
// absurdly far ahead.  Its own type so that a transport can drop the offending link instead of letting one bogus frame end the
// process; errors that are this worker's own (the data source returning the wrong size, AllreduceWorker.scala:174-176) stay
// ordinary exceptions and propagate.
struct ProtocolError : std::runtime_error {
  using std::runtime_error::runtime_error;
};


// Opaque address of an actor inside a transport (index into its routing table).
using Address = int64_t;
constexpr Address kNoAddress = -1;

struct InitWorkers {
  std::map<int, Address> workers;  // dense id -> address, including the receiver itself
  Address master = kNoAddress;
  int destId = -1;
  float thReduce = 1.f;
  float thComplete = 1.f;
  int maxLag = 0;
  int dataSize = 0;
  int maxChunkSize = 1024;
  int generation = 0;  // (re)admission epoch; the reference has none (B7)
  int startRound = 0;  // first round of this incarnation: 0 like the reference, > 0 when the master resumed from a checkpoint
};

struct StartAllreduce {
  int round = 0;
  int generation = 0;  // 0 = unspecified (accepted by any incarnation)
};

struct ScatterBlock {
  std::vector<float> value;
  int srcId = 0, destId = 0, chunkId = 0, round = 0;
  int generation = 0;
};

struct ReduceBlock {
  std::vector<float> value;
  int srcId = 0, destId = 0, chunkId = 0, round = 0, count = 0;
  int generation = 0;
};

struct CompleteAllreduce {
  int srcId = 0;
  int round = 0;
  int generation = 0;
};

using Message = std::variant<InitWorkers, StartAllreduce, ScatterBlock, ReduceBlock, CompleteAllreduce>;

inline const char* messageName(const Message& m) {
  static const char* names[] = {"InitWorkers", "StartAllreduce", "ScatterBlock", "ReduceBlock", "CompleteAllreduce"};
  return names[m.index()];
}

// Caller API (DataWrapper.scala:3-7).  `count` is really filled in here: one entry per
// (owner, chunk) in flush order, 0 for chunks that never arrived.
struct AllReduceInputRequest {
  int iteration = 0;
};
struct AllReduceInput {
  std::vector<float> data;
};
struct AllReduceOutput {
  std::vector<float> data;
  std::vector<int> count;
  int iteration = 0;
};

}  // namespace aar
