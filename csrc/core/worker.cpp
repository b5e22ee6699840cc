// See worker.h.  Reference call sites are cited per function.
#include "worker.h"

#include <stdexcept>
#include <string>

#include "logging.h"

namespace aar {

void Worker::receive(const Message& m) {
  if (std::holds_alternative<InitWorkers>(m)) {
    onInit(std::get<InitWorkers>(m));
    return;
  }
  if (!initialized()) {
    // Reference: `self ! msg` busy loop (AllreduceWorker.scala:86-88,108-110,131-133); we park it.
    deferred_.push_back(m);
    metrics_.deferredBeforeInit++;
    return;
  }
  // Traffic of a previous incarnation (the master re-initialised everyone, AllreduceMaster.scala:40-47) is
  // dropped; the reference has no such tag and would mix it into the new round 0.
  int gen = 0;
  switch (m.index()) {
    case 1: gen = std::get<StartAllreduce>(m).generation; break;
    case 2: gen = std::get<ScatterBlock>(m).generation; break;
    case 3: gen = std::get<ReduceBlock>(m).generation; break;
    default: break;
  }
  if (gen != 0 && generation_ != 0 && gen != generation_) {
    metrics_.staleDrops++;
    return;
  }
  switch (m.index()) {
    case 1: onStart(std::get<StartAllreduce>(m)); break;
    case 2: onScatter(std::get<ScatterBlock>(m)); break;
    case 3: onReduce(std::get<ReduceBlock>(m)); break;
    default: break;  // CompleteAllreduce is for the master
  }
}

// AllreduceWorker.scala:37-82 -- (re)initialisation wipes all round state.
void Worker::onInit(const InitWorkers& m) {
  if (m.workers.find(m.destId) == m.workers.end()) throw std::invalid_argument("InitWorkers: destId not in worker map");
  if (m.maxLag < 0 || m.maxLag > (1 << 20)) throw std::invalid_argument("InitWorkers: maxLag out of range");
  id_ = m.destId;
  master_ = m.master;
  peers_ = m.workers;
  thReduce_ = m.thReduce;
  thComplete_ = m.thComplete;
  maxLag_ = m.maxLag;
  generation_ = m.generation;
  if (m.startRound < 0) throw std::invalid_argument("InitWorkers: startRound must be >= 0");
  round_ = m.startRound;  // 0 in the reference; a resumed master continues its round numbering
  maxRound_ = m.startRound - 1;
  maxScattered_ = m.startRound - 1;
  completed_.clear();
  geo_ = Geometry((int)peers_.size(), m.dataSize, m.maxChunkSize);
  for (int i = 0; i < geo_.P; ++i)
    if (!peers_.count(i)) throw std::invalid_argument("InitWorkers: worker ids must be dense 0..P-1");
  allocate();
  AAR_INFO("worker %d: init P=%d dataSize=%d maxChunk=%d thReduce=%.3f thComplete=%.3f maxLag=%d gen=%d", id_, geo_.P,
           geo_.dataSize, geo_.maxChunkSize, thReduce_, thComplete_, maxLag_, generation_);
  std::deque<Message> pending;
  pending.swap(deferred_);
  for (auto& p : pending) receive(p);
}

// AllreduceWorker.scala:52-73 -- two rings of maxLag+1 slots.
void Worker::allocate() {
  data_.assign(geo_.dataSize, 0.f);
  int P = geo_.P;
  std::vector<int> mine(P, geo_.numChunks(id_));
  std::vector<int> perOwner(P);
  for (int i = 0; i < P; ++i) perOwner[i] = geo_.numChunks(i);
  scatterBuf_ = DataBuffer(geo_.blockLen(id_), P, maxLag_ + 1, geo_.maxChunkSize, mine);
  reduceBuf_ = DataBuffer(geo_.maxBlockLen(), P, maxLag_ + 1, geo_.maxChunkSize, perOwner);
  minRequired_ = Geometry::minRequired(thReduce_, P);
  minChunksRequired_ = geo_.minChunksRequired(thComplete_);
}

// AllreduceWorker.scala:84-104
void Worker::onStart(const StartAllreduce& m) {
  // The catch-up below fetches and scatters EVERY round up to m.round (AllreduceWorker.scala:98-102); a round number that
  // is absurdly far ahead (corrupt frame, wrong master) would keep this worker busy for hours.  The master is never more
  // than pipelineDepth + 1 rounds ahead of a quorum, and a resumed master sends its start round in InitWorkers.
  constexpr int kMaxCatchUp = 1 << 20;
  if (m.round < 0 || (long long)m.round - maxScattered_ > kMaxCatchUp)
    throw std::invalid_argument("StartAllreduce: round " + std::to_string(m.round) + " is out of range (last scattered " +
                                std::to_string(maxScattered_) + ")");
  maxRound_ = std::max(maxRound_, m.round);
  metrics_.maxLagSeen = std::max<int64_t>(metrics_.maxLagSeen, maxRound_ - round_);
  // Fell out of the maxLag window: force the oldest rounds through with whatever has arrived.
  std::vector<float> reduced;
  while (round_ < maxRound_ - maxLag_) {
    int r = round_;
    for (int k = 0; k < geo_.numChunks(id_); ++k) {
      int cnt = reduce(r, k, reduced);
      broadcast(reduced, k, r, cnt);
    }
    complete(r, /*forced=*/true);
  }
  // Every round number is fetched and scattered exactly once, in order.
  while (maxScattered_ < maxRound_) {
    const int r = maxScattered_ + 1;
    fetch(r);
    scatter(r);
    maxScattered_ += 1;
    // An empty vector has no chunk whose arrival could complete the round: it is complete as soon as it is scattered
    // (found by csrc/tests/fuzz_cluster.cpp: dataSize = 0 used to wait forever).
    if (geo_.totalChunks() == 0 && r >= round_ && !completed_.count(r)) complete(r, /*forced=*/false);
  }
  for (auto it = completed_.begin(); it != completed_.end();) it = (*it < round_) ? completed_.erase(it) : std::next(it);
}

// How far ahead of this worker's newest round a peer's block may be: the master is never more than maxLag + 1 rounds ahead of a
// quorum, so a peer cannot legitimately be further ahead than a few windows.
static constexpr long long kMaxPeerAhead = 1 << 16;

// AllreduceWorker.scala:106-127
void Worker::onScatter(const ScatterBlock& s) {
  // Everything a peer controls is checked BEFORE the frame can make this worker do anything (the self-start below fetches and
  // scatters every round up to s.round): a bogus frame costs its sender the link, not this worker hours of catch-up.
  if (s.destId != id_) throw ProtocolError("ScatterBlock delivered to the wrong worker");
  if (s.srcId < 0 || s.srcId >= geo_.P || s.chunkId < 0 || s.chunkId >= geo_.numChunks(id_)) throw ProtocolError("ScatterBlock: bad srcId/chunkId");
  if ((int)s.value.size() > geo_.chunkLen(id_, s.chunkId)) throw ProtocolError("ScatterBlock: chunk larger than its slot");
  if (s.round < 0 || (long long)s.round - maxRound_ > kMaxPeerAhead) throw ProtocolError("ScatterBlock: round " + std::to_string(s.round) + " is out of range");
  metrics_.scatterReceived++;
  if (s.round < round_ || completed_.count(s.round)) {
    metrics_.staleDrops++;
    return;
  }
  if (s.round > maxRound_) {  // a peer is ahead of the master's message to us: self-start
    metrics_.selfStarts++;
    onStart(StartAllreduce{s.round, 0});
    if (s.round < round_ || completed_.count(s.round)) {
      metrics_.staleDrops++;
      return;
    }
  }
  bool fresh = scatterBuf_.store(s.round, s.srcId, s.chunkId, s.value.data(), (int)s.value.size());
  if (!fresh) {
    metrics_.duplicateDrops++;
    return;
  }
  if (scatterBuf_.countChunk(s.round, s.chunkId) == minRequired_) {  // fires exactly once
    std::vector<float> reduced;
    int cnt = reduce(s.round, s.chunkId, reduced);
    broadcast(reduced, s.chunkId, s.round, cnt);
  }
}

// AllreduceWorker.scala:129-150
void Worker::onReduce(const ReduceBlock& r) {
  if ((int)r.value.size() > geo_.maxChunkSize)
    throw ProtocolError("Reduced block of size " + std::to_string(r.value.size()) + " is larger than maxChunkSize " + std::to_string(geo_.maxChunkSize));
  if (r.destId != id_) throw ProtocolError("ReduceBlock delivered to the wrong worker");
  if (r.srcId < 0 || r.srcId >= geo_.P || r.chunkId < 0 || r.chunkId >= geo_.numChunks(r.srcId)) throw ProtocolError("ReduceBlock: bad srcId/chunkId");
  if ((int)r.value.size() > geo_.chunkLen(r.srcId, r.chunkId)) throw ProtocolError("ReduceBlock: chunk larger than its slot");
  if (r.round < 0 || (long long)r.round - maxRound_ > kMaxPeerAhead) throw ProtocolError("ReduceBlock: round " + std::to_string(r.round) + " is out of range");
  metrics_.reduceReceived++;
  if (r.round < round_ || completed_.count(r.round)) {
    metrics_.staleDrops++;
    return;
  }
  if (r.round > maxRound_) {
    metrics_.selfStarts++;
    onStart(StartAllreduce{r.round, 0});
    if (r.round < round_ || completed_.count(r.round)) {
      metrics_.staleDrops++;
      return;
    }
  }
  bool fresh = reduceBuf_.store(r.round, r.srcId, r.chunkId, r.value.data(), (int)r.value.size(), r.count);
  if (!fresh) {
    metrics_.duplicateDrops++;
    return;
  }
  if (reduceBuf_.countRound(r.round) == minChunksRequired_) complete(r.round, /*forced=*/false);
}

// AllreduceWorker.scala:171-178
void Worker::fetch(int round) {
  AllReduceInput in = source_(AllReduceInputRequest{round});
  if ((int)in.data.size() != geo_.dataSize)
    throw std::invalid_argument("Input data size " + std::to_string(in.data.size()) + " is different from initialization time " +
                                std::to_string(geo_.dataSize) + "!");
  data_ = std::move(in.data);
}

// AllreduceWorker.scala:194-209 -- staggered start at self; chunked by the destination block (B1).
void Worker::scatter(int round) {
  int P = geo_.P;
  for (int i = 0; i < P; ++i) {
    int idx = (i + id_) % P;
    Address to = peers_.at(idx);
    const float* block = data_.data() + geo_.blockStart(idx);
    int nc = geo_.numChunks(idx);
    for (int c = 0; c < nc; ++c) {
      int len = geo_.chunkLen(idx, c);
      ScatterBlock s;
      s.value.assign(block + (size_t)c * geo_.maxChunkSize, block + (size_t)c * geo_.maxChunkSize + len);
      s.srcId = id_;
      s.destId = idx;
      s.chunkId = c;
      s.round = round;
      s.generation = generation_;
      metrics_.scatterSent++;
      port_->send(to, std::move(s));
    }
  }
}

// AllreduceWorker.scala:240-251 -- sequential fp32 sum in peer-index order starting from 0.0f.
int Worker::reduce(int round, int c, std::vector<float>& out) {
  int len = geo_.chunkLen(id_, c);
  out.assign(len, 0.f);
  uint64_t mask = scatterBuf_.maskChunk(round, c);
  int count = 0;
  for (int p = 0; p < geo_.P; ++p) {
    if (!(mask >> p & 1)) continue;
    ++count;
    const float* src = scatterBuf_.row(round, p) + (size_t)c * geo_.maxChunkSize;
    for (int j = 0; j < len; ++j) out[j] += src[j];
  }
  metrics_.chunksReduced++;
  metrics_.contributorsSum += count;
  metrics_.contributorHist[std::min(count, 64)]++;
  return count;
}

// AllreduceWorker.scala:230-238
void Worker::broadcast(const std::vector<float>& reduced, int c, int round, int count) {
  int P = geo_.P;
  for (int i = 0; i < P; ++i) {
    int idx = (i + id_) % P;
    ReduceBlock r;
    r.value = reduced;
    r.srcId = id_;
    r.destId = idx;
    r.chunkId = c;
    r.round = round;
    r.count = count;
    r.generation = generation_;
    metrics_.reduceSent++;
    port_->send(peers_.at(idx), std::move(r));
  }
}

// AllreduceWorker.scala:253-268
void Worker::complete(int r, bool forced) {
  flush(r);
  data_.clear();
  port_->send(master_, CompleteAllreduce{id_, r, generation_});
  completed_.insert(r);
  metrics_.roundsCompleted++;
  if (forced) metrics_.roundsForced++;
  if (round_ == r) {
    do {
      scatterBuf_.retire(round_);
      reduceBuf_.retire(round_);
      round_ += 1;
    } while (completed_.count(round_));
  } else {
    metrics_.roundsOutOfOrder++;
  }
}

// AllreduceWorker.scala:180-192 -- owner-major concatenation truncated to dataSize; chunks that never
// arrived stay zero.  Unlike the reference (count = Array(0)) the per-chunk contributor counts are delivered.
void Worker::flush(int r) {
  AllReduceOutput out;
  out.data.assign(geo_.dataSize, 0.f);
  out.count.assign(geo_.totalChunks(), 0);
  out.iteration = r;
  int ci = 0;
  for (int p = 0; p < geo_.P; ++p) {
    for (int c = 0; c < geo_.numChunks(p); ++c, ++ci) {
      if (!reduceBuf_.has(r, p, c)) continue;
      int len = geo_.chunkLen(p, c);
      const float* src = reduceBuf_.row(r, p) + (size_t)c * geo_.maxChunkSize;
      std::copy(src, src + len, out.data.begin() + geo_.blockStart(p) + (size_t)c * geo_.maxChunkSize);
      out.count[ci] = reduceBuf_.aux(r, p, c);
    }
  }
  sink_(out);
}

WorkerState Worker::state() const {
  WorkerState s;
  s.id = id_;
  s.thReduce = thReduce_;
  s.thComplete = thComplete_;
  s.maxLag = maxLag_;
  s.round = round_;
  s.maxRound = maxRound_;
  s.maxScattered = maxScattered_;
  s.completed.assign(completed_.begin(), completed_.end());
  s.dataSize = geo_.dataSize;
  s.maxChunkSize = geo_.maxChunkSize;
  s.numPeers = geo_.P;
  s.generation = generation_;
  return s;
}

void Worker::restore(const WorkerState& s, const std::map<int, Address>& peers, Address master) {
  InitWorkers init;
  init.workers = peers;
  init.master = master;
  init.destId = s.id;
  init.thReduce = s.thReduce;
  init.thComplete = s.thComplete;
  init.maxLag = s.maxLag;
  init.dataSize = s.dataSize;
  init.maxChunkSize = s.maxChunkSize;
  init.generation = s.generation;
  onInit(init);
  round_ = s.round;
  maxRound_ = s.maxRound;
  maxScattered_ = s.maxScattered;
  completed_ = std::set<int>(s.completed.begin(), s.completed.end());
}

}  // namespace aar
