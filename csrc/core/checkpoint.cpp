// See checkpoint.h.  Format: "AARCKPT1" u32 kind {1 worker, 2 master, 3 ring}, then fields in the
// order written below; vectors are u64 length + raw little-endian elements.
#include "checkpoint.h"

#include <cstring>
#include <fstream>
#include <stdexcept>

namespace aar {
namespace {

constexpr char kMagic[8] = {'A', 'A', 'R', 'C', 'K', 'P', 'T', '1'};

struct Out {
  std::ofstream f;
  explicit Out(const std::string& path, uint32_t kind) : f(path, std::ios::binary | std::ios::trunc) {
    if (!f) throw std::runtime_error("checkpoint: cannot write " + path);
    f.write(kMagic, 8);
    pod(kind);
  }
  template <class T>
  void pod(const T& v) { f.write(reinterpret_cast<const char*>(&v), sizeof(T)); }
  template <class T>
  void vec(const std::vector<T>& v) {
    pod<uint64_t>(v.size());
    if (!v.empty()) f.write(reinterpret_cast<const char*>(v.data()), sizeof(T) * v.size());
  }
  void done() {
    f.flush();
    if (!f) throw std::runtime_error("checkpoint: write failed");
  }
};

struct In {
  std::ifstream f;
  uint64_t size = 0;
  In(const std::string& path, uint32_t kind) : f(path, std::ios::binary) {
    if (!f) throw std::runtime_error("checkpoint: cannot read " + path);
    f.seekg(0, std::ios::end);
    size = (uint64_t)f.tellg();
    f.seekg(0, std::ios::beg);
    char m[8];
    f.read(m, 8);
    if (!f || std::memcmp(m, kMagic, 8) != 0) throw std::runtime_error("checkpoint: bad magic in " + path);
    if (pod<uint32_t>() != kind) throw std::runtime_error("checkpoint: wrong section kind in " + path);
  }
  template <class T>
  T pod() {
    T v{};
    f.read(reinterpret_cast<char*>(&v), sizeof(T));
    if (!f) throw std::runtime_error("checkpoint: truncated file");
    return v;
  }
  template <class T>
  std::vector<T> vec() {
    uint64_t n = pod<uint64_t>();
    // a length that the rest of the file cannot hold is corruption: refuse before allocating for it
    uint64_t here = (uint64_t)f.tellg();
    if (n > (size - here) / sizeof(T)) throw std::runtime_error("checkpoint: vector length exceeds the file");
    std::vector<T> v(n);
    if (n) f.read(reinterpret_cast<char*>(v.data()), sizeof(T) * n);
    if (!f) throw std::runtime_error("checkpoint: truncated file");
    return v;
  }
};

void writeBuffer(Out& o, DataBuffer& b, int round) {
  o.pod<int32_t>(b.depth());
  o.pod<int32_t>(b.peers());
  o.pod<int32_t>(b.rowLen());
  o.pod<int32_t>(b.depth() ? round % b.depth() : 0);  // temporalOffset of the reference layout
  // countFilled[L][numChunks]: derived, written for readability of the snapshot next to DataBuffer.scala:27
  std::vector<int32_t> countFilled;
  for (int s = 0; s < b.depth(); ++s) {
    int r = b.rawSlotRound()[s];
    for (int c = 0; c < std::max(1, b.maxChunks()); ++c) countFilled.push_back(r >= 0 ? b.countChunk(r, c) : 0);
  }
  o.vec(countFilled);
  o.vec(b.rawSlotRound());
  o.vec(b.rawArrived());
  o.vec(b.rawAux());
  o.vec(b.rawData());
}

void readBuffer(In& in, DataBuffer& b) {
  int depth = in.pod<int32_t>(), peers = in.pod<int32_t>(), rowLen = in.pod<int32_t>();
  (void)in.pod<int32_t>();
  (void)in.vec<int32_t>();
  if (depth != b.depth() || peers != b.peers() || rowLen != b.rowLen()) throw std::runtime_error("checkpoint: ring shape mismatch");
  auto slotRound = in.vec<int>();
  auto arrived = in.vec<uint8_t>();
  auto aux = in.vec<int>();
  auto data = in.vec<float>();
  if (slotRound.size() != b.rawSlotRound().size() || arrived.size() != b.rawArrived().size() || aux.size() != b.rawAux().size() ||
      data.size() != b.rawData().size())
    throw std::runtime_error("checkpoint: ring payload mismatch");
  b.rawSlotRound() = std::move(slotRound);
  b.rawArrived() = std::move(arrived);
  b.rawAux() = std::move(aux);
  b.rawData() = std::move(data);
}

}  // namespace

void saveWorkerCheckpoint(const std::string& path, Worker& w) {
  if (!w.initialized()) throw std::runtime_error("checkpoint: worker is not initialised");
  Out o(path, 1);
  WorkerState s = w.state();
  o.pod<int32_t>(s.id);
  o.pod<int32_t>(s.round);
  o.pod<int32_t>(s.maxRound);
  o.pod<int32_t>(s.maxScattered);
  o.vec(s.completed);
  o.pod<float>(s.thReduce);
  o.pod<float>(s.thComplete);
  o.pod<int32_t>(s.maxLag);
  o.pod<int32_t>(s.dataSize);
  o.pod<int32_t>(s.maxChunkSize);
  o.pod<int32_t>(s.numPeers);
  o.pod<int32_t>(s.generation);
  std::vector<int64_t> peers;
  for (auto& kv : w.peers()) {
    peers.push_back(kv.first);
    peers.push_back(kv.second);
  }
  o.vec(peers);
  o.pod<int64_t>(w.masterAddress());
  writeBuffer(o, w.scatterBuffer(), s.round);
  writeBuffer(o, w.reduceBuffer(), s.round);
  o.done();
}

void loadWorkerCheckpoint(const std::string& path, Worker& w) {
  In in(path, 1);
  WorkerState s;
  s.id = in.pod<int32_t>();
  s.round = in.pod<int32_t>();
  s.maxRound = in.pod<int32_t>();
  s.maxScattered = in.pod<int32_t>();
  s.completed = in.vec<int>();
  s.thReduce = in.pod<float>();
  s.thComplete = in.pod<float>();
  s.maxLag = in.pod<int32_t>();
  s.dataSize = in.pod<int32_t>();
  s.maxChunkSize = in.pod<int32_t>();
  s.numPeers = in.pod<int32_t>();
  s.generation = in.pod<int32_t>();
  auto flat = in.vec<int64_t>();
  std::map<int, Address> peers;
  for (size_t i = 0; i + 1 < flat.size(); i += 2) peers[(int)flat[i]] = flat[i + 1];
  Address master = in.pod<int64_t>();
  // The rings follow in the file: L x P x block floats each.  Fields that imply rings larger than the file are corrupt;
  // refuse before restore() allocates for them.
  if (s.maxLag < 0 || s.dataSize < 0 || s.maxChunkSize <= 0 || s.numPeers <= 0 || (size_t)s.numPeers != peers.size() ||
      (uint64_t)(s.maxLag + 1) * (uint64_t)s.dataSize * sizeof(float) > in.size)
    throw std::runtime_error("checkpoint: worker fields are inconsistent with the file");
  // round bookkeeping as the state machine leaves it after any message: round <= maxScattered + 1 <= maxRound + 1, and the
  // catch-up loop keeps round within maxLag of maxRound.  A file that says otherwise would send the restored worker into a
  // catch-up of billions of rounds on its next StartAllreduce.
  if (s.round < 0 || s.maxScattered < -1 || s.maxRound < -1 || s.maxScattered > s.maxRound || s.round > s.maxScattered + 1 ||
      (long long)s.maxRound - s.round > (long long)s.maxLag + 1 || s.id < 0 || s.id >= s.numPeers)
    throw std::runtime_error("checkpoint: worker round state is inconsistent");
  for (int r : s.completed)
    if (r < s.round || r > s.maxRound) throw std::runtime_error("checkpoint: completed set is inconsistent with the round state");
  w.restore(s, peers, master);
  readBuffer(in, w.scatterBuffer());
  readBuffer(in, w.reduceBuffer());
}

void saveMasterCheckpoint(const std::string& path, const Master& m) {
  Out o(path, 2);
  const MasterConfig& c = m.config();
  o.pod<int32_t>(c.totalWorkers);
  o.pod<float>(c.thAllreduce);
  o.pod<float>(c.thReduce);
  o.pod<float>(c.thComplete);
  o.pod<int32_t>(c.maxLag);
  o.pod<int32_t>(c.dataSize);
  o.pod<int32_t>(c.maxRound);
  o.pod<int32_t>(c.maxChunkSize);
  o.pod<int32_t>(c.pipelineDepth);
  o.pod<int32_t>(m.round());
  o.pod<int32_t>(m.quorumRound());
  o.pod<int32_t>(m.generation());
  std::vector<int64_t> workers;
  for (auto& kv : m.workers()) {
    workers.push_back(kv.first);
    workers.push_back(kv.second);
  }
  o.vec(workers);
  std::vector<int32_t> completes;
  for (auto& kv : m.completes()) {
    completes.push_back(kv.first);
    completes.push_back(kv.second);
  }
  o.vec(completes);
  o.done();
}

int resumeMasterFromCheckpoint(const std::string& path, Master& m) {
  Master scratch(m.config(), nullptr);
  loadMasterCheckpoint(path, scratch);  // validates the geometry against m's configuration
  if (m.started()) throw std::runtime_error("checkpoint: resume must happen before the first admission");
  int start = scratch.quorumRound() + 1;
  m.resumeAt(start, scratch.generation());
  return start;
}

void loadMasterCheckpoint(const std::string& path, Master& m) {
  In in(path, 2);
  MasterConfig c;
  c.totalWorkers = in.pod<int32_t>();
  c.thAllreduce = in.pod<float>();
  c.thReduce = in.pod<float>();
  c.thComplete = in.pod<float>();
  c.maxLag = in.pod<int32_t>();
  c.dataSize = in.pod<int32_t>();
  c.maxRound = in.pod<int32_t>();
  c.maxChunkSize = in.pod<int32_t>();
  c.pipelineDepth = in.pod<int32_t>();
  const MasterConfig& have = m.config();
  if (c.totalWorkers != have.totalWorkers || c.dataSize != have.dataSize || c.maxChunkSize != have.maxChunkSize || c.maxLag != have.maxLag)
    throw std::runtime_error("checkpoint: master configuration differs from the snapshot");
  int round = in.pod<int32_t>(), quorum = in.pod<int32_t>(), gen = in.pod<int32_t>();
  auto flat = in.vec<int64_t>();
  std::map<int, Address> workers;
  for (size_t i = 0; i + 1 < flat.size(); i += 2) workers[(int)flat[i]] = flat[i + 1];
  auto cf = in.vec<int32_t>();
  std::map<int, int> completes;
  for (size_t i = 0; i + 1 < cf.size(); i += 2) completes[cf[i]] = cf[i + 1];
  m.restore(round, quorum, gen, workers, completes);
}

void saveRingSnapshot(const std::string& path, const RingSnapshot& s) {
  Out o(path, 3);
  o.pod<int32_t>(s.rank);
  o.pod<int32_t>(s.P);
  o.pod<int32_t>(s.L);
  o.pod<int32_t>(s.dataSize);
  o.pod<int32_t>(s.maxChunkSize);
  o.pod<int32_t>(s.round);
  o.vec(s.scatterRing);
  o.vec(s.reduceRing);
  o.vec(s.scatterFlags);
  o.vec(s.reduceFlags);
  o.done();
}

RingSnapshot loadRingSnapshot(const std::string& path) {
  In in(path, 3);
  RingSnapshot s;
  s.rank = in.pod<int32_t>();
  s.P = in.pod<int32_t>();
  s.L = in.pod<int32_t>();
  s.dataSize = in.pod<int32_t>();
  s.maxChunkSize = in.pod<int32_t>();
  s.round = in.pod<int32_t>();
  s.scatterRing = in.vec<float>();
  s.reduceRing = in.vec<float>();
  s.scatterFlags = in.vec<uint32_t>();
  s.reduceFlags = in.vec<uint64_t>();
  return s;
}

}  // namespace aar
