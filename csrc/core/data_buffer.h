// Host-memory ring of in-flight round slots (the CPU engine's counterpart of the
// device-resident ring managed by the CUDA kernels).
//
// Capability parity with buffer/DataBuffer.scala:3-83: `depth` temporal slots, each holding
// one row of `rowLen` floats per peer and an arrival counter per chunk; two threshold
// predicates; retiring the oldest slot.  Differences by design:
//   * a slot is addressed by the absolute round (slot = round % depth) instead
//     of a rotating offset, which is the layout the device ring and the checkpoint use;
//   * arrivals are tracked as a per-chunk bitmask of sources, so duplicates cannot overshoot
//     the firing count (B4) and the reducer can sum exactly the arrived set (B6);
//   * retiring a slot clears counters only; data is never bulk-zeroed (absent sources are
//     skipped by mask, the flush zero-fills chunks that never arrived).
#pragma once
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <vector>

namespace aar {

class DataBuffer {
 public:
  DataBuffer() = default;
  // rowLen: floats per peer row; peers: rows per slot; depth: maxLag + 1; chunksOf[p]: number of
  // chunks of row p (the scatter buffer has the same count for all rows, the reduce buffer one per owner).
  DataBuffer(int rowLen, int peers, int depth, int maxChunkSize, std::vector<int> chunksPerRow)
      : rowLen_(rowLen), peers_(peers), depth_(depth), maxChunk_(maxChunkSize), chunksPerRow_(std::move(chunksPerRow)) {
    if (depth_ <= 0 || peers_ <= 0) throw std::invalid_argument("DataBuffer: bad shape");
    if (peers_ > 64) throw std::invalid_argument("DataBuffer: at most 64 peers");
    maxChunks_ = 0;
    for (int c : chunksPerRow_) maxChunks_ = std::max(maxChunks_, c);
    data_.assign((size_t)depth_ * peers_ * rowLen_, 0.f);
    arrived_.assign((size_t)depth_ * peers_ * std::max(1, maxChunks_), 0);
    aux_.assign(arrived_.size(), 0);
    slotRound_.assign(depth_, -1);
  }

  int depth() const { return depth_; }
  int peers() const { return peers_; }
  int rowLen() const { return rowLen_; }
  int maxChunks() const { return maxChunks_; }
  int slotOf(int round) const { return ((round % depth_) + depth_) % depth_; }

  // Bind a slot to a round (first touch after the previous tenant was retired).
  void open(int round) {
    int s = slotOf(round);
    if (slotRound_[s] == round) return;
    clearSlot(s);
    slotRound_[s] = round;
  }
  void retire(int round) {
    int s = slotOf(round);
    if (slotRound_[s] == round) {
      clearSlot(s);
      slotRound_[s] = -1;
    }
  }

  // Store one chunk from `row` (source for the scatter buffer, owner for the reduce buffer).
  // Returns true when this is the first arrival of (row, chunk) for that round.
  bool store(int round, int row, int chunk, const float* v, int n, int aux = 0) {
    open(round);
    if (row < 0 || row >= peers_ || chunk < 0 || chunk >= std::max(1, maxChunks_)) throw std::out_of_range("DataBuffer::store index");
    if ((int64_t)chunk * maxChunk_ + n > rowLen_) throw std::out_of_range("DataBuffer::store overflows the row");
    int s = slotOf(round);
    if (n > 0) std::memcpy(rowPtr(s, row) + (size_t)chunk * maxChunk_, v, sizeof(float) * (size_t)n);  // an empty payload may come with a null pointer
    uint8_t& a = arrived_[flagIndex(s, row, chunk)];
    aux_[flagIndex(s, row, chunk)] = aux;
    bool fresh = !a;
    a = 1;
    return fresh;
  }

  bool has(int round, int row, int chunk) const {
    int s = slotOf(round);
    return slotRound_[s] == round && arrived_[flagIndex(s, row, chunk)];
  }
  int aux(int round, int row, int chunk) const { return has(round, row, chunk) ? aux_[flagIndex(slotOf(round), row, chunk)] : 0; }

  // Number of rows that delivered `chunk` for `round` (reference: count(row, chunkId)).
  int countChunk(int round, int chunk) const {
    int n = 0;
    for (int p = 0; p < peers_; ++p) n += has(round, p, chunk) ? 1 : 0;
    return n;
  }
  uint64_t maskChunk(int round, int chunk) const {
    uint64_t m = 0;
    for (int p = 0; p < peers_; ++p)
      if (has(round, p, chunk)) m |= (1ull << p);
    return m;
  }
  // Total (row, chunk) arrivals in the round (reference: the sum inside reachRoundThreshold).
  int countRound(int round) const {
    int n = 0;
    for (int p = 0; p < peers_; ++p)
      for (int c = 0; c < chunksPerRow_[p]; ++c) n += has(round, p, c) ? 1 : 0;
    return n;
  }

  const float* row(int round, int p) const { return rowPtr(slotOf(round), p); }
  float* rowPtr(int slot, int p) { return data_.data() + ((size_t)slot * peers_ + p) * rowLen_; }
  const float* rowPtr(int slot, int p) const { return data_.data() + ((size_t)slot * peers_ + p) * rowLen_; }

  // Raw state for checkpoint / restore.
  std::vector<float>& rawData() { return data_; }
  std::vector<uint8_t>& rawArrived() { return arrived_; }
  std::vector<int>& rawAux() { return aux_; }
  std::vector<int>& rawSlotRound() { return slotRound_; }
  const std::vector<int>& chunksPerRow() const { return chunksPerRow_; }

 private:
  size_t flagIndex(int slot, int row, int chunk) const { return ((size_t)slot * peers_ + row) * std::max(1, maxChunks_) + chunk; }
  void clearSlot(int s) {
    size_t per = (size_t)peers_ * std::max(1, maxChunks_);
    std::fill(arrived_.begin() + s * per, arrived_.begin() + (s + 1) * per, 0);
    std::fill(aux_.begin() + s * per, aux_.begin() + (s + 1) * per, 0);
  }

  int rowLen_ = 0, peers_ = 0, depth_ = 0, maxChunk_ = 1024, maxChunks_ = 0;
  std::vector<int> chunksPerRow_;
  std::vector<float> data_;
  std::vector<uint8_t> arrived_;
  std::vector<int> aux_;  // per arrival payload (the reduce buffer keeps the contributor count here, B5)
  std::vector<int> slotRound_;
};

}  // namespace aar
