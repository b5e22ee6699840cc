// Block / chunk partitioning and threshold arithmetic shared by the host protocol
// engine, the CPU oracle and the CUDA round engine.
//
// Reference behaviour being restated (not copied):
//   partitioning   AllreduceWorker.scala:211-228 (stepSize = ceil(dataSize / P), last block clipped)
//   chunking       AllreduceWorker.scala:55-57, DataBuffer.scala:12
//   thresholds     DataBuffer.scala:28-29 (float32 multiply, then truncation)
// Deliberate fixes: B1 chunk by destination block, B2 empty tail
// blocks are legal, B3 the round threshold counts the true number of chunks, B4 floor of 1.
#pragma once
#include <algorithm>
#include <cstdint>
#include <stdexcept>
#include <vector>

namespace aar {

struct Geometry {
  int P = 0;             // number of peers (workers)
  int dataSize = 0;      // elements per worker per round
  int maxChunkSize = 1024;
  int stepSize = 0;      // nominal block length = ceil(dataSize / P)

  Geometry() = default;
  Geometry(int peers, int data_size, int max_chunk) : P(peers), dataSize(data_size), maxChunkSize(max_chunk) {
    if (P <= 0) throw std::invalid_argument("Geometry: P must be positive");
    if (dataSize < 0) throw std::invalid_argument("Geometry: dataSize must be >= 0");
    if (maxChunkSize <= 0) throw std::invalid_argument("Geometry: maxChunkSize must be positive");
    stepSize = (int)(((int64_t)dataSize + P - 1) / P);
  }

  int blockStart(int i) const { return (int)std::min<int64_t>((int64_t)i * stepSize, dataSize); }
  int blockEnd(int i) const { return (int)std::min<int64_t>((int64_t)(i + 1) * stepSize, dataSize); }
  int blockLen(int i) const { return blockEnd(i) - blockStart(i); }
  int maxBlockLen() const { return blockLen(0); }
  int numChunks(int i) const { return (blockLen(i) + maxChunkSize - 1) / maxChunkSize; }
  int maxNumChunks() const { return numChunks(0); }
  int chunkStart(int /*block*/, int c) const { return c * maxChunkSize; }  // offset inside the block
  int chunkLen(int block, int c) const {
    int len = blockLen(block);
    int s = c * maxChunkSize;
    return std::max(0, std::min(maxChunkSize, len - s));
  }
  int totalChunks() const {
    int t = 0;
    for (int i = 0; i < P; ++i) t += numChunks(i);
    return t;
  }
  // Position of (owner, chunk) in the flattened per-round chunk list (owner-major).
  int chunkIndex(int owner, int c) const {
    int t = 0;
    for (int i = 0; i < owner; ++i) t += numChunks(i);
    return t + c;
  }

  // (thReduce * P).toInt evaluated in float32, floored at 1 (B4).
  static int minRequired(float th, int P) {
    int v = (int)((float)th * (float)P);
    return std::max(1, std::min(v, P));
  }
  // (thComplete * P * numChunks).toInt in float32, left-to-right like the reference when
  // every block has the same chunk count; otherwise over the true chunk total (B3).
  int minChunksRequired(float th) const {
    int total = totalChunks();
    if (total == 0) return 0;
    int v;
    if (total == P * maxNumChunks())
      v = (int)(((float)th * (float)P) * (float)maxNumChunks());
    else
      v = (int)((float)th * (float)total);
    return std::max(1, std::min(v, total));
  }
};

}  // namespace aar
