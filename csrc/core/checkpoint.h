// Save / restore of round state and configuration.
//
// The reference has no checkpointing: all state lives in actor fields and InitWorkers wipes it
// (AllreduceWorker.scala:37-47).  The saved layout mirrors those fields so a snapshot is readable
// next to the reference source:
//   worker  {id, round, maxRound, maxScattered, completed[], thReduce, thComplete, maxLag, dataSize,
//            maxChunkSize}                                   (AllreduceWorker.scala:12-33)
//   buffer  {temporalOffset, countFilled[L][numChunks], temporalBuffer[L][P][rowLen]} + the arrival
//            bitmap that replaces the reference's bare counters  (DataBuffer.scala:10,14,27)
//   master  {round, numComplete (per open round), workers}   (AllreduceMaster.scala:26-30)
// File format: little-endian, magic "AARCKPT1", then tagged sections; see checkpoint.cpp.
#pragma once
#include <string>

#include "config.h"
#include "master.h"
#include "worker.h"

namespace aar {

void saveWorkerCheckpoint(const std::string& path, Worker& w);
// Re-creates the worker's state (bookkeeping + both rings). The worker must have been constructed with its
// source/sink/port; peers/master come from the file.
void loadWorkerCheckpoint(const std::string& path, Worker& w);

void saveMasterCheckpoint(const std::string& path, const Master& m);
void loadMasterCheckpoint(const std::string& path, Master& m);
// Restart after the coordinator process died: the saved worker addresses are gone with it, so instead of restoring the
// member table a NEW master continues the round numbering -- the first admission starts at the first round that had not
// reached its completion quorum, with a generation above the saved one.  Rounds in flight at the save are run again
// (workers hold no state across a restart; the data source is asked for those iterations again).  Returns that round.
int resumeMasterFromCheckpoint(const std::string& path, Master& m);

// Device-ring snapshot used by the CUDA engine (host copies of the ring slots and flag words).
struct RingSnapshot {
  int rank = 0, P = 0, L = 0, dataSize = 0, maxChunkSize = 0, round = 0;
  std::vector<float> scatterRing, reduceRing;
  std::vector<uint32_t> scatterFlags;
  std::vector<uint64_t> reduceFlags;
};
void saveRingSnapshot(const std::string& path, const RingSnapshot& s);
RingSnapshot loadRingSnapshot(const std::string& path);

}  // namespace aar
