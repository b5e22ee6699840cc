// Native round engine (see include/akka_allreduce_b200.h): the per-round host work of parallel/engine.py behind one C call.
//
// Reference steps covered: StartAllreduce -> fetch -> scatter (AllreduceWorker.scala:84-104) and complete (:253-268) as far
// as they touch the data plane; the actor mailbox is replaced by stream order.  Rounds of one rank alternate between two
// internal streams when the ring has at least two slots (the maxLag window), exactly like the Python engine.
// The default enqueue path of parallel/engine.py (AAR_NATIVE_ENGINE=0 switches back to the Python choreography).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../../include/akka_allreduce_b200.h"
#include "aar_host.h"

struct AarEngine {
  AarLayout lay;
  AarRankParams p;
  unsigned baseFlags = 0;
  int algo = 0, grid = 0, split = 32 | (48 << 16), nvlsReduce = 8, nvlsStage = 48, bulkSplit = 0;
  int nStreams = 0;   // internal streams the rounds rotate over (0: rounds run on the caller's stream)
  int device = 0;
  cudaStream_t streams[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  // One event per round in a ring of L + streams entries: what a later round, a consumer or synchronize() waits for.
  // Round R is launched behind (a) the caller's stream, (b) round R - L (same output slot) and (c) EVERY round <= R - L of this
  // rank: retirement is per ring slot, so nothing else keeps one stream from running ahead of the other, and a peer that sees
  // round R retired must be able to conclude that no rank is still in a round <= R - L (the catch-up predicate relies on it).
  std::vector<cudaEvent_t> slotDone;
  std::vector<int> slotRound;          // round whose completion the event holds, -1 = none
  int newest = -1;
  cudaEvent_t inputReady = nullptr;    // records the caller's stream before a round moves to an internal stream
  unsigned* hostWords = nullptr;       // [0] doorbell, [kDoneWord0 + slot] round+1 of the last round flushed in that slot (pinned, mapped)
  int inOrder = 0;                     // rounds known flushed in order (host-side scan of the slot words)
  unsigned* devWords = nullptr;
  long long launches = 0;
};

namespace {

using aar_host::fail;
constexpr int kDoneWord0 = 16, kHostWords = 16 + 64;  // room for a ring of up to 64 slots

#define AAR_CUDA(call, what)                                                                     \
  do {                                                                                           \
    cudaError_t e_ = (call);                                                                     \
    if (e_ != cudaSuccess) return fail(AAR_ECUDA, "%s: %s", what, cudaGetErrorString(e_));       \
  } while (0)

void destroyParts(AarEngine* e) {
  for (cudaEvent_t ev : e->slotDone)
    if (ev) cudaEventDestroy(ev);
  if (e->inputReady) cudaEventDestroy(e->inputReady);
  for (cudaStream_t s : e->streams)
    if (s) cudaStreamDestroy(s);
  if (e->hostWords) cudaFreeHost(e->hostWords);
}

}  // namespace

AAR_API int aar_engine_create(const AarLayout* layout, int rank, void* const* regions, void* mcRegion, int minRequired, int minChunksRequired,
                              unsigned flags, unsigned long long timeoutNs, int algo, int grid, int overlap, AarEngine** out) {
  const char* what = "aar_engine_create";
  if (!layout || !regions || !out) return fail(AAR_EINVAL, "%s: null argument", what);
  if (rank < 0 || rank >= layout->P || layout->P > AAR_MAXP) return fail(AAR_EINVAL, "%s: bad rank %d for P=%d", what, rank, layout->P);
  if (algo < 0 || algo > 6) return fail(AAR_EINVAL, "%s: algo %d (0 two-shot, 1 pipelined, 2 one-shot, 3 NVLS, 4 bulk P2P, 5 bulk NVLS, 6 bulk NVLS one hop)", what, algo);
  if (algo == 2 && !layout->oneshotOff) return fail(AAR_EINVAL, "%s: the layout has no one-shot buffers", what);
  if ((algo == 3 || algo == 5 || algo == 6) && !(mcRegion && layout->stageOff)) return fail(AAR_EINVAL, "%s: NVLS needs a multicast mapping and the staging area", what);
  AarEngine* e = new AarEngine();
  e->lay = *layout;
  memset(&e->p, 0, sizeof e->p);
  e->p.rank = rank;
  e->p.minRequired = minRequired;
  e->p.minChunksRequired = minChunksRequired;
  e->p.maxLag = layout->L - 1;
  e->p.timeoutNs = timeoutNs;
  e->p.mcRegion = mcRegion;
  for (int j = 0; j < layout->P; ++j) e->p.region[j] = regions[j];
  e->baseFlags = flags;
  e->algo = algo;
  e->grid = grid;
  // overlap: 0 / 1 = one round at a time, n >= 2 = up to n rounds of this rank in flight (never more than the ring has slots)
  e->nStreams = overlap >= 2 ? overlap : (overlap == 1 ? 2 : 0);
  if (e->nStreams > layout->L) e->nStreams = layout->L;
  if (e->nStreams > 8) e->nStreams = 8;
  if (e->nStreams < 2) e->nStreams = 0;
  auto bail = [&](cudaError_t err, const char* step) {
    int rc = fail(AAR_ECUDA, "%s: %s: %s", what, step, cudaGetErrorString(err));
    destroyParts(e);
    delete e;
    return rc;
  };
  cudaError_t err = cudaGetDevice(&e->device);
  if (err != cudaSuccess) return bail(err, "cudaGetDevice");
  void* h = nullptr;
  if (layout->L > kHostWords - kDoneWord0) return bail(cudaErrorInvalidValue, "ring deeper than 64 slots");
  err = cudaHostAlloc(&h, kHostWords * sizeof(unsigned), cudaHostAllocMapped | cudaHostAllocPortable);
  if (err != cudaSuccess) return bail(err, "cudaHostAlloc");
  memset(h, 0, kHostWords * sizeof(unsigned));
  e->hostWords = static_cast<unsigned*>(h);
  void* d = nullptr;
  err = cudaHostGetDevicePointer(&d, h, 0);
  if (err != cudaSuccess) return bail(err, "cudaHostGetDevicePointer");
  e->devWords = static_cast<unsigned*>(d);
  e->p.doorbell = reinterpret_cast<const int*>(e->devWords);
  e->p.hostDone = e->devWords + kDoneWord0;
  if (e->nStreams) {
    // The round streams get the device's greatest priority unless AAR_STREAM_PRIORITY=default: a bulk CTA needs a whole SM's
    // shared memory, so when the caller's own kernels (a backward pass) keep every SM busy, a round only starts promptly if
    // its CTAs are placed ahead of the caller's pending ones.
    int least = 0, greatest = 0;
    err = cudaDeviceGetStreamPriorityRange(&least, &greatest);
    if (err != cudaSuccess) return bail(err, "cudaDeviceGetStreamPriorityRange");
    const char* pe = getenv("AAR_STREAM_PRIORITY");
    const bool plain = pe && (!strcmp(pe, "default") || !strcmp(pe, "0"));
    for (int i = 0; i < e->nStreams; ++i) {
      err = cudaStreamCreateWithPriority(&e->streams[i], cudaStreamNonBlocking, plain ? 0 : greatest);
      if (err != cudaSuccess) return bail(err, "cudaStreamCreateWithPriority");
    }
  }
  const int ring = layout->L + (e->nStreams > 1 ? e->nStreams : 1);
  e->slotDone.assign((size_t)ring, nullptr);
  e->slotRound.assign((size_t)ring, -1);
  for (int s = 0; s < ring; ++s) {
    err = cudaEventCreateWithFlags(&e->slotDone[(size_t)s], cudaEventDisableTiming);
    if (err != cudaSuccess) return bail(err, "cudaEventCreateWithFlags");
  }
  err = cudaEventCreateWithFlags(&e->inputReady, cudaEventDisableTiming);
  if (err != cudaSuccess) return bail(err, "cudaEventCreateWithFlags");
  *out = e;
  return AAR_OK;
}

AAR_API int aar_engine_tune(AarEngine* e, int split, int nvlsCtas) {
  if (!e) return fail(AAR_EINVAL, "aar_engine_tune: null engine");
  if (split > 0) e->split = split;
  if (nvlsCtas > 0) {
    e->nvlsReduce = nvlsCtas & 0xff;
    e->nvlsStage = (nvlsCtas >> 8) & 0xff;
  }
  return AAR_OK;
}

AAR_API int aar_engine_enqueue(AarEngine* e, int round, const void* input, void* output, int* outCounts, unsigned extraFlags, void* callerStream) {
  return aar_engine_enqueue_ex(e, round, input, output, outCounts, extraFlags, callerStream, 0ull);
}

AAR_API int aar_engine_tune_bulk(AarEngine* e, int split) {
  if (!e) return fail(AAR_EINVAL, "aar_engine_tune_bulk: null engine");
  e->bulkSplit = split;
  return AAR_OK;
}

// Round numbering that does not begin at 0 (a master resumed from a checkpoint): the completion word starts at `round`.
AAR_API int aar_engine_start_at(AarEngine* e, int round) {
  if (!e || round < 0) return fail(AAR_EINVAL, "aar_engine_start_at: bad argument");
  volatile unsigned* w = e->hostWords;
  for (int r = round - e->lay.L; r < round; ++r)
    if (r >= 0) w[kDoneWord0 + r % e->lay.L] = (unsigned)(r + 1);   // every earlier round counts as flushed
  e->inOrder = round;
  return AAR_OK;
}

AAR_API int aar_engine_enqueue_ex(AarEngine* e, int round, const void* input, void* output, int* outCounts, unsigned extraFlags, void* callerStream,
                                  unsigned long long delayNs) {
  const char* what = "aar_engine_enqueue";
  if (!e || !input || !output) return fail(AAR_EINVAL, "%s: null argument", what);
  if (round < 0) return fail(AAR_EINVAL, "%s: round %d", what, round);
  const int ring = (int)e->slotDone.size();
  cudaStream_t caller = static_cast<cudaStream_t>(callerStream);
  cudaStream_t run = caller;
  AarRankParams p = e->p;  // by value: the launchers copy it into the kernel parameters before they return
  p.round = round;
  p.input = input;
  p.output = output;
  p.outCounts = outCounts;
  p.flags = e->baseFlags | extraFlags;
  if (e->nStreams) {
    run = e->streams[round % e->nStreams];
    AAR_CUDA(cudaEventRecord(e->inputReady, caller), "cudaEventRecord(input)");   // the input was produced on the caller's stream
    AAR_CUDA(cudaStreamWaitEvent(run, e->inputReady, 0), "cudaStreamWaitEvent(input)");
    p.flags |= AAR_FLAG_OVERLAPPED;
    // every round <= R - L of this rank must be done: one wait per other stream, on that stream's newest such round (the
    // round's own stream is in order already)
    for (int s = 1; s < e->nStreams; ++s) {
      int q = round - e->lay.L;
      while (q >= 0 && q % e->nStreams != (round + s) % e->nStreams) --q;
      if (q >= 0 && e->slotRound[(size_t)(q % ring)] == q)
        AAR_CUDA(cudaStreamWaitEvent(run, e->slotDone[(size_t)(q % ring)], 0), "cudaStreamWaitEvent(window)");
    }
  }
  {  // the output / count buffers of this slot belong to round R - L until it is done
    const int q = round - e->lay.L;
    if (q >= 0 && e->slotRound[(size_t)(q % ring)] == q)
      AAR_CUDA(cudaStreamWaitEvent(run, e->slotDone[(size_t)(q % ring)], 0), "cudaStreamWaitEvent(slot)");
  }
  int rc;
  if (delayNs) {  // straggler injection: this rank's round starts late
    rc = aar_launch_delay(delayNs, run);
    if (rc) return rc;
  }
  switch (e->algo) {
    case 0:
      rc = aar_launch_round_fused(&p, &e->lay, e->grid, run);
      break;
    case 1:
      rc = aar_launch_round_pipelined(&p, &e->lay, e->grid, e->split, run);
      break;
    case 2:
      rc = aar_launch_oneshot(&p, &e->lay, 2, e->grid, run);
      break;
    case 3:
      rc = aar_launch_nvls(&p, &e->lay, 7 | (e->nvlsReduce << 8) | (e->nvlsStage << 16), e->grid, run);
      break;
    default:
      rc = aar_launch_round_bulk(&p, &e->lay, e->algo == 6 ? 2 : (e->algo == 5 ? 1 : 0), 7, e->grid, e->bulkSplit, run);
      break;
  }
  if (rc) return rc;
  AAR_CUDA(cudaEventRecord(e->slotDone[(size_t)(round % ring)], run), "cudaEventRecord(round)");
  e->slotRound[(size_t)(round % ring)] = round;
  if (round > e->newest) e->newest = round;
  e->launches++;
  return AAR_OK;
}

AAR_API int aar_engine_wait(AarEngine* e, int round, void* stream) {
  if (!e || round < 0) return fail(AAR_EINVAL, "aar_engine_wait: bad argument");
  const size_t at = (size_t)(round % (int)e->slotDone.size());
  if (e->slotRound[at] < round) return AAR_OK;  // not launched (yet)
  // == round: that round's event.  > round: the entry was reused by a later round, which was launched behind this one's completion
  AAR_CUDA(cudaStreamWaitEvent(static_cast<cudaStream_t>(stream), e->slotDone[at], 0), "cudaStreamWaitEvent(round)");
  return AAR_OK;
}

AAR_API int aar_engine_synchronize(AarEngine* e) {
  if (!e) return fail(AAR_EINVAL, "aar_engine_synchronize: null engine");
  for (size_t s = 0; s < e->slotDone.size(); ++s)
    if (e->slotRound[s] >= 0) AAR_CUDA(cudaEventSynchronize(e->slotDone[s]), "cudaEventSynchronize");
  return AAR_OK;
}

AAR_API int aar_engine_set_doorbell(AarEngine* e, int newestRound) {
  if (!e) return fail(AAR_EINVAL, "aar_engine_set_doorbell: null engine");
  volatile unsigned* w = e->hostWords;
  if (newestRound >= 0 && (unsigned)newestRound > w[0]) w[0] = (unsigned)newestRound;
  return AAR_OK;
}

AAR_API int aar_engine_round_flushed(AarEngine* e, int round) {
  if (!e || round < 0) return 0;
  // slot words only grow (round + 1, then round + L + 1, ...): >= covers a later tenant of the slot as well
  return static_cast<volatile unsigned*>(e->hostWords)[kDoneWord0 + round % e->lay.L] >= (unsigned)(round + 1) ? 1 : 0;
}

AAR_API int aar_engine_rounds_flushed(AarEngine* e) {
  if (!e) return -1;
  while (aar_engine_round_flushed(e, e->inOrder)) e->inOrder++;
  return e->inOrder;
}

AAR_API long long aar_engine_launches(AarEngine* e) { return e ? e->launches : -1; }

// Kernels of this engine may still be waiting for peers: ring the doorbell far ahead so that every device-side wait gives up
// (the catch-up predicate), let them finish, then free what they were polling.
AAR_API int aar_engine_destroy(AarEngine* e) {
  if (!e) return AAR_OK;
  static_cast<volatile unsigned*>(e->hostWords)[0] = 1u << 30;
  for (size_t s = 0; s < e->slotDone.size(); ++s)
    if (e->slotRound[s] >= 0) cudaEventSynchronize(e->slotDone[s]);
  destroyParts(e);
  delete e;
  return AAR_OK;
}
