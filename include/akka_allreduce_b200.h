/* C ABI of the sm_100a data plane (libaar_cuda.so).
 *
 * The reference moves ScatterBlock / ReduceBlock actor messages over TCP
 * (AllreduceWorker.scala:194-251).  Here the same steps are CUDA kernels that store straight into
 * peer GPUs' ring slots over NVLink and publish arrival with release flags; see csrc/cuda/.
 * Every function returns 0 on success or a negative AAR_E* code; aar_last_error() gives the text.
 * `stream` arguments are cudaStream_t passed as void*.
 */
#ifndef AKKA_ALLREDUCE_B200_H
#define AKKA_ALLREDUCE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AAR_MAXP 16 /* peers addressable by one round engine (one NVSwitch domain) */

#define AAR_OK 0
#define AAR_EINVAL -1
#define AAR_ECUDA -2
#define AAR_EUNSUPPORTED -3

/* element types on the wire / in the rings (accumulation is always fp32) */
#define AAR_F32 0
#define AAR_BF16 1

/* AarRankParams.flags */
#define AAR_FLAG_NORMALIZE 1u   /* divide each reduced chunk by its contributor count */
#define AAR_FLAG_EXACT_QUORUM 2u /* use exactly minRequired contributors (lowest ids), like the reference's == firing */
#define AAR_FLAG_FORCE 4u       /* K8: do not wait, take whatever has arrived (catch-up, AllreduceWorker.scala:91-97) */
#define AAR_FLAG_SKIP_SCATTER 8u /* fused kernel: this rank contributes nothing this round (fault injection) */
#define AAR_FLAG_VIEW_OUTPUT 16u /* K3 does not copy: the caller reads the reduce-ring slot itself, stream-ordered */
#define AAR_FLAG_OVERLAPPED 64u /* consecutive rounds of this rank run on different streams (informational since retirement is per slot) */
#define AAR_FLAG_ELIDE_LATE 512u /* thresholds < 1: a rank that is a whole round late skips pushes whose owner has decided without it
                                   and broadcasts to peers that have retired the round; needs the strict kernel variants (set
                                   together with AAR_FLAG_BCAST_FC); opt-in until verified on hardware */
#define AAR_FLAG_EAGER_DEC 256u /* thComplete = 1: the eager gather takes ONE decision per (round, chunk) -- arrived with this count, or
                                   abandoned -- so a wait that gives up can never leave a chunk half copied, half zero-filled
                                   (docs/PROTOCOL.md "Known limits"); opt-in until verified on hardware */
#define AAR_FLAG_BCAST_FC 128u  /* thresholds < 1: the reduce-broadcast adopts the scatter's per-destination slot flow control
                                   (docs/PROTOCOL.md "Known limits"); opt-in until verified on hardware */
#define AAR_FLAG_SELF_DIRECT 1024u /* aar_launch_round_bulk: a rank's own contribution is read from its input (no copy into its own
                                   scatter ring) and, when every chunk is required, its own reduced block goes straight to the output */
#define AAR_FLAG_TRACE 32u      /* every CTA stamps %globaltimer at its phase boundaries into the trace section */
#define AAR_TRACE_CTAS 512
#define AAR_TRACE_MARKS 8

/* bits of the device status word (AarLayout.statusOff, word 0) */
#define AAR_ST_TIMEOUT_SCATTER 1u
#define AAR_ST_TIMEOUT_REDUCE 2u
#define AAR_ST_TIMEOUT_GATHER 4u
#define AAR_ST_PUSH_DROPPED 8u
#define AAR_ST_BULK_FAULT 16u /* a bulk load never completed (bad address): the kernel trapped instead of hanging */

/* Byte offsets of the sections of one rank's symmetric region; identical on every rank. */
typedef struct AarLayout {
  int P, L, C;               /* peers, ring depth (maxLag+1), max chunks per block */
  int dataSize, maxChunk, stepSize, blkStride, totalChunks;
  int elemSize, dtype;
  int tileElems, tilesPerChunk;
  uint64_t scatterOff;       /* T   [L][P(src)][blkStride]      written by peers (K1) */
  uint64_t reduceOff;        /* T   [L][P(owner)][blkStride]    written by peers (K2); slot row == flushed vector */
  uint64_t sflagOff;         /* u32 [L][C][AAR_MAXP]            epoch = round+1, written by src */
  uint64_t rflagOff;         /* u64 [L][P(owner)][C]            epoch<<32 | contributor count, written by owner */
  uint64_t peerBaseOff;      /* u32 [AAR_MAXP]                  rounds retired by each peer (flow control) */
  uint64_t k1DecOff;         /* u64 [L][AAR_MAXP]               local: per-destination go/skip decision */
  uint64_t k1CtrOff;         /* u32 [L][P][C]                   local: last-CTA election, scatter */
  uint64_t k2DecOff;         /* u64 [L][C]                      local: epoch<<32 | contributor mask */
  uint64_t k2CtrOff;         /* u32 [L][C]                      local: last-CTA election, broadcast */
  uint64_t k3DecOff;         /* u64 [L]                         local: epoch<<32 | state */
  uint64_t k3CtrOff;         /* u32 [L]                         local: last-CTA election, retire (per slot: rounds overlap) */
  uint64_t countsOff;        /* i32 [L][totalChunks]            local: per-chunk contributor count, -1 = missing */
  uint64_t statusOff;        /* u32 [16]                        local: status bits, counters */
  uint64_t oneshotOff;       /* T   [L][P(src)][dataSize] + u32 [L][AAR_MAXP] flags; 0 if not allocated */
  uint64_t oneshotFlagOff;
  uint64_t stageOff;         /* T   [L][dataSize] staging for the NVLS path; 0 if not allocated */
  uint64_t barrierOff;       /* reserved */
  uint64_t traceOff;         /* u64 [AAR_TRACE_CTAS][AAR_TRACE_MARKS] phase timestamps (AAR_FLAG_TRACE) */
  uint64_t ctrlBytes;        /* first byte of the control area that must be zeroed at allocation */
  uint64_t regionBytes;      /* total, rounded up to 2 MiB */
  uint64_t egDecOff;         /* u64 [L][P][C]                   local: eager gather, epoch<<32 | count+1 (0 = abandoned); AAR_FLAG_EAGER_DEC */
  /* pieces: sub-chunk granules (<= 8 tiles) in which the kernels of aar_launch_round_bulk publish arrival.  Exactly one CTA moves
   * a piece, so its flag needs no election; a chunk has arrived from a source when all its pieces are flagged. */
  int tilesPerPiece, piecesPerChunk;
  uint64_t pflagSOff;        /* u32 [L][C][piecesPerChunk][AAR_MAXP]   epoch, written by src: "my piece of your block is staged" */
  uint64_t pflagROff;        /* u64 [L][P(owner)][C][piecesPerChunk]   epoch<<32 | contributor count, written by owner: "reduced piece has landed" */
  uint64_t slotBaseOff;      /* u32 [L][AAR_MAXP]   per ring slot: round+1 of the last round each peer retired IN THAT SLOT.  Flow control for
                                slot reuse reads this (round R may overwrite slot R % L of a peer once that peer retired round R - L), so rounds
                                need not retire in order: round r+1 completes before round r when its quorum lands first
                                (AllreduceWorker.scala:253-268, AllreduceSpec.scala:675-677).  peerBase keeps the NEWEST retired round per peer. */
} AarLayout;

/* Per-rank launch parameters (passed to the kernels by value). */
typedef struct AarRankParams {
  int rank;
  int round;                 /* absolute round; slot = round % L, epoch = round + 1 */
  int minRequired;           /* floor(thReduce * P), >= 1 */
  int minChunksRequired;     /* floor(thComplete * sum of chunks) */
  int maxLag;
  unsigned flags;
  unsigned long long timeoutNs; /* liveness guard for every device-side wait; 0 = unbounded */
  const void* input;         /* dataSize elements (this round's source data) */
  void* output;              /* dataSize elements */
  int* outCounts;            /* totalChunks ints (AllReduceOutput.count), may be NULL */
  const int* doorbell;       /* host-mapped: newest round announced by the master, may be NULL */
  unsigned* hostDone;        /* host-mapped u32[L]: word (round % L) receives round+1 when the round is flushed, may be NULL */
  void* region[AAR_MAXP];    /* every rank's region as mapped in this rank's address space */
  void* mcRegion;            /* multicast mapping of the regions (NVLS), may be NULL */
} AarRankParams;

const char* aar_last_error(void);
/* sizeof(AarLayout) / sizeof(AarRankParams) as this library was compiled: language bindings compare them with their own
   declarations before the first call (a field added on one side only would otherwise corrupt memory silently). */
int aar_abi_sizes(int* layoutBytes, int* rankParamsBytes);

/* Geometry -> layout.  maxLag >= 0; dtype AAR_F32 | AAR_BF16; oneshotMaxElems = 0 disables the one-shot buffers;
 * withStage != 0 adds the NVLS staging area. */
int aar_layout_compute(int P, int dataSize, int maxChunkSize, int maxLag, int dtype, int oneshotMaxElems, int withStage, AarLayout* out);

/* K1 scatter_push (AllreduceWorker.scala:194-209): copy my block j into peer j's scatter ring + arrival flags. */
int aar_launch_scatter(const AarRankParams* p, const AarLayout* lay, int grid, void* stream);
/* K2 reduce_bcast (AllreduceWorker.scala:106-127,230-251): wait for thReduce contributors per chunk, ordered fp32
 * sum of exactly the arrived set, store to every peer's reduce ring + (epoch,count) flags. */
int aar_launch_reduce_bcast(const AarRankParams* p, const AarLayout* lay, int grid, void* stream);
/* K3 gather_flush (AllreduceWorker.scala:129-150,180-192,253-268 + DataBuffer.up :63-67): wait for thComplete chunks,
 * copy the slot row to the caller's output (zeros for missing chunks), write counts, retire the slot. */
int aar_launch_gather(const AarRankParams* p, const AarLayout* lay, int grid, void* stream);
/* K1+K2+K3 in one launch (one kernel per round per rank; requires peers on other GPUs or an earlier launch). */
int aar_launch_round_fused(const AarRankParams* p, const AarLayout* lay, int grid, void* stream);
/* Same round with the CTAs specialised by role (push / reduce+broadcast / gather) so the phases overlap chunk by chunk.
 * split = nPush | nReduce << 16, 0 = heuristic; the remaining CTAs gather. */
int aar_launch_round_pipelined(const AarRankParams* p, const AarLayout* lay, int grid, int split, void* stream);
/* The same round on the bulk async-copy engine (csrc/cuda/aar_bulk.cu): every copy is cp.async.bulk through a shared-memory
 * ring driven by one thread per CTA, every completion is a bulk-group wait (no store-queue drain), roles sweep the block
 * chunk by chunk so push, reduce and gather of one round overlap.  Same flags, decisions and thresholds as the kernels above
 * (a rank may use either for any round); slot flow control for the broadcast and one eager-gather decision per chunk are
 * always on.  algo 0 = two-shot P2P, 1 = NVLS (thresholds must be 1.0; piece-granular flags, every rank must use it),
 * 2 = NVLS in one hop (small / medium messages: every rank reduces the whole vector through the switch into its own output).
 * phases: 7 = one launch, CTAs specialised by role (split = nA | nB << 16: push/stage CTAs, reduce CTAs; 0 = heuristic);
 * 1 / 2 / 4 = one phase with the whole grid (single-device emulation).  Needs 16-byte aligned geometry and pointers
 * (AAR_EUNSUPPORTED otherwise: use the kernels above). */
int aar_launch_round_bulk(const AarRankParams* p, const AarLayout* lay, int algo, int phases, int grid, int split, void* stream);
/* 1 if aar_launch_round_bulk can run this geometry / these pointers */
int aar_bulk_supported(const AarRankParams* p, const AarLayout* lay, int algo);

/* K5 one-shot: push the whole vector to every peer, then each rank sums the arrived set locally. Two phases so the
 * single-GPU emulation can order them; phase = 0 push, 1 reduce, 2 both. */
int aar_launch_oneshot(const AarRankParams* p, const AarLayout* lay, int phase, int grid, void* stream);
/* K4 NVLS: multimem.ld_reduce over the staged inputs + multimem.st of the result (thresholds must be 1.0).
 * phases is a bit mask: 1 stage + arrival flags, 2 in-switch reduce + multicast store, 4 gather; 7 = one launch with the CTAs
 * specialised by role (| nReduceCtas << 8 | nStageCtas << 16, 0 = defaults). */
int aar_launch_nvls(const AarRankParams* p, const AarLayout* lay, int phases, int grid, void* stream);

/* Single-device cluster emulation: P ranks' regions live on the current device; runs one full round for all
 * ranks in phase order (all K1, all K2, all K3) on `stream`.  params[r] describes rank r.  skipMask bit r set =
 * rank r does not scatter this round (missing contributor). */
int aar_local_round(const AarRankParams* params, int nranks, const AarLayout* lay, unsigned skipMask, int grid, void* stream);

/* No counterpart in the reference.  aar_local_round() for callers that may not allocate device memory (code running
 * inside a host application's fixed memory plan, or on a path where cudaMalloc's implicit synchronisation or failure is not
 * acceptable): one plain allreduce round (all thresholds 1.0) of `nranks` emulated ranks on the current device, with the
 * ring buffers and flag words taken from an array built into the library instead of caller-allocated regions.
 * inputs[r] / outputs[r]: dataSize elements of `dtype` each, device memory owned by the caller.
 * Limit: nranks x per-rank ring bytes must fit AAR_STATIC_SCRATCH_BYTES, otherwise AAR_EINVAL with the sizes in
 * aar_last_error().  No cudaMalloc / cudaFree / allocating routine is called.  The array is shared by every call on a
 * device and keeps the previous call's contents; each call clears its control words on `stream` first.  Calls that overlap
 * on one GPU race on it.  The first call on a device also looks up the array's device address; make that call before any
 * stream capture. */
#define AAR_STATIC_SCRATCH_BYTES (16u << 20)
int aar_local_allreduce_static(const void* const* inputs, void* const* outputs, int nranks, int dataSize, int maxChunkSize, int dtype,
                               void* stream);

/* Plain device allocation helpers for the emulation mode and tests (zero-filled). */
int aar_region_alloc_local(size_t bytes, void** ptr);
int aar_region_free_local(void* ptr);

/* Symmetric memory over CUDA VMM (cuMemCreate + POSIX fd export/import) and multicast objects. */
typedef struct AarVmm {
  uint64_t handle;   /* CUmemGenericAllocationHandle */
  void* ptr;         /* mapped address in this process */
  size_t bytes;      /* rounded up to the allocation granularity */
  int fd;            /* exportable POSIX fd (-1 for imports) */
  int device;
} AarVmm;
int aar_vmm_granularity(int device, int multicast, int nDevices, size_t* out);
int aar_vmm_create(int device, size_t bytes, AarVmm* out);               /* allocate + map + zero + export fd */
int aar_vmm_import(int device, int fd, size_t bytes, AarVmm* out);        /* map a peer's allocation */
int aar_vmm_release(AarVmm* v);
int aar_mc_supported(int device);                                         /* 1 if multicast objects are available */
int aar_mc_create(int nDevices, size_t bytes, uint64_t* mcHandle, int* fd); /* rank 0 */
int aar_mc_import(int fd, uint64_t* mcHandle);                            /* other ranks */
int aar_mc_add_device(uint64_t mcHandle, int device);
int aar_mc_bind_and_map(uint64_t mcHandle, int device, const AarVmm* mem, size_t bytes, void** mcPtr);
int aar_mc_release(uint64_t mcHandle, void* mcPtr, size_t bytes);
/* Host-mapped words (doorbell / completion): pinned + device-visible. */
int aar_host_words_alloc(size_t nWords, unsigned** hostPtr, unsigned** devPtr);
int aar_host_words_free(unsigned* hostPtr);

/* Device properties the selector needs. */
int aar_device_sm_count(int device, int* out);
int aar_enable_peer_access(int device, int peer);

/* ---------------------------------------------------------------------------------------------------------------------
 * Native round engine: what parallel/engine.py does per round (order after the caller's stream, pick the kernel, launch,
 * record the slot's event; two internal streams so that consecutive rounds overlap) behind one C call, for C/C++ callers and
 * as the enqueue path of the Python engine (AAR_NATIVE_ENGINE=0 switches back to the Python choreography).  algo: 0 two-shot fused, 1 pipelined P2P, 2 one-shot, 3 NVLS, 4 bulk P2P, 5 bulk NVLS.  The engine owns its doorbell and
 * completion words (pinned host memory); regions[] are the peers' regions as mapped here (AarRankParams.region). */
typedef struct AarEngine AarEngine;
int aar_engine_create(const AarLayout* layout, int rank, void* const* regions, void* mcRegion, int minRequired, int minChunksRequired,
                      unsigned flags, unsigned long long timeoutNs, int algo, int grid, int overlap, AarEngine** out);
/* overlap: 0 = one round at a time on the caller's stream; 1 = two rounds in flight; n >= 2 = up to n (capped by the ring depth and 8) */
/* knobs of the role-specialised kernels (0 keeps the default): split = nPush | nReduce << 16, nvlsCtas = nReduce | nStage << 8 */
int aar_engine_tune(AarEngine* e, int split, int nvlsCtas);
/* Launch round `round`: input / output / outCounts as in AarRankParams; extraFlags are OR-ed into the engine's flags for this
 * round (AAR_FLAG_FORCE, AAR_FLAG_SKIP_SCATTER).  callerStream: the stream that produced `input` (the round is ordered after
 * what it holds now).  With overlap the round runs on an internal stream; use aar_engine_wait before consuming the output. */
int aar_engine_enqueue(AarEngine* e, int round, const void* input, void* output, int* outCounts, unsigned extraFlags, void* callerStream);
/* The same with a straggler injector: the round starts `delayNs` late on this rank (0 = aar_engine_enqueue). */
int aar_engine_enqueue_ex(AarEngine* e, int round, const void* input, void* output, int* outCounts, unsigned extraFlags, void* callerStream,
                          unsigned long long delayNs);
int aar_engine_tune_bulk(AarEngine* e, int split);            /* role split of the bulk kernels (algo 4 / 5), 0 = heuristic */
int aar_engine_start_at(AarEngine* e, int round);             /* round numbering begins at `round` (resumed master) */
int aar_engine_wait(AarEngine* e, int round, void* stream);   /* `stream` waits for the round (no host blocking) */
int aar_engine_synchronize(AarEngine* e);                      /* host waits for every launched round */
int aar_engine_set_doorbell(AarEngine* e, int newestRound);    /* StartAllreduce from the master: monotone */
int aar_engine_rounds_flushed(AarEngine* e);                   /* rounds flushed IN ORDER from the start (every round below this is done) */
int aar_engine_round_flushed(AarEngine* e, int round);         /* 1 if this round has been flushed (rounds may complete out of order) */
long long aar_engine_launches(AarEngine* e);
int aar_engine_destroy(AarEngine* e);

/* Link micro-benchmark (roofline denominators): mode 0 multimem.ld_reduce, 1 multimem.st, 2 both, 3 peer load, 4 peer
 * store, 5 local copy; buffers hold 2 * bytes; `iters` launches on `stream`, timed by the caller. */
int aar_bench_link(int mode, int grid, int unroll, void* mc, void* peer, void* local, size_t bytes, int iters, void* stream);

/* All-to-all variant: mode 0 each CTA stores to one peer, 1 each CTA stores to all peers, 2 each CTA loads from one
 * peer, 3 like 0 with 32-byte stores.  Every peer buffer holds 2 * bytesPerPeer. */
int aar_bench_all2all(int mode, int grid, int unroll, void** peers, int nPeers, void* local, size_t bytesPerPeer, int iters, void* stream);

/* Copy by the bulk async-copy engine (cp.async.bulk + mbarrier), one issuing thread per CTA: what a register-free copy role
 * could move.  stages in {2,4,8}, tileBytes % 128 == 0; *errOut (device word, may be NULL) is set to 1 if a wait gave up.
 * Bench only -- written without GPU access, not used by any allreduce path. */
int aar_bench_bulk_copy(int grid, int stages, int tileBytes, const void* src, void* dst, size_t bytes, int iters, void* stream, unsigned* errOut);
/* The same with `warps` (1..8) issuing threads per CTA, each with its own ring of `stages` tiles. */
int aar_bench_bulk_copy_multi(int grid, int warps, int stages, int tileBytes, const void* src, void* dst, size_t bytes, int iters, void* stream,
                              unsigned* errOut);

/* Two-GPU litmus (bench/litmus_bulk_flag.py): is a flag written after cp.async.bulk.wait_group -- without a fence -- ever seen
 * by the peer before the bulk-stored data?  sender = 1 on one GPU, 0 on the other, launched at the same time. */
int aar_bench_litmus_bulk_flag(int sender, int mode, int grid, void* peerBuf, void* myBuf, void* peerWords, void* myWords, int bytes, int iters,
                               unsigned long long* violations, unsigned* gaveUp, void* stream);

/* Spin kernel used by the straggler injector: busy-waits `ns` nanoseconds on `stream`. */
int aar_launch_delay(unsigned long long ns, void* stream);

#ifdef __cplusplus
}
#endif
#endif
