#include "device_loop.h"

#include <sched.h>

#include <chrono>
#include <stdexcept>
#include <string>

#include "logging.h"
#include "tcp_transport.h"

namespace aar {

namespace {
inline void cpuRelax() {
#if defined(__x86_64__) || defined(__i386__)
  __builtin_ia32_pause();
#elif defined(__aarch64__)
  asm volatile("yield");
#endif
}
}  // namespace

void DeviceLoop::setPlane(const DevicePlane& plane, int id, int generation, Address master, int startRound) {
  if (!plane.valid() || plane.slots < 1) throw std::invalid_argument("DeviceLoop: incomplete data plane");
  std::lock_guard<std::mutex> g(mu_);
  plane_ = plane;
  id_ = id;
  generation_ = generation;
  master_ = master;
  startRound_ = startRound;
  maxRound_ = maxScattered_ = startRound - 1;
  round_ = startRound;
  completed_.clear();
  stats_ = DeviceLoopStats();
}

void DeviceLoop::receive(const Message& msg) {
  if (auto* init = std::get_if<InitWorkers>(&msg)) {
    if (init_) init_(*init);
  } else if (auto* start = std::get_if<StartAllreduce>(&msg)) {
    if (!dropStart_ && (start->generation == 0 || start->generation == generation_)) onStart(start->round);
  }
  // ScatterBlock / ReduceBlock never travel on the control channel of the device plane: they are NVLink stores
}

void DeviceLoop::onStart(int round) {
  if (!plane_.valid()) return;   // before InitWorkers: nothing to start (the master re-sends after initialisation)
  if (round > maxRound_) {
    std::lock_guard<std::mutex> g(mu_);
    maxRound_ = round;
  }
  plane_.doorbell(plane_.engine, maxRound_);   // kernels more than maxLag behind stop waiting: the catch-up loop without a host round-trip
  if (boardPost_) boardPost_(maxRound_);
  const int L = plane_.slots;
  while (maxScattered_ < maxRound_) {
    const int r = maxScattered_ + 1;
    // the output / count buffers of round r - L must have reached the sink before round r reuses them
    while (r - L >= round_ && !completed_.count(r - L)) deliverFlushed(true);
    const RoundBuffers b = source_(r);
    if (b.numel != plane_.dataSize)
      throw std::invalid_argument("Input data size " + std::to_string(b.numel) + " is different from initialization time " +
                                  std::to_string(plane_.dataSize) + "!");
    const int rc = plane_.enqueue(plane_.engine, r, reinterpret_cast<const void*>(b.input), reinterpret_cast<void*>(b.output),
                                  reinterpret_cast<int*>(b.counts), plane_.extraFlags, reinterpret_cast<void*>(b.stream), 0ull);
    if (rc != 0) throw std::runtime_error("DeviceLoop: enqueue of round " + std::to_string(r) + " failed with code " + std::to_string(rc));
    std::lock_guard<std::mutex> g(mu_);
    ++stats_.launched;
    maxScattered_ = r;
  }
}

bool DeviceLoop::peerAhead() {
  if (!boardNewest_ || !plane_.valid()) return false;
  const int newest = boardNewest_();
  if (newest <= maxRound_) return false;
  {
    std::lock_guard<std::mutex> g(mu_);
    stats_.selfStarts += newest - maxRound_;
  }
  onStart(newest);
  return true;
}

int DeviceLoop::deliverFlushed(bool block) {
  if (!plane_.valid()) return 0;
  int total = 0;
  for (;;) {
    int progressed = 0;
    for (int r = round_; r <= maxScattered_; ++r) {
      if (r < round_ || completed_.count(r)) continue;   // round_ moves inside this loop
      if (!plane_.flushed(plane_.engine, r)) continue;
      sink_(r, r % plane_.slots);
      if (port_) port_->send(master_, CompleteAllreduce{id_, r, generation_});
      ++progressed;
      std::lock_guard<std::mutex> g(mu_);
      const int lag = maxRound_ - r;
      ++stats_.lagHist[lag];
      if (lag > plane_.maxLag) ++stats_.forced;
      if (r != round_) ++stats_.outOfOrder;
      completed_.insert(r);
      ++stats_.delivered;
      while (completed_.count(round_)) {   // AllreduceWorker.scala:261-267: past every already-completed round
        completed_.erase(round_);
        ++round_;
      }
    }
    total += progressed;
    if (!block || progressed || !busy()) return total;
    sched_yield();   // blocking mode: the slot this caller needs is still in flight on the device
  }
}

void DeviceLoop::drain() {
  if (!plane_.valid()) return;
  // Rounds still in flight can only be ones the master no longer waits for, or ones whose peers are gone: their device-side
  // waits give up once the doorbell is far ahead.
  plane_.doorbell(plane_.engine, 1 << 30);
  if (plane_.synchronize) plane_.synchronize(plane_.engine);
  deliverFlushed(false);
}

bool DeviceLoop::run(WorkerNode& node, double timeoutSec) {
  using Clock = std::chrono::steady_clock;
  const auto deadline = Clock::now() + std::chrono::duration<double>(timeoutSec);
  node.setHandler([this](const Message& m) { receive(m); });
  unsigned idle = 0;
  while (!node.shutdownRequested() && Clock::now() < deadline) {
    // Nothing in flight: sleep in poll() until the master (or a heartbeat deadline) wakes us.  Rounds in flight: watch the
    // completion words closely for a short while (small rounds finish within microseconds), then back off -- the socket and the
    // board are looked at every eighth pass, with a pause of about two microseconds between passes -- so that a long round does
    // not cost a core hammering system calls.
    const bool inFlight = busy();
    const bool lookAround = !inFlight || idle < 64 || (idle & 7u) == 0;
    const int handled = lookAround ? node.poll(inFlight ? 0 : 1) : 0;
    const bool started = lookAround && peerAhead();
    const int delivered = deliverFlushed(false);
    if (handled || started || delivered) {
      idle = 0;
    } else if (inFlight && ++idle >= 64) {
      for (int i = 0; i < 48; ++i) cpuRelax();
    }
  }
  drain();
  return node.shutdownRequested();
}

}  // namespace aar
