// Control (and CPU data) transport over TCP sockets, one poll()-driven event loop per process.
//
// Stands in for what the reference gets from Akka remoting + cluster membership (build.sbt:16-22,
// application.conf:5-21): a master that listens on a seed port, workers that join it with a role,
// death-watch (socket close / missed heartbeats -> Terminated, AllreduceMaster.scala:50-56,74), and
// worker-to-worker links for ScatterBlock / ReduceBlock when the data plane is the CPU engine.  With
// the CUDA data plane only InitWorkers / StartAllreduce / CompleteAllreduce travel here; the payload
// moves over NVLink.
//
// Wire format: u32 length | u8 type | body (little-endian PODs; float vectors as u32 n + raw).
#pragma once
#include <chrono>
#include <functional>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "master.h"
#include "worker.h"

namespace aar {

struct Endpoint {
  std::string host;
  int port = 0;
};

namespace wire {
enum Type : uint8_t { Join = 1, Init = 2, Start = 3, Scatter = 4, Reduce = 5, Complete = 6, Heartbeat = 7, Hello = 8, Shutdown = 9 };
std::string encode(const Message& m, const std::map<int, Endpoint>* endpoints = nullptr);
std::string encodeJoin(const std::string& role, const Endpoint& dataEndpoint);
std::string encodeSimple(Type t, int32_t arg = 0);
}  // namespace wire

class Connection;  // non-blocking framed socket
struct Frame;      // view of one received frame

// Master side: accepts members, feeds a Master, forwards its sends.
class MasterNode : public WorkerPort {
 public:
  MasterNode(MasterConfig cfg, const std::string& host, int port, double autoDownAfterMs = 10000);
  ~MasterNode() override;
  int port() const { return port_; }
  Master& master() { return *master_; }
  // One loop iteration; returns number of frames handled.
  int poll(int timeoutMs);
  // Run until every round reached its quorum (then broadcast Shutdown) or `timeoutSec` elapsed. Returns finished().
  bool run(double timeoutSec);
  void broadcastShutdown();
  // Write the coordinator's state (checkpoint.h, master section) to `path` whenever the quorum clock has advanced by at
  // least `everyRounds` since the last save, and once more when run() ends.  The file is replaced atomically.
  void setCheckpoint(const std::string& path, int everyRounds = 100);
  void checkpointNow();
  void send(Address to, Message msg) override;

 private:
  struct Member;
  void acceptAll();
  void dropMember(Address a, const char* why);
  MasterConfig cfg_;
  int listenFd_ = -1, port_ = 0;
  double autoDownAfterMs_;
  std::unique_ptr<Master> master_;
  std::map<Address, std::unique_ptr<Member>> members_;
  Address nextAddr_ = 1;
  std::string ckptPath_;
  int ckptEvery_ = 100, ckptLast_ = -1;
};

// Worker side: joins the master, listens for peers, routes inbound messages to `handler`.
class WorkerNode : public WorkerPort {
 public:
  static constexpr Address kMaster = -2;
  using Handler = std::function<void(const Message&)>;
  WorkerNode(const std::string& masterHost, int masterPort, const std::string& listenHost = "127.0.0.1", int listenPort = 0,
             double heartbeatMs = 1000, const std::string& role = "worker");
  ~WorkerNode() override;
  void setHandler(Handler h) { handler_ = std::move(h); }
  int listenPort() const { return listenPort_; }
  bool shutdownRequested() const { return shutdown_; }
  bool masterAlive() const { return masterConn_ != nullptr; }
  // true: the master sent Shutdown (job done); false with shutdownRequested(): its link closed without one (coordinator gone)
  bool shutdownByMaster() const { return shutdownByMaster_; }
  int poll(int timeoutMs);
  // Loop until the master says Shutdown, the master link drops, or the timeout expires.
  bool run(double timeoutSec);
  void send(Address to, Message msg) override;  // to = worker id, or kMaster

 private:
  void dispatch(const Frame& frame);
  Connection* peer(int id);
  Handler handler_;
  std::unique_ptr<Connection> masterConn_;
  int listenFd_ = -1, listenPort_ = 0;
  double heartbeatMs_;
  std::chrono::steady_clock::time_point lastBeat_;
  std::map<int, Endpoint> endpoints_;
  std::map<int, std::unique_ptr<Connection>> outPeers_;
  std::vector<std::unique_ptr<Connection>> inPeers_;
  bool shutdown_ = false;
  bool shutdownByMaster_ = false;
};

}  // namespace aar
