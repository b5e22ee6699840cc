// See tcp_transport.h.
#include "tcp_transport.h"

#include <arpa/inet.h>
#include <errno.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <signal.h>
#include <sys/socket.h>
#include <unistd.h>

#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <thread>

#include "checkpoint.h"
#include "logging.h"

namespace aar {

// One frame body (type byte + payload) as it sits in a connection's receive buffer; valid until the connection is read
// or asked for the next frame again.
struct Frame {
  const char* data = nullptr;
  size_t size = 0;
};

namespace {

using Clock = std::chrono::steady_clock;

// Malformed bytes from the network.  Its own type so that the poll loops can drop the offending link without swallowing
// errors raised by the message handlers (e.g. the input-size check of AllreduceWorker.scala:174-176).
struct WireError : std::runtime_error {
  using std::runtime_error::runtime_error;
};

struct Writer {
  std::string b = std::string(5, '\0');  // u32 length + type byte, filled in by frame()
  template <class T>
  void pod(T v) { b.append(reinterpret_cast<const char*>(&v), sizeof(T)); }
  void str(const std::string& s) {
    pod<uint32_t>((uint32_t)s.size());
    b += s;
  }
  void floats(const std::vector<float>& v) {
    b.reserve(b.size() + 4 + v.size() * sizeof(float));
    pod<uint32_t>((uint32_t)v.size());
    b.append(reinterpret_cast<const char*>(v.data()), v.size() * sizeof(float));
  }
  std::string frame(uint8_t type) {
    uint32_t len = (uint32_t)b.size() - 4;  // type byte + payload
    std::memcpy(&b[0], &len, 4);
    b[4] = (char)type;
    return std::move(b);
  }
};

struct Reader {
  const char* b;
  size_t n, i;
  explicit Reader(const Frame& f, size_t start = 0) : b(f.data), n(f.size), i(start) {}
  template <class T>
  T pod() {
    if (i + sizeof(T) > n) throw WireError("wire: truncated frame");
    T v;
    std::memcpy(&v, b + i, sizeof(T));
    i += sizeof(T);
    return v;
  }
  std::string str() {
    uint32_t len = pod<uint32_t>();
    if (i + len > n) throw WireError("wire: truncated string");
    std::string s(b + i, len);
    i += len;
    return s;
  }
  std::vector<float> floats() {
    uint32_t len = pod<uint32_t>();
    if (i + (size_t)len * 4 > n) throw WireError("wire: truncated vector");
    const float* p = reinterpret_cast<const float*>(b + i);  // memcpy below: the payload need not be aligned
    std::vector<float> v(len);
    std::memcpy(v.data(), p, (size_t)len * 4);
    i += (size_t)len * 4;
    return v;
  }
};

void setNonBlocking(int fd) {
  int fl = fcntl(fd, F_GETFL, 0);
  fcntl(fd, F_SETFL, fl | O_NONBLOCK);
  int one = 1;
  setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
}

int listenOn(const std::string& host, int port, int* boundPort) {
  int fd = socket(AF_INET, SOCK_STREAM, 0);
  if (fd < 0) throw std::runtime_error("socket() failed");
  int one = 1;
  setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
  sockaddr_in a{};
  a.sin_family = AF_INET;
  a.sin_port = htons((uint16_t)port);
  if (inet_pton(AF_INET, host.c_str(), &a.sin_addr) != 1) a.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
  if (bind(fd, (sockaddr*)&a, sizeof a) != 0) {
    close(fd);
    throw std::runtime_error("bind(" + host + ":" + std::to_string(port) + ") failed: " + strerror(errno));
  }
  if (listen(fd, 128) != 0) {
    close(fd);
    throw std::runtime_error("listen() failed");
  }
  socklen_t len = sizeof a;
  getsockname(fd, (sockaddr*)&a, &len);
  *boundPort = ntohs(a.sin_port);
  setNonBlocking(fd);
  return fd;
}

int connectTo(const std::string& host, int port, double timeoutSec) {
  auto deadline = Clock::now() + std::chrono::duration<double>(timeoutSec);
  for (;;) {
    int fd = socket(AF_INET, SOCK_STREAM, 0);
    if (fd < 0) throw std::runtime_error("socket() failed");
    sockaddr_in a{};
    a.sin_family = AF_INET;
    a.sin_port = htons((uint16_t)port);
    if (inet_pton(AF_INET, host.c_str(), &a.sin_addr) != 1) a.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
    if (connect(fd, (sockaddr*)&a, sizeof a) == 0) {
      setNonBlocking(fd);
      return fd;
    }
    close(fd);
    if (Clock::now() > deadline) throw std::runtime_error("connect(" + host + ":" + std::to_string(port) + ") timed out: " + strerror(errno));
    std::this_thread::sleep_for(std::chrono::milliseconds(50));
  }
}

struct IgnoreSigpipe {
  IgnoreSigpipe() { signal(SIGPIPE, SIG_IGN); }
} ignoreSigpipe;

}  // namespace

class Connection {
 public:
  explicit Connection(int fd) : fd_(fd) {}
  ~Connection() {
    if (fd_ >= 0) close(fd_);
  }
  int fd() const { return fd_; }
  bool open() const { return fd_ >= 0; }
  bool wantsWrite() const { return outPos_ < out_.size(); }
  void queue(std::string frame) {
    if (outPos_ == out_.size()) {  // nothing pending: adopt the frame's buffer instead of copying it
      out_ = std::move(frame);
      outPos_ = 0;
    } else {
      if (outPos_ > (4u << 20)) {  // reclaim the sent prefix now and then, never once per send()
        out_.erase(0, outPos_);
        outPos_ = 0;
      }
      out_ += frame;
    }
    flush();
  }
  // false when the peer closed or an error occurred
  bool flush() {
    while (outPos_ < out_.size()) {
      ssize_t n = ::send(fd_, out_.data() + outPos_, out_.size() - outPos_, MSG_NOSIGNAL);
      if (n > 0) {
        outPos_ += (size_t)n;
      } else if (n < 0 && (errno == EAGAIN || errno == EWOULDBLOCK)) {
        return true;
      } else if (n < 0 && errno == EINTR) {
        continue;
      } else {
        return false;
      }
    }
    out_.clear();
    outPos_ = 0;
    return true;
  }
  bool readAvailable() {
    size_t got = 0;
    for (;;) {
      if (got >= (32u << 20)) return true;  // let the frames be consumed; poll() is level-triggered and comes back
      // receive straight into the frame buffer (no bounce buffer, no zero-fill): keep at least 64 KiB of room, and once a
      // frame header is known, room for the whole frame
      size_t want = 65536;
      if (len_ - pos_ >= 4) {
        uint32_t flen;
        std::memcpy(&flen, in_.get() + pos_, 4);
        if (flen <= kMaxFrameBytes && (size_t)flen + 4 > len_ - pos_) want = std::max(want, (size_t)flen + 4 - (len_ - pos_));
      }
      reserve(want);
      ssize_t n = ::recv(fd_, in_.get() + len_, cap_ - len_, 0);
      if (n > 0) {
        len_ += (size_t)n;
        got += (size_t)n;
      } else if (n == 0) {
        return false;
      } else if (errno == EAGAIN || errno == EWOULDBLOCK) {
        return true;
      } else if (errno == EINTR) {
        continue;
      } else {
        return false;
      }
    }
  }
  // Pops one complete frame body (type byte + payload) as a view into the receive buffer.
  bool next(Frame& body) {
    if (len_ - pos_ < 4) {
      if (pos_ == len_) pos_ = len_ = 0;
      return false;
    }
    uint32_t len;
    std::memcpy(&len, in_.get() + pos_, 4);
    if (len == 0 || len > kMaxFrameBytes) throw WireError("wire: frame length " + std::to_string(len) + " out of range");
    if (len_ - pos_ - 4 < len) return false;
    body.data = in_.get() + pos_ + 4;
    body.size = len;
    pos_ += 4 + (size_t)len;
    return true;
  }
  // Blocking drain (used before close so queued frames are not lost).
  void drain(double sec) {
    auto deadline = Clock::now() + std::chrono::duration<double>(sec);
    while (wantsWrite() && Clock::now() < deadline) {
      pollfd p{fd_, POLLOUT, 0};
      ::poll(&p, 1, 20);
      if (!flush()) return;
    }
  }
  // Orderly goodbye: no more writes from this side, and whatever the peer still sends is read and thrown away.  Closing a
  // socket with unread input makes the kernel answer with RST, and an RST lets the peer's kernel discard data it has not
  // read yet -- such as the Shutdown frame just sent.  Returns false once the peer has closed its side.
  void finishWrites() { ::shutdown(fd_, SHUT_WR); }
  bool discardInput() {
    char buf[65536];
    for (;;) {
      ssize_t n = ::recv(fd_, buf, sizeof buf, 0);
      if (n > 0) continue;
      if (n == 0) return false;
      if (errno == EINTR) continue;
      return errno == EAGAIN || errno == EWOULDBLOCK;
    }
  }
  Clock::time_point lastSeen = Clock::now();

 private:
  // Make room for `want` more bytes behind len_: slide the unread tail to the front when that is enough (it is at most one
  // partial frame), otherwise grow geometrically.
  void reserve(size_t want) {
    if (cap_ - len_ >= want) return;
    const size_t live = len_ - pos_;
    if (pos_ > 0 && cap_ - live >= want) {
      std::memmove(in_.get(), in_.get() + pos_, live);
    } else {
      size_t ncap = std::max(cap_ * 2, live + want);
      std::unique_ptr<char[]> nb(new char[ncap]);
      if (live) std::memcpy(nb.get(), in_.get() + pos_, live);
      in_ = std::move(nb);
      cap_ = ncap;
    }
    pos_ = 0;
    len_ = live;
  }
  // One chunk of floats plus a header; nothing legitimate is larger, and a garbage length must not make us buffer 4 GB.
  static constexpr uint32_t kMaxFrameBytes = (1u << 30) + 64;
  int fd_;
  std::unique_ptr<char[]> in_;  // receive buffer: [pos_, len_) is unread, capacity cap_
  size_t cap_ = 0, len_ = 0, pos_ = 0;
  std::string out_;
  size_t outPos_ = 0;
};

namespace wire {

std::string encode(const Message& m, const std::map<int, Endpoint>* endpoints) {
  Writer w;
  switch (m.index()) {
    case 0: {
      const auto& i = std::get<InitWorkers>(m);
      w.pod<int32_t>(i.destId);
      w.pod<float>(i.thReduce);
      w.pod<float>(i.thComplete);
      w.pod<int32_t>(i.maxLag);
      w.pod<int32_t>(i.dataSize);
      w.pod<int32_t>(i.maxChunkSize);
      w.pod<int32_t>(i.generation);
      w.pod<int32_t>(i.startRound);
      w.pod<uint32_t>((uint32_t)i.workers.size());
      for (auto& kv : i.workers) {
        w.pod<int32_t>(kv.first);
        Endpoint e;
        if (endpoints) {
          auto it = endpoints->find(kv.first);
          if (it != endpoints->end()) e = it->second;
        }
        w.str(e.host);
        w.pod<int32_t>(e.port);
      }
      return w.frame(Init);
    }
    case 1:
      w.pod<int32_t>(std::get<StartAllreduce>(m).round);
      w.pod<int32_t>(std::get<StartAllreduce>(m).generation);
      return w.frame(Start);
    case 2: {
      const auto& s = std::get<ScatterBlock>(m);
      w.pod<int32_t>(s.srcId);
      w.pod<int32_t>(s.destId);
      w.pod<int32_t>(s.chunkId);
      w.pod<int32_t>(s.round);
      w.pod<int32_t>(s.generation);
      w.floats(s.value);
      return w.frame(Scatter);
    }
    case 3: {
      const auto& r = std::get<ReduceBlock>(m);
      w.pod<int32_t>(r.srcId);
      w.pod<int32_t>(r.destId);
      w.pod<int32_t>(r.chunkId);
      w.pod<int32_t>(r.round);
      w.pod<int32_t>(r.count);
      w.pod<int32_t>(r.generation);
      w.floats(r.value);
      return w.frame(Reduce);
    }
    default: {
      const auto& c = std::get<CompleteAllreduce>(m);
      w.pod<int32_t>(c.srcId);
      w.pod<int32_t>(c.round);
      w.pod<int32_t>(c.generation);
      return w.frame(Complete);
    }
  }
}

std::string encodeJoin(const std::string& role, const Endpoint& e) {
  Writer w;
  w.str(role);
  w.str(e.host);
  w.pod<int32_t>(e.port);
  return w.frame(Join);
}

std::string encodeSimple(Type t, int32_t arg) {
  Writer w;
  w.pod<int32_t>(arg);
  return w.frame(t);
}

}  // namespace wire

namespace {

// Decodes a frame body into a Message. Returns false for non-message frames (join/heartbeat/...).
bool decode(const Frame& body, Message& out, std::map<int, Endpoint>* endpoints) {
  if (body.size == 0) return false;
  Reader r(body, 1);
  switch ((uint8_t)body.data[0]) {
    case wire::Init: {
      InitWorkers i;
      i.destId = r.pod<int32_t>();
      i.thReduce = r.pod<float>();
      i.thComplete = r.pod<float>();
      i.maxLag = r.pod<int32_t>();
      i.dataSize = r.pod<int32_t>();
      i.maxChunkSize = r.pod<int32_t>();
      i.generation = r.pod<int32_t>();
      i.startRound = r.pod<int32_t>();
      uint32_t n = r.pod<uint32_t>();
      if (endpoints) endpoints->clear();
      for (uint32_t k = 0; k < n; ++k) {
        int id = r.pod<int32_t>();
        Endpoint e;
        e.host = r.str();
        e.port = r.pod<int32_t>();
        i.workers[id] = id;  // on a worker node the address of peer `id` is its id
        if (endpoints) (*endpoints)[id] = e;
      }
      i.master = WorkerNode::kMaster;
      out = std::move(i);
      return true;
    }
    case wire::Start:
    {
      StartAllreduce st;
      st.round = r.pod<int32_t>();
      st.generation = r.pod<int32_t>();
      out = st;
    }
      return true;
    case wire::Scatter: {
      ScatterBlock s;
      s.srcId = r.pod<int32_t>();
      s.destId = r.pod<int32_t>();
      s.chunkId = r.pod<int32_t>();
      s.round = r.pod<int32_t>();
      s.generation = r.pod<int32_t>();
      s.value = r.floats();
      out = std::move(s);
      return true;
    }
    case wire::Reduce: {
      ReduceBlock b;
      b.srcId = r.pod<int32_t>();
      b.destId = r.pod<int32_t>();
      b.chunkId = r.pod<int32_t>();
      b.round = r.pod<int32_t>();
      b.count = r.pod<int32_t>();
      b.generation = r.pod<int32_t>();
      b.value = r.floats();
      out = std::move(b);
      return true;
    }
    case wire::Complete: {
      CompleteAllreduce c;
      c.srcId = r.pod<int32_t>();
      c.round = r.pod<int32_t>();
      c.generation = r.pod<int32_t>();
      out = c;
      return true;
    }
    default:
      return false;
  }
}

}  // namespace

// ------------------------------------------------------------------ MasterNode

struct MasterNode::Member {
  std::unique_ptr<Connection> conn;
  std::string role;
  Endpoint data;
  bool joined = false;
};

MasterNode::MasterNode(MasterConfig cfg, const std::string& host, int port, double autoDownAfterMs)
    : cfg_(cfg), autoDownAfterMs_(autoDownAfterMs) {
  listenFd_ = listenOn(host, port, &port_);
  master_ = std::make_unique<Master>(cfg_, this, 0);
  AAR_INFO("master: listening on %s:%d (totalWorkers=%d dataSize=%d maxChunkSize=%d)", host.c_str(), port_, cfg.totalWorkers, cfg.dataSize,
           cfg.maxChunkSize);
}

MasterNode::~MasterNode() {
  for (auto& kv : members_)
    if (kv.second->conn) kv.second->conn->drain(0.5);
  if (listenFd_ >= 0) close(listenFd_);
}

void MasterNode::acceptAll() {
  for (;;) {
    int fd = accept(listenFd_, nullptr, nullptr);
    if (fd < 0) return;
    setNonBlocking(fd);
    auto m = std::make_unique<Member>();
    m->conn = std::make_unique<Connection>(fd);
    members_[nextAddr_++] = std::move(m);
  }
}

void MasterNode::dropMember(Address a, const char* why) {
  auto it = members_.find(a);
  if (it == members_.end()) return;
  AAR_INFO("master: member %ld down (%s)", (long)a, why);
  bool wasWorker = it->second->joined && it->second->role == "worker";
  members_.erase(it);
  if (wasWorker) master_->terminated(a);
}

void MasterNode::send(Address to, Message msg) {
  auto it = members_.find(to);
  if (it == members_.end()) return;  // dead letter
  std::map<int, Endpoint> eps;
  if (std::holds_alternative<InitWorkers>(msg)) {
    for (auto& kv : std::get<InitWorkers>(msg).workers) {
      auto mt = members_.find(kv.second);
      if (mt != members_.end()) eps[kv.first] = mt->second->data;
    }
  }
  it->second->conn->queue(wire::encode(msg, &eps));
}

int MasterNode::poll(int timeoutMs) {
  std::vector<pollfd> fds;
  std::vector<Address> addrs;
  fds.push_back({listenFd_, POLLIN, 0});
  for (auto& kv : members_) {
    short ev = POLLIN;
    if (kv.second->conn->wantsWrite()) ev |= POLLOUT;
    fds.push_back({kv.second->conn->fd(), ev, 0});
    addrs.push_back(kv.first);
  }
  int rc = ::poll(fds.data(), fds.size(), timeoutMs);
  int handled = 0;
  if (rc > 0) {
    if (fds[0].revents & POLLIN) acceptAll();
    for (size_t k = 1; k < fds.size(); ++k) {
      Address a = addrs[k - 1];
      auto it = members_.find(a);
      if (it == members_.end()) continue;
      Member& m = *it->second;
      bool alive = true;
      if (fds[k].revents & POLLOUT) alive = m.conn->flush();
      if (alive && (fds[k].revents & (POLLIN | POLLHUP | POLLERR))) alive = m.conn->readAvailable();
      Frame body;
      try {  // a peer that speaks garbage is disconnected; it must not take the coordinator down
        while (m.conn->next(body)) {
          ++handled;
          m.conn->lastSeen = Clock::now();
          uint8_t type = (uint8_t)body.data[0];
          if (type == wire::Join) {
            Reader r(body, 1);
            m.role = r.str();
            m.data.host = r.str();
            m.data.port = r.pod<int32_t>();
            m.joined = true;
            master_->memberUp(a, m.role);
          } else if (type == wire::Heartbeat) {
            // lastSeen already refreshed
          } else {
            Message msg;
            if (decode(body, msg, nullptr)) master_->receive(msg);
          }
          if (members_.find(a) == members_.end()) break;
        }
      } catch (const WireError& e) {
        AAR_WARN("master: protocol error from member %ld: %s", (long)a, e.what());
        dropMember(a, "protocol error");
        continue;
      }
      if (!alive) dropMember(a, "connection closed");
    }
  }
  // phi-accrual stand-in: a fixed silence window (application.conf:20 auto-down-unreachable-after)
  auto now = Clock::now();
  std::vector<Address> dead;
  for (auto& kv : members_)
    if (kv.second->joined && std::chrono::duration<double, std::milli>(now - kv.second->conn->lastSeen).count() > autoDownAfterMs_) dead.push_back(kv.first);
  for (Address a : dead) dropMember(a, "unreachable");
  return handled;
}

void MasterNode::broadcastShutdown() {
  for (auto& kv : members_) {
    kv.second->conn->queue(wire::encodeSimple(wire::Shutdown));
    kv.second->conn->drain(1.0);
    kv.second->conn->finishWrites();
  }
  // Workers that lag (thAllreduce < 1) are still sending completions.  Keep reading until every peer has seen the Shutdown and
  // closed its side (or two seconds have passed): closing with unread input would turn into an RST that can cost the peer
  // the Shutdown frame itself -- it would then report a lost master although the job finished.
  std::map<Address, bool> open;
  for (auto& kv : members_) open[kv.first] = true;
  auto deadline = Clock::now() + std::chrono::seconds(2);
  size_t left = open.size();
  while (left > 0 && Clock::now() < deadline) {
    std::vector<pollfd> fds;
    std::vector<Address> addrs;
    for (auto& kv : members_)
      if (open[kv.first]) {
        fds.push_back({kv.second->conn->fd(), POLLIN, 0});
        addrs.push_back(kv.first);
      }
    if (::poll(fds.data(), fds.size(), 50) <= 0) continue;
    for (size_t k = 0; k < fds.size(); ++k)
      if ((fds[k].revents & (POLLIN | POLLHUP | POLLERR)) && !members_[addrs[k]]->conn->discardInput()) {
        open[addrs[k]] = false;
        --left;
      }
  }
}

void MasterNode::setCheckpoint(const std::string& path, int everyRounds) {
  ckptPath_ = path;
  ckptEvery_ = everyRounds < 1 ? 1 : everyRounds;
  ckptLast_ = master_->quorumRound();
}

void MasterNode::checkpointNow() {
  if (ckptPath_.empty()) return;
  const std::string tmp = ckptPath_ + ".tmp";
  saveMasterCheckpoint(tmp, *master_);
  if (std::rename(tmp.c_str(), ckptPath_.c_str()) != 0) throw std::runtime_error("checkpoint: cannot replace " + ckptPath_);
  ckptLast_ = master_->quorumRound();
}

bool MasterNode::run(double timeoutSec) {
  auto deadline = Clock::now() + std::chrono::duration<double>(timeoutSec);
  while (!master_->finished() && Clock::now() < deadline) {
    poll(50);
    if (!ckptPath_.empty() && master_->quorumRound() >= ckptLast_ + ckptEvery_) checkpointNow();
  }
  bool ok = master_->finished();
  if (master_->started()) checkpointNow();
  broadcastShutdown();
  return ok;
}

// ------------------------------------------------------------------ WorkerNode

WorkerNode::WorkerNode(const std::string& masterHost, int masterPort, const std::string& listenHost, int listenPort, double heartbeatMs,
                       const std::string& role)
    : heartbeatMs_(heartbeatMs), lastBeat_(Clock::now()) {
  listenFd_ = listenOn(listenHost, listenPort, &listenPort_);
  masterConn_ = std::make_unique<Connection>(connectTo(masterHost, masterPort, 30.0));
  masterConn_->queue(wire::encodeJoin(role, Endpoint{listenHost, listenPort_}));
}

WorkerNode::~WorkerNode() {
  if (masterConn_) masterConn_->drain(0.5);
  for (auto& kv : outPeers_) kv.second->drain(0.5);
  if (listenFd_ >= 0) close(listenFd_);
}

Connection* WorkerNode::peer(int id) {
  auto it = outPeers_.find(id);
  if (it != outPeers_.end()) return it->second.get();
  auto ep = endpoints_.find(id);
  if (ep == endpoints_.end()) return nullptr;
  try {
    // Every worker listens before it joins the master, so by the time InitWorkers names a peer that peer accepts connections:
    // a refusal means it is gone.  A short patience (not the 10 s of a start-up race) keeps the survivors of a crash moving.
    auto c = std::make_unique<Connection>(connectTo(ep->second.host, ep->second.port, 1.0));
    Connection* raw = c.get();
    outPeers_[id] = std::move(c);
    return raw;
  } catch (const std::runtime_error& e) {
    // A peer that cannot be reached is a dead letter box, as with Akka (the thresholds absorb its absence); forget the
    // endpoint so that later sends do not pay the connect timeout again.  The next InitWorkers brings a fresh table.
    AAR_WARN("worker: peer %d unreachable (%s); its messages are dropped", id, e.what());
    endpoints_.erase(ep);
    return nullptr;
  }
}

void WorkerNode::send(Address to, Message msg) {
  if (to == kMaster) {
    if (masterConn_) masterConn_->queue(wire::encode(msg));
    return;
  }
  Connection* c = peer((int)to);
  if (!c) return;  // dead letter
  c->queue(wire::encode(msg));
}

void WorkerNode::dispatch(const Frame& body) {
  uint8_t type = (uint8_t)body.data[0];
  if (type == wire::Shutdown) {
    shutdown_ = true;
    shutdownByMaster_ = true;
    if (masterConn_) {  // tell the master we are done talking: it waits for this before it closes (see broadcastShutdown)
      masterConn_->drain(0.2);
      masterConn_->finishWrites();
    }
    return;
  }
  Message msg;
  std::map<int, Endpoint> eps;
  if (!decode(body, msg, &eps)) return;
  if (std::holds_alternative<InitWorkers>(msg)) {
    endpoints_ = eps;
    outPeers_.clear();  // re-initialisation: peers may have changed
  }
  if (handler_) handler_(msg);
}

int WorkerNode::poll(int timeoutMs) {
  std::vector<pollfd> fds;
  fds.push_back({listenFd_, POLLIN, 0});
  if (masterConn_) fds.push_back({masterConn_->fd(), (short)(POLLIN | (masterConn_->wantsWrite() ? POLLOUT : 0)), 0});
  size_t inBase = fds.size();
  for (auto& c : inPeers_) fds.push_back({c->fd(), POLLIN, 0});
  size_t outBase = fds.size();
  std::vector<int> outIds;
  for (auto& kv : outPeers_) {
    fds.push_back({kv.second->fd(), (short)(kv.second->wantsWrite() ? POLLOUT : 0), 0});
    outIds.push_back(kv.first);
  }
  int rc = ::poll(fds.data(), fds.size(), timeoutMs);
  int handled = 0;
  Frame body;
  if (rc > 0) {
    if (fds[0].revents & POLLIN) {
      for (;;) {
        int fd = accept(listenFd_, nullptr, nullptr);
        if (fd < 0) break;
        setNonBlocking(fd);
        inPeers_.push_back(std::make_unique<Connection>(fd));
      }
    }
    if (masterConn_ && inBase == 2) {
      bool alive = true;
      if (fds[1].revents & POLLOUT) alive = masterConn_->flush();
      if (alive && (fds[1].revents & (POLLIN | POLLHUP | POLLERR))) alive = masterConn_->readAvailable();
      try {
        while (masterConn_->next(body)) {
          ++handled;
          dispatch(body);
        }
      } catch (const WireError& e) {  // a coordinator that speaks garbage is as good as gone (handler errors still propagate)
        AAR_WARN("worker: protocol error on the master link: %s", e.what());
        alive = false;
      }
      if (!alive) {
        AAR_WARN("worker: master link closed");
        masterConn_.reset();
        shutdown_ = true;
      }
    }
    size_t nIn = outBase - inBase;
    std::vector<size_t> deadIn;
    for (size_t k = 0; k < nIn; ++k) {
      if (!(fds[inBase + k].revents & (POLLIN | POLLHUP | POLLERR))) continue;
      Connection& c = *inPeers_[k];
      bool alive = c.readAvailable();
      try {
        while (c.next(body)) {
          ++handled;
          dispatch(body);
        }
      } catch (const WireError& e) {
        AAR_WARN("worker: protocol error on a peer link: %s", e.what());
        alive = false;
      } catch (const ProtocolError& e) {  // well-formed frame, impossible content: that peer's link goes, this worker stays
        AAR_WARN("worker: peer sent an impossible message, dropping its link: %s", e.what());
        alive = false;
      }
      if (!alive) deadIn.push_back(k);
    }
    for (auto it = deadIn.rbegin(); it != deadIn.rend(); ++it) inPeers_.erase(inPeers_.begin() + *it);
    for (size_t k = 0; k < outIds.size(); ++k) {
      if (!(fds[outBase + k].revents & (POLLOUT | POLLHUP | POLLERR))) continue;
      auto it = outPeers_.find(outIds[k]);
      if (it != outPeers_.end() && !it->second->flush()) outPeers_.erase(it);
    }
  }
  auto now = Clock::now();
  if (masterConn_ && std::chrono::duration<double, std::milli>(now - lastBeat_).count() >= heartbeatMs_) {
    masterConn_->queue(wire::encodeSimple(wire::Heartbeat));
    lastBeat_ = now;
  }
  return handled;
}

bool WorkerNode::run(double timeoutSec) {
  auto deadline = Clock::now() + std::chrono::duration<double>(timeoutSec);
  while (!shutdown_ && Clock::now() < deadline) poll(50);
  return shutdown_;
}

}  // namespace aar
