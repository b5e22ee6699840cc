// Garbage on live sockets, meant to be built with -fsanitize=address,undefined: raw TCP clients throw random frames at a
// MasterNode (valid length prefixes with random types and bodies, plausible Join frames, absurd element counts), and a fake
// master does the same to a WorkerNode.  Contract: the offending link is dropped (a worker whose master speaks garbage
// shuts down as if the link had closed) -- no crash, no out-of-bounds read in the decoder, no unbounded buffering.
//
//   fuzz_wire [seed] [connections]
#include <arpa/inet.h>
#include <netinet/in.h>
#include <sys/socket.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include "../core/logging.h"
#include "../core/tcp_transport.h"

using namespace aar;

namespace {
int connectRaw(int port) {
  int fd = socket(AF_INET, SOCK_STREAM, 0);
  sockaddr_in a{};
  a.sin_family = AF_INET;
  a.sin_port = htons((uint16_t)port);
  inet_pton(AF_INET, "127.0.0.1", &a.sin_addr);
  if (connect(fd, (sockaddr*)&a, sizeof a) != 0) {
    close(fd);
    return -1;
  }
  return fd;
}

std::string randomFrames(std::mt19937& rng) {
  std::string out;
  const int frames = 1 + (int)(rng() % 6);
  for (int f = 0; f < frames; ++f) {
    std::string body;
    uint8_t type = (uint8_t)(rng() % 12);
    const unsigned kind = rng() % 10;
    if (kind < 3) {  // plausible Join
      type = 1;
      const char* roles[] = {"worker", "master", "x"};
      std::string role = roles[rng() % 3], host = "127.0.0.1";
      uint32_t n = (uint32_t)role.size();
      body.append((char*)&n, 4).append(role);
      n = (uint32_t)host.size();
      body.append((char*)&n, 4).append(host);
      int32_t port = (int32_t)(rng() % 70000);
      body.append((char*)&port, 4);
    } else if (kind == 3) {  // Init / Scatter / Reduce header followed by an absurd element count
      type = (uint8_t)(2 + rng() % 4);
      for (int i = 0; i < 8; ++i) {
        int32_t v = (int32_t)rng();
        body.append((char*)&v, 4);
      }
      uint32_t n = 0xfffffff0u;
      body.append((char*)&n, 4);
    } else {
      const int len = (int)(rng() % 96);
      for (int i = 0; i < len; ++i) body.push_back((char)rng());
    }
    uint32_t len = (uint32_t)body.size() + 1;
    if (kind == 9) len = (rng() % 2) ? 0u : 0xffffffffu;  // impossible length prefixes
    out.append((char*)&len, 4);
    out.push_back((char)type);
    out += body;
  }
  return out;
}
}  // namespace

int main(int argc, char** argv) {
  const unsigned seed = argc > 1 ? (unsigned)atoi(argv[1]) : 1;
  const int conns = argc > 2 ? atoi(argv[2]) : 300;
  std::mt19937 rng(seed);
  Logger::instance().setLevel(LogLevel::Error);

  // 1. raw clients against the coordinator
  MasterConfig cfg;
  cfg.totalWorkers = 3;
  cfg.dataSize = 10;
  cfg.maxChunkSize = 2;
  cfg.maxRound = 5;
  MasterNode node(cfg, "127.0.0.1", 0);
  std::vector<int> open;
  for (int k = 0; k < conns; ++k) {
    int fd = connectRaw(node.port());
    if (fd < 0) continue;
    std::string bytes = randomFrames(rng);
    (void)!write(fd, bytes.data(), bytes.size());
    open.push_back(fd);
    for (int i = 0; i < 3; ++i) node.poll(1);
    if (open.size() > 40) {
      close(open.front());
      open.erase(open.begin());
    }
  }
  for (int fd : open) close(fd);
  for (int i = 0; i < 5; ++i) node.poll(1);

  // 2. a fake master that speaks garbage to a worker node
  int threw = 0, survived = 0;
  for (int k = 0; k < conns / 10 + 1; ++k) {
    int lfd = socket(AF_INET, SOCK_STREAM, 0);
    sockaddr_in a{};
    a.sin_family = AF_INET;
    a.sin_port = 0;
    inet_pton(AF_INET, "127.0.0.1", &a.sin_addr);
    if (bind(lfd, (sockaddr*)&a, sizeof a) != 0 || listen(lfd, 4) != 0) return 2;
    socklen_t len = sizeof a;
    getsockname(lfd, (sockaddr*)&a, &len);
    const int port = ntohs(a.sin_port);
    std::string bytes = randomFrames(rng);
    std::thread fake([&] {
      int c = accept(lfd, nullptr, nullptr);
      if (c >= 0) {
        (void)!write(c, bytes.data(), bytes.size());
        std::this_thread::sleep_for(std::chrono::milliseconds(20));
        close(c);
      }
    });
    try {
      WorkerNode w("127.0.0.1", port);
      w.setHandler([](const Message&) {});
      for (int i = 0; i < 20 && !w.shutdownRequested(); ++i) w.poll(5);
      ++survived;
    } catch (const std::exception&) {
      ++threw;
    }
    fake.join();
    close(lfd);
  }
  printf("fuzz_wire seed=%u: %d raw connections against the master, workers: %d survived, %d threw\n", seed, conns, survived, threw);
  return 0;
}
