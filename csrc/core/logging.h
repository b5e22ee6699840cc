// Levelled logger for the host runtime.  Replaces the reference's ActorLogging calls
// (e.g. AllreduceWorker.scala:75-82, AllreduceMaster.scala:39-63; level from application.conf:22).
// Payload vectors are never formatted here: the reference prints whole arrays at INFO
// (AllreduceWorker.scala:190), which is a throughput killer on a hot path.
#pragma once
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <ctime>
#include <functional>
#include <mutex>
#include <string>

namespace aar {

enum class LogLevel : int { Debug = 0, Info = 1, Warning = 2, Error = 3, Off = 4 };

class Logger {
 public:
  static Logger& instance() {
    static Logger l;
    return l;
  }
  void setLevel(LogLevel l) { level_.store((int)l); }
  LogLevel level() const { return (LogLevel)level_.load(); }
  static LogLevel parse(const std::string& s) {
    std::string u;
    for (char c : s) u += (char)toupper((unsigned char)c);
    if (u == "DEBUG") return LogLevel::Debug;
    if (u == "INFO") return LogLevel::Info;
    if (u == "WARNING" || u == "WARN") return LogLevel::Warning;
    if (u == "ERROR") return LogLevel::Error;
    if (u == "OFF") return LogLevel::Off;
    return LogLevel::Info;
  }
  // Optional sink (tests capture lines); default is stderr.
  void setSink(std::function<void(LogLevel, const std::string&)> s) {
    std::lock_guard<std::mutex> g(mu_);
    sink_ = std::move(s);
  }
  bool enabled(LogLevel l) const { return (int)l >= level_.load(); }
  void log(LogLevel l, const char* fmt, ...) __attribute__((format(printf, 3, 4))) {
    if (!enabled(l)) return;
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    std::lock_guard<std::mutex> g(mu_);
    if (sink_) {
      sink_(l, buf);
      return;
    }
    static const char* names[] = {"DEBUG", "INFO", "WARN", "ERROR"};
    timespec ts;
    clock_gettime(CLOCK_REALTIME, &ts);
    fprintf(stderr, "[%ld.%03ld] [%s] %s\n", (long)ts.tv_sec, ts.tv_nsec / 1000000, names[(int)l], buf);
  }

 private:
  Logger() {
    const char* e = getenv("AAR_LOGLEVEL");
    level_.store((int)(e ? parse(e) : LogLevel::Warning));
  }
  std::atomic<int> level_{(int)LogLevel::Warning};
  std::mutex mu_;
  std::function<void(LogLevel, const std::string&)> sink_;
};

#define AAR_LOG(lvl, ...)                                                       \
  do {                                                                          \
    if (::aar::Logger::instance().enabled(lvl)) ::aar::Logger::instance().log(lvl, __VA_ARGS__); \
  } while (0)
#define AAR_DEBUG(...) AAR_LOG(::aar::LogLevel::Debug, __VA_ARGS__)
#define AAR_INFO(...) AAR_LOG(::aar::LogLevel::Info, __VA_ARGS__)
#define AAR_WARN(...) AAR_LOG(::aar::LogLevel::Warning, __VA_ARGS__)
#define AAR_ERROR(...) AAR_LOG(::aar::LogLevel::Error, __VA_ARGS__)

}  // namespace aar
