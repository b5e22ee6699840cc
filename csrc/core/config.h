// application.conf-equivalent parser (a HOCON subset) and the typed allreduce settings.
//
// The reference reads src/main/resources/application.conf:1-34 through Typesafe Config and layers
// `parseString(port override).withFallback(role).withFallback(load())` on top
// (AllreduceMaster.scala:116-118, AllreduceWorker.scala:276-278).  The algorithm knobs are hard-coded
// in AllreduceMaster.scala:105-109; README.md:5,13 says they were meant to move into the file.  Here
// they live in an `allreduce { ... }` block.
//
// Supported syntax: nested `{}` blocks, dotted keys, `=` or `:` separators, optional commas,
// `#` and `//` comments, quoted / unquoted scalars, `[a, b, ...]` lists (may span lines),
// `${path}` and `${?path}` substitution (config path first, then environment; `user.dir` = cwd),
// durations (`10s`, `250ms`), booleans (`true/false/on/off/yes/no`).
#pragma once
#include <map>
#include <string>
#include <vector>

#include "master.h"

namespace aar {

class Config {
 public:
  struct Value {
    bool isList = false;
    std::string scalar;
    std::vector<std::string> list;
  };

  static Config parseString(const std::string& text);
  static Config parseFile(const std::string& path);

  // Keys present in *this win; missing keys come from `fallback` (Typesafe's withFallback).
  Config withFallback(const Config& fallback) const;

  bool has(const std::string& path) const { return values_.count(path) != 0; }
  std::string getString(const std::string& path) const;
  std::string getString(const std::string& path, const std::string& dflt) const { return has(path) ? getString(path) : dflt; }
  long getInt(const std::string& path) const;
  long getInt(const std::string& path, long dflt) const { return has(path) ? getInt(path) : dflt; }
  double getDouble(const std::string& path) const;
  double getDouble(const std::string& path, double dflt) const { return has(path) ? getDouble(path) : dflt; }
  bool getBool(const std::string& path) const;
  bool getBool(const std::string& path, bool dflt) const { return has(path) ? getBool(path) : dflt; }
  std::vector<std::string> getList(const std::string& path) const;
  // Duration in milliseconds ("10s", "250ms", "2m", plain number = ms).
  double getDurationMs(const std::string& path) const;
  double getDurationMs(const std::string& path, double dflt) const { return has(path) ? getDurationMs(path) : dflt; }

  void set(const std::string& path, const std::string& scalar);
  void setList(const std::string& path, const std::vector<std::string>& list);
  std::vector<std::string> keys() const;
  // Flat `path = value` rendering that parseString() reads back (used by checkpoints).
  std::string render() const;

 private:
  void resolveSubstitutions();
  std::map<std::string, Value> values_;
};

// Typed view of the `allreduce { }` block plus transport settings.
struct AllreduceSettings {
  MasterConfig master;           // total-workers, th-*, max-lag, data-size, max-round, max-chunk-size, pipeline-depth
  std::string dtype = "fp32";    // wire/output dtype of the CUDA data plane: fp32 | bf16
  std::string algo = "auto";     // auto | twoshot | pipelined | oneshot | nvls
  bool normalize = false;        // divide by the actual contributor count (the intent behind AllReduceOutput.count)
  std::string quorumPolicy = "all-arrived";  // all-arrived | exact
  double spinTimeoutUs = 0;      // device-side liveness guard; 0 = wait for the host doorbell only
  std::string hostname = "127.0.0.1";
  int port = 2551;               // master control port (seed node)
  double heartbeatMs = 1000;
  double autoDownAfterMs = 10000;  // akka.cluster.auto-down-unreachable-after (application.conf:20)
  std::string loglevel = "INFO";

  static AllreduceSettings fromConfig(const Config& c);
};

}  // namespace aar
