// Shared argument handling of the two CLI mains.
#pragma once
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../core/config.h"
#include "../core/logging.h"

namespace aar {

struct CliArgs {
  bool help = false;
  std::vector<std::string> positional;
  std::string conf;
  std::vector<std::string> overrides;  // --set path=value
  double timeoutSec = 3600;
  std::string checkpoint;  // master only: state file, replaced atomically every `checkpointEvery` quorum rounds and at exit
  int checkpointEvery = 100;
  bool resume = false;     // master only: continue the round numbering of `checkpoint` if the file exists
};

inline CliArgs parseCli(int argc, char** argv) {
  CliArgs a;
  for (int i = 1; i < argc; ++i) {
    std::string s = argv[i];
    if (s == "--help" || s == "-h") {
      a.help = true;
    } else if (s == "--conf" && i + 1 < argc) {
      a.conf = argv[++i];
    } else if (s == "--set" && i + 1 < argc) {
      a.overrides.push_back(argv[++i]);
    } else if (s == "--checkpoint" && i + 1 < argc) {
      a.checkpoint = argv[++i];
    } else if (s == "--checkpoint-every" && i + 1 < argc) {
      a.checkpointEvery = atoi(argv[++i]);
    } else if (s == "--resume") {
      a.resume = true;
    } else if (s == "--timeout" && i + 1 < argc) {
      a.timeoutSec = atof(argv[++i]);
    } else {
      a.positional.push_back(s);
    }
  }
  return a;
}

// Layering like the reference mains: explicit overrides, then the role, then the file
// (AllreduceMaster.scala:116-118, AllreduceWorker.scala:276-278).
inline Config loadLayered(const CliArgs& a, const std::string& role) {
  std::string over;
  for (auto& o : a.overrides) over += o + "\n";
  Config c = Config::parseString(over).withFallback(Config::parseString("akka.cluster.roles = [" + role + "]"));
  std::string path = a.conf;
  if (path.empty()) {
    if (const char* e = getenv("AAR_CONF")) path = e;
  }
  if (path.empty()) {
    for (const char* cand : {"application.conf", "conf/application.conf"}) {
      if (FILE* f = fopen(cand, "r")) {
        fclose(f);
        path = cand;
        break;
      }
    }
  }
  if (!path.empty()) c = c.withFallback(Config::parseFile(path));
  return c;
}

}  // namespace aar
