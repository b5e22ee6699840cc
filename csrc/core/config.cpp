// See config.h.
#include "config.h"

#include <unistd.h>

#include <cctype>
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <stdexcept>

namespace aar {
namespace {

struct Parser {
  const std::string& s;
  size_t i = 0;
  int line = 1;
  std::map<std::string, Config::Value>& out;

  Parser(const std::string& text, std::map<std::string, Config::Value>& o) : s(text), out(o) {}

  [[noreturn]] void fail(const std::string& why) const { throw std::runtime_error("config parse error at line " + std::to_string(line) + ": " + why); }
  bool eof() const { return i >= s.size(); }
  char peek() const { return eof() ? '\0' : s[i]; }
  char get() {
    char c = s[i++];
    if (c == '\n') ++line;
    return c;
  }
  bool atComment() const { return peek() == '#' || (peek() == '/' && i + 1 < s.size() && s[i + 1] == '/'); }
  void skipLine() {
    while (!eof() && peek() != '\n') get();
  }
  // whitespace, newlines, commas and comments between members
  void skipFiller(bool commas) {
    for (;;) {
      if (eof()) return;
      char c = peek();
      if (isspace((unsigned char)c) || (commas && c == ',')) {
        get();
      } else if (atComment()) {
        skipLine();
      } else {
        return;
      }
    }
  }
  void skipInlineSpace() {
    while (!eof() && (peek() == ' ' || peek() == '\t' || peek() == '\r')) get();
  }
  std::string quoted() {
    get();  // opening quote
    std::string r;
    while (!eof() && peek() != '"') {
      char c = get();
      if (c == '\\' && !eof()) {
        char e = get();
        switch (e) {
          case 'n': r += '\n'; break;
          case 't': r += '\t'; break;
          default: r += e;
        }
      } else {
        r += c;
      }
    }
    if (eof()) fail("unterminated string");
    get();
    return r;
  }
  static std::string trim(const std::string& v) {
    size_t a = 0, b = v.size();
    while (a < b && isspace((unsigned char)v[a])) ++a;
    while (b > a && isspace((unsigned char)v[b - 1])) --b;
    return v.substr(a, b - a);
  }
  std::string key() {
    std::string k;
    for (;;) {
      if (peek() == '"') {
        k += quoted();
      } else {
        while (!eof() && !isspace((unsigned char)peek()) && peek() != '=' && peek() != ':' && peek() != '{' && peek() != '"') k += get();
      }
      if (peek() != '"') break;
    }
    if (k.empty()) fail("expected a key");
    return k;
  }
  // scalar up to end of line / ',' / '}' / ']' / comment; quoted pieces are concatenated verbatim
  std::string scalar(bool inList) {
    std::string v;
    bool lastQuoted = false;
    for (;;) {
      if (eof()) break;
      char c = peek();
      if (c == '"') {
        v += quoted();
        lastQuoted = true;
        continue;
      }
      if (c == '$' && i + 1 < s.size() && s[i + 1] == '{') {  // keep ${path} intact (it contains '}')
        while (!eof() && peek() != '}' && peek() != '\n') v += get();
        if (peek() == '}') v += get();
        continue;
      }
      if (c == '\n' || c == ',' || c == '}' || atComment()) break;
      if (inList && c == ']') break;
      v += get();
      lastQuoted = false;
    }
    (void)lastQuoted;
    return trim(v);
  }
  Config::Value list() {
    get();  // [
    Config::Value val;
    val.isList = true;
    for (;;) {
      skipFiller(true);
      if (eof()) fail("unterminated list");
      if (peek() == ']') {
        get();
        break;
      }
      const size_t before = i;
      val.list.push_back(scalar(true));
      // scalar() stops in front of '}' (the end of an enclosing object) without consuming it: inside a list that is a
      // stray brace, and looping on it would append empty elements until memory runs out
      if (i == before) fail(std::string("unexpected '") + peek() + "' in list");
    }
    return val;
  }
  void object(const std::string& prefix, bool braced) {
    for (;;) {
      skipFiller(true);
      if (eof()) {
        if (braced) fail("missing '}'");
        return;
      }
      if (peek() == '}') {
        if (!braced) fail("unexpected '}'");
        get();
        return;
      }
      std::string k = prefix + key();
      skipInlineSpace();
      char c = peek();
      if (c == '{') {
        get();
        object(k + ".", true);
        continue;
      }
      if (c != '=' && c != ':') fail("expected '=', ':' or '{' after key '" + k + "'");
      get();
      skipInlineSpace();
      if (peek() == '{') {
        get();
        object(k + ".", true);
      } else if (peek() == '[') {
        out[k] = list();
      } else {
        Config::Value v;
        v.scalar = scalar(false);
        out[k] = v;
      }
    }
  }
};

std::string lower(std::string v) {
  for (auto& c : v) c = (char)tolower((unsigned char)c);
  return v;
}

}  // namespace

Config Config::parseString(const std::string& text) {
  Config c;
  Parser p(text, c.values_);
  p.object("", false);
  c.resolveSubstitutions();
  return c;
}

Config Config::parseFile(const std::string& path) {
  std::ifstream f(path);
  if (!f) throw std::runtime_error("cannot open config file: " + path);
  std::stringstream ss;
  ss << f.rdbuf();
  return parseString(ss.str());
}

void Config::resolveSubstitutions() {
  auto resolveOne = [&](const std::string& in, int depth, auto&& self) -> std::string {
    if (depth > 16) throw std::runtime_error("config: substitution cycle in '" + in + "'");
    std::string out;
    size_t pos = 0;
    for (;;) {
      size_t a = in.find("${", pos);
      if (a == std::string::npos) {
        out += in.substr(pos);
        break;
      }
      size_t b = in.find('}', a);
      if (b == std::string::npos) throw std::runtime_error("config: unterminated substitution in '" + in + "'");
      out += in.substr(pos, a - pos);
      std::string path = in.substr(a + 2, b - a - 2);
      bool optional = !path.empty() && path[0] == '?';
      if (optional) path = path.substr(1);
      auto it = values_.find(path);
      if (it != values_.end() && !it->second.isList) {
        out += self(it->second.scalar, depth + 1, self);
      } else if (path == "user.dir") {
        char buf[4096];
        out += getcwd(buf, sizeof buf) ? buf : ".";
      } else if (const char* e = getenv(path.c_str())) {
        out += e;
      } else if (!optional) {
        throw std::runtime_error("config: could not resolve substitution ${" + path + "}");
      }
      pos = b + 1;
    }
    return out;
  };
  for (auto& kv : values_) {
    if (kv.second.isList) {
      for (auto& e : kv.second.list) e = resolveOne(e, 0, resolveOne);
    } else {
      kv.second.scalar = resolveOne(kv.second.scalar, 0, resolveOne);
    }
  }
}

Config Config::withFallback(const Config& fallback) const {
  Config c = *this;
  for (auto& kv : fallback.values_)
    if (!c.values_.count(kv.first)) c.values_[kv.first] = kv.second;
  return c;
}

std::string Config::getString(const std::string& path) const {
  auto it = values_.find(path);
  if (it == values_.end()) throw std::out_of_range("config: no setting at '" + path + "'");
  if (it->second.isList) throw std::runtime_error("config: '" + path + "' is a list");
  return it->second.scalar;
}

long Config::getInt(const std::string& path) const {
  std::string v = getString(path);
  try {
    size_t n = 0;
    long r = std::stol(v, &n);
    if (n != v.size()) {
      double d = std::stod(v, &n);
      if (n != v.size()) throw std::invalid_argument(v);
      return (long)d;
    }
    return r;
  } catch (const std::exception&) {
    throw std::runtime_error("config: '" + path + "' = '" + v + "' is not an integer");
  }
}

double Config::getDouble(const std::string& path) const {
  std::string v = getString(path);
  try {
    size_t n = 0;
    double d = std::stod(v, &n);
    if (n != v.size()) throw std::invalid_argument(v);
    return d;
  } catch (const std::exception&) {
    throw std::runtime_error("config: '" + path + "' = '" + v + "' is not a number");
  }
}

bool Config::getBool(const std::string& path) const {
  std::string v = lower(getString(path));
  if (v == "true" || v == "on" || v == "yes") return true;
  if (v == "false" || v == "off" || v == "no") return false;
  throw std::runtime_error("config: '" + path + "' = '" + v + "' is not a boolean");
}

std::vector<std::string> Config::getList(const std::string& path) const {
  auto it = values_.find(path);
  if (it == values_.end()) throw std::out_of_range("config: no setting at '" + path + "'");
  if (!it->second.isList) return {it->second.scalar};
  return it->second.list;
}

double Config::getDurationMs(const std::string& path) const {
  std::string v = getString(path);
  size_t n = 0;
  double d;
  try {
    d = std::stod(v, &n);
  } catch (const std::exception&) {
    throw std::runtime_error("config: '" + path + "' = '" + v + "' is not a duration");
  }
  std::string unit = lower(Parser::trim(v.substr(n)));
  if (unit.empty() || unit == "ms" || unit == "millis" || unit == "milliseconds") return d;
  if (unit == "s" || unit == "second" || unit == "seconds") return d * 1e3;
  if (unit == "us" || unit == "micros" || unit == "microseconds") return d * 1e-3;
  if (unit == "m" || unit == "minute" || unit == "minutes") return d * 6e4;
  if (unit == "h" || unit == "hour" || unit == "hours") return d * 3.6e6;
  throw std::runtime_error("config: unknown duration unit '" + unit + "' at '" + path + "'");
}

void Config::set(const std::string& path, const std::string& scalar) {
  Value v;
  v.scalar = scalar;
  values_[path] = v;
}

void Config::setList(const std::string& path, const std::vector<std::string>& list) {
  Value v;
  v.isList = true;
  v.list = list;
  values_[path] = v;
}

std::vector<std::string> Config::keys() const {
  std::vector<std::string> k;
  for (auto& kv : values_) k.push_back(kv.first);
  return k;
}

std::string Config::render() const {
  auto quote = [](const std::string& v) {
    std::string r = "\"";
    for (char c : v) {
      if (c == '"' || c == '\\') r += '\\';
      if (c == '\n') {
        r += "\\n";
        continue;
      }
      r += c;
    }
    return r + "\"";
  };
  std::string out;
  for (auto& kv : values_) {
    out += "\"" + kv.first + "\" = ";
    if (kv.second.isList) {
      out += "[";
      for (size_t i = 0; i < kv.second.list.size(); ++i) out += (i ? ", " : "") + quote(kv.second.list[i]);
      out += "]";
    } else {
      out += quote(kv.second.scalar);
    }
    out += "\n";
  }
  return out;
}

AllreduceSettings AllreduceSettings::fromConfig(const Config& c) {
  AllreduceSettings s;
  MasterConfig& m = s.master;
  // every integer is range-checked BEFORE it is narrowed to int (a value like 99999999999 must be an error, not garbage)
  auto intIn = [&](const char* key, long def, long lo, long hi) {
    const long v = c.getInt(key, def);
    if (v < lo || v > hi)
      throw std::runtime_error(std::string("config: ") + key + " = " + std::to_string(v) + " is outside " + std::to_string(lo) + ".." + std::to_string(hi));
    return (int)v;
  };
  constexpr long kIntMax = 2147483647L;
  m.totalWorkers = intIn("allreduce.total-workers", m.totalWorkers, 1, 1 << 16);
  m.thAllreduce = (float)c.getDouble("allreduce.th-allreduce", m.thAllreduce);
  m.thReduce = (float)c.getDouble("allreduce.th-reduce", m.thReduce);
  m.thComplete = (float)c.getDouble("allreduce.th-complete", m.thComplete);
  m.maxLag = intIn("allreduce.max-lag", m.maxLag, 0, 1 << 20);
  m.maxRound = intIn("allreduce.max-round", m.maxRound, -1, kIntMax - 1);
  m.maxChunkSize = intIn("allreduce.max-chunk-size", m.maxChunkSize, 1, kIntMax);
  // AllreduceMaster.scala:113 -- dataSize defaults to totalWorkers * 5
  m.dataSize = intIn("allreduce.data-size", (long)m.totalWorkers * 5, 0, kIntMax);
  m.pipelineDepth = intIn("allreduce.pipeline-depth", m.pipelineDepth, 0, 1 << 20);
  m.startRound = intIn("allreduce.start-round", m.startRound, 0, kIntMax - 1);
  s.dtype = c.getString("allreduce.dtype", s.dtype);
  s.algo = c.getString("allreduce.algo", s.algo);
  s.normalize = c.getBool("allreduce.normalize", s.normalize);
  s.quorumPolicy = c.getString("allreduce.quorum-policy", s.quorumPolicy);
  s.spinTimeoutUs = c.has("allreduce.spin-timeout") ? c.getDurationMs("allreduce.spin-timeout") * 1e3 : s.spinTimeoutUs;
  s.hostname = c.getString("allreduce.transport.hostname", c.getString("akka.remote.netty.tcp.hostname", s.hostname));
  s.port = intIn("allreduce.transport.port", s.port, 0, 65535);
  s.heartbeatMs = c.getDurationMs("allreduce.transport.heartbeat-interval", s.heartbeatMs);
  s.autoDownAfterMs = c.getDurationMs("allreduce.transport.auto-down-unreachable-after",
                                      c.getDurationMs("akka.cluster.auto-down-unreachable-after", s.autoDownAfterMs));
  s.loglevel = c.getString("allreduce.loglevel", c.getString("akka.loglevel", s.loglevel));
  if (!(s.heartbeatMs > 0) || !(s.autoDownAfterMs > 0)) throw std::runtime_error("config: transport intervals must be positive");
  if (s.quorumPolicy != "all-arrived" && s.quorumPolicy != "exact")
    throw std::runtime_error("config: allreduce.quorum-policy must be all-arrived or exact");
  bool algoOk = false;
  for (const char* a : {"auto", "bulk", "nvls_bulk", "nvls_onehop", "oneshot", "twoshot", "pipelined", "nvls"}) algoOk = algoOk || s.algo == a;
  if (!algoOk)
    throw std::runtime_error("config: allreduce.algo must be one of auto, bulk, nvls_bulk, nvls_onehop, oneshot, twoshot, pipelined, nvls");
  for (float th : {m.thAllreduce, m.thReduce, m.thComplete})
    if (!(th > 0.f && th <= 1.f)) throw std::runtime_error("config: thresholds must be in (0, 1]");
  if (s.dtype != "fp32" && s.dtype != "bf16") throw std::runtime_error("config: allreduce.dtype must be fp32 or bf16");
  return s;
}

}  // namespace aar
