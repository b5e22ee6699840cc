// Host-side helpers shared by the translation units of libaar_cuda.so.
#pragma once
#include <cstdarg>
#include <cstdio>
#include <string>

#define AAR_API extern "C" __attribute__((visibility("default")))

namespace aar_host {

inline thread_local std::string g_lastError;

inline int fail(int code, const char* fmt, ...) __attribute__((format(printf, 2, 3)));
inline int fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_lastError = buf;
  return code;
}

}  // namespace aar_host
