// glog/logging.h — the two glog macros the reference API surface uses (LOG(severity), CHECK(cond)), so that
// scenes build without the vendored glog.  LOG lines go to stderr as "[I file:line] message"; CHECK failure and
// LOG(FATAL) abort like glog.  RT_LOG=0 silences INFO.
#pragma once
#include <cstdlib>
#include <functional>
#include <iostream>
#include <sstream>
namespace rtapi {
class LogMessage {
 public:
  LogMessage(const char *file, int line, int severity) : severity_(severity) {
    static const char tag[] = {'I', 'W', 'E', 'F'};
    const char *base = file;
    for (const char *p = file; *p; ++p)
      if (*p == '/') base = p + 1;
    os_ << "[" << tag[severity & 3] << " " << base << ":" << line << "] ";
  }
  ~LogMessage() {
    const char *q = std::getenv("RT_LOG");
    if (severity_ > 0 || !q || q[0] != '0') std::cerr << os_.str() << std::endl;
    if (severity_ >= 3) std::abort();
  }
  std::ostream &stream() { return os_; }

 private:
  std::ostringstream os_;
  int severity_;
};
struct LogVoidify {
  void operator&(std::ostream &) {}
};
}  // namespace rtapi
#define RTAPI_SEV_INFO 0
#define RTAPI_SEV_WARNING 1
#define RTAPI_SEV_ERROR 2
#define RTAPI_SEV_FATAL 3
#define LOG(severity) ::rtapi::LogMessage(__FILE__, __LINE__, RTAPI_SEV_##severity).stream()
#define CHECK(cond) \
  (cond) ? (void)0 : ::rtapi::LogVoidify() & ::rtapi::LogMessage(__FILE__, __LINE__, 3).stream() << "Check failed: " #cond " "
