// Test-infrastructure shim: minimal glog surface (LOG(INFO), CHECK) for compiling the
// reference sources on the host.  Not product code.
#pragma once
#include <cstdlib>
#include <functional>
#include <iostream>
#include <sstream>
namespace rt_shim {
struct LogLine {
  std::ostringstream os;
  bool fatal;
  explicit LogLine(bool f) : fatal(f) {}
  ~LogLine() {
    if (fatal) { std::cerr << os.str() << std::endl; std::abort(); }
    if (std::getenv("RT_SHIM_VERBOSE")) std::cerr << os.str() << std::endl;
  }
  template <typename T> LogLine &operator<<(const T &v) { os << v; return *this; }
};
struct Voidify { void operator&(const LogLine &) {} };
}  // namespace rt_shim
#define INFO 0
#define WARNING 1
#define ERROR 2
#define FATAL 3
#define LOG(sev) ::rt_shim::LogLine((sev) == 3)
#define CHECK(cond) (cond) ? (void)0 : ::rt_shim::Voidify() & ::rt_shim::LogLine(true) << "CHECK failed: " #cond " "
