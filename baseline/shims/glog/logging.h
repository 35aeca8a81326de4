// Shim for the reference arm (baseline/README.md): the slice of glog the reference's
// heter_ps sources use — VLOG(n) / LOG(sev) / CHECK(cond) as ostream-style sinks.
// VLOG is silent unless REF_VLOG=<level> is set; CHECK aborts with the message.
#pragma once
#include <cstdlib>
#include <iostream>
#include <sstream>

namespace ref_shim {
inline int vlog_level() {
  static int lv = [] { const char* e = getenv("REF_VLOG"); return e ? atoi(e) : -1; }();
  return lv;
}
class LogLine {
 public:
  LogLine(bool on, bool fatal, const char* file, int line) : on_(on), fatal_(fatal) {
    if (on_ || fatal_) ss_ << "[ref " << file << ":" << line << "] ";
  }
  ~LogLine() {
    if (on_ || fatal_) std::cerr << ss_.str() << std::endl;
    if (fatal_) abort();
  }
  template <typename T>
  LogLine& operator<<(const T& v) {
    if (on_ || fatal_) ss_ << v;
    return *this;
  }
  LogLine& operator<<(std::ostream& (*f)(std::ostream&)) {
    if (on_ || fatal_) f(ss_);
    return *this;
  }
 private:
  bool on_, fatal_;
  std::ostringstream ss_;
};
struct Voidify { void operator&(LogLine&) {} };
}  // namespace ref_shim

#define VLOG(n) ::ref_shim::LogLine(::ref_shim::vlog_level() >= (n), false, __FILE__, __LINE__)
#define REF_LOG_INFO ::ref_shim::LogLine(::ref_shim::vlog_level() >= 0, false, __FILE__, __LINE__)
#define REF_LOG_WARNING ::ref_shim::LogLine(true, false, __FILE__, __LINE__)
#define REF_LOG_ERROR ::ref_shim::LogLine(true, false, __FILE__, __LINE__)
#define REF_LOG_FATAL ::ref_shim::LogLine(true, true, __FILE__, __LINE__)
#define LOG(sev) REF_LOG_##sev
#define CHECK(cond) \
  (cond) ? (void)0 : ::ref_shim::Voidify() & ::ref_shim::LogLine(true, true, __FILE__, __LINE__) << "CHECK failed: " #cond " "
