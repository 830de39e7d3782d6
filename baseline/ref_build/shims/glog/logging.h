// glog/logging.h shim — LOG / VLOG / CHECK* with glog's streaming syntax, stderr sink.
#ifndef MINIPS_SHIM_GLOG_H_
#define MINIPS_SHIM_GLOG_H_
#include <cstdlib>
#include <iostream>
#include <sstream>
#include <string>

namespace google {
enum LogSeverity { GLOG_INFO = 0, GLOG_WARNING = 1, GLOG_ERROR = 2, GLOG_FATAL = 3 };
inline int& MinLogLevel() {
  static int lvl = [] {
    const char* e = std::getenv("GLOG_minloglevel");
    return e ? std::atoi(e) : 1;  // quiet by default: INFO lines cost time in a benchmark
  }();
  return lvl;
}
inline int& VLevel() {
  static int v = [] {
    const char* e = std::getenv("GLOG_v");
    return e ? std::atoi(e) : 0;
  }();
  return v;
}
inline void InitGoogleLogging(const char*) {}
inline void ShutdownGoogleLogging() {}
inline void InstallFailureSignalHandler() {}
class LogMessage {
 public:
  LogMessage(const char* file, int line, int severity) : severity_(severity) {
    static const char tag[] = "IWEF";
    os_ << tag[severity] << " " << file << ":" << line << "] ";
  }
  ~LogMessage() {
    os_ << "\n";
    if (severity_ >= MinLogLevel() || severity_ == GLOG_FATAL) std::cerr << os_.str() << std::flush;
    if (severity_ == GLOG_FATAL) std::abort();
  }
  std::ostream& stream() { return os_; }

 private:
  std::ostringstream os_;
  int severity_;
};
struct LogMessageVoidify {
  void operator&(std::ostream&) {}
};
}  // namespace google

#define LOG(sev) ::google::LogMessage(__FILE__, __LINE__, ::google::GLOG_##sev).stream()
#define LOG_IF(sev, cond) !(cond) ? (void)0 : ::google::LogMessageVoidify() & LOG(sev)
#define VLOG(n) !(::google::VLevel() >= (n)) ? (void)0 : ::google::LogMessageVoidify() & LOG(INFO)
#define DLOG(sev) LOG(sev)
#define CHECK(cond) (cond) ? (void)0 : ::google::LogMessageVoidify() & LOG(FATAL) << "Check failed: " #cond " "
#define MINIPS_CHECK_OP(a, b, op)                                                                      \
  ((a)op(b)) ? (void)0                                                                                 \
             : ::google::LogMessageVoidify() & LOG(FATAL) << "Check failed: " #a " " #op " " #b " (" << (a) \
                                                          << " vs " << (b) << ") "
#define CHECK_EQ(a, b) MINIPS_CHECK_OP(a, b, ==)
#define CHECK_NE(a, b) MINIPS_CHECK_OP(a, b, !=)
#define CHECK_LT(a, b) MINIPS_CHECK_OP(a, b, <)
#define CHECK_LE(a, b) MINIPS_CHECK_OP(a, b, <=)
#define CHECK_GT(a, b) MINIPS_CHECK_OP(a, b, >)
#define CHECK_GE(a, b) MINIPS_CHECK_OP(a, b, >=)
#define CHECK_NOTNULL(p) (p)
#define DCHECK(c) CHECK(c)
#define DCHECK_EQ(a, b) CHECK_EQ(a, b)
#define DCHECK_LT(a, b) CHECK_LT(a, b)
#endif
