// Minimal stand-in for <glog/logging.h>, written for the oracle build only.
//
// TEST INFRASTRUCTURE -- not part of the product.  The reference (mfouda/epsilon)
// uses seven glog macros (LOG, VLOG, CHECK, CHECK_EQ/NE/GE, InitGoogleLogging);
// building the vendored glog needs cmake-generated headers, so the oracle recipe
// (oracle/Makefile) puts this directory on the include path instead.  Semantics
// kept: a failed CHECK / LOG(FATAL) prints the message and calls the installed failure function, else aborts (the reference
// longjmps out of glog's failure function, python/epopt/solvemodule.cc:245-263;
// stand-alone C++ aborts, which is what the harness wants).
#ifndef ORACLE_SHIM_GLOG_LOGGING_H
#define ORACLE_SHIM_GLOG_LOGGING_H

#include <cstdlib>
#include <iostream>
#include <sstream>
#include <string>

namespace google {

enum LogSeverity { INFO = 0, WARNING = 1, ERROR = 2, FATAL = 3 };

inline int& ShimVerbosity() {
  static int v = [] {
    const char* s = std::getenv("GLOG_v");
    return s ? std::atoi(s) : 0;
  }();
  return v;
}

// InstallFailureFunction (glog): called instead of abort() when a CHECK / LOG(FATAL) fires.  The reference's
// Python binding longjmps out of it (python/epopt/solvemodule.cc:245-248); so does adapter/solvemodule_py3.cc.
inline void (*&ShimFailureFunction())() {
  static void (*f)() = nullptr;
  return f;
}

class ShimLogMessage {
 public:
  ShimLogMessage(const char* file, int line, int severity)
      : severity_(severity) {
    stream_ << "[" << "IWEF"[severity] << " " << file << ":" << line << "] ";
  }
  ~ShimLogMessage() {
    stream_ << "\n";
    std::cerr << stream_.str();
    if (severity_ == FATAL) {
      std::cerr.flush();
      if (ShimFailureFunction() != nullptr) ShimFailureFunction()();
      std::abort();
    }
  }
  std::ostream& stream() { return stream_; }

 private:
  int severity_;
  std::ostringstream stream_;
};

// Swallows a stream expression so `cond ? (void)0 : Voidify() & stream` types.
struct ShimVoidify {
  void operator&(std::ostream&) {}
};

inline void InitGoogleLogging(const char*) {}
inline void InstallFailureFunction(void (*f)()) { ShimFailureFunction() = f; }

}  // namespace google

#define SHIM_LOG_INFO ::google::ShimLogMessage(__FILE__, __LINE__, ::google::INFO)
#define SHIM_LOG_WARNING ::google::ShimLogMessage(__FILE__, __LINE__, ::google::WARNING)
#define SHIM_LOG_ERROR ::google::ShimLogMessage(__FILE__, __LINE__, ::google::ERROR)
#define SHIM_LOG_FATAL ::google::ShimLogMessage(__FILE__, __LINE__, ::google::FATAL)

#define LOG(severity) SHIM_LOG_##severity.stream()
#define VLOG_IS_ON(n) (::google::ShimVerbosity() >= (n))
#define VLOG(n) \
  !VLOG_IS_ON(n) ? (void)0 : ::google::ShimVoidify() & SHIM_LOG_INFO.stream()
#define CHECK(cond)                 \
  (cond) ? (void)0                  \
         : ::google::ShimVoidify() & \
               SHIM_LOG_FATAL.stream() << "Check failed: " #cond " "
#define SHIM_CHECK_OP(a, b, op)                                          \
  ((a)op(b)) ? (void)0                                                   \
             : ::google::ShimVoidify() &                                 \
                   SHIM_LOG_FATAL.stream()                               \
                       << "Check failed: " #a " " #op " " #b " (" << (a) \
                       << " vs. " << (b) << ") "
#define CHECK_EQ(a, b) SHIM_CHECK_OP(a, b, ==)
#define CHECK_NE(a, b) SHIM_CHECK_OP(a, b, !=)
#define CHECK_GE(a, b) SHIM_CHECK_OP(a, b, >=)
#define CHECK_GT(a, b) SHIM_CHECK_OP(a, b, >)
#define CHECK_LE(a, b) SHIM_CHECK_OP(a, b, <=)
#define CHECK_LT(a, b) SHIM_CHECK_OP(a, b, <)

#endif  // ORACLE_SHIM_GLOG_LOGGING_H
