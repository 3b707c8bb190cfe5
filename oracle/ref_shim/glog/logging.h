// TEST INFRASTRUCTURE (oracle/_ref build only): glog is not in this image.  CHECK* abort with a
// message like glog's; LOG(INFO/WARNING) are swallowed, LOG(FATAL) aborts.
#ifndef ORACLE_REF_SHIM_GLOG_LOGGING_H_
#define ORACLE_REF_SHIM_GLOG_LOGGING_H_

#include <cstdlib>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <utility>
#include <vector>

namespace oracle_ref_shim {
class LogSink {
 public:
  LogSink(const char* file, int line, bool fatal, bool print) : fatal_(fatal), print_(print || fatal) {
    if (print_) stream_ << file << ":" << line << ": ";
  }
  ~LogSink() {
    if (print_) std::cerr << stream_.str() << std::endl;
    if (fatal_) std::abort();
  }
  std::ostream& stream() { return stream_; }

 private:
  std::ostringstream stream_;
  bool fatal_, print_;
};
struct Voidify {
  void operator&(std::ostream&) {}
};
template <typename T>
T CheckNotNull(const char* file, int line, const char* what, T&& t) {
  if (t == nullptr) LogSink(file, line, true, true).stream() << what;
  return std::forward<T>(t);
}
}  // namespace oracle_ref_shim

#define ORACLE_SHIM_LOG_INFO ::oracle_ref_shim::LogSink(__FILE__, __LINE__, false, false).stream()
#define ORACLE_SHIM_LOG_WARNING ::oracle_ref_shim::LogSink(__FILE__, __LINE__, false, false).stream()
#define ORACLE_SHIM_LOG_ERROR ::oracle_ref_shim::LogSink(__FILE__, __LINE__, false, true).stream()
#define ORACLE_SHIM_LOG_FATAL ::oracle_ref_shim::LogSink(__FILE__, __LINE__, true, true).stream()
#define LOG(severity) ORACLE_SHIM_LOG_##severity
#define VLOG(level) ORACLE_SHIM_LOG_INFO
#define LOG_IF(severity, cond) \
  !(cond) ? (void)0 : ::oracle_ref_shim::Voidify() & LOG(severity)
#define LOG_EVERY_N(severity, n) LOG(severity)

#define CHECK(cond) \
  (cond) ? (void)0  \
         : ::oracle_ref_shim::Voidify() & ORACLE_SHIM_LOG_FATAL << "Check failed: " #cond " "
#define ORACLE_SHIM_CHECK_OP(a, op, b) \
  ((a)op(b)) ? (void)0                 \
             : ::oracle_ref_shim::Voidify() & ORACLE_SHIM_LOG_FATAL << "Check failed: " #a " " #op " " #b " "
#define CHECK_EQ(a, b) ORACLE_SHIM_CHECK_OP(a, ==, b)
#define CHECK_NE(a, b) ORACLE_SHIM_CHECK_OP(a, !=, b)
#define CHECK_LT(a, b) ORACLE_SHIM_CHECK_OP(a, <, b)
#define CHECK_LE(a, b) ORACLE_SHIM_CHECK_OP(a, <=, b)
#define CHECK_GT(a, b) ORACLE_SHIM_CHECK_OP(a, >, b)
#define CHECK_GE(a, b) ORACLE_SHIM_CHECK_OP(a, >=, b)
#define CHECK_NOTNULL(p) ::oracle_ref_shim::CheckNotNull(__FILE__, __LINE__, "'" #p "' Must be non NULL", (p))
// Release-mode glog compiles DCHECKs away (the reference's benchmarks are release builds).
#define ORACLE_SHIM_DCHECK(cond) \
  true ? (void)0 : ::oracle_ref_shim::Voidify() & ORACLE_SHIM_LOG_INFO << (cond)
#define DCHECK(cond) ORACLE_SHIM_DCHECK(cond)
#define DCHECK_EQ(a, b) ORACLE_SHIM_DCHECK((a) == (b))
#define DCHECK_NE(a, b) ORACLE_SHIM_DCHECK((a) != (b))
#define DCHECK_LT(a, b) ORACLE_SHIM_DCHECK((a) < (b))
#define DCHECK_LE(a, b) ORACLE_SHIM_DCHECK((a) <= (b))
#define DCHECK_GT(a, b) ORACLE_SHIM_DCHECK((a) > (b))
#define DCHECK_GE(a, b) ORACLE_SHIM_DCHECK((a) >= (b))

#endif  // ORACLE_REF_SHIM_GLOG_LOGGING_H_
