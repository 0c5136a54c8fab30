/*
 * CHECK / LOG with the reference's observable behaviour on this path: a failed CHECK or
 * LOG(FATAL) throws dmlc::Error (reference include/dmlc/logging.h:31-37,379-405 with the default
 * DMLC_LOG_FATAL_THROW=1, base.h:20-22); LOG(INFO/WARNING/ERROR) print one line to stderr.
 */
#ifndef DMLC_LOGGING_H_
#define DMLC_LOGGING_H_

#include <cstdio>
#include <sstream>
#include <stdexcept>
#include <string>

namespace dmlc {

/* the exception every fatal condition raises */
struct Error : public std::runtime_error {
  explicit Error(const std::string &s) : std::runtime_error(s) {}
};

namespace log_detail {

class Line {
 public:
  Line(const char *file, int line, const char *tag) { os_ << "[" << tag << "] " << file << ":" << line << ": "; }
  ~Line() { std::fprintf(stderr, "%s\n", os_.str().c_str()); }
  std::ostringstream &stream() { return os_; }
 private:
  std::ostringstream os_;
};

class Fatal {
 public:
  Fatal(const char *file, int line) { os_ << file << ":" << line << ": "; }
  std::ostringstream &stream() { return os_; }
  ~Fatal() noexcept(false) { throw Error(os_.str()); }
 private:
  std::ostringstream os_;
};

/* swallows the stream expression in `cond ? (void)0 : Voidify() & stream` */
struct Voidify {
  void operator&(std::ostream &) {}
};

}  // namespace log_detail
}  // namespace dmlc

#define LOG_INFO ::dmlc::log_detail::Line(__FILE__, __LINE__, "INFO").stream()
#define LOG_WARNING ::dmlc::log_detail::Line(__FILE__, __LINE__, "WARNING").stream()
#define LOG_ERROR ::dmlc::log_detail::Line(__FILE__, __LINE__, "ERROR").stream()
#define LOG_FATAL ::dmlc::log_detail::Fatal(__FILE__, __LINE__).stream()
#define LOG(severity) LOG_##severity

#define CHECK(x) \
  (x) ? (void)0 : ::dmlc::log_detail::Voidify() & LOG_FATAL << "Check failed: " #x << ' '
#define DMLC_CHECK_OP_(a, b, op) \
  ((a)op(b)) ? (void)0 : ::dmlc::log_detail::Voidify() & LOG_FATAL << "Check failed: " #a " " #op " " #b << ' '
#define CHECK_EQ(a, b) DMLC_CHECK_OP_(a, b, ==)
#define CHECK_NE(a, b) DMLC_CHECK_OP_(a, b, !=)
#define CHECK_LT(a, b) DMLC_CHECK_OP_(a, b, <)
#define CHECK_LE(a, b) DMLC_CHECK_OP_(a, b, <=)
#define CHECK_GT(a, b) DMLC_CHECK_OP_(a, b, >)
#define CHECK_GE(a, b) DMLC_CHECK_OP_(a, b, >=)
#define CHECK_NOTNULL(p) \
  ((p) == NULL ? (LOG_FATAL << "Check notnull: " #p << ' ', (p)) : (p))

#endif  // DMLC_LOGGING_H_
