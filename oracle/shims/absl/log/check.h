// TEST INFRASTRUCTURE ONLY.  Stand-in for absl/log/check.h (abseil 20230125.3): a failed
// CHECK throws shim::CheckFailure (the real one aborts the process) so that the oracle/_ref
// driver can report the failed precondition as a status code.
#ifndef ORACLE_SHIMS_ABSL_LOG_CHECK_H_
#define ORACLE_SHIMS_ABSL_LOG_CHECK_H_
#include <optional>
#include <sstream>
#include <stdexcept>
#include <string>
namespace shim {
struct CheckFailure : std::runtime_error {
  using std::runtime_error::runtime_error;
};
class CheckSink {
 public:
  CheckSink(const char* file, int line, const char* what) { s_ << file << ":" << line << " CHECK failed: " << what << " "; }
  ~CheckSink() noexcept(false) { throw CheckFailure(s_.str()); }
  template <typename T>
  CheckSink& operator<<(const T& v) {
    s_ << v;
    return *this;
  }
 private:
  std::ostringstream s_;
};
}  // namespace shim
namespace absl {
using std::optional;
using std::nullopt;
}  // namespace absl
#define SHIM_CHECK_IMPL(cond, text) \
  if (cond) {                       \
  } else                            \
    ::shim::CheckSink(__FILE__, __LINE__, text)
#define CHECK(c) SHIM_CHECK_IMPL((c), #c)
#define CHECK_EQ(a, b) SHIM_CHECK_IMPL((a) == (b), #a " == " #b)
#define CHECK_NE(a, b) SHIM_CHECK_IMPL((a) != (b), #a " != " #b)
#define CHECK_LT(a, b) SHIM_CHECK_IMPL((a) < (b), #a " < " #b)
#define CHECK_LE(a, b) SHIM_CHECK_IMPL((a) <= (b), #a " <= " #b)
#define CHECK_GT(a, b) SHIM_CHECK_IMPL((a) > (b), #a " > " #b)
#define CHECK_GE(a, b) SHIM_CHECK_IMPL((a) >= (b), #a " >= " #b)
#endif
