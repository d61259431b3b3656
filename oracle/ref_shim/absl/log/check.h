// ORACLE BUILD SHIM (test infrastructure): CHECK* macros with absl's streaming syntax.
#ifndef ORACLE_SHIM_ABSL_CHECK_H_
#define ORACLE_SHIM_ABSL_CHECK_H_
#include <cstdlib>
#include <iostream>
#include <sstream>
namespace absl_shim {
struct Fatal {
  std::ostringstream s; bool on;
  explicit Fatal(bool failed) : on(failed) {}
  template <class T> Fatal& operator<<(const T& v) { if (on) s << v; return *this; }
  ~Fatal() { if (on) { std::cerr << "CHECK failed: " << s.str() << std::endl; std::abort(); } }
};
}  // namespace absl_shim
#define CHECK(c) ::absl_shim::Fatal(!(c)) << #c << " "
#define CHECK_EQ(a, b) CHECK((a) == (b))
#define CHECK_NE(a, b) CHECK((a) != (b))
#define CHECK_LE(a, b) CHECK((a) <= (b))
#define CHECK_LT(a, b) CHECK((a) < (b))
#define CHECK_GE(a, b) CHECK((a) >= (b))
#define CHECK_GT(a, b) CHECK((a) > (b))
#endif
