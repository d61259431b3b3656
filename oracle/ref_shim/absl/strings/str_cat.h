// ORACLE BUILD SHIM (test infrastructure): absl::StrCat for strings / string views / numbers.
#ifndef ORACLE_SHIM_ABSL_STR_CAT_H_
#define ORACLE_SHIM_ABSL_STR_CAT_H_
#include <sstream>
#include <string>
#include <string_view>
namespace absl {
template <class... A> std::string StrCat(const A&... a) { std::ostringstream s; (void)std::initializer_list<int>{((s << a), 0)...}; return s.str(); }
}  // namespace absl
#endif
