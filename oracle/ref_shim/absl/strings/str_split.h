// ORACLE BUILD SHIM (test infrastructure): absl::StrSplit(text, char) -> a range of string_views that also converts to
// std::vector<std::string> (the two ways mjpc/utilities.cc consumes it).
#ifndef ORACLE_SHIM_ABSL_STR_SPLIT_H_
#define ORACLE_SHIM_ABSL_STR_SPLIT_H_
#include <string>
#include <string_view>
#include <vector>
#include "match.h"
namespace absl {
struct SplitResult {
  std::vector<std::string_view> parts;
  std::vector<std::string_view>::const_iterator begin() const { return parts.begin(); }
  std::vector<std::string_view>::const_iterator end() const { return parts.end(); }
  operator std::vector<std::string>() const { return std::vector<std::string>(parts.begin(), parts.end()); }
  operator std::vector<std::string_view>() const { return parts; }
};
inline SplitResult StrSplit(std::string_view text, char d) {
  SplitResult r; size_t a = 0;
  for (;;) { size_t b = text.find(d, a); if (b == std::string_view::npos) { r.parts.push_back(text.substr(a)); break; } r.parts.push_back(text.substr(a, b - a)); a = b + 1; }
  return r;
}
}  // namespace absl
#endif
