// ORACLE BUILD SHIM (test infrastructure): absl::StartsWith / EndsWith / StrContains over std::string_view.
#ifndef ORACLE_SHIM_ABSL_MATCH_H_
#define ORACLE_SHIM_ABSL_MATCH_H_
#include <string_view>
namespace absl {
inline bool StartsWith(std::string_view s, std::string_view p) { return s.size() >= p.size() && s.compare(0, p.size(), p) == 0; }
inline bool EndsWith(std::string_view s, std::string_view p) { return s.size() >= p.size() && s.compare(s.size() - p.size(), p.size(), p) == 0; }
inline bool StrContains(std::string_view s, std::string_view p) { return s.find(p) != std::string_view::npos; }
inline std::string_view StripPrefix(std::string_view s, std::string_view p) { return StartsWith(s, p) ? s.substr(p.size()) : s; }
}  // namespace absl
#endif
