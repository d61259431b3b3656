// ORACLE BUILD SHIM (test infrastructure): absl::flat_hash_map over std::unordered_map, with the string_view key lookup
// (heterogeneous operator[]) mjpc/utilities.cc relies on.
#ifndef ORACLE_SHIM_ABSL_FLAT_HASH_MAP_H_
#define ORACLE_SHIM_ABSL_FLAT_HASH_MAP_H_
#include <string>
#include <string_view>
#include <unordered_map>
namespace absl {
template <class K, class V>
class flat_hash_map : public std::unordered_map<K, V> {
 public:
  using std::unordered_map<K, V>::unordered_map;
  using std::unordered_map<K, V>::operator[];
  template <class Q = K, class = std::enable_if_t<std::is_same<Q, std::string>::value>>
  V& operator[](std::string_view k) { return std::unordered_map<K, V>::operator[](std::string(k)); }
  template <class Q = K, class = std::enable_if_t<std::is_same<Q, std::string>::value>>
  auto find(std::string_view k) { return std::unordered_map<K, V>::find(std::string(k)); }
  using std::unordered_map<K, V>::find;
};
}  // namespace absl
#endif
