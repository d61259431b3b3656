// ORACLE BUILD SHIM (test infrastructure): the subset of absl::Span the reference's
// mjpc/spline/spline.{h,cc} uses, so those files compile unmodified into oracle/_ref/.
#ifndef ORACLE_SHIM_ABSL_SPAN_H_
#define ORACLE_SHIM_ABSL_SPAN_H_
#include <cstddef>
#include <initializer_list>
#include <type_traits>
#include <vector>
namespace absl {
template <class T>
class Span {
 public:
  using value_type = std::remove_cv_t<T>;
  Span() : p_(nullptr), n_(0) {}
  Span(T* p, size_t n) : p_(p), n_(n) {}
  template <class V, class = std::enable_if_t<std::is_same<std::remove_cv_t<typename V::value_type>, value_type>::value>>
  Span(V& v) : p_(v.data()), n_(v.size()) {}
  template <class V, class = std::enable_if_t<std::is_const<T>::value && std::is_same<typename V::value_type, value_type>::value>, class = void>
  Span(const V& v) : p_(v.data()), n_(v.size()) {}
  template <class U = T, class = std::enable_if_t<std::is_const<U>::value>>
  Span(std::initializer_list<value_type> l) : p_(l.begin()), n_(l.size()) {}
  T* data() const { return p_; }
  size_t size() const { return n_; }
  bool empty() const { return n_ == 0; }
  T* begin() const { return p_; }
  T* end() const { return p_ + n_; }
  T& operator[](size_t i) const { return p_[i]; }
  T& at(size_t i) const { return p_[i]; }
 private:
  T* p_; size_t n_;
};
template <class T> Span<T> MakeSpan(T* p, size_t n) { return Span<T>(p, n); }
template <class V> Span<typename V::value_type> MakeSpan(V& v) { return Span<typename V::value_type>(v.data(), v.size()); }
template <class T> Span<const T> MakeConstSpan(const T* p, size_t n) { return Span<const T>(p, n); }
template <class V> Span<const typename V::value_type> MakeConstSpan(const V& v) { return Span<const typename V::value_type>(v.data(), v.size()); }
}  // namespace absl
#endif
