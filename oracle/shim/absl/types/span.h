// TEST INFRASTRUCTURE ONLY -- minimal stand-in for absl/types/span.h.
#ifndef ORACLE_SHIM_ABSL_TYPES_SPAN_H_
#define ORACLE_SHIM_ABSL_TYPES_SPAN_H_

#include <cstddef>
#include <type_traits>
#include <vector>

namespace absl {

template <typename T>
class Span {
 public:
  using value_type = typename std::remove_cv<T>::type;
  Span() = default;
  Span(T* p, size_t n) : p_(p), n_(n) {}
  Span(const std::vector<value_type>& v) : p_(v.data()), n_(v.size()) {}  // NOLINT
  T* data() const { return p_; }
  size_t size() const { return n_; }
  bool empty() const { return n_ == 0; }
  T& operator[](size_t i) const { return p_[i]; }
  T* begin() const { return p_; }
  T* end() const { return p_ + n_; }
  T& front() const { return p_[0]; }
  T& back() const { return p_[n_ - 1]; }
  Span subspan(size_t pos, size_t len) const { return Span(p_ + pos, len); }

 private:
  T* p_ = nullptr;
  size_t n_ = 0;
};

template <typename T>
Span<const T> MakeConstSpan(const T* p, size_t n) {
  return Span<const T>(p, n);
}

}  // namespace absl

#endif  // ORACLE_SHIM_ABSL_TYPES_SPAN_H_
