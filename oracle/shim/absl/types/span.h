// TEST INFRASTRUCTURE ONLY -- minimal stand-in for absl/types/span.h.
#ifndef ORACLE_SHIM_ABSL_TYPES_SPAN_H_
#define ORACLE_SHIM_ABSL_TYPES_SPAN_H_

#include <cstddef>
#include <type_traits>
#include <utility>
#include <vector>

namespace absl {

template <typename T>
class Span {
 public:
  using value_type = typename std::remove_cv<T>::type;
  Span() = default;
  Span(T* p, size_t n) : p_(p), n_(n) {}
  // any contiguous container with data() / size() (std::vector, Eigen vectors)
  template <typename C, typename = typename std::enable_if<
                            std::is_convertible<decltype(std::declval<C&>().data()), T*>::value &&
                            !std::is_same<typename std::decay<C>::type, Span>::value>::type>
  Span(C& c) : p_(c.data()), n_(static_cast<size_t>(c.size())) {}  // NOLINT
  template <typename C, typename = typename std::enable_if<
                            std::is_const<T>::value &&
                            std::is_convertible<decltype(std::declval<const C&>().data()), T*>::value &&
                            !std::is_same<typename std::decay<C>::type, Span>::value>::type>
  Span(const C& c) : p_(c.data()), n_(static_cast<size_t>(c.size())) {}  // NOLINT
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
template <typename C>
auto MakeConstSpan(const C& c)
    -> Span<const typename std::remove_pointer<decltype(c.data())>::type> {
  return Span<const typename std::remove_pointer<decltype(c.data())>::type>(
      c.data(), static_cast<size_t>(c.size()));
}
template <typename C>
auto MakeSpan(C& c) -> Span<typename std::remove_pointer<decltype(c.data())>::type> {
  return Span<typename std::remove_pointer<decltype(c.data())>::type>(
      c.data(), static_cast<size_t>(c.size()));
}

}  // namespace absl

#endif  // ORACLE_SHIM_ABSL_TYPES_SPAN_H_
