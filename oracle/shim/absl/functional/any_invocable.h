// TEST INFRASTRUCTURE ONLY -- minimal stand-in for absl/functional/any_invocable.h:
// AnyInvocable<R(Args...) const> and AnyInvocable<R(Args...)> over std::function
// (every callable the reference stores is copyable).
#ifndef ORACLE_SHIM_ABSL_FUNCTIONAL_ANY_INVOCABLE_H_
#define ORACLE_SHIM_ABSL_FUNCTIONAL_ANY_INVOCABLE_H_

#include <functional>
#include <utility>

namespace absl {
template <typename Sig>
class AnyInvocable;
template <typename R, typename... Args>
class AnyInvocable<R(Args...) const> {
 public:
  AnyInvocable() = default;
  template <typename F>
  AnyInvocable(F f) : f_(std::move(f)) {}  // NOLINT
  R operator()(Args... args) const { return f_(std::forward<Args>(args)...); }

 private:
  std::function<R(Args...)> f_;
};
template <typename R, typename... Args>
class AnyInvocable<R(Args...)> {
 public:
  AnyInvocable() = default;
  template <typename F>
  AnyInvocable(F f) : f_(std::move(f)) {}  // NOLINT
  R operator()(Args... args) { return f_(std::forward<Args>(args)...); }

 private:
  std::function<R(Args...)> f_;
};
}  // namespace absl

#endif  // ORACLE_SHIM_ABSL_FUNCTIONAL_ANY_INVOCABLE_H_
