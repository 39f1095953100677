// TEST INFRASTRUCTURE ONLY -- stand-in for absl/functional/function_ref.h: a
// non-owning reference to a callable, valid while the callable lives.
#ifndef ORACLE_SHIM_ABSL_FUNCTIONAL_FUNCTION_REF_H_
#define ORACLE_SHIM_ABSL_FUNCTIONAL_FUNCTION_REF_H_

#include <type_traits>
#include <utility>

namespace absl {

template <typename Sig>
class FunctionRef;

template <typename R, typename... Args>
class FunctionRef<R(Args...)> {
 public:
  template <typename F, typename = std::enable_if_t<!std::is_same<std::decay_t<F>, FunctionRef>::value>>
  FunctionRef(const F& f)  // NOLINT
      : obj_(&f), call_([](const void* o, Args... args) -> R {
          return (*static_cast<const F*>(o))(std::forward<Args>(args)...);
        }) {}
  R operator()(Args... args) const { return call_(obj_, std::forward<Args>(args)...); }

 private:
  const void* obj_;
  R (*call_)(const void*, Args...);
};

}  // namespace absl

#endif  // ORACLE_SHIM_ABSL_FUNCTIONAL_FUNCTION_REF_H_
