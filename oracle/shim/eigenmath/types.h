// TEST INFRASTRUCTURE ONLY -- not part of the product.
//
// Stand-in for eigenmath/types.h (x_edr_eigenmath v1.0.0 is an un-vendored Bazel
// module; Eigen is not installed in this image).  The eigenmath names are assumed
// to be plain Eigen typedefs; the Eigen stand-in is
// oracle/shim/Eigen/Dense, which evaluates every operator per coefficient in the
// order Eigen's expression templates do.
#ifndef ORACLE_SHIM_EIGENMATH_TYPES_H_
#define ORACLE_SHIM_EIGENMATH_TYPES_H_

#include "Eigen/Dense"

namespace eigenmath {

using VectorXd = Eigen::VectorXd;
using MatrixXd = Eigen::MatrixXd;
using Vector2d = Eigen::Vector2d;
using Vector3d = Eigen::Vector3d;
using Quaterniond = Eigen::Quaterniond;
using Vector6d = Eigen::Vector6d;
using Matrix6Xd = Eigen::Matrix6Xd;

// VectorNd is Eigen's dynamic vector with a fixed maximum capacity; the capacity
// (un-vendored) only bounds BSplineNd's point size, which nothing here instantiates.
constexpr int kMaxEigenVectorCapacity = 16;
using VectorNd = Eigen::VectorXd;

namespace types_detail {
template <typename Scalar, int N>
struct VectorSel;
template <>
struct VectorSel<double, Eigen::Dynamic> {
  using type = Eigen::VectorXd;
};
template <>
struct VectorSel<double, 2> {
  using type = Eigen::Vector2d;
};
template <>
struct VectorSel<double, 3> {
  using type = Eigen::Vector3d;
};
}  // namespace types_detail
template <typename Scalar, int N>
using Vector = typename types_detail::VectorSel<Scalar, N>::type;

}  // namespace eigenmath

#endif  // ORACLE_SHIM_EIGENMATH_TYPES_H_
