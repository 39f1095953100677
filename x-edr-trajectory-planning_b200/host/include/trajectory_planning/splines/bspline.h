// Host-side mirror of BSplineBase / BSplineT (trajectory_planning/splines/
// bspline_base.h:31-111, bspline.h:56-120): knot/control-point management with
// the reference's checks and status codes on the host, curve evaluation on the
// GPU (tpb_spline_eval_host).  EvalCurveBatch / EvalCurveAndDerivativesBatch are
// the calls the GPU is for; the scalar Eval* methods are a batch of one.
// Knot insertion / truncation / extension (replanning helpers) run on
// the host copy for a single spline; TimeableJointSplinePath::SwitchToWaypointPath uses
// their batched device form (tpb_switch_to_waypoint_path).
#ifndef TRAJECTORY_PLANNING_SPLINES_BSPLINE_H_
#define TRAJECTORY_PLANNING_SPLINES_BSPLINE_H_

#include <algorithm>
#include <cmath>
#include <limits>
#include <string>
#include <vector>

#include "tp/base.h"
#include "tpb.h"

namespace trajectory_planning {

class BSplineXd {
 public:
  using Point = tp::VectorXd;
  static constexpr size_t kMaxDegree = 5;  // device evaluator limit (reference: 128)

  // BSplineBase::Init + BSplineT::Init (bspline_base.cc:33-79, bspline.h:143-175)
  tp::Status Init(size_t degree, size_t max_num_knots, size_t points_dim) {
    init_ = false;
    if (max_num_knots < 3) return tp::OutOfRangeError("max_num_knots must be >= 3");
    if (degree > kMaxDegree || degree < 1) return tp::OutOfRangeError("degree out of range for the device evaluator");
    if (static_cast<long>(max_num_knots) - static_cast<long>(degree) - 1 < 1)
      return tp::InvalidArgumentError("Degree and knot capacity inconsistent");
    if (points_dim < 1) return tp::InvalidArgumentError("Point size not valid");
    degree_ = degree;
    max_num_knots_ = max_num_knots;
    points_dim_ = points_dim;
    num_knots_ = num_points_ = 0;
    knots_.assign(max_num_knots, 0.0);
    init_ = true;
    return tp::OkStatus();
  }
  tp::Status Init(size_t degree, tp::Span<const Point> points, tp::Span<const double> knots) {
    if (points.empty()) return tp::InvalidArgumentError("Control point vector empty.");
    if (auto s = Init(degree, knots.size(), points.front().size()); !s.ok()) return s;
    if (auto s = SetKnotVector(knots); !s.ok()) return s;
    return SetControlPoints(points);
  }
  double GetDegree() const { return static_cast<double>(degree_); }
  int NumPoints() const { return static_cast<int>(num_points_); }
  int MaxNumPoints() const { return static_cast<int>(NumPoints(max_num_knots_, degree_)); }
  int GetPointsDim() const { return static_cast<int>(points_dim_); }
  static constexpr long NumPoints(size_t knots, size_t degree) {
    return static_cast<long>(knots) - static_cast<long>(degree) - 1;
  }
  static constexpr size_t NumKnots(size_t points, size_t degree) { return points + degree + 1; }
  static constexpr size_t MinNumPoints(size_t degree) { return degree + 1; }
  static constexpr size_t MinNumKnots(size_t degree) { return 2 * degree + 2; }

  // SetKnotVector (bspline_base.cc:106-160)
  tp::Status SetKnotVector(tp::Span<const double> knots,
                           double umin = std::numeric_limits<double>::quiet_NaN(),
                           double umax = std::numeric_limits<double>::quiet_NaN()) {
    if (knots.size() > max_num_knots_) return tp::OutOfRangeError("knots.size() > max_num_knots");
    if (knots.size() < (degree_ + 1) * 2) return tp::OutOfRangeError("Knot vector too short");
    for (size_t i = 1; i < knots.size(); i++) {
      if (knots[i - 1] > knots[i]) return tp::InvalidArgumentError("knot vector not increasing");
    }
    for (size_t i = 1; i < degree_; ++i) {
      if (knots[i] != knots.front()) return tp::InvalidArgumentError("First degree_+1 knots must be equal.");
      if (knots[knots.size() - degree_ - 1 + i] != knots.back())
        return tp::InvalidArgumentError("Last degree_+1 knots must be equal.");
    }
    num_knots_ = knots.size();
    std::copy(knots.begin(), knots.end(), knots_.begin());
    num_points_ = NumPoints(num_knots_, degree_);
    umin_ = (!std::isfinite(umin) || umin < knots[0]) ? knots[0] : umin;
    umax_ = (!std::isfinite(umax) || umax > knots.back()) ? knots.back() : umax;
    return tp::OkStatus();
  }
  // SetUniformKnotVector (bspline_base.cc:81-104)
  tp::Status SetUniformKnotVector(size_t num_knots) {
    umin_ = 0;
    umax_ = 1.0;
    if (num_knots > max_num_knots_) return tp::OutOfRangeError("num_knots > max_num_knots");
    if (num_knots < (degree_ + 1) * 2) return tp::OutOfRangeError("Knot vector too short");
    for (size_t i = 0; i <= degree_; ++i) knots_[i] = 0.0;
    for (size_t i = num_knots - 1 - degree_; i < num_knots; ++i) knots_[i] = 1.0;
    const double spacing = 1.0 / (num_knots - 2 * (degree_ + 1) + 1);
    for (size_t idx = degree_ + 1; idx < num_knots - degree_ - 1; idx++) knots_[idx] = knots_[idx - 1] + spacing;
    num_knots_ = num_knots;
    num_points_ = NumPoints(num_knots_, degree_);
    return tp::OkStatus();
  }
  tp::Span<const double> GetKnotVector() const { return tp::Span<const double>(knots_.data(), num_knots_); }
  // MakeUniformKnotVector (bspline_base.cc:356-399)
  static tp::Status MakeUniformKnotVector(int num_points, size_t degree, double lo, double hi,
                                          tp::Span<double> knots) {
    if (hi <= lo) return tp::InvalidArgumentError("high_knot_value must be > low_knot_value");
    if (num_points < static_cast<int>(MinNumPoints(degree))) return tp::OutOfRangeError("Too few points");
    const size_t nknots = NumKnots(num_points, degree);
    if (nknots != knots.size()) return tp::OutOfRangeError("knots has wrong size");
    for (size_t i = 0; i <= degree; ++i) knots[i] = lo;
    const double spacing = (1.0 / (nknots - 2.0 * (degree + 1.0) + 1.0)) * (hi - lo);
    for (size_t idx = degree + 1; idx < nknots - degree - 1; idx++) knots[idx] = knots[idx - 1] + spacing;
    for (size_t i = nknots - degree - 1; i < nknots; ++i) knots[i] = hi;
    return tp::OkStatus();
  }
  static tp::Status MakeUniformKnotVector(int num_points, std::vector<double>* knots, size_t degree) {
    if (num_points < static_cast<int>(MinNumPoints(degree))) return tp::OutOfRangeError("Too few points");
    knots->resize(NumKnots(num_points, degree));
    return MakeUniformKnotVector(num_points, degree, 0.0, 1.0, tp::Span<double>(knots->data(), knots->size()));
  }

  // SetControlPoints (bspline.h:190-216)
  tp::Status SetControlPoints(tp::Span<const Point> points) {
    if (!init_) return tp::FailedPreconditionError("Call Init first.");
    if (static_cast<long>(points.size()) != NumPoints(num_knots_, degree_))
      return tp::InvalidArgumentError("points.size() inconsistent, got: " + std::to_string(points.size()));
    for (const auto& p : points) {
      if (p.size() != points_dim_) return tp::InvalidArgumentError("Invalid point size");
    }
    num_points_ = points.size();
    ctrl_.resize(num_points_ * points_dim_);
    for (size_t i = 0; i < num_points_; ++i) std::copy(points[i].begin(), points[i].end(), ctrl_.begin() + i * points_dim_);
    return tp::OkStatus();
  }
  std::vector<Point> GetControlPoints() const {
    std::vector<Point> out(num_points_);
    for (size_t i = 0; i < num_points_; ++i) out[i] = Point(&ctrl_[i * points_dim_], points_dim_);
    return out;
  }

  // ---- replanning helpers on the host copy of the spline (a handful of control
  // points; the batched device form is tpb_switch_to_waypoint_path).
  // InsertKnotAndUpdateControlPointsRef (bspline.h:234-278, NURBS Alg. 5.1) with
  // CanInsertKnot (bspline_base.cc:166-195).  InsertKnotAndUpdateControlPoints is the
  // same curve; the reference solves its equations by QR (bspline.h:281-399).
  tp::Status InsertKnotAndUpdateControlPointsRef(double knot, int multiplicity = 1) {
    if (multiplicity > static_cast<int>(degree_) + 1) {
      return tp::InvalidArgumentError("Knot multiplicity > degree + 1 not supported.");
    }
    if (num_knots_ + multiplicity > max_num_knots_) {
      return tp::FailedPreconditionError("Knot capacity too small, required: " +
                                         std::to_string(num_knots_ + multiplicity));
    }
    if (num_knots_ < 2) return tp::FailedPreconditionError("Set initial knot vector first.");
    if (knot <= knots_[0] || knot >= knots_[num_knots_ - 1]) {
      return tp::InvalidArgumentError("knot not in range of current knots, knot=" + std::to_string(knot));
    }
    const size_t p = degree_, D = points_dim_;
    ctrl_.resize(static_cast<size_t>(MaxNumPoints()) * D, 0.0);
    std::vector<double> scratch(p * D);
    for (int m = 0; m < multiplicity; ++m) {
      const size_t span = KnotSpan(knot);
      std::copy_backward(ctrl_.begin() + span * D, ctrl_.begin() + num_points_ * D,
                         ctrl_.begin() + (num_points_ + 1) * D);
      for (size_t i = 0; i < p; ++i) {
        const size_t k = span + i - p + 1;
        const double alpha = (knot - knots_[k]) / (knots_[k + p] - knots_[k]);
        for (size_t d = 0; d < D; ++d) {
          scratch[i * D + d] = alpha * ctrl_[k * D + d] + (1.0 - alpha) * ctrl_[(k - 1) * D + d];
        }
      }
      std::copy(scratch.begin(), scratch.end(), ctrl_.begin() + (span - p + 1) * D);
      std::copy_backward(knots_.begin() + span + 1, knots_.begin() + num_knots_,
                         knots_.begin() + num_knots_ + 1);
      knots_[span + 1] = knot;
      num_knots_ += 1;
      num_points_ += 1;
    }
    return tp::OkStatus();
  }
  tp::Status InsertKnotAndUpdateControlPoints(double knot, int multiplicity = 1) {
    return InsertKnotAndUpdateControlPointsRef(knot, multiplicity);
  }
  // TruncateSplineAt (bspline.h:402-427)
  tp::Status TruncateSplineAt(double u_end) {
    if (u_end >= umax_) return tp::OkStatus();
    if (u_end <= umin_) {
      umin_ = std::numeric_limits<double>::infinity();
      umax_ = -std::numeric_limits<double>::infinity();
      num_knots_ = num_points_ = 0;
      return tp::OkStatus();
    }
    if (auto s = InsertKnotAndUpdateControlPointsRef(u_end, static_cast<int>(degree_) + 1); !s.ok()) return s;
    num_knots_ = KnotSpan(u_end) + 1;
    num_points_ = static_cast<size_t>(NumPoints(num_knots_, degree_));
    umax_ = u_end;
    return tp::OkStatus();
  }
  // ExtendWithControlPoints (bspline.h:429-511), degree 2 only
  tp::Status ExtendWithControlPoints(tp::Span<const Point> points) {
    if (degree_ != 2) return tp::UnimplementedError("Only implemented for 2nd order splines.");
    const size_t p = degree_, D = points_dim_, n = points.size();
    const size_t new_num_points = num_points_ + n;
    const size_t new_num_knots = num_knots_ + NumKnots(n + 1, p) - 2 * p;
    if (num_knots_ < MinNumKnots(p)) return tp::FailedPreconditionError("Spline is empty or invalid.");
    if (new_num_points > static_cast<size_t>(MaxNumPoints())) {
      return tp::FailedPreconditionError("Point capacity too small to append all points.");
    }
    if (new_num_knots > max_num_knots_) {
      return tp::FailedPreconditionError("Knot capacity too small to append all points.");
    }
    if (n < 2) return tp::UnimplementedError("Only implemented for >= 2 points.");
    const double u_join = knots_[num_knots_ - 1];
    const double old_range = knots_[num_knots_ - 1] - knots_[0];
    const double new_range = (old_range * static_cast<int>(new_num_knots - 2 * p - 1)) /
                             static_cast<int>(num_knots_ - 2 * p - 1);
    const size_t start = num_knots_ - p - 1, size = (new_num_knots - p) - start;
    const double step = (size == 1) ? 0.0 : (new_range - old_range) / static_cast<double>(size - 1);
    for (size_t i = 0; i < size; ++i) {  // setLinSpaced(old_range, new_range)
      knots_[start + i] = (i == (size == 1 ? 1 : size - 1)) ? new_range : old_range + static_cast<double>(i) * step;
    }
    for (size_t i = 0; i < p + 1; ++i) knots_[new_num_knots - p - 1 + i] = knots_[0] + new_range;
    const size_t old_points = num_points_;
    num_knots_ = new_num_knots;
    umax_ = knots_[new_num_knots - 1];
    // the old end point moves so that the curve keeps its value at the join
    const size_t span = KnotSpan(u_join), mod = old_points - 1;
    const double left1 = u_join - knots_[span], right1 = knots_[span + 1] - u_join;
    const double left2 = u_join - knots_[span - 1], right2 = knots_[span + 2] - u_join;
    double b0 = 1.0, b1, tmp = b0 / (right1 + left1);  // UpdateBasis, degree 2 (A2.2)
    b0 = right1 * tmp;
    b1 = left1 * tmp;
    tmp = b0 / (right1 + left2);
    b0 = right1 * tmp;
    double saved = left2 * tmp;
    tmp = b1 / (right2 + left1);
    b1 = saved + right2 * tmp;
    if (!(b1 > 0)) return tp::InternalError("basis[1] should be > 0 (0.5)");
    ctrl_.resize(static_cast<size_t>(MaxNumPoints()) * D, 0.0);
    for (size_t d = 0; d < D; ++d) {
      ctrl_[mod * D + d] = 1.0 / (b1) * (ctrl_[mod * D + d] - b0 * ctrl_[(mod - 1) * D + d]);
    }
    for (size_t i = 0; i < n; ++i) {
      if (points[i].size() != D) return tp::InvalidArgumentError("Invalid point size");
      std::copy(points[i].begin(), points[i].end(), ctrl_.begin() + (old_points + i) * D);
    }
    num_points_ = new_num_points;
    return tp::OkStatus();
  }
  // KnotSpan (bspline_base.cc:218-246)
  size_t KnotSpan(double u) const {
    if (num_knots_ == 0) return 0;
    if (u == knots_[num_knots_ - 1]) return num_points_ - 1;
    size_t lo = degree_, hi = num_knots_ - degree_;
    while (lo < hi) {
      const size_t mid = lo + (hi - lo) / 2;
      if (knots_[mid] <= u) {
        lo = mid + 1;
      } else {
        hi = mid;
      }
    }
    return lo - 1;
  }

  // EvalCurve (bspline.h:514-536)
  tp::Status EvalCurve(double u, Point& value) {
    std::vector<double> out(points_dim_);
    value.resize(points_dim_);
    std::copy(value.begin(), value.end(), out.begin());
    int32_t st = 0;
    if (auto s = EvalBatch(&u, 1, 0, out.data(), &st); !s.ok()) return s;
    if (st != 0) return tp::OutOfRangeError("Spline parameter " + std::to_string(u) + " outside valid range");
    value = Point(out.data(), points_dim_);
    return tp::OkStatus();
  }
  // EvalCurveAndDerivatives (bspline.h:539-568): values.size()-1 derivatives.
  tp::Status EvalCurveAndDerivatives(double u, tp::Span<Point> values) {
    if (values.empty()) return tp::InvalidArgumentError("values must not be empty.");
    const int nder = static_cast<int>(values.size()) - 1;
    std::vector<double> out(values.size() * points_dim_);
    int32_t st = 0;
    if (auto s = EvalBatch(&u, 1, nder, out.data(), &st); !s.ok()) return s;
    if (st != 0) return tp::OutOfRangeError("Spline parameter " + std::to_string(u) + " outside valid range");
    for (size_t k = 0; k < values.size(); ++k) values[k] = Point(&out[k * points_dim_], points_dim_);
    return tp::OkStatus();
  }
  // Batched forms: out is [n][nder+1][dim]; status (may be null) is [n] with 0 or
  // kOutOfRange (11) per parameter.
  tp::Status EvalCurveAndDerivativesBatch(tp::Span<const double> u, int nder, double* out, int32_t* status) {
    return EvalBatch(u.data(), u.size(), nder, out, status);
  }
  tp::Status EvalCurveBatch(tp::Span<const double> u, double* out, int32_t* status) {
    return EvalBatch(u.data(), u.size(), 0, out, status);
  }
  static void SetDevice(int device) { Device() = device; }

 private:
  static int& Device() {
    static int d = 0;
    return d;
  }
  tp::Status EvalBatch(const double* u, size_t n, int nder, double* out, int32_t* status) {
    if (!init_) return tp::FailedPreconditionError("Call Init first.");
    if (ctrl_.empty() || num_knots_ == 0) return tp::FailedPreconditionError("Set knots and control points first.");
    if (nder > static_cast<int>(degree_)) {
      // the reference CHECK-aborts here (bspline_base.cc:270); report instead
      return tp::InvalidArgumentError("number of derivatives exceeds the degree");
    }
    // the reference range-checks against [umin_, umax_], which may be tighter than
    // the knots (bspline.h:519-522); the device checks against the knots
    std::vector<int32_t> st(n, 0);
    tpb_spline_eval_desc d{};
    d.degree = static_cast<int32_t>(degree_);
    d.num_points = static_cast<int32_t>(num_points_);
    d.dim = static_cast<int32_t>(points_dim_);
    d.nder = nder;
    d.num_splines = 1;
    d.evals_per_spline = static_cast<int64_t>(n);
    d.knots = knots_.data();
    d.ctrl = ctrl_.data();
    d.u = u;
    d.out = out;
    d.status = st.data();
    d.exact = 1;
    const int32_t rc = tpb_spline_eval_host(&d, Device());
    if (rc != TPB_OK) return tp::FromCode(rc, "tpb_spline_eval_host failed (no CUDA device?)");
    for (size_t i = 0; i < n; ++i) {
      if (u[i] < umin_ || u[i] > umax_) st[i] = TPB_OUT_OF_RANGE;
      if (status) status[i] = st[i];
    }
    return tp::OkStatus();
  }

  bool init_ = false;
  size_t degree_ = 0, max_num_knots_ = 0, points_dim_ = 0, num_knots_ = 0, num_points_ = 0;
  double umin_ = 0, umax_ = 0;
  std::vector<double> knots_, ctrl_;
};

// Fixed-size variants of the reference are the same evaluator with a fixed dim.
using BSplineNd = BSplineXd;

}  // namespace trajectory_planning
#endif  // TRAJECTORY_PLANNING_SPLINES_BSPLINE_H_
