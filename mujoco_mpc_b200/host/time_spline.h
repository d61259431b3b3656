// Host-side TimeSpline for the planner's nominal / previous policies.
//
// Same observable behaviour as mjpc::spline::TimeSpline (mjpc/spline/spline.h:37-300,
// spline.cc:103-277) for the operations the sampling planner uses -- Sample (zero / linear / cubic
// Hermite with finite-difference slopes), AddNode at either end, Clear, DiscardBefore, Size --
// but stored as two flat vectors (times, values) because the plan is re-uploaded to the GPU as a
// contiguous [P][dim] block every iteration; the reference's ring buffer buys nothing here.
#ifndef MJPC_B200_TIME_SPLINE_H_
#define MJPC_B200_TIME_SPLINE_H_
#include <algorithm>
#include <vector>

namespace mjpc_b200 {

enum SplineInterpolation : int { kZeroSpline = 0, kLinearSpline = 1, kCubicSpline = 2 };

class TimeSpline {
 public:
  explicit TimeSpline(int dim = 0, SplineInterpolation interp = kZeroSpline) : dim_(dim), interp_(interp) {}
  int Dim() const { return dim_; }
  int Size() const { return (int)times_.size(); }
  SplineInterpolation Interpolation() const { return interp_; }
  void SetInterpolation(SplineInterpolation i) { interp_ = i; }
  void Clear() { times_.clear(); values_.clear(); }
  const std::vector<double>& times() const { return times_; }
  const std::vector<double>& values() const { return values_; }   // [Size][Dim]
  double* NodeValues(int i) { return values_.data() + (size_t)i * dim_; }
  const double* NodeValues(int i) const { return values_.data() + (size_t)i * dim_; }
  double NodeTime(int i) const { return times_[i]; }

  // adds a node after the last or before the first node; returns its index, -1 if time is interior
  int AddNode(double time, const double* values = nullptr) {
    if (times_.empty() || time > times_.back()) {
      times_.push_back(time); values_.insert(values_.end(), dim_, 0.0);
      if (values) std::copy(values, values + dim_, values_.end() - dim_);
      return Size() - 1;
    }
    if (time < times_.front()) {
      times_.insert(times_.begin(), time); values_.insert(values_.begin(), dim_, 0.0);
      if (values) std::copy(values, values + dim_, values_.begin());
      return 0;
    }
    return -1;
  }

  // removes nodes that cannot influence Sample(t >= time); the cubic spline keeps one extra node
  int DiscardBefore(double time) {
    int first = (int)(std::upper_bound(times_.begin(), times_.end(), time) - times_.begin());
    int keep_from = std::max(0, first - 1);                  // last node with t <= time
    if (interp_ == kCubicSpline) keep_from = std::max(0, keep_from - 1);
    times_.erase(times_.begin(), times_.begin() + keep_from);
    values_.erase(values_.begin(), values_.begin() + (size_t)keep_from * dim_);
    return keep_from;
  }

  void Sample(double time, double* out) const {
    const int P = Size();
    if (P == 0) { std::fill(out, out + dim_, 0.0); return; }
    const int up = (int)(std::upper_bound(times_.begin(), times_.end(), time) - times_.begin());
    if (up == P) { std::copy(NodeValues(P - 1), NodeValues(P - 1) + dim_, out); return; }
    if (up == 0) { std::copy(NodeValues(0), NodeValues(0) + dim_, out); return; }
    const int lo = up - 1;
    const double* pl = NodeValues(lo); const double* pu = NodeValues(up);
    if (interp_ == kZeroSpline) { std::copy(pl, pl + dim_, out); return; }
    const double dt = times_[up] - times_[lo];
    const double t = (time - times_[lo]) / dt;
    if (interp_ == kLinearSpline) { for (int i = 0; i < dim_; i++) out[i] = pl[i] * (1 - t) + pu[i] * t; return; }
    const double c0 = 2.0 * t * t * t - 3.0 * t * t + 1.0, c1 = (t * t * t - 2.0 * t * t + t) * dt;
    const double c2 = -2.0 * t * t * t + 3 * t * t, c3 = (t * t * t - t * t) * dt;
    for (int i = 0; i < dim_; i++) out[i] = c0 * pl[i] + c1 * Slope(lo, i) + c2 * pu[i] + c3 * Slope(up, i);
  }

 private:
  double Slope(int n, int i) const {
    const double v = NodeValues(n)[i];
    if (n == 0) return (NodeValues(1)[i] - v) / (times_[1] - times_[0]);
    const double pv = NodeValues(n - 1)[i];
    if (n == Size() - 1) return (v - pv) / (times_[n] - times_[n - 1]);
    return 0.5 * (NodeValues(n + 1)[i] - v) / (times_[n + 1] - times_[n]) + 0.5 * (v - pv) / (times_[n] - times_[n - 1]);
  }
  int dim_; SplineInterpolation interp_;
  std::vector<double> times_, values_;
};

}  // namespace mjpc_b200
#endif
