// ORACLE (test infrastructure): C entry points over the reference's OWN norm.cc and
// spline/spline.cc, compiled unmodified from /root/reference (recipe: oracle/Makefile target
// _ref/libmjpc_ref.so).  Used to pin the restatement in mm_cost.cc and to generate tests/golden/.
#include <vector>
#include "mjpc/norm.h"
#include "mjpc/spline/spline.h"

extern "C" {
double ref_norm(const double* x, const double* params, int n, int type) {
  return mjpc::Norm(nullptr, nullptr, x, params, n, (mjpc::NormType)type);
}
int ref_norm_parameter_dimension(int type) { return mjpc::NormParameterDimension(type); }
// TimeSpline built by AddNode in time order, then Sample (mjpc/spline/spline.cc:103-156)
void ref_spline_sample(const double* times, const double* knots, int P, int dim, int interp, double t, double* out) {
  mjpc::spline::TimeSpline s(dim, (mjpc::spline::SplineInterpolation)interp);
  s.Reserve(P);
  for (int i = 0; i < P; i++) s.AddNode(times[i], absl::Span<const double>(knots + (size_t)i * dim, dim));
  s.Sample(t, absl::Span<double>(out, dim));
}
}
