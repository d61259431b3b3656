// Small fixed-size FP64 helpers for the device rigid-body step (row-major 3x3, quaternions w-x-y-z,
// spatial vectors [angular; linear] in the subtree-CoM frame -- the conventions of MuJoCo's
// engine_util_{blas,spatial}, which the reference's rollouts run on via mj_step,
// mjpc/trajectory.cc:158).
#pragma once
#include "ps_types.cuh"

namespace ps {

PS_HD void vcopy(double* r, const double* a, int n) { PS_SPEC_UNROLL for (int i = 0; i < n; i++) r[i] = a[i]; }
PS_HD void vzero(double* r, int n) { PS_SPEC_UNROLL for (int i = 0; i < n; i++) r[i] = 0.0; }
PS_HD double vdot(const double* a, const double* b, int n) { double s = 0; PS_SPEC_UNROLL for (int i = 0; i < n; i++) s += a[i] * b[i]; return s; }
PS_HD double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
PS_HD void cross3(double* r, const double* a, const double* b) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
PS_HD double norm3(const double* a) { return sqrt(dot3(a, a)); }
PS_HD double unit3(double* a) {
  double n = norm3(a);
  if (n < kMinVal) { a[0] = 1; a[1] = 0; a[2] = 0; } else { const double inv = 1 / n; a[0] *= inv; a[1] *= inv; a[2] *= inv; }
  return n;
}
PS_HD void unit4(double* q) {
  double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < kMinVal) { q[0] = 1; q[1] = 0; q[2] = 0; q[3] = 0; }
  else if (fabs(n - 1) > kMinVal) { const double inv = 1 / n; q[0] *= inv; q[1] *= inv; q[2] *= inv; q[3] *= inv; }
}
PS_HD void mat_vec3(double* r, const double* M, const double* v) {
  double x = M[0] * v[0] + M[1] * v[1] + M[2] * v[2];
  double y = M[3] * v[0] + M[4] * v[1] + M[5] * v[2];
  double z = M[6] * v[0] + M[7] * v[1] + M[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
PS_HD void matT_vec3(double* r, const double* M, const double* v) {
  double x = M[0] * v[0] + M[3] * v[1] + M[6] * v[2];
  double y = M[1] * v[0] + M[4] * v[1] + M[7] * v[2];
  double z = M[2] * v[0] + M[5] * v[1] + M[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
PS_HD void quat_mul(double* r, const double* a, const double* b) {
  double w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  double x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  double y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  double z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
PS_HD void quat_to_mat(double* m, const double* q) {
  double q00 = q[0] * q[0], q01 = q[0] * q[1], q02 = q[0] * q[2], q03 = q[0] * q[3];
  double q11 = q[1] * q[1], q12 = q[1] * q[2], q13 = q[1] * q[3];
  double q22 = q[2] * q[2], q23 = q[2] * q[3], q33 = q[3] * q[3];
  m[0] = q00 + q11 - q22 - q33; m[4] = q00 - q11 + q22 - q33; m[8] = q00 - q11 - q22 + q33;
  m[1] = 2 * (q12 - q03); m[2] = 2 * (q13 + q02); m[3] = 2 * (q12 + q03);
  m[5] = 2 * (q23 - q01); m[6] = 2 * (q13 - q02); m[7] = 2 * (q23 + q01);
}
PS_HD void quat_rotate(double* r, const double* v, const double* q) {
  double m[9]; quat_to_mat(m, q); mat_vec3(r, m, v);
}
// sin and cos from IEEE +,-,* only (two-term Cody-Waite reduction by pi/2 carried to ~118 bits, then
// the classic minimax kernels on [-pi/4, pi/4]; error < 1 ulp).  Vendor libm sin/cos differ from
// one another in the last bit for a few percent of arguments, and contact-rich rollouts amplify
// a one-ulp difference by ten orders of magnitude; evaluating the pair this way makes the
// rollout independent of the math library underneath it.
PS_NT void sincos_det(double x, double* sn, double* cs) {
  if (!(fabs(x) < 1.0e6)) { *sn = sin(x); *cs = cos(x); return; }   // out of the reduction's range (never an angle)
  const double fn = floor(x * 6.36619772367581382433e-01 + 0.5);
  const double t = x - fn * 1.57079632673412561417e+00;
  double w = fn * 6.07710050630396597660e-11;
  const double r = t - w;
  w = fn * 2.02226624879595063154e-21 - ((t - r) - w);
  const double y0 = r - w, y1 = (r - y0) - w;
  const double z = y0 * y0;
  // sine kernel
  const double v = z * y0;
  const double rs = 8.33333333332248946124e-03 + z * (-1.98412698298579493134e-04 + z * (2.75573137070700676789e-06 +
                    z * (-2.50507602534068634195e-08 + z * 1.58969099521155010221e-10)));
  const double ks = y0 - ((z * (0.5 * y1 - v * rs) - y1) - v * -1.66666666666666324348e-01);
  // cosine kernel
  const double ww = z * z;
  const double rc = z * (4.16666666666666019037e-02 + z * (-1.38888888888741095749e-03 + z * 2.48015872894767294178e-05)) +
                    (ww * ww) * (-2.75573143513906633035e-07 + z * (2.08757232129817482790e-09 + z * -1.13596475577881948265e-11));
  const double hz = 0.5 * z, wc = 1.0 - hz;
  const double kc = wc + (((1.0 - wc) - hz) + (z * rc - y0 * y1));
  const long long n = (long long)fn;
  switch ((int)(n & 3)) {
    case 0: *sn = ks; *cs = kc; break;
    case 1: *sn = kc; *cs = -ks; break;
    case 2: *sn = -ks; *cs = -kc; break;
    default: *sn = -kc; *cs = ks; break;
  }
}
PS_HD double sin_det(double x) { double s, c; sincos_det(x, &s, &c); return s; }
PS_HD double cos_det(double x) { double s, c; sincos_det(x, &s, &c); return c; }

PS_HD void axis_angle_quat(double* q, const double* axis, double angle) {
  if (angle == 0) { q[0] = 1; q[1] = 0; q[2] = 0; q[3] = 0; return; }
  double s, c; sincos_det(angle * 0.5, &s, &c);
  q[0] = c; q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
}
// q <- q * exp(dt * w / 2), w in the body frame (free / ball joint position update)
PS_HD void quat_integrate(double* q, const double* w, double dt) {
  double ax[3] = {w[0], w[1], w[2]};
  double ang = dt * unit3(ax);
  double dq[4]; axis_angle_quat(dq, ax, ang);
  unit4(q); quat_mul(q, q, dq);
}
// spatial motion cross product  r = vel x v
PS_HD void cross_motion(double* r, const double* vel, const double* v) {
  r[0] = -vel[2] * v[1] + vel[1] * v[2];
  r[1] = vel[2] * v[0] - vel[0] * v[2];
  r[2] = -vel[1] * v[0] + vel[0] * v[1];
  r[3] = -vel[2] * v[4] + vel[1] * v[5];
  r[4] = vel[2] * v[3] - vel[0] * v[5];
  r[5] = -vel[1] * v[3] + vel[0] * v[4];
  r[3] += -vel[5] * v[1] + vel[4] * v[2];
  r[4] += vel[5] * v[0] - vel[3] * v[2];
  r[5] += -vel[4] * v[0] + vel[3] * v[1];
}
// spatial force cross product  r = vel x* f
PS_HD void cross_force(double* r, const double* vel, const double* f) {
  r[0] = -vel[2] * f[1] + vel[1] * f[2];
  r[1] = vel[2] * f[0] - vel[0] * f[2];
  r[2] = -vel[1] * f[0] + vel[0] * f[1];
  r[3] = -vel[2] * f[4] + vel[1] * f[5];
  r[4] = vel[2] * f[3] - vel[0] * f[5];
  r[5] = -vel[1] * f[3] + vel[0] * f[4];
  r[0] += -vel[5] * f[4] + vel[4] * f[5];
  r[1] += vel[5] * f[3] - vel[3] * f[5];
  r[2] += -vel[4] * f[3] + vel[3] * f[4];
}
// 10-number spatial inertia (Ixx Iyy Izz Ixy Ixz Iyz mx my mz m) times spatial vector
PS_HD void inert_vec(double* r, const double* i, const double* v) {
  r[0] = i[0] * v[0] + i[3] * v[1] + i[4] * v[2] - i[8] * v[4] + i[7] * v[5];
  r[1] = i[3] * v[0] + i[1] * v[1] + i[5] * v[2] + i[8] * v[3] - i[6] * v[5];
  r[2] = i[4] * v[0] + i[5] * v[1] + i[2] * v[2] - i[7] * v[3] + i[6] * v[4];
  r[3] = i[8] * v[1] - i[7] * v[2] + i[9] * v[3];
  r[4] = i[6] * v[2] - i[8] * v[0] + i[9] * v[4];
  r[5] = i[7] * v[0] - i[6] * v[1] + i[9] * v[5];
}
PS_HD double clampd(double x, double lo, double hi) { return x < lo ? lo : (x > hi ? hi : x); }
PS_HD bool bad_number(double x) { return !(x == x) || x > kMaxVal || x < -kMaxVal; }

}  // namespace ps
