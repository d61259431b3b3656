// ORACLE (test infrastructure, not product): small fixed-size math helpers for the CPU
// restatement of the MuJoCo 3.1.4 forward step.  MuJoCo's source is NOT under /root/reference
// (fetched at configure time, /root/reference/CMakeLists.txt:58-88); these follow the public
// engine_util_blas / engine_util_spatial conventions (row-major 3x3, quaternion w-x-y-z,
// spatial vectors [angular; linear]).  Parity vs real MuJoCo: UNPINNED.
#ifndef ORACLE_MM_MATH_H_
#define ORACLE_MM_MATH_H_
#include <cmath>

namespace mm {
constexpr double kMinVal = 1e-15;   // mjMINVAL
constexpr double kMaxVal = 1e10;    // mjMAXVAL
constexpr double kPi = 3.14159265358979323846;

inline void copy(double* r, const double* a, int n) { for (int i = 0; i < n; i++) r[i] = a[i]; }
inline void zero(double* r, int n) { for (int i = 0; i < n; i++) r[i] = 0; }
inline double dot(const double* a, const double* b, int n) { double s = 0; for (int i = 0; i < n; i++) s += a[i] * b[i]; return s; }
inline double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline void cross(double* r, const double* a, const double* b) {
  r[0] = a[1] * b[2] - a[2] * b[1]; r[1] = a[2] * b[0] - a[0] * b[2]; r[2] = a[0] * b[1] - a[1] * b[0];
}
inline double norm3(const double* a) { return std::sqrt(dot3(a, a)); }
inline double normalize3(double* a) {
  double n = norm3(a);
  if (n < kMinVal) { a[0] = 1; a[1] = 0; a[2] = 0; } else { const double inv = 1 / n; a[0] *= inv; a[1] *= inv; a[2] *= inv; }
  return n;
}
inline double normalize4(double* q) {
  double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < kMinVal) { q[0] = 1; q[1] = q[2] = q[3] = 0; }
  else if (std::fabs(n - 1) > kMinVal) { const double inv = 1 / n; for (int i = 0; i < 4; i++) q[i] *= inv; }
  return n;
}
// r = M * v (M row-major 3x3)
inline void mulMatVec3(double* r, const double* M, const double* v) {
  r[0] = M[0] * v[0] + M[1] * v[1] + M[2] * v[2];
  r[1] = M[3] * v[0] + M[4] * v[1] + M[5] * v[2];
  r[2] = M[6] * v[0] + M[7] * v[1] + M[8] * v[2];
}
// r = M' * v
inline void mulMatTVec3(double* r, const double* M, const double* v) {
  r[0] = M[0] * v[0] + M[3] * v[1] + M[6] * v[2];
  r[1] = M[1] * v[0] + M[4] * v[1] + M[7] * v[2];
  r[2] = M[2] * v[0] + M[5] * v[1] + M[8] * v[2];
}
inline void mulQuat(double* r, const double* a, const double* b) {
  double t[4] = {a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3],
                 a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
                 a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1],
                 a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]};
  copy(r, t, 4);
}
inline void quat2Mat(double* m, const double* q) {
  double q00 = q[0] * q[0], q01 = q[0] * q[1], q02 = q[0] * q[2], q03 = q[0] * q[3];
  double q11 = q[1] * q[1], q12 = q[1] * q[2], q13 = q[1] * q[3];
  double q22 = q[2] * q[2], q23 = q[2] * q[3], q33 = q[3] * q[3];
  m[0] = q00 + q11 - q22 - q33; m[4] = q00 - q11 + q22 - q33; m[8] = q00 - q11 - q22 + q33;
  m[1] = 2 * (q12 - q03); m[2] = 2 * (q13 + q02); m[3] = 2 * (q12 + q03);
  m[5] = 2 * (q23 - q01); m[6] = 2 * (q13 - q02); m[7] = 2 * (q23 + q01);
}
// rotate vector by quaternion
inline void rotVecQuat(double* r, const double* v, const double* q) {
  double m[9]; quat2Mat(m, q); double t[3]; mulMatVec3(t, m, v); copy(r, t, 3);
}
// sin/cos restated with IEEE +,-,* only (fdlibm-style: Cody-Waite reduction by pi/2 with the
// hi/mid/lo split of pi/2, then the k_sin / k_cos minimax kernels; < 1 ulp).  The reference calls
// libm's sin/cos (via MuJoCo's mju_sin/mju_cos); libm implementations disagree with each other in
// the last bit for a few percent of arguments, so the oracle pins one published algorithm instead
// of whatever libm the host has.  DEVIATION from the reference: <= 1 ulp per call.
inline void sincosDet(double x, double* sn, double* cs) {
  if (!(std::fabs(x) < 1.0e6)) { *sn = std::sin(x); *cs = std::cos(x); return; }
  static const double invpio2 = 6.36619772367581382433e-01, pio2_1 = 1.57079632673412561417e+00,
                      pio2_2 = 6.07710050630396597660e-11, pio2_2t = 2.02226624879595063154e-21;
  static const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03, S3 = -1.98412698298579493134e-04,
                      S4 = 2.75573137070700676789e-06, S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
  static const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03, C3 = 2.48015872894767294178e-05,
                      C4 = -2.75573143513906633035e-07, C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
  double fn = std::floor(x * invpio2 + 0.5);
  double t = x - fn * pio2_1;
  double w = fn * pio2_2;
  double r = t - w;
  w = fn * pio2_2t - ((t - r) - w);
  double y0 = r - w, y1 = (r - y0) - w;
  double z = y0 * y0;
  double v = z * y0;
  double rs = S2 + z * (S3 + z * (S4 + z * (S5 + z * S6)));
  double ks = y0 - ((z * (0.5 * y1 - v * rs) - y1) - v * S1);
  double ww = z * z;
  double rc = z * (C1 + z * (C2 + z * C3)) + (ww * ww) * (C4 + z * (C5 + z * C6));
  double hz = 0.5 * z, wc = 1.0 - hz;
  double kc = wc + (((1.0 - wc) - hz) + (z * rc - y0 * y1));
  long long n = (long long)fn;
  switch ((int)(n & 3)) {
    case 0: *sn = ks; *cs = kc; break;
    case 1: *sn = kc; *cs = -ks; break;
    case 2: *sn = -ks; *cs = -kc; break;
    default: *sn = -kc; *cs = ks; break;
  }
}
inline double sinDet(double x) { double s, c; sincosDet(x, &s, &c); return s; }
inline double cosDet(double x) { double s, c; sincosDet(x, &s, &c); return c; }
inline void axisAngle2Quat(double* q, const double* axis, double angle) {
  if (angle == 0) { q[0] = 1; q[1] = q[2] = q[3] = 0; return; }
  double s, c; sincosDet(angle * 0.5, &s, &c);
  q[0] = c; q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
}
// integrate quaternion by angular velocity (body frame) * dt   (mju_quatIntegrate)
inline void quatIntegrate(double* q, const double* vel, double dt) {
  double ax[3] = {vel[0], vel[1], vel[2]};
  double ang = dt * normalize3(ax);
  double qr[4]; axisAngle2Quat(qr, ax, ang);
  normalize4(q); mulQuat(q, q, qr);
}
// spatial cross products (mju_crossMotion / mju_crossForce)
inline void crossMotion(double* r, const double* vel, const double* v) {
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
inline void crossForce(double* r, const double* vel, const double* f) {
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
// 10-number spatial inertia times spatial vector (mju_mulInertVec)
inline void mulInertVec(double* r, const double* i, const double* v) {
  r[0] = i[0] * v[0] + i[3] * v[1] + i[4] * v[2] - i[8] * v[4] + i[7] * v[5];
  r[1] = i[3] * v[0] + i[1] * v[1] + i[5] * v[2] + i[8] * v[3] - i[6] * v[5];
  r[2] = i[4] * v[0] + i[5] * v[1] + i[2] * v[2] - i[7] * v[3] + i[6] * v[4];
  r[3] = i[8] * v[1] - i[7] * v[2] + i[9] * v[3];
  r[4] = i[6] * v[2] - i[8] * v[0] + i[9] * v[4];
  r[5] = i[7] * v[0] - i[6] * v[1] + i[9] * v[5];
}
inline double clip(double x, double lo, double hi) { return x < lo ? lo : (x > hi ? hi : x); }
}  // namespace mm
#endif
