// ORACLE BUILD SHIM (test infrastructure).  Minimal stand-in for <mujoco/mujoco.h> so that the
// reference's own mjpc/norm.cc compiles UNMODIFIED from /root/reference into oracle/_ref/.  Only the
// mju_* scalar helpers that file uses are declared; semantics follow MuJoCo 3.1.4's public
// engine_util_blas / engine_util_misc (mju_pow = pow, mju_sign = sign with sign(0) = 0, ...).
#ifndef ORACLE_SHIM_MUJOCO_H_
#define ORACLE_SHIM_MUJOCO_H_
#include <cmath>
#include <cstdio>
#include <cstdlib>
typedef double mjtNum;
#define mjMINVAL 1E-15
inline mjtNum mju_abs(mjtNum x) { return std::fabs(x); }
inline mjtNum mju_pow(mjtNum x, mjtNum y) { return std::pow(x, y); }
inline mjtNum mju_exp(mjtNum x) { return std::exp(x); }
inline mjtNum mju_log(mjtNum x) { return std::log(x); }
inline mjtNum mju_sqrt(mjtNum x) { return std::sqrt(x); }
inline mjtNum mju_sign(mjtNum x) { return x < 0 ? -1 : (x > 0 ? 1 : 0); }
inline mjtNum mju_dot(const mjtNum* a, const mjtNum* b, int n) { mjtNum s = 0; for (int i = 0; i < n; i++) s += a[i] * b[i]; return s; }
inline void mju_zero(mjtNum* r, int n) { for (int i = 0; i < n; i++) r[i] = 0; }
inline void mju_scl(mjtNum* r, const mjtNum* v, mjtNum s, int n) { for (int i = 0; i < n; i++) r[i] = v[i] * s; }
inline mjtNum mju_norm(const mjtNum* v, int n) { return std::sqrt(mju_dot(v, v, n)); }
inline void mju_error(const char* msg, ...) { std::fprintf(stderr, "mju_error: %s\n", msg); std::abort(); }
#endif
