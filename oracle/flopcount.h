// ORACLE (test infrastructure): counting stand-in for `double`.  Force-included (g++ -include) when building
// oracle/libmm_oracle_flops.so (Makefile target `flops`): every `double` of the oracle sources becomes a `cdouble` whose
// +, -, *, /, sqrt are tallied, so that one oracle rollout yields the ALGORITHMIC FP64 operation count of a sample-step
//  -- the numerator of the roofline fraction bench.py reports (tools/count_flops.py ->
// baseline/flops_<task>.json).  Transcendentals are charged a nominal operation count (exp / log 20, sin / cos 30,
// pow / cosh / atan2 / acos 40) and reported separately.  Same layout as double (the C ABI is unchanged).
#pragma once
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <algorithm>
#include <thread>
#include <mutex>
#include <atomic>
#include <functional>
#include <condition_variable>
#include <numeric>
#include <limits>
#include <memory>
#include <queue>
struct FlopCounter { long long add = 0, mul = 0, div = 0, sqrt_ = 0, special = 0; };
extern FlopCounter g_flops;     // single-threaded use (nthread = 1)
struct cdouble {
  double v;
  cdouble() = default;
  constexpr cdouble(double x) : v(x) {}
  constexpr cdouble(int x) : v(x) {}
  constexpr cdouble(long long x) : v((double)x) {}
  constexpr cdouble(long x) : v((double)x) {}
  constexpr cdouble(unsigned x) : v((double)x) {}
  constexpr cdouble(unsigned long x) : v((double)x) {}
  constexpr cdouble(float x) : v(x) {}
  constexpr operator double() const { return v; }
  cdouble& operator+=(cdouble o) { g_flops.add++; v += o.v; return *this; }
  cdouble& operator-=(cdouble o) { g_flops.add++; v -= o.v; return *this; }
  cdouble& operator*=(cdouble o) { g_flops.mul++; v *= o.v; return *this; }
  cdouble& operator/=(cdouble o) { g_flops.div++; v /= o.v; return *this; }
};
#define CD_BIN(op, ctr) \
  inline cdouble operator op(cdouble a, cdouble b) { g_flops.ctr++; return cdouble(a.v op b.v); } \
  inline cdouble operator op(cdouble a, double b) { g_flops.ctr++; return cdouble(a.v op b); } \
  inline cdouble operator op(double a, cdouble b) { g_flops.ctr++; return cdouble(a op b.v); } \
  inline cdouble operator op(cdouble a, int b) { g_flops.ctr++; return cdouble(a.v op b); } \
  inline cdouble operator op(int a, cdouble b) { g_flops.ctr++; return cdouble(a op b.v); }
CD_BIN(+, add) CD_BIN(-, add) CD_BIN(*, mul) CD_BIN(/, div)
inline cdouble operator-(cdouble a) { return cdouble(-a.v); }
inline cdouble operator+(cdouble a) { return a; }
#define CD_CMP(op) \
  inline bool operator op(cdouble a, cdouble b) { return a.v op b.v; } \
  inline bool operator op(cdouble a, double b) { return a.v op b; } \
  inline bool operator op(double a, cdouble b) { return a op b.v; } \
  inline bool operator op(cdouble a, int b) { return a.v op b; } \
  inline bool operator op(int a, cdouble b) { return a op b.v; }
CD_CMP(<) CD_CMP(>) CD_CMP(<=) CD_CMP(>=) CD_CMP(==) CD_CMP(!=)
namespace std {
inline cdouble sqrt(cdouble a) { g_flops.sqrt_++; return cdouble(::sqrt(a.v)); }
inline cdouble fabs(cdouble a) { return cdouble(::fabs(a.v)); }
inline cdouble abs(cdouble a) { return cdouble(::fabs(a.v)); }
inline cdouble fmax(cdouble a, cdouble b) { return cdouble(::fmax(a.v, b.v)); }
inline cdouble fmin(cdouble a, cdouble b) { return cdouble(::fmin(a.v, b.v)); }
inline cdouble fmax(cdouble a, double b) { return cdouble(::fmax(a.v, b)); }
inline cdouble fmax(double a, cdouble b) { return cdouble(::fmax(a, b.v)); }
inline cdouble fmin(cdouble a, double b) { return cdouble(::fmin(a.v, b)); }
inline cdouble fmin(double a, cdouble b) { return cdouble(::fmin(a, b.v)); }
inline cdouble exp(cdouble a) { g_flops.special += 20; return cdouble(::exp(a.v)); }
inline cdouble log(cdouble a) { g_flops.special += 20; return cdouble(::log(a.v)); }
inline cdouble pow(cdouble a, cdouble b) { g_flops.special += 40; return cdouble(::pow(a.v, b.v)); }
inline cdouble pow(cdouble a, double b) { g_flops.special += 40; return cdouble(::pow(a.v, b)); }
inline cdouble pow(cdouble a, int b) { g_flops.special += 40; return cdouble(::pow(a.v, b)); }
inline cdouble cosh(cdouble a) { g_flops.special += 40; return cdouble(::cosh(a.v)); }
inline cdouble floor(cdouble a) { return cdouble(::floor(a.v)); }
inline cdouble sin(cdouble a) { g_flops.special += 30; return cdouble(::sin(a.v)); }
inline cdouble cos(cdouble a) { g_flops.special += 30; return cdouble(::cos(a.v)); }
inline cdouble atan2(cdouble a, cdouble b) { g_flops.special += 40; return cdouble(::atan2(a.v, b.v)); }
inline cdouble acos(cdouble a) { g_flops.special += 40; return cdouble(::acos(a.v)); }
inline bool isnan(cdouble a) { return a.v != a.v; }
inline bool isfinite(cdouble a) { return a.v - a.v == 0; }
inline cdouble max(cdouble a, double b) { return a.v < b ? cdouble(b) : a; }
inline cdouble max(double a, cdouble b) { return a < b.v ? b : cdouble(a); }
inline cdouble min(cdouble a, double b) { return b < a.v ? cdouble(b) : a; }
inline cdouble min(double a, cdouble b) { return b.v < a ? b : cdouble(a); }
}
static_assert(sizeof(cdouble) == 8, "layout");
#define double cdouble
