/* count.h — CPU ORACLE instrumentation (test infrastructure): an operation-counting stand-in for `double`.
 * oracle/count/pbo_count.cpp includes the oracle's C sources with `double` redefined to this type, so every floating-point
 * operation the FP64 restatement executes is counted ("the official figure is counted, not assumed").
 * Counted: add/sub, mul, div, sqrt, transcendental (sin, cos, pow, log, exp) and comparisons separately; FLOPs = add + mul +
 * div + sqrt + transcendental (one each).  Thread-local counters, summed by pbo_count_get(). */
#pragma once
#include <math.h>
#include <stdint.h>

struct pbo_opcount { uint64_t add, mul, div, sqrt_, trans, cmp; };
extern thread_local pbo_opcount g_ops;

struct cdouble {
  double v;
  cdouble() = default;
  cdouble(double x) : v(x) {}
  cdouble(float x) : v(x) {}
  cdouble(int x) : v(x) {}
  cdouble(long x) : v((double)x) {}
  cdouble(unsigned x) : v(x) {}
  cdouble(unsigned long x) : v((double)x) {}
  explicit operator double() const { return v; }
  explicit operator float() const { return (float)v; }
  explicit operator int() const { return (int)v; }
  explicit operator long() const { return (long)v; }
  explicit operator bool() const { return v != 0; }
  cdouble &operator+=(cdouble o) { g_ops.add++; v += o.v; return *this; }
  cdouble &operator-=(cdouble o) { g_ops.add++; v -= o.v; return *this; }
  cdouble &operator*=(cdouble o) { g_ops.mul++; v *= o.v; return *this; }
  cdouble &operator/=(cdouble o) { g_ops.div++; v /= o.v; return *this; }
};
inline cdouble operator+(cdouble a, cdouble b) { g_ops.add++; return cdouble(a.v + b.v); }
inline cdouble operator-(cdouble a, cdouble b) { g_ops.add++; return cdouble(a.v - b.v); }
inline cdouble operator*(cdouble a, cdouble b) { g_ops.mul++; return cdouble(a.v * b.v); }
inline cdouble operator/(cdouble a, cdouble b) { g_ops.div++; return cdouble(a.v / b.v); }
inline cdouble operator-(cdouble a) { return cdouble(-a.v); }
inline cdouble operator+(cdouble a) { return a; }
#define PBO_MIX(op) \
  inline cdouble operator op(cdouble a, double b) { return a op cdouble(b); } inline cdouble operator op(double a, cdouble b) { return cdouble(a) op b; } \
  inline cdouble operator op(cdouble a, int b) { return a op cdouble(b); } inline cdouble operator op(int a, cdouble b) { return cdouble(a) op b; } \
  inline cdouble operator op(cdouble a, float b) { return a op cdouble(b); } inline cdouble operator op(float a, cdouble b) { return cdouble(a) op b; } \
  inline cdouble operator op(cdouble a, long b) { return a op cdouble(b); } inline cdouble operator op(long a, cdouble b) { return cdouble(a) op b; }
PBO_MIX(+) PBO_MIX(-) PBO_MIX(*) PBO_MIX(/)
#define PBO_CMP(op) \
  inline bool operator op(cdouble a, cdouble b) { g_ops.cmp++; return a.v op b.v; } inline bool operator op(cdouble a, double b) { g_ops.cmp++; return a.v op b; } \
  inline bool operator op(double a, cdouble b) { g_ops.cmp++; return a op b.v; } inline bool operator op(cdouble a, int b) { g_ops.cmp++; return a.v op b; } \
  inline bool operator op(int a, cdouble b) { g_ops.cmp++; return a op b.v; } inline bool operator op(cdouble a, float b) { g_ops.cmp++; return a.v op b; } \
  inline bool operator op(float a, cdouble b) { g_ops.cmp++; return a op b.v; }
PBO_CMP(<) PBO_CMP(>) PBO_CMP(<=) PBO_CMP(>=) PBO_CMP(==) PBO_CMP(!=)
inline cdouble sqrt(cdouble a) { g_ops.sqrt_++; return cdouble(::sqrt(a.v)); }
inline cdouble fabs(cdouble a) { return cdouble(::fabs(a.v)); }
inline cdouble fmax(cdouble a, cdouble b) { g_ops.cmp++; return cdouble(::fmax(a.v, b.v)); }
inline cdouble fmin(cdouble a, cdouble b) { g_ops.cmp++; return cdouble(::fmin(a.v, b.v)); }
inline cdouble fmax(cdouble a, double b) { return fmax(a, cdouble(b)); }
inline cdouble fmax(double a, cdouble b) { return fmax(cdouble(a), b); }
inline cdouble sin(cdouble a) { g_ops.trans++; return cdouble(::sin(a.v)); }
inline cdouble cos(cdouble a) { g_ops.trans++; return cdouble(::cos(a.v)); }
inline cdouble log(cdouble a) { g_ops.trans++; return cdouble(::log(a.v)); }
inline cdouble exp(cdouble a) { g_ops.trans++; return cdouble(::exp(a.v)); }
inline cdouble pow(cdouble a, cdouble b) { g_ops.trans++; return cdouble(::pow(a.v, b.v)); }
inline cdouble pow(cdouble a, double b) { return pow(a, cdouble(b)); }
inline cdouble nextafter(cdouble a, cdouble b) { return cdouble(::nextafter(a.v, b.v)); }
inline cdouble nextafter(double a, cdouble b) { return cdouble(::nextafter(a, b.v)); }
inline bool isfinite(cdouble a) { return ::isfinite(a.v); }
inline long lround(cdouble a) { return ::lround(a.v); }
