// ORACLE — TEST INFRASTRUCTURE ONLY (never shipped, never on the product path).
//
// Dual numbers in the style of ceres::Jet<double, K>.  Ceres itself is a third-party
// dependency of the reference that is NOT present under /root/reference
// (CMakeLists.txt:68 `find_package(Ceres REQUIRED)`, version unpinned, API use implies
// Ceres <= 2.1).  This header restates the published forward-mode rules used by
// ceres/jet.h: value `a`, infinitesimal part `v[K]`, comparisons act on `a` only.
#pragma once
#include <cmath>

namespace oracle {

template <int K>
struct Jet {
  double a;
  double v[K];
  Jet() : a(0.0) { for (int i = 0; i < K; ++i) v[i] = 0.0; }
  Jet(double s) : a(s) { for (int i = 0; i < K; ++i) v[i] = 0.0; }  // NOLINT
  Jet(double s, int k) : a(s) { for (int i = 0; i < K; ++i) v[i] = 0.0; v[k] = 1.0; }
};

#define ORACLE_JET_BIN(op, expr_a, expr_v)                                      \
  template <int K> inline Jet<K> operator op(const Jet<K>& f, const Jet<K>& g) { \
    Jet<K> h; h.a = expr_a; for (int i = 0; i < K; ++i) h.v[i] = expr_v; return h; }

ORACLE_JET_BIN(+, f.a + g.a, f.v[i] + g.v[i])
ORACLE_JET_BIN(-, f.a - g.a, f.v[i] - g.v[i])
ORACLE_JET_BIN(*, f.a * g.a, f.a * g.v[i] + f.v[i] * g.a)
#undef ORACLE_JET_BIN

template <int K> inline Jet<K> operator/(const Jet<K>& f, const Jet<K>& g) {
  // ceres/jet.h: h = f/g ; dh = (df - h dg)/g
  Jet<K> h; const double ginv = 1.0 / g.a; h.a = f.a * ginv;
  for (int i = 0; i < K; ++i) h.v[i] = (f.v[i] - h.a * g.v[i]) * ginv;
  return h;
}
template <int K> inline Jet<K> operator-(const Jet<K>& f) {
  Jet<K> h; h.a = -f.a; for (int i = 0; i < K; ++i) h.v[i] = -f.v[i]; return h;
}
template <int K> inline Jet<K> operator+(const Jet<K>& f, double s) { Jet<K> h = f; h.a += s; return h; }
template <int K> inline Jet<K> operator+(double s, const Jet<K>& f) { return f + s; }
template <int K> inline Jet<K> operator-(const Jet<K>& f, double s) { Jet<K> h = f; h.a -= s; return h; }
template <int K> inline Jet<K> operator-(double s, const Jet<K>& f) { Jet<K> h = -f; h.a += s; return h; }
template <int K> inline Jet<K> operator*(const Jet<K>& f, double s) {
  Jet<K> h; h.a = f.a * s; for (int i = 0; i < K; ++i) h.v[i] = f.v[i] * s; return h;
}
template <int K> inline Jet<K> operator*(double s, const Jet<K>& f) { return f * s; }
template <int K> inline Jet<K> operator/(const Jet<K>& f, double s) { return f * (1.0 / s); }
template <int K> inline Jet<K> operator/(double s, const Jet<K>& g) { return Jet<K>(s) / g; }
template <int K> inline Jet<K>& operator+=(Jet<K>& f, const Jet<K>& g) { f = f + g; return f; }
template <int K> inline Jet<K>& operator-=(Jet<K>& f, const Jet<K>& g) { f = f - g; return f; }
template <int K> inline Jet<K>& operator*=(Jet<K>& f, const Jet<K>& g) { f = f * g; return f; }
template <int K> inline Jet<K>& operator/=(Jet<K>& f, const Jet<K>& g) { f = f / g; return f; }

#define ORACLE_JET_CMP(op)                                                                  \
  template <int K> inline bool operator op(const Jet<K>& f, const Jet<K>& g) { return f.a op g.a; } \
  template <int K> inline bool operator op(const Jet<K>& f, double g) { return f.a op g; }   \
  template <int K> inline bool operator op(double f, const Jet<K>& g) { return f op g.a; }
ORACLE_JET_CMP(<) ORACLE_JET_CMP(<=) ORACLE_JET_CMP(>) ORACLE_JET_CMP(>=) ORACLE_JET_CMP(==) ORACLE_JET_CMP(!=)
#undef ORACLE_JET_CMP

template <int K> inline Jet<K> chain(double fa, double dfa, const Jet<K>& x) {
  Jet<K> h; h.a = fa; for (int i = 0; i < K; ++i) h.v[i] = dfa * x.v[i]; return h;
}
template <int K> inline Jet<K> sqrt(const Jet<K>& f) { double s = std::sqrt(f.a); return chain(s, 1.0 / (2.0 * s), f); }
template <int K> inline Jet<K> sin(const Jet<K>& f) { return chain(std::sin(f.a), std::cos(f.a), f); }
template <int K> inline Jet<K> cos(const Jet<K>& f) { return chain(std::cos(f.a), -std::sin(f.a), f); }
template <int K> inline Jet<K> tan(const Jet<K>& f) { double t = std::tan(f.a); return chain(t, 1.0 + t * t, f); }
template <int K> inline Jet<K> atan(const Jet<K>& f) { return chain(std::atan(f.a), 1.0 / (1.0 + f.a * f.a), f); }
template <int K> inline Jet<K> abs(const Jet<K>& f) { return f.a < 0.0 ? -f : f; }

inline double sqrt(double x) { return std::sqrt(x); }
inline double sin(double x) { return std::sin(x); }
inline double cos(double x) { return std::cos(x); }
inline double tan(double x) { return std::tan(x); }
inline double atan(double x) { return std::atan(x); }
inline double abs(double x) { return std::fabs(x); }

inline double value_of(double x) { return x; }
template <int K> inline double value_of(const Jet<K>& x) { return x.a; }

}  // namespace oracle
