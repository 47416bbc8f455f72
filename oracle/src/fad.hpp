// ORACLE (test infrastructure, NOT product code).
// Fixed-width forward-mode AD scalar restating Sacado::Fad::SFad<double,N> as the
// reference uses it (typedefs AD2..AD32, AD: /root/reference/src/preferences.hpp:75-98).
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference legs may use oracle/.
#pragma once
#include <cmath>

namespace orc {

template <int N>
struct SFad {
  double v;
  double d[N];
  SFad() : v(0.0) { for (int i = 0; i < N; ++i) d[i] = 0.0; }
  SFad(double x) : v(x) { for (int i = 0; i < N; ++i) d[i] = 0.0; }
  // seeded constructor: SFad(size, lane, value)  (workset.cpp:842)
  SFad(int /*sz*/, int lane, double x) : v(x) { for (int i = 0; i < N; ++i) d[i] = 0.0; d[lane] = 1.0; }
  static constexpr int size() { return N; }
  double val() const { return v; }
  double& val() { return v; }
  double dx(int i) const { return d[i]; }
  double& fastAccessDx(int i) { return d[i]; }
  double fastAccessDx(int i) const { return d[i]; }

  SFad& operator+=(const SFad& o) { v += o.v; for (int i = 0; i < N; ++i) d[i] += o.d[i]; return *this; }
  SFad& operator-=(const SFad& o) { v -= o.v; for (int i = 0; i < N; ++i) d[i] -= o.d[i]; return *this; }
  SFad& operator*=(const SFad& o) {
    for (int i = 0; i < N; ++i) d[i] = d[i] * o.v + v * o.d[i];
    v *= o.v; return *this;
  }
  SFad& operator/=(const SFad& o) {
    for (int i = 0; i < N; ++i) d[i] = (d[i] * o.v - v * o.d[i]) / (o.v * o.v);
    v /= o.v; return *this;
  }
  SFad& operator+=(double o) { v += o; return *this; }
  SFad& operator-=(double o) { v -= o; return *this; }
  SFad& operator*=(double o) { v *= o; for (int i = 0; i < N; ++i) d[i] *= o; return *this; }
  SFad& operator/=(double o) { v /= o; for (int i = 0; i < N; ++i) d[i] /= o; return *this; }
};

template <int N> inline SFad<N> operator+(const SFad<N>& a, const SFad<N>& b) { SFad<N> r; r.v = a.v + b.v; for (int i = 0; i < N; ++i) r.d[i] = a.d[i] + b.d[i]; return r; }
template <int N> inline SFad<N> operator-(const SFad<N>& a, const SFad<N>& b) { SFad<N> r; r.v = a.v - b.v; for (int i = 0; i < N; ++i) r.d[i] = a.d[i] - b.d[i]; return r; }
template <int N> inline SFad<N> operator*(const SFad<N>& a, const SFad<N>& b) { SFad<N> r; r.v = a.v * b.v; for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * b.v + a.v * b.d[i]; return r; }
template <int N> inline SFad<N> operator/(const SFad<N>& a, const SFad<N>& b) { SFad<N> r; r.v = a.v / b.v; for (int i = 0; i < N; ++i) r.d[i] = (a.d[i] * b.v - a.v * b.d[i]) / (b.v * b.v); return r; }
template <int N> inline SFad<N> operator-(const SFad<N>& a) { SFad<N> r; r.v = -a.v; for (int i = 0; i < N; ++i) r.d[i] = -a.d[i]; return r; }

template <int N> inline SFad<N> operator+(const SFad<N>& a, double b) { SFad<N> r = a; r.v += b; return r; }
template <int N> inline SFad<N> operator+(double b, const SFad<N>& a) { SFad<N> r = a; r.v = b + a.v; return r; }
template <int N> inline SFad<N> operator-(const SFad<N>& a, double b) { SFad<N> r = a; r.v -= b; return r; }
template <int N> inline SFad<N> operator-(double b, const SFad<N>& a) { SFad<N> r; r.v = b - a.v; for (int i = 0; i < N; ++i) r.d[i] = -a.d[i]; return r; }
template <int N> inline SFad<N> operator*(const SFad<N>& a, double b) { SFad<N> r; r.v = a.v * b; for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * b; return r; }
template <int N> inline SFad<N> operator*(double b, const SFad<N>& a) { SFad<N> r; r.v = b * a.v; for (int i = 0; i < N; ++i) r.d[i] = b * a.d[i]; return r; }
template <int N> inline SFad<N> operator/(const SFad<N>& a, double b) { SFad<N> r; r.v = a.v / b; for (int i = 0; i < N; ++i) r.d[i] = a.d[i] / b; return r; }
template <int N> inline SFad<N> operator/(double b, const SFad<N>& a) { SFad<N> r; r.v = b / a.v; for (int i = 0; i < N; ++i) r.d[i] = -b * a.d[i] / (a.v * a.v); return r; }

template <int N> inline bool operator<(const SFad<N>& a, const SFad<N>& b) { return a.v < b.v; }
template <int N> inline bool operator>(const SFad<N>& a, const SFad<N>& b) { return a.v > b.v; }
template <int N> inline bool operator<=(const SFad<N>& a, const SFad<N>& b) { return a.v <= b.v; }
template <int N> inline bool operator>=(const SFad<N>& a, const SFad<N>& b) { return a.v >= b.v; }
template <int N> inline bool operator<(const SFad<N>& a, double b) { return a.v < b; }
template <int N> inline bool operator>(const SFad<N>& a, double b) { return a.v > b; }
template <int N> inline bool operator<=(const SFad<N>& a, double b) { return a.v <= b; }
template <int N> inline bool operator>=(const SFad<N>& a, double b) { return a.v >= b; }

#define ORC_UNARY(NAME, VAL, DER)                                              \
  template <int N> inline SFad<N> NAME(const SFad<N>& a) {                      \
    SFad<N> r; const double x = a.v; r.v = VAL; const double g = DER;           \
    for (int i = 0; i < N; ++i) r.d[i] = g * a.d[i]; return r; }
ORC_UNARY(sin, std::sin(x), std::cos(x))
ORC_UNARY(cos, std::cos(x), -std::sin(x))
ORC_UNARY(tan, std::tan(x), 1.0 + std::tan(x) * std::tan(x))
ORC_UNARY(exp, std::exp(x), std::exp(x))
ORC_UNARY(log, std::log(x), 1.0 / x)
ORC_UNARY(sqrt, std::sqrt(x), 0.5 / std::sqrt(x))
ORC_UNARY(sinh, std::sinh(x), std::cosh(x))
ORC_UNARY(cosh, std::cosh(x), std::sinh(x))
#undef ORC_UNARY

// pow(a,b), both possibly active.  d/da = b a^(b-1), d/db = a^b log(a) (only used when b is active).
template <int N> inline SFad<N> pow(const SFad<N>& a, const SFad<N>& b) {
  SFad<N> r; r.v = std::pow(a.v, b.v);
  const double ga = (b.v == 0.0) ? 0.0 : b.v * std::pow(a.v, b.v - 1.0);
  bool bactive = false; for (int i = 0; i < N; ++i) if (b.d[i] != 0.0) bactive = true;
  const double gb = bactive ? r.v * std::log(a.v) : 0.0;
  for (int i = 0; i < N; ++i) r.d[i] = ga * a.d[i] + (bactive ? gb * b.d[i] : 0.0);
  return r;
}
template <int N> inline SFad<N> pow(const SFad<N>& a, double b) {
  SFad<N> r; r.v = std::pow(a.v, b);
  const double ga = (b == 0.0) ? 0.0 : b * std::pow(a.v, b - 1.0);
  for (int i = 0; i < N; ++i) r.d[i] = ga * a.d[i];
  return r;
}

// helpers usable for both double and SFad
inline double value_of(double x) { return x; }
template <int N> inline double value_of(const SFad<N>& x) { return x.v; }

using std::sin; using std::cos; using std::tan; using std::exp; using std::log;
using std::sqrt; using std::sinh; using std::cosh; using std::pow;

template <class T> struct ad_traits { static constexpr int lanes = 0; };
template <int N> struct ad_traits<SFad<N>> { static constexpr int lanes = N; };

}  // namespace orc
