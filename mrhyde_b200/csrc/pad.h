// Point-level forward-mode derivative lanes ("explicit derivative lanes held in registers").
//
// The reference differentiates the element residual with Sacado::Fad::SFad<double,N>, N = element DOFs
// (/root/reference/src/preferences.hpp:75-98), i.e. every weak-form operation drags N partials through
// global memory.  Here the chain rule is cut at the quadrature point: a physics routine sees the point
// state {u_m, du_m/dx_c, du_m/dt} and produces F_(v,a); PAD<P> carries the P = nvar*(1+dim) partials of F
// with respect to that point state.  The element Jacobian is then the dense contraction
//   K_(v,i),(m,j) = sum_q sum_a sum_b T_ia(q) dF_(v,a)/d(m,b) T_jb(q)            (assemble.h, phase 2).
#pragma once
#if defined(__CUDACC_RTC__)
// NVRTC build of the specialised kernels (jit.hpp): no host headers; the fixed-width types are spelled out
typedef unsigned char uint8_t; typedef unsigned short uint16_t; typedef int int32_t; typedef unsigned int uint32_t;
typedef long long int64_t; typedef unsigned long long uint64_t; typedef unsigned long size_t;
#else
#include <cmath>
#endif

#if defined(__CUDACC__)
#define MRH_HD __host__ __device__ __forceinline__
#define MRH_HDN __host__ __device__
#else
#define MRH_HD inline
#define MRH_HDN
#endif

#if !defined(__CUDACC__)
struct int4 { int x, y, z, w; };
struct uint4 { unsigned x, y, z, w; };
#endif

namespace mrh {

template <int P>
struct PAD {
  double v;
  double d[P];
  MRH_HD PAD() {}
  MRH_HD PAD(double x) : v(x) {
#pragma unroll
    for (int i = 0; i < P; ++i) d[i] = 0.0;
  }
  MRH_HD static PAD seed(double x, int lane, double dv) {
    PAD r(x);
#pragma unroll
    for (int i = 0; i < P; ++i) if (i == lane) r.d[i] = dv;
    return r;
  }
};

#define MRH_LANES for (int i = 0; i < P; ++i)

template <int P> MRH_HD PAD<P> operator+(const PAD<P>& a, const PAD<P>& b) { PAD<P> r; r.v = a.v + b.v;
#pragma unroll
  MRH_LANES r.d[i] = a.d[i] + b.d[i]; return r; }
template <int P> MRH_HD PAD<P> operator-(const PAD<P>& a, const PAD<P>& b) { PAD<P> r; r.v = a.v - b.v;
#pragma unroll
  MRH_LANES r.d[i] = a.d[i] - b.d[i]; return r; }
template <int P> MRH_HD PAD<P> operator*(const PAD<P>& a, const PAD<P>& b) { PAD<P> r; r.v = a.v * b.v;
#pragma unroll
  MRH_LANES r.d[i] = a.d[i] * b.v + a.v * b.d[i]; return r; }
template <int P> MRH_HD PAD<P> operator/(const PAD<P>& a, const PAD<P>& b) { PAD<P> r; const double ib = 1.0 / b.v; r.v = a.v * ib;
#pragma unroll
  MRH_LANES r.d[i] = (a.d[i] - r.v * b.d[i]) * ib; return r; }
template <int P> MRH_HD PAD<P> operator-(const PAD<P>& a) { PAD<P> r; r.v = -a.v;
#pragma unroll
  MRH_LANES r.d[i] = -a.d[i]; return r; }

template <int P> MRH_HD PAD<P> operator+(const PAD<P>& a, double b) { PAD<P> r = a; r.v = a.v + b; return r; }
template <int P> MRH_HD PAD<P> operator+(double b, const PAD<P>& a) { PAD<P> r = a; r.v = b + a.v; return r; }
template <int P> MRH_HD PAD<P> operator-(const PAD<P>& a, double b) { PAD<P> r = a; r.v = a.v - b; return r; }
template <int P> MRH_HD PAD<P> operator-(double b, const PAD<P>& a) { PAD<P> r; r.v = b - a.v;
#pragma unroll
  MRH_LANES r.d[i] = -a.d[i]; return r; }
template <int P> MRH_HD PAD<P> operator*(const PAD<P>& a, double b) { PAD<P> r; r.v = a.v * b;
#pragma unroll
  MRH_LANES r.d[i] = a.d[i] * b; return r; }
template <int P> MRH_HD PAD<P> operator*(double b, const PAD<P>& a) { return a * b; }
template <int P> MRH_HD PAD<P> operator/(const PAD<P>& a, double b) { const double ib = 1.0 / b; PAD<P> r; r.v = a.v / b;
#pragma unroll
  MRH_LANES r.d[i] = a.d[i] * ib; return r; }
template <int P> MRH_HD PAD<P> operator/(double b, const PAD<P>& a) { PAD<P> r; r.v = b / a.v; const double g = -r.v / a.v;
#pragma unroll
  MRH_LANES r.d[i] = g * a.d[i]; return r; }
template <int P> MRH_HD PAD<P>& operator+=(PAD<P>& a, const PAD<P>& b) { a = a + b; return a; }
template <int P> MRH_HD PAD<P>& operator-=(PAD<P>& a, const PAD<P>& b) { a = a - b; return a; }
template <int P> MRH_HD PAD<P>& operator*=(PAD<P>& a, const PAD<P>& b) { a = a * b; return a; }
template <int P> MRH_HD PAD<P>& operator+=(PAD<P>& a, double b) { a.v += b; return a; }
template <int P> MRH_HD PAD<P>& operator*=(PAD<P>& a, double b) { a = a * b; return a; }

MRH_HD double val(double x) { return x; }
template <int P> MRH_HD double val(const PAD<P>& x) { return x.v; }

// chain rule helper: r = f(a) with f'(a) = g
template <int P> MRH_HD PAD<P> chain(const PAD<P>& a, double f, double g) { PAD<P> r; r.v = f;
#pragma unroll
  MRH_LANES r.d[i] = g * a.d[i]; return r; }

#if defined(__CUDACC__)
// sin(x) (shift 0) or cos(x) (shift 1) for |x| < 105615: three-term Cody-Waite reduction by pi/2 with FMAs and the
// fdlibm kernel polynomials (error below 1 ulp on the reduced range); straight-line code with immediate
// coefficients, unlike the library routine whose table loads and slow-path branch serialise neighbouring calls.
// `ok` is cleared for arguments outside the range (and NaN); the caller then re-evaluates with sin() / cos().
__device__ __forceinline__ double trig_fast(double x, int shift, bool& ok) {
  ok = ok && (fabs(x) < 105615.0);
  const double t = fma(x, 6.3661977236758138e-01, 6755399441055744.0);
  const int k = __double2loint(t) + shift;
  const double q = t - 6755399441055744.0;
  double r = fma(-q, 1.5707963267948966e+00, x);
  r = fma(-q, 6.1232339957367574e-17, r);
  r = fma(-q, 8.4784276603688985e-32, r);
  const double z = r * r;
  const bool c = (k & 1) != 0;
  double p = c ? -1.13596475577881948265e-11 : 1.58969099521155010221e-10;
  p = fma(p, z, c ? 2.08757232129817482790e-09 : -2.50507602534068634195e-08);
  p = fma(p, z, c ? -2.75573143513906633035e-07 : 2.75573137070700676789e-06);
  p = fma(p, z, c ? 2.48015872894767294178e-05 : -1.98412698298579493134e-04);
  p = fma(p, z, c ? -1.38888888888741095749e-03 : 8.33333333332248946124e-03);
  p = fma(p, z, c ? 4.16666666666666019037e-02 : -1.66666666666666324348e-01);
  const double base = c ? fma(-0.5, z, 1.0) : r;
  const double m = c ? z * z : r * z;
  const double v = fma(m, p, base);
  return (k & 2) ? -v : v;
}
#endif
MRH_HD double f_sin(double x) { return sin(x); }
MRH_HD double f_cos(double x) { return cos(x); }
MRH_HD double f_tan(double x) { return tan(x); }
MRH_HD double f_exp(double x) { return exp(x); }
MRH_HD double f_log(double x) { return log(x); }
MRH_HD double f_sqrt(double x) { return sqrt(x); }
MRH_HD double f_sinh(double x) { return sinh(x); }
MRH_HD double f_cosh(double x) { return cosh(x); }
MRH_HD double f_pow(double a, double b) { return pow(a, b); }
template <int P> MRH_HD PAD<P> f_sin(const PAD<P>& a) { return chain(a, sin(a.v), cos(a.v)); }
template <int P> MRH_HD PAD<P> f_cos(const PAD<P>& a) { return chain(a, cos(a.v), -sin(a.v)); }
template <int P> MRH_HD PAD<P> f_tan(const PAD<P>& a) { const double t = tan(a.v); return chain(a, t, 1.0 + t * t); }
template <int P> MRH_HD PAD<P> f_exp(const PAD<P>& a) { const double e = exp(a.v); return chain(a, e, e); }
template <int P> MRH_HD PAD<P> f_log(const PAD<P>& a) { return chain(a, log(a.v), 1.0 / a.v); }
template <int P> MRH_HD PAD<P> f_sqrt(const PAD<P>& a) { const double s = sqrt(a.v); return chain(a, s, 0.5 / s); }
// "sqrt" of the coefficient language: 0 with zero derivatives for arguments <= 0
// (/root/reference/src/managers/functionManager.cpp:1062-1067, 1258-1262, 1573-1581)
MRH_HD double f_sqrt0(double x) { return x <= 0.0 ? 0.0 : sqrt(x); }
template <int P> MRH_HD PAD<P> f_sqrt0(const PAD<P>& a) { if (a.v <= 0.0) return PAD<P>(0.0); return f_sqrt(a); }
template <int P> MRH_HD PAD<P> f_sinh(const PAD<P>& a) { return chain(a, sinh(a.v), cosh(a.v)); }
template <int P> MRH_HD PAD<P> f_cosh(const PAD<P>& a) { return chain(a, cosh(a.v), sinh(a.v)); }
// pow with both arguments possibly active (d/db only when b carries derivatives, like Sacado's pow)
template <int P> MRH_HD PAD<P> f_pow(const PAD<P>& a, const PAD<P>& b) {
  PAD<P> r; r.v = pow(a.v, b.v);
  const double ga = (b.v == 0.0) ? 0.0 : b.v * pow(a.v, b.v - 1.0);
  bool bactive = false;
#pragma unroll
  MRH_LANES if (b.d[i] != 0.0) bactive = true;
  const double gb = bactive ? r.v * log(a.v) : 0.0;
#pragma unroll
  MRH_LANES r.d[i] = ga * a.d[i] + (bactive ? gb * b.d[i] : 0.0);
  return r;
}
#undef MRH_LANES

template <class S> struct lanes_of { static constexpr int value = 0; };
template <int P> struct lanes_of<PAD<P>> { static constexpr int value = P; };
MRH_HD double lane(double, int) { return 0.0; }
template <int P> MRH_HD double lane(const PAD<P>& x, int i) { return x.d[i]; }

}  // namespace mrh
