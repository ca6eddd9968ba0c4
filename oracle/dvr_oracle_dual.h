// TEST INFRASTRUCTURE ONLY -- never imported, linked or executed by the product.
//
// Dual number with a run-time number of tangent slots, used by the CPU restatement
// (oracle/dvr_oracle.cpp) for the forward-mode path. It restates the published forward-mode
// AD rules that the un-vendored dependency `cudad` @ f00bc425e8a5abd84a0690a69f2b36535454849c
// (reference .SUBMODULES.json:16-22) provides to the reference; see
// oracle/cudad_restated/forward.h for the call sites and the conventions chosen.
#pragma once
#include <cmath>
#include <vector>

namespace dvro {

template <typename R>
struct Dual {
  R v;
  std::vector<R> d;
  Dual() : v(0) {}
  Dual(R value, int D) : v(value), d(size_t(D), R(0)) {}
  static Dual input(R value, int idx, int D) {
    Dual r(value, D);
    if (idx >= 0 && idx < D) r.d[size_t(idx)] = R(1);
    return r;
  }
  int D() const { return int(d.size()); }
};

template <typename R>
inline Dual<R> operator+(const Dual<R>& a, const Dual<R>& b) {
  Dual<R> r(a.v + b.v, a.D());
  for (int i = 0; i < a.D(); ++i) r.d[i] = a.d[i] + b.d[i];
  return r;
}
template <typename R>
inline Dual<R> operator+(const Dual<R>& a, R b) {
  Dual<R> r = a;
  r.v = a.v + b;
  return r;
}
template <typename R>
inline Dual<R> operator+(R a, const Dual<R>& b) {
  Dual<R> r = b;
  r.v = a + b.v;
  return r;
}
template <typename R>
inline Dual<R> operator-(const Dual<R>& a, const Dual<R>& b) {
  Dual<R> r(a.v - b.v, a.D());
  for (int i = 0; i < a.D(); ++i) r.d[i] = a.d[i] - b.d[i];
  return r;
}
template <typename R>
inline Dual<R> operator-(const Dual<R>& a, R b) {
  Dual<R> r = a;
  r.v = a.v - b;
  return r;
}
template <typename R>
inline Dual<R> operator-(R a, const Dual<R>& b) {
  Dual<R> r(a - b.v, b.D());
  for (int i = 0; i < b.D(); ++i) r.d[i] = -b.d[i];
  return r;
}
template <typename R>
inline Dual<R> operator-(const Dual<R>& a) {
  Dual<R> r(-a.v, a.D());
  for (int i = 0; i < a.D(); ++i) r.d[i] = -a.d[i];
  return r;
}
template <typename R>
inline Dual<R> operator*(const Dual<R>& a, const Dual<R>& b) {
  Dual<R> r(a.v * b.v, a.D());
  for (int i = 0; i < a.D(); ++i) r.d[i] = a.d[i] * b.v + a.v * b.d[i];
  return r;
}
template <typename R>
inline Dual<R> operator*(const Dual<R>& a, R b) {
  Dual<R> r(a.v * b, a.D());
  for (int i = 0; i < a.D(); ++i) r.d[i] = a.d[i] * b;
  return r;
}
template <typename R>
inline Dual<R> operator*(R a, const Dual<R>& b) {
  Dual<R> r(a * b.v, b.D());
  for (int i = 0; i < b.D(); ++i) r.d[i] = a * b.d[i];
  return r;
}
template <typename R>
inline Dual<R> operator/(const Dual<R>& a, const Dual<R>& b) {
  Dual<R> r(a.v / b.v, a.D());
  const R bb = b.v * b.v;
  for (int i = 0; i < a.D(); ++i) r.d[i] = (a.d[i] * b.v - a.v * b.d[i]) / bb;
  return r;
}
template <typename R>
inline Dual<R> operator/(const Dual<R>& a, R b) {
  Dual<R> r(a.v / b, a.D());
  for (int i = 0; i < a.D(); ++i) r.d[i] = a.d[i] / b;
  return r;
}
template <typename R>
inline Dual<R> dexp(const Dual<R>& a) {
  const R e = std::exp(a.v);
  Dual<R> r(e, a.D());
  for (int i = 0; i < a.D(); ++i) r.d[i] = a.d[i] * e;
  return r;
}
// selection rules: tangent of the selected argument (a<b ? a : b / a>b ? a : b)
template <typename R>
inline Dual<R> dmin(const Dual<R>& a, const Dual<R>& b) { return a.v < b.v ? a : b; }
template <typename R>
inline Dual<R> dmax(const Dual<R>& a, const Dual<R>& b) { return a.v > b.v ? a : b; }
template <typename R>
inline Dual<R> dclamp(const Dual<R>& x, const Dual<R>& lo, const Dual<R>& hi) {
  return dmax(lo, dmin(x, hi));
}
template <typename R>
inline Dual<R> dlerp(const Dual<R>& a, const Dual<R>& b, const Dual<R>& t) {
  return a + t * (b - a);
}

}  // namespace dvro
