// TEST INFRASTRUCTURE ONLY -- never linked into the product.
//
// Restatement of the forward-mode dual-number type the reference's forward-differentiation
// path is written against: module `cudad` (https://gitlab.com/shaman42/cudad.git) pinned at
// f00bc425e8a5abd84a0690a69f2b36535454849c (reference .SUBMODULES.json:16-22). Its sources
// are NOT part of /root/reference (un-vendored submodule), so this header restates the
// published forward-mode algorithm (value + D tangent slots, sum/product/quotient/chain
// rules) behind exactly the API the reference call sites use:
//   renderer/renderer_kernels.cuh:151-157,180-185,211,222,236-246
//   renderer/renderer_camera.cuh:82,128-129
//   renderer/renderer_blending.cuh:40-51,105-116
//   renderer/renderer_interpolation.cuh:52-65,145-177,210-240
//   renderer/renderer_tf.cuh:121-154,255-288,436-485,521-560
//   unittests/testNonConvexDensity.cpp:23-42
// With it the reference's own DvrKernelForwardGradientsHost compiles unchanged
// (oracle/ref_shim_fwdgrad.cpp). Parity of this restatement is pinned by
// unittests/NonConvexDensity.tsv (tests/test_oracle_golden.py) and by forward==adjoint
// consistency against the reference's in-tree adjoint code.
//
// Conventions chosen where cudAD's behaviour cannot be inferred from the call sites:
//   fmin(a,b) = a<b ? a : b, fmax(a,b) = a>b ? a : b (tangent of the selected argument),
//   clamp(x,lo,hi) = fmax(lo, fmin(x,hi)) -- same selection rule as helper_math.cuh:1898-1901.
#pragma once

#include <cmath>
#include <type_traits>
#include <cuda_runtime.h>
#include "helper_math.cuh"

namespace cudAD {

namespace internal {
template <typename T, int D>
struct array {
  T data[D];
  __host__ __device__ T& operator[](int i) { return data[i]; }
  __host__ __device__ const T& operator[](int i) const { return data[i]; }
};
}  // namespace internal

template <typename T, int D>
class fvar {
 public:
  typedef T value_type;
  static constexpr int NumDerivatives = D;

  __host__ __device__ fvar() : v_{}, d_{} {}
  // implicit promotion of a plain value to a constant (zero tangents)
  __host__ __device__ fvar(const T& v) : v_(v), d_{} {}
  // value + all tangents (call site: renderer_compareToImage.cuh:19-22)
  __host__ __device__ fvar(const T& v, const internal::array<T, D>& d) : v_(v), d_{} {
    for (int i = 0; i < D; ++i) d_[i] = d[i];
  }

  static __host__ __device__ fvar constant(const T& v) { return fvar(v); }
  // seeds the unit tangent e_idx if idx is a valid slot, otherwise a constant
  static __host__ __device__ fvar input(const T& v, int idx) {
    fvar r(v);
    if (idx >= 0 && idx < D) r.d_[idx] = T(1);
    return r;
  }
  template <int I>
  static __host__ __device__ fvar input(const T& v) {
    static_assert(I >= 0 && I < D, "derivative slot out of range");
    fvar r(v);
    r.d_[I] = T(1);
    return r;
  }

  __host__ __device__ const T& value() const { return v_; }
  __host__ __device__ T& value() { return v_; }
  __host__ __device__ const T& derivative(int i) const { return d_[i]; }
  __host__ __device__ T& derivative(int i) { return d_[i]; }
  template <int I>
  __host__ __device__ const T& derivative() const { return d_[I]; }

  __host__ __device__ explicit operator T() const { return v_; }

 private:
  T v_;
  T d_[D];
};

// ---- traits ---------------------------------------------------------------------------
template <typename X> struct is_fvar : std::false_type {};
template <typename T, int D> struct is_fvar<fvar<T, D>> : std::true_type {};
template <typename X> struct fvar_dim { static constexpr int value = 0; };
template <typename T, int D> struct fvar_dim<fvar<T, D>> { static constexpr int value = D; };

template <typename T, int D>
__host__ __device__ inline const T& val(const fvar<T, D>& a) { return a.value(); }
template <typename P, typename = typename std::enable_if<!is_fvar<P>::value>::type>
__host__ __device__ inline const P& val(const P& a) { return a; }

#define CUDAD_ENABLE_PLAIN(P) typename = typename std::enable_if<!is_fvar<P>::value>::type

// ---- addition / subtraction -------------------------------------------------------------
template <typename A, typename B, int D>
__host__ __device__ inline auto operator+(const fvar<A, D>& a, const fvar<B, D>& b)
    -> fvar<decltype(A() + B()), D> {
  fvar<decltype(A() + B()), D> r(a.value() + b.value());
  for (int i = 0; i < D; ++i) r.derivative(i) = a.derivative(i) + b.derivative(i);
  return r;
}
template <typename A, typename P, int D, CUDAD_ENABLE_PLAIN(P)>
__host__ __device__ inline auto operator+(const fvar<A, D>& a, const P& b)
    -> fvar<decltype(A() + P()), D> {
  fvar<decltype(A() + P()), D> r(a.value() + b);
  for (int i = 0; i < D; ++i) r.derivative(i) = a.derivative(i);
  return r;
}
template <typename P, typename B, int D, CUDAD_ENABLE_PLAIN(P)>
__host__ __device__ inline auto operator+(const P& a, const fvar<B, D>& b)
    -> fvar<decltype(P() + B()), D> {
  fvar<decltype(P() + B()), D> r(a + b.value());
  for (int i = 0; i < D; ++i) r.derivative(i) = b.derivative(i);
  return r;
}
template <typename A, typename B, int D>
__host__ __device__ inline auto operator-(const fvar<A, D>& a, const fvar<B, D>& b)
    -> fvar<decltype(A() - B()), D> {
  fvar<decltype(A() - B()), D> r(a.value() - b.value());
  for (int i = 0; i < D; ++i) r.derivative(i) = a.derivative(i) - b.derivative(i);
  return r;
}
template <typename A, typename P, int D, CUDAD_ENABLE_PLAIN(P)>
__host__ __device__ inline auto operator-(const fvar<A, D>& a, const P& b)
    -> fvar<decltype(A() - P()), D> {
  fvar<decltype(A() - P()), D> r(a.value() - b);
  for (int i = 0; i < D; ++i) r.derivative(i) = a.derivative(i);
  return r;
}
template <typename P, typename B, int D, CUDAD_ENABLE_PLAIN(P)>
__host__ __device__ inline auto operator-(const P& a, const fvar<B, D>& b)
    -> fvar<decltype(P() - B()), D> {
  fvar<decltype(P() - B()), D> r(a - b.value());
  for (int i = 0; i < D; ++i) r.derivative(i) = -b.derivative(i);
  return r;
}
template <typename A, int D>
__host__ __device__ inline fvar<A, D> operator-(const fvar<A, D>& a) {
  fvar<A, D> r(-a.value());
  for (int i = 0; i < D; ++i) r.derivative(i) = -a.derivative(i);
  return r;
}

// ---- product rule -----------------------------------------------------------------------
template <typename A, typename B, int D>
__host__ __device__ inline auto operator*(const fvar<A, D>& a, const fvar<B, D>& b)
    -> fvar<decltype(A() * B()), D> {
  fvar<decltype(A() * B()), D> r(a.value() * b.value());
  for (int i = 0; i < D; ++i)
    r.derivative(i) = a.derivative(i) * b.value() + a.value() * b.derivative(i);
  return r;
}
template <typename A, typename P, int D, CUDAD_ENABLE_PLAIN(P)>
__host__ __device__ inline auto operator*(const fvar<A, D>& a, const P& b)
    -> fvar<decltype(A() * P()), D> {
  fvar<decltype(A() * P()), D> r(a.value() * b);
  for (int i = 0; i < D; ++i) r.derivative(i) = a.derivative(i) * b;
  return r;
}
template <typename P, typename B, int D, CUDAD_ENABLE_PLAIN(P)>
__host__ __device__ inline auto operator*(const P& a, const fvar<B, D>& b)
    -> fvar<decltype(P() * B()), D> {
  fvar<decltype(P() * B()), D> r(a * b.value());
  for (int i = 0; i < D; ++i) r.derivative(i) = a * b.derivative(i);
  return r;
}

// ---- quotient rule ----------------------------------------------------------------------
template <typename A, typename B, int D>
__host__ __device__ inline auto operator/(const fvar<A, D>& a, const fvar<B, D>& b)
    -> fvar<decltype(A() / B()), D> {
  fvar<decltype(A() / B()), D> r(a.value() / b.value());
  const B bb = b.value() * b.value();
  for (int i = 0; i < D; ++i)
    r.derivative(i) = (a.derivative(i) * b.value() - a.value() * b.derivative(i)) / bb;
  return r;
}
template <typename A, typename P, int D, CUDAD_ENABLE_PLAIN(P)>
__host__ __device__ inline auto operator/(const fvar<A, D>& a, const P& b)
    -> fvar<decltype(A() / P()), D> {
  fvar<decltype(A() / P()), D> r(a.value() / b);
  for (int i = 0; i < D; ++i) r.derivative(i) = a.derivative(i) / b;
  return r;
}
template <typename P, typename B, int D, CUDAD_ENABLE_PLAIN(P)>
__host__ __device__ inline auto operator/(const P& a, const fvar<B, D>& b)
    -> fvar<decltype(P() / B()), D> {
  fvar<decltype(P() / B()), D> r(a / b.value());
  const B bb = b.value() * b.value();
  for (int i = 0; i < D; ++i) r.derivative(i) = -(a * b.derivative(i)) / bb;
  return r;
}

// ---- compound assignment ----------------------------------------------------------------
template <typename A, typename X, int D>
__host__ __device__ inline fvar<A, D>& operator+=(fvar<A, D>& a, const X& b) {
  a = a + b;
  return a;
}
template <typename A, typename X, int D>
__host__ __device__ inline fvar<A, D>& operator-=(fvar<A, D>& a, const X& b) {
  a = a - b;
  return a;
}
template <typename A, typename X, int D>
__host__ __device__ inline fvar<A, D>& operator*=(fvar<A, D>& a, const X& b) {
  a = a * b;
  return a;
}

// ---- elementary functions (chain rule) ----------------------------------------------------
template <typename A, int D>
__host__ __device__ inline fvar<A, D> exp(const fvar<A, D>& a) {
  using std::exp;
  const A e = exp(a.value());
  fvar<A, D> r(e);
  for (int i = 0; i < D; ++i) r.derivative(i) = a.derivative(i) * e;
  return r;
}
template <typename A, int D>
__host__ __device__ inline fvar<A, D> log(const fvar<A, D>& a) {
  using std::log;
  fvar<A, D> r(log(a.value()));
  for (int i = 0; i < D; ++i) r.derivative(i) = a.derivative(i) / a.value();
  return r;
}
template <typename A, int D>
__host__ __device__ inline fvar<A, D> sqrt(const fvar<A, D>& a) {
  using std::sqrt;
  const A s = sqrt(a.value());
  fvar<A, D> r(s);
  for (int i = 0; i < D; ++i) r.derivative(i) = a.derivative(i) / (2 * s);
  return r;
}

// ---- selection functions ------------------------------------------------------------------
template <typename A, int D>
__host__ __device__ inline fvar<A, D> fmin(const fvar<A, D>& a, const fvar<A, D>& b) {
  return a.value() < b.value() ? a : b;
}
template <typename A, typename P, int D, CUDAD_ENABLE_PLAIN(P)>
__host__ __device__ inline fvar<A, D> fmin(const fvar<A, D>& a, const P& b) {
  return a.value() < A(b) ? a : fvar<A, D>(A(b));
}
template <typename P, typename A, int D, CUDAD_ENABLE_PLAIN(P)>
__host__ __device__ inline fvar<A, D> fmin(const P& a, const fvar<A, D>& b) {
  return A(a) < b.value() ? fvar<A, D>(A(a)) : b;
}
template <typename A, int D>
__host__ __device__ inline fvar<A, D> fmax(const fvar<A, D>& a, const fvar<A, D>& b) {
  return a.value() > b.value() ? a : b;
}
template <typename A, typename P, int D, CUDAD_ENABLE_PLAIN(P)>
__host__ __device__ inline fvar<A, D> fmax(const fvar<A, D>& a, const P& b) {
  return a.value() > A(b) ? a : fvar<A, D>(A(b));
}
template <typename P, typename A, int D, CUDAD_ENABLE_PLAIN(P)>
__host__ __device__ inline fvar<A, D> fmax(const P& a, const fvar<A, D>& b) {
  return A(a) > b.value() ? fvar<A, D>(A(a)) : b;
}
// clamp(x, lo, hi) = fmax(lo, fmin(x, hi)); any argument may be plain, at least one is an fvar
template <typename X, typename L, typename H,
          typename = typename std::enable_if<is_fvar<X>::value || is_fvar<L>::value ||
                                             is_fvar<H>::value>::type>
__host__ __device__ inline auto clamp(const X& x, const L& lo, const H& hi) {
  constexpr int D = fvar_dim<X>::value > 0
                        ? fvar_dim<X>::value
                        : (fvar_dim<L>::value > 0 ? fvar_dim<L>::value : fvar_dim<H>::value);
  typedef typename std::decay<decltype(val(x))>::type S;
  typedef fvar<S, D> R;
  const R xx(x), ll(lo), hh(hi);
  return fmax(ll, fmin(xx, hh));
}

// lerp(a, b, t) = a + t*(b-a), same operation order as helper_math.cuh:1835-1838
template <typename A, typename B, typename T,
          typename = typename std::enable_if<is_fvar<A>::value || is_fvar<B>::value ||
                                             is_fvar<T>::value>::type>
__host__ __device__ inline auto lerp(const A& a, const B& b, const T& t) {
  return a + t * (b - a);
}

#undef CUDAD_ENABLE_PLAIN

}  // namespace cudAD
