// TEST INFRASTRUCTURE ONLY -- never linked into the product.
// Vector helpers of the restated cudAD (see forward.h for provenance): component access,
// packing, seeding and broadcasting for fvar<float3|float4|double3|double4, D>.
#pragma once

#include "forward.h"

namespace cudAD {

// ---- plain component access (so generic code can call getX on either kind) ----------------
#define CUDAD_PLAIN_GET(V3, V4, S)                                         \
  __host__ __device__ inline S getX(const V3& v) { return v.x; }           \
  __host__ __device__ inline S getY(const V3& v) { return v.y; }           \
  __host__ __device__ inline S getZ(const V3& v) { return v.z; }           \
  __host__ __device__ inline S getX(const V4& v) { return v.x; }           \
  __host__ __device__ inline S getY(const V4& v) { return v.y; }           \
  __host__ __device__ inline S getZ(const V4& v) { return v.z; }           \
  __host__ __device__ inline S getW(const V4& v) { return v.w; }
CUDAD_PLAIN_GET(float3, float4, float)
CUDAD_PLAIN_GET(double3, double4, double)
#undef CUDAD_PLAIN_GET

template <typename V> struct scalar_of;
template <> struct scalar_of<float3> { typedef float type; };
template <> struct scalar_of<float4> { typedef float type; };
template <> struct scalar_of<double3> { typedef double type; };
template <> struct scalar_of<double4> { typedef double type; };

#define CUDAD_FVAR_GET(NAME, MEMBER)                                               \
  template <typename V, int D>                                                     \
  __host__ __device__ inline fvar<typename scalar_of<V>::type, D> NAME(            \
      const fvar<V, D>& v) {                                                       \
    fvar<typename scalar_of<V>::type, D> r(v.value().MEMBER);                      \
    for (int i = 0; i < D; ++i) r.derivative(i) = v.derivative(i).MEMBER;          \
    return r;                                                                      \
  }
CUDAD_FVAR_GET(getX, x)
CUDAD_FVAR_GET(getY, y)
CUDAD_FVAR_GET(getZ, z)
CUDAD_FVAR_GET(getW, w)
#undef CUDAD_FVAR_GET

// ---- packing --------------------------------------------------------------------------------
template <int D>
__host__ __device__ inline fvar<float3, D> make_float3(const fvar<float, D>& x,
                                                        const fvar<float, D>& y,
                                                        const fvar<float, D>& z) {
  fvar<float3, D> r(::make_float3(x.value(), y.value(), z.value()));
  for (int i = 0; i < D; ++i)
    r.derivative(i) = ::make_float3(x.derivative(i), y.derivative(i), z.derivative(i));
  return r;
}
template <int D>
__host__ __device__ inline fvar<float4, D> make_float4(const fvar<float, D>& x,
                                                        const fvar<float, D>& y,
                                                        const fvar<float, D>& z,
                                                        const fvar<float, D>& w) {
  fvar<float4, D> r(::make_float4(x.value(), y.value(), z.value(), w.value()));
  for (int i = 0; i < D; ++i)
    r.derivative(i) =
        ::make_float4(x.derivative(i), y.derivative(i), z.derivative(i), w.derivative(i));
  return r;
}
template <int D>
__host__ __device__ inline fvar<double3, D> make_double3(const fvar<double, D>& x,
                                                          const fvar<double, D>& y,
                                                          const fvar<double, D>& z) {
  fvar<double3, D> r(::make_double3(x.value(), y.value(), z.value()));
  for (int i = 0; i < D; ++i)
    r.derivative(i) = ::make_double3(x.derivative(i), y.derivative(i), z.derivative(i));
  return r;
}
template <int D>
__host__ __device__ inline fvar<double4, D> make_double4(const fvar<double, D>& x,
                                                          const fvar<double, D>& y,
                                                          const fvar<double, D>& z,
                                                          const fvar<double, D>& w) {
  fvar<double4, D> r(::make_double4(x.value(), y.value(), z.value(), w.value()));
  for (int i = 0; i < D; ++i)
    r.derivative(i) =
        ::make_double4(x.derivative(i), y.derivative(i), z.derivative(i), w.derivative(i));
  return r;
}

// ---- seeding: component c gets the unit tangent of slot idx.c (if idx.c is a valid slot) ----
template <int D>
__host__ __device__ inline fvar<float3, D> make_float3in(const float3& v, const int3& idx) {
  return make_float3(fvar<float, D>::input(v.x, idx.x), fvar<float, D>::input(v.y, idx.y),
                     fvar<float, D>::input(v.z, idx.z));
}
template <int D>
__host__ __device__ inline fvar<float4, D> make_float4in(const float4& v, const int4& idx) {
  return make_float4(fvar<float, D>::input(v.x, idx.x), fvar<float, D>::input(v.y, idx.y),
                     fvar<float, D>::input(v.z, idx.z), fvar<float, D>::input(v.w, idx.w));
}
template <int D>
__host__ __device__ inline fvar<double3, D> make_double3in(const double3& v, const int3& idx) {
  return make_double3(fvar<double, D>::input(v.x, idx.x), fvar<double, D>::input(v.y, idx.y),
                      fvar<double, D>::input(v.z, idx.z));
}
template <int D>
__host__ __device__ inline fvar<double4, D> make_double4in(const double4& v, const int4& idx) {
  return make_double4(fvar<double, D>::input(v.x, idx.x), fvar<double, D>::input(v.y, idx.y),
                      fvar<double, D>::input(v.z, idx.z), fvar<double, D>::input(v.w, idx.w));
}

// ---- broadcasting a scalar to all components -------------------------------------------------
__host__ __device__ inline float3 broadcast3(float s) { return ::make_float3(s, s, s); }
__host__ __device__ inline float4 broadcast4(float s) { return ::make_float4(s, s, s, s); }
__host__ __device__ inline double3 broadcast3(double s) { return ::make_double3(s, s, s); }
__host__ __device__ inline double4 broadcast4(double s) { return ::make_double4(s, s, s, s); }
template <int D>
__host__ __device__ inline fvar<float3, D> broadcast3(const fvar<float, D>& s) {
  return make_float3(s, s, s);
}
template <int D>
__host__ __device__ inline fvar<float4, D> broadcast4(const fvar<float, D>& s) {
  return make_float4(s, s, s, s);
}
template <int D>
__host__ __device__ inline fvar<double3, D> broadcast3(const fvar<double, D>& s) {
  return make_double3(s, s, s);
}
template <int D>
__host__ __device__ inline fvar<double4, D> broadcast4(const fvar<double, D>& s) {
  return make_double4(s, s, s, s);
}

// ---- reductions --------------------------------------------------------------------------------
template <typename V, int D>
__host__ __device__ inline fvar<typename scalar_of<V>::type, D> dot(const fvar<V, D>& a,
                                                                    const fvar<V, D>& b) {
  fvar<typename scalar_of<V>::type, D> r(::dot(a.value(), b.value()));
  for (int i = 0; i < D; ++i)
    r.derivative(i) = ::dot(a.derivative(i), b.value()) + ::dot(a.value(), b.derivative(i));
  return r;
}
// |v| with d|v| = <v, dv> / |v|
template <typename V, int D>
__host__ __device__ inline fvar<typename scalar_of<V>::type, D> length(const fvar<V, D>& a) {
  typedef typename scalar_of<V>::type S;
  const S l = ::length(a.value());
  fvar<S, D> r(l);
  for (int i = 0; i < D; ++i) r.derivative(i) = ::dot(a.value(), a.derivative(i)) / l;
  return r;
}

}  // namespace cudAD
