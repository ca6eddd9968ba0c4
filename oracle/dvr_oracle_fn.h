// TEST INFRASTRUCTURE ONLY -- never linked into the product.
//
// The port's device functions one call at a time, in the form the reference's unit tests drive the reference's
// (unittests/testAdjointBlending.cpp, testAdjointTF.cpp, testAdjointInterpolation.cpp): blend / blend_adjoint,
// tf_eval / tf_adjoint (with TF derivatives), trilinear fetch / adjoint (with volume derivatives) of dvr_oracle.cpp.
// Same signatures as oracle/ref_shim_fn.cpp (the reference's own functions) and tests/csrc/devfn_harness.cu (the
// product's CUDA device functions). Included by dvr_oracle.cpp. Pinned by tests/test_oracle_devfn.py: against the
// reference build on the unit tests' seeded inputs, and by the unit tests' own criterion (adjoint == finite differences
// of the forward function, check_adjoint.h:36-88) in double precision.
#pragma once

namespace dvro {

template <typename R>
static void fn_blend(int mode, int n, const R* x, R* out) {
  for (int i = 0; i < n; ++i) {
    const R* v = x + 9 * i;
    const V4<R> r = blend<R>(mode, {v[0], v[1], v[2], v[3]}, {v[4], v[5], v[6], v[7]}, v[8]);
    out[4 * i] = r.x; out[4 * i + 1] = r.y; out[4 * i + 2] = r.z; out[4 * i + 3] = r.w;
  }
}

template <typename R>
static void fn_blend_adjoint(int mode, int n, const R* x, const R* e, const R* g, R* z, R* accIn) {
  for (int i = 0; i < n; ++i) {
    const R* v = x + 9 * i;
    V4<R> adjIn, adjCur;
    R adjStep;
    const V4<R> in = blend_adjoint<R>(mode, {e[4 * i], e[4 * i + 1], e[4 * i + 2], e[4 * i + 3]}, {v[4], v[5], v[6], v[7]},
                                      v[8], {g[4 * i], g[4 * i + 1], g[4 * i + 2], g[4 * i + 3]}, adjIn, adjCur, adjStep);
    R* o = z + 9 * i;
    o[0] = adjIn.x; o[1] = adjIn.y; o[2] = adjIn.z; o[3] = adjIn.w;
    o[4] = adjCur.x; o[5] = adjCur.y; o[6] = adjCur.z; o[7] = adjCur.w;
    o[8] = adjStep;
    accIn[4 * i] = in.x; accIn[4 * i + 1] = in.y; accIn[4 * i + 2] = in.z; accIn[4 * i + 3] = in.w;
  }
}

template <typename R>
static DvrTensor fn_tf_tensor(const R* tf, int Rn, int C) {
  return DvrTensor{const_cast<R*>(tf), 3, {1, Rn, C}, {int64_t(Rn) * C, C, 1}};
}

template <typename R>
static void fn_tf(int mode, int Rn, int C, const R* tf, int n, const R* density, R* out) {
  const Scene<R> s(fn_tf_tensor(tf, Rn, C), mode);
  for (int i = 0; i < n; ++i) {
    const V4<R> c = tf_eval(s, 0, density[i]);
    out[4 * i] = c.x; out[4 * i + 1] = c.y; out[4 * i + 2] = c.z; out[4 * i + 3] = c.w;
  }
}

template <typename R>
static void fn_tf_adjoint(int mode, int Rn, int C, const R* tf, int n, const R* density, const R* g, R* adjDensity,
                          R* adjTf) {
  const Scene<R> s(fn_tf_tensor(tf, Rn, C), mode);
  for (int i = 0; i < n; ++i) {
    R* table = adjTf + size_t(i) * Rn * C;
    auto acc = [&](int r, int c, R v) { table[r * C + c] += v; };
    adjDensity[i] = tf_adjoint(s, 0, density[i], V4<R>{g[4 * i], g[4 * i + 1], g[4 * i + 2], g[4 * i + 3]}, true, acc);
  }
}

template <typename R>
static Scene<R> fn_volume_scene(const R* volume, int X, int Y, int Z, R* dummyTf) {
  Scene<R> s(fn_tf_tensor(dummyTf, 1, 2), DVR_TF_IDENTITY);
  const DvrTensor v{const_cast<R*>(volume), 4, {1, X, Y, Z}, {int64_t(X) * Y * Z, int64_t(Y) * Z, Z, 1}};
  s.volume = View<R>(v, 4);
  s.X = X; s.Y = Y; s.Z = Z;
  s.filter = DVR_FILTER_TRILINEAR;
  return s;
}

template <typename R>
static void fn_interp(int X, int Y, int Z, const R* volume, int n, const R* pos, R* out) {
  R dummy[2] = {1, 1};
  const Scene<R> s = fn_volume_scene(volume, X, Y, Z, dummy);
  for (int i = 0; i < n; ++i) {
    Sampler<R, 1> smp(s, 0);
    smp.fetch({pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]}, out + i);
  }
}

template <typename R>
static void fn_interp_adjoint(int X, int Y, int Z, const R* volume, int n, const R* pos, const R* g, R* adjPos,
                              R* adjVolume) {
  R dummy[2] = {1, 1};
  const Scene<R> s = fn_volume_scene(volume, X, Y, Z, dummy);
  const int64_t N = int64_t(X) * Y * Z;
  for (int i = 0; i < n; ++i) {
    const DvrTensor a{adjVolume + i * N, 4, {1, X, Y, Z}, {N, int64_t(Y) * Z, Z, 1}};
    Sampler<R, 1> smp(s, 0, View<R>(a, 4, false), true);
    R d;
    smp.fetch({pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]}, &d);
    V3<R> ap;
    smp.adjoint(g + i, ap);
    smp.finish_adjoint();
    adjPos[3 * i] = ap.x; adjPos[3 * i + 1] = ap.y; adjPos[3 * i + 2] = ap.z;
  }
}

}  // namespace dvro
