// TEST INFRASTRUCTURE ONLY -- never linked into the product.
//
// extern "C" access to the REFERENCE's own device functions, one call per function, in the form the reference's
// unit tests drive them (unittests/testAdjointBlending.cpp, testAdjointTF.cpp, testAdjointInterpolation.cpp):
//   fn_blend / fn_blend_adjoint    kernel::Blending<mode>::blend / adjoint          (renderer/renderer_blending.cuh)
//   fn_tf / fn_tf_adjoint          kernel::TransferFunctionEval<mode>::eval / adjoint<true, false> (renderer_tf.cuh)
//   fn_interp / fn_interp_adjoint  kernel::VolumeInterpolation<FilterTrilinear, true, true>::fetch / adjoint
//                                  (renderer/renderer_interpolation.cuh:270-392)
// compiled from the headers where they lie. Arrays are dense, `n` independent test vectors per call; the same
// signatures are exported by the CPU port (oracle/dvr_oracle_fn.h) and, on device pointers, by the harness over the
// product's CUDA device functions (tests/csrc/devfn_harness.cu).
#include "ref_shim_common.h"

using namespace refshim;

namespace {

template <kernel::BlendMode M>
void blend_impl(int n, const real_t* x, real_t* out) {
  for (int i = 0; i < n; ++i) {
    const real_t* v = x + 9 * i;
    const real4 r = kernel::Blending<M>::blend(make_real4(v[0], v[1], v[2], v[3]), make_real4(v[4], v[5], v[6], v[7]), v[8]);
    out[4 * i] = r.x; out[4 * i + 1] = r.y; out[4 * i + 2] = r.z; out[4 * i + 3] = r.w;
  }
}

template <kernel::BlendMode M>
void blend_adj_impl(int n, const real_t* x, const real_t* e, const real_t* g, real_t* z, real_t* accIn) {
  for (int i = 0; i < n; ++i) {
    const real_t* v = x + 9 * i;
    real4 adjAcc, adjCur;
    real_t adjStep;
    const real4 in = kernel::Blending<M>::adjoint(make_real4(e[4 * i], e[4 * i + 1], e[4 * i + 2], e[4 * i + 3]),
                                                  make_real4(v[4], v[5], v[6], v[7]), v[8],
                                                  make_real4(g[4 * i], g[4 * i + 1], g[4 * i + 2], g[4 * i + 3]), adjAcc,
                                                  adjCur, adjStep);
    real_t* o = z + 9 * i;
    o[0] = adjAcc.x; o[1] = adjAcc.y; o[2] = adjAcc.z; o[3] = adjAcc.w;
    o[4] = adjCur.x; o[5] = adjCur.y; o[6] = adjCur.z; o[7] = adjCur.w;
    o[8] = adjStep;
    accIn[4 * i] = in.x; accIn[4 * i + 1] = in.y; accIn[4 * i + 2] = in.z; accIn[4 * i + 3] = in.w;
  }
}

template <kernel::TFMode M>
void tf_impl(const kernel::Tensor3Read& tf, int n, const real_t* density, real_t* out) {
  kernel::TransferFunctionEval<M> tfEval;
  for (int i = 0; i < n; ++i) {
    const real4 c = tfEval.eval(tf, 0, density[i]);
    out[4 * i] = c.x; out[4 * i + 1] = c.y; out[4 * i + 2] = c.z; out[4 * i + 3] = c.w;
  }
}

template <kernel::TFMode M>
void tf_adj_impl(const kernel::Tensor3Read& tf, int R, int C, int n, const real_t* density, const real_t* g,
                 real_t* adjDensity, real_t* adjTf) {
  kernel::TransferFunctionEval<M> tfEval;
  for (int i = 0; i < n; ++i) {
    DvrTensor t{adjTf + size_t(i) * R * C, 3, {1, R, C}, {int64_t(R) * C, C, 1}};
    kernel::BTensor3RW acc = make_acc<kernel::BTensor3RW, 3>(t);
    real_t ad = 0;
    tfEval.template adjoint<true, false>(tf, 0, density[i], make_real4(g[4 * i], g[4 * i + 1], g[4 * i + 2], g[4 * i + 3]),
                                         ad, acc, nullptr);
    adjDensity[i] = ad;
  }
}

}  // namespace

extern "C" int REFSYM(fn_blend)(int32_t mode, int32_t n, const real_t* x, real_t* out) {
  return guarded([&] {
    if (mode == DVR_BLEND_BEER_LAMBERT) blend_impl<kernel::BlendBeerLambert>(n, x, out);
    else blend_impl<kernel::BlendAlpha>(n, x, out);
  });
}

extern "C" int REFSYM(fn_blend_adjoint)(int32_t mode, int32_t n, const real_t* x, const real_t* e, const real_t* g,
                                        real_t* z, real_t* accIn) {
  return guarded([&] {
    if (mode == DVR_BLEND_BEER_LAMBERT) blend_adj_impl<kernel::BlendBeerLambert>(n, x, e, g, z, accIn);
    else blend_adj_impl<kernel::BlendAlpha>(n, x, e, g, z, accIn);
  });
}

#define TF_SWITCH(mode, CALL)                                                     \
  switch (mode) {                                                                 \
    case DVR_TF_IDENTITY: { constexpr auto M = kernel::TFIdentity; CALL; } break;  \
    case DVR_TF_TEXTURE: { constexpr auto M = kernel::TFTexture; CALL; } break;    \
    case DVR_TF_LINEAR: { constexpr auto M = kernel::TFLinear; CALL; } break;      \
    case DVR_TF_GAUSSIAN: { constexpr auto M = kernel::TFGaussian; CALL; } break;  \
    case DVR_TF_GAUSSIAN_LOG: { constexpr auto M = kernel::TFGaussianLog; CALL; } break; \
    default: throw std::runtime_error("unknown tf mode");                        \
  }

extern "C" int REFSYM(fn_tf)(int32_t mode, int32_t R, int32_t C, const real_t* tf, int32_t n, const real_t* density,
                             real_t* out) {
  return guarded([&] {
    DvrTensor t{const_cast<real_t*>(tf), 3, {1, R, C}, {int64_t(R) * C, C, 1}};
    auto acc = make_acc<kernel::Tensor3Read, 3>(t);
    TF_SWITCH(mode, tf_impl<M>(acc, n, density, out));
  });
}

extern "C" int REFSYM(fn_tf_adjoint)(int32_t mode, int32_t R, int32_t C, const real_t* tf, int32_t n,
                                     const real_t* density, const real_t* g, real_t* adjDensity, real_t* adjTf) {
  return guarded([&] {
    DvrTensor t{const_cast<real_t*>(tf), 3, {1, R, C}, {int64_t(R) * C, C, 1}};
    auto acc = make_acc<kernel::Tensor3Read, 3>(t);
    TF_SWITCH(mode, tf_adj_impl<M>(acc, R, C, n, density, g, adjDensity, adjTf));
  });
}

extern "C" int REFSYM(fn_interp)(int32_t X, int32_t Y, int32_t Z, const real_t* volume, int32_t n, const real_t* pos,
                                 real_t* out) {
  return guarded([&] {
    DvrTensor t{const_cast<real_t*>(volume), 4, {1, X, Y, Z}, {int64_t(X) * Y * Z, int64_t(Y) * Z, Z, 1}};
    auto vol = make_acc<kernel::Tensor4Read, 4>(t);
    const int3 size = make_int3(X, Y, Z);
    for (int i = 0; i < n; ++i) {
      kernel::VolumeInterpolation<kernel::FilterTrilinear> vi;
      out[i] = vi.fetch(vol, size, 0, make_real3(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]));
    }
  });
}

extern "C" int REFSYM(fn_interp_adjoint)(int32_t X, int32_t Y, int32_t Z, const real_t* volume, int32_t n,
                                         const real_t* pos, const real_t* g, real_t* adjPos, real_t* adjVolume) {
  return guarded([&] {
    const int64_t N = int64_t(X) * Y * Z;
    DvrTensor t{const_cast<real_t*>(volume), 4, {1, X, Y, Z}, {N, int64_t(Y) * Z, Z, 1}};
    auto vol = make_acc<kernel::Tensor4Read, 4>(t);
    const int3 size = make_int3(X, Y, Z);
    for (int i = 0; i < n; ++i) {
      DvrTensor a{adjVolume + i * N, 4, {1, X, Y, Z}, {N, int64_t(Y) * Z, Z, 1}};
      auto adj = make_acc<kernel::BTensor4RW, 4>(a);
      real3 ap;
      {
        kernel::VolumeInterpolation<kernel::FilterTrilinear, true, true> vi(0, size, adj);
        vi.fetch(vol, size, 0, make_real3(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]));
        vi.adjoint(g[i], ap);
      }  // the destructor flushes the cell's adjoints (renderer_interpolation.cuh:329)
      adjPos[3 * i] = ap.x; adjPos[3 * i + 1] = ap.y; adjPos[3 * i + 2] = ap.z;
    }
  });
}
