// TEST HARNESS -- not part of the product library.
//
// Runs the PRODUCT's CUDA device functions (diffdvr_b200/csrc/dvr_device.cuh, dvr_kernels.cuh, included as they are)
// one call per test vector, in the form the reference's unit tests drive the reference's device functions
// (unittests/testAdjointBlending.cpp, testAdjointTF.cpp, testAdjointInterpolation.cpp), so that tests/test_gpu_devfn.py
// can compare them function by function with the CPU oracle and the reference build on the same seeded inputs:
//   devfn_blend / devfn_blend_adjoint    blend_alpha + blend_over / blend_adjoint
//   devfn_tf / devfn_tf_adjoint          stage_tf + tf_eval(_cached) / tf_adjoint with every TF-gradient accumulator
//                                        (register table, cached rows + atomics)
//   devfn_interp / devfn_interp_adjoint  Sampler<Trilinear | TrilinearQuad>::fetch / adjoint / flush
// Same signatures as oracle/ref_shim_fn.cpp and oracle/dvr_oracle_fn.h, on DEVICE pointers, float only.
// Built by tests/csrc/Makefile into tests/_build/libdvr_devfn.so (__graft_entry__.build()).
#include <cstdint>

#include "dvr_kernels.cuh"

using namespace dvr;

namespace {

template <int BLEND>
__global__ void k_blend(int n, const float* x, float* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* v = x + 9 * i;
  const f4 acc = {v[0], v[1], v[2], v[3]}, cur = {v[4], v[5], v[6], v[7]};
  const f4 r = blend_over(acc, cur, blend_alpha<BLEND>(cur.w, v[8]));
  out[4 * i] = r.x; out[4 * i + 1] = r.y; out[4 * i + 2] = r.z; out[4 * i + 3] = r.w;
}

template <int BLEND>
__global__ void k_blend_adjoint(int n, const float* x, const float* e, const float* g, float* z, float* accIn) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* v = x + 9 * i;
  f4 acc = {e[4 * i], e[4 * i + 1], e[4 * i + 2], e[4 * i + 3]};
  f4 adjAcc = {g[4 * i], g[4 * i + 1], g[4 * i + 2], g[4 * i + 3]};
  const f4 cur = {v[4], v[5], v[6], v[7]};
  f4 adjCur;
  float adjStep;
  blend_adjoint<BLEND>(acc, adjAcc, cur, v[8], adjCur, adjStep);
  float* o = z + 9 * i;
  o[0] = adjAcc.x; o[1] = adjAcc.y; o[2] = adjAcc.z; o[3] = adjAcc.w;
  o[4] = adjCur.x; o[5] = adjCur.y; o[6] = adjCur.z; o[7] = adjCur.w;
  o[8] = adjStep;
  accIn[4 * i] = acc.x; accIn[4 * i + 1] = acc.y; accIn[4 * i + 2] = acc.z; accIn[4 * i + 3] = acc.w;
}

// the transmittance form of the Beer-Lambert adjoint step the adjoint kernels use (blend_adjoint_T, DVR_BLEND_T): same
// inputs / outputs; the colour channels of the reconstructed accumulator are not produced (the kernels never read them)
__global__ void k_blend_adjoint_T(int n, const float* x, const float* e, const float* g, float* z, float* accIn) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* v = x + 9 * i;
  float T = 1.f - e[4 * i + 3];
  f4 adjAcc = {g[4 * i], g[4 * i + 1], g[4 * i + 2], g[4 * i + 3]};
  const f4 cur = {v[4], v[5], v[6], v[7]};
  f4 adjCur;
  float adjStep;
  blend_adjoint_T(T, adjAcc, cur, v[8], adjCur, adjStep);
  float* o = z + 9 * i;
  o[0] = adjAcc.x; o[1] = adjAcc.y; o[2] = adjAcc.z; o[3] = adjAcc.w;
  o[4] = adjCur.x; o[5] = adjCur.y; o[6] = adjCur.z; o[7] = adjCur.w;
  o[8] = adjStep;
  accIn[4 * i] = accIn[4 * i + 1] = accIn[4 * i + 2] = 0.f;
  accIn[4 * i + 3] = 1.f - T;
}

RayInputs tf_inputs(const float* tf, int R, int C) {
  RayInputs in{};
  in.tf = {tf, 0, (long long)C, 1};
  in.R = R;
  in.C = C;
  return in;
}

template <int TF>
__global__ void k_tf(const RayInputs in, int n, const float* density, float* out) {
  extern __shared__ __align__(16) float s_tf[];
  stage_tf<TF>(in, 0, s_tf);
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const f4 c = tf_eval<TF>(s_tf, in.R, density[i]);
  out[4 * i] = c.x; out[4 * i + 1] = c.y; out[4 * i + 2] = c.z; out[4 * i + 3] = c.w;
}

// adjTf: n tables of R * C floats, zeroed by the caller; every thread owns one
template <int TF, int TFACC>
__global__ void k_tf_adjoint(const RayInputs in, int n, const float* density, const float* g, float* adjDensity,
                             float* adjTf) {
  extern __shared__ __align__(16) float s_tf[];
  stage_tf<TF>(in, 0, s_tf);
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  constexpr int CH = TfChannels<TF>::v;
  float* table = adjTf + (long long)i * in.R * CH;
  TfGrad<TF, TFACC> tfg;
  tfg.init(table, in.R);
  TfCache<TF> cache;
  tf_eval_cached<TF>(s_tf, in.R, density[i], cache);
  const f4 adjC = {g[4 * i], g[4 * i + 1], g[4 * i + 2], g[4 * i + 3]};
  adjDensity[i] = tf_adjoint<TF, TFACC>(s_tf, in.R, density[i], adjC, tfg, cache);
  if constexpr (TFACC == kTfAccReg) {
    for (int k = 0; k < in.R * CH; ++k) {
      float v = 0.f;
#pragma unroll
      for (int r = 0; r < TfGrad<TF, TFACC>::NREG; ++r)
        if (r == k) v = tfg.g[r];
      table[k] = v;
    }
  } else if constexpr (TFACC == kTfAccSmem) {
    tfg.flush_rows();
  }
}

RayInputs volume_inputs(const float* volume, int X, int Y, int Z, const float4* quads) {
  RayInputs in{};
  in.X = X; in.Y = Y; in.Z = Z;
  in.volume = volume;
  in.sx = Y * Z; in.sy = Z; in.sz = 1;
  in.quads = quads;
#if DVR_QUAD_XINNER  // as set_quad_indexing() of dvr_capi.cu
  in.quadSZ = unsigned(X) + 1u;
  in.quadYZ = in.quadSZ * unsigned(Z);
  in.quadBias = unsigned(kTwo23Bits) * (in.quadYZ + in.quadSZ + 1u);
  in.quadMax = in.quadYZ * unsigned(Y) - 2u;
#else
  in.quadYZ = unsigned(Y) * unsigned(Z);
  in.quadSZ = 1u;
  in.quadBias = unsigned(kTwo23Bits) * (in.quadYZ + unsigned(Z) + 1u);
  in.quadMax = unsigned(X) * in.quadYZ - 1u;
#endif
  return in;
}

template <int FILTER>
__global__ void k_interp(const RayInputs in, int n, const float* pos, float* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Sampler<FILTER, false> smp;
  smp.init(in, 0, nullptr, 0);
  out[i] = smp.fetch(in, nullptr, pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]);
}

// adjVolume: n volumes (FILTER = Trilinear: X*Y*Z floats each; TrilinearQuad: (X+1)*Y*Z float4 each), zeroed by the caller
template <int FILTER>
__global__ void k_interp_adjoint(const RayInputs in, const AdjointArgs aa, int n, const float* pos, const float* g,
                                 float* adjPos, float* adjVolume) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const long long N = FILTER == kFilterTrilinearQuad ? (long long)(in.X + 1) * in.Y * in.Z * 4 : (long long)in.X * in.Y * in.Z;
  Sampler<FILTER, true, DVR_BRANCHFREE_NORAY != 0> smp;  // as the adjoint kernels without ray gradients
  smp.init(in, 0, adjVolume + i * N, 0);
  smp.fetch(in, &aa, pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]);
  const f3 ap = smp.template adjoint<true>(g[i]);
  smp.flush(&aa);
  adjPos[3 * i] = ap.x; adjPos[3 * i + 1] = ap.y; adjPos[3 * i + 2] = ap.z;
}

int done(cudaStream_t s = nullptr) {
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  return e == cudaSuccess ? 0 : int(e);
}
inline int blocks(int n) { return (n + 127) / 128; }

}  // namespace

#define DEVFN_API extern "C" __attribute__((visibility("default")))

DEVFN_API int devfn_blend(int32_t mode, int32_t n, const float* x, float* out) {
  if (mode == kBlendBeer) k_blend<kBlendBeer><<<blocks(n), 128>>>(n, x, out);
  else k_blend<kBlendAlpha><<<blocks(n), 128>>>(n, x, out);
  return done();
}

DEVFN_API int devfn_blend_adjoint(int32_t mode, int32_t n, const float* x, const float* e, const float* g, float* z,
                                  float* accIn) {
  if (mode == 100) k_blend_adjoint_T<<<blocks(n), 128>>>(n, x, e, g, z, accIn);  // Beer-Lambert, transmittance form
  else if (mode == kBlendBeer) k_blend_adjoint<kBlendBeer><<<blocks(n), 128>>>(n, x, e, g, z, accIn);
  else k_blend_adjoint<kBlendAlpha><<<blocks(n), 128>>>(n, x, e, g, z, accIn);
  return done();
}

#define TF_SWITCH(mode, CALL)                                          \
  switch (mode) {                                                      \
    case kTfIdentity: { constexpr int M = kTfIdentity; CALL; } break;   \
    case kTfTexture: { constexpr int M = kTfTexture; CALL; } break;     \
    case kTfLinear: { constexpr int M = kTfLinear; CALL; } break;       \
    case kTfGaussian: { constexpr int M = kTfGaussian; CALL; } break;   \
    case kTfGaussianLog: { constexpr int M = kTfGaussianLog; CALL; } break; \
    default: return -1;                                                \
  }

DEVFN_API int devfn_tf(int32_t mode, int32_t R, int32_t C, const float* tf, int32_t n, const float* density,
                       float* out) {
  const RayInputs in = tf_inputs(tf, R, C);
  TF_SWITCH(mode, (k_tf<M><<<blocks(n), 128, size_t(R) * TfLayout<M>::kRow * 4>>>(in, n, density, out)));
  return done();
}

// acc: 1 = register table (R <= 6), 2 = the shared-memory-table code path (cached rows / atomics)
DEVFN_API int devfn_tf_adjoint(int32_t mode, int32_t R, int32_t C, const float* tf, int32_t n, const float* density,
                               const float* g, float* adjDensity, float* adjTf, int32_t acc) {
  const RayInputs in = tf_inputs(tf, R, C);
  if (acc == kTfAccReg) {
    if (R > kRegRows) return -2;
    TF_SWITCH(mode, (k_tf_adjoint<M, kTfAccReg><<<blocks(n), 128, size_t(R) * TfLayout<M>::kRow * 4>>>(
                        in, n, density, g, adjDensity, adjTf)));
  } else {
    TF_SWITCH(mode, (k_tf_adjoint<M, kTfAccSmem><<<blocks(n), 128, size_t(R) * TfLayout<M>::kRow * 4>>>(
                        in, n, density, g, adjDensity, adjTf)));
  }
  return done();
}

// layout of the QUAD buffers this build expects: 0 = [x][y][z] quads, 1 = [y][z][x] quads (DVR_QUAD_XINNER)
DEVFN_API int devfn_quad_xinner() { return DVR_QUAD_XINNER; }

// quads: nullptr = sample the linear volume in place; else the QUAD copy of it ((X+1)*Y*Z float4, dvr_misc.cu layout)
DEVFN_API int devfn_interp(int32_t X, int32_t Y, int32_t Z, const float* volume, const float* quads, int32_t n,
                           const float* pos, float* out) {
  const RayInputs in = volume_inputs(volume, X, Y, Z, reinterpret_cast<const float4*>(quads));
  if (quads) k_interp<kFilterTrilinearQuad><<<blocks(n), 128>>>(in, n, pos, out);
  else k_interp<kFilterTrilinear><<<blocks(n), 128>>>(in, n, pos, out);
  return done();
}

DEVFN_API int devfn_interp_adjoint(int32_t X, int32_t Y, int32_t Z, const float* volume, const float* quads, int32_t n,
                                   const float* pos, const float* g, float* adjPos, float* adjVolume) {
  const RayInputs in = volume_inputs(volume, X, Y, Z, reinterpret_cast<const float4*>(quads));
  AdjointArgs aa{};
  aa.avx = Y * Z; aa.avy = Z; aa.avz = 1;
  aa.sameStrides = 1;
  if (quads) k_interp_adjoint<kFilterTrilinearQuad><<<blocks(n), 128>>>(in, aa, n, pos, g, adjPos, adjVolume);
  else k_interp_adjoint<kFilterTrilinear><<<blocks(n), 128>>>(in, aa, n, pos, g, adjPos, adjVolume);
  return done();
}
