// Instances of the adjoint kernel for ONE (tf, blend) pair, selected with -DDVR_TF=.. -DDVR_BL=..
// (one translation unit per pair so the ~20 instances per pair compile in parallel).
// Flags: hasVol x {no TF grad, register TF grad, shared-memory TF grad} x {no ray gradients, stepsize only, camera
// (+ stepsize as a run-time flag)}.
#include "dvr_kernels.cuh"
#include "dvr_launch.h"

#ifndef DVR_TF
#error "compile with -DDVR_TF=<tf mode> -DDVR_BL=<blend mode>"
#endif
#define DVR_CAT3(a, b, c) a##b##_##c
#define DVR_NAME(tf, bl) DVR_CAT3(launch_adjoint_, tf, bl)

namespace dvr {

template <int F, bool V, int A, int R>
static cudaError_t go(const RayInputs& in, const AdjointArgs& aa, dim3 grid, size_t smem, cudaStream_t s) {
  if constexpr (!V && A == kTfAccNone && R == kRayNone) {
    return kErrNoInstance;  // nothing to compute
  } else if constexpr (DVR_TF == kTfIdentity && A == kTfAccSmem) {
    return kErrNoInstance;  // the 2-entry identity table always lives in registers
  } else if constexpr ((DVR_TF == kTfTexture || DVR_TF == kTfLinear) && A == kTfAccReg) {
    return kErrNoInstance;  // two-row TFs always use the row window + per-warp shared table (tfacc_for)
  } else {
    return launch_kernel(k_adjoint<F, DVR_TF, DVR_BL, V, A, R>, grid, kThreads, smem, s, in, aa);
  }
}
template <int F, bool V, int A>
static cudaError_t by_ray(int r, const RayInputs& in, const AdjointArgs& aa, dim3 g, size_t sm, cudaStream_t s) {
  switch (r) {
    case kRayNone: return go<F, V, A, kRayNone>(in, aa, g, sm, s);
    case kRayStep: return go<F, V, A, kRayStep>(in, aa, g, sm, s);
    case kRayCam: return go<F, V, A, kRayCam>(in, aa, g, sm, s);
    default: return kErrNoInstance;
  }
}
template <int F, bool V>
static cudaError_t by_acc(int a, int r, const RayInputs& in, const AdjointArgs& aa, dim3 g, size_t sm, cudaStream_t s) {
  switch (a) {
    case kTfAccNone: return by_ray<F, V, kTfAccNone>(r, in, aa, g, sm, s);
    case kTfAccReg: return by_ray<F, V, kTfAccReg>(r, in, aa, g, sm, s);
    case kTfAccSmem: return by_ray<F, V, kTfAccSmem>(r, in, aa, g, sm, s);
    default: return kErrNoInstance;
  }
}

cudaError_t DVR_NAME(DVR_TF, DVR_BL)(int filter, bool hasVol, int tfacc, int ray, const RayInputs& in,
                                     const AdjointArgs& aa, dim3 grid, size_t smem, cudaStream_t s) {
  if (filter == kFilterNearest)
    return hasVol ? by_acc<kFilterNearest, true>(tfacc, ray, in, aa, grid, smem, s)
                  : by_acc<kFilterNearest, false>(tfacc, ray, in, aa, grid, smem, s);
  if (filter == kFilterTrilinear)
    return hasVol ? by_acc<kFilterTrilinear, true>(tfacc, ray, in, aa, grid, smem, s)
                  : by_acc<kFilterTrilinear, false>(tfacc, ray, in, aa, grid, smem, s);
  if (filter == kFilterTrilinearQuad)
    return hasVol ? by_acc<kFilterTrilinearQuad, true>(tfacc, ray, in, aa, grid, smem, s)
                  : by_acc<kFilterTrilinearQuad, false>(tfacc, ray, in, aa, grid, smem, s);
  return kErrNoInstance;
}

}  // namespace dvr
