// Instances of the forward-mode kernel for ONE (tf, blend) pair (-DDVR_TF=.. -DDVR_BL=..):
// {Nearest, Trilinear, Trilinear on the QUAD copy} x ({input-major} + {sample-major TF Jacobian} + chunk {1, 6, 8} x
// {tf, pos+tf}).
#include "dvr_fwdgrad.cuh"
#include "dvr_tfjac.cuh"
#include "dvr_launch.h"

#ifndef DVR_TF
#error "compile with -DDVR_TF=<tf mode> -DDVR_BL=<blend mode>"
#endif
#define DVR_CAT3(a, b, c) a##b##_##c
#define DVR_NAME(tf, bl) DVR_CAT3(launch_fwdgrad_, tf, bl)

namespace dvr {

template <int F, int N>
static cudaError_t by_flags(bool p, bool t, const RayInputs& in, const ForwardOutputs& out,
                            const ForwardGradArgs& ga, dim3 g, size_t sm, cudaStream_t s) {
  if (p && t) return launch_kernel(k_forward_gradients<F, DVR_TF, DVR_BL, N, true, true>, g, kThreads, sm, s, in, out, ga);
  // ray / stepsize tangents without TF tangents go through k_forward_gradients_inputs (chunk 0)
  if (t) return launch_kernel(k_forward_gradients<F, DVR_TF, DVR_BL, N, false, true>, g, kThreads, sm, s, in, out, ga);
  return kErrNoInstance;
}
template <int F>
static cudaError_t by_chunk(int n, bool p, bool t, const RayInputs& in, const ForwardOutputs& out,
                            const ForwardGradArgs& ga, dim3 g, size_t sm, cudaStream_t s) {
  switch (n) {
    case -1:  // sample-major single pass: TF tangents (k_tf_jacobian)
      if constexpr (DVR_TF == kTfIdentity || DVR_TF == kTfTexture || DVR_TF == kTfLinear || DVR_TF == kTfGaussian) {
        if (!t) return kErrNoInstance;
        return launch_kernel(k_tf_jacobian<F, DVR_TF, DVR_BL>, g, kThreads, sm, s, in, out, ga);
      } else {
        return kErrNoInstance;
      }
    case 0:  // input-major single pass: ray / stepsize tangents only (k_forward_gradients_inputs)
      if (!p || t) return kErrNoInstance;
      return launch_kernel(k_forward_gradients_inputs<F, DVR_TF, DVR_BL>, g, kThreads, sm, s, in, out, ga);
    case 1: return by_flags<F, 1>(p, t, in, out, ga, g, sm, s);
    case 6: return by_flags<F, 6>(p, t, in, out, ga, g, sm, s);
    case 8: return by_flags<F, 8>(p, t, in, out, ga, g, sm, s);
    default: return kErrNoInstance;
  }
}

cudaError_t DVR_NAME(DVR_TF, DVR_BL)(int filter, int chunk, bool hasPos, bool hasTf, const RayInputs& in,
                                     const ForwardOutputs& out, const ForwardGradArgs& ga, dim3 grid,
                                     size_t smem, cudaStream_t s) {
  if (filter == kFilterNearest) return by_chunk<kFilterNearest>(chunk, hasPos, hasTf, in, out, ga, grid, smem, s);
  if (filter == kFilterTrilinear) return by_chunk<kFilterTrilinear>(chunk, hasPos, hasTf, in, out, ga, grid, smem, s);
  if (filter == kFilterTrilinearQuad) return by_chunk<kFilterTrilinearQuad>(chunk, hasPos, hasTf, in, out, ga, grid, smem, s);
  return kErrNoInstance;
}

}  // namespace dvr
