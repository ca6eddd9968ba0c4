// Instances of the plain forward kernel: {Nearest, Trilinear, Trilinear on the QUAD copy} x {Identity, Texture, Linear,
// Gaussian, GaussianLog} x {BeerLambert, Alpha}. The camera mode is a run-time switch outside
// the marching loop, so the reference's 3 camera instantiations collapse into one.
#include "dvr_kernels.cuh"
#include "dvr_launch.h"

namespace dvr {

template <int F, int B>
static cudaError_t by_tf(int tf, const RayInputs& in, const ForwardOutputs& out, dim3 grid, size_t smem,
                         cudaStream_t s) {
  // one case per (tf, blend) pair of DVR_FOR_EACH_FORWARD_INSTANCE (dvr_launch.h)
#define DVR_CASE(TF, BL) \
  if (tf == TF && B == BL) return launch_kernel(k_forward<F, TF, BL>, grid, kThreads, smem, s, in, out);
  DVR_FOR_EACH_FORWARD_INSTANCE(DVR_CASE)
#undef DVR_CASE
  return kErrNoInstance;
}

cudaError_t launch_forward(int filter, int tf, int blend, const RayInputs& in, const ForwardOutputs& out,
                           dim3 grid, size_t smem, cudaStream_t s) {
  if (filter == kFilterNearest && blend == kBlendBeer) return by_tf<kFilterNearest, kBlendBeer>(tf, in, out, grid, smem, s);
  if (filter == kFilterNearest && blend == kBlendAlpha) return by_tf<kFilterNearest, kBlendAlpha>(tf, in, out, grid, smem, s);
  if (filter == kFilterTrilinear && blend == kBlendBeer) return by_tf<kFilterTrilinear, kBlendBeer>(tf, in, out, grid, smem, s);
  if (filter == kFilterTrilinear && blend == kBlendAlpha) return by_tf<kFilterTrilinear, kBlendAlpha>(tf, in, out, grid, smem, s);
  if (filter == kFilterTrilinearQuad && blend == kBlendBeer) return by_tf<kFilterTrilinearQuad, kBlendBeer>(tf, in, out, grid, smem, s);
  if (filter == kFilterTrilinearQuad && blend == kBlendAlpha) return by_tf<kFilterTrilinearQuad, kBlendAlpha>(tf, in, out, grid, smem, s);
  return kErrNoInstance;
}

}  // namespace dvr
