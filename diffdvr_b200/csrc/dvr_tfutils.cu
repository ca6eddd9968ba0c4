// TFUtils::preshadeVolume: the producer of the 4*X*Y*Z colour volume that the
// pre-shaded marcher (dvr_preshaded.cu) consumes, and its adjoint with respect to the density volume
// (renderer/tf_utils_impl.cu:16-92 forward, :98-183 adjoint; the reference has no TF adjoint here either,
// ":159 TODO"). One thread per voxel, Z fastest; the TF is staged in shared memory once per CTA and evaluated by
// the same tf_eval / tf_adjoint device functions as the marcher. Streaming kernels: 4 B read + 16 B written per
// voxel forward, 20 B read + 4 B written per voxel adjoint.
#include "dvr_kernels.cuh"
#include "dvr_launch.h"

namespace dvr {

namespace {

template <int TF, bool ADJOINT>
__global__ void __launch_bounds__(256) k_preshade(const RayInputs in, const PreshadeArgs a) {
  extern __shared__ __align__(16) float s_tf[];
  stage_tf<TF>(in, 0, s_tf);
  __syncthreads();
  const unsigned n = unsigned(in.X) * unsigned(in.Y) * unsigned(in.Z);
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const unsigned z = i % unsigned(in.Z), xy = i / unsigned(in.Z);
    const unsigned y = xy % unsigned(in.Y), x = xy / unsigned(in.Y);
    const float density = __ldg(in.volume + (long long)x * in.sx + (long long)y * in.sy + (long long)z * in.sz);
    float* c = a.color + (long long)x * a.cX + (long long)y * a.cY + (long long)z * a.cZ;
    if constexpr (!ADJOINT) {
      const f4 v = tf_eval<TF>(s_tf, in.R, density);
      c[0] = v.x; c[a.cC] = v.y; c[2 * a.cC] = v.z; c[3 * a.cC] = v.w;
    } else {
      const f4 adjC = {__ldg(c), __ldg(c + a.cC), __ldg(c + 2 * a.cC), __ldg(c + 3 * a.cC)};
      TfCache<TF> cache;
      tf_eval_cached<TF>(s_tf, in.R, density, cache);
      TfGrad<TF, kTfAccNone> none;
      a.adjVolume[(long long)x * a.ax + (long long)y * a.ay + (long long)z * a.az] =
          tf_adjoint<TF, kTfAccNone>(s_tf, in.R, density, adjC, none, cache);
    }
  }
}

template <int TF>
cudaError_t go(bool adjoint, const RayInputs& in, const PreshadeArgs& a, size_t smem, cudaStream_t s) {
  const unsigned n = unsigned(in.X) * unsigned(in.Y) * unsigned(in.Z);
  unsigned blocks = (n + 255) / 256;
  if (blocks > 148u * 32u) blocks = 148u * 32u;  // grid-stride: the TF is staged once per CTA
  if (adjoint) return launch_kernel(k_preshade<TF, true>, dim3(blocks), 256, smem, s, in, a);
  return launch_kernel(k_preshade<TF, false>, dim3(blocks), 256, smem, s, in, a);
}

}  // namespace

cudaError_t launch_preshade(int tf, bool adjoint, const RayInputs& in, const PreshadeArgs& a, size_t smem,
                            cudaStream_t s) {
  switch (tf) {
    case kTfIdentity: return go<kTfIdentity>(adjoint, in, a, smem, s);
    case kTfTexture: return go<kTfTexture>(adjoint, in, a, smem, s);
    case kTfLinear: return go<kTfLinear>(adjoint, in, a, smem, s);
    case kTfGaussian: return go<kTfGaussian>(adjoint, in, a, smem, s);
    case kTfGaussianLog: return go<kTfGaussianLog>(adjoint, in, a, smem, s);
    default: return kErrNoInstance;
  }
}

}  // namespace dvr
