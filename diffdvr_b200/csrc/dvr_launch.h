// Host-visible launchers of the precompiled sm_100a kernel instances. They replace the
// reference's NVRTC kernel loader (renderer/kernel_loader.cpp:193-261,356-457): the
// (filter, tf, blend, flags) tuple selects an instance from a static switch instead of
// assembling a template-name string and compiling it at run time.
#pragma once
#include <cuda_runtime.h>
#include "dvr_params.h"

// The (tf, blend) pairs that have instance translation units: ONE list per pass declares the launchers below, drives
// the dispatch switches of dvr_capi.cu and answers dvr_has_variant_pass(), so the three cannot drift apart (a pair
// missing from the Makefile's TFS_* lists fails at link time). TF ids: 0 Identity, 1 Texture, 2 Linear, 3 Gaussian,
// 6 GaussianLog (no forward-mode: its forward derivatives are deleted in the reference too, renderer_tf.cuh:692-702).
#define DVR_FOR_EACH_ADJOINT_INSTANCE(X) X(0, 0) X(0, 1) X(1, 0) X(1, 1) X(2, 0) X(2, 1) X(3, 0) X(3, 1) X(6, 0) X(6, 1)
#define DVR_FOR_EACH_FWDGRAD_INSTANCE(X) X(0, 0) X(0, 1) X(1, 0) X(1, 1) X(2, 0) X(2, 1) X(3, 0) X(3, 1)
// the plain forward kernels live in one translation unit (dvr_forward.cu) that switches over the same TF set
#define DVR_FOR_EACH_FORWARD_INSTANCE(X) DVR_FOR_EACH_ADJOINT_INSTANCE(X)

namespace dvr {

// the launchers' own "no precompiled instance for this combination" (never returned by the CUDA runtime)
constexpr cudaError_t kErrNoInstance = static_cast<cudaError_t>(20001);

cudaError_t launch_forward(int filter, int tf, int blend, const RayInputs& in, const ForwardOutputs& out,
                           dim3 grid, size_t smem, cudaStream_t stream);

// tfacc: 0 none, 1 register accumulators, 2 shared-memory tables (see dvr_kernels.cuh)
// ray: 0 no per-ray gradients, 1 stepsize only, 2 camera (+ stepsize)
cudaError_t launch_adjoint(int filter, int tf, int blend, bool hasVol, int tfacc, int ray,
                           const RayInputs& in, const AdjointArgs& aa, dim3 grid, size_t smem,
                           cudaStream_t stream);

// chunk: number of tangent slots per pass (1, 6 or 8); 0 = the input-major single pass (ray / stepsize tangents only);
// -1 = the sample-major single pass for TF tangents (Identity / Texture / Linear; reads `out` as written by the plain
// forward kernel and accumulates into ga.grad)
cudaError_t launch_forward_gradients(int filter, int tf, int blend, int chunk, bool hasPos, bool hasTf,
                                     const RayInputs& in, const ForwardOutputs& out,
                                     const ForwardGradArgs& ga, dim3 grid, size_t smem,
                                     cudaStream_t stream);

// pre-shaded volumes (FilterPreshaded + TFPreshaded), dvr_preshaded.cu
cudaError_t launch_forward_preshaded(int blend, const RayInputs& in, const ForwardOutputs& out, dim3 grid,
                                     cudaStream_t stream);
cudaError_t launch_adjoint_preshaded(int blend, bool hasVol, bool hasRay, const RayInputs& in,
                                     const AdjointArgs& aa, dim3 grid, cudaStream_t stream);

cudaError_t launch_fwdvars_to_gradients(const FwdVarsArgs& a, cudaStream_t stream);
// zero-fill of a strided B*H*W*D*4 Jacobian (dvr_misc.cu)
cudaError_t launch_zero_jacobian(const ForwardGradArgs& ga, int W, int H, int B, cudaStream_t stream);

cudaError_t launch_count_cells(const RayInputs& in, const int* term, long long tB, long long tY, long long tX,
                               unsigned long long* counts, dim3 grid, cudaStream_t stream);

// QUAD copy of the volume for kFilterTrilinearQuad (dvr_misc.cu): nb batch entries of X*Y*Z float4
cudaError_t launch_build_quads(const RayInputs& in, int nb, float4* quads, cudaStream_t stream);
// adjVolume[b][x][y][z] += the four quad components that target voxel (x, y, z); nb gradient volumes
cudaError_t launch_fold_quads(const RayInputs& in, const AdjointArgs& aa, int nb, cudaStream_t stream);

// Camera::generateRays forward / adjoint and Renderer::compareToImage (dvr_camera.cu)
cudaError_t launch_generate_rays(const GenRaysArgs& a, cudaStream_t stream);
cudaError_t launch_generate_rays_adjoint(const GenRaysArgs& a, cudaStream_t stream);  // 2 launches
cudaError_t launch_compare_to_image(const CompareArgs& a, cudaStream_t stream);       // reduce: 2 launches

// TFUtils::preshadeVolume forward / adjoint (dvr_tfutils.cu)
cudaError_t launch_preshade(int tf, bool adjoint, const RayInputs& in, const PreshadeArgs& a, size_t smem,
                            cudaStream_t stream);

// renderer::Volume on the device (dvr_volume.cu)
cudaError_t launch_implicit_volume(const ImplicitArgs& a, cudaStream_t stream);
cudaError_t launch_volume_raw(int dtype, bool toRaw, const RawVolumeArgs& a, cudaStream_t stream);
cudaError_t launch_downsample_volume(const DownsampleArgs& a, cudaStream_t stream);

// per-(tf, blend) translation units (parallel compilation)
#define DVR_DECL_ADJ(TF, BL)                                                                          \
  cudaError_t launch_adjoint_##TF##_##BL(int filter, bool hasVol, int tfacc, int ray,                 \
                                         const RayInputs& in, const AdjointArgs& aa, dim3 grid,       \
                                         size_t smem, cudaStream_t stream);
DVR_FOR_EACH_ADJOINT_INSTANCE(DVR_DECL_ADJ)
#undef DVR_DECL_ADJ
#define DVR_DECL_FG(TF, BL)                                                                           \
  cudaError_t launch_fwdgrad_##TF##_##BL(int filter, int chunk, bool hasPos, bool hasTf,              \
                                         const RayInputs& in, const ForwardOutputs& out,              \
                                         const ForwardGradArgs& ga, dim3 grid, size_t smem,           \
                                         cudaStream_t stream);
DVR_FOR_EACH_FWDGRAD_INSTANCE(DVR_DECL_FG)
#undef DVR_DECL_FG

template <typename K, typename... Args>
inline cudaError_t launch_kernel(K kernel, dim3 grid, int threads, size_t smem, cudaStream_t stream,
                                 const Args&... args) {
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (e != cudaSuccess) return e;
  }
  kernel<<<grid, threads, smem, stream>>>(args...);
  return cudaGetLastError();
}

}  // namespace dvr
