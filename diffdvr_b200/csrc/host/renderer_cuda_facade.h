// renderer::RendererCuda over the C-ABI (include/dvr_b200.h): the back-end interface of the reference
// (renderer/renderer_cuda.h:14-128) with the same method names, argument lists and bool returns, so that the
// reference's own host layer (renderer/renderer.cpp:405-415, 470-482, 548-556, 722-739) can keep calling
// `RendererCuda::Instance().renderForward(...)` & co unchanged while the NVRTC kernel loader underneath is gone
// (SURVEY.md 8(b) "Backend interface", INTEGRATION.md option C).
//
// Torch-free: plain pointers in, plain C calls out. The accessor structs below have the memory layout of the
// reference's kernel::GenericPackedTensorAccessor<T, N, ..., int32> (renderer/renderer_tensor.cuh:122-269: data pointer,
// int32 sizes[N], int32 strides[N]) and the aggregate structs the field order of renderer/renderer_settings.cuh:150-244,
// so a maintainer who links this facade into the reference tree passes the reference's own structs by reinterpretation
// or compiles the .cpp against the reference's headers with -DDVR_FACADE_USE_REFERENCE_TYPES (then nothing is mirrored).
#pragma once
#include <cstdint>
#include <filesystem>
#include <string>

#include "../../../include/dvr_b200.h"
#include "kernel_types.h"

#ifndef DVR_FACADE_USE_REFERENCE_TYPES
namespace kernel {

// layout of GenericPackedTensorAccessorBase<T, N, PtrTraits, int32, Broadcast> (renderer/renderer_tensor.cuh:122-181)
template <typename T, int N>
struct PackedAccessor32 {
  T* data_ = nullptr;
  int32_t sizes_[N] = {};
  int32_t strides_[N] = {};
  PackedAccessor32() = default;
  PackedAccessor32(T* data, const int64_t* sizes, const int64_t* strides, bool broadcastEmptyDimensions = false)
      : data_(data) {
    for (int i = 0; i < N; ++i) {
      sizes_[i] = int32_t(sizes[i]);
      strides_[i] = (broadcastEmptyDimensions && sizes[i] == 1) ? 0 : int32_t(strides[i]);
    }
  }
  int32_t size(int i) const { return sizes_[i]; }
  int32_t stride(int i) const { return strides_[i]; }
  T* data() const { return data_; }
};

// renderer/renderer_settings.cuh:50-78
typedef PackedAccessor32<const real_t, 1> Tensor1Read;
typedef PackedAccessor32<const real_t, 2> Tensor2Read;
typedef PackedAccessor32<const real_t, 3> Tensor3Read;
typedef PackedAccessor32<const real_t, 4> Tensor4Read;
typedef PackedAccessor32<const real_t, 5> Tensor5Read;
typedef PackedAccessor32<real_t, 1> Tensor1RW;
typedef PackedAccessor32<real_t, 4> Tensor4RW;
typedef PackedAccessor32<real_t, 5> Tensor5RW;
typedef PackedAccessor32<real_t, 3> BTensor3RW;
typedef PackedAccessor32<real_t, 4> BTensor4RW;
typedef PackedAccessor32<const int, 3> ITensor3Read;
typedef PackedAccessor32<int, 3> ITensor3RW;

// renderer/renderer_settings.cuh:150-180
struct RendererInputs {
  int2 screenSize{0, 0};
  Tensor4Read volume;  // B*X*Y*Z
  int3 volumeSize{0, 0, 0};
  Tensor2Read boxMin;   // B*3
  Tensor2Read boxSize;  // B*3
  Tensor4Read cameraRayStart;  // B*H*W*3
  Tensor4Read cameraRayDir;    // B*H*W*3
  Tensor3Read cameraMatrix;    // B*4*4 | B*3*3
  real_t cameraFovYRadians = 0;
  Tensor3Read stepSize;  // B*H*W
  Tensor3Read tf;        // B*R*C
  real_t blendingEarlyOut = real_t(1) - real_t(1e-5);
};
// :183-187
struct RendererOutputs {
  Tensor4RW color;
  ITensor3RW terminationIndex;
};
// :189-204 (the type of this name in kernel_types.h only carries NOT_ENABLED; the full struct lives here)
struct ForwardDifferencesSettingsFull {
  static constexpr const int NOT_ENABLED = -1;
  int d_stepsize = NOT_ENABLED;
  int3 d_rayStart{NOT_ENABLED, NOT_ENABLED, NOT_ENABLED};
  int3 d_rayDir{NOT_ENABLED, NOT_ENABLED, NOT_ENABLED};
  ITensor3Read d_tf;
  bool hasTFDerivatives = false;
};
// :209-216
struct ForwardDifferencesOutput {
  Tensor5RW gradients;  // B*H*W*D*4
};
// :219-223
struct RendererOutputsAsInput {
  Tensor4Read color;
  ITensor3Read terminationIndex;
};
typedef Tensor4Read AdjointColor_t;
// :229-244
struct AdjointOutputs {
  BTensor4RW adj_volume;
  bool stepSizeHasBroadcasting = false;
  BTensor3RW adj_stepSize;
  bool cameraHasBroadcasting = false;
  BTensor4RW adj_cameraRayStart;
  BTensor4RW adj_cameraRayDir;
  BTensor3RW adj_tf;
};

}  // namespace kernel
#define DVR_FACADE_FD_SETTINGS kernel::ForwardDifferencesSettingsFull
#else
#define DVR_FACADE_FD_SETTINGS kernel::ForwardDifferencesSettings
#endif

namespace renderer {

// The conversions the methods below apply, exposed for tests/csrc/facade_shim.cpp (and for integrators who want to call
// the C-ABI themselves, e.g. with a DvrCallOptions carrying a workspace).
namespace facade {
DvrInputs to_c(const kernel::RendererInputs& in, kernel::VolumeFilterMode f, kernel::CameraMode c, kernel::TFMode t,
               kernel::BlendMode b);
DvrOutputs to_c(const kernel::RendererOutputs& out);
DvrOutputs to_c(const kernel::RendererOutputsAsInput& out);
DvrForwardDifferencesSettings to_c(const DVR_FACADE_FD_SETTINGS& s, int numDerivatives);
DvrAdjointOutputs to_c(const kernel::AdjointOutputs& a, bool hasStepsizeDerivative, bool hasCameraDerivative,
                       bool hasTFDerivative, bool tfDelayedAccumulation, bool hasVolumeDerivative);
}  // namespace facade

struct RendererCuda {
  RendererCuda(const RendererCuda&) = delete;
  RendererCuda& operator=(const RendererCuda&) = delete;
  static RendererCuda& Instance();

  // renderer_cuda.h:18-31. Nothing is compiled or cached at run time here, so the cache controls are no-ops.
  bool initCuda();
  void setCudaCacheDir(const std::filesystem::path& path);
  void disableCudaCache();
  void reloadCudaKernels();
  void cleanup();            // releases the library's scratch pool
  void setSyncMode(bool sync);
  bool getSyncMode() const;
  void setDebugMode(bool debug);

  // The reference launches on torch's current stream (renderer_cuda.cpp:105); this facade is torch-free, so the
  // integrator hands it over: setStream(c10::cuda::getCurrentCUDAStream().stream()) before a call. Default: stream 0.
  void setStream(void* cudaStream) { stream_ = cudaStream; }
  // Message of the last failed call (the reference prints the CUDA error and returns false, kernel_loader.cpp:43-49).
  const std::string& lastError() const { return lastError_; }

  bool renderForward(const kernel::RendererInputs& inputs, kernel::RendererOutputs& outputs, int B, int W, int H,
                     kernel::VolumeFilterMode volumeFilterMode, kernel::CameraMode cameraMode, kernel::TFMode tfMode,
                     kernel::BlendMode blendMode);
  std::string getForwardKernelName(kernel::VolumeFilterMode volumeFilterMode, kernel::CameraMode cameraMode,
                                   kernel::TFMode tfMode, kernel::BlendMode blendMode);

  bool renderForwardGradients(const kernel::RendererInputs& inputs, const DVR_FACADE_FD_SETTINGS& settings,
                              kernel::RendererOutputs& outputs, kernel::ForwardDifferencesOutput& gradients, int B, int W,
                              int H, kernel::VolumeFilterMode volumeFilterMode, kernel::CameraMode cameraMode,
                              kernel::TFMode tfMode, kernel::BlendMode blendMode, int numDerivatives,
                              bool hasStepsizeDerivative, bool hasCameraDerivative, bool hasTFDerivative);
  std::string getForwardGradientsKernelName(kernel::VolumeFilterMode volumeFilterMode, kernel::CameraMode cameraMode,
                                            kernel::TFMode tfMode, kernel::BlendMode blendMode, int numDerivatives,
                                            bool hasStepsizeDerivative, bool hasCameraDerivative, bool hasTFDerivative);

  bool forwardVariablesToGradients(const kernel::ForwardDifferencesOutput& forwardVariables,
                                   const kernel::AdjointColor_t& gradientOutputColor,
                                   const DVR_FACADE_FD_SETTINGS& differencesSettings, kernel::AdjointOutputs& adj_outputs,
                                   int B, int W, int H);
  std::string getForwardVariablesToGradientsName();

  void compareToImage_NoReduce(const kernel::Tensor4Read& colorInput, const kernel::Tensor5Read& gradientsInput,
                               const kernel::Tensor4Read& colorReference, kernel::Tensor4RW& differenceOut,
                               kernel::Tensor4RW& gradientsOut, int B, int W, int H, int D);
  void compareToImage_WithReduce(const kernel::Tensor4Read& colorInput, const kernel::Tensor5Read& gradientsInput,
                                 const kernel::Tensor4Read& colorReference, kernel::Tensor1RW& differenceOut,
                                 kernel::Tensor1RW& gradientsOut, int B, int W, int H, int D);

  bool renderAdjoint(const kernel::RendererInputs& inputs, const kernel::RendererOutputsAsInput& outputsFromForward,
                     const kernel::AdjointColor_t& adj_color, kernel::AdjointOutputs& adj_outputs, int B, int W, int H,
                     kernel::VolumeFilterMode volumeFilterMode, kernel::CameraMode cameraMode, kernel::TFMode tfMode,
                     kernel::BlendMode blendMode, bool hasStepsizeDerivative, bool hasCameraDerivative,
                     bool hasTFDerivative, bool tfDelayedAccumulation, bool hasVolumeDerivative);
  std::string getAdjointKernelName(kernel::VolumeFilterMode volumeFilterMode, kernel::CameraMode cameraMode,
                                   kernel::TFMode tfMode, kernel::BlendMode blendMode, bool hasStepsizeDerivative,
                                   bool hasCameraDerivative, bool hasTFDerivative, bool tfDelayedAccumulation,
                                   bool hasVolumeDerivative);

  // "does this variant exist" -- what unittests/testCudaCompilation.cpp:109-287 finds out by compiling every kernel
  // name: pass 0 forward, 1 forward-mode, 2 adjoint (dvr_has_variant_pass)
  bool hasVariant(int pass, kernel::VolumeFilterMode volumeFilterMode, kernel::CameraMode cameraMode, kernel::TFMode tfMode,
                  kernel::BlendMode blendMode) const;

 private:
  RendererCuda() = default;
  bool done(int rc);
  bool syncMode = false;
  void* stream_ = nullptr;
  std::string lastError_;
};

}  // namespace renderer
