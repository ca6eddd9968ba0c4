// C++ host interface of the B200 ray marcher: the reference's `renderer::Renderer` API
// (renderer/renderer.h:20-269) re-created over the C-ABI of include/dvr_b200.h.
//
// Same class, static methods, host structs and field names as the reference, so code written
// against renderer.h (and, through bindings.cpp, against the `pyrenderer` module) compiles and
// runs unchanged. What is different underneath:
//  * there is no RendererCuda / KernelLoader / NVRTC: every method validates its tensors, fills
//    the plain-C structs and calls libdvr_b200.so on the current torch CUDA stream;
//  * scalar variants (real3 box, scalar stepsize, mat4 camera) are passed by value, not uploaded
//    as 1-element tensors (renderer/renderer.cpp:101-102,117-118,217-218);
//  * there is NO CPU back-end: CPU tensors raise (the reference forks to RendererCpu,
//    renderer/renderer.cpp:405-415);
//  * glm::mat4 (not installed here) is replaced by the layout-compatible POD `mat4`.
#pragma once
#include <torch/types.h>

#include <filesystem>
#include <string>
#include <tuple>
#include <variant>
#include <vector>

#include "kernel_types.h"

namespace renderer {

// Stand-in for glm::mat4: 4 columns of 4 floats, m[i][j] addressed like glm (column i, row j).
// The reference turns it into a 1*4*4 tensor by reinterpreting this memory row-major
// (renderer/renderer.cpp:170-179), i.e. tensor[0][i][j] == m[i][j]; so does this code.
struct mat4 {
  float m[4][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}};
  float* operator[](int i) { return m[i]; }
  const float* operator[](int i) const { return m[i]; }
};

// renderer/renderer.h:20-64
struct RendererInputsHost {
  int2 screenSize{0, 0};  // W*H
  kernel::VolumeFilterMode volumeFilterMode = kernel::FilterTrilinear;
  torch::Tensor volume;  // B*X*Y*Z

  std::variant<torch::Tensor, real3> boxMin;   // (B)*3
  std::variant<torch::Tensor, real3> boxSize;  // (B)*3

  kernel::CameraMode cameraMode = kernel::CameraRayStartDir;
  struct CameraPerPixelRays {
    torch::Tensor cameraRayStart;  // B*H*W*3
    torch::Tensor cameraRayDir;    // B*H*W*3
  };
  typedef torch::Tensor CameraInverseViewMatrixTensor;  // B*4*4
  typedef mat4 CameraInverseViewMatrixScalar;
  struct CameraReferenceFrame {
    torch::Tensor viewport;  // B*3*3: eye, right, up
    real_t fovYRadians = 0;
  };
  std::variant<CameraPerPixelRays, CameraInverseViewMatrixTensor, CameraInverseViewMatrixScalar, CameraReferenceFrame>
      camera;

  std::variant<torch::Tensor, real_t> stepSize;  // (B)*(H)*(W)

  kernel::TFMode tfMode = kernel::TFIdentity;
  torch::Tensor tf;  // (B)*R*C

  kernel::BlendMode blendMode = kernel::BlendBeerLambert;
  real_t blendingEarlyOut = real_t(1) - real_t(1e-5);
};

// renderer/renderer.h:67-71
struct RendererOutputsHost {
  torch::Tensor color;             // B*H*W*4
  torch::Tensor terminationIndex;  // B*H*W int32
};

// renderer/renderer.h:74-93
struct ForwardDifferencesSettingsHost {
  int D = 0;
  static constexpr const int NOT_ENABLED = kernel::ForwardDifferencesSettings::NOT_ENABLED;
  int d_stepsize = NOT_ENABLED;
  int3 d_rayStart{NOT_ENABLED, NOT_ENABLED, NOT_ENABLED};
  int3 d_rayDir{NOT_ENABLED, NOT_ENABLED, NOT_ENABLED};
  torch::Tensor d_tf;  // B*R*C int32
  bool hasTfDerivatives = false;
};

// renderer/renderer.h:95-111 (field names kept, including the reference's spelling)
struct AdjointOutputsHost {
  bool hasVolumeDerivatives = false;
  torch::Tensor adj_volume;  // (B)*X*Y*Z
  bool hasStepSizeDerivatives = false;
  torch::Tensor adj_stepSize;  // (B)*(H)*(W)
  bool hasCameraDerivatives = false;
  torch::Tensor adj_cameraRayStart;  // (B)*H*W*3
  torch::Tensor adj_cameraRayDir;    // (B)*H*W*3
  bool hasTFDerivatives = false;
  bool tfDelayedAcummulation = false;  // accepted; both settings give the same result here
  torch::Tensor adj_tf;                // (B)*R*C
};

// renderer/renderer.h:113-269
class Renderer {
 public:
  // Validates like Renderer::checkInputs (renderer/renderer.cpp:49-263) and returns the broadcast
  // batch size and volume resolution. (The reference returns kernel accessor structs; callers
  // outside the library only use the sizes.)
  static void checkInputs(const RendererInputsHost& inputsHost, int& B, int& W, int& H, int& X, int& Y, int& Z,
                          bool& cuda, bool ignoreVolume = false);

  static bool initCuda();
  static void setCudaCacheDir(const std::filesystem::path& path);  // no-op: nothing is compiled at run time
  static void reloadCudaKernels();                                  // no-op
  static void disableCudaCache();                                   // no-op
  static void cleanupCuda();  // frees the volume caches and the library's scratch pool (nothing else to clean up)
  static void setCudaSyncMode(bool sync);
  static bool getCudaSyncMode();
  static void setCudaDebugMode(bool debug);                         // no-op (kernels carry -lineinfo)

  static void renderForward(const RendererInputsHost& inputsHost, RendererOutputsHost& outputsHost);
  static void renderForwardGradients(const RendererInputsHost& inputsHost,
                                     ForwardDifferencesSettingsHost& differencesSettingsHost,
                                     RendererOutputsHost& outputsHost, torch::Tensor& gradientsOut);
  static void forwardVariablesToGradients(const torch::Tensor& forwardVariables,
                                          const torch::Tensor& gradientOutputColor,
                                          ForwardDifferencesSettingsHost& differencesSettingsHost,
                                          AdjointOutputsHost& adj_outputs);
  static void renderAdjoint(const RendererInputsHost& inputsHost, const RendererOutputsHost& outputsFromForwardHost,
                            const torch::Tensor& adj_color, AdjointOutputsHost& adj_outputs);

  // renderer/renderer.h:228-250: L2 norm between colorInput and colorReference with its D forward
  // derivatives, per pixel (differenceOut B*H*W*1, gradientsOut B*H*W*D) or averaged over the image
  // (shapes 1 and D). Any D (the reference's CUDA path only runs D == 1).
  static void compareToImage(const torch::Tensor& colorInput, const torch::Tensor& gradientsInput,
                             const torch::Tensor& colorReference, torch::Tensor& differenceOut,
                             torch::Tensor& gradientsOut, bool reduce = false);

  // extensions (not in the reference). Options of this static facade, passed to the C-ABI per call (DvrCallOptions):
  //   "compat_stepsize_overwrite" 0/1, "volume_layout" -1/0/1 (include/dvr_b200.h),
  //   "volume_cache" 1 (default): one QUAD copy per device cached on the volume tensor and torch's version counter, so
  //   forward, forward-mode and adjoint calls on an unchanged volume share it; 0 disables and frees the cache.
  // "sync_mode" is the library's own process-wide switch.
  static void setOption(const std::string& name, int value);
  static int getOption(const std::string& name);
  static long long launchCount();
  static long long stat(const std::string& name);  // dvr_get_stat: "quad_builds", "quad_reuses", ...
};

}  // namespace renderer
