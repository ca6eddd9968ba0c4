// renderer::RendererCuda over the C-ABI: see renderer_cuda_facade.h. Every method converts the reference's accessor
// structs into the plain-C structs of include/dvr_b200.h and forwards; no CUDA call is made here.
#include "renderer_cuda_facade.h"

#include <sstream>

namespace renderer {
namespace {

template <typename A>
DvrTensor view(const A& a, int n) {
  DvrTensor t{};
  t.ptr = const_cast<void*>(static_cast<const void*>(a.data()));
  t.ndim = n;
  for (int i = 0; i < n; ++i) {
    t.size[i] = a.size(i);
    t.stride[i] = a.stride(i);
  }
  return t;
}

}  // namespace

namespace facade {

// kernel::RendererInputs (renderer_settings.cuh:150-180) + the four mode enums -> DvrInputs. The reference always
// passes tensors here (its host layer uploads scalar variants first, renderer.cpp:101-118,217-218), so no *_value
// field is used except for an undefined accessor (data == nullptr), which the C-ABI then reports.
DvrInputs to_c(const kernel::RendererInputs& in, kernel::VolumeFilterMode f, kernel::CameraMode c, kernel::TFMode t,
               kernel::BlendMode b) {
  DvrInputs o{};
  o.screen_width = in.screenSize.x;
  o.screen_height = in.screenSize.y;
  o.volume_filter_mode = int32_t(f);
  o.volume = view(in.volume, 4);
  o.box_min = view(in.boxMin, 2);
  o.box_size = view(in.boxSize, 2);
  o.camera_mode = int32_t(c);
  if (c == kernel::CameraRayStartDir) {
    o.camera_ray_start = view(in.cameraRayStart, 4);
    o.camera_ray_dir = view(in.cameraRayDir, 4);
  } else {
    o.camera_matrix = view(in.cameraMatrix, 3);
    o.camera_fov_y_radians = in.cameraFovYRadians;
  }
  o.step_size = view(in.stepSize, 3);
  o.tf_mode = int32_t(t);
  o.tf = view(in.tf, 3);
  o.blend_mode = int32_t(b);
  o.blending_early_out = in.blendingEarlyOut;
  return o;
}

DvrOutputs to_c(const kernel::RendererOutputs& out) {
  DvrOutputs o{};
  o.color = view(out.color, 4);
  o.termination_index = view(out.terminationIndex, 3);
  return o;
}
DvrOutputs to_c(const kernel::RendererOutputsAsInput& out) {
  DvrOutputs o{};
  o.color = view(out.color, 4);
  o.termination_index = view(out.terminationIndex, 3);
  return o;
}

DvrForwardDifferencesSettings to_c(const DVR_FACADE_FD_SETTINGS& s, int D) {
  DvrForwardDifferencesSettings o{};
  o.D = D;
  o.d_stepsize = s.d_stepsize;
  o.d_ray_start[0] = s.d_rayStart.x; o.d_ray_start[1] = s.d_rayStart.y; o.d_ray_start[2] = s.d_rayStart.z;
  o.d_ray_dir[0] = s.d_rayDir.x; o.d_ray_dir[1] = s.d_rayDir.y; o.d_ray_dir[2] = s.d_rayDir.z;
  o.has_tf_derivatives = s.hasTFDerivatives ? 1 : 0;
  if (s.hasTFDerivatives) o.d_tf = view(s.d_tf, 3);
  return o;
}

DvrAdjointOutputs to_c(const kernel::AdjointOutputs& a, bool hasStep, bool hasCam, bool hasTf, bool delayed, bool hasVol) {
  DvrAdjointOutputs o{};
  o.has_volume_derivatives = hasVol;
  if (hasVol) o.adj_volume = view(a.adj_volume, 4);
  o.has_stepsize_derivatives = hasStep;
  if (hasStep) o.adj_step_size = view(a.adj_stepSize, 3);
  o.has_camera_derivatives = hasCam;
  if (hasCam) {
    o.adj_camera_ray_start = view(a.adj_cameraRayStart, 4);
    o.adj_camera_ray_dir = view(a.adj_cameraRayDir, 4);
  }
  o.has_tf_derivatives = hasTf;
  o.tf_delayed_accumulation = delayed;
  if (hasTf) o.adj_tf = view(a.adj_tf, 3);
  return o;
}

}  // namespace facade

namespace {
// names as the reference spells them through magic_enum (renderer_cuda.cpp:64-83)
const char* name(kernel::VolumeFilterMode m) {
  switch (m) {
    case kernel::FilterNearest: return "FilterNearest";
    case kernel::FilterTrilinear: return "FilterTrilinear";
    case kernel::FilterPreshaded: return "FilterPreshaded";
    case kernel::FilterNetwork: return "FilterNetwork";
    default: return "?";
  }
}
const char* name(kernel::CameraMode m) {
  switch (m) {
    case kernel::CameraRayStartDir: return "CameraRayStartDir";
    case kernel::CameraInverseViewMatrix: return "CameraInverseViewMatrix";
    case kernel::CameraReferenceFrame: return "CameraReferenceFrame";
    default: return "?";
  }
}
const char* name(kernel::TFMode m) {
  switch (m) {
    case kernel::TFIdentity: return "TFIdentity";
    case kernel::TFTexture: return "TFTexture";
    case kernel::TFLinear: return "TFLinear";
    case kernel::TFGaussian: return "TFGaussian";
    case kernel::TFPreshaded: return "TFPreshaded";
    case kernel::TFGaussianLog: return "TFGaussianLog";
    default: return "?";
  }
}
const char* name(kernel::BlendMode m) { return m == kernel::BlendBeerLambert ? "BlendBeerLambert" : "BlendAlpha"; }

std::string modes(kernel::VolumeFilterMode f, kernel::CameraMode c, kernel::TFMode t, kernel::BlendMode b) {
  std::stringstream ss;
  ss << "kernel::VolumeFilterMode::" << name(f) << ", kernel::CameraMode::" << name(c) << ", kernel::TFMode::" << name(t)
     << ", kernel::BlendMode::" << name(b);
  return ss.str();
}

}  // namespace

using namespace facade;

RendererCuda& RendererCuda::Instance() {
  static RendererCuda INSTANCE;
  return INSTANCE;
}

bool RendererCuda::done(int rc) {
  if (rc == DVR_OK) return true;
  lastError_ = dvr_last_error();
  return false;
}

bool RendererCuda::initCuda() { return dvr_abi_version() == DVR_ABI_VERSION; }
void RendererCuda::setCudaCacheDir(const std::filesystem::path&) {}
void RendererCuda::disableCudaCache() {}
void RendererCuda::reloadCudaKernels() {}
void RendererCuda::cleanup() { dvr_trim_scratch(); }
void RendererCuda::setSyncMode(bool sync) {
  syncMode = sync;
  dvr_set_option("sync_mode", sync ? 1 : 0);
}
bool RendererCuda::getSyncMode() const { return syncMode; }
void RendererCuda::setDebugMode(bool) {}

bool RendererCuda::hasVariant(int pass, kernel::VolumeFilterMode f, kernel::CameraMode c, kernel::TFMode t,
                              kernel::BlendMode b) const {
  return dvr_has_variant_pass(pass, int(f), int(c), int(t), int(b)) != 0;
}

bool RendererCuda::renderForward(const kernel::RendererInputs& inputs, kernel::RendererOutputs& outputs, int, int, int,
                                 kernel::VolumeFilterMode f, kernel::CameraMode c, kernel::TFMode t, kernel::BlendMode b) {
  const DvrInputs in = to_c(inputs, f, c, t, b);
  const DvrOutputs out = to_c(outputs);
  return done(dvr_render_forward(&in, &out, stream_));
}

std::string RendererCuda::getForwardKernelName(kernel::VolumeFilterMode f, kernel::CameraMode c, kernel::TFMode t,
                                               kernel::BlendMode b) {
  return "kernel::DvrKernelForwardDevice<" + modes(f, c, t, b) + ">";
}

bool RendererCuda::renderForwardGradients(const kernel::RendererInputs& inputs, const DVR_FACADE_FD_SETTINGS& settings,
                                          kernel::RendererOutputs& outputs, kernel::ForwardDifferencesOutput& gradients, int,
                                          int, int, kernel::VolumeFilterMode f, kernel::CameraMode c, kernel::TFMode t,
                                          kernel::BlendMode b, int numDerivatives, bool, bool, bool) {
  const DvrInputs in = to_c(inputs, f, c, t, b);
  const DvrForwardDifferencesSettings st = to_c(settings, numDerivatives);
  const DvrOutputs out = to_c(outputs);
  const DvrTensor g = view(gradients.gradients, 5);
  return done(dvr_render_forward_gradients(&in, &st, &out, &g, stream_));
}

std::string RendererCuda::getForwardGradientsKernelName(kernel::VolumeFilterMode f, kernel::CameraMode c, kernel::TFMode t,
                                                        kernel::BlendMode b, int numDerivatives, bool hasStepsizeDerivative,
                                                        bool hasCameraDerivative, bool hasTFDerivative) {
  std::stringstream ss;  // renderer_cuda.cpp:120-147
  ss << "kernel::DvrKernelForwardGradientsDevice<" << modes(f, c, t, b) << ", " << numDerivatives << ", "
     << (hasStepsizeDerivative ? "true" : "false") << ", " << (hasCameraDerivative ? "true" : "false") << ", "
     << (hasTFDerivative ? "true" : "false") << ">";
  return ss.str();
}

bool RendererCuda::forwardVariablesToGradients(const kernel::ForwardDifferencesOutput& forwardVariables,
                                               const kernel::AdjointColor_t& gradientOutputColor,
                                               const DVR_FACADE_FD_SETTINGS& s, kernel::AdjointOutputs& adj, int, int, int) {
  const DvrTensor fv = view(forwardVariables.gradients, 5);
  const DvrTensor gc = view(gradientOutputColor, 4);
  const DvrForwardDifferencesSettings st = to_c(s, forwardVariables.gradients.size(3));
  const bool hasStep = s.d_stepsize >= 0;
  const bool hasCam = s.d_rayStart.x >= 0 || s.d_rayStart.y >= 0 || s.d_rayStart.z >= 0 || s.d_rayDir.x >= 0 ||
                      s.d_rayDir.y >= 0 || s.d_rayDir.z >= 0;
  const DvrAdjointOutputs ao = to_c(adj, hasStep, hasCam, s.hasTFDerivatives, false, false);
  return done(dvr_forward_variables_to_gradients(&fv, &gc, &st, &ao, stream_));
}

std::string RendererCuda::getForwardVariablesToGradientsName() { return "kernel::ForwardVariablesToGradientsDevice"; }

void RendererCuda::compareToImage_NoReduce(const kernel::Tensor4Read& colorInput, const kernel::Tensor5Read& gradientsInput,
                                           const kernel::Tensor4Read& colorReference, kernel::Tensor4RW& differenceOut,
                                           kernel::Tensor4RW& gradientsOut, int, int, int, int) {
  const DvrTensor c = view(colorInput, 4), g = view(gradientsInput, 5), r = view(colorReference, 4),
                  d = view(differenceOut, 4), o = view(gradientsOut, 4);
  done(dvr_compare_to_image(&c, &g, &r, &d, &o, 0, stream_));
}

void RendererCuda::compareToImage_WithReduce(const kernel::Tensor4Read& colorInput, const kernel::Tensor5Read& gradientsInput,
                                             const kernel::Tensor4Read& colorReference, kernel::Tensor1RW& differenceOut,
                                             kernel::Tensor1RW& gradientsOut, int, int, int, int) {
  const DvrTensor c = view(colorInput, 4), g = view(gradientsInput, 5), r = view(colorReference, 4),
                  d = view(differenceOut, 1), o = view(gradientsOut, 1);
  done(dvr_compare_to_image(&c, &g, &r, &d, &o, 1, stream_));
}

bool RendererCuda::renderAdjoint(const kernel::RendererInputs& inputs, const kernel::RendererOutputsAsInput& fwd,
                                 const kernel::AdjointColor_t& adj_color, kernel::AdjointOutputs& adj_outputs, int, int, int,
                                 kernel::VolumeFilterMode f, kernel::CameraMode c, kernel::TFMode t, kernel::BlendMode b,
                                 bool hasStepsizeDerivative, bool hasCameraDerivative, bool hasTFDerivative,
                                 bool tfDelayedAccumulation, bool hasVolumeDerivative) {
  const DvrInputs in = to_c(inputs, f, c, t, b);
  const DvrOutputs out = to_c(fwd);
  const DvrTensor ac = view(adj_color, 4);
  const DvrAdjointOutputs ao = to_c(adj_outputs, hasStepsizeDerivative, hasCameraDerivative, hasTFDerivative,
                                    tfDelayedAccumulation, hasVolumeDerivative);
  // the reference's recurrence for the stepsize adjoint, bit for bit: this facade stands in for the reference's kernels
  DvrCallOptions opt;
  dvr_call_options_init(&opt);
  opt.compat_stepsize_overwrite = 1;
  return done(dvr_render_adjoint_ex(&in, &out, &ac, &ao, &opt, stream_));
}

std::string RendererCuda::getAdjointKernelName(kernel::VolumeFilterMode f, kernel::CameraMode c, kernel::TFMode t,
                                               kernel::BlendMode b, bool hasStepsizeDerivative, bool hasCameraDerivative,
                                               bool hasTFDerivative, bool tfDelayedAccumulation, bool hasVolumeDerivative) {
  std::stringstream ss;  // renderer_cuda.cpp:226-258
  ss << "kernel::DvrKernelAdjointDevice<" << modes(f, c, t, b) << ", " << (hasStepsizeDerivative ? "true" : "false") << ", "
     << (hasCameraDerivative ? "true" : "false") << ", " << (hasTFDerivative ? "true" : "false") << ", "
     << (tfDelayedAccumulation ? "true" : "false") << ", " << (hasVolumeDerivative ? "true" : "false") << ">";
  return ss.str();
}

}  // namespace renderer
