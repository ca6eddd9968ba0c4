// renderer::Renderer over the C-ABI (include/dvr_b200.h). Mirrors the validation of
// renderer/renderer.cpp:49-742 (defined / dtype / device here, shapes + broadcast batch in the
// C-ABI with the reference's wording) and launches on the current torch CUDA stream like
// renderer/renderer_cuda.cpp:105.
#include "renderer.h"

#include <c10/cuda/CUDAFunctions.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>

#include <atomic>
#include <mutex>
#include <stdexcept>

#include "../../../include/dvr_b200.h"

namespace renderer {
namespace {

// wording of renderer/pytorch_utils.h:5-11
#define DVR_CHECK_DEFINED(x, name) TORCH_CHECK((x).defined(), name " is not defined!")
#define DVR_CHECK_CUDA(x, name) \
  TORCH_CHECK((x).device().is_cuda(), "All tensors must be on the same device, but " name " is different, is_cuda=", (x).device().is_cuda())
#define DVR_CHECK_DTYPE(x, name, t) TORCH_CHECK((x).dtype() == (t), name " must be of type " #t ", but is ", (x).dtype())

DvrTensor view(const torch::Tensor& t) {
  DvrTensor v{};
  if (!t.defined()) return v;
  TORCH_CHECK(t.dim() <= DVR_MAX_DIMS, "tensor has too many dimensions");
  v.ptr = t.data_ptr();
  v.ndim = int32_t(t.dim());
  for (int64_t i = 0; i < t.dim(); ++i) {
    v.size[i] = t.size(i);
    v.stride[i] = t.stride(i);
  }
  return v;
}

// float tensor living on the GPU of the volume
DvrTensor real_view(const torch::Tensor& t, const char* name, const torch::Device& dev) {
  TORCH_CHECK(t.defined(), name, " is not defined!");
  TORCH_CHECK(t.device().is_cuda(), "All tensors must be on the same device, but ", name,
              " is different, is_cuda=", t.device().is_cuda());
  TORCH_CHECK(t.device() == dev, "All tensors must be on the same device, but ", name, " is on ", t.device(),
              " and the volume on ", dev);
  TORCH_CHECK(t.dtype() == c10::ScalarType::Float, name, " must be of type c10::ScalarType::Float, but is ", t.dtype());
  return view(t);
}
DvrTensor int_view(const torch::Tensor& t, const char* name, const torch::Device& dev) {
  TORCH_CHECK(t.defined(), name, " is not defined!");
  TORCH_CHECK(t.device() == dev, "All tensors must be on the same device, but ", name, " is different, is_cuda=",
              t.device().is_cuda());
  TORCH_CHECK(t.dtype() == c10::kInt, name, " must be of type c10::kInt, but is ", t.dtype());
  return view(t);
}

// C-ABI error code -> the exception the reference would throw: TORCH_CHECK failures are c10::Error
// (-> RuntimeError in Python), enum / flag / launch problems std::runtime_error.
void raise_on_error(int rc) {
  if (rc == DVR_OK) return;
  const std::string msg = dvr_last_error();
  if (rc == DVR_ERR_INVALID_ARGUMENT) TORCH_CHECK(false, msg);
  throw std::runtime_error(msg);
}

torch::Device device_of_volume(const RendererInputsHost& in) {
  DVR_CHECK_DEFINED(in.volume, "volume");
  TORCH_CHECK(in.volume.device().is_cuda(),
              "volume is a CPU tensor: this build has no CPU back-end (the hot path is CUDA for sm_100a only)");
  return in.volume.device();
}

// Fills the plain-C inputs; every tensor referenced stays alive in `in` for the duration of the call.
DvrInputs to_c(const RendererInputsHost& in, const torch::Device& dev) {
  DvrInputs c{};
  c.screen_width = in.screenSize.x;
  c.screen_height = in.screenSize.y;
  c.volume_filter_mode = int32_t(in.volumeFilterMode);
  c.volume = real_view(in.volume, "volume", dev);
  if (in.boxMin.index() == 0) {
    c.box_min = real_view(std::get<0>(in.boxMin), "boxMin", dev);
  } else {
    const real3 v = std::get<1>(in.boxMin);
    c.box_min_value[0] = v.x; c.box_min_value[1] = v.y; c.box_min_value[2] = v.z;
  }
  if (in.boxSize.index() == 0) {
    c.box_size = real_view(std::get<0>(in.boxSize), "boxSize", dev);
  } else {
    const real3 v = std::get<1>(in.boxSize);
    c.box_size_value[0] = v.x; c.box_size_value[1] = v.y; c.box_size_value[2] = v.z;
  }
  c.camera_mode = int32_t(in.cameraMode);
  switch (in.cameraMode) {
    case kernel::CameraRayStartDir: {
      TORCH_CHECK(in.camera.index() == 0,
                  "'CameraRayStartDir' specified in 'cameraMode', but 'camera' does not hold the first variant 'CameraPerPixelRays'");
      const auto& rays = std::get<0>(in.camera);
      c.camera_ray_start = real_view(rays.cameraRayStart, "cameraRayStart", dev);
      c.camera_ray_dir = real_view(rays.cameraRayDir, "cameraRayDir", dev);
    } break;
    case kernel::CameraInverseViewMatrix: {
      TORCH_CHECK(in.camera.index() > 0 && in.camera.index() < 3,
                  "'CameraInverseViewMatrix' specified in 'cameraMode', but 'camera' holds the first variant 'CameraPerPixelRays'");
      if (in.camera.index() == 1) {
        c.camera_matrix = real_view(std::get<1>(in.camera), "cameraMatrix", dev);
      } else {
        const mat4& m = std::get<2>(in.camera);
        for (int i = 0; i < 4; ++i)
          for (int j = 0; j < 4; ++j) c.camera_matrix_value[i * 4 + j] = m[i][j];
      }
    } break;
    case kernel::CameraReferenceFrame: {
      TORCH_CHECK(in.camera.index() == 3,
                  "'CameraReferenceFrame' specified in 'cameraMode', but 'camera' does not hold the fourth variant 'CameraReferenceFrame'");
      const auto& frame = std::get<3>(in.camera);
      c.camera_matrix = real_view(frame.viewport, "cameraMatrix", dev);
      c.camera_fov_y_radians = frame.fovYRadians;
    } break;
    default: throw std::runtime_error("unknown camera enum value");
  }
  if (in.stepSize.index() == 0) c.step_size = real_view(std::get<0>(in.stepSize), "stepSize", dev);
  else c.step_size_value = std::get<1>(in.stepSize);
  c.tf_mode = int32_t(in.tfMode);
  c.tf = real_view(in.tf, "tf", dev);
  c.blend_mode = int32_t(in.blendMode);
  c.blending_early_out = in.blendingEarlyOut;
  return c;
}

DvrOutputs to_c(const RendererOutputsHost& out, const torch::Device& dev, const char* colorName) {
  DvrOutputs o{};
  o.color = real_view(out.color, colorName, dev);
  o.termination_index = int_view(out.terminationIndex, "terminationIndex", dev);
  return o;
}

DvrForwardDifferencesSettings to_c(const ForwardDifferencesSettingsHost& s, const torch::Device& dev) {
  DvrForwardDifferencesSettings c{};
  c.D = s.D;
  c.d_stepsize = s.d_stepsize;
  c.d_ray_start[0] = s.d_rayStart.x; c.d_ray_start[1] = s.d_rayStart.y; c.d_ray_start[2] = s.d_rayStart.z;
  c.d_ray_dir[0] = s.d_rayDir.x; c.d_ray_dir[1] = s.d_rayDir.y; c.d_ray_dir[2] = s.d_rayDir.z;
  c.has_tf_derivatives = s.hasTfDerivatives ? 1 : 0;
  if (s.d_tf.defined()) {
    c.d_tf = int_view(s.d_tf, "d_tf", dev);
  } else {
    // renderer/renderer.cpp:445-447
    TORCH_CHECK(!s.hasTfDerivatives, "d_tf not defined, hence hasTfDerivatives must be false but was set to true");
  }
  return c;
}

DvrAdjointOutputs to_c(const AdjointOutputsHost& a, const torch::Device& dev) {
  DvrAdjointOutputs c{};
  c.has_volume_derivatives = a.hasVolumeDerivatives;
  if (a.hasVolumeDerivatives) c.adj_volume = real_view(a.adj_volume, "adj_volume", dev);
  c.has_stepsize_derivatives = a.hasStepSizeDerivatives;
  if (a.hasStepSizeDerivatives) c.adj_step_size = real_view(a.adj_stepSize, "adj_stepSize", dev);
  c.has_camera_derivatives = a.hasCameraDerivatives;
  if (a.hasCameraDerivatives) {
    c.adj_camera_ray_start = real_view(a.adj_cameraRayStart, "adj_cameraRayStart", dev);
    c.adj_camera_ray_dir = real_view(a.adj_cameraRayDir, "adj_cameraRayDir", dev);
  }
  c.has_tf_derivatives = a.hasTFDerivatives;
  c.tf_delayed_accumulation = a.tfDelayedAcummulation;
  if (a.hasTFDerivatives) c.adj_tf = real_view(a.adj_tf, "adj_tf", dev);
  return c;
}

void* current_stream(const torch::Device& dev) { return c10::cuda::getCurrentCUDAStream(dev.index()).stream(); }

// ---- host-layer options and the volume cache --------------------------------------------------------------------
// The C-ABI is reentrant (options travel per call); this class is, like the reference's, a static facade with
// process-wide settings (Renderer::setCudaSyncMode & co), so its defaults live here and are passed per call.
std::atomic<int> g_compat_stepsize{0}, g_volume_layout{-1}, g_volume_cache{1}, g_forward_mode_tf{0}, g_march_zero_rays{0};

// One workspace per device, keyed on the volume TENSOR: (data_ptr, sizes, strides) are checked by the library,
// the contents by torch's version counter (every in-place operation bumps it). The cache holds a handle of the
// tensor, so its address cannot be recycled for another tensor while the entry lives. Writes that bypass the
// version counter (raw kernels, .data) need Renderer::setOption("volume_cache", 0) or cleanupCuda().
struct VolumeCache {
  std::mutex m;
  DvrWorkspace* ws = nullptr;
  torch::Tensor volume;
  uint32_t version = 0;
  void* stream = nullptr;
};
constexpr int kMaxDevices = 64;
VolumeCache g_cache[kMaxDevices];

struct CallOptions {
  DvrCallOptions o;
  std::unique_lock<std::mutex> lock;  // held while a cached workspace is in use by this call
};

void fill_options(CallOptions& c, const RendererInputsHost& in, const torch::Device& dev, void* stream) {
  dvr_call_options_init(&c.o);
  c.o.volume_layout = g_volume_layout.load();
  c.o.compat_stepsize_overwrite = g_compat_stepsize.load();
  c.o.forward_mode_tf = g_forward_mode_tf.load();
  c.o.adjoint_march_zero_rays = g_march_zero_rays.load();
  if (!g_volume_cache.load() || c.o.volume_layout == 0 || in.volumeFilterMode != kernel::FilterTrilinear) return;
  const int d = dev.index();
  if (d < 0 || d >= kMaxDevices) return;
  VolumeCache& vc = g_cache[d];
  c.lock = std::unique_lock<std::mutex>(vc.m);
  uint32_t version = 0;
  try {
    version = in.volume._version();
  } catch (const c10::Error&) {  // inference tensors carry no version counter: do not cache them
    c.lock.unlock();
    return;
  }
  if (!vc.ws) vc.ws = dvr_workspace_create();
  if (!vc.ws) {
    c.lock.unlock();
    return;
  }
  const bool same = vc.volume.defined() && vc.volume.is_same(in.volume) && vc.version == version;
  if (same && vc.stream != stream) {  // the copy was built on another stream: stay off it (no cross-stream ordering)
    c.lock.unlock();
    return;
  }
  if (!same) {
    dvr_workspace_invalidate(vc.ws);
    vc.volume = in.volume;
    vc.version = version;
    vc.stream = stream;
  }
  c.o.workspace = vc.ws;
}

void drop_volume_caches() {
  for (VolumeCache& vc : g_cache) {
    std::lock_guard<std::mutex> lock(vc.m);
    if (vc.ws) dvr_workspace_destroy(vc.ws);
    vc.ws = nullptr;
    vc.volume = torch::Tensor();
  }
}

}  // namespace

void Renderer::checkInputs(const RendererInputsHost& inputsHost, int& B, int& W, int& H, int& X, int& Y, int& Z,
                           bool& cuda, bool ignoreVolume) {
  const torch::Device dev = device_of_volume(inputsHost);
  cuda = true;
  const DvrInputs c = to_c(inputsHost, dev);
  int32_t b = 1, x = 0, y = 0, z = 0;
  raise_on_error(dvr_check_inputs(&c, &b, &x, &y, &z));
  W = inputsHost.screenSize.x;
  H = inputsHost.screenSize.y;
  B = b;
  X = ignoreVolume ? 0 : x;
  Y = ignoreVolume ? 0 : y;
  Z = ignoreVolume ? 0 : z;
}

bool Renderer::initCuda() { return c10::cuda::device_count() > 0; }
void Renderer::setCudaCacheDir(const std::filesystem::path&) {}
void Renderer::reloadCudaKernels() {}
void Renderer::disableCudaCache() {}
void Renderer::cleanupCuda() {  // frees what the library holds on the device: volume caches and scratch pool pages
  drop_volume_caches();
  dvr_trim_scratch();
}
void Renderer::setCudaSyncMode(bool sync) { raise_on_error(dvr_set_option("sync_mode", sync ? 1 : 0)); }
bool Renderer::getCudaSyncMode() { return getOption("sync_mode") != 0; }
void Renderer::setCudaDebugMode(bool) {}

void Renderer::setOption(const std::string& name, int value) {
  if (name == "compat_stepsize_overwrite") { g_compat_stepsize = value != 0; return; }
  if (name == "volume_layout") {
    TORCH_CHECK(value >= -1 && value <= 1, "volume_layout must be -1, 0 or 1");
    g_volume_layout = value;
    return;
  }
  if (name == "forward_mode_tf") {
    TORCH_CHECK(value == 0 || value == 1, "forward_mode_tf must be 0 or 1");
    g_forward_mode_tf = value;
    return;
  }
  if (name == "adjoint_march_zero_rays") { g_march_zero_rays = value != 0; return; }
  if (name == "volume_cache") {
    g_volume_cache = value != 0;
    if (!value) drop_volume_caches();
    return;
  }
  raise_on_error(dvr_set_option(name.c_str(), value));
}
int Renderer::getOption(const std::string& name) {
  if (name == "compat_stepsize_overwrite") return g_compat_stepsize.load();
  if (name == "volume_layout") return g_volume_layout.load();
  if (name == "volume_cache") return g_volume_cache.load();
  if (name == "forward_mode_tf") return g_forward_mode_tf.load();
  if (name == "adjoint_march_zero_rays") return g_march_zero_rays.load();
  int32_t v = 0;
  raise_on_error(dvr_get_option(name.c_str(), &v));
  return v;
}
long long Renderer::stat(const std::string& name) {
  int64_t v = 0;
  raise_on_error(dvr_get_stat(name.c_str(), &v));
  return v;
}
long long Renderer::launchCount() { return dvr_launch_count(); }

void Renderer::renderForward(const RendererInputsHost& inputsHost, RendererOutputsHost& outputsHost) {
  const torch::Device dev = device_of_volume(inputsHost);
  const c10::cuda::CUDAGuard guard(dev);
  const DvrInputs in = to_c(inputsHost, dev);
  const DvrOutputs out = to_c(outputsHost, dev, "color");
  void* stream = current_stream(dev);
  CallOptions opt;
  fill_options(opt, inputsHost, dev, stream);
  raise_on_error(dvr_render_forward_ex(&in, &out, &opt.o, stream));
}

void Renderer::renderForwardGradients(const RendererInputsHost& inputsHost,
                                      ForwardDifferencesSettingsHost& differencesSettingsHost,
                                      RendererOutputsHost& outputsHost, torch::Tensor& gradientsOut) {
  const torch::Device dev = device_of_volume(inputsHost);
  const c10::cuda::CUDAGuard guard(dev);
  const DvrInputs in = to_c(inputsHost, dev);
  TORCH_CHECK(differencesSettingsHost.D > 0, "number of derivatives must be positive");
  const DvrForwardDifferencesSettings st = to_c(differencesSettingsHost, dev);
  const DvrOutputs out = to_c(outputsHost, dev, "color");
  const DvrTensor g = real_view(gradientsOut, "gradientsOut", dev);
  void* stream = current_stream(dev);
  CallOptions opt;
  fill_options(opt, inputsHost, dev, stream);
  raise_on_error(dvr_render_forward_gradients_ex(&in, &st, &out, &g, &opt.o, stream));
}

void Renderer::forwardVariablesToGradients(const torch::Tensor& forwardVariables,
                                           const torch::Tensor& gradientOutputColor,
                                           ForwardDifferencesSettingsHost& differencesSettingsHost,
                                           AdjointOutputsHost& adj_outputs) {
  DVR_CHECK_DEFINED(forwardVariables, "forwardVariables");
  TORCH_CHECK(forwardVariables.device().is_cuda(),
              "forwardVariables is a CPU tensor: this build has no CPU back-end");
  const torch::Device dev = forwardVariables.device();
  const c10::cuda::CUDAGuard guard(dev);
  const DvrTensor fv = real_view(forwardVariables, "forwardVariables", dev);
  const DvrTensor gc = real_view(gradientOutputColor, "gradientOutputColor", dev);
  // renderer/renderer.cpp:541-542
  TORCH_CHECK(!adj_outputs.hasVolumeDerivatives, "volume derivatives are not supported with forward variables");
  const DvrForwardDifferencesSettings st = to_c(differencesSettingsHost, dev);
  const DvrAdjointOutputs adj = to_c(adj_outputs, dev);
  raise_on_error(dvr_forward_variables_to_gradients(&fv, &gc, &st, &adj, current_stream(dev)));
}

void Renderer::renderAdjoint(const RendererInputsHost& inputsHost, const RendererOutputsHost& outputsFromForwardHost,
                             const torch::Tensor& adj_color, AdjointOutputsHost& adj_outputs) {
  const torch::Device dev = device_of_volume(inputsHost);
  const c10::cuda::CUDAGuard guard(dev);
  const DvrInputs in = to_c(inputsHost, dev);
  const DvrOutputs fwd = to_c(outputsFromForwardHost, dev, "output_color");
  const DvrTensor ac = real_view(adj_color, "adj_color", dev);
  const DvrAdjointOutputs adj = to_c(adj_outputs, dev);
  if (!adj_outputs.hasTFDerivatives && !adj_outputs.hasCameraDerivatives && !adj_outputs.hasStepSizeDerivatives &&
      !adj_outputs.hasVolumeDerivatives)
    throw std::runtime_error("At least one adjoint variable must be activated!");  // renderer.cpp:711-716
  void* stream = current_stream(dev);
  CallOptions opt;
  fill_options(opt, inputsHost, dev, stream);
  raise_on_error(dvr_render_adjoint_ex(&in, &fwd, &ac, &adj, &opt.o, stream));
}

void Renderer::compareToImage(const torch::Tensor& colorInput, const torch::Tensor& gradientsInput,
                              const torch::Tensor& colorReference, torch::Tensor& differenceOut,
                              torch::Tensor& gradientsOut, bool reduce) {
  DVR_CHECK_DEFINED(colorInput, "colorInput");
  TORCH_CHECK(colorInput.device().is_cuda(), "colorInput is a CPU tensor: this build has no CPU back-end");
  const torch::Device dev = colorInput.device();
  const c10::cuda::CUDAGuard guard(dev);
  const DvrTensor c = real_view(colorInput, "colorInput", dev);
  const DvrTensor g = real_view(gradientsInput, "gradientsInput", dev);
  const DvrTensor r = real_view(colorReference, "colorReference", dev);
  const DvrTensor d = real_view(differenceOut, "differenceOut", dev);
  const DvrTensor o = real_view(gradientsOut, "gradientsOut", dev);
  raise_on_error(dvr_compare_to_image(&c, &g, &r, &d, &o, reduce ? 1 : 0, current_stream(dev)));
}

}  // namespace renderer
