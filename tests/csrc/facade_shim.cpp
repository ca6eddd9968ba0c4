// Test shim for diffdvr_b200/csrc/host/renderer_cuda_facade.{h,cpp}: plays the part of the reference's host layer
// (renderer/renderer.cpp:405-415, 722-739), i.e. fills kernel::RendererInputs & co -- the reference's packed 32-bit
// accessors -- and calls renderer::RendererCuda. The tests hand the tensors over as the C-ABI's DvrTensor structs
// (ctypes has those already), so every call is DvrTensor -> accessor (here) -> DvrTensor (facade) -> dvr_*.
#include <cstddef>
#include <cstring>
#include <string>

#include "../../diffdvr_b200/csrc/host/renderer_cuda_facade.h"

#define SHIM_API extern "C" __attribute__((visibility("default")))

namespace {

using namespace kernel;

// the layout the facade relies on: renderer/renderer_tensor.cuh:122-181 with index_t = int32
static_assert(offsetof(Tensor4Read, data_) == 0 && offsetof(Tensor4Read, sizes_) == 8 &&
                  offsetof(Tensor4Read, strides_) == 8 + 4 * 4 && sizeof(Tensor4Read) == 40,
              "packed accessor layout");
static_assert(sizeof(Tensor3Read) == 32 && sizeof(Tensor2Read) == 24 && sizeof(Tensor5RW) == 48, "packed accessor layout");

template <typename A, int N>
A acc(const DvrTensor& t, bool broadcast = false) {
  typedef decltype(A().data()) P;
  if (!t.ptr) return A();
  return A(static_cast<P>(t.ptr), t.size, t.stride, broadcast);
}

RendererInputs inputs(const DvrInputs& in) {
  RendererInputs r;
  r.screenSize = int2{in.screen_width, in.screen_height};
  r.volume = acc<Tensor4Read, 4>(in.volume, true);
  r.volumeSize = int3{int(in.volume.size[1]), int(in.volume.size[2]), int(in.volume.size[3])};  // B|4 * X*Y*Z
  r.boxMin = acc<Tensor2Read, 2>(in.box_min, true);
  r.boxSize = acc<Tensor2Read, 2>(in.box_size, true);
  r.cameraRayStart = acc<Tensor4Read, 4>(in.camera_ray_start, true);
  r.cameraRayDir = acc<Tensor4Read, 4>(in.camera_ray_dir, true);
  r.cameraMatrix = acc<Tensor3Read, 3>(in.camera_matrix, true);
  r.cameraFovYRadians = in.camera_fov_y_radians;
  r.stepSize = acc<Tensor3Read, 3>(in.step_size, true);
  r.tf = acc<Tensor3Read, 3>(in.tf, true);
  r.blendingEarlyOut = in.blending_early_out;
  return r;
}

ForwardDifferencesSettingsFull settings(const DvrForwardDifferencesSettings& s) {
  ForwardDifferencesSettingsFull r;
  r.d_stepsize = s.d_stepsize;
  r.d_rayStart = int3{s.d_ray_start[0], s.d_ray_start[1], s.d_ray_start[2]};
  r.d_rayDir = int3{s.d_ray_dir[0], s.d_ray_dir[1], s.d_ray_dir[2]};
  r.hasTFDerivatives = s.has_tf_derivatives != 0;
  r.d_tf = acc<ITensor3Read, 3>(s.d_tf, true);
  return r;
}

AdjointOutputs adjoint_outputs(const DvrAdjointOutputs& a) {
  AdjointOutputs r;
  r.adj_volume = acc<BTensor4RW, 4>(a.adj_volume, true);
  r.adj_stepSize = acc<BTensor3RW, 3>(a.adj_step_size, true);
  r.adj_cameraRayStart = acc<BTensor4RW, 4>(a.adj_camera_ray_start, true);
  r.adj_cameraRayDir = acc<BTensor4RW, 4>(a.adj_camera_ray_dir, true);
  r.adj_tf = acc<BTensor3RW, 3>(a.adj_tf, true);
  return r;
}

VolumeFilterMode F(const DvrInputs& in) { return VolumeFilterMode(in.volume_filter_mode); }
CameraMode C(const DvrInputs& in) { return CameraMode(in.camera_mode); }
TFMode T(const DvrInputs& in) { return TFMode(in.tf_mode); }
BlendMode Bl(const DvrInputs& in) { return BlendMode(in.blend_mode); }

renderer::RendererCuda& R(void* stream) {
  renderer::RendererCuda& r = renderer::RendererCuda::Instance();
  r.setStream(stream);
  return r;
}

int copy_out(const std::string& s, char* buf, int n) {
  if (n <= 0) return int(s.size());
  std::strncpy(buf, s.c_str(), size_t(n) - 1);
  buf[n - 1] = 0;
  return int(s.size());
}

}  // namespace

// sizeof the aggregate mirrors, for the layout test: RendererInputs, RendererOutputs, AdjointOutputs
SHIM_API void shim_sizes(int32_t* out) {
  out[0] = int32_t(sizeof(RendererInputs));
  out[1] = int32_t(sizeof(RendererOutputs));
  out[2] = int32_t(sizeof(AdjointOutputs));
  out[3] = int32_t(offsetof(RendererInputs, volume));
  out[4] = int32_t(offsetof(RendererInputs, tf));
  out[5] = int32_t(offsetof(AdjointOutputs, adj_tf));
}

// DvrInputs -> reference accessors -> DvrInputs, no device access
SHIM_API void shim_inputs_roundtrip(const DvrInputs* in, DvrInputs* out) {
  *out = renderer::facade::to_c(inputs(*in), F(*in), C(*in), T(*in), Bl(*in));
}
SHIM_API void shim_adjoint_outputs_roundtrip(const DvrAdjointOutputs* in, DvrAdjointOutputs* out) {
  *out = renderer::facade::to_c(adjoint_outputs(*in), in->has_stepsize_derivatives, in->has_camera_derivatives,
                                in->has_tf_derivatives, in->tf_delayed_accumulation, in->has_volume_derivatives);
}
SHIM_API void shim_settings_roundtrip(const DvrForwardDifferencesSettings* in, DvrForwardDifferencesSettings* out) {
  *out = renderer::facade::to_c(settings(*in), in->D);
}

SHIM_API int shim_init() { return renderer::RendererCuda::Instance().initCuda() ? 1 : 0; }
SHIM_API const char* shim_last_error() { return renderer::RendererCuda::Instance().lastError().c_str(); }
SHIM_API void shim_set_sync_mode(int sync) { renderer::RendererCuda::Instance().setSyncMode(sync != 0); }
SHIM_API int shim_get_sync_mode() { return renderer::RendererCuda::Instance().getSyncMode() ? 1 : 0; }
SHIM_API void shim_cleanup() { renderer::RendererCuda::Instance().cleanup(); }

SHIM_API int shim_has_variant(int pass, int f, int c, int t, int b) {
  return renderer::RendererCuda::Instance().hasVariant(pass, VolumeFilterMode(f), CameraMode(c), TFMode(t), BlendMode(b)) ? 1
                                                                                                                          : 0;
}

// pass 0: forward, 1: forward-mode (D, hasStep, hasCam, hasTf), 2: adjoint (hasStep, hasCam, hasTf, delayed, hasVol),
// 3: forwardVariablesToGradients
SHIM_API int shim_kernel_name(int pass, int f, int c, int t, int b, int D, int hasStep, int hasCam, int hasTf, int delayed,
                              int hasVol, char* buf, int n) {
  renderer::RendererCuda& r = renderer::RendererCuda::Instance();
  const VolumeFilterMode vf = VolumeFilterMode(f);
  const CameraMode cm = CameraMode(c);
  const TFMode tm = TFMode(t);
  const BlendMode bm = BlendMode(b);
  switch (pass) {
    case 0: return copy_out(r.getForwardKernelName(vf, cm, tm, bm), buf, n);
    case 1: return copy_out(r.getForwardGradientsKernelName(vf, cm, tm, bm, D, hasStep, hasCam, hasTf), buf, n);
    case 2: return copy_out(r.getAdjointKernelName(vf, cm, tm, bm, hasStep, hasCam, hasTf, delayed, hasVol), buf, n);
    default: return copy_out(r.getForwardVariablesToGradientsName(), buf, n);
  }
}

SHIM_API int shim_render_forward(const DvrInputs* in, const DvrOutputs* out, void* stream) {
  const RendererInputs ri = inputs(*in);
  RendererOutputs ro{acc<Tensor4RW, 4>(out->color), acc<ITensor3RW, 3>(out->termination_index)};
  const int B = int(out->color.size[0]), H = int(out->color.size[1]), W = int(out->color.size[2]);
  return R(stream).renderForward(ri, ro, B, W, H, F(*in), C(*in), T(*in), Bl(*in)) ? 1 : 0;
}

SHIM_API int shim_render_forward_gradients(const DvrInputs* in, const DvrForwardDifferencesSettings* s, const DvrOutputs* out,
                                           const DvrTensor* gradients, void* stream) {
  const RendererInputs ri = inputs(*in);
  const ForwardDifferencesSettingsFull st = settings(*s);
  RendererOutputs ro{acc<Tensor4RW, 4>(out->color), acc<ITensor3RW, 3>(out->termination_index)};
  ForwardDifferencesOutput go{acc<Tensor5RW, 5>(*gradients)};
  const int B = int(out->color.size[0]), H = int(out->color.size[1]), W = int(out->color.size[2]);
  const bool hasCam = s->d_ray_start[0] >= 0 || s->d_ray_start[1] >= 0 || s->d_ray_start[2] >= 0 || s->d_ray_dir[0] >= 0 ||
                      s->d_ray_dir[1] >= 0 || s->d_ray_dir[2] >= 0;
  return R(stream).renderForwardGradients(ri, st, ro, go, B, W, H, F(*in), C(*in), T(*in), Bl(*in), s->D, s->d_stepsize >= 0,
                                          hasCam, s->has_tf_derivatives != 0)
             ? 1
             : 0;
}

SHIM_API int shim_forward_variables_to_gradients(const DvrTensor* forwardVariables, const DvrTensor* gradientOutputColor,
                                                 const DvrForwardDifferencesSettings* s, const DvrAdjointOutputs* adj,
                                                 void* stream) {
  ForwardDifferencesOutput fv{acc<Tensor5RW, 5>(*forwardVariables)};
  const AdjointColor_t gc = acc<Tensor4Read, 4>(*gradientOutputColor);
  const ForwardDifferencesSettingsFull st = settings(*s);
  AdjointOutputs ao = adjoint_outputs(*adj);
  const int B = int(gradientOutputColor->size[0]), H = int(gradientOutputColor->size[1]), W = int(gradientOutputColor->size[2]);
  return R(stream).forwardVariablesToGradients(fv, gc, st, ao, B, W, H) ? 1 : 0;
}

SHIM_API int shim_render_adjoint(const DvrInputs* in, const DvrOutputs* fromForward, const DvrTensor* adjColor,
                                 const DvrAdjointOutputs* adj, void* stream) {
  const RendererInputs ri = inputs(*in);
  const RendererOutputsAsInput ro{acc<Tensor4Read, 4>(fromForward->color), acc<ITensor3Read, 3>(fromForward->termination_index)};
  const AdjointColor_t ac = acc<Tensor4Read, 4>(*adjColor);
  AdjointOutputs ao = adjoint_outputs(*adj);
  const int B = int(adjColor->size[0]), H = int(adjColor->size[1]), W = int(adjColor->size[2]);
  return R(stream).renderAdjoint(ri, ro, ac, ao, B, W, H, F(*in), C(*in), T(*in), Bl(*in), adj->has_stepsize_derivatives != 0,
                                 adj->has_camera_derivatives != 0, adj->has_tf_derivatives != 0,
                                 adj->tf_delayed_accumulation != 0, adj->has_volume_derivatives != 0)
             ? 1
             : 0;
}
