// pybind11 module `pyrenderer`: the renderer surface of the reference's bindings
// (bindings/bindings.cpp:132-232, 306-511, 554-558) over renderer::Renderer of renderer.h.
// Same module name, class names, attribute names and enum value names, so
// pytests/diffdvr/renderer.py and the experiment scripts that only use the renderer run unchanged.
// Side utilities of the reference module (Volume, Camera, TFUtils, losses) are not part of the hot
// path.
#include <pybind11/stl.h>
#include <torch/extension.h>

#include <cstdio>
#include <sstream>
#include <stdexcept>

#include "../../../include/dvr_b200.h"
#include "renderer.h"
#include "camera.h"
#include "tf_utils.h"
#include "volume.h"

namespace py = pybind11;
using namespace renderer;

namespace {

std::string fmt(const char* f, double a, double b, double c = 0, double d = 0) {
  char buf[160];
  std::snprintf(buf, sizeof(buf), f, a, b, c, d);
  return buf;
}

// GPUTimer of bindings/bindings.cpp:99-129 (events recorded on the legacy default stream, as there)
class GPUTimer {
  DvrTimer* t_;

 public:
  GPUTimer() : t_(dvr_timer_create()) {
    if (!t_) throw std::runtime_error(dvr_last_error());
  }
  ~GPUTimer() { dvr_timer_destroy(t_); }
  GPUTimer(const GPUTimer&) = delete;
  GPUTimer& operator=(const GPUTimer&) = delete;
  void start() {
    if (dvr_timer_start(t_, nullptr) != DVR_OK) throw std::runtime_error(dvr_last_error());
  }
  void stop() {
    if (dvr_timer_stop(t_, nullptr) != DVR_OK) throw std::runtime_error(dvr_last_error());
  }
  float getElapsedMilliseconds() {
    float ms = 0;
    if (dvr_timer_elapsed_ms(t_, &ms) != DVR_OK) throw std::runtime_error(dvr_last_error());
    return ms;
  }
};

}  // namespace

PYBIND11_MODULE(pyrenderer, m) {
  // the structs this module passes by pointer are those of the header it was compiled against
  if (dvr_abi_version() != DVR_ABI_VERSION)
    throw std::runtime_error("pyrenderer was built against ABI version " + std::to_string(DVR_ABI_VERSION) +
                             " of libdvr_b200.so, the loaded library has version " + std::to_string(dvr_abi_version()) +
                             ": rebuild (diffdvr_b200/csrc/host/build_pyrenderer.py)");
  Renderer::initCuda();
  m.doc() = "python bindings for the differentiable volume renderer (B200-native build)";

  // ---- control (bindings.cpp:143-165, 306) ----------------------------------------------------
  m.def("use_double_precision", []() { return 0; },
        py::doc("Checks if the library was compiled with double-precision (returns 1) or single-precision (returns 0)."));
  m.def("reload_kernels", &Renderer::reloadCudaKernels, py::doc("no-op: kernels are precompiled sm_100a instances"));
  m.def("set_cuda_cache_dir", [](const std::string& path) { Renderer::setCudaCacheDir(path); });
  m.def("disable_cuda_cache", &Renderer::disableCudaCache);
  m.def("set_cuda_sync_mode", &Renderer::setCudaSyncMode);
  m.def("get_cuda_sync_mode", &Renderer::getCudaSyncMode);
  m.def("set_cuda_debug_mode", &Renderer::setCudaDebugMode);
  m.def("set_kernel_cache_file", [](const std::string&) {}, py::doc("no-op: there is no kernel cache"));
  m.def("cleanup", &Renderer::cleanupCuda, py::doc("Explicit cleanup of all CUDA references"));
  m.def("sync", []() {
    if (dvr_device_synchronize() != DVR_OK) throw std::runtime_error(dvr_last_error());
  });
  // extensions of this build
  m.def("set_option", &Renderer::setOption, py::arg("name"), py::arg("value"),
        py::doc("library option of include/dvr_b200.h, e.g. 'compat_stepsize_overwrite'"));
  m.def("get_option", &Renderer::getOption, py::arg("name"));
  m.def("launch_count", &Renderer::launchCount, py::doc("number of CUDA kernels launched by the library so far"));
  m.def("stat", &Renderer::stat, py::arg("name"),
        py::doc("library counters since load: launches, quad_builds, quad_reuses, quad_fallbacks, folds"));
  m.def("has_variant_pass", [](int pass, kernel::VolumeFilterMode f, kernel::CameraMode c, kernel::TFMode t, kernel::BlendMode b) {
    return dvr_has_variant_pass(pass, int(f), int(c), int(t), int(b)) != 0;
  }, py::arg("pass_id"), py::arg("filter"), py::arg("camera"), py::arg("tf"), py::arg("blend"),
        py::doc("pass_id: 0 forward, 1 forward-mode, 2 adjoint"));
  m.def("has_variant", [](kernel::VolumeFilterMode f, kernel::CameraMode c, kernel::TFMode t, kernel::BlendMode b) {
    return dvr_has_variant(int(f), int(c), int(t), int(b)) != 0;
  }, py::doc("True iff a precompiled kernel instance exists for this combination"));

  // ---- small vector types (bindings.cpp:168-213) ------------------------------------------------
  py::class_<real3>(m, "real3")
      .def(py::init<>())
      .def(py::init([](real_t x, real_t y, real_t z) { return make_real3(x, y, z); }))
      .def_readwrite("x", &real3::x)
      .def_readwrite("y", &real3::y)
      .def_readwrite("z", &real3::z)
      .def("__str__", [](const real3& v) { return fmt("(%f, %f, %f)", v.x, v.y, v.z); });
  py::class_<real4>(m, "real4")
      .def(py::init<>())
      .def(py::init([](real_t x, real_t y, real_t z, real_t w) { return make_real4(x, y, z, w); }))
      .def_readwrite("x", &real4::x)
      .def_readwrite("y", &real4::y)
      .def_readwrite("z", &real4::z)
      .def_readwrite("w", &real4::w)
      .def("__str__", [](const real4& v) { return fmt("(%f, %f, %f, %f)", v.x, v.y, v.z, v.w); });
  py::class_<int2>(m, "int2")
      .def(py::init<>())
      .def(py::init([](int x, int y) { return int2{x, y}; }))
      .def_readwrite("x", &int2::x)
      .def_readwrite("y", &int2::y)
      .def("__str__", [](const int2& v) { return fmt("(%.0f, %.0f)", v.x, v.y); });
  py::class_<int3>(m, "int3")
      .def(py::init<>())
      .def(py::init([](int x, int y, int z) { return int3{x, y, z}; }))
      .def_readwrite("x", &int3::x)
      .def_readwrite("y", &int3::y)
      .def_readwrite("z", &int3::z)
      .def("__str__", [](const int3& v) { return fmt("(%.0f, %.0f, %.0f)", v.x, v.y, v.z); });

  // ---- enums (bindings.cpp:215-232) -------------------------------------------------------------
  py::enum_<kernel::VolumeFilterMode>(m, "VolumeFilterMode")
      .value("Nearest", kernel::FilterNearest)
      .value("Trilinear", kernel::FilterTrilinear)
      .value("Preshaded", kernel::FilterPreshaded);
  py::enum_<kernel::CameraMode>(m, "CameraMode")
      .value("RayStartDir", kernel::CameraRayStartDir)
      .value("InverseViewMatrix", kernel::CameraInverseViewMatrix)
      .value("ReferenceFrame", kernel::CameraReferenceFrame);
  py::enum_<kernel::TFMode>(m, "TFMode")
      .value("Identity", kernel::TFIdentity)
      .value("Texture", kernel::TFTexture)
      .value("Linear", kernel::TFLinear)
      .value("Gaussian", kernel::TFGaussian)
      .value("GaussianLog", kernel::TFGaussianLog)
      .value("Preshaded", kernel::TFPreshaded);
  py::enum_<kernel::BlendMode>(m, "BlendMode")
      .value("BeerLambert", kernel::BlendBeerLambert)
      .value("Alpha", kernel::BlendAlpha);

  // ---- camera variants (bindings.cpp:309-379) ---------------------------------------------------
  py::class_<RendererInputsHost::CameraPerPixelRays>(m, "CameraPerPixelRays")
      .def(py::init<>())
      .def(py::init([](const torch::Tensor& s, const torch::Tensor& d) {
        return RendererInputsHost::CameraPerPixelRays{s, d};
      }))
      .def_readwrite("ray_start", &RendererInputsHost::CameraPerPixelRays::cameraRayStart,
                     py::doc("Tensor of ray starting positions of shape B*H*W*3"))
      .def_readwrite("ray_dir", &RendererInputsHost::CameraPerPixelRays::cameraRayDir,
                     py::doc("Tensor of ray direction of shape B*H*W*3"));
  py::class_<RendererInputsHost::CameraReferenceFrame>(m, "CameraReferenceFrame")
      .def(py::init<>())
      .def(py::init([](const torch::Tensor& s, real_t fovY) {
        return RendererInputsHost::CameraReferenceFrame{s, fovY};
      }))
      .def_readwrite("viewport", &RendererInputsHost::CameraReferenceFrame::viewport,
                     py::doc("Tensor of the viewport matrix of shape B*3*3 (eye, right, up)"))
      .def_readwrite("fov_y_radians", &RendererInputsHost::CameraReferenceFrame::fovYRadians,
                     py::doc("field-of-view for the y-axis in radians"));
  py::class_<mat4>(m, "mat4", py::buffer_protocol())
      .def(py::init([](py::buffer b) {
        py::buffer_info info = b.request();
        if (info.format != py::format_descriptor<float>::format())
          throw std::runtime_error("Incompatible format: expected a float array!");
        if (info.ndim != 2) throw std::runtime_error("Incompatible buffer dimension, expected 2D!");
        if (info.shape[0] != 4 || info.shape[1] != 4)
          throw std::runtime_error("Incompatible buffer sizes! Expected a shape of 4x4");
        mat4 r;
        const char* base = static_cast<const char*>(info.ptr);
        for (int i = 0; i < 4; ++i)
          for (int j = 0; j < 4; ++j)
            r[i][j] = *reinterpret_cast<const float*>(base + info.strides[0] * i + info.strides[1] * j);
        return r;
      }))
      .def("__str__", [](const mat4& v) {
        std::string s = "mat4x4(";
        for (int i = 0; i < 4; ++i)
          s += fmt("(%f, %f, %f, %f)", v[i][0], v[i][1], v[i][2], v[i][3]) + (i < 3 ? ", " : ")");
        return s;
      })
      .def_buffer([](mat4& v) -> py::buffer_info {
        return py::buffer_info(&v.m[0][0], sizeof(float), py::format_descriptor<float>::format(), 2, {4, 4},
                               {sizeof(float) * 4, sizeof(float)});
      })
      .def("to_torch", [](const mat4& v) {
        return torch::from_blob(const_cast<float*>(&v.m[0][0]), {1, 4, 4},
                                at::TensorOptions().device(c10::kCPU).dtype(c10::ScalarType::Float)).clone();
      });

  // ---- inputs / outputs (bindings.cpp:382-449) --------------------------------------------------
  py::class_<RendererInputsHost>(m, "RendererInputs")
      .def(py::init<>())
      .def("clone", [](const RendererInputsHost& args) -> RendererInputsHost { return args; },
           py::doc("performs a shallow clone"))
      .def_readwrite("screen_size", &RendererInputsHost::screenSize, py::doc("the screen size (width, height) of type int2"))
      .def_readwrite("volume", &RendererInputsHost::volume, py::doc("the volume tensor of shape B*X*Y*Z"))
      .def_readwrite("volume_filter_mode", &RendererInputsHost::volumeFilterMode)
      .def_readwrite("box_min", &RendererInputsHost::boxMin, py::doc("real3 or float tensor of shape B*3"))
      .def_readwrite("box_size", &RendererInputsHost::boxSize, py::doc("real3 or float tensor of shape B*3"))
      .def_readwrite("camera_mode", &RendererInputsHost::cameraMode)
      .def_readwrite("camera", &RendererInputsHost::camera,
                     py::doc("CameraPerPixelRays | tensor B*4*4 | mat4 | CameraReferenceFrame, matching camera_mode"))
      .def_readwrite("step_size", &RendererInputsHost::stepSize, py::doc("float or float tensor of shape B*H*W"))
      .def_readwrite("tf_mode", &RendererInputsHost::tfMode)
      .def_readwrite("tf", &RendererInputsHost::tf, py::doc("the transfer function, a float tensor of shape B*R*C"))
      .def_readwrite("blend_mode", &RendererInputsHost::blendMode)
      .def_readwrite("blending_early_out", &RendererInputsHost::blendingEarlyOut, py::doc("Default = 1 - 1e-5"));

  py::class_<RendererOutputsHost>(m, "RendererOutputs")
      .def(py::init<>())
      .def(py::init([](const torch::Tensor& c, const torch::Tensor& t) { return RendererOutputsHost{c, t}; }))
      .def_readwrite("color", &RendererOutputsHost::color)
      .def_readwrite("terminationIndex", &RendererOutputsHost::terminationIndex);

  py::class_<ForwardDifferencesSettingsHost>(m, "ForwardDifferencesSettings")
      .def(py::init<>())
      .def_readwrite("D", &ForwardDifferencesSettingsHost::D, py::doc("The number of derivatives to trace"))
      .def_readwrite("d_stepsize", &ForwardDifferencesSettingsHost::d_stepsize)
      .def_readwrite("d_rayStart", &ForwardDifferencesSettingsHost::d_rayStart)
      .def_readwrite("d_rayDir", &ForwardDifferencesSettingsHost::d_rayDir)
      .def_readwrite("d_tf", &ForwardDifferencesSettingsHost::d_tf, py::doc("shape B*R*C, dtype=int32"))
      .def_readwrite("has_tf_derivatives", &ForwardDifferencesSettingsHost::hasTfDerivatives);

  py::class_<AdjointOutputsHost>(m, "AdjointOutputs")
      .def(py::init<>())
      .def_readwrite("has_volume_derivatives", &AdjointOutputsHost::hasVolumeDerivatives)
      .def_readwrite("adj_volume", &AdjointOutputsHost::adj_volume)
      .def_readwrite("has_stepsize_derivatives", &AdjointOutputsHost::hasStepSizeDerivatives)
      .def_readwrite("adj_stepsize", &AdjointOutputsHost::adj_stepSize)
      .def_readwrite("has_camera_derivatives", &AdjointOutputsHost::hasCameraDerivatives)
      .def_readwrite("adj_camera_ray_start", &AdjointOutputsHost::adj_cameraRayStart)
      .def_readwrite("adj_camera_ray_dir", &AdjointOutputsHost::adj_cameraRayDir)
      .def_readwrite("has_tf_derivatives", &AdjointOutputsHost::hasTFDerivatives)
      .def_readwrite("tf_delayed_accumulation", &AdjointOutputsHost::tfDelayedAcummulation)
      .def_readwrite("adj_tf", &AdjointOutputsHost::adj_tf);

  // ---- the renderer (bindings.cpp:451-511) ------------------------------------------------------
  py::class_<Renderer>(m, "Renderer")
      .def_static("render_forward", &Renderer::renderForward,
                  py::doc("Performs a simple forward rendering without gradients."), py::arg("inputs"), py::arg("outputs"))
      .def_static("render_forward_gradients", &Renderer::renderForwardGradients,
                  py::doc("Forward rendering with forward-mode derivatives; gradients_out has shape B*H*W*D*4."),
                  py::arg("inputs"), py::arg("differences_settings"), py::arg("outputs"), py::arg("gradients_out"))
      .def_static("render_adjoint", &Renderer::renderAdjoint,
                  py::doc("Adjoint/backward rendering; adjoint variables are accumulated in adj_outputs."),
                  py::arg("inputs"), py::arg("outputs_from_forward"), py::arg("adj_color"), py::arg("adj_outputs"))
      .def_static("forward_variables_to_gradients", &Renderer::forwardVariablesToGradients,
                  py::doc("Contracts the forward variables with the adjoint colour into camera/stepsize/TF adjoints."),
                  py::arg("forward_variables"), py::arg("adj_color"), py::arg("differences_settings"),
                  py::arg("adj_outputs"))
      // C++-only in the reference (renderer/renderer.h:228-250); exposed here for scripting and tests
      .def_static("compare_to_image", &Renderer::compareToImage,
                  py::doc("L2 norm |color_input - color_reference| and its forward derivatives, per pixel or reduced."),
                  py::arg("color_input"), py::arg("gradients_input"), py::arg("color_reference"),
                  py::arg("difference_out"), py::arg("gradients_out"), py::arg("reduce") = false);

  // ---- camera (bindings.cpp:514-551) ------------------------------------------------------------
  py::enum_<Camera::Orientation>(m, "Orientation")
      .value("Xp", Camera::Xp)
      .value("Xm", Camera::Xm)
      .value("Yp", Camera::Yp)
      .value("Ym", Camera::Ym)
      .value("Zp", Camera::Zp)
      .value("Zm", Camera::Zm);

  py::class_<Camera>(m, "Camera")
      .def_static("compute_matrix", &Camera::computeInverseViewProjectionMatrix,
                  py::doc("Computes the inverse view-projection matrix. Not differentiable!"), py::arg("origin"),
                  py::arg("look_at"), py::arg("up"), py::arg("fov_degrees"), py::arg("width"), py::arg("height"),
                  py::arg("near_clip") = 0.1f, py::arg("far_clip") = 10.0f)
      .def_static("generate_rays", &Camera::generateRays,
                  py::doc("Per-pixel rays (ray_start, ray_dir: B*H*W*3) from the viewport matrix B*3*3 "
                          "(eye, right, up). Differentiable with respect to the viewport."),
                  py::arg("viewport"), py::arg("fovY"), py::arg("screen_width"), py::arg("screen_height"))
      .def_static("viewport_from_lookat", &Camera::viewportFromLookAt,
                  py::doc("Viewport matrix for generate_rays() from origin, look-at target and up vector."),
                  py::arg("origin"), py::arg("look_at"), py::arg("camera_up"))
      .def_static("viewport_from_sphere", &Camera::viewportFromSphere,
                  py::doc("Viewport matrix for generate_rays() from a position on the sphere around a center."),
                  py::arg("center"), py::arg("yaw_radians"), py::arg("pitch_radians"), py::arg("distance"),
                  py::arg("orientation"));

  // ---- transfer-function utilities (bindings.cpp:560-571) -----------------------------------------
  py::class_<TFPoint>(m, "TFPoint")
      .def(py::init<>())
      .def(py::init([](real_t pos, real4 val) { return TFPoint{pos, val}; }))
      .def_readwrite("pos", &TFPoint::pos)
      .def_readwrite("val", &TFPoint::val);

  py::class_<TFUtils>(m, "TFUtils")
      .def_static("assemble_from_settings", &TFUtils::assembleFromSettings, py::arg("color_values_lab"),
                  py::arg("color_positions"), py::arg("opacity_values"), py::arg("opacity_positions"),
                  py::arg("min_density"), py::arg("max_density"), py::arg("opacity_scaling"),
                  py::arg("purge_zero_opacity_regions") = true)
      .def_static("get_piecewise_tensor", &TFUtils::getPiecewiseTensor)
      .def_static("get_texture_tensor", &TFUtils::getTextureTensor)
      .def_static("preshade_volume", &TFUtils::preshadeVolume,
                  py::doc("Pre-shades a density volume 1*X*Y*Z with the TF into a colour volume 4*X*Y*Z "
                          "(r, g, b, opacity). Differentiable with respect to the density."),
                  py::arg("density"), py::arg("tf"), py::arg("tfMode"));
  m.def("lab_to_rgb", &labToRgb);
  m.def("rgb_to_lab", &rgbToLab);

  // ---- volumes (bindings.cpp:234-304) ------------------------------------------------------------
  py::enum_<Volume::ImplicitEquation>(m, "ImplicitEquation")
      .value("MarschnerLobb", Volume::ImplicitEquation::MARSCHNER_LOBB)
      .value("Cube", Volume::ImplicitEquation::CUBE)
      .value("Sphere", Volume::ImplicitEquation::SPHERE)
      .value("InverseSphere", Volume::ImplicitEquation::INVERSE_SPHERE)
      .value("DingDong", Volume::ImplicitEquation::DING_DONG)
      .value("Endrass", Volume::ImplicitEquation::ENDRASS)
      .value("Barth", Volume::ImplicitEquation::BARTH)
      .value("Heart", Volume::ImplicitEquation::HEART)
      .value("Kleine", Volume::ImplicitEquation::KLEINE)
      .value("Cassini", Volume::ImplicitEquation::CASSINI)
      .value("Steiner", Volume::ImplicitEquation::STEINER)
      .value("CrossCap", Volume::ImplicitEquation::CROSS_CAP)
      .value("Kummer", Volume::ImplicitEquation::KUMMER)
      .value("Blobbly", Volume::ImplicitEquation::BLOBBY)  // sic, bindings.cpp:248
      .value("Tube", Volume::ImplicitEquation::TUBE);
  py::class_<Volume, std::shared_ptr<Volume>>(m, "Volume")
      .def(py::init<std::string>())
      .def("create_mipmap_level", &Volume::createMipmapLevel)
      .def_property_readonly("world_size",
                             [](const Volume& v) { return make_real3(v.worldSizeX(), v.worldSizeY(), v.worldSizeZ()); })
      .def("set_world_size", [](Volume& v, float x, float y, float z) { v.setWorldSize(x, y, z); })
      .def_property_readonly("resolution", &Volume::baseResolution)
      .def_static("create_implicit",
                  [](Volume::ImplicitEquation name, int resolution, py::kwargs kwargs) {
                    std::unordered_map<std::string, float> args;
                    for (const auto& e : kwargs) args.insert({e.first.cast<std::string>(), e.second.cast<float>()});
                    return Volume::createImplicitDataset(resolution, name, args);
                  })
      .def_static("from_numpy",
                  [](py::buffer b) {
                    const py::buffer_info info = b.request();
                    if (info.format != py::format_descriptor<real_t>::format())
                      throw std::runtime_error("Incompatible format: expected a real array!");
                    if (info.ndim != 3) throw std::runtime_error("Incompatible buffer dimension, expected 3!");
                    const long long sizes[] = {info.shape[0], info.shape[1], info.shape[2]};
                    const long long strides[] = {(long long)(info.strides[0] / ssize_t(sizeof(real_t))),
                                                 (long long)(info.strides[1] / ssize_t(sizeof(real_t))),
                                                 (long long)(info.strides[2] / ssize_t(sizeof(real_t)))};
                    return Volume::createFromBuffer(static_cast<const real_t*>(info.ptr), sizes, strides);
                  },
                  py::doc("Create a volume from a 3D numpy array of type 'real'"))
      .def_static("from_tensor", &Volume::createFromTensor,
                  py::doc("Wrap a float tensor of shape X*Y*Z or 1*X*Y*Z (a CUDA tensor is not copied)"))
      .def("copy_to_gpu", [](const Volume& v) { v.copyCpuToGpu(); },
           py::doc("No-op: the device copy is the primary one here"))
      .def("save", [](const Volume& v, const std::string& filename, int compression) { v.save(filename, compression); },
           py::arg("filename"), py::arg("compression") = 0)
      .def("getDataCpu", &Volume::dataCpu,
           py::doc("Returns the density tensor of shape 1*X*Y*Z for the specified layer with the data located on the CPU"),
           py::arg("level"))
      .def("getDataGpu", &Volume::dataGpu,
           py::doc("Returns the density tensor of shape 1*X*Y*Z for the specified layer with the data located on the GPU"),
           py::arg("level"));

  // the .cvol body codec on its own (host code; lets the CPU test-suite pin it against the reference's LZ4)
  m.def("cvol_compress_body", [](py::bytes data, int level) {
    const std::string in = data;
    const std::vector<uint8_t> out = cvol::compressBody(reinterpret_cast<const uint8_t*>(in.data()), in.size(), level);
    return py::bytes(reinterpret_cast<const char*>(out.data()), out.size());
  }, py::arg("data"), py::arg("level"));
  m.def("cvol_decompress_body", [](py::bytes data, size_t size) {
    std::istringstream in{std::string(data)};
    std::string out(size, '\0');
    cvol::decompressBody(in, reinterpret_cast<uint8_t*>(&out[0]), size, "buffer");
    return py::bytes(out);
  }, py::arg("data"), py::arg("size"));

  // ---- timing (bindings.cpp:554-558) ------------------------------------------------------------
  py::class_<GPUTimer>(m, "GpuTimer")
      .def(py::init<>())
      .def("start", &GPUTimer::start)
      .def("stop", &GPUTimer::stop)
      .def("elapsed_ms", &GPUTimer::getElapsedMilliseconds);
}
