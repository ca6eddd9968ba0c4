// TEST INFRASTRUCTURE ONLY -- never linked into the product.
//
// Glue that lets the tests drive the REFERENCE's own host kernels
// (/root/reference/renderer/renderer_kernels.cuh: DvrKernelForwardHost :108-128,
// DvrKernelForwardGradientsHost :281-308, DvrKernelAdjointHost :470-497,
// ForwardVariablesToGradientsHost :580-598) through the same plain-C structs as the
// product's C-ABI (include/dvr_b200.h). Compiled by oracle/Makefile from the reference
// sources where they lie (-I/root/reference/renderer); nothing from the reference is copied.
// All pointers are HOST pointers here. Built twice: REAL=float (symbols dvrref32_*) and
// REAL=double (symbols dvrref64_*, the precision the reference authors ran their unit tests
// in, unittests/main.cpp:22-24).
#pragma once

#include <cstdint>
#include <cstring>
#include <string>
#include <stdexcept>
#include <vector>
#include <algorithm>

#include <cuda_runtime.h>
#include "renderer_kernels.cuh"  // from /root/reference/renderer (include path)
#include "../include/dvr_b200.h"

#if USE_DOUBLE_PRECISION == 0
#define REFSYM(name) dvrref32_##name
#else
#define REFSYM(name) dvrref64_##name
#endif

namespace refshim {

extern thread_local std::string g_error;

template <typename Acc, int N>
inline Acc make_acc(const DvrTensor& t) {
  typedef typename Acc::PtrType P;
  int64_t zeros[N];
  for (int i = 0; i < N; ++i) zeros[i] = 0;
  if (t.ptr == nullptr) return Acc(static_cast<P>(nullptr), zeros, zeros);
  if (t.ndim != N) throw std::runtime_error("tensor has wrong number of dimensions");
  return Acc(static_cast<P>(t.ptr), t.size, t.stride);
}

inline int bcast(int B, const DvrTensor& t) {
  if (t.ptr == nullptr) return B;
  int b2 = int(t.size[0]);
  if (b2 == 1) return B;
  if (B == 1) return b2;
  if (b2 != B) throw std::runtime_error("batch size mismatch");
  return B;
}

// Builds kernel::RendererInputs the way Renderer::checkInputs does
// (renderer/renderer.cpp:49-263); scalar variants become 1-element host tensors.
struct InputsHolder {
  real_t boxMin[3], boxSize[3], step[1], mat[16];
  kernel::RendererInputs in;
  int B, W, H;
};

inline void build_inputs(const DvrInputs& a, InputsHolder& h) {
  h.W = a.screen_width;
  h.H = a.screen_height;
  int B = 1;
  const bool preshaded = a.volume_filter_mode == DVR_FILTER_PRESHADED;
  if (a.volume.ptr == nullptr || a.volume.ndim != 4) throw std::runtime_error("volume must be 4-D");
  if (!preshaded) B = int(a.volume.size[0]);
  DvrTensor boxMin = a.box_min, boxSize = a.box_size, step = a.step_size, mat = a.camera_matrix;
  if (!boxMin.ptr) {
    for (int i = 0; i < 3; ++i) h.boxMin[i] = real_t(a.box_min_value[i]);
    boxMin = DvrTensor{h.boxMin, 2, {1, 3}, {3, 1}};
  } else B = bcast(B, boxMin);
  if (!boxSize.ptr) {
    for (int i = 0; i < 3; ++i) h.boxSize[i] = real_t(a.box_size_value[i]);
    boxSize = DvrTensor{h.boxSize, 2, {1, 3}, {3, 1}};
  } else B = bcast(B, boxSize);
  if (!step.ptr) {
    h.step[0] = real_t(a.step_size_value);
    step = DvrTensor{h.step, 3, {1, 1, 1}, {1, 1, 1}};
  } else B = bcast(B, step);
  switch (a.camera_mode) {
    case DVR_CAMERA_RAY_START_DIR:
      B = bcast(B, a.camera_ray_start);
      B = bcast(B, a.camera_ray_dir);
      mat.ptr = nullptr;
      break;
    case DVR_CAMERA_INVERSE_VIEW_MATRIX:
      if (!mat.ptr) {
        for (int i = 0; i < 16; ++i) h.mat[i] = real_t(a.camera_matrix_value[i]);
        mat = DvrTensor{h.mat, 3, {1, 4, 4}, {16, 4, 1}};
      } else B = bcast(B, mat);
      break;
    case DVR_CAMERA_REFERENCE_FRAME:
      B = bcast(B, mat);
      break;
    default:
      throw std::runtime_error("unknown camera enum value");
  }
  B = bcast(B, a.tf);
  h.B = B;
  const bool rays = a.camera_mode == DVR_CAMERA_RAY_START_DIR;
  DvrTensor none{};
  h.in = kernel::RendererInputs{
      make_int2(h.W, h.H),
      make_acc<kernel::Tensor4Read, 4>(a.volume),
      make_int3(int(a.volume.size[1]), int(a.volume.size[2]), int(a.volume.size[3])),
      make_acc<kernel::Tensor2Read, 2>(boxMin),
      make_acc<kernel::Tensor2Read, 2>(boxSize),
      make_acc<kernel::Tensor4Read, 4>(rays ? a.camera_ray_start : none),
      make_acc<kernel::Tensor4Read, 4>(rays ? a.camera_ray_dir : none),
      make_acc<kernel::Tensor3Read, 3>(mat),
      real_t(a.camera_fov_y_radians),
      make_acc<kernel::Tensor3Read, 3>(step),
      make_acc<kernel::Tensor3Read, 3>(a.tf),
      real_t(a.blending_early_out)};
}

// Same decisions as checkAdjointOutput (renderer/renderer.cpp:265-365).
inline kernel::AdjointOutputs build_adjoint_outputs(const DvrAdjointOutputs& o, int B, int H,
                                                    int W) {
  kernel::AdjointOutputs r;
  DvrTensor none{};
  r.adj_volume = make_acc<kernel::BTensor4RW, 4>(o.has_volume_derivatives ? o.adj_volume : none);
  r.adj_stepSize =
      make_acc<kernel::BTensor3RW, 3>(o.has_stepsize_derivatives ? o.adj_step_size : none);
  r.stepSizeHasBroadcasting =
      o.has_stepsize_derivatives && (B != o.adj_step_size.size[0] || H != o.adj_step_size.size[1] ||
                                     W != o.adj_step_size.size[2]);
  r.adj_cameraRayStart =
      make_acc<kernel::BTensor4RW, 4>(o.has_camera_derivatives ? o.adj_camera_ray_start : none);
  r.adj_cameraRayDir =
      make_acc<kernel::BTensor4RW, 4>(o.has_camera_derivatives ? o.adj_camera_ray_dir : none);
  r.cameraHasBroadcasting =
      o.has_camera_derivatives &&
      (B != o.adj_camera_ray_start.size[0] || B != o.adj_camera_ray_dir.size[0]);
  r.adj_tf = make_acc<kernel::BTensor3RW, 3>(o.has_tf_derivatives ? o.adj_tf : none);
  return r;
}

// ---- runtime enum -> template parameter ---------------------------------------------------
template <typename E, E V>
struct EC {
  static constexpr E value = V;
};

template <typename F>
inline void with_filter(int m, F&& f) {
  switch (m) {
    case DVR_FILTER_NEAREST: f(EC<kernel::VolumeFilterMode, kernel::FilterNearest>()); break;
    case DVR_FILTER_TRILINEAR: f(EC<kernel::VolumeFilterMode, kernel::FilterTrilinear>()); break;
    default: throw std::runtime_error("unsupported volume filter mode");
  }
}
template <typename F>
inline void with_camera(int m, F&& f) {
  switch (m) {
    case DVR_CAMERA_RAY_START_DIR: f(EC<kernel::CameraMode, kernel::CameraRayStartDir>()); break;
    case DVR_CAMERA_INVERSE_VIEW_MATRIX:
      f(EC<kernel::CameraMode, kernel::CameraInverseViewMatrix>());
      break;
    case DVR_CAMERA_REFERENCE_FRAME: f(EC<kernel::CameraMode, kernel::CameraReferenceFrame>()); break;
    default: throw std::runtime_error("unknown camera enum value");
  }
}
template <typename F>
inline void with_blend(int m, F&& f) {
  switch (m) {
    case DVR_BLEND_BEER_LAMBERT: f(EC<kernel::BlendMode, kernel::BlendBeerLambert>()); break;
    case DVR_BLEND_ALPHA: f(EC<kernel::BlendMode, kernel::BlendAlpha>()); break;
    default: throw std::runtime_error("unknown blend mode");
  }
}
template <bool WithLog, typename F>
inline void with_tf(int m, F&& f) {
  switch (m) {
    case DVR_TF_IDENTITY: f(EC<kernel::TFMode, kernel::TFIdentity>()); break;
    case DVR_TF_TEXTURE: f(EC<kernel::TFMode, kernel::TFTexture>()); break;
    case DVR_TF_LINEAR: f(EC<kernel::TFMode, kernel::TFLinear>()); break;
    case DVR_TF_GAUSSIAN: f(EC<kernel::TFMode, kernel::TFGaussian>()); break;
    case DVR_TF_GAUSSIAN_LOG:
      if constexpr (WithLog) {
        f(EC<kernel::TFMode, kernel::TFGaussianLog>());
        break;
      }
      throw std::runtime_error("GaussianLog TF not supported here");
    default: throw std::runtime_error("unknown tf enum value");
  }
}
template <typename F>
inline void with_bool(bool b, F&& f) {
  if (b) f(std::true_type());
  else f(std::false_type());
}

template <typename F>
inline int guarded(F&& f) {
  try {
    f();
    return DVR_OK;
  } catch (const std::exception& e) {
    g_error = e.what();
    return DVR_ERR_INVALID_ARGUMENT;
  }
}

}  // namespace refshim
