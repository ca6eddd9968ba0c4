// renderer::Camera over the C-ABI; see camera.h. Behaviour follows renderer/camera.cpp:45-252.
#include "camera.h"

#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/torch.h>

#include <cmath>
#include <stdexcept>

#include "../../../include/dvr_b200.h"

namespace renderer {
namespace {

DvrTensor view_of(const torch::Tensor& t) {
  DvrTensor v{};
  v.ptr = t.data_ptr();
  v.ndim = int32_t(t.dim());
  for (int64_t i = 0; i < t.dim(); ++i) {
    v.size[i] = t.size(i);
    v.stride[i] = t.stride(i);
  }
  return v;
}

void raise_on_error(int rc) {
  if (rc == DVR_OK) return;
  const std::string msg = dvr_last_error();
  if (rc == DVR_ERR_INVALID_ARGUMENT) TORCH_CHECK(false, msg);
  throw std::runtime_error(msg);
}

void* stream_of(const torch::Device& dev) { return c10::cuda::getCurrentCUDAStream(dev.index()).stream(); }

// wording of renderer/pytorch_utils.h:5-19
void check_dim(const torch::Tensor& t, const char* name, int d) {
  TORCH_CHECK(t.defined(), name, " is not defined!");
  TORCH_CHECK(t.dim() == d, name, " must be a tensor with ", d, " dimensions, but has ", t.dim());
}
void check_size(const torch::Tensor& t, const char* name, int d, int64_t s) {
  TORCH_CHECK(t.size(d) == s, name, " must have ", s, " entries at dimension ", d, ", but has ", t.size(d), " entries");
}
int check_batch(const torch::Tensor& t, const char* name, int B1) {
  const int B2 = int(t.size(0));
  if (B2 == 1) return B1;
  if (B1 == 1) return B2;
  TORCH_CHECK(B1 == B2, name, " must be have the same batch size as other tensors, if batch>1. actual=", B2,
              ", expected=", B1);
  return B1;
}

struct V3 {
  float x, y, z;
};
V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
V3 normalize(V3 a) {
  const float inv = 1.0f / std::sqrt(dot(a, a));
  return {a.x * inv, a.y * inv, a.z * inv};
}

// 4x4 matrices in mathematical notation, M[row][col]
struct M4 {
  double m[4][4];
};
M4 mul(const M4& a, const M4& b) {
  M4 r{};
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j)
      for (int k = 0; k < 4; ++k) r.m[i][j] += a.m[i][k] * b.m[k][j];
  return r;
}
// Gauss-Jordan with partial pivoting (the view-projection matrix is well conditioned)
M4 inverse(M4 a) {
  M4 inv{};
  for (int i = 0; i < 4; ++i) inv.m[i][i] = 1.0;
  for (int c = 0; c < 4; ++c) {
    int p = c;
    for (int r = c + 1; r < 4; ++r)
      if (std::fabs(a.m[r][c]) > std::fabs(a.m[p][c])) p = r;
    if (a.m[p][c] == 0.0) throw std::runtime_error("view-projection matrix is singular");
    for (int k = 0; k < 4; ++k) {
      std::swap(a.m[c][k], a.m[p][k]);
      std::swap(inv.m[c][k], inv.m[p][k]);
    }
    const double d = 1.0 / a.m[c][c];
    for (int k = 0; k < 4; ++k) {
      a.m[c][k] *= d;
      inv.m[c][k] *= d;
    }
    for (int r = 0; r < 4; ++r) {
      if (r == c) continue;
      const double f = a.m[r][c];
      for (int k = 0; k < 4; ++k) {
        a.m[r][k] -= f * a.m[c][k];
        inv.m[r][k] -= f * inv.m[c][k];
      }
    }
  }
  return inv;
}

}  // namespace

const char* Camera::OrientationNames[6] = {"Xp", "Xm", "Yp", "Ym", "Zp", "Zm"};
const float3 Camera::OrientationUp[6] = {float3{1, 0, 0},  float3{-1, 0, 0}, float3{0, 1, 0},
                                         float3{0, -1, 0}, float3{0, 0, 1},  float3{0, 0, -1}};
const int3 Camera::OrientationPermutation[6] = {int3{2, -1, -3}, int3{-2, 1, 3},   int3{1, 2, 3},
                                                int3{-1, -2, -3}, int3{-3, -1, 2}, int3{3, 1, -2}};
const bool Camera::OrientationInvertYaw[6] = {true, false, false, true, false, true};
const bool Camera::OrientationInvertPitch[6] = {false, false, false, false, false, false};

// Camera::computeMatrices (renderer/camera.cpp:86-112): left-handed look-at (glm::lookAtLH), zero-to-one
// left-handed perspective (camera.cpp:47-64), inverse of proj * view. The reference returns the
// TRANSPOSE of glm's column-major inverse, i.e. result[i] holds row i of the inverse, which is what the
// marcher's InverseViewMatrix camera reads as tensor[0][i][j] (renderer/renderer.cpp:170-179).
mat4 Camera::computeInverseViewProjectionMatrix(float3 origin, float3 lookAt, float3 upIn, float fovDegrees,
                                                int width, int height, float nearClip, float farClip) {
  const V3 eye = {origin.x, origin.y, origin.z};
  const V3 f = normalize(sub({lookAt.x, lookAt.y, lookAt.z}, eye));
  const V3 s = normalize(cross(normalize({upIn.x, upIn.y, upIn.z}), f));
  const V3 u = cross(f, s);
  M4 viewM{};
  viewM.m[0][0] = s.x; viewM.m[0][1] = s.y; viewM.m[0][2] = s.z; viewM.m[0][3] = -dot(s, eye);
  viewM.m[1][0] = u.x; viewM.m[1][1] = u.y; viewM.m[1][2] = u.z; viewM.m[1][3] = -dot(u, eye);
  viewM.m[2][0] = f.x; viewM.m[2][1] = f.y; viewM.m[2][2] = f.z; viewM.m[2][3] = -dot(f, eye);
  viewM.m[3][3] = 1.0;
  const float rad = fovDegrees * 0.01745329251994329576923690768489f;  // glm::radians
  const float h = std::cos(0.5f * rad) / std::sin(0.5f * rad);
  const float w = h * float(height) / float(width);
  M4 proj{};
  proj.m[0][0] = w;
  proj.m[1][1] = h;
  proj.m[2][2] = farClip / (farClip - nearClip);
  proj.m[3][2] = 1.0;
  proj.m[2][3] = -(farClip * nearClip) / (farClip - nearClip);
  const M4 inv = inverse(mul(proj, viewM));
  mat4 r;
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) r[i][j] = float(inv.m[i][j]);
  return r;
}

torch::Tensor Camera::viewportFromLookAt(const torch::Tensor& origin, const torch::Tensor& lookAt,
                                         const torch::Tensor& cameraUp) {
  check_dim(origin, "origin", 2);
  check_size(origin, "origin", 1, 3);
  int B = int(origin.size(0));
  check_dim(lookAt, "lookAt", 2);
  check_size(lookAt, "lookAt", 1, 3);
  B = check_batch(lookAt, "lookAt", B);
  check_dim(cameraUp, "cameraUp", 2);
  check_size(cameraUp, "cameraUp", 1, 3);
  check_batch(cameraUp, "cameraUp", B);
  namespace F = torch::nn::functional;
  const auto front = F::normalize(lookAt - origin);
  const auto right = F::normalize(torch::cross(front, cameraUp, 1));
  const auto up2 = F::normalize(torch::cross(right, front, 1));
  return torch::stack({origin.expand_as(right), right, up2}, 1);
}

torch::Tensor Camera::viewportFromSphere(const torch::Tensor& center, const torch::Tensor& yawRadians,
                                         const torch::Tensor& pitchRadians, const torch::Tensor& distance,
                                         Orientation orientation) {
  check_dim(center, "center", 2);
  check_size(center, "center", 1, 3);
  int B = int(center.size(0));
  const torch::Tensor* ts[3] = {&yawRadians, &pitchRadians, &distance};
  const char* names[3] = {"yawRadians", "pitchRadians", "distance"};
  for (int k = 0; k < 3; ++k) {
    check_dim(*ts[k], names[k], 2);
    check_size(*ts[k], names[k], 1, 1);
    B = check_batch(*ts[k], names[k], B);
  }
  TORCH_CHECK(int(orientation) >= 0 && int(orientation) < 6, "unknown orientation");
  const auto yaw = yawRadians * (OrientationInvertYaw[orientation] ? +1.0f : -1.0f);
  const auto pitch = pitchRadians * (OrientationInvertPitch[orientation] ? +1.0f : -1.0f);
  const torch::Tensor pos[3] = {torch::cos(pitch) * torch::cos(yaw) * distance, torch::sin(pitch) * distance,
                                torch::cos(pitch) * torch::sin(yaw) * distance};
  const int3 perm = OrientationPermutation[orientation];
  const int p[3] = {perm.x, perm.y, perm.z};
  std::vector<torch::Tensor> axes;
  for (int i = 0; i < 3; ++i) {
    const torch::Tensor& a = pos[std::abs(p[i]) - 1];
    axes.push_back(p[i] > 0 ? a : -a);
  }
  const auto origin = torch::cat(axes, 1);
  const float3 upv = OrientationUp[orientation];
  const auto up = torch::tensor({{upv.x, upv.y, upv.z}}, center.options()).expand({B, 3});
  return viewportFromLookAt(origin + center, center.expand({B, 3}), up);
}

namespace {

// GenerateRayFunction of renderer/camera.cpp:182-246
class GenerateRaysFn : public torch::autograd::Function<GenerateRaysFn> {
 public:
  static torch::autograd::variable_list forward(torch::autograd::AutogradContext* ctx, const torch::Tensor& viewport,
                                                double fovY, int64_t W, int64_t H) {
    check_dim(viewport, "viewport", 3);
    check_size(viewport, "viewport", 1, 3);
    check_size(viewport, "viewport", 2, 3);
    TORCH_CHECK(viewport.dtype() == c10::ScalarType::Float, "viewport must be of type c10::ScalarType::Float, but is ",
                viewport.dtype());
    TORCH_CHECK(viewport.device().is_cuda(),
                "viewport is a CPU tensor: this build has no CPU back-end (the hot path is CUDA for sm_100a only)");
    const torch::Device dev = viewport.device();
    const c10::cuda::CUDAGuard guard(dev);
    const int64_t B = viewport.size(0);
    ctx->save_for_backward({viewport});
    ctx->saved_data["fovY"] = fovY;
    torch::Tensor rayStart = torch::empty({B, H, W, 3}, viewport.options());
    torch::Tensor rayDir = torch::empty({B, H, W, 3}, viewport.options());
    const DvrTensor vp = view_of(viewport), rs = view_of(rayStart), rd = view_of(rayDir);
    raise_on_error(dvr_generate_rays(&vp, float(fovY), int32_t(W), int32_t(H), &rs, &rd, stream_of(dev)));
    return {rayStart, rayDir};
  }

  static torch::autograd::variable_list backward(torch::autograd::AutogradContext* ctx,
                                                 torch::autograd::variable_list grad) {
    const auto saved = ctx->get_saved_variables();
    const torch::Tensor viewport = saved[0];
    const double fovY = ctx->saved_data["fovY"].toDouble();
    const torch::Device dev = viewport.device();
    const c10::cuda::CUDAGuard guard(dev);
    // undefined incoming gradients (an output that does not reach the loss) count as zeros
    torch::Tensor gs = grad[0], gd = grad[1];
    TORCH_CHECK(gs.defined() || gd.defined(), "generateRays backward: no incoming gradient");
    if (!gs.defined()) gs = torch::zeros_like(gd);
    if (!gd.defined()) gd = torch::zeros_like(gs);
    torch::Tensor adjViewport = torch::empty({viewport.size(0), 3, 3}, gs.options());
    const DvrTensor vp = view_of(viewport), a = view_of(gs), b = view_of(gd), out = view_of(adjViewport);
    raise_on_error(dvr_generate_rays_adjoint(&vp, float(fovY), &a, &b, &out, stream_of(dev)));
    return {adjViewport, torch::Tensor(), torch::Tensor(), torch::Tensor()};
  }
};

}  // namespace

std::tuple<torch::Tensor, torch::Tensor> Camera::generateRays(const torch::Tensor& viewport, float fovY,
                                                              int screenWidth, int screenHeight) {
  const auto ret = GenerateRaysFn::apply(viewport, double(fovY), int64_t(screenWidth), int64_t(screenHeight));
  return std::make_tuple(ret[0], ret[1]);
}

}  // namespace renderer
