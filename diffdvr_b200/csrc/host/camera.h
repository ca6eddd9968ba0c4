// renderer::Camera of the reference (renderer/camera.h:14-135) over the C-ABI: the producer of the
// per-pixel rays the marcher consumes in CameraRayStartDir mode and the consumer of
// adj_cameraRayStart / adj_cameraRayDir.
//  * generateRays is a torch::autograd::Function like the reference's GenerateRayFunction
//    (renderer/camera.cpp:182-252); forward and backward are the sm_100a kernels of dvr_camera.cu
//    (dvr_generate_rays / dvr_generate_rays_adjoint). CUDA tensors only: there is no CPU back-end.
//  * viewportFromLookAt / viewportFromSphere are compositions of differentiable torch operators,
//    as in the reference (camera.cpp:118-180).
//  * computeInverseViewProjectionMatrix needs glm in the reference; here the same left-handed
//    look-at and zero-to-one perspective matrices are written out on the POD `mat4`.
#pragma once
#include "renderer.h"

namespace renderer {

class Camera {
 public:
  // renderer/camera.h:39-41 -> inverse view-projection, row i of the inverse in result[i]
  static mat4 computeInverseViewProjectionMatrix(float3 cameraOrigin, float3 cameraLookAt, float3 cameraUp,
                                                 float fovDegrees, int width, int height, float nearClip,
                                                 float farClip);

  enum Orientation { Xp, Xm, Yp, Ym, Zp, Zm };
  static const char* OrientationNames[6];
  static const float3 OrientationUp[6];
  static const int3 OrientationPermutation[6];
  static const bool OrientationInvertYaw[6];
  static const bool OrientationInvertPitch[6];

  // B*3 each -> B*3*3 (eye, right, up)
  static torch::Tensor viewportFromLookAt(const torch::Tensor& origin, const torch::Tensor& lookAt,
                                          const torch::Tensor& cameraUp);
  // center B*3, yaw / pitch / distance B*1 -> B*3*3
  static torch::Tensor viewportFromSphere(const torch::Tensor& center, const torch::Tensor& yawRadians,
                                          const torch::Tensor& pitchRadians, const torch::Tensor& distance,
                                          Orientation orientation);
  // viewport B*3*3 -> (ray_start, ray_dir), both B*H*W*3; differentiable with respect to viewport
  static std::tuple<torch::Tensor, torch::Tensor> generateRays(const torch::Tensor& viewport, float fovY,
                                                               int screenWidth, int screenHeight);
};

}  // namespace renderer
