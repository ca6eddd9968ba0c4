// Scalar / enum types shared by the host layers over the C-ABI: the torch-based renderer::Renderer (renderer.h) and the
// torch-free renderer::RendererCuda facade (renderer_cuda_facade.h). Mirrors renderer/renderer_commons.cuh:13-17 and the
// numeric enum values of renderer/renderer_settings.cuh:81-146.
#pragma once
#include <vector_types.h>  // int2, int3, float3, float4 (CUDA toolkit headers, host-only use)

typedef float real_t;  // renderer/renderer_commons.cuh:13-17 with USE_DOUBLE_PRECISION=0
typedef float3 real3;
typedef float4 real4;
inline real3 make_real3(real_t x, real_t y, real_t z) { return real3{x, y, z}; }
inline real4 make_real4(real_t x, real_t y, real_t z, real_t w) { return real4{x, y, z, w}; }

namespace kernel {
// numeric values of renderer/renderer_settings.cuh:81-146
enum VolumeFilterMode { FilterNearest = 0, FilterTrilinear = 1, __VolumeFilterModeCount__ = 2, FilterPreshaded = 3, FilterNetwork = 4 };
enum CameraMode { CameraRayStartDir = 0, CameraInverseViewMatrix = 1, CameraReferenceFrame = 2, __CameraModeCount__ = 3 };
enum TFMode { TFIdentity = 0, TFTexture = 1, TFLinear = 2, TFGaussian = 3, __TFModeCount__ = 4, TFPreshaded = 5, TFGaussianLog = 6 };
enum BlendMode { BlendBeerLambert = 0, BlendAlpha = 1, __BlendModeCount__ = 2 };
struct ForwardDifferencesSettings { static constexpr int NOT_ENABLED = -1; };  // renderer_settings.cuh:189-204
}  // namespace kernel

