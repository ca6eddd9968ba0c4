// renderer::TFUtils / TFPoint of the reference (renderer/tf_utils.h:93-152): builds the TF tensors the marcher
// consumes from editor-style control points (CIE-Lab colours + opacities) and pre-shades a density volume for
// the FilterPreshaded / TFPreshaded marcher.
//  * assembleFromSettings / getPiecewiseTensor / getTextureTensor are host functions, as in the reference;
//  * preshadeVolume is a torch::autograd::Function over dvr_preshade_volume{,_adjoint} (CUDA tensors only);
//  * findBestFit (random search for densities matching a colour volume) is not part of the differentiable path
//    and is not provided.
#pragma once
#include <vector>

#include "renderer.h"

namespace renderer {

struct TFPoint {
  real_t pos;
  real4 val;  // rgb, opacity
};

// colour conversions of renderer/tf_utils.h:19-91 (D65 white point)
real3 labToRgb(const real3& lab);
real3 rgbToLab(const real3& rgb);

struct TFUtils {
  TFUtils() = delete;
  static std::vector<TFPoint> assembleFromSettings(const std::vector<real3>& colorValuesLab,
                                                   const std::vector<real_t>& colorPositions,
                                                   const std::vector<real_t>& opacityValues,
                                                   const std::vector<real_t>& opacityPositions, real_t minDensity,
                                                   real_t maxDensity, real_t opacityScaling,
                                                   bool purgeZeroOpacityRegions = true);
  // 1*R*5 (r, g, b, opacity, position) for kernel::TFLinear; float32 on the CPU
  static torch::Tensor getPiecewiseTensor(const std::vector<TFPoint>& points);
  // 1*resolution*4 for kernel::TFTexture; float32 on the CPU
  static torch::Tensor getTextureTensor(const std::vector<TFPoint>& points, int resolution);
  // density 1*X*Y*Z -> colour 4*X*Y*Z; differentiable with respect to the density
  static torch::Tensor preshadeVolume(const torch::Tensor& densityVolume, const torch::Tensor& tf,
                                      kernel::TFMode tfMode);
};

}  // namespace renderer
