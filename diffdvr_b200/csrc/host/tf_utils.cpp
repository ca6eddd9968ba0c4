// renderer::TFUtils over the C-ABI; see tf_utils.h. Behaviour follows renderer/tf_utils.cpp:12-300.
#include "tf_utils.h"

#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/torch.h>

#include <cfloat>
#include <cmath>
#include <stdexcept>

#include "../../../include/dvr_b200.h"

namespace renderer {
namespace {

real3 lerp3(const real3& a, const real3& b, real_t t) {
  return {a.x + t * (b.x - a.x), a.y + t * (b.y - a.y), a.z + t * (b.z - a.z)};
}
real_t lerp1(real_t a, real_t b, real_t t) { return a + t * (b - a); }
real_t clamp1(real_t v, real_t lo, real_t hi) { return std::max(lo, std::min(v, hi)); }

real3 xyzToRgb(const real3& xyz) {  // tf_utils.h:37-55
  const float x = xyz.x / 100.0f, y = xyz.y / 100.0f, z = xyz.z / 100.0f;
  float r = x * 3.2404542f + y * -1.5371385f + z * -0.4985314f;
  float g = x * -0.9692660f + y * 1.8760108f + z * 0.0415560f;
  float b = x * 0.0556434f + y * -0.2040259f + z * 1.0572252f;
  auto gamma = [](float c) { return c > 0.0031308f ? float(1.055f * std::pow(c, 1.0f / 2.4f) - 0.055f) : 12.92f * c; };
  return {gamma(r), gamma(g), gamma(b)};
}
real3 rgbToXyz(const real3& rgb) {  // tf_utils.h:21-33
  auto lin = [](float c) { return (c > 0.04045f ? std::pow((c + 0.055f) / 1.055f, 2.4f) : c / 12.92f) * 100.0f; };
  const float r = lin(rgb.x), g = lin(rgb.y), b = lin(rgb.z);
  return {r * 0.4124564f + g * 0.3575761f + b * 0.1804375f, r * 0.2126729f + g * 0.7151522f + b * 0.0721750f,
          r * 0.0193339f + g * 0.1191920f + b * 0.9503041f};
}

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

}  // namespace

real3 labToRgb(const real3& lab) {  // tf_utils.h:74-91
  float y = (lab.x + 16.0f) / 116.0f;
  float x = lab.y / 500.0f + y;
  float z = y - lab.z / 200.0f;
  const float x3 = x * x * x, y3 = y * y * y, z3 = z * z * z;
  x = (x3 > 0.008856f ? x3 : (x - 16.0f / 116.0f) / 7.787f) * 95.047f;
  y = (y3 > 0.008856f ? y3 : (y - 16.0f / 116.0f) / 7.787f) * 100.0f;
  z = (z3 > 0.008856f ? z3 : (z - 16.0f / 116.0f) / 7.787f) * 108.883f;
  return xyzToRgb({x, y, z});
}

real3 rgbToLab(const real3& rgb) {  // tf_utils.h:57-72
  const real3 xyz = rgbToXyz(rgb);
  auto f = [](float v) { return v > 0.008856f ? std::cbrt(v) : 7.787f * v + 16.0f / 116.0f; };
  const float x = f(xyz.x / 95.047f), y = f(xyz.y / 100.00f), z = f(xyz.z / 108.883f);
  return {116.0f * y - 16.0f, 500.0f * (x - y), 200.0f * (y - z)};
}

// Merges the colour and the opacity control-point lists into one list sorted by density
// (renderer/tf_utils.cpp:12-127): positions are mapped into [minDensity, maxDensity]; guard points at -1 and 2
// extend both lists beyond [0, 1]; at every position of either list the other quantity is interpolated
// (colours in Lab space, then converted); optionally points inside zero-opacity stretches and coincident
// points are dropped; finally colours are clamped to [0, 1) and the opacity is scaled.
std::vector<TFPoint> TFUtils::assembleFromSettings(const std::vector<real3>& colorValuesLab_,
                                                   const std::vector<real_t>& colorPositions_,
                                                   const std::vector<real_t>& opacityValues_,
                                                   const std::vector<real_t>& opacityPositions_, real_t minDensity,
                                                   real_t maxDensity, real_t opacityScaling,
                                                   bool purgeZeroOpacityRegions) {
  TORCH_CHECK(!colorValuesLab_.empty() && colorValuesLab_.size() == colorPositions_.size(),
              "colour values and positions must be non-empty lists of the same length");
  TORCH_CHECK(!opacityValues_.empty() && opacityValues_.size() == opacityPositions_.size(),
              "opacity values and positions must be non-empty lists of the same length");
  std::vector<real3> cv = colorValuesLab_;
  std::vector<real_t> cp = colorPositions_, ov = opacityValues_, op = opacityPositions_;
  for (real_t& d : cp) d = minDensity + d * (maxDensity - minDensity);
  for (real_t& d : op) d = minDensity + d * (maxDensity - minDensity);
  if (cp.front() > 0) { cp.insert(cp.begin(), -1.0f); cv.insert(cv.begin(), cv.front()); }
  if (op.front() > 0) { op.insert(op.begin(), -1.0f); ov.insert(ov.begin(), ov.front()); }
  if (cp.back() < 1) { cp.push_back(2); cv.push_back(cv.back()); }
  if (op.back() < 1) { op.push_back(2); ov.push_back(ov.back()); }

  auto point = [](real_t pos, const real3& rgb, real_t a) { return TFPoint{pos, real4{rgb.x, rgb.y, rgb.z, a}}; };
  std::vector<TFPoint> points;
  points.push_back(point(cp[0], labToRgb(cv[0]), ov[0]));
  size_t io = 0, ic = 0;
  while (io + 1 < op.size() && ic + 1 < cp.size()) {
    if (op[io + 1] < cp[ic + 1]) {  // next point comes from the opacity list: interpolate the colour
      const real_t t = (op[io + 1] - cp[ic]) / (cp[ic + 1] - cp[ic]);
      points.push_back(point(op[io + 1], labToRgb(lerp3(cv[ic], cv[ic + 1], t)), ov[io + 1]));
      ++io;
    } else {  // next point comes from the colour list: interpolate the opacity
      const real_t t = (cp[ic + 1] - op[io]) / (op[io + 1] - op[io]);
      points.push_back(point(cp[ic + 1], labToRgb(cv[ic + 1]), lerp1(ov[io], ov[io + 1], t)));
      ++ic;
    }
  }
  if (purgeZeroOpacityRegions) {
    constexpr real_t EPS = 1e-7f;
    for (int64_t i = 0; i < int64_t(points.size()) - 2;) {
      const bool zero = points[i].val.w < EPS && points[i + 1].val.w < EPS && points[i + 2].val.w < EPS;
      if (zero || points[i + 1].pos - points[i].pos < EPS) points.erase(points.begin() + (i + 1));
      else ++i;
    }
  }
  for (TFPoint& p : points) {
    const real_t hi = 1.0f - FLT_EPSILON;
    p.val = real4{clamp1(p.val.x, 0.f, hi), clamp1(p.val.y, 0.f, hi), clamp1(p.val.z, 0.f, hi),
                  clamp1(p.val.w, 0.f, hi) * opacityScaling};
  }
  return points;
}

torch::Tensor TFUtils::getPiecewiseTensor(const std::vector<TFPoint>& points) {
  const int64_t n = int64_t(points.size());
  torch::Tensor t = torch::empty({1, n, 5}, at::TensorOptions().dtype(c10::ScalarType::Float));
  auto acc = t.accessor<float, 3>();
  for (int64_t i = 0; i < n; ++i) {
    acc[0][i][0] = points[i].val.x; acc[0][i][1] = points[i].val.y; acc[0][i][2] = points[i].val.z;
    acc[0][i][3] = points[i].val.w; acc[0][i][4] = points[i].pos;
  }
  return t;
}

// texel i holds the piecewise-linear TF at density (i + 0.5) / resolution (renderer/tf_utils.cpp:146-178)
torch::Tensor TFUtils::getTextureTensor(const std::vector<TFPoint>& points, int resolution) {
  TORCH_CHECK(points.size() >= 2, "at least two control points are needed");
  TORCH_CHECK(resolution >= 1, "resolution must be positive");
  torch::Tensor t = torch::empty({1, resolution, 4}, at::TensorOptions().dtype(c10::ScalarType::Float));
  auto acc = t.accessor<float, 3>();
  const int n = int(points.size());
  for (int i = 0; i < resolution; ++i) {
    const real_t density = (i + 0.5f) / resolution;
    int idx;
    for (idx = 0; idx < n - 2; ++idx)
      if (points[idx + 1].pos > density) break;
    const real_t lo = points[idx].pos, hi = points[idx + 1].pos;
    const real_t f = (density - lo) / (hi - lo);
    const real4 a = points[idx].val, b = points[idx + 1].val;
    acc[0][i][0] = (1 - f) * a.x + f * b.x; acc[0][i][1] = (1 - f) * a.y + f * b.y;
    acc[0][i][2] = (1 - f) * a.z + f * b.z; acc[0][i][3] = (1 - f) * a.w + f * b.w;
  }
  return t;
}

namespace {

// PreshadeVolumeFunction of renderer/tf_utils.cpp:180-286
class PreshadeFn : public torch::autograd::Function<PreshadeFn> {
 public:
  static torch::autograd::variable_list forward(torch::autograd::AutogradContext* ctx, const torch::Tensor& density,
                                                const torch::Tensor& tf, int64_t tfMode) {
    TORCH_CHECK(density.defined() && tf.defined(), "densityVolume / tf is not defined!");
    TORCH_CHECK(density.device().is_cuda(),
                "densityVolume is a CPU tensor: this build has no CPU back-end (the hot path is CUDA for sm_100a only)");
    TORCH_CHECK(tf.device() == density.device(), "All tensors must be on the same device, but tf is different, is_cuda=",
                tf.device().is_cuda());
    TORCH_CHECK(density.dtype() == c10::ScalarType::Float && tf.dtype() == c10::ScalarType::Float,
                "densityVolume and tf must be of type c10::ScalarType::Float");
    TORCH_CHECK(density.dim() == 4, "densityVolume must be a tensor with 4 dimensions, but has ", density.dim());
    const torch::Device dev = density.device();
    const c10::cuda::CUDAGuard guard(dev);
    torch::Tensor out = torch::empty({4, density.size(1), density.size(2), density.size(3)}, density.options());
    const DvrTensor v = view_of(density), t = view_of(tf), c = view_of(out);
    raise_on_error(dvr_preshade_volume(&v, &t, int32_t(tfMode), &c, stream_of(dev)));
    ctx->save_for_backward({density, tf});
    ctx->saved_data["tfMode"] = tfMode;
    return {out};
  }

  static torch::autograd::variable_list backward(torch::autograd::AutogradContext* ctx,
                                                 torch::autograd::variable_list grad) {
    const auto saved = ctx->get_saved_variables();
    const torch::Tensor density = saved[0], tf = saved[1];
    const int64_t tfMode = ctx->saved_data["tfMode"].toInt();
    const torch::Device dev = density.device();
    const c10::cuda::CUDAGuard guard(dev);
    torch::Tensor adjVolume = torch::empty_like(density);
    const DvrTensor v = view_of(density), t = view_of(tf), g = view_of(grad[0]), o = view_of(adjVolume);
    raise_on_error(dvr_preshade_volume_adjoint(&v, &t, int32_t(tfMode), &g, &o, stream_of(dev)));
    return {adjVolume, torch::Tensor(), torch::Tensor()};
  }
};

}  // namespace

torch::Tensor TFUtils::preshadeVolume(const torch::Tensor& densityVolume, const torch::Tensor& tf,
                                      kernel::TFMode tfMode) {
  return PreshadeFn::apply(densityVolume, tf, int64_t(tfMode))[0];
}

}  // namespace renderer
