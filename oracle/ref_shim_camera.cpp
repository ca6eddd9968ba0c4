// TEST INFRASTRUCTURE ONLY -- never linked into the product.
//
// extern "C" entry points for Camera::generateRays (forward + backward) and
// Renderer::compareToImage built on the REFERENCE's own headers where they lie
// (-I/root/reference/renderer), see ref_shim_common.h:
//   * forward rays: kernel::CameraEval<CameraReferenceFrame>::computeRays
//     (renderer/renderer_camera.cuh:38-58), the same arithmetic as generateRaysForwardImpl
//     (renderer/camera_impl.cu:17-37; that file itself needs cuMat + cub and cannot be compiled);
//   * backward: the per-pixel term of GenerateRaysAdjoint_InputOp (camera_impl.cu:134-178)
//     re-assembled from the reference's kernel::adjNormalize / kernel::adjCross
//     (renderer/renderer_adjoint.cuh:28-64), summed per view in pixel order like the reference's
//     CPU branch (std::accumulate, camera_impl.cu:347-354);
//   * comparison: kernel::CompareToImage<D> (renderer/renderer_compareToImage.cuh:9-65) over the
//     restated cudAD, looped like CompareToImage_{NoReduce,WithReduce}_CPU_Kernel
//     (renderer/renderer_cpu.cpp:156-243) minus its debug printf.
#include "ref_shim_common.h"
#include "renderer_adjoint.cuh"
#include "renderer_compareToImage.cuh"

using namespace refshim;

extern "C" int REFSYM(generate_rays)(const DvrTensor* viewport, float fovY, int32_t W, int32_t H,
                                     const DvrTensor* rayStart, const DvrTensor* rayDir, void*) {
  return guarded([&] {
    const int B = int(viewport->size[0]);
    DvrTensor none{};
    kernel::RendererInputs in{
        make_int2(W, H),
        make_acc<kernel::Tensor4Read, 4>(none),
        make_int3(0, 0, 0),
        make_acc<kernel::Tensor2Read, 2>(none),
        make_acc<kernel::Tensor2Read, 2>(none),
        make_acc<kernel::Tensor4Read, 4>(none),
        make_acc<kernel::Tensor4Read, 4>(none),
        make_acc<kernel::Tensor3Read, 3>(*viewport),
        real_t(fovY),
        make_acc<kernel::Tensor3Read, 3>(none),
        make_acc<kernel::Tensor3Read, 3>(none),
        real_t(1)};
    auto rs = make_acc<kernel::Tensor4RW, 4>(*rayStart);
    auto rd = make_acc<kernel::Tensor4RW, 4>(*rayDir);
    for (int b = 0; b < B; ++b)
      for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
          real3 s, d;
          kernel::CameraEval<kernel::CameraReferenceFrame>::computeRays(x, y, b, in, s, d);
          kernel::writeReal3(s, rs, b, y, x);
          kernel::writeReal3(d, rd, b, y, x);
        }
  });
}

extern "C" int REFSYM(generate_rays_adjoint)(const DvrTensor* viewportT, float fovY,
                                             const DvrTensor* adjStart, const DvrTensor* adjDir,
                                             const DvrTensor* adjViewport, void*) {
  return guarded([&] {
    using namespace kernel;
    const int B = int(viewportT->size[0]), H = int(adjStart->size[1]), W = int(adjStart->size[2]);
    auto viewport = make_acc<Tensor3Read, 3>(*viewportT);
    auto adj_rayStart = make_acc<Tensor4Read, 4>(*adjStart);
    auto adj_rayDir = make_acc<Tensor4Read, 4>(*adjDir);
    auto out = make_acc<Tensor3RW, 3>(*adjViewport);
    const float fovX = fovY * float(W) / float(H);
    const float tanFovX = tanf(fovX / 2);
    const float tanFovY = tanf(fovY / 2);
    for (int b = 0; b < B; ++b) {
      real3 sumEye = make_real3(0), sumRight = make_real3(0), sumUp = make_real3(0);
      for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
          real3 eye = fetchReal3(viewport, b, 0);
          real3 right = fetchReal3(viewport, b, 1);
          real3 up = fetchReal3(viewport, b, 2);
          real3 front = cross(up, right);
          const real_t fx = 2 * (x + 0.5f) / real_t(W) - 1;
          const real_t fy = 2 * (y + 0.5f) / real_t(H) - 1;
          real3 dir = front + fx * tanFovX * right + fy * tanFovY * up;
          real3 adj_eye = fetchReal3(adj_rayStart, b, y, x);
          real3 adj_dirNormalized = fetchReal3(adj_rayDir, b, y, x);
          real3 adj_dir = adjNormalize(dir, adj_dirNormalized);
          real3 adj_front = adj_dir;
          real3 adj_right = adj_dir * (fx * tanFovX);
          real3 adj_up = adj_dir * (fy * tanFovY);
          adjCross(up, right, adj_front, adj_up, adj_right);
          sumEye = sumEye + adj_eye;
          sumRight = sumRight + adj_right;
          sumUp = sumUp + adj_up;
          (void)eye;
        }
      writeReal3(sumEye, out, b, 0);
      writeReal3(sumRight, out, b, 1);
      writeReal3(sumUp, out, b, 2);
    }
  });
}

namespace {

template <int D>
void compare_impl(const kernel::Tensor4Read& color, const kernel::Tensor5Read& grads,
                  const kernel::Tensor4Read& reference, const DvrTensor& diffOut,
                  const DvrTensor& gradsOut, int B, int H, int W, bool reduce) {
  using namespace kernel;
  if (!reduce) {
    auto d = make_acc<Tensor4RW, 4>(diffOut);
    auto g = make_acc<Tensor4RW, 4>(gradsOut);
    for (int b = 0; b < B; ++b)
      for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
          cudAD::fvar<real4, D> input = CompareToImage<D>::readInput(color, grads, b, y, x);
          real4 ref = CompareToImage<D>::readReference(reference, b, y, x);
          cudAD::fvar<real_t, D> result = CompareToImage<D>::compare(input, ref);
          CompareToImage<D>::writeOutput(result, d, g, b, y, x);
        }
  } else {
    auto d = make_acc<Tensor1RW, 1>(diffOut);
    auto g = make_acc<Tensor1RW, 1>(gradsOut);
    cudAD::fvar<real_t, D> acc;
    const real_t normalization = real_t(1) / real_t(B * W * H);
    for (int b = 0; b < B; ++b)
      for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
          cudAD::fvar<real4, D> input = CompareToImage<D>::readInput(color, grads, b, y, x);
          real4 ref = CompareToImage<D>::readReference(reference, b, y, x);
          cudAD::fvar<real_t, D> result = CompareToImage<D>::compare(input, ref);
          acc += result * normalization;
        }
    CompareToImage<D>::writeOutput(acc, d, g);
  }
}

}  // namespace

extern "C" int REFSYM(compare_to_image)(const DvrTensor* colorT, const DvrTensor* gradsT,
                                        const DvrTensor* referenceT, const DvrTensor* diffOut,
                                        const DvrTensor* gradsOut, int32_t reduce, void*) {
  return guarded([&] {
    const int B = int(colorT->size[0]), H = int(colorT->size[1]), W = int(colorT->size[2]);
    const int D = int(gradsT->size[3]);
    auto color = make_acc<kernel::Tensor4Read, 4>(*colorT);
    auto grads = make_acc<kernel::Tensor5Read, 5>(*gradsT);
    auto reference = make_acc<kernel::Tensor4Read, 4>(*referenceT);
    switch (D) {
#define CASE(N) \
  case N: compare_impl<N>(color, grads, reference, *diffOut, *gradsOut, B, H, W, reduce != 0); break;
      CASE(1) CASE(2) CASE(3) CASE(4) CASE(6) CASE(8)
#undef CASE
      default: throw std::runtime_error("ref shim: compare_to_image is instantiated for D in {1,2,3,4,6,8}");
    }
  });
}

// ---- TFUtils::preshadeVolume forward / density adjoint ------------------------------------------------
// PreshadeVolumeImpl / PreshadeVolumeAdjImpl of renderer/tf_utils_impl.cu:16-33, 98-122 re-assembled from the
// reference's own TransferFunctionEval<mode>::eval / adjoint<false,false> (renderer/renderer_tf.cuh); that file
// itself needs cuMat and cannot be compiled here.
namespace {

template <kernel::TFMode MODE>
void preshade_impl(const kernel::Tensor4Read& volume, const kernel::Tensor3Read& tf, kernel::Tensor4RW& color) {
  for (int x = 0; x < volume.size(1); ++x)
    for (int y = 0; y < volume.size(2); ++y)
      for (int z = 0; z < volume.size(3); ++z) {
        real_t density = volume[0][x][y][z];
        kernel::TransferFunctionEval<MODE> tfEval;
        real4 c = tfEval.eval(tf, 0, density);
        color[0][x][y][z] = c.x;
        color[1][x][y][z] = c.y;
        color[2][x][y][z] = c.z;
        color[3][x][y][z] = c.w;
      }
}

template <kernel::TFMode MODE>
void preshade_adj_impl(const kernel::Tensor4Read& volume, const kernel::Tensor3Read& tf,
                       const kernel::Tensor4Read& adj_color, kernel::Tensor4RW& adj_volume) {
  DvrTensor none{};
  kernel::BTensor3RW dummy_adj_tf = make_acc<kernel::BTensor3RW, 3>(none);
  for (int x = 0; x < volume.size(1); ++x)
    for (int y = 0; y < volume.size(2); ++y)
      for (int z = 0; z < volume.size(3); ++z) {
        real_t density = volume[0][x][y][z];
        kernel::TransferFunctionEval<MODE> tfEval;
        real4 adj_c = make_real4(adj_color[0][x][y][z], adj_color[1][x][y][z], adj_color[2][x][y][z],
                                 adj_color[3][x][y][z]);
        real_t adj_density = 0;
        tfEval.template adjoint<false, false>(tf, 0, density, adj_c, adj_density, dummy_adj_tf, nullptr);
        adj_volume[0][x][y][z] = adj_density;
      }
}

}  // namespace

#define TF_SWITCH(mode, CALL)                                                     \
  switch (mode) {                                                                 \
    case DVR_TF_IDENTITY: { constexpr auto M = kernel::TFIdentity; CALL; } break;  \
    case DVR_TF_TEXTURE: { constexpr auto M = kernel::TFTexture; CALL; } break;    \
    case DVR_TF_LINEAR: { constexpr auto M = kernel::TFLinear; CALL; } break;      \
    case DVR_TF_GAUSSIAN: { constexpr auto M = kernel::TFGaussian; CALL; } break;  \
    case DVR_TF_GAUSSIAN_LOG: { constexpr auto M = kernel::TFGaussianLog; CALL; } break; \
    default: throw std::runtime_error("unknown tf mode");                        \
  }

extern "C" int REFSYM(preshade_volume)(const DvrTensor* volumeT, const DvrTensor* tfT, int32_t mode,
                                       const DvrTensor* colorT, void*) {
  return guarded([&] {
    auto volume = make_acc<kernel::Tensor4Read, 4>(*volumeT);
    auto tf = make_acc<kernel::Tensor3Read, 3>(*tfT);
    auto color = make_acc<kernel::Tensor4RW, 4>(*colorT);
    TF_SWITCH(mode, preshade_impl<M>(volume, tf, color));
  });
}

extern "C" int REFSYM(preshade_volume_adjoint)(const DvrTensor* volumeT, const DvrTensor* tfT, int32_t mode,
                                               const DvrTensor* adjColorT, const DvrTensor* adjVolumeT, void*) {
  return guarded([&] {
    auto volume = make_acc<kernel::Tensor4Read, 4>(*volumeT);
    auto tf = make_acc<kernel::Tensor3Read, 3>(*tfT);
    auto adj_color = make_acc<kernel::Tensor4Read, 4>(*adjColorT);
    auto adj_volume = make_acc<kernel::Tensor4RW, 4>(*adjVolumeT);
    TF_SWITCH(mode, preshade_adj_impl<M>(volume, tf, adj_color, adj_volume));
  });
}
