// TEST INFRASTRUCTURE ONLY -- never linked into the product.
//
// CPU restatement of TFUtils::preshadeVolume and its density adjoint (renderer/tf_utils_impl.cu:16-33,
// 98-122): per voxel, the TF evaluation / adjoint of dvr_oracle.cpp (tf_eval, tf_adjoint) on batch entry 0.
// Included by dvr_oracle.cpp. Pinned by tests/test_oracle_tfutils.py against oracle/_ref
// (ref_shim_camera.cpp: the reference's own TransferFunctionEval<mode>::eval / adjoint) and the fixtures
// tests/golden/ref_preshade_*.npz.
#pragma once

namespace dvro {

template <typename R>
static void preshade_volume(const DvrTensor& volume, const DvrTensor& tf, int mode, const DvrTensor& color) {
  const Scene<R> s(tf, mode);
  const View<const R> v(volume, 4, false);
  const View<R> c(color, 4, false);
  for (int64_t x = 0; x < v.size[1]; ++x)
    for (int64_t y = 0; y < v.size[2]; ++y)
      for (int64_t z = 0; z < v.size[3]; ++z) {
        const V4<R> col = tf_eval(s, 0, v.at(0, x, y, z));
        c.at(0, x, y, z) = col.x; c.at(1, x, y, z) = col.y; c.at(2, x, y, z) = col.z; c.at(3, x, y, z) = col.w;
      }
}

template <typename R>
static void preshade_volume_adjoint(const DvrTensor& volume, const DvrTensor& tf, int mode, const DvrTensor& adjColor,
                                    const DvrTensor& adjVolume) {
  const Scene<R> s(tf, mode);
  const View<const R> v(volume, 4, false), ac(adjColor, 4, false);
  const View<R> av(adjVolume, 4, false);
  auto none = [](int, int, R) {};
  for (int64_t x = 0; x < v.size[1]; ++x)
    for (int64_t y = 0; y < v.size[2]; ++y)
      for (int64_t z = 0; z < v.size[3]; ++z) {
        const V4<R> a = {ac.at(0, x, y, z), ac.at(1, x, y, z), ac.at(2, x, y, z), ac.at(3, x, y, z)};
        av.at(0, x, y, z) = tf_adjoint(s, 0, v.at(0, x, y, z), a, false, none);
      }
}

}  // namespace dvro
