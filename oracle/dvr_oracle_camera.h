// TEST INFRASTRUCTURE ONLY -- never linked into the product.
//
// CPU restatement of the callers either side of the marcher;
// included by dvr_oracle.cpp (namespace dvro, View<>, V3<>, V4<> come from there):
//   * Camera::generateRays forward  (renderer/camera_impl.cu:17-94)
//   * Camera::generateRays backward (renderer/camera_impl.cu:116-181 per-pixel term, :339-355
//     sequential sum per view; adjNormalize / adjCross renderer/renderer_adjoint.cuh:28-64)
//   * Renderer::compareToImage      (renderer/renderer_compareToImage.cuh:9-65,
//     renderer/renderer_cpu.cpp:156-243), derivative of the norm as in the restated
//     cudAD::length (oracle/cudad_restated/forward_vector.h): <v, dv> / |v|
// Pinned by tests/test_oracle_camera.py against oracle/_ref (ref_shim_camera.cpp: the same
// functions built on the reference's own CameraEval / adjNormalize / adjCross / CompareToImage
// headers) and by the committed fixtures tests/golden/ref_camera_*.npz.
#pragma once

namespace dvro {

template <typename R>
struct FrameCam {
  V3<R> eye, right, up, front;
  R tanX, tanY;
  int W, H;
};

template <typename R>
static FrameCam<R> frame_cam(const View<const R>& vp, int b, R fovY, int W, int H) {
  FrameCam<R> f;
  f.eye = {vp.at(b, 0, 0), vp.at(b, 0, 1), vp.at(b, 0, 2)};
  f.right = {vp.at(b, 1, 0), vp.at(b, 1, 1), vp.at(b, 1, 2)};
  f.up = {vp.at(b, 2, 0), vp.at(b, 2, 1), vp.at(b, 2, 2)};
  f.front = {f.up.y * f.right.z - f.up.z * f.right.y, f.up.z * f.right.x - f.up.x * f.right.z,
             f.up.x * f.right.y - f.up.y * f.right.x};
  // camera_impl.cu:76-78: float arithmetic whatever real_t is
  const float fovX = float(fovY) * float(W) / float(H);
  f.tanX = R(std::tan(fovX / 2));
  f.tanY = R(std::tan(float(fovY) / 2));
  f.W = W;
  f.H = H;
  return f;
}

template <typename R>
static V3<R> frame_dir(const FrameCam<R>& f, int x, int y, R& ax, R& ay) {
  const R fx = 2 * (x + 0.5f) / R(f.W) - 1;
  const R fy = 2 * (y + 0.5f) / R(f.H) - 1;
  ax = fx * f.tanX;
  ay = fy * f.tanY;
  return {f.front.x + ax * f.right.x + ay * f.up.x, f.front.y + ax * f.right.y + ay * f.up.y,
          f.front.z + ax * f.right.z + ay * f.up.z};
}

template <typename R>
static void generate_rays(const DvrTensor& viewport, float fovY, int W, int H, const DvrTensor& start,
                          const DvrTensor& dir) {
  const View<const R> vp(viewport, 3, false);
  const View<R> rs(start, 4, false), rd(dir, 4, false);
  const int B = int(vp.size[0]);
  for (int b = 0; b < B; ++b) {
    const FrameCam<R> f = frame_cam<R>(vp, b, R(fovY), W, H);
    for (int y = 0; y < H; ++y)
      for (int x = 0; x < W; ++x) {
        R ax, ay;
        const V3<R> d = frame_dir(f, x, y, ax, ay);
        const R inv = R(1) / std::sqrt(d.x * d.x + d.y * d.y + d.z * d.z);
        rs.at(b, y, x, 0) = f.eye.x; rs.at(b, y, x, 1) = f.eye.y; rs.at(b, y, x, 2) = f.eye.z;
        rd.at(b, y, x, 0) = d.x * inv; rd.at(b, y, x, 1) = d.y * inv; rd.at(b, y, x, 2) = d.z * inv;
      }
  }
}

template <typename R>
static void generate_rays_adjoint(const DvrTensor& viewport, float fovY, const DvrTensor& adjStart,
                                  const DvrTensor& adjDir, const DvrTensor& adjViewport) {
  const View<const R> vp(viewport, 3, false), gs(adjStart, 4, false), gd(adjDir, 4, false);
  const View<R> out(adjViewport, 3, false);
  const int B = int(vp.size[0]), H = int(gs.size[1]), W = int(gs.size[2]);
  for (int b = 0; b < B; ++b) {
    const FrameCam<R> f = frame_cam<R>(vp, b, R(fovY), W, H);
    V3<R> sumEye = {0, 0, 0}, sumRight = {0, 0, 0}, sumUp = {0, 0, 0};
    for (int y = 0; y < H; ++y)
      for (int x = 0; x < W; ++x) {
        R ax, ay;
        const V3<R> v = frame_dir(f, x, y, ax, ay);
        const V3<R> g = {gd.at(b, y, x, 0), gd.at(b, y, x, 1), gd.at(b, y, x, 2)};
        // adjNormalize
        const R len = std::sqrt(v.x * v.x + v.y * v.y + v.z * v.z);
        V3<R> ad = {g.x / len, g.y / len, g.z / len};
        const R l2 = len * len;
        const R adjLen = g.x * (-v.x / l2) + g.y * (-v.y / l2) + g.z * (-v.z / l2);
        ad.x += adjLen * v.x / len; ad.y += adjLen * v.y / len; ad.z += adjLen * v.z / len;
        const V3<R> c = ad;  // adj_front
        V3<R> ar = {ad.x * ax, ad.y * ax, ad.z * ax};
        V3<R> au = {ad.x * ay, ad.y * ay, ad.z * ay};
        // adjCross(a = up, b = right, adj_c = adj_front, adj_a = adj_up, adj_b = adj_right)
        const V3<R>& A = f.up;
        const V3<R>& Bv = f.right;
        au.y += Bv.z * c.x; ar.z += A.y * c.x; au.z -= Bv.y * c.x; ar.y -= A.z * c.x;
        au.z += Bv.x * c.y; ar.x += A.z * c.y; au.x -= Bv.z * c.y; ar.z -= A.x * c.y;
        au.x += Bv.y * c.z; ar.y += A.x * c.z; au.y -= Bv.x * c.z; ar.x -= A.y * c.z;
        sumEye.x += gs.at(b, y, x, 0); sumEye.y += gs.at(b, y, x, 1); sumEye.z += gs.at(b, y, x, 2);
        sumRight.x += ar.x; sumRight.y += ar.y; sumRight.z += ar.z;
        sumUp.x += au.x; sumUp.y += au.y; sumUp.z += au.z;
      }
    out.at(b, 0, 0) = sumEye.x; out.at(b, 0, 1) = sumEye.y; out.at(b, 0, 2) = sumEye.z;
    out.at(b, 1, 0) = sumRight.x; out.at(b, 1, 1) = sumRight.y; out.at(b, 1, 2) = sumRight.z;
    out.at(b, 2, 0) = sumUp.x; out.at(b, 2, 1) = sumUp.y; out.at(b, 2, 2) = sumUp.z;
  }
}

template <typename R>
static void compare_to_image(const DvrTensor& color, const DvrTensor& grads, const DvrTensor& reference,
                             const DvrTensor& diffOut, const DvrTensor& gradsOut, int reduce) {
  const View<const R> c(color, 4, false), q(reference, 4, false), g(grads, 5, false);
  const int B = int(c.size[0]), H = int(c.size[1]), W = int(c.size[2]), D = int(g.size[3]);
  const R normalization = R(1) / R(B * W * H);
  std::vector<R> acc(D + 1, R(0));
  const View<R> d1(diffOut, reduce ? 1 : 4, false), g1(gradsOut, reduce ? 1 : 4, false);
  for (int b = 0; b < B; ++b)
    for (int y = 0; y < H; ++y)
      for (int x = 0; x < W; ++x) {
        R v[4];
        for (int k = 0; k < 4; ++k) v[k] = c.at(b, y, x, k) - q.at(b, y, x, k);
        const R len = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2] + v[3] * v[3]);
        if (reduce) acc[0] += len * normalization;
        else d1.at(b, y, x, 0) = len;
        for (int d = 0; d < D; ++d) {
          const R dv = (v[0] * g.at(b, y, x, d, 0) + v[1] * g.at(b, y, x, d, 1) + v[2] * g.at(b, y, x, d, 2) +
                        v[3] * g.at(b, y, x, d, 3)) / len;
          if (reduce) acc[1 + d] += dv * normalization;
          else g1.at(b, y, x, d) = dv;
        }
      }
  if (reduce) {
    d1.at(0) = acc[0];
    for (int d = 0; d < D; ++d) g1.at(d) = acc[1 + d];
  }
}

}  // namespace dvro
