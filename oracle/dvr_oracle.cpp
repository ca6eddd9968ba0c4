// TEST INFRASTRUCTURE ONLY.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library; the product (diffdvr_b200/) never imports, links or executes it.
//
// CPU restatement ("port") of the hot path of shamanDevel/DiffDVR: the per-ray algorithms of
// renderer/renderer_kernels.cuh (forward :27-90, forward-mode :161-255, adjoint :324-444,
// forward-variables->gradients :504-561) and their building blocks
// (renderer_camera.cuh, renderer_interpolation.cuh, renderer_tf.cuh, renderer_blending.cuh,
// renderer_utils.cuh). Written from the behaviour of those files, operation order kept where
// it decides fp32 rounding, so that in fp32 it agrees with the reference's own host kernels
// (oracle/_ref, built from the reference sources) to round-off; tests/test_oracle_*.py pin it
// against oracle/_ref, against the reference's golden vector unittests/NonConvexDensity.tsv
// and against the committed fixtures under tests/golden/.
//
// Parity status: PINNED (forward/adjoint against the reference's compiled host kernels and
// golden fixtures; forward-mode against NonConvexDensity.tsv + forward==adjoint consistency;
// the dual-number arithmetic itself restates the un-vendored module `cudad` @
// f00bc425e8a5abd84a0690a69f2b36535454849c, see dvr_oracle_dual.h).
//
// Exports the same plain-C interface as include/dvr_b200.h with HOST pointers:
//   dvro32_* : fp32 (real_t = float, the reference's default build, renderer_commons.cuh:13-17)
//   dvro64_* : fp64 (tensors hold doubles; used as ground truth for finite differences)
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "../include/dvr_b200.h"
#include "dvr_oracle_dual.h"

namespace dvro {

static thread_local std::string g_error;
// see include/dvr_b200.h "compat_stepsize_overwrite"; the oracle defaults to the REFERENCE
// behaviour (1) because that is what it restates.
static int g_compat_stepsize_overwrite = 1;

template <typename R> struct V3 { R x, y, z; };
template <typename R> struct V4 { R x, y, z, w; };

template <typename R> inline R rmin(R a, R b) { return a < b ? a : b; }  // helper_math.cuh:86-104
template <typename R> inline R rmax(R a, R b) { return a > b ? a : b; }
template <typename R> inline R clampr(R f, R a, R b) { return rmax(a, rmin(f, b)); }  // :1898-1901
inline int clampi(int f, int a, int b) { return std::max(a, std::min(f, b)); }
template <typename R> inline R lerp(R a, R b, R t) { return a + t * (b - a); }  // :1835-1838

// Strided view with the reference's broadcasting rule: size-1 dims get stride 0
// (renderer_tensor.cuh:140-144).
template <typename T>
struct View {
  T* p = nullptr;
  int64_t size[DVR_MAX_DIMS] = {0, 0, 0, 0, 0};
  int64_t stride[DVR_MAX_DIMS] = {0, 0, 0, 0, 0};
  View() {}
  View(const DvrTensor& t, int ndim, bool broadcast = true) {
    if (!t.ptr) return;
    if (t.ndim != ndim) throw std::runtime_error("tensor has the wrong number of dimensions");
    p = static_cast<T*>(t.ptr);
    for (int i = 0; i < ndim; ++i) {
      size[i] = t.size[i];
      stride[i] = (broadcast && t.size[i] == 1) ? 0 : t.stride[i];
    }
  }
  T& at(int64_t a) const { return p[a * stride[0]]; }
  T& at(int64_t a, int64_t b) const { return p[a * stride[0] + b * stride[1]]; }
  T& at(int64_t a, int64_t b, int64_t c) const { return p[a * stride[0] + b * stride[1] + c * stride[2]]; }
  T& at(int64_t a, int64_t b, int64_t c, int64_t d) const {
    return p[a * stride[0] + b * stride[1] + c * stride[2] + d * stride[3]];
  }
  T& at(int64_t a, int64_t b, int64_t c, int64_t d, int64_t e) const {
    return p[a * stride[0] + b * stride[1] + c * stride[2] + d * stride[3] + e * stride[4]];
  }
};

inline int bcast(int B, const DvrTensor& t) {  // pytorch_utils.h:12-19
  if (!t.ptr) return B;
  const int b2 = int(t.size[0]);
  if (b2 == 1) return B;
  if (B == 1) return b2;
  if (b2 != B) throw std::runtime_error("batch size mismatch");
  return B;
}

// Inputs resolved like Renderer::checkInputs (renderer.cpp:49-263).
template <typename R>
struct Scene {
  int W, H, B, X, Y, Z;
  int filter, camera, tfMode, blend;
  View<R> volume;  // (B)*X*Y*Z
  View<R> boxMin, boxSize, rayStart, rayDir, camMat, step, tf;
  R boxMinV[3], boxSizeV[3], stepV[1], matV[16];
  R fovY, earlyOut;
  int R_, C_;  // tf resolution and channels

  // TF only (preshade_volume): tf_eval / tf_adjoint read tfMode, tf, R_, C_
  Scene(const DvrTensor& tfT, int mode) {
    W = H = X = Y = Z = 0;
    B = 1;
    filter = camera = blend = 0;
    tfMode = mode;
    tf = View<R>(tfT, 3);
    if (!tf.p) throw std::runtime_error("tf missing");
    R_ = int(tfT.size[1]);
    C_ = int(tfT.size[2]);
    fovY = earlyOut = R(0);
  }

  explicit Scene(const DvrInputs& a) {
    W = a.screen_width;
    H = a.screen_height;
    filter = a.volume_filter_mode;
    camera = a.camera_mode;
    tfMode = a.tf_mode;
    blend = a.blend_mode;
    if (!a.volume.ptr || a.volume.ndim != 4) throw std::runtime_error("volume must be a 4-D tensor");
    if ((filter == DVR_FILTER_PRESHADED) != (tfMode == DVR_TF_PRESHADED))
      throw std::runtime_error(
          "if pre-shading is selected, both the filter mode and TF mode must be set to preshaded");
    volume = View<R>(a.volume, 4, filter != DVR_FILTER_PRESHADED);
    B = filter == DVR_FILTER_PRESHADED ? 1 : int(a.volume.size[0]);
    X = int(a.volume.size[1]);
    Y = int(a.volume.size[2]);
    Z = int(a.volume.size[3]);
    DvrTensor t;
    if (a.box_min.ptr) { boxMin = View<R>(a.box_min, 2); B = bcast(B, a.box_min); }
    else { for (int i = 0; i < 3; ++i) boxMinV[i] = R(a.box_min_value[i]); t = DvrTensor{boxMinV, 2, {1, 3}, {3, 1}}; boxMin = View<R>(t, 2); }
    if (a.box_size.ptr) { boxSize = View<R>(a.box_size, 2); B = bcast(B, a.box_size); }
    else { for (int i = 0; i < 3; ++i) boxSizeV[i] = R(a.box_size_value[i]); t = DvrTensor{boxSizeV, 2, {1, 3}, {3, 1}}; boxSize = View<R>(t, 2); }
    switch (camera) {
      case DVR_CAMERA_RAY_START_DIR:
        rayStart = View<R>(a.camera_ray_start, 4);
        rayDir = View<R>(a.camera_ray_dir, 4);
        if (!rayStart.p || !rayDir.p) throw std::runtime_error("per-pixel rays missing");
        B = bcast(B, a.camera_ray_start);
        B = bcast(B, a.camera_ray_dir);
        break;
      case DVR_CAMERA_INVERSE_VIEW_MATRIX:
        if (a.camera_matrix.ptr) { camMat = View<R>(a.camera_matrix, 3); B = bcast(B, a.camera_matrix); }
        else { for (int i = 0; i < 16; ++i) matV[i] = R(a.camera_matrix_value[i]); t = DvrTensor{matV, 3, {1, 4, 4}, {16, 4, 1}}; camMat = View<R>(t, 3); }
        break;
      case DVR_CAMERA_REFERENCE_FRAME:
        camMat = View<R>(a.camera_matrix, 3);
        if (!camMat.p) throw std::runtime_error("camera viewport missing");
        B = bcast(B, a.camera_matrix);
        break;
      default: throw std::runtime_error("unknown camera enum value");
    }
    fovY = R(a.camera_fov_y_radians);
    if (a.step_size.ptr) { step = View<R>(a.step_size, 3); B = bcast(B, a.step_size); }
    else { stepV[0] = R(a.step_size_value); t = DvrTensor{stepV, 3, {1, 1, 1}, {1, 1, 1}}; step = View<R>(t, 3); }
    tf = View<R>(a.tf, 3);
    if (!tf.p) throw std::runtime_error("tf missing");
    B = bcast(B, a.tf);
    R_ = int(a.tf.size[1]);
    C_ = int(a.tf.size[2]);
    switch (tfMode) {  // renderer.cpp:226-247
      case DVR_TF_IDENTITY: if (R_ != 1 || C_ != 2) throw std::runtime_error("identity tf must be (B)*1*2"); break;
      case DVR_TF_TEXTURE: if (C_ != 4) throw std::runtime_error("texture tf must be (B)*R*4"); break;
      case DVR_TF_LINEAR: if (C_ != 5 || R_ < 2) throw std::runtime_error("linear tf must be (B)*R*5 with R>1"); break;
      case DVR_TF_GAUSSIAN: case DVR_TF_GAUSSIAN_LOG: if (C_ != 6) throw std::runtime_error("gaussian tf must be (B)*R*6"); break;
      case DVR_TF_PRESHADED: break;
      default: throw std::runtime_error("unknown tf enum value");
    }
    if (blend != DVR_BLEND_BEER_LAMBERT && blend != DVR_BLEND_ALPHA) throw std::runtime_error("unknown blend mode");
    if (filter != DVR_FILTER_NEAREST && filter != DVR_FILTER_TRILINEAR && filter != DVR_FILTER_PRESHADED)
      throw std::runtime_error("unsupported volume filter mode");
    earlyOut = R(a.blending_early_out);
  }
};

template <typename R>
struct Ray {
  V3<R> start, dir, boxMin, voxelSize;
  R step, tmin, tmax;
};

// CameraEval::computeRays (renderer_camera.cuh:15-60) + ray/box setup (renderer_kernels.cuh:31-48)
template <typename R>
static Ray<R> setup_ray(const Scene<R>& s, int x, int y, int b) {
  Ray<R> r;
  r.boxMin = {s.boxMin.at(b, 0), s.boxMin.at(b, 1), s.boxMin.at(b, 2)};
  const V3<R> boxSize = {s.boxSize.at(b, 0), s.boxSize.at(b, 1), s.boxSize.at(b, 2)};
  r.voxelSize = {boxSize.x / R(s.X - 1), boxSize.y / R(s.Y - 1), boxSize.z / R(s.Z - 1)};
  if (s.camera == DVR_CAMERA_RAY_START_DIR) {
    r.start = {s.rayStart.at(b, y, x, 0), s.rayStart.at(b, y, x, 1), s.rayStart.at(b, y, x, 2)};
    r.dir = {s.rayDir.at(b, y, x, 0), s.rayDir.at(b, y, x, 1), s.rayDir.at(b, y, x, 2)};
  } else if (s.camera == DVR_CAMERA_INVERSE_VIEW_MATRIX) {
    const R fx = 2 * (x + R(0.5)) / R(s.W) - 1;
    const R fy = 2 * (y + R(0.5)) / R(s.H) - 1;
    auto proj = [&](R z) {  // matmul_proj, renderer_utils.cuh:43-58
      R w[4];
      for (int i = 0; i < 4; ++i)  // dot(float4,float4) = x*x + y*y + z*z + w*w, left to right
        w[i] = s.camMat.at(b, i, 0) * fx + s.camMat.at(b, i, 1) * fy + s.camMat.at(b, i, 2) * z +
               s.camMat.at(b, i, 3) * R(1);
      return V3<R>{w[0] / w[3], w[1] / w[3], w[2] / w[3]};
    };
    r.start = proj(R(0));
    const V3<R> end = proj(R(0.9));
    const V3<R> d = {end.x - r.start.x, end.y - r.start.y, end.z - r.start.z};
    const R inv = R(1) / std::sqrt(d.x * d.x + d.y * d.y + d.z * d.z);  // helper_math.cuh:74-82,2237-2241
    r.dir = {d.x * inv, d.y * inv, d.z * inv};
  } else {
    const R fovX = s.fovY * R(s.W) / R(s.H);  // the ANGLE is scaled, not its tangent
    const R tanX = std::tan(fovX / 2), tanY = std::tan(s.fovY / 2);
    const V3<R> eye = {s.camMat.at(b, 0, 0), s.camMat.at(b, 0, 1), s.camMat.at(b, 0, 2)};
    const V3<R> right = {s.camMat.at(b, 1, 0), s.camMat.at(b, 1, 1), s.camMat.at(b, 1, 2)};
    const V3<R> up = {s.camMat.at(b, 2, 0), s.camMat.at(b, 2, 1), s.camMat.at(b, 2, 2)};
    const V3<R> front = {up.y * right.z - up.z * right.y, up.z * right.x - up.x * right.z,
                         up.x * right.y - up.y * right.x};  // cross(up, right)
    // the reference writes (x + 0.5f): float in the fp32 build, promoted to double in fp64
    const R fx = 2 * (x + R(0.5)) / R(s.W) - 1;
    const R fy = 2 * (y + R(0.5)) / R(s.H) - 1;
    const R ax = fx * tanX, ay = fy * tanY;
    const V3<R> d = {front.x + ax * right.x + ay * up.x, front.y + ax * right.y + ay * up.y,
                     front.z + ax * right.z + ay * up.z};
    const R inv = R(1) / std::sqrt(d.x * d.x + d.y * d.y + d.z * d.z);
    r.start = eye;
    r.dir = {d.x * inv, d.y * inv, d.z * inv};
  }
  r.step = s.step.at(b, y, x);
  // intersectionRayAABB (renderer_utils.cuh:60-74)
  const V3<R> inv = {R(1) / r.dir.x, R(1) / r.dir.y, R(1) / r.dir.z};
  const R t1 = (r.boxMin.x - r.start.x) * inv.x, t2 = (r.boxMin.x + boxSize.x - r.start.x) * inv.x;
  const R t3 = (r.boxMin.y - r.start.y) * inv.y, t4 = (r.boxMin.y + boxSize.y - r.start.y) * inv.y;
  const R t5 = (r.boxMin.z - r.start.z) * inv.z, t6 = (r.boxMin.z + boxSize.z - r.start.z) * inv.z;
  r.tmin = rmax(rmax(rmin(t1, t2), rmin(t3, t4)), rmin(t5, t6));
  r.tmax = rmin(rmin(rmax(t1, t2), rmax(t3, t4)), rmax(t5, t6));
  r.tmin = rmax(r.tmin, R(0));
  return r;
}

template <typename R>
inline V3<R> volume_pos(const Ray<R>& r, R t) {  // renderer_kernels.cuh:67-72
  const V3<R> w = {r.start.x + t * r.dir.x, r.start.y + t * r.dir.y, r.start.z + t * r.dir.z};
  return {(w.x - r.boxMin.x) / r.voxelSize.x, (w.y - r.boxMin.y) / r.voxelSize.y,
          (w.z - r.boxMin.z) / r.voxelSize.z};
}

// ---- volume sampling (renderer_interpolation.cuh) -------------------------------------------
// NC = 1 (density) or 4 (pre-shaded RGBA held in the batch dimension, :393-579).
template <typename R, int NC>
struct Sampler {
  const Scene<R>& s;
  int b;
  View<R> adjVolume;
  bool hasVolumeAdjoint;
  int last[3] = {-1, -1, -1};
  bool haveCell = false;  // forward cache key includes nothing else (one batch per ray)
  R dens[NC][8];
  R adjDens[NC][8];
  R nearest = 0, adjNearest = 0;
  V3<R> fpos{0, 0, 0};

  Sampler(const Scene<R>& s_, int b_, const View<R>& adj = View<R>(), bool hasAdj = false)
      : s(s_), b(b_), adjVolume(adj), hasVolumeAdjoint(hasAdj) {
    for (int c = 0; c < NC; ++c) for (int i = 0; i < 8; ++i) dens[c][i] = adjDens[c][i] = 0;
  }
  static void atomic_add(R* p, R v) {
#pragma omp atomic
    *p += v;
  }
  void corners(const int ip[3], int lo[3], int hi[3]) const {
    const int sz[3] = {s.X - 1, s.Y - 1, s.Z - 1};
    for (int i = 0; i < 3; ++i) {
      lo[i] = clampi(ip[i], 0, sz[i]);
      hi[i] = clampi(ip[i] + 1, 0, sz[i]);
    }
  }
  // flush of the per-cell accumulated adjoints (:299-321 trilinear, :95-106 nearest)
  void finish_adjoint() {
    if (!hasVolumeAdjoint) return;
    if (s.filter == DVR_FILTER_NEAREST) {
      if (adjNearest != 0) {
        atomic_add(&adjVolume.at(b, last[0], last[1], last[2]), adjNearest);
        adjNearest = 0;
      }
      return;
    }
    if (last[0] >= 0) {  // "not in uninitialized state" (:302)
      int lo[3], hi[3];
      corners(last, lo, hi);
      for (int c = 0; c < NC; ++c) {
        const int bb = NC == 1 ? b : c;
        for (int i = 0; i < 8; ++i) {  // corner index bit2 = x, bit1 = y, bit0 = z
          const int xx = (i & 4) ? hi[0] : lo[0], yy = (i & 2) ? hi[1] : lo[1], zz = (i & 1) ? hi[2] : lo[2];
          atomic_add(&adjVolume.at(bb, xx, yy, zz), adjDens[c][i]);
          adjDens[c][i] = 0;
        }
      }
    }
  }
  // returns NC channels
  void fetch(const V3<R>& pos, R out[NC]) {
    if (s.filter == DVR_FILTER_NEAREST) {  // :38-50, :111-132
      const R eps = R(1e-7);
      // clamp(v, EPS, size - (1 + EPS)) then roundf (round half away from zero)
      const R cx = clampr(pos.x, eps, R(s.X) - (1 + eps));
      const R cy = clampr(pos.y, eps, R(s.Y) - (1 + eps));
      const R cz = clampr(pos.z, eps, R(s.Z) - (1 + eps));
      const int ip[3] = {int(std::round(cx)), int(std::round(cy)), int(std::round(cz))};
      if (ip[0] != last[0] || ip[1] != last[1] || ip[2] != last[2]) {
        finish_adjoint();
        last[0] = ip[0]; last[1] = ip[1]; last[2] = ip[2];
        nearest = s.volume.at(b, ip[0], ip[1], ip[2]);
      }
      out[0] = nearest;
      return;
    }
    const int ip[3] = {int(pos.x), int(pos.y), int(pos.z)};  // truncation, helper_math.cuh:245-248
    if (ip[0] != last[0] || ip[1] != last[1] || ip[2] != last[2]) {
      finish_adjoint();
      last[0] = ip[0]; last[1] = ip[1]; last[2] = ip[2];
      int lo[3], hi[3];
      corners(ip, lo, hi);
      for (int c = 0; c < NC; ++c) {
        const int bb = NC == 1 ? b : c;
        for (int i = 0; i < 8; ++i) {
          const int xx = (i & 4) ? hi[0] : lo[0], yy = (i & 2) ? hi[1] : lo[1], zz = (i & 1) ? hi[2] : lo[2];
          dens[c][i] = s.volume.at(bb, xx, yy, zz);
        }
      }
    }
    fpos = {pos.x - R(ip[0]), pos.y - R(ip[1]), pos.z - R(ip[2])};  // unclamped fraction
    for (int c = 0; c < NC; ++c) {
      const R* d = dens[c];
      out[c] = lerp(lerp(lerp(d[0], d[4], fpos.x), lerp(d[2], d[6], fpos.x), fpos.y),
                    lerp(lerp(d[1], d[5], fpos.x), lerp(d[3], d[7], fpos.x), fpos.y), fpos.z);
    }
  }
  // adjoint of fetch at the same position (:360-387; pre-shaded :548-578)
  void adjoint(const R adjOut[NC], V3<R>& adjPos) {
    if (s.filter == DVR_FILTER_NEAREST) {
      adjNearest += adjOut[0];
      adjPos = {0, 0, 0};
      return;
    }
    const R cw[3] = {fpos.x, fpos.y, fpos.z};
    const R fw[3] = {1 - fpos.x, 1 - fpos.y, 1 - fpos.z};
    R adj_cw[3] = {0, 0, 0}, adj_fw[3] = {0, 0, 0};
    // the reference visits the corners in this order: 000,100,010,110,001,101,011,111 (xyz bits)
    static const int order[8] = {0, 4, 2, 6, 1, 5, 3, 7};
    for (int c = 0; c < NC; ++c) {
      const R a = adjOut[c];
      for (int k = 0; k < 8; ++k) {
        const int i = order[k];
        const bool hx = i & 4, hy = i & 2, hz = i & 1;
        const R wx = hx ? cw[0] : fw[0], wy = hy ? cw[1] : fw[1], wz = hz ? cw[2] : fw[2];
        adjDens[c][i] += a * (wx * wy * wz);
        (hx ? adj_cw : adj_fw)[0] += a * dens[c][i] * wy * wz;
        (hy ? adj_cw : adj_fw)[1] += a * dens[c][i] * wx * wz;
        if (NC == 1) {
          (hz ? adj_cw : adj_fw)[2] += a * dens[c][i] * wx * wy;
        } else {
          // reference typo (renderer_interpolation.cuh:561): the z-derivative is added to
          // adj_<Y>.z, i.e. selected by the corner's Y weight instead of its Z weight
          (hy ? adj_cw : adj_fw)[2] += a * dens[c][i] * wx * wy;
        }
      }
    }
    adjPos = {adj_cw[0] - adj_fw[0], adj_cw[1] - adj_fw[1], adj_cw[2] - adj_fw[2]};
  }
};

// ---- transfer functions (renderer_tf.cuh) ----------------------------------------------------
template <typename R>
inline V4<R> tf_row(const Scene<R>& s, int b, int i) {
  return {s.tf.at(b, i, 0), s.tf.at(b, i, 1), s.tf.at(b, i, 2), s.tf.at(b, i, 3)};
}
template <typename R>
inline R normal_pdf(R d, R mu, R sigma) {  // :492-496
  return std::exp(-(d - mu) * (d - mu) / (2 * sigma * sigma));
}
template <typename R>
inline R log_normal(R d, R mu, R sigma) {  // :657-661
  return -(d - mu) * (d - mu) / (2 * sigma * sigma);
}

template <typename R>
static V4<R> tf_eval(const Scene<R>& s, int b, R density) {
  switch (s.tfMode) {
    case DVR_TF_IDENTITY: {  // :62-73
      density = clampr(density, R(0), R(1));
      const R sa = s.tf.at(b, 0, 0), sc = s.tf.at(b, 0, 1);
      return {density * sc, density * sc, density * sc, density * sa};
    }
    case DVR_TF_TEXTURE: {  // :160-171
      const int Rn = s.R_;
      const R d = density * Rn - R(0.5);
      const int di = int(std::floor(float(d)));  // floorf on a float argument
      const R df = d - di;
      const V4<R> v0 = tf_row(s, b, clampi(di, 0, Rn - 1)), v1 = tf_row(s, b, clampi(di + 1, 0, Rn - 1));
      return {lerp(v0.x, v1.x, df), lerp(v0.y, v1.y, df), lerp(v0.z, v1.z, df), lerp(v0.w, v1.w, df)};
    }
    case DVR_TF_LINEAR: {  // :294-313
      int i;
      for (i = 0; i < s.R_ - 2; ++i)
        if (s.tf.at(b, i + 1, 4) > density) break;
      const V4<R> v0 = tf_row(s, b, i), v1 = tf_row(s, b, i + 1);
      const R p0 = s.tf.at(b, i, 4), p1 = s.tf.at(b, i + 1, 4);
      density = clampr(density, p0, p1);
      const R frac = (density - p0) / (p1 - p0);
      return {lerp(v0.x, v1.x, frac), lerp(v0.y, v1.y, frac), lerp(v0.z, v1.z, frac), lerp(v0.w, v1.w, frac)};
    }
    case DVR_TF_GAUSSIAN: {  // :507-519
      V4<R> c{0, 0, 0, 0};
      for (int i = 0; i < s.R_; ++i) {
        const V4<R> ci = tf_row(s, b, i);
        const R ni = normal_pdf(density, s.tf.at(b, i, 4), s.tf.at(b, i, 5));
        c.x += ci.x * ni; c.y += ci.y * ni; c.z += ci.z * ni; c.w += ci.w * ni;
      }
      return c;
    }
    case DVR_TF_GAUSSIAN_LOG: {  // :679-690 with logSumExp (renderer_utils.cuh:163-175)
      auto value = [&](int i) {
        const V4<R> ci = tf_row(s, b, i);
        const R ln = log_normal(density, s.tf.at(b, i, 4), s.tf.at(b, i, 5));
        return V4<R>{std::log(ci.x) + ln, std::log(ci.y) + ln, std::log(ci.z) + ln, std::log(ci.w) + ln};
      };
      V4<R> mx = value(0);
      for (int i = 1; i < s.R_; ++i) {
        const V4<R> v = value(i);
        mx = {rmax(mx.x, v.x), rmax(mx.y, v.y), rmax(mx.z, v.z), rmax(mx.w, v.w)};
      }
      V4<R> v0 = value(0);
      V4<R> sum = {std::exp(v0.x - mx.x), std::exp(v0.y - mx.y), std::exp(v0.z - mx.z), std::exp(v0.w - mx.w)};
      for (int i = 1; i < s.R_; ++i) {
        const V4<R> v = value(i);
        sum.x += std::exp(v.x - mx.x); sum.y += std::exp(v.y - mx.y);
        sum.z += std::exp(v.z - mx.z); sum.w += std::exp(v.w - mx.w);
      }
      return {mx.x + std::log(sum.x), mx.y + std::log(sum.y), mx.z + std::log(sum.z), mx.w + std::log(sum.w)};
    }
  }
  throw std::runtime_error("unknown tf enum value");
}

template <typename R>
inline R dot4(const V4<R>& a, const V4<R>& b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }

// TF adjoint; contributions to adj_tf go through `acc(r, c, value)`.
template <typename R, typename Acc>
static R tf_adjoint(const Scene<R>& s, int b, R density, const V4<R>& adjColor, bool hasTF, Acc&& acc) {
  switch (s.tfMode) {
    case DVR_TF_IDENTITY: {  // :75-104
      const R dc = clampr(density, R(0), R(1));
      const R sa = s.tf.at(b, 0, 0), sc = s.tf.at(b, 0, 1);
      const R adjDc = (adjColor.x * sc + adjColor.y * sc + adjColor.z * sc) + adjColor.w * sa;
      if (hasTF) {
        const R adjSc = adjColor.x * dc + adjColor.y * dc + adjColor.z * dc;
        const R adjSa = adjColor.w * dc;
        acc(0, 0, adjSa);
        acc(0, 1, adjSc);
      }
      return (density > 0 && density < 1) ? adjDc : R(0);
    }
    case DVR_TF_TEXTURE: {  // :173-220
      const int Rn = s.R_;
      const R d = density * Rn - R(0.5);
      const int di = int(std::floor(float(d)));
      const R df = d - di;
      const int i0 = clampi(di, 0, Rn - 1), i1 = clampi(di + 1, 0, Rn - 1);
      const V4<R> v0 = tf_row(s, b, i0), v1 = tf_row(s, b, i1);
      const V4<R> diff = {v1.x - v0.x, v1.y - v0.y, v1.z - v0.z, v1.w - v0.w};
      const R adjDf = dot4(adjColor, diff);
      if (hasTF) {
        const R w0 = 1 - df;
        acc(i0, 0, adjColor.x * w0); acc(i0, 1, adjColor.y * w0); acc(i0, 2, adjColor.z * w0); acc(i0, 3, adjColor.w * w0);
        acc(i1, 0, adjColor.x * df); acc(i1, 1, adjColor.y * df); acc(i1, 2, adjColor.z * df); acc(i1, 3, adjColor.w * df);
      }
      return adjDf * Rn;
    }
    case DVR_TF_LINEAR: {  // :315-400
      int i;
      for (i = 0; i < s.R_ - 2; ++i)
        if (s.tf.at(b, i + 1, 4) > density) break;
      const V4<R> v0 = tf_row(s, b, i), v1 = tf_row(s, b, i + 1);
      const R p0 = s.tf.at(b, i, 4), p1 = s.tf.at(b, i + 1, 4);
      const R d2 = clampr(density, p0, p1);
      const R frac = (d2 - p0) / (p1 - p0);
      const V4<R> diff = {v1.x - v0.x, v1.y - v0.y, v1.z - v0.z, v1.w - v0.w};
      const R adjFrac = dot4(adjColor, diff);
      const R denom = (p1 - p0) * (p1 - p0);
      const R adjP0 = adjFrac * (d2 - p1) / denom;
      const R adjP1 = adjFrac * (p0 - d2) / denom;
      const R adjD2 = adjFrac / (p1 - p0);
      if (hasTF) {
        const R w0 = 1 - frac;
        acc(i, 0, adjColor.x * w0); acc(i, 1, adjColor.y * w0); acc(i, 2, adjColor.z * w0); acc(i, 3, adjColor.w * w0);
        acc(i, 4, adjP0);
        acc(i + 1, 0, adjColor.x * frac); acc(i + 1, 1, adjColor.y * frac); acc(i + 1, 2, adjColor.z * frac); acc(i + 1, 3, adjColor.w * frac);
        acc(i + 1, 4, adjP1);
      }
      return adjD2;  // passed through the clamp unconditionally ("#if 1", :350-352)
    }
    case DVR_TF_GAUSSIAN: {  // :562-614
      R adjDensity = 0;
      for (int i = 0; i < s.R_; ++i) {
        const V4<R> ci = tf_row(s, b, i);
        const R mu = s.tf.at(b, i, 4), sigma = s.tf.at(b, i, 5);
        const R sigma2 = sigma * sigma;
        const R ni = normal_pdf(density, mu, sigma);
        const R f1 = (mu - density) / sigma2 * ni;
        adjDensity += dot4(adjColor, V4<R>{ci.x * f1, ci.y * f1, ci.z * f1, ci.w * f1});
        if (hasTF) {
          const R nf1 = -f1;
          const R adjMu = dot4(adjColor, V4<R>{ci.x * nf1, ci.y * nf1, ci.z * nf1, ci.w * nf1});
          const float f2 = float(f1 * (mu - density) / sigma);  // "const float f2" (:591)
          const R adjSigma = dot4(adjColor, V4<R>{ci.x * f2, ci.y * f2, ci.z * f2, ci.w * f2});
          acc(i, 0, adjColor.x * ni); acc(i, 1, adjColor.y * ni); acc(i, 2, adjColor.z * ni); acc(i, 3, adjColor.w * ni);
          acc(i, 4, adjMu);
          acc(i, 5, adjSigma);
        }
      }
      return adjDensity;
    }
    case DVR_TF_GAUSSIAN_LOG: {  // :705-780
      R adjDensity = 0;
      auto tval = [&](int i) {
        const V4<R> ci = tf_row(s, b, i);
        const R ln = log_normal(density, s.tf.at(b, i, 4), s.tf.at(b, i, 5));
        return V4<R>{std::log(ci.x) + ln, std::log(ci.y) + ln, std::log(ci.z) + ln, std::log(ci.w) + ln};
      };
      for (int i = 0; i < s.R_; ++i) {
        const V4<R> ci = tf_row(s, b, i);
        const R mu = s.tf.at(b, i, 4), sigma = s.tf.at(b, i, 5);
        const R sigma2 = sigma * sigma;
        const V4<R> ti = tval(i);
        V4<R> denom{1, 1, 1, 1};
        for (int j = 0; j < s.R_; ++j)
          if (j != i) {
            const V4<R> tj = tval(j);
            denom.x += std::exp(tj.x - ti.x); denom.y += std::exp(tj.y - ti.y);
            denom.z += std::exp(tj.z - ti.z); denom.w += std::exp(tj.w - ti.w);
          }
        const V4<R> adjTi = {adjColor.x / denom.x, adjColor.y / denom.y, adjColor.z / denom.z, adjColor.w / denom.w};
        const R f1 = (mu - density) / sigma2;
        const R adjXmu = adjTi.x * f1 + adjTi.y * f1 + adjTi.z * f1 + adjTi.w * f1;  // sum(adj_ti * f1)
        adjDensity += adjXmu;
        if (hasTF) {
          const float f2 = float(f1 * (mu - density) / sigma);
          const R adjSigma = adjTi.x * f2 + adjTi.y * f2 + adjTi.z * f2 + adjTi.w * f2;
          acc(i, 0, adjTi.x / ci.x); acc(i, 1, adjTi.y / ci.y); acc(i, 2, adjTi.z / ci.z); acc(i, 3, adjTi.w / ci.w);
          acc(i, 4, -adjXmu);
          acc(i, 5, adjSigma);
        }
      }
      return adjDensity;
    }
  }
  throw std::runtime_error("unknown tf enum value");
}

// ---- blending (renderer_blending.cuh) ----------------------------------------------------------
template <typename R>
inline V4<R> blend(int mode, const V4<R>& acc, const V4<R>& cur, R step) {  // :54-61, :118-125
  const R alpha = mode == DVR_BLEND_BEER_LAMBERT ? 1 - R(std::exp(-cur.w * step)) : rmin(R(1), cur.w * step);
  const R om = 1 - acc.w;
  return {acc.x + om * cur.x * alpha, acc.y + om * cur.y * alpha, acc.z + om * cur.z * alpha, acc.w + om * alpha};
}

// Inverts the blend and returns the accumulator before the sample; :63-98 and :128-164.
template <typename R>
inline V4<R> blend_adjoint(int mode, const V4<R>& accOut, const V4<R>& cur, R step, const V4<R>& adjOut,
                           V4<R>& adjIn, V4<R>& adjCur, R& adjStep) {
  const bool bl = mode == DVR_BLEND_BEER_LAMBERT;
  // Alpha mode clamps to 1-1e-5 here although the forward pass clamps to 1 (:136)
  const R alpha = bl ? 1 - R(std::exp(-cur.w * step)) : rmin(R(1 - R(1e-5)), cur.w * step);
  const R alphaIn = (alpha - accOut.w) / (alpha - 1);
  const R k = 1 - alphaIn;
  const V4<R> accIn = {accOut.x - k * cur.x * alpha, accOut.y - k * cur.y * alpha, accOut.z - k * cur.z * alpha, alphaIn};
  R adjAlpha = adjOut.w * (1 - alphaIn);
  R adjInAlpha = adjOut.w * (1 - alpha);
  adjAlpha += adjOut.x * (cur.x - cur.x * alphaIn) + adjOut.y * (cur.y - cur.y * alphaIn) + adjOut.z * (cur.z - cur.z * alphaIn);
  adjInAlpha += adjOut.x * (-cur.x * alpha) + adjOut.y * (-cur.y * alpha) + adjOut.z * (-cur.z * alpha);
  const R kc = alpha * (1 - alphaIn);
  R adjW;
  if (bl) {
    const R e = R(std::exp(-cur.w * step));
    adjW = adjAlpha * step * e;
    adjStep = adjAlpha * cur.w * e;
  } else {
    const bool active = cur.w * step < 1;
    adjStep = active ? adjAlpha * cur.w : R(0);
    adjW = active ? adjAlpha * step : R(0);
  }
  adjCur = {adjOut.x * kc, adjOut.y * kc, adjOut.z * kc, adjW};
  adjIn = {adjOut.x, adjOut.y, adjOut.z, adjInAlpha};
  return accIn;
}

// ---- forward (renderer_kernels.cuh:27-90) ---------------------------------------------------------
template <typename R>
static void render_forward(const DvrInputs& in, const DvrOutputs& out) {
  Scene<R> s(in);
  View<R> color(out.color, 4, false);
  View<int32_t> term(out.termination_index, 3, false);
  if (!color.p || !term.p) throw std::runtime_error("outputs missing");
  const int64_t count = int64_t(s.W) * s.H * s.B;
#pragma omp parallel for schedule(guided)
  for (int64_t idx = 0; idx < count; ++idx) {
    const int b = int(idx / (int64_t(s.W) * s.H));
    const int y = int((idx - int64_t(b) * s.W * s.H) / s.W);
    const int x = int(idx - int64_t(s.W) * (y + int64_t(s.H) * b));
    const Ray<R> r = setup_ray(s, x, y, b);
    V4<R> acc{0, 0, 0, 0};
    int i;
    if (s.filter == DVR_FILTER_PRESHADED) {
      Sampler<R, 4> smp(s, b);
      for (i = 0;; ++i) {
        const R t = r.tmin + i * r.step;
        if (t > r.tmax) break;
        if (acc.w > s.earlyOut) break;
        R c[4];
        smp.fetch(volume_pos(r, t), c);
        acc = blend(s.blend, acc, V4<R>{c[0], c[1], c[2], c[3]}, r.step);
      }
    } else {
      Sampler<R, 1> smp(s, b);
      for (i = 0;; ++i) {
        const R t = r.tmin + i * r.step;
        if (t > r.tmax) break;
        if (acc.w > s.earlyOut) break;
        R d[1];
        smp.fetch(volume_pos(r, t), d);
        acc = blend(s.blend, acc, tf_eval(s, b, d[0]), r.step);
      }
    }
    color.at(b, y, x, 0) = acc.x;
    color.at(b, y, x, 1) = acc.y;
    color.at(b, y, x, 2) = acc.z;
    color.at(b, y, x, 3) = acc.w;
    term.at(b, y, x) = i;
  }
}

// ---- adjoint (renderer_kernels.cuh:324-444) ---------------------------------------------------------
template <typename R>
static void render_adjoint(const DvrInputs& in, const DvrOutputs& fwd, const DvrTensor& adjColorT,
                           const DvrAdjointOutputs& ao) {
  Scene<R> s(in);
  if (!ao.has_volume_derivatives && !ao.has_stepsize_derivatives && !ao.has_camera_derivatives &&
      !ao.has_tf_derivatives)
    throw std::runtime_error("At least one adjoint variable must be activated!");
  View<R> color(fwd.color, 4), adjColor(adjColorT, 4);
  View<int32_t> term(fwd.termination_index, 3);
  const bool hasVol = ao.has_volume_derivatives, hasStep = ao.has_stepsize_derivatives,
             hasCam = ao.has_camera_derivatives, hasTF = ao.has_tf_derivatives;
  View<R> adjVolume = hasVol ? View<R>(ao.adj_volume, 4, s.filter != DVR_FILTER_PRESHADED) : View<R>();
  View<R> adjStep = hasStep ? View<R>(ao.adj_step_size, 3) : View<R>();
  View<R> adjStart = hasCam ? View<R>(ao.adj_camera_ray_start, 4) : View<R>();
  View<R> adjDir = hasCam ? View<R>(ao.adj_camera_ray_dir, 4) : View<R>();
  View<R> adjTf = hasTF ? View<R>(ao.adj_tf, 3) : View<R>();
  // renderer.cpp:303-306, 335-337
  const bool stepBroadcast = hasStep && (s.B != ao.adj_step_size.size[0] || s.H != ao.adj_step_size.size[1] || s.W != ao.adj_step_size.size[2]);
  const bool camBroadcast = hasCam && (s.B != ao.adj_camera_ray_start.size[0] || s.B != ao.adj_camera_ray_dir.size[0]);
  const bool compat = g_compat_stepsize_overwrite != 0;
  auto atomic_add = [](R* p, R v) {
#pragma omp atomic
    *p += v;
  };
  const int64_t count = int64_t(s.W) * s.H * s.B;
#pragma omp parallel for schedule(guided)
  for (int64_t idx = 0; idx < count; ++idx) {
    const int b = int(idx / (int64_t(s.W) * s.H));
    const int y = int((idx - int64_t(b) * s.W * s.H) / s.W);
    const int x = int(idx - int64_t(s.W) * (y + int64_t(s.H) * b));
    const Ray<R> r = setup_ray(s, x, y, b);
    V4<R> acc = {color.at(b, y, x, 0), color.at(b, y, x, 1), color.at(b, y, x, 2), color.at(b, y, x, 3)};
    V4<R> adjAcc = {adjColor.at(b, y, x, 0), adjColor.at(b, y, x, 1), adjColor.at(b, y, x, 2), adjColor.at(b, y, x, 3)};
    R adjStepSize = 0;
    V3<R> adjRayStart{0, 0, 0}, adjRayDir{0, 0, 0};
    auto tfAcc = [&](int rr, int cc, R v) { atomic_add(&adjTf.at(b, rr, cc), v); };
    auto body = [&](auto& smp, auto nc) {
      constexpr int NC = decltype(nc)::value;
      for (int i = term.at(b, y, x) - 1; i >= 0; --i) {
        const R t = r.tmin + i * r.step;
        const V3<R> pos = volume_pos(r, t);
        R d[NC];
        smp.fetch(pos, d);
        V4<R> cur;
        if constexpr (NC == 1) cur = tf_eval(s, b, d[0]);
        else cur = {d[0], d[1], d[2], d[3]};
        V4<R> adjCur, adjIn;
        R adjStepBlend;
        acc = blend_adjoint(s.blend, acc, cur, r.step, adjAcc, adjIn, adjCur, adjStepBlend);
        adjAcc = adjIn;
        // F4: the reference ASSIGNS the blend term to its running accumulator every sample
        // (kernels.cuh:393-395 passes adj_stepSize as the out-parameter that blending.cuh:93,158
        // overwrites); the complete gradient accumulates it.
        if (compat) adjStepSize = adjStepBlend;
        else adjStepSize += adjStepBlend;
        R adjD[NC];
        if constexpr (NC == 1) adjD[0] = tf_adjoint(s, b, d[0], adjCur, hasTF, tfAcc);
        else { adjD[0] = adjCur.x; adjD[1] = adjCur.y; adjD[2] = adjCur.z; adjD[3] = adjCur.w; }
        V3<R> adjPos;
        smp.adjoint(adjD, adjPos);
        const V3<R> adjWorld = {adjPos.x / r.voxelSize.x, adjPos.y / r.voxelSize.y, adjPos.z / r.voxelSize.z};
        adjRayStart.x += adjWorld.x; adjRayStart.y += adjWorld.y; adjRayStart.z += adjWorld.z;
        adjRayDir.x += adjWorld.x * t; adjRayDir.y += adjWorld.y * t; adjRayDir.z += adjWorld.z * t;
        const R adjT = adjWorld.x * r.dir.x + adjWorld.y * r.dir.y + adjWorld.z * r.dir.z;
        adjStepSize += adjT * i;
      }
      smp.finish_adjoint();  // destructor flush (:324)
    };
    if (s.filter == DVR_FILTER_PRESHADED) {
      Sampler<R, 4> smp(s, b, adjVolume, hasVol);
      body(smp, std::integral_constant<int, 4>());
    } else {
      Sampler<R, 1> smp(s, b, adjVolume, hasVol);
      body(smp, std::integral_constant<int, 1>());
    }
    if (hasStep) {
      if (stepBroadcast) atomic_add(&adjStep.at(b, y, x), adjStepSize);
      else adjStep.at(b, y, x) = adjStepSize;
    }
    if (hasCam) {
      const R vs[3] = {adjRayStart.x, adjRayStart.y, adjRayStart.z}, vd[3] = {adjRayDir.x, adjRayDir.y, adjRayDir.z};
      for (int c = 0; c < 3; ++c) {
        if (camBroadcast) { atomic_add(&adjStart.at(b, y, x, c), vs[c]); atomic_add(&adjDir.at(b, y, x, c), vd[c]); }
        else { adjStart.at(b, y, x, c) = vs[c]; adjDir.at(b, y, x, c) = vd[c]; }
      }
    }
  }
}

// ---- forward-mode (renderer_kernels.cuh:161-255) -----------------------------------------------------
template <typename R>
struct D4 { Dual<R> x, y, z, w; };

template <typename R>
static D4<R> tf_eval_dual(const Scene<R>& s, int b, Dual<R> density, const View<int32_t>& dtf, bool hasTF, int D) {
  auto seed = [&](int r, int c) { return Dual<R>::input(s.tf.at(b, r, c), hasTF ? dtf.at(b, r, c) : -1, D); };
  auto row = [&](int r) { return D4<R>{seed(r, 0), seed(r, 1), seed(r, 2), seed(r, 3)}; };
  auto cst = [&](R v) { return Dual<R>(v, D); };
  switch (s.tfMode) {
    case DVR_TF_IDENTITY: {  // :121-154
      density = dclamp(density, cst(0), cst(1));
      const Dual<R> sa = seed(0, 0), sc = seed(0, 1);
      return {density * sc, density * sc, density * sc, density * sa};
    }
    case DVR_TF_TEXTURE: {  // :255-288
      const int Rn = s.R_;
      const Dual<R> d = density * R(Rn) - R(0.5);
      const int di = int(std::floor(float(d.v)));
      const Dual<R> df = d - R(di);
      const D4<R> v0 = row(clampi(di, 0, Rn - 1)), v1 = row(clampi(di + 1, 0, Rn - 1));
      return {dlerp(v0.x, v1.x, df), dlerp(v0.y, v1.y, df), dlerp(v0.z, v1.z, df), dlerp(v0.w, v1.w, df)};
    }
    case DVR_TF_LINEAR: {  // :436-485
      int i;
      for (i = 0; i < s.R_ - 2; ++i)
        if (s.tf.at(b, i + 1, 4) > density.v) break;
      const D4<R> v0 = row(i), v1 = row(i + 1);
      const Dual<R> p0 = seed(i, 4), p1 = seed(i + 1, 4);
      const Dual<R> d2 = dclamp(density, p0, p1);
      const Dual<R> frac = (d2 - p0) / (p1 - p0);
      return {dlerp(v0.x, v1.x, frac), dlerp(v0.y, v1.y, frac), dlerp(v0.z, v1.z, frac), dlerp(v0.w, v1.w, frac)};
    }
    case DVR_TF_GAUSSIAN: {  // :521-560
      D4<R> c{cst(0), cst(0), cst(0), cst(0)};
      for (int i = 0; i < s.R_; ++i) {
        const D4<R> ci = row(i);
        const Dual<R> mu = seed(i, 4), sigma = seed(i, 5);
        const Dual<R> ni = dexp(-(density - mu) * (density - mu) / (R(2) * sigma * sigma));
        c = {c.x + ci.x * ni, c.y + ci.y * ni, c.z + ci.z * ni, c.w + ci.w * ni};
      }
      return c;
    }
  }
  throw std::runtime_error("forward-mode: unsupported tf mode");
}

template <typename R>
static void render_forward_gradients(const DvrInputs& in, const DvrForwardDifferencesSettings& fs,
                                     const DvrOutputs& out, const DvrTensor& gradT) {
  Scene<R> s(in);
  if (s.filter == DVR_FILTER_PRESHADED) throw std::runtime_error("forward-mode: preshaded volumes unsupported");
  const int D = fs.D;
  if (D <= 0) throw std::runtime_error("number of derivatives must be positive");
  const bool hasStep = fs.d_stepsize >= 0;
  const bool hasCam = fs.d_ray_start[0] >= 0 || fs.d_ray_start[1] >= 0 || fs.d_ray_start[2] >= 0 ||
                      fs.d_ray_dir[0] >= 0 || fs.d_ray_dir[1] >= 0 || fs.d_ray_dir[2] >= 0;
  const bool hasTF = fs.has_tf_derivatives != 0;
  if (!hasStep && !hasCam && !hasTF) throw std::runtime_error("at least one derivative must be activated");
  View<int32_t> dtf = hasTF ? View<int32_t>(fs.d_tf, 3) : View<int32_t>();
  if (hasTF && !dtf.p) throw std::runtime_error("d_tf not defined, hence hasTfDerivatives must be false but was set to true");
  View<R> color(out.color, 4, false), grad(gradT, 5, false);
  View<int32_t> term(out.termination_index, 3, false);
  const int64_t count = int64_t(s.W) * s.H * s.B;
#pragma omp parallel for schedule(guided)
  for (int64_t idx = 0; idx < count; ++idx) {
    const int b = int(idx / (int64_t(s.W) * s.H));
    const int y = int((idx - int64_t(b) * s.W * s.H) / s.W);
    const int x = int(idx - int64_t(s.W) * (y + int64_t(s.H) * b));
    const Ray<R> r = setup_ray(s, x, y, b);  // box intersection on VALUES only (:201-208)
    auto cst = [&](R v) { return Dual<R>(v, D); };
    const Dual<R> sx = Dual<R>::input(r.start.x, fs.d_ray_start[0], D), sy = Dual<R>::input(r.start.y, fs.d_ray_start[1], D),
                  sz = Dual<R>::input(r.start.z, fs.d_ray_start[2], D);
    const Dual<R> dx = Dual<R>::input(r.dir.x, fs.d_ray_dir[0], D), dy = Dual<R>::input(r.dir.y, fs.d_ray_dir[1], D),
                  dz = Dual<R>::input(r.dir.z, fs.d_ray_dir[2], D);
    const Dual<R> step = Dual<R>::input(r.step, fs.d_stepsize, D);
    D4<R> acc{cst(0), cst(0), cst(0), cst(0)};
    int last[3] = {-1, -1, -1};
    R dens[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int i;
    for (i = 0;; ++i) {
      const Dual<R> t = r.tmin + R(i) * step;
      if (t.v > r.tmax) break;
      if (acc.w.v > s.earlyOut) break;
      const Dual<R> px = ((sx + t * dx) - r.boxMin.x) / r.voxelSize.x;
      const Dual<R> py = ((sy + t * dy) - r.boxMin.y) / r.voxelSize.y;
      const Dual<R> pz = ((sz + t * dz) - r.boxMin.z) / r.voxelSize.z;
      Dual<R> density;
      if (s.filter == DVR_FILTER_NEAREST) {  // constant: no positional tangent (interpolation.cuh:52-65)
        const R eps = R(1e-7);
        const int ix = int(std::round(clampr(px.v, eps, R(s.X) - (1 + eps))));
        const int iy = int(std::round(clampr(py.v, eps, R(s.Y) - (1 + eps))));
        const int iz = int(std::round(clampr(pz.v, eps, R(s.Z) - (1 + eps))));
        density = cst(s.volume.at(b, ix, iy, iz));
      } else {  // interpolation.cuh:210-240
        const int ip[3] = {int(px.v), int(py.v), int(pz.v)};
        if (ip[0] != last[0] || ip[1] != last[1] || ip[2] != last[2]) {
          last[0] = ip[0]; last[1] = ip[1]; last[2] = ip[2];
          const int sz3[3] = {s.X - 1, s.Y - 1, s.Z - 1};
          int lo[3], hi[3];
          for (int k = 0; k < 3; ++k) { lo[k] = clampi(ip[k], 0, sz3[k]); hi[k] = clampi(ip[k] + 1, 0, sz3[k]); }
          for (int k = 0; k < 8; ++k)
            dens[k] = s.volume.at(b, (k & 4) ? hi[0] : lo[0], (k & 2) ? hi[1] : lo[1], (k & 1) ? hi[2] : lo[2]);
        }
        const Dual<R> fx = px - R(ip[0]), fy = py - R(ip[1]), fz = pz - R(ip[2]);
        auto l = [&](R a, R bb, const Dual<R>& tt) { return a + tt * (bb - a); };
        density = dlerp(dlerp(l(dens[0], dens[4], fx), l(dens[2], dens[6], fx), fy),
                        dlerp(l(dens[1], dens[5], fx), l(dens[3], dens[7], fx), fy), fz);
      }
      const D4<R> cur = tf_eval_dual(s, b, density, dtf, hasTF, D);
      // Blending::blend generic (renderer_blending.cuh:40-51, 105-116)
      const Dual<R> alpha = s.blend == DVR_BLEND_BEER_LAMBERT ? R(1) - dexp(-cur.w * step) : dmin(cst(1), cur.w * step);
      const Dual<R> om = R(1) - acc.w;
      acc = {acc.x + om * cur.x * alpha, acc.y + om * cur.y * alpha, acc.z + om * cur.z * alpha, acc.w + om * alpha};
    }
    const Dual<R>* ch[4] = {&acc.x, &acc.y, &acc.z, &acc.w};
    for (int c = 0; c < 4; ++c) {
      color.at(b, y, x, c) = ch[c]->v;
      for (int d = 0; d < D; ++d) grad.at(b, y, x, d, c) = ch[c]->d[d];
    }
    term.at(b, y, x) = i;
  }
}

// ---- forward variables -> gradients (renderer_kernels.cuh:504-561) -----------------------------------
template <typename R>
static void forward_variables_to_gradients(const DvrTensor& fvT, const DvrTensor& gcT,
                                           const DvrForwardDifferencesSettings& fs, const DvrAdjointOutputs& ao) {
  if (ao.has_volume_derivatives) throw std::runtime_error("volume derivatives are not supported with forward variables");
  View<R> fv(fvT, 5, false), gc(gcT, 4);
  if (!fv.p || !gc.p) throw std::runtime_error("inputs missing");
  const int B = int(fvT.size[0]), H = int(fvT.size[1]), W = int(fvT.size[2]);
  View<R> adjStep = ao.has_stepsize_derivatives ? View<R>(ao.adj_step_size, 3) : View<R>();
  View<R> adjStart = ao.has_camera_derivatives ? View<R>(ao.adj_camera_ray_start, 4) : View<R>();
  View<R> adjDir = ao.has_camera_derivatives ? View<R>(ao.adj_camera_ray_dir, 4) : View<R>();
  View<R> adjTf = ao.has_tf_derivatives ? View<R>(ao.adj_tf, 3) : View<R>();
  View<int32_t> dtf = fs.has_tf_derivatives ? View<int32_t>(fs.d_tf, 3) : View<int32_t>();
  const int64_t count = int64_t(W) * H * B;
#pragma omp parallel for schedule(guided)
  for (int64_t idx = 0; idx < count; ++idx) {
    const int b = int(idx / (int64_t(W) * H));
    const int y = int((idx - int64_t(b) * W * H) / W);
    const int x = int(idx - int64_t(W) * (y + int64_t(H) * b));
    const V4<R> g = {gc.at(b, y, x, 0), gc.at(b, y, x, 1), gc.at(b, y, x, 2), gc.at(b, y, x, 3)};
    auto contract = [&](int d) {
      return g.x * fv.at(b, y, x, d, 0) + g.y * fv.at(b, y, x, d, 1) + g.z * fv.at(b, y, x, d, 2) + g.w * fv.at(b, y, x, d, 3);
    };
    if (fs.d_stepsize >= 0) adjStep.at(b, y, x) = contract(fs.d_stepsize);  // STORED, not accumulated
    for (int c = 0; c < 3; ++c) {
      if (fs.d_ray_start[c] >= 0) adjStart.at(b, y, x, c) = contract(fs.d_ray_start[c]);
      if (fs.d_ray_dir[c] >= 0) adjDir.at(b, y, x, c) = contract(fs.d_ray_dir[c]);
    }
    if (fs.has_tf_derivatives) {
      const int Rn = int(fs.d_tf.size[1]), Cn = int(fs.d_tf.size[2]);
      for (int rr = 0; rr < Rn; ++rr)
        for (int cc = 0; cc < Cn; ++cc) {
          const int d = dtf.at(b, rr, cc);
          if (d >= 0) {
            const R v = contract(d);
#pragma omp atomic
            adjTf.at(b, rr, cc) += v;
          }
        }
    }
  }
}

// ---- test diagnostic: distance of a ray's samples to the nearest cell face --------------------------
// d(trilinear)/d(position) is discontinuous where the integer cell changes (renderer_interpolation.cuh
// :221-228: ipos = int(volumePos), corners clamped to size-1), i.e. at integer coordinates 1..size-1 of
// every axis. A sample within round-off of such a face flips cells between ANY two fp32 evaluation
// orders (the reference's own fp32 and fp64 builds disagree on such rays; rays that enter the box
// through a "max" face always start exactly on one). The parity tests use this per-ray minimum (in
// voxels, over the samples 0..terminationIndex-1) to separate face-sensitive rays from the rest.
template <typename R>
static void face_distance(const DvrInputs& in, const DvrTensor& termT, const DvrTensor& outT) {
  Scene<R> s(in);
  View<int32_t> term(termT, 3);
  View<R> out(outT, 3, false);
  const int64_t count = int64_t(s.W) * s.H * s.B;
  const int size[3] = {s.X, s.Y, s.Z};
#pragma omp parallel for schedule(guided)
  for (int64_t idx = 0; idx < count; ++idx) {
    const int b = int(idx / (int64_t(s.W) * s.H));
    const int y = int((idx - int64_t(b) * s.W * s.H) / s.W);
    const int x = int(idx - int64_t(s.W) * (y + int64_t(s.H) * b));
    const Ray<R> r = setup_ray(s, x, y, b);
    const int n = term.at(b, y, x);
    R best = R(1);
    for (int i = 0; i < n; ++i) {
      const V3<R> p = volume_pos(r, r.tmin + i * r.step);
      const R pc[3] = {p.x, p.y, p.z};
      for (int a = 0; a < 3; ++a) {
        const R k = clampr(std::round(pc[a]), R(1), R(size[a] - 1));
        best = rmin(best, std::abs(pc[a] - k));
      }
    }
    out.at(b, y, x) = best;
  }
}

}  // namespace dvro
#include "dvr_oracle_camera.h"
#include "dvr_oracle_tfutils.h"
#include "dvr_oracle_volume.h"
#include "dvr_oracle_fn.h"
namespace dvro {

template <typename F>
static int guarded(F&& f) {
  try {
    f();
    return DVR_OK;
  } catch (const std::exception& e) {
    g_error = e.what();
    return DVR_ERR_INVALID_ARGUMENT;
  }
}

}  // namespace dvro

using namespace dvro;

#define DVRO_EXPORTS(P, R)                                                                          \
  extern "C" int dvro##P##_render_forward(const DvrInputs* i, const DvrOutputs* o, void*) {         \
    return guarded([&] { render_forward<R>(*i, *o); });                                             \
  }                                                                                                 \
  extern "C" int dvro##P##_render_forward_gradients(const DvrInputs* i,                             \
                                                    const DvrForwardDifferencesSettings* s,        \
                                                    const DvrOutputs* o, const DvrTensor* g,       \
                                                    void*) {                                        \
    return guarded([&] { render_forward_gradients<R>(*i, *s, *o, *g); });                           \
  }                                                                                                 \
  extern "C" int dvro##P##_forward_variables_to_gradients(const DvrTensor* fv, const DvrTensor* gc, \
                                                          const DvrForwardDifferencesSettings* s,  \
                                                          const DvrAdjointOutputs* ao, void*) {    \
    return guarded([&] { forward_variables_to_gradients<R>(*fv, *gc, *s, *ao); });                  \
  }                                                                                                 \
  extern "C" int dvro##P##_render_adjoint(const DvrInputs* i, const DvrOutputs* f,                  \
                                          const DvrTensor* ac, const DvrAdjointOutputs* ao,        \
                                          void*) {                                                  \
    return guarded([&] { render_adjoint<R>(*i, *f, *ac, *ao); });                                   \
  }                                                                                                 \
  extern "C" int dvro##P##_face_distance(const DvrInputs* i, const DvrTensor* t, const DvrTensor* o) { \
    return guarded([&] { face_distance<R>(*i, *t, *o); });                                          \
  }                                                                                                 \
  extern "C" int dvro##P##_generate_rays(const DvrTensor* vp, float fovY, int32_t W, int32_t H,     \
                                         const DvrTensor* rs, const DvrTensor* rd, void*) {         \
    return guarded([&] { generate_rays<R>(*vp, fovY, W, H, *rs, *rd); });                           \
  }                                                                                                 \
  extern "C" int dvro##P##_generate_rays_adjoint(const DvrTensor* vp, float fovY,                   \
                                                 const DvrTensor* as, const DvrTensor* ad,         \
                                                 const DvrTensor* av, void*) {                      \
    return guarded([&] { generate_rays_adjoint<R>(*vp, fovY, *as, *ad, *av); });                    \
  }                                                                                                 \
  extern "C" int dvro##P##_compare_to_image(const DvrTensor* c, const DvrTensor* g,                 \
                                            const DvrTensor* q, const DvrTensor* d,                \
                                            const DvrTensor* go, int32_t reduce, void*) {           \
    return guarded([&] { compare_to_image<R>(*c, *g, *q, *d, *go, reduce); });                      \
  }                                                                                                 \
  extern "C" int dvro##P##_preshade_volume(const DvrTensor* v, const DvrTensor* tf, int32_t mode,   \
                                           const DvrTensor* c, void*) {                             \
    return guarded([&] { preshade_volume<R>(*v, *tf, mode, *c); });                                 \
  }                                                                                                 \
  extern "C" int dvro##P##_preshade_volume_adjoint(const DvrTensor* v, const DvrTensor* tf,         \
                                                   int32_t mode, const DvrTensor* ac,               \
                                                   const DvrTensor* av, void*) {                    \
    return guarded([&] { preshade_volume_adjoint<R>(*v, *tf, mode, *ac, *av); });                   \
  }                                                                                                 \
  extern "C" int dvro##P##_create_implicit_volume(int32_t eq, const float* params,                  \
                                                  const DvrTensor* v, void*) {                      \
    return guarded([&] { create_implicit_volume<R>(eq, params, *v); });                             \
  }                                                                                                 \
  extern "C" int dvro##P##_volume_from_raw(const void* raw, int32_t dtype, const DvrTensor* v,      \
                                           void*) {                                                 \
    return guarded([&] { volume_from_raw<R>(raw, dtype, *v); });                                    \
  }                                                                                                 \
  extern "C" int dvro##P##_volume_to_raw(const DvrTensor* v, float* raw, void*) {                   \
    return guarded([&] { volume_to_raw<R>(*v, raw); });                                             \
  }                                                                                                 \
  extern "C" int dvro##P##_downsample_volume(const DvrTensor* v, const DvrTensor* o, void*) {       \
    return guarded([&] { downsample_volume<R>(*v, *o); });                                          \
  }                                                                                                 \
  extern "C" int dvro##P##_fn_blend(int32_t mode, int32_t n, const R* x, R* out) {                  \
    return guarded([&] { fn_blend<R>(mode, n, x, out); });                                          \
  }                                                                                                 \
  extern "C" int dvro##P##_fn_blend_adjoint(int32_t mode, int32_t n, const R* x, const R* e,        \
                                            const R* g, R* z, R* accIn) {                           \
    return guarded([&] { fn_blend_adjoint<R>(mode, n, x, e, g, z, accIn); });                       \
  }                                                                                                 \
  extern "C" int dvro##P##_fn_tf(int32_t mode, int32_t Rn, int32_t C, const R* tf, int32_t n,       \
                                 const R* density, R* out) {                                        \
    return guarded([&] { fn_tf<R>(mode, Rn, C, tf, n, density, out); });                            \
  }                                                                                                 \
  extern "C" int dvro##P##_fn_tf_adjoint(int32_t mode, int32_t Rn, int32_t C, const R* tf,          \
                                         int32_t n, const R* density, const R* g, R* adjDensity,    \
                                         R* adjTf) {                                                \
    return guarded([&] { fn_tf_adjoint<R>(mode, Rn, C, tf, n, density, g, adjDensity, adjTf); });   \
  }                                                                                                 \
  extern "C" int dvro##P##_fn_interp(int32_t X, int32_t Y, int32_t Z, const R* volume, int32_t n,   \
                                     const R* pos, R* out) {                                        \
    return guarded([&] { fn_interp<R>(X, Y, Z, volume, n, pos, out); });                            \
  }                                                                                                 \
  extern "C" int dvro##P##_fn_interp_adjoint(int32_t X, int32_t Y, int32_t Z, const R* volume,      \
                                             int32_t n, const R* pos, const R* g, R* adjPos,        \
                                             R* adjVolume) {                                        \
    return guarded([&] { fn_interp_adjoint<R>(X, Y, Z, volume, n, pos, g, adjPos, adjVolume); });   \
  }                                                                                                 \
  extern "C" const char* dvro##P##_last_error() { return g_error.c_str(); }                         \
  extern "C" int dvro##P##_real_size() { return int(sizeof(R)); }

DVRO_EXPORTS(32, float)
DVRO_EXPORTS(64, double)

extern "C" int dvro_set_option(const char* name, int32_t value) {
  if (std::strcmp(name, "compat_stepsize_overwrite") == 0) {
    g_compat_stepsize_overwrite = value;
    return DVR_OK;
  }
  return DVR_ERR_INVALID_ARGUMENT;
}

// Known-answer fixture: the driver loop of unittests/testNonConvexDensity.cpp:8-80 restated over
// this file's Gaussian TF + Beer-Lambert blend with one tangent (d/d d1).
// out = {red, alpha, loss, dloss/dd1}; compared with unittests/NonConvexDensity.tsv.
extern "C" int dvro32_nonconvex_density(double d1_in, double* out) {
  typedef float R;
  return guarded([&] {
    const R mean = 0, sigma = R(0.5), stepsize = 1 / 32.f, d0 = -1;
    auto eval = [&](R d1v, Dual<R> acc[4]) {
      const int D = 1;
      for (int c = 0; c < 4; ++c) acc[c] = Dual<R>(0, D);
      const Dual<R> d1 = Dual<R>::input(d1v, 0, D);
      const int n = std::max(32, int(std::abs((d1v - d0) * stepsize)));
      const R local = 1.0f / n;
      for (int i = 0; i <= n; ++i) {
        const R f = i / R(n);
        const Dual<R> d = (1 - f) * d0 + f * d1;
        const Dual<R> ni = dexp(-(d - mean) * (d - mean) / (2 * sigma * sigma));
        const Dual<R> cur[4] = {R(1) * ni, R(0) * ni, R(0) * ni, R(1) * ni};  // (red,green,blue,alpha)=(1,0,0,1)
        const Dual<R> alpha = R(1) - dexp(-cur[3] * local);
        const Dual<R> om = R(1) - acc[3];
        for (int c = 0; c < 3; ++c) acc[c] = acc[c] + om * cur[c] * alpha;
        acc[3] = acc[3] + om * alpha;
      }
    };
    Dual<R> gt[4], c[4];
    eval(d0, gt);
    eval(R(d1_in), c);
    // loss = |gt - c| with d|v| = <v, dv> / |v|
    R v[4], dv[4], l2 = 0;
    for (int k = 0; k < 4; ++k) { v[k] = gt[k].v - c[k].v; dv[k] = -c[k].d[0]; l2 += v[k] * v[k]; }
    const R l = std::sqrt(l2);
    R dl = 0;
    for (int k = 0; k < 4; ++k) dl += v[k] * dv[k];
    out[0] = c[0].v; out[1] = c[3].v; out[2] = l; out[3] = dl / l;
  });
}
