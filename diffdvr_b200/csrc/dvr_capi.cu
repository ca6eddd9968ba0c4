// C-ABI of the B200 ray marcher (include/dvr_b200.h): validation, stride resolution and
// dispatch to the precompiled kernel instances. Plays the role of Renderer::checkInputs /
// checkOutput / checkAdjointOutput (renderer/renderer.cpp:49-391) for shapes and of
// RendererCuda (renderer/renderer_cuda.cpp:85-317) for launching. No torch types, no device
// allocations, no host<->device copies, no CPU fallback.
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstddef>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>

#include "../../include/dvr_b200.h"
#include "dvr_launch.h"
#include "dvr_kernels.cuh"  // kRegRows, TfLayout, TfChannels (host-side constants only)

namespace {

using namespace dvr;

thread_local std::string g_error;
thread_local int t_call_sync = 0;  // DvrCallOptions::sync of the call running on this thread
std::atomic<long long> g_launches{0}, g_quad_builds{0}, g_quad_reuses{0}, g_quad_fallbacks{0}, g_folds{0};
std::atomic<int> g_sync_mode{0};  // the one process-wide switch (Renderer::setCudaSyncMode)

// per-call options resolved to plain values (DvrCallOptions, include/dvr_b200.h)
struct CallCtx {
  int layout = -1;
  int compat = 0;
  long long cap = 16ll << 30;
  DvrWorkspace* ws = nullptr;
  bool fwdTfSlotMajor = false;
  bool marchZeroRays = false;
};

struct Fail {
  int code;
};

[[noreturn]] void fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_error = buf;
  throw Fail{code};
}

template <typename F>
int guarded(F&& f) {
  try {
    f();
    t_call_sync = 0;
    return DVR_OK;
  } catch (const Fail& e) {
    return e.code;
  } catch (const std::exception& e) {
    g_error = e.what();
    return DVR_ERR_INVALID_ARGUMENT;
  }
}

void cuda_check(cudaError_t e, const char* what) {
  // kErrNoInstance is the launchers' own sentinel (dvr_launch.h); CUDA never returns it
  if (e == kErrNoInstance) fail(DVR_ERR_UNSUPPORTED, "%s: no precompiled kernel instance for this combination", what);
  if (e != cudaSuccess) {
    cudaGetLastError();  // clear the sticky-less error state of the runtime
    fail(DVR_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
  }
}

void after_launch(cudaError_t e, const char* what, cudaStream_t s) {
  cuda_check(e, what);
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (g_sync_mode.load() || t_call_sync) cuda_check(cudaStreamSynchronize(s), what);  // Renderer::setCudaSyncMode
}

CallCtx resolve_options(const DvrCallOptions* o) {
  CallCtx c;
  t_call_sync = 0;
  if (!o) return c;
  // ABI 2.0 shipped a 32-byte struct (up to `workspace`); later fields are read only when struct_size covers them
  if (o->struct_size < offsetof(DvrCallOptions, workspace) + sizeof(void*))
    fail(DVR_ERR_INVALID_ARGUMENT, "DvrCallOptions::struct_size is %u, expected %zu (use dvr_call_options_init)",
         o->struct_size, sizeof(DvrCallOptions));
  if (o->struct_size >= offsetof(DvrCallOptions, forward_mode_tf) + sizeof(int32_t)) {
    if (o->forward_mode_tf < 0 || o->forward_mode_tf > 1)
      fail(DVR_ERR_INVALID_ARGUMENT, "forward_mode_tf must be 0 or 1, but is %d", o->forward_mode_tf);
    c.fwdTfSlotMajor = o->forward_mode_tf == 1;
  }
  if (o->struct_size >= offsetof(DvrCallOptions, adjoint_march_zero_rays) + sizeof(int32_t))
    c.marchZeroRays = o->adjoint_march_zero_rays != 0;
  if (o->volume_layout < -1 || o->volume_layout > 1)
    fail(DVR_ERR_INVALID_ARGUMENT, "volume_layout must be -1, 0 or 1, but is %d", o->volume_layout);
  c.layout = o->volume_layout;
  c.compat = o->compat_stepsize_overwrite != 0;
  t_call_sync = o->sync != 0;
  if (o->max_scratch_bytes > 0) c.cap = o->max_scratch_bytes;
  c.ws = o->workspace;
  return c;
}

// ---- shape checks, with the reference's wording (renderer/pytorch_utils.h:5-19) ---------------
void check_dim(const DvrTensor& t, const char* name, int d) {
  if (!t.ptr) fail(DVR_ERR_INVALID_ARGUMENT, "%s is not defined!", name);
  if (t.ndim != d) fail(DVR_ERR_INVALID_ARGUMENT, "%s must be a tensor with %d dimensions, but has %d", name, d, t.ndim);
}
void check_size(const DvrTensor& t, const char* name, int d, long long s) {
  if (t.size[d] != s)
    fail(DVR_ERR_INVALID_ARGUMENT, "%s must have %lld entries at dimension %d, but has %lld entries", name, s, d,
         (long long)t.size[d]);
}
int check_batch(const DvrTensor& t, const char* name, int B1, int dim = 0) {
  const int B2 = int(t.size[dim]);
  if (B2 == 1) return B1;
  if (B1 == 1) return B2;
  if (B1 != B2)
    fail(DVR_ERR_INVALID_ARGUMENT, "%s must be have the same batch size as other tensors, if batch>1. actual=%d, expected=%d",
         name, B2, B1);
  return B1;
}
long long bstride(const DvrTensor& t, int d) { return t.size[d] == 1 ? 0 : t.stride[d]; }

struct Checked {
  int B = 1, W = 0, H = 0, X = 0, Y = 0, Z = 0;
  RayInputs ri;
};

// Renderer::checkInputs (renderer/renderer.cpp:49-263)
void check_inputs(const DvrInputs& a, Checked& c) {
  RayInputs& r = c.ri;
  std::memset(&r, 0, sizeof(r));
  c.W = a.screen_width;
  c.H = a.screen_height;
  if (c.W <= 0 || c.H <= 0) fail(DVR_ERR_INVALID_ARGUMENT, "screenSize must be positive, but is (%d, %d)", c.W, c.H);
  check_dim(a.volume, "volume", 4);
  const bool preshaded = a.volume_filter_mode == DVR_FILTER_PRESHADED;
  if (preshaded != (a.tf_mode == DVR_TF_PRESHADED))
    fail(DVR_ERR_INVALID_ARGUMENT, "if pre-shading is selected, both the filter mode and TF mode must be set to preshaded");
  if (a.volume_filter_mode != DVR_FILTER_NEAREST && a.volume_filter_mode != DVR_FILTER_TRILINEAR && !preshaded)
    fail(DVR_ERR_UNSUPPORTED, "volume filter mode %d is not supported", a.volume_filter_mode);
  int B = int(a.volume.size[0]);
  if (preshaded) {  // renderer.cpp:78-83: the four colour channels sit in the batch dimension
    if (a.volume.size[0] != 4)
      fail(DVR_ERR_INVALID_ARGUMENT,
           "In pre-shaded mode, the volume colors are encoded in the batch dimension of the volume. It must be 4, but is %lld",
           (long long)a.volume.size[0]);
    B = 1;
  }
  c.X = int(a.volume.size[1]);
  c.Y = int(a.volume.size[2]);
  c.Z = int(a.volume.size[3]);
  if (c.X < 1 || c.Y < 1 || c.Z < 1) fail(DVR_ERR_INVALID_ARGUMENT, "volume must not be empty");
  {
    long long span = 0;
    for (int d = 1; d < 4; ++d) span += (a.volume.size[d] - 1) * bstride(a.volume, d);
    if (span >= (1ll << 31) || c.X >= (1 << 22) || c.Y >= (1 << 22) || c.Z >= (1 << 22))
      fail(DVR_ERR_UNSUPPORTED, "one volume batch entry must span fewer than 2^31 elements");
  }
  r.volume = static_cast<const float*>(a.volume.ptr);
  r.volB = preshaded ? a.volume.stride[0] : bstride(a.volume, 0);
  r.sx = int(bstride(a.volume, 1));
  r.sy = int(bstride(a.volume, 2));
  r.sz = int(bstride(a.volume, 3));
  r.X = c.X; r.Y = c.Y; r.Z = c.Z;

  if (a.box_min.ptr) {
    check_dim(a.box_min, "boxMin", 2);
    check_size(a.box_min, "boxMin", 1, 3);
    B = check_batch(a.box_min, "boxMin", B);
    r.boxMin = {static_cast<const float*>(a.box_min.ptr), bstride(a.box_min, 0), a.box_min.stride[1]};
  } else {
    for (int i = 0; i < 3; ++i) r.boxMinV[i] = a.box_min_value[i];
  }
  if (a.box_size.ptr) {
    check_dim(a.box_size, "boxSize", 2);
    check_size(a.box_size, "boxSize", 1, 3);
    B = check_batch(a.box_size, "boxSize", B);
    r.boxSize = {static_cast<const float*>(a.box_size.ptr), bstride(a.box_size, 0), a.box_size.stride[1]};
  } else {
    for (int i = 0; i < 3; ++i) r.boxSizeV[i] = a.box_size_value[i];
  }

  r.cameraMode = a.camera_mode;
  switch (a.camera_mode) {
    case DVR_CAMERA_RAY_START_DIR: {
      const DvrTensor* ts[2] = {&a.camera_ray_start, &a.camera_ray_dir};
      const char* names[2] = {"cameraRayStart", "cameraRayDir"};
      Strided4* dst[2] = {&r.rayStart, &r.rayDir};
      for (int k = 0; k < 2; ++k) {
        check_dim(*ts[k], names[k], 4);
        check_size(*ts[k], names[k], 1, c.H);
        check_size(*ts[k], names[k], 2, c.W);
        check_size(*ts[k], names[k], 3, 3);
        B = check_batch(*ts[k], names[k], B);
        *dst[k] = {static_cast<const float*>(ts[k]->ptr), bstride(*ts[k], 0), bstride(*ts[k], 1), bstride(*ts[k], 2),
                   ts[k]->stride[3]};
      }
    } break;
    case DVR_CAMERA_INVERSE_VIEW_MATRIX:
      if (a.camera_matrix.ptr) {
        check_dim(a.camera_matrix, "cameraMatrix", 3);
        check_size(a.camera_matrix, "cameraMatrix", 1, 4);
        check_size(a.camera_matrix, "cameraMatrix", 2, 4);
        B = check_batch(a.camera_matrix, "cameraMatrix", B);
        r.camMat = {static_cast<const float*>(a.camera_matrix.ptr), bstride(a.camera_matrix, 0),
                    a.camera_matrix.stride[1], a.camera_matrix.stride[2]};
      } else {
        for (int i = 0; i < 16; ++i) r.camMatV[i] = a.camera_matrix_value[i];
      }
      break;
    case DVR_CAMERA_REFERENCE_FRAME: {
      check_dim(a.camera_matrix, "cameraMatrix", 3);
      check_size(a.camera_matrix, "cameraMatrix", 1, 3);
      check_size(a.camera_matrix, "cameraMatrix", 2, 3);
      B = check_batch(a.camera_matrix, "cameraMatrix", B);
      r.camMat = {static_cast<const float*>(a.camera_matrix.ptr), bstride(a.camera_matrix, 0),
                  a.camera_matrix.stride[1], a.camera_matrix.stride[2]};
      // renderer_camera.cuh:42-45: the ANGLE is scaled by the aspect ratio, then halved
      const float fovX = a.camera_fov_y_radians * float(c.W) / float(c.H);
      r.tanFovX = std::tan(fovX / 2);
      r.tanFovY = std::tan(a.camera_fov_y_radians / 2);
    } break;
    default: fail(DVR_ERR_INVALID_ARGUMENT, "unknown camera enum value");
  }

  if (a.step_size.ptr) {
    check_dim(a.step_size, "stepSize", 3);
    B = check_batch(a.step_size, "stepSize", B);
    check_batch(a.step_size, "stepSize-Height", c.H, 1);
    check_batch(a.step_size, "stepSize-Width", c.W, 2);
    r.stepSize = {static_cast<const float*>(a.step_size.ptr), bstride(a.step_size, 0), bstride(a.step_size, 1),
                  bstride(a.step_size, 2)};
  } else {
    r.stepV = a.step_size_value;
  }

  check_dim(a.tf, "tf", 3);
  B = check_batch(a.tf, "tf", B);
  switch (a.tf_mode) {
    case DVR_TF_IDENTITY:
      check_size(a.tf, "tf", 1, 1);
      check_size(a.tf, "tf", 2, 2);
      break;
    case DVR_TF_TEXTURE: check_size(a.tf, "tf", 2, 4); break;
    case DVR_TF_LINEAR:
      check_size(a.tf, "tf", 2, 5);
      if (a.tf.size[1] <= 1)
        fail(DVR_ERR_INVALID_ARGUMENT, "tensor 'tf' must have at least two control points in TFLinear-mode");
      break;
    case DVR_TF_GAUSSIAN:
    case DVR_TF_GAUSSIAN_LOG: check_size(a.tf, "tf", 2, 6); break;
    case DVR_TF_PRESHADED: break;  // any size is allowed, the TF is ignored (renderer.cpp:240-242)
    default: fail(DVR_ERR_INVALID_ARGUMENT, "unknown tf enum value");
  }
  if (a.tf.size[1] < 1 && !preshaded) fail(DVR_ERR_INVALID_ARGUMENT, "tensor 'tf' must have at least one row");
  r.tf = {static_cast<const float*>(a.tf.ptr), bstride(a.tf, 0), a.tf.stride[1], a.tf.stride[2]};
  r.R = int(a.tf.size[1]);
  r.C = int(a.tf.size[2]);
  if (a.blend_mode != DVR_BLEND_BEER_LAMBERT && a.blend_mode != DVR_BLEND_ALPHA)
    fail(DVR_ERR_INVALID_ARGUMENT, "unknown blend mode");
  r.earlyOut = a.blending_early_out;
  if (B > 65535) fail(DVR_ERR_UNSUPPORTED, "batch size %d exceeds 65535", B);
  c.B = B;
  r.W = c.W; r.H = c.H; r.B = B;
}

// Renderer::checkOutput (renderer/renderer.cpp:367-391)
ForwardOutputs check_outputs(const DvrOutputs& o, const Checked& c) {
  check_dim(o.color, "color", 4);
  check_size(o.color, "color", 0, c.B);
  check_size(o.color, "color", 1, c.H);
  check_size(o.color, "color", 2, c.W);
  check_size(o.color, "color", 3, 4);
  check_dim(o.termination_index, "terminationIndex", 3);
  check_size(o.termination_index, "terminationIndex", 0, c.B);
  check_size(o.termination_index, "terminationIndex", 1, c.H);
  check_size(o.termination_index, "terminationIndex", 2, c.W);
  ForwardOutputs f;
  f.color = static_cast<float*>(o.color.ptr);
  f.cB = o.color.stride[0]; f.cY = o.color.stride[1]; f.cX = o.color.stride[2]; f.cC = o.color.stride[3];
  f.colorVec4 = f.cC == 1 && (reinterpret_cast<uintptr_t>(f.color) % 16 == 0) && f.cB % 4 == 0 && f.cY % 4 == 0 &&
                f.cX % 4 == 0;
  f.term = static_cast<int*>(o.termination_index.ptr);
  f.tB = o.termination_index.stride[0]; f.tY = o.termination_index.stride[1]; f.tX = o.termination_index.stride[2];
  return f;
}

dim3 tile_grid(const Checked& c) {
  return dim3((c.W + kCtaW - 1) / kCtaW, (c.H + kCtaH - 1) / kCtaH, c.B);
}

template <int TFMODE> constexpr int row_of() { return TfLayout<TFMODE>::kRow; }
int tf_row(int tf) { return tf == kTfIdentity ? 2 : tf == kTfTexture ? 4 : 8; }
int tf_channels(int tf) { return tf == kTfIdentity ? 2 : tf == kTfTexture ? 4 : tf == kTfLinear ? 5 : 6; }
size_t check_smem(size_t bytes) {
  if (bytes > 200 * 1024) fail(DVR_ERR_UNSUPPORTED, "transfer function too large for shared memory (%zu bytes)", bytes);
  return bytes;
}

// checkAdjointOutput (renderer/renderer.cpp:265-365); X<0 skips the volume checks
void check_adjoint_outputs(const DvrAdjointOutputs& o, int B, int H, int W, int X, int Y, int Z, const DvrTensor* tf,
                           AdjointArgs& aa, bool preshaded = false) {
  if (o.has_volume_derivatives) {
    check_dim(o.adj_volume, "adj_volume", 4);
    if (preshaded) check_size(o.adj_volume, "adj_volume", 0, 4);  // renderer.cpp:277-279
    else check_batch(o.adj_volume, "adj_volume", B);
    check_size(o.adj_volume, "adj_volume", 1, X);
    check_size(o.adj_volume, "adj_volume", 2, Y);
    check_size(o.adj_volume, "adj_volume", 3, Z);
    long long span = 0;
    for (int d = 1; d < 4; ++d) span += (o.adj_volume.size[d] - 1) * bstride(o.adj_volume, d);
    if (span >= (1ll << 31)) fail(DVR_ERR_UNSUPPORTED, "one adj_volume batch entry must span fewer than 2^31 elements");
    aa.adjVolume = static_cast<float*>(o.adj_volume.ptr);
    aa.avB = preshaded ? o.adj_volume.stride[0] : bstride(o.adj_volume, 0);
    aa.avx = int(bstride(o.adj_volume, 1));
    aa.avy = int(bstride(o.adj_volume, 2));
    aa.avz = int(bstride(o.adj_volume, 3));
  }
  aa.hasStep = o.has_stepsize_derivatives != 0;
  if (aa.hasStep) {
    check_dim(o.adj_step_size, "adj_stepSize", 3);
    check_batch(o.adj_step_size, "adj_stepSize", B);
    check_batch(o.adj_step_size, "adj_stepSize", H, 1);
    check_batch(o.adj_step_size, "adj_stepSize", W, 2);
    aa.adjStep = static_cast<float*>(o.adj_step_size.ptr);
    aa.asB = bstride(o.adj_step_size, 0); aa.asY = bstride(o.adj_step_size, 1); aa.asX = bstride(o.adj_step_size, 2);
    aa.stepAtomic = B != o.adj_step_size.size[0] || H != o.adj_step_size.size[1] || W != o.adj_step_size.size[2];
  }
  aa.hasCam = o.has_camera_derivatives != 0;
  if (aa.hasCam) {
    const DvrTensor* ts[2] = {&o.adj_camera_ray_start, &o.adj_camera_ray_dir};
    const char* names[2] = {"adj_cameraRayStart", "adj_cameraRayDir"};
    for (int k = 0; k < 2; ++k) {
      check_dim(*ts[k], names[k], 4);
      check_batch(*ts[k], names[k], B);
      check_size(*ts[k], names[k], 1, H);
      check_size(*ts[k], names[k], 2, W);
      check_size(*ts[k], names[k], 3, 3);
    }
    aa.adjRayStart = static_cast<float*>(ts[0]->ptr);
    aa.rsB = bstride(*ts[0], 0); aa.rsY = bstride(*ts[0], 1); aa.rsX = bstride(*ts[0], 2); aa.rsC = ts[0]->stride[3];
    aa.adjRayDir = static_cast<float*>(ts[1]->ptr);
    aa.rdB = bstride(*ts[1], 0); aa.rdY = bstride(*ts[1], 1); aa.rdX = bstride(*ts[1], 2); aa.rdC = ts[1]->stride[3];
    aa.camAtomic = B != ts[0]->size[0] || B != ts[1]->size[0];
  }
  if (o.has_tf_derivatives) {
    check_dim(o.adj_tf, "adj_tf", 3);
    check_batch(o.adj_tf, "adj_tf", B);
    if (tf && !preshaded) {
      check_size(o.adj_tf, "adj_tf", 1, tf->size[1]);
      check_size(o.adj_tf, "adj_tf", 2, tf->size[2]);
    }
    aa.adjTf = static_cast<float*>(o.adj_tf.ptr);
    aa.atB = bstride(o.adj_tf, 0); aa.atR = o.adj_tf.stride[1]; aa.atC = o.adj_tf.stride[2];
  }
}

Strided4 strided4(const DvrTensor& t) {
  return {static_cast<const float*>(t.ptr), bstride(t, 0), bstride(t, 1), bstride(t, 2), t.stride[3]};
}

void fill_fd_settings(const DvrForwardDifferencesSettings& s, int* dStep, int* dStart, int* dDir) {
  *dStep = s.d_stepsize;
  for (int i = 0; i < 3; ++i) {
    dStart[i] = s.d_ray_start[i];
    dDir[i] = s.d_ray_dir[i];
  }
}

// shared by dvr_generate_rays / dvr_generate_rays_adjoint: GenerateRayFunction::forward checks
// (renderer/camera.cpp:189-193) and the shapes of the two B*H*W*3 ray tensors
void fill_gen_rays(const DvrTensor& viewport, float fovY, int W, int H, const DvrTensor& start, const DvrTensor& dir,
                   const char* nameStart, const char* nameDir, GenRaysArgs& a) {
  std::memset(&a, 0, sizeof(a));
  check_dim(viewport, "viewport", 3);
  check_size(viewport, "viewport", 1, 3);
  check_size(viewport, "viewport", 2, 3);
  const int B = int(viewport.size[0]);
  if (W <= 0 || H <= 0) fail(DVR_ERR_INVALID_ARGUMENT, "screen size must be positive, but is (%d, %d)", W, H);
  if (B < 1 || B > 65535) fail(DVR_ERR_UNSUPPORTED, "batch size %d outside [1, 65535]", B);
  if ((long long)W * H >= (1ll << 31)) fail(DVR_ERR_UNSUPPORTED, "screen size %d x %d exceeds 2^31 pixels", W, H);
  const DvrTensor* ts[2] = {&start, &dir};
  const char* names[2] = {nameStart, nameDir};
  for (int k = 0; k < 2; ++k) {
    check_dim(*ts[k], names[k], 4);
    check_size(*ts[k], names[k], 0, B);
    check_size(*ts[k], names[k], 1, H);
    check_size(*ts[k], names[k], 2, W);
    check_size(*ts[k], names[k], 3, 3);
  }
  a.W = W; a.H = H; a.B = B;
  a.viewport = {static_cast<const float*>(viewport.ptr), viewport.stride[0], viewport.stride[1], viewport.stride[2]};
  // camera_impl.cu:76-78
  const float fovX = fovY * float(W) / float(H);
  a.tanFovX = tanf(fovX / 2);
  a.tanFovY = tanf(fovY / 2);
  a.rayStart = static_cast<float*>(start.ptr);
  a.sB = start.stride[0]; a.sY = start.stride[1]; a.sX = start.stride[2]; a.sC = start.stride[3];
  a.rayDir = static_cast<float*>(dir.ptr);
  a.dB = dir.stride[0]; a.dY = dir.stride[1]; a.dX = dir.stride[2]; a.dC = dir.stride[3];
  auto packed = [&](const DvrTensor& t) {
    return t.stride[3] == 1 && t.stride[2] == 3 && t.stride[1] == 3ll * W && t.stride[0] == 3ll * W * H &&
           reinterpret_cast<uintptr_t>(t.ptr) % 16 == 0;
  };
  a.packed = W % 4 == 0 && packed(start) && packed(dir);
}

// ---- QUAD buffers: copy of the volume (RayInputs::quads) and gradient scratch (AdjointArgs::gradQuads) --------
// Per-call scratch is stream-ordered memory from a PRIVATE pool (one per device, created on first use): the
// device's default pool -- which other cudaMallocAsync users share -- is left alone. The private pool keeps its
// pages between calls (release threshold = max) until dvr_trim_scratch().
struct QuadScratch {
  float4* ptr = nullptr;
  cudaStream_t stream = nullptr;
  ~QuadScratch() {
    if (ptr) cudaFreeAsync(ptr, stream);
  }
};

constexpr int kMaxDevices = 64;
cudaMemPool_t g_pools[kMaxDevices] = {};
std::mutex g_pool_mutex;

cudaMemPool_t scratch_pool() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return nullptr;
  std::lock_guard<std::mutex> lock(g_pool_mutex);
  if (!g_pools[dev]) {
    cudaMemPoolProps props = {};
    props.allocType = cudaMemAllocationTypePinned;
    props.handleTypes = cudaMemHandleTypeNone;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = dev;
    cudaMemPool_t pool = nullptr;
    if (cudaMemPoolCreate(&pool, &props) != cudaSuccess) {
      cudaGetLastError();
      return nullptr;
    }
    unsigned long long keep = ~0ull;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    g_pools[dev] = pool;
  }
  return g_pools[dev];
}

bool scratch_alloc(QuadScratch& q, size_t bytes, cudaStream_t s) {
  cudaMemPool_t pool = scratch_pool();
  void* p = nullptr;
  if (!pool || cudaMallocFromPoolAsync(&p, bytes, pool, s) != cudaSuccess) {
    cudaGetLastError();  // out of memory: the caller takes the in-place path (counted)
    return false;
  }
  q.ptr = static_cast<float4*>(p);
  q.stream = s;
  return true;
}

}  // namespace

// Persistent QUAD buffers (include/dvr_b200.h "DvrWorkspace"). Plain cudaMalloc: long-lived, grown on demand.
struct DvrWorkspace {
  std::mutex m;
  int device = 0;
  // copy of the volume, valid for the tensor it was built from
  float4* quads = nullptr;
  size_t quadBytes = 0;
  bool quadValid = false;
  const void* volPtr = nullptr;
  long long volSize[4] = {0, 0, 0, 0}, volStride[4] = {0, 0, 0, 0};
  // gradient scratch
  float4* grad = nullptr;
  size_t gradBytes = 0;
  bool accumulating = false;
  const void* adjPtr = nullptr;  // adj_volume of the open accumulation scope
  long long adjVoxels = 0;
  int adjNb = 0;
};

namespace {

bool ws_reserve(float4*& buf, size_t& have, size_t want) {
  if (have >= want) return true;
  if (buf) cudaFree(buf);  // synchronises: no kernel of an earlier call still reads it
  buf = nullptr;
  have = 0;
  void* p = nullptr;
  if (cudaMalloc(&p, want) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  buf = static_cast<float4*>(p);
  have = want;
  return true;
}

bool ws_volume_matches(const DvrWorkspace& w, const RayInputs& r, int nb) {
  return w.quadValid && w.volPtr == r.volume && w.volSize[0] == nb && w.volSize[1] == r.X && w.volSize[2] == r.Y &&
         w.volSize[3] == r.Z && w.volStride[0] == r.volB && w.volStride[1] == r.sx && w.volStride[2] == r.sy &&
         w.volStride[3] == r.sz;
}

void ws_remember_volume(DvrWorkspace& w, const RayInputs& r, int nb) {
  w.volPtr = r.volume;
  w.volSize[0] = nb; w.volSize[1] = r.X; w.volSize[2] = r.Y; w.volSize[3] = r.Z;
  w.volStride[0] = r.volB; w.volStride[1] = r.sx; w.volStride[2] = r.sy; w.volStride[3] = r.sz;
  w.quadValid = true;
}

bool quad_eligible(const Checked& c) {
  const long long voxels = (long long)c.X * c.Y * c.Z;
  return voxels < (1ll << 31) && c.X >= 2 && c.Y >= 2 && c.Z >= 2;
}

// quads per batch entry of a QUAD buffer: X + 1 slices (the last one duplicates slice X-1, RayInputs::quads)
long long quad_count(int X, int Y, int Z) { return ((long long)X + 1) * Y * Z; }

void set_quad_indexing(RayInputs& r) {
#if DVR_QUAD_XINNER
  const unsigned SZ = unsigned(r.X) + 1u, SY = SZ * unsigned(r.Z);
  r.quadYZ = SY;
  r.quadSZ = SZ;
  r.quadBias = unsigned(kTwo23Bits) * (SY + SZ + 1u);
  r.quadMax = SY * unsigned(r.Y) - 2u;
#else
  const unsigned YZ = unsigned(r.Y) * unsigned(r.Z);
  r.quadYZ = YZ;
  r.quadSZ = 1u;
  r.quadBias = unsigned(kTwo23Bits) * (YZ + unsigned(r.Z) + 1u);
  r.quadMax = unsigned(r.X) * YZ - 1u;
#endif
}

// returns the filter id to dispatch on (kFilterTrilinearQuad if a QUAD copy of the volume serves this call)
int maybe_build_quads(int filter, Checked& c, QuadScratch& q, const CallCtx& ctx, cudaStream_t s) {
  if (filter != DVR_FILTER_TRILINEAR || ctx.layout == 0 || !quad_eligible(c)) return filter;
  const long long voxels = (long long)c.X * c.Y * c.Z;
  const long long quadsPer = quad_count(c.X, c.Y, c.Z);
  const int nb = c.ri.volB == 0 ? 1 : c.B;
  const size_t bytes = size_t(quadsPer) * nb * sizeof(float4);
  // automatic: the copy costs ~5 volume sweeps; worth it once the rays visit about as many cells as there are voxels
  const bool worth = ctx.layout == 1 || (long long)c.B * c.W * c.H * 64 >= voxels * nb;
  float4* quads = nullptr;
  if (ctx.ws) {
    DvrWorkspace& w = *ctx.ws;
    std::lock_guard<std::mutex> lock(w.m);
    if (ws_volume_matches(w, c.ri, nb)) {
      g_quad_reuses.fetch_add(1, std::memory_order_relaxed);
      quads = w.quads;  // free of charge: always used
    } else {
      if (!worth) return filter;
      if ((long long)bytes > ctx.cap || !ws_reserve(w.quads, w.quadBytes, bytes)) {
        w.quadValid = false;
        g_quad_fallbacks.fetch_add(1, std::memory_order_relaxed);
        return filter;
      }
      w.quadValid = false;
      after_launch(launch_build_quads(c.ri, nb, w.quads, s), "dvr_build_quads", s);
      g_quad_builds.fetch_add(1, std::memory_order_relaxed);
      ws_remember_volume(w, c.ri, nb);
      quads = w.quads;
    }
  } else {
    if (!worth) return filter;
    if ((long long)bytes > ctx.cap || !scratch_alloc(q, bytes, s)) {
      g_quad_fallbacks.fetch_add(1, std::memory_order_relaxed);
      return filter;
    }
    after_launch(launch_build_quads(c.ri, nb, q.ptr, s), "dvr_build_quads", s);
    g_quad_builds.fetch_add(1, std::memory_order_relaxed);
    quads = q.ptr;
  }
  c.ri.quads = quads;
  c.ri.quadB = nb == 1 ? 0 : quadsPer;
  set_quad_indexing(c.ri);
  return kFilterTrilinearQuad;
}

// (B)*X*Y*Z tensor -> the volume fields of RayInputs (shared by the workspace entry points)
void volume_fields(const DvrTensor& t, const char* name, RayInputs& r, int& nb) {
  check_dim(t, name, 4);
  std::memset(&r, 0, sizeof(r));
  r.X = int(t.size[1]); r.Y = int(t.size[2]); r.Z = int(t.size[3]);
  if (r.X < 1 || r.Y < 1 || r.Z < 1 || t.size[0] < 1) fail(DVR_ERR_INVALID_ARGUMENT, "%s must not be empty", name);
  long long span = 0;
  for (int d = 1; d < 4; ++d) span += (t.size[d] - 1) * bstride(t, d);
  if (span >= (1ll << 31) || t.size[1] >= (1 << 22) || t.size[2] >= (1 << 22) || t.size[3] >= (1 << 22))
    fail(DVR_ERR_UNSUPPORTED, "one %s batch entry must span fewer than 2^31 elements", name);
  if (t.size[0] > 65535) fail(DVR_ERR_UNSUPPORTED, "batch size %lld exceeds 65535", (long long)t.size[0]);
  r.volume = static_cast<const float*>(t.ptr);
  r.volB = bstride(t, 0);
  r.sx = int(bstride(t, 1)); r.sy = int(bstride(t, 2)); r.sz = int(bstride(t, 3));
  nb = int(t.size[0]);
}

// a spatial dimension of size > 1 with stride 0 (an expanded view): several voxels share one address, which
// only the atomics of the in-place path accumulate correctly
bool aliased_spatial_strides(const DvrTensor& t) {
  for (int d = 1; d < 4; ++d)
    if (t.size[d] > 1 && t.stride[d] == 0) return true;
  return false;
}

}  // namespace

namespace dvr {

cudaError_t launch_adjoint(int filter, int tf, int blend, bool hasVol, int tfacc, int ray, const RayInputs& in,
                           const AdjointArgs& aa, dim3 grid, size_t smem, cudaStream_t s) {
#define DVR_CASE(TF, BL) \
  if (tf == TF && blend == BL) return launch_adjoint_##TF##_##BL(filter, hasVol, tfacc, ray, in, aa, grid, smem, s);
  DVR_FOR_EACH_ADJOINT_INSTANCE(DVR_CASE)
#undef DVR_CASE
  return kErrNoInstance;
}

cudaError_t launch_forward_gradients(int filter, int tf, int blend, int chunk, bool hasPos, bool hasTf,
                                     const RayInputs& in, const ForwardOutputs& out, const ForwardGradArgs& ga,
                                     dim3 grid, size_t smem, cudaStream_t s) {
#define DVR_CASE(TF, BL) \
  if (tf == TF && blend == BL) return launch_fwdgrad_##TF##_##BL(filter, chunk, hasPos, hasTf, in, out, ga, grid, smem, s);
  DVR_FOR_EACH_FWDGRAD_INSTANCE(DVR_CASE)
#undef DVR_CASE
  return kErrNoInstance;
}

}  // namespace dvr

// ---- dvr_render_forward_noalloc: the QUAD copy in static device memory -----------------------------------------
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc): the QUAD copy
// of the volume lives in this file-scope array instead of pool / workspace memory. Every GPU that loads the library
// holds the array (DVR_NOALLOC_MAX_QUADS x 16 bytes = 33.8 MB) until the process exits. The name is a host-side
// stand-in in host code: the device address comes from cudaGetSymbolAddress, once per device, with that device current.
static __device__ __align__(256) float4 g_noalloc_quads[DVR_NOALLOC_MAX_QUADS];

namespace {

float4* g_noalloc_ptr[kMaxDevices] = {};
std::mutex g_noalloc_mutex;

float4* noalloc_quads() {
  int dev = 0;
  cuda_check(cudaGetDevice(&dev), "dvr_render_forward_noalloc: cudaGetDevice");
  if (dev < 0 || dev >= kMaxDevices) fail(DVR_ERR_UNSUPPORTED, "dvr_render_forward_noalloc: device %d outside [0, %d)", dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_ptr[dev]) {
    void* p = nullptr;
    cuda_check(cudaGetSymbolAddress(&p, g_noalloc_quads), "dvr_render_forward_noalloc: cudaGetSymbolAddress(g_noalloc_quads)");
    if (!p) fail(DVR_ERR_CUDA, "dvr_render_forward_noalloc: cudaGetSymbolAddress returned NULL");
    g_noalloc_ptr[dev] = static_cast<float4*>(p);
  }
  return g_noalloc_ptr[dev];
}

}  // namespace

// ================================================================================================
extern "C" {

int dvr_check_inputs(const DvrInputs* inputs, int32_t* B, int32_t* X, int32_t* Y, int32_t* Z) {
  return guarded([&] {
    if (!inputs) fail(DVR_ERR_INVALID_ARGUMENT, "inputs is NULL");
    Checked c;
    check_inputs(*inputs, c);
    if (B) *B = c.B;
    if (X) *X = c.X;
    if (Y) *Y = c.Y;
    if (Z) *Z = c.Z;
  });
}

int dvr_render_forward(const DvrInputs* inputs, const DvrOutputs* outputs, void* stream) {
  return dvr_render_forward_ex(inputs, outputs, nullptr, stream);
}

int dvr_render_forward_ex(const DvrInputs* inputs, const DvrOutputs* outputs, const DvrCallOptions* options,
                          void* stream) {
  return guarded([&] {
    if (!inputs || !outputs) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    const CallCtx ctx = resolve_options(options);
    Checked c;
    check_inputs(*inputs, c);
    const ForwardOutputs fo = check_outputs(*outputs, c);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (inputs->tf_mode == DVR_TF_PRESHADED) {
      after_launch(launch_forward_preshaded(inputs->blend_mode, c.ri, fo, tile_grid(c), s), "dvr_render_forward", s);
      return;
    }
    const size_t smem = check_smem(size_t(c.ri.R) * tf_row(inputs->tf_mode) * sizeof(float));
    QuadScratch quads;
    const int filter = maybe_build_quads(inputs->volume_filter_mode, c, quads, ctx, s);
    after_launch(launch_forward(filter, inputs->tf_mode, inputs->blend_mode, c.ri, fo,
                                tile_grid(c), smem, s),
                 "dvr_render_forward", s);
  });
}

// Forward render that allocates nothing (include/dvr_b200.h): same checks, same kernels as dvr_render_forward_ex with
// volume_layout = 1, the QUAD copy built into g_noalloc_quads. No cudaMalloc / cudaFree / pool call on this path.
int dvr_render_forward_noalloc(const DvrInputs* inputs, const DvrOutputs* outputs, void* stream) {
  return guarded([&] {
    if (!inputs || !outputs) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    t_call_sync = 0;
    Checked c;
    check_inputs(*inputs, c);
    const ForwardOutputs fo = check_outputs(*outputs, c);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (inputs->tf_mode == DVR_TF_PRESHADED) {  // no scratch in this kernel
      after_launch(launch_forward_preshaded(inputs->blend_mode, c.ri, fo, tile_grid(c), s), "dvr_render_forward_noalloc", s);
      return;
    }
    const size_t smem = check_smem(size_t(c.ri.R) * tf_row(inputs->tf_mode) * sizeof(float));
    int filter = inputs->volume_filter_mode;
    if (filter == DVR_FILTER_TRILINEAR && quad_eligible(c)) {  // (volumes thinner than 2 voxels: in place, no scratch)
      const long long quadsPer = quad_count(c.X, c.Y, c.Z);
      const int nb = c.ri.volB == 0 ? 1 : c.B;
      if (quadsPer * nb > (long long)DVR_NOALLOC_MAX_QUADS)
        fail(DVR_ERR_UNSUPPORTED,
             "dvr_render_forward_noalloc: the QUAD copy of %d volume(s) of %d x %d x %d needs %lld entries ((X + 1) * Y * Z "
             "each), the static scratch holds DVR_NOALLOC_MAX_QUADS = %lld (one 128^3 volume); use dvr_render_forward_ex",
             nb, c.X, c.Y, c.Z, quadsPer * nb, (long long)DVR_NOALLOC_MAX_QUADS);
      float4* quads = noalloc_quads();
      // the build kernel overwrites every entry the march reads ((X + 1) * Y * Z per volume): nothing to clear
      after_launch(launch_build_quads(c.ri, nb, quads, s), "dvr_build_quads", s);
      g_quad_builds.fetch_add(1, std::memory_order_relaxed);
      c.ri.quads = quads;
      c.ri.quadB = nb == 1 ? 0 : quadsPer;
      set_quad_indexing(c.ri);
      filter = kFilterTrilinearQuad;
    }
    after_launch(launch_forward(filter, inputs->tf_mode, inputs->blend_mode, c.ri, fo, tile_grid(c), smem, s),
                 "dvr_render_forward_noalloc", s);
  });
}

int dvr_render_adjoint(const DvrInputs* inputs, const DvrOutputs* fwd, const DvrTensor* adj_color,
                       const DvrAdjointOutputs* adj, void* stream) {
  return dvr_render_adjoint_ex(inputs, fwd, adj_color, adj, nullptr, stream);
}

int dvr_render_adjoint_ex(const DvrInputs* inputs, const DvrOutputs* fwd, const DvrTensor* adj_color,
                          const DvrAdjointOutputs* adj, const DvrCallOptions* options, void* stream) {
  return guarded([&] {
    if (!inputs || !fwd || !adj_color || !adj) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    const CallCtx ctx = resolve_options(options);
    Checked c;
    check_inputs(*inputs, c);
    // outputs of the forward pass (renderer.cpp:670-691) and adjoint colour (:693-701)
    const DvrTensor* col[2] = {&fwd->color, adj_color};
    const char* names[2] = {"output_color", "adj_color"};
    for (int k = 0; k < 2; ++k) {
      check_dim(*col[k], names[k], 4);
      check_size(*col[k], names[k], 0, c.B);
      check_size(*col[k], names[k], 1, c.H);
      check_size(*col[k], names[k], 2, c.W);
      check_size(*col[k], names[k], 3, 4);
    }
    check_dim(fwd->termination_index, "terminationIndex", 3);
    check_size(fwd->termination_index, "terminationIndex", 0, c.B);
    check_size(fwd->termination_index, "terminationIndex", 1, c.H);
    check_size(fwd->termination_index, "terminationIndex", 2, c.W);
    AdjointArgs aa;
    std::memset(&aa, 0, sizeof(aa));
    aa.color = strided4(fwd->color);
    aa.adjColor = strided4(*adj_color);
    aa.term = static_cast<const int*>(fwd->termination_index.ptr);
    aa.tB = bstride(fwd->termination_index, 0);
    aa.tY = bstride(fwd->termination_index, 1);
    aa.tX = bstride(fwd->termination_index, 2);
    const bool preshaded = inputs->tf_mode == DVR_TF_PRESHADED;
    check_adjoint_outputs(*adj, c.B, c.H, c.W, c.X, c.Y, c.Z, &inputs->tf, aa, preshaded);
    const bool hasVol = adj->has_volume_derivatives != 0, hasTf = adj->has_tf_derivatives != 0;
    const bool hasRay = aa.hasStep || aa.hasCam;
    if (!hasVol && !hasTf && !hasRay) fail(DVR_ERR_INVALID_ARGUMENT, "At least one adjoint variable must be activated!");
    aa.compatStepOverwrite = ctx.compat;
    aa.marchZeroRays = ctx.marchZeroRays;
    aa.sameStrides = hasVol && aa.avx == c.ri.sx && aa.avy == c.ri.sy && aa.avz == c.ri.sz;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (preshaded) {
      // the pre-shaded TF has no parameters: a request for adj_tf alone is valid and a no-op
      if (hasVol || hasRay)
        after_launch(launch_adjoint_preshaded(inputs->blend_mode, hasVol, hasRay, c.ri, aa, tile_grid(c), s),
                     "dvr_render_adjoint", s);
      return;
    }
    const int tf = inputs->tf_mode;
    int tfacc = kTfAccNone;
    // Texture / Linear touch two neighbouring rows per sample: a two-row register window flushed into a per-warp
    // shared table, whatever R. Identity / Gaussian touch every row: register accumulators while they fit.
    if (hasTf) {
      if (tf == kTfTexture || tf == kTfLinear) tfacc = kTfAccSmem;
      else tfacc = (tf == kTfIdentity || c.ri.R <= kRegRows) ? kTfAccReg : kTfAccSmem;
    }
    const int rayMode = aa.hasCam ? kRayCam : (aa.hasStep ? kRayStep : kRayNone);
    const int entries = c.ri.R * tf_channels(tf);
    const int ntab = tfacc == kTfAccSmem ? kThreads / 32 : (tfacc == kTfAccReg ? 1 : 0);
    const size_t smem =
        check_smem((size_t((c.ri.R * tf_row(tf) + 3) & ~3) + size_t(ntab) * entries) * sizeof(float));
    QuadScratch quads, gradQuads;
    int filter = maybe_build_quads(inputs->volume_filter_mode, c, quads, ctx, s);
    int ng = 0;
    bool fold = false;
    if (filter == kFilterTrilinearQuad && hasVol) {
      // gradient scratch in the same layout: 2 x red.v4 per cell instead of 8 scalar REDs, folded afterwards
      const long long voxels = (long long)c.X * c.Y * c.Z;
      const long long quadsPer = quad_count(c.X, c.Y, c.Z);
      ng = aa.avB == 0 ? 1 : c.B;
      const size_t bytes = size_t(quadsPer) * ng * sizeof(float4);
      float4* grad = nullptr;
      if (aliased_spatial_strides(adj->adj_volume) || (long long)bytes > ctx.cap) {
        // expanded adj_volume views need the atomics of the in-place path (a fold would lose updates)
      } else if (ctx.ws) {
        DvrWorkspace& w = *ctx.ws;
        std::lock_guard<std::mutex> lock(w.m);
        if (w.accumulating && w.adjPtr == adj->adj_volume.ptr && w.adjVoxels == voxels && w.adjNb == ng) {
          grad = w.grad;  // open accumulation scope: zeroed by begin_volume_gradient, folded by end_volume_gradient
        } else if (!w.accumulating && ws_reserve(w.grad, w.gradBytes, bytes)) {
          cuda_check(cudaMemsetAsync(w.grad, 0, bytes, s), "dvr_render_adjoint");
          grad = w.grad;
          fold = true;
        }
      }
      if (!grad && !(ctx.ws && ctx.ws->accumulating) && !aliased_spatial_strides(adj->adj_volume) &&
          (long long)bytes <= ctx.cap && scratch_alloc(gradQuads, bytes, s)) {
        cuda_check(cudaMemsetAsync(gradQuads.ptr, 0, bytes, s), "dvr_render_adjoint");
        grad = gradQuads.ptr;
        fold = true;
      }
      if (grad) {
        aa.gradQuads = grad;
        aa.gradQuadB = ng == 1 ? 0 : quadsPer;
      } else {
        g_quad_fallbacks.fetch_add(1, std::memory_order_relaxed);
        filter = inputs->volume_filter_mode;  // no scratch: the in-place path accumulates straight into adj_volume
      }
    }
    after_launch(launch_adjoint(filter, tf, inputs->blend_mode, hasVol, tfacc, rayMode, c.ri, aa,
                                tile_grid(c), smem, s),
                 "dvr_render_adjoint", s);
    if (fold) {
      after_launch(launch_fold_quads(c.ri, aa, ng, s), "dvr_fold_quads", s);
      g_folds.fetch_add(1, std::memory_order_relaxed);
    }
  });
}

int dvr_render_forward_gradients(const DvrInputs* inputs, const DvrForwardDifferencesSettings* st,
                                 const DvrOutputs* outputs, const DvrTensor* gradients_out, void* stream) {
  return dvr_render_forward_gradients_ex(inputs, st, outputs, gradients_out, nullptr, stream);
}

int dvr_render_forward_gradients_ex(const DvrInputs* inputs, const DvrForwardDifferencesSettings* st,
                                    const DvrOutputs* outputs, const DvrTensor* gradients_out,
                                    const DvrCallOptions* options, void* stream) {
  return guarded([&] {
    if (!inputs || !st || !outputs || !gradients_out) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    const CallCtx ctx = resolve_options(options);
    Checked c;
    check_inputs(*inputs, c);
    if (st->D <= 0) fail(DVR_ERR_INVALID_ARGUMENT, "number of derivatives must be positive");
    const bool hasStep = st->d_stepsize >= 0;
    bool hasCam = false;
    for (int i = 0; i < 3; ++i) hasCam = hasCam || st->d_ray_start[i] >= 0 || st->d_ray_dir[i] >= 0;
    const bool hasTf = st->has_tf_derivatives != 0;
    if (!hasStep && !hasCam && !hasTf) fail(DVR_ERR_INVALID_ARGUMENT, "at least one derivative must be activated");
    if (inputs->tf_mode == DVR_TF_PRESHADED)
      fail(DVR_ERR_UNSUPPORTED, "forward derivatives are not available for pre-shaded volumes");
    if (inputs->tf_mode == DVR_TF_GAUSSIAN_LOG)
      fail(DVR_ERR_UNSUPPORTED, "forward derivatives are not supported for TFGaussianLog (deleted in the reference, renderer_tf.cuh:692-702)");
    ForwardGradArgs ga;
    std::memset(&ga, 0, sizeof(ga));
    if (hasTf) {
      if (!st->d_tf.ptr)
        fail(DVR_ERR_INVALID_ARGUMENT, "d_tf not defined, hence hasTfDerivatives must be false but was set to true");
      check_dim(st->d_tf, "d_tf", 3);
      check_size(st->d_tf, "d_tf", 0, inputs->tf.size[0]);
      check_size(st->d_tf, "d_tf", 1, inputs->tf.size[1]);
      check_size(st->d_tf, "d_tf", 2, inputs->tf.size[2]);
      ga.dTf = static_cast<const int*>(st->d_tf.ptr);
      ga.dtB = bstride(st->d_tf, 0); ga.dtR = st->d_tf.stride[1]; ga.dtC = st->d_tf.stride[2];
    }
    const ForwardOutputs fo = check_outputs(*outputs, c);
    check_dim(*gradients_out, "gradientsOut", 5);
    check_size(*gradients_out, "gradientsOut", 0, c.B);
    check_size(*gradients_out, "gradientsOut", 1, c.H);
    check_size(*gradients_out, "gradientsOut", 2, c.W);
    check_size(*gradients_out, "gradientsOut", 3, st->D);
    check_size(*gradients_out, "gradientsOut", 4, 4);
    ga.D = st->D;
    fill_fd_settings(*st, &ga.dStep, ga.dRayStart, ga.dRayDir);
    ga.grad = static_cast<float*>(gradients_out->ptr);
    ga.gB = gradients_out->stride[0]; ga.gY = gradients_out->stride[1]; ga.gX = gradients_out->stride[2];
    ga.gD = gradients_out->stride[3]; ga.gC = gradients_out->stride[4];
    // the four channels of a Jacobian slot as one 16-byte access where the layout allows it
    ga.gradVec4 = ga.gC == 1 && ga.gD % 4 == 0 && ga.gX % 4 == 0 && ga.gY % 4 == 0 && ga.gB % 4 == 0 &&
                  (reinterpret_cast<uintptr_t>(ga.grad) & 15) == 0;
    const int tf = inputs->tf_mode;
    const size_t smem = check_smem(size_t(((c.ri.R * tf_row(tf) + 3) & ~3) + c.ri.R * tf_row(tf)) * sizeof(float));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    QuadScratch quads;  // one QUAD copy of the volume serves all passes
    const int filter = maybe_build_quads(inputs->volume_filter_mode, c, quads, ctx, s);
    if (!hasTf) {
      // ray / stepsize tangents only: the slots are a linear map of <= 7 input tangents carried per input, ONE
      // marching pass for any D (k_forward_gradients_inputs)
      ga.d0 = 0;
      ga.writeColor = 1;
      after_launch(launch_forward_gradients(filter, tf, inputs->blend_mode, 0, true, false, c.ri, fo, ga, tile_grid(c),
                                            smem, s),
                   "dvr_render_forward_gradients", s);
      return;
    }
    // Gaussian tables: the one-pass kernel keeps 15 moments per row and thread in shared memory (R <= 28 fits);
    // larger tables take the slot-major passes
    const size_t smemGauss = smem + size_t(15) * c.ri.R * kThreads * sizeof(float);
    const bool gaussOnePass = tf == kTfGaussian && smemGauss <= 200 * 1024;
    if (!ctx.fwdTfSlotMajor && (tf == kTfIdentity || tf == kTfTexture || tf == kTfLinear || gaussOnePass)) {
      // TF tangents in ONE sample-major pass for any D (dvr_tfjac.cuh). It needs the
      // final pixel values, so the colour comes first: from the input-major pass when ray / stepsize tangents are
      // traced as well (it writes all D slots, zeros where no input maps), else from the plain forward kernel
      // followed by a zero-fill of the Jacobian.
      ga.d0 = 0;
      ga.writeColor = 1;
      if (hasStep || hasCam) {
        after_launch(launch_forward_gradients(filter, tf, inputs->blend_mode, 0, true, false, c.ri, fo, ga,
                                              tile_grid(c), smem, s),
                     "dvr_render_forward_gradients", s);
      } else {
        after_launch(launch_forward(filter, tf, inputs->blend_mode, c.ri, fo, tile_grid(c), smem, s),
                     "dvr_render_forward_gradients", s);
        const long long D4 = 4ll * st->D;
        const bool dense = ga.gC == 1 && ga.gD == 4 && ga.gX == D4 && ga.gY == D4 * c.W &&
                           (c.B == 1 || ga.gB == D4 * c.W * c.H);
        if (dense) {
          cuda_check(cudaMemsetAsync(ga.grad, 0, size_t(D4) * c.W * c.H * c.B * sizeof(float), s),
                     "dvr_render_forward_gradients");
        } else {
          after_launch(launch_zero_jacobian(ga, c.W, c.H, c.B, s), "dvr_render_forward_gradients", s);
        }
      }
      ga.exactPos = hasStep || hasCam;
      after_launch(launch_forward_gradients(filter, tf, inputs->blend_mode, -1, hasStep || hasCam, true, c.ri, fo, ga,
                                            tile_grid(c), gaussOnePass ? smemGauss : smem, s),
                   "dvr_render_forward_gradients", s);
      return;
    }
    // TF tangents, slot-major: ceil(D / chunk) marching passes over disjoint slot ranges
    int d0 = 0;
    while (d0 < st->D) {
      const int left = st->D - d0;
      const int chunk = left == 1 ? 1 : (left <= 6 ? 6 : 8);
      ga.d0 = d0;
      ga.writeColor = d0 == 0;
      after_launch(launch_forward_gradients(filter, tf, inputs->blend_mode, chunk,
                                            hasStep || hasCam, hasTf, c.ri, fo, ga, tile_grid(c), smem, s),
                   "dvr_render_forward_gradients", s);
      d0 += chunk;
    }
  });
}

int dvr_forward_variables_to_gradients(const DvrTensor* fv, const DvrTensor* gc, const DvrForwardDifferencesSettings* st,
                                       const DvrAdjointOutputs* adj, void* stream) {
  return guarded([&] {
    if (!fv || !gc || !st || !adj) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    check_dim(*fv, "forwardVariables", 5);
    const int B = int(fv->size[0]), H = int(fv->size[1]), W = int(fv->size[2]), D = int(fv->size[3]);
    check_size(*fv, "forwardVariables", 4, 4);
    check_dim(*gc, "gradientOutputColor", 4);
    check_size(*gc, "gradientOutputColor", 0, B);
    check_size(*gc, "gradientOutputColor", 1, H);
    check_size(*gc, "gradientOutputColor", 2, W);
    check_size(*gc, "gradientOutputColor", 3, 4);
    if (adj->has_volume_derivatives)
      fail(DVR_ERR_INVALID_ARGUMENT, "volume derivatives are not supported with forward variables");
    if (B > 65535) fail(DVR_ERR_UNSUPPORTED, "batch size %d exceeds 65535", B);
    AdjointArgs aa;
    std::memset(&aa, 0, sizeof(aa));
    check_adjoint_outputs(*adj, B, H, W, -1, -1, -1, nullptr, aa);
    FwdVarsArgs a;
    std::memset(&a, 0, sizeof(a));
    a.W = W; a.H = H; a.B = B; a.D = D;
    a.fv = static_cast<const float*>(fv->ptr);
    a.fB = fv->stride[0]; a.fY = fv->stride[1]; a.fX = fv->stride[2]; a.fD = fv->stride[3]; a.fC = fv->stride[4];
    a.gradColor = strided4(*gc);
    fill_fd_settings(*st, &a.dStep, a.dRayStart, a.dRayDir);
    auto need = [&](int d, bool have, const char* what) {
      if (d >= D) fail(DVR_ERR_INVALID_ARGUMENT, "derivative slot %d of %s exceeds D=%d", d, what, D);
      if (d >= 0 && !have) fail(DVR_ERR_INVALID_ARGUMENT, "%s derivative requested but the adjoint output is not enabled", what);
    };
    need(a.dStep, aa.hasStep, "stepsize");
    for (int i = 0; i < 3; ++i) {
      need(a.dRayStart[i], aa.hasCam, "rayStart");
      need(a.dRayDir[i], aa.hasCam, "rayDir");
    }
    a.hasTf = st->has_tf_derivatives != 0;
    if (a.hasTf) {
      check_dim(st->d_tf, "d_tf", 3);
      if (!adj->has_tf_derivatives) fail(DVR_ERR_INVALID_ARGUMENT, "TF derivative requested but adj_tf is not enabled");
      check_size(adj->adj_tf, "adj_tf", 1, st->d_tf.size[1]);
      check_size(adj->adj_tf, "adj_tf", 2, st->d_tf.size[2]);
      a.dTf = static_cast<const int*>(st->d_tf.ptr);
      a.dtB = bstride(st->d_tf, 0); a.dtR = st->d_tf.stride[1]; a.dtC = st->d_tf.stride[2];
      a.R = int(st->d_tf.size[1]);
      a.C = int(st->d_tf.size[2]);
    }
    a.adjStep = aa.adjStep; a.asB = aa.asB; a.asY = aa.asY; a.asX = aa.asX;
    a.adjRayStart = aa.adjRayStart; a.rsB = aa.rsB; a.rsY = aa.rsY; a.rsX = aa.rsX; a.rsC = aa.rsC;
    a.adjRayDir = aa.adjRayDir; a.rdB = aa.rdB; a.rdY = aa.rdY; a.rdX = aa.rdX; a.rdC = aa.rdC;
    a.adjTf = aa.adjTf; a.atB = aa.atB; a.atR = aa.atR; a.atC = aa.atC;
    check_smem(size_t(a.R) * a.C * sizeof(float));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    after_launch(launch_fwdvars_to_gradients(a, s), "dvr_forward_variables_to_gradients", s);
  });
}

int dvr_generate_rays(const DvrTensor* viewport, float fov_y, int32_t W, int32_t H, const DvrTensor* ray_start,
                      const DvrTensor* ray_dir, void* stream) {
  return guarded([&] {
    if (!viewport || !ray_start || !ray_dir) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    GenRaysArgs a;
    fill_gen_rays(*viewport, fov_y, W, H, *ray_start, *ray_dir, "rayStart", "rayDir", a);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    after_launch(launch_generate_rays(a, s), "dvr_generate_rays", s);
  });
}

int dvr_generate_rays_adjoint(const DvrTensor* viewport, float fov_y, const DvrTensor* adj_ray_start,
                              const DvrTensor* adj_ray_dir, const DvrTensor* adj_viewport, void* stream) {
  return guarded([&] {
    if (!viewport || !adj_ray_start || !adj_ray_dir || !adj_viewport) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    check_dim(*adj_ray_start, "adj_rayStart", 4);
    GenRaysArgs a;
    fill_gen_rays(*viewport, fov_y, int(adj_ray_start->size[2]), int(adj_ray_start->size[1]), *adj_ray_start,
                  *adj_ray_dir, "adj_rayStart", "adj_rayDir", a);
    check_dim(*adj_viewport, "adj_viewport", 3);
    check_size(*adj_viewport, "adj_viewport", 0, a.B);
    check_size(*adj_viewport, "adj_viewport", 1, 3);
    check_size(*adj_viewport, "adj_viewport", 2, 3);
    a.adjViewport = static_cast<float*>(adj_viewport->ptr);
    a.avB = adj_viewport->stride[0]; a.avR = adj_viewport->stride[1]; a.avC = adj_viewport->stride[2];
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    cuda_check(launch_generate_rays_adjoint(a, s), "dvr_generate_rays_adjoint");
    g_launches.fetch_add(1, std::memory_order_relaxed);  // zero + reduce = 2 launches
    after_launch(cudaSuccess, "dvr_generate_rays_adjoint", s);
  });
}

int dvr_compare_to_image(const DvrTensor* color, const DvrTensor* grads, const DvrTensor* reference,
                         const DvrTensor* diff_out, const DvrTensor* grads_out, int32_t reduce, void* stream) {
  return guarded([&] {
    if (!color || !grads || !reference || !diff_out || !grads_out) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    // Renderer::compareToImage (renderer/renderer.cpp:566-593)
    check_dim(*color, "colorInput", 4);
    const int B = int(color->size[0]), H = int(color->size[1]), W = int(color->size[2]);
    check_size(*color, "colorInput", 3, 4);
    check_dim(*grads, "gradientsInput", 5);
    check_size(*grads, "gradientsInput", 0, B);
    check_size(*grads, "gradientsInput", 1, H);
    check_size(*grads, "gradientsInput", 2, W);
    const int D = int(grads->size[3]);
    check_size(*grads, "gradientsInput", 4, 4);
    check_dim(*reference, "colorReference", 4);
    check_size(*reference, "colorReference", 0, B);
    check_size(*reference, "colorReference", 1, H);
    check_size(*reference, "colorReference", 2, W);
    check_size(*reference, "colorReference", 3, 4);
    CompareArgs a;
    std::memset(&a, 0, sizeof(a));
    a.W = W; a.H = H; a.B = B; a.D = D;
    a.color = {static_cast<const float*>(color->ptr), color->stride[0], color->stride[1], color->stride[2], color->stride[3]};
    a.reference = {static_cast<const float*>(reference->ptr), reference->stride[0], reference->stride[1],
                   reference->stride[2], reference->stride[3]};
    a.grad = static_cast<const float*>(grads->ptr);
    a.gB = grads->stride[0]; a.gY = grads->stride[1]; a.gX = grads->stride[2]; a.gD = grads->stride[3]; a.gC = grads->stride[4];
    a.reduce = reduce != 0;
    a.diffOut = static_cast<float*>(diff_out->ptr);
    a.gradOut = static_cast<float*>(grads_out->ptr);
    if (a.reduce) {  // renderer.cpp:598-608
      check_dim(*diff_out, "differenceOut", 1);
      check_size(*diff_out, "differenceOut", 0, 1);
      check_dim(*grads_out, "gradientsOut", 1);
      check_size(*grads_out, "gradientsOut", 0, D);
      a.qD = grads_out->stride[0];
      a.normalization = 1.0f / float((long long)B * H * W);
      if (D > 8192) fail(DVR_ERR_UNSUPPORTED, "more than 8192 derivatives in the reduced comparison");
    } else {  // renderer.cpp:625-641
      check_dim(*diff_out, "differenceOut", 4);
      check_size(*diff_out, "differenceOut", 0, B);
      check_size(*diff_out, "differenceOut", 1, H);
      check_size(*diff_out, "differenceOut", 2, W);
      check_size(*diff_out, "differenceOut", 3, 1);
      check_dim(*grads_out, "gradientsOut", 4);
      check_size(*grads_out, "gradientsOut", 0, B);
      check_size(*grads_out, "gradientsOut", 1, H);
      check_size(*grads_out, "gradientsOut", 2, W);
      check_size(*grads_out, "gradientsOut", 3, D);
      a.oB = diff_out->stride[0]; a.oY = diff_out->stride[1]; a.oX = diff_out->stride[2];
      a.qB = grads_out->stride[0]; a.qY = grads_out->stride[1]; a.qX = grads_out->stride[2]; a.qD = grads_out->stride[3];
    }
    if ((long long)B * H * W == 0) return;
    if (B > 65535) fail(DVR_ERR_UNSUPPORTED, "batch size %d exceeds 65535", B);
    if ((long long)W * H >= (1ll << 31)) fail(DVR_ERR_UNSUPPORTED, "screen size %d x %d exceeds 2^31 pixels", W, H);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    cuda_check(launch_compare_to_image(a, s), "dvr_compare_to_image");
    if (a.reduce) g_launches.fetch_add(1, std::memory_order_relaxed);  // zero + reduce
    after_launch(cudaSuccess, "dvr_compare_to_image", s);
  });
}

// shared by the two preshade entry points: PreshadeVolumeFunction::forward checks (renderer/tf_utils.cpp:180-214)
static void fill_preshade(const DvrTensor& vol, const DvrTensor& tf, int tfMode, const DvrTensor& color,
                          const char* colorName, RayInputs& in, PreshadeArgs& a, size_t& smem) {
  std::memset(&in, 0, sizeof(in));
  std::memset(&a, 0, sizeof(a));
  check_dim(vol, "densityVolume", 4);
  check_size(vol, "densityVolume", 0, 1);
  const long long X = vol.size[1], Y = vol.size[2], Z = vol.size[3];
  if (X < 1 || Y < 1 || Z < 1) fail(DVR_ERR_INVALID_ARGUMENT, "densityVolume must not be empty");
  if (X * Y * Z >= (1ll << 31)) fail(DVR_ERR_UNSUPPORTED, "densityVolume must have fewer than 2^31 voxels");
  check_dim(tf, "tf", 3);
  check_size(tf, "tf", 0, 1);
  switch (tfMode) {
    case DVR_TF_IDENTITY: check_size(tf, "tf", 1, 1); check_size(tf, "tf", 2, 2); break;
    case DVR_TF_TEXTURE: check_size(tf, "tf", 2, 4); break;
    case DVR_TF_LINEAR: check_size(tf, "tf", 2, 5); break;
    case DVR_TF_GAUSSIAN:
    case DVR_TF_GAUSSIAN_LOG: check_size(tf, "tf", 2, 6); break;
    default: fail(DVR_ERR_INVALID_ARGUMENT, "unknown tf enum value");
  }
  if (tf.size[1] < (tfMode == DVR_TF_LINEAR ? 2 : 1)) fail(DVR_ERR_INVALID_ARGUMENT, "tensor 'tf' has too few rows");
  check_dim(color, colorName, 4);
  check_size(color, colorName, 0, 4);
  check_size(color, colorName, 1, X);
  check_size(color, colorName, 2, Y);
  check_size(color, colorName, 3, Z);
  in.X = int(X); in.Y = int(Y); in.Z = int(Z);
  in.volume = static_cast<const float*>(vol.ptr);
  in.sx = int(vol.stride[1]); in.sy = int(vol.stride[2]); in.sz = int(vol.stride[3]);
  in.tf = {static_cast<const float*>(tf.ptr), 0, tf.stride[1], tf.stride[2]};
  in.R = int(tf.size[1]);
  in.C = int(tf.size[2]);
  a.color = static_cast<float*>(color.ptr);
  a.cC = color.stride[0]; a.cX = color.stride[1]; a.cY = color.stride[2]; a.cZ = color.stride[3];
  smem = check_smem(size_t(in.R) * tf_row(tfMode) * sizeof(float));
}

int dvr_preshade_volume(const DvrTensor* vol, const DvrTensor* tf, int32_t tf_mode, const DvrTensor* color_out,
                        void* stream) {
  return guarded([&] {
    if (!vol || !tf || !color_out) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    RayInputs in;
    PreshadeArgs a;
    size_t smem;
    fill_preshade(*vol, *tf, tf_mode, *color_out, "colorVolume", in, a, smem);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    after_launch(launch_preshade(tf_mode, false, in, a, smem, s), "dvr_preshade_volume", s);
  });
}

int dvr_preshade_volume_adjoint(const DvrTensor* vol, const DvrTensor* tf, int32_t tf_mode, const DvrTensor* adj_color,
                                const DvrTensor* adj_volume_out, void* stream) {
  return guarded([&] {
    if (!vol || !tf || !adj_color || !adj_volume_out) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    RayInputs in;
    PreshadeArgs a;
    size_t smem;
    fill_preshade(*vol, *tf, tf_mode, *adj_color, "adj_colorVolume", in, a, smem);
    check_dim(*adj_volume_out, "adj_densityVolume", 4);
    check_size(*adj_volume_out, "adj_densityVolume", 0, 1);
    for (int d = 1; d < 4; ++d) check_size(*adj_volume_out, "adj_densityVolume", d, vol->size[d]);
    a.adjVolume = static_cast<float*>(adj_volume_out->ptr);
    a.ax = adj_volume_out->stride[1]; a.ay = adj_volume_out->stride[2]; a.az = adj_volume_out->stride[3];
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    after_launch(launch_preshade(tf_mode, true, in, a, smem, s), "dvr_preshade_volume_adjoint", s);
  });
}

// ---- renderer::Volume on the device --------------------------------------------
namespace {
void volume_dims(const DvrTensor& t, const char* name, int& X, int& Y, int& Z) {
  check_dim(t, name, 4);
  check_size(t, name, 0, 1);
  for (int d = 1; d < 4; ++d)
    if (t.size[d] <= 0) fail(DVR_ERR_INVALID_ARGUMENT, "%s must not be empty", name);
  if (t.size[1] * t.size[2] * t.size[3] > 0x7fffffffll)  // renderer/volume.cpp:192-193
    fail(DVR_ERR_INVALID_ARGUMENT, "Dataset is too big, must fit into an 32-bit signed int");
  X = int(t.size[1]); Y = int(t.size[2]); Z = int(t.size[3]);
}
}  // namespace

int dvr_create_implicit_volume(int32_t equation, const float* params, const DvrTensor* volume_out, void* stream) {
  return guarded([&] {
    if (!volume_out) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    if (equation < 0 || equation >= kEqCount) fail(DVR_ERR_INVALID_ARGUMENT, "unknown implicit equation %d", equation);
    ImplicitArgs a{};
    volume_dims(*volume_out, "volume", a.X, a.Y, a.Z);
    if (a.X != a.Y || a.Y != a.Z) fail(DVR_ERR_INVALID_ARGUMENT, "implicit volumes are cubes: X == Y == Z expected");
    // sampling box per equation (boxMin / boxMax members, renderer/volume_implicit.cpp:30-215)
    static const float kBox[kEqCount] = {1, 1, 1, 1, 2, 2, 1.5f, 1, 5, 1, 0.5f, 1, 2, 2, 1};
    a.equation = equation;
    a.boxMin = -kBox[equation];
    a.scaleStep = (kBox[equation] - a.boxMin) / float(a.X - 1);  // renderer/volume.cpp:375
    a.fM = params ? params[0] : 6.0f;
    a.alpha = params ? params[1] : 0.25f;
    a.scale = params ? params[2] : 0.5f;
    a.a = params ? params[3] : 0.25f;
    a.out = static_cast<float*>(volume_out->ptr);
    a.sx = volume_out->stride[1]; a.sy = volume_out->stride[2]; a.sz = volume_out->stride[3];
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    after_launch(launch_implicit_volume(a, s), "dvr_create_implicit_volume", s);
  });
}

int dvr_volume_from_raw(const void* raw, int32_t dtype, const DvrTensor* volume_out, void* stream) {
  return guarded([&] {
    if (!raw || !volume_out) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    if (dtype < 0 || dtype > 2) fail(DVR_ERR_INVALID_ARGUMENT, "unknown .cvol data type %d", dtype);
    RawVolumeArgs a{};
    volume_dims(*volume_out, "volume", a.X, a.Y, a.Z);
    if (a.Y > 65535) fail(DVR_ERR_UNSUPPORTED, "volume conversion supports Y <= 65535, but Y is %d", a.Y);
    a.raw = const_cast<void*>(raw);
    a.vol = static_cast<float*>(volume_out->ptr);
    a.sx = volume_out->stride[1]; a.sy = volume_out->stride[2]; a.sz = volume_out->stride[3];
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    after_launch(launch_volume_raw(dtype, false, a, s), "dvr_volume_from_raw", s);
  });
}

int dvr_volume_to_raw(const DvrTensor* volume, float* raw_out, void* stream) {
  return guarded([&] {
    if (!raw_out || !volume) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    RawVolumeArgs a{};
    volume_dims(*volume, "volume", a.X, a.Y, a.Z);
    if (a.Y > 65535) fail(DVR_ERR_UNSUPPORTED, "volume conversion supports Y <= 65535, but Y is %d", a.Y);
    a.raw = raw_out;
    a.vol = static_cast<float*>(volume->ptr);
    a.sx = volume->stride[1]; a.sy = volume->stride[2]; a.sz = volume->stride[3];
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    after_launch(launch_volume_raw(2, true, a, s), "dvr_volume_to_raw", s);
  });
}

int dvr_downsample_volume(const DvrTensor* volume, const DvrTensor* volume_out, void* stream) {
  return guarded([&] {
    if (!volume || !volume_out) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    DownsampleArgs a{};
    volume_dims(*volume, "volume", a.X, a.Y, a.Z);
    volume_dims(*volume_out, "mipmap level", a.oX, a.oY, a.oZ);
    if (a.oX > a.X || a.oY > a.Y || a.oZ > a.Z) fail(DVR_ERR_INVALID_ARGUMENT, "a mipmap level cannot be larger than the volume");
    a.in = static_cast<const float*>(volume->ptr);
    a.sx = volume->stride[1]; a.sy = volume->stride[2]; a.sz = volume->stride[3];
    a.out = static_cast<float*>(volume_out->ptr);
    a.ox = volume_out->stride[1]; a.oy = volume_out->stride[2]; a.oz = volume_out->stride[3];
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    after_launch(launch_downsample_volume(a, s), "dvr_downsample_volume", s);
  });
}

int dvr_count_cell_visits(const DvrInputs* inputs, const DvrTensor* termination_index, uint64_t* counts,
                          void* stream) {
  return guarded([&] {
    if (!inputs || !termination_index || !counts) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    Checked c;
    check_inputs(*inputs, c);
    check_dim(*termination_index, "terminationIndex", 3);
    check_size(*termination_index, "terminationIndex", 0, c.B);
    check_size(*termination_index, "terminationIndex", 1, c.H);
    check_size(*termination_index, "terminationIndex", 2, c.W);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    after_launch(launch_count_cells(c.ri, static_cast<const int*>(termination_index->ptr),
                                    termination_index->stride[0], termination_index->stride[1],
                                    termination_index->stride[2],
                                    reinterpret_cast<unsigned long long*>(counts), tile_grid(c), s),
                 "dvr_count_cell_visits", s);
  });
}

int dvr_has_variant_pass(int32_t pass, int32_t filter, int32_t camera, int32_t tf, int32_t blend) {
  if (camera < 0 || camera > 2) return 0;  // the camera mode is a run-time switch inside every instance
  if (blend != DVR_BLEND_BEER_LAMBERT && blend != DVR_BLEND_ALPHA) return 0;
  if (filter == DVR_FILTER_PRESHADED || tf == DVR_TF_PRESHADED)  // only valid together; dvr_preshaded.cu, no forward-mode
    return filter == DVR_FILTER_PRESHADED && tf == DVR_TF_PRESHADED && pass != DVR_PASS_FORWARD_GRADIENTS &&
           (pass == DVR_PASS_FORWARD || pass == DVR_PASS_ADJOINT);
  if (filter != DVR_FILTER_NEAREST && filter != DVR_FILTER_TRILINEAR) return 0;
  // membership in the SAME lists that declare and dispatch the instance translation units (dvr_launch.h)
#define DVR_MATCH(TF, BL) if (tf == TF && blend == BL) return 1;
  switch (pass) {
    case DVR_PASS_FORWARD: { DVR_FOR_EACH_FORWARD_INSTANCE(DVR_MATCH) } break;
    case DVR_PASS_FORWARD_GRADIENTS: { DVR_FOR_EACH_FWDGRAD_INSTANCE(DVR_MATCH) } break;
    case DVR_PASS_ADJOINT: { DVR_FOR_EACH_ADJOINT_INSTANCE(DVR_MATCH) } break;
    default: break;
  }
#undef DVR_MATCH
  return 0;
}

int dvr_has_variant(int32_t filter, int32_t camera, int32_t tf, int32_t blend) {
  return dvr_has_variant_pass(DVR_PASS_FORWARD, filter, camera, tf, blend) &&
         dvr_has_variant_pass(DVR_PASS_ADJOINT, filter, camera, tf, blend);
}

int dvr_set_option(const char* name, int32_t value) {
  if (!name) return DVR_ERR_INVALID_ARGUMENT;
  if (!std::strcmp(name, "sync_mode")) { g_sync_mode = value; return DVR_OK; }
  g_error = std::string("unknown option '") + name + "' (results-changing options are per call: DvrCallOptions)";
  return DVR_ERR_INVALID_ARGUMENT;
}

int dvr_get_option(const char* name, int32_t* value) {
  if (!name || !value) return DVR_ERR_INVALID_ARGUMENT;
  if (!std::strcmp(name, "sync_mode")) { *value = g_sync_mode; return DVR_OK; }
  g_error = std::string("unknown option '") + name + "'";
  return DVR_ERR_INVALID_ARGUMENT;
}

int dvr_get_stat(const char* name, int64_t* value) {
  if (!name || !value) return DVR_ERR_INVALID_ARGUMENT;
  if (!std::strcmp(name, "launches")) { *value = g_launches.load(); return DVR_OK; }
  if (!std::strcmp(name, "quad_builds")) { *value = g_quad_builds.load(); return DVR_OK; }
  if (!std::strcmp(name, "quad_reuses")) { *value = g_quad_reuses.load(); return DVR_OK; }
  if (!std::strcmp(name, "quad_fallbacks")) { *value = g_quad_fallbacks.load(); return DVR_OK; }
  if (!std::strcmp(name, "folds")) { *value = g_folds.load(); return DVR_OK; }
  g_error = std::string("unknown counter '") + name + "'";
  return DVR_ERR_INVALID_ARGUMENT;
}

void dvr_call_options_init(DvrCallOptions* o) {
  if (!o) return;
  std::memset(o, 0, sizeof(*o));
  o->struct_size = sizeof(DvrCallOptions);
  o->volume_layout = -1;
}

DvrWorkspace* dvr_workspace_create(void) {
  DvrWorkspace* w = new (std::nothrow) DvrWorkspace();
  if (!w || cudaGetDevice(&w->device) != cudaSuccess) {
    g_error = "dvr_workspace_create: no CUDA device";
    delete w;
    return nullptr;
  }
  return w;
}

void dvr_workspace_destroy(DvrWorkspace* w) {
  if (!w) return;
  if (w->quads) cudaFree(w->quads);
  if (w->grad) cudaFree(w->grad);
  delete w;
}

int64_t dvr_workspace_bytes(const DvrWorkspace* w) { return w ? int64_t(w->quadBytes + w->gradBytes) : 0; }

int dvr_workspace_invalidate(DvrWorkspace* w) {
  return guarded([&] {
    if (!w) fail(DVR_ERR_INVALID_ARGUMENT, "NULL workspace");
    std::lock_guard<std::mutex> lock(w->m);
    w->quadValid = false;
  });
}

int dvr_workspace_set_volume(DvrWorkspace* w, const DvrTensor* volume, void* stream) {
  return guarded([&] {
    if (!w || !volume) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    RayInputs r;
    int nb = 1;
    volume_fields(*volume, "volume", r, nb);
    if (r.X < 2 || r.Y < 2 || r.Z < 2 || (long long)r.X * r.Y * r.Z >= (1ll << 31))
      fail(DVR_ERR_UNSUPPORTED, "the QUAD layout needs 2 <= X, Y, Z and fewer than 2^31 voxels");
    if (r.volB == 0) nb = 1;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    std::lock_guard<std::mutex> lock(w->m);
    w->quadValid = false;
    if (!ws_reserve(w->quads, w->quadBytes, size_t(quad_count(r.X, r.Y, r.Z)) * nb * sizeof(float4)))
      fail(DVR_ERR_CUDA, "dvr_workspace_set_volume: out of device memory for the QUAD copy");
    after_launch(launch_build_quads(r, nb, w->quads, s), "dvr_build_quads", s);
    g_quad_builds.fetch_add(1, std::memory_order_relaxed);
    ws_remember_volume(*w, r, nb);
  });
}

int dvr_workspace_begin_volume_gradient(DvrWorkspace* w, const DvrTensor* adj_volume, void* stream) {
  return guarded([&] {
    if (!w || !adj_volume) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    RayInputs r;
    int nb = 1;
    volume_fields(*adj_volume, "adj_volume", r, nb);
    if (r.volB == 0) nb = 1;
    if (aliased_spatial_strides(*adj_volume))
      fail(DVR_ERR_UNSUPPORTED, "adj_volume views with stride-0 spatial dimensions accumulate in place only");
    if (r.X < 2 || r.Y < 2 || r.Z < 2 || (long long)r.X * r.Y * r.Z >= (1ll << 31))
      fail(DVR_ERR_UNSUPPORTED, "the QUAD layout needs 2 <= X, Y, Z and fewer than 2^31 voxels");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    std::lock_guard<std::mutex> lock(w->m);
    if (w->accumulating) fail(DVR_ERR_INVALID_ARGUMENT, "a volume-gradient scope is already open on this workspace");
    const long long voxels = (long long)r.X * r.Y * r.Z;
    const size_t bytes = size_t(quad_count(r.X, r.Y, r.Z)) * nb * sizeof(float4);
    if (!ws_reserve(w->grad, w->gradBytes, bytes))
      fail(DVR_ERR_CUDA, "dvr_workspace_begin_volume_gradient: out of device memory for the gradient scratch");
    cuda_check(cudaMemsetAsync(w->grad, 0, bytes, s), "dvr_workspace_begin_volume_gradient");
    w->accumulating = true;
    w->adjPtr = adj_volume->ptr;
    w->adjVoxels = voxels;
    w->adjNb = nb;
  });
}

int dvr_workspace_end_volume_gradient(DvrWorkspace* w, const DvrTensor* adj_volume, void* stream) {
  return guarded([&] {
    if (!w || !adj_volume) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    RayInputs r;
    int nb = 1;
    volume_fields(*adj_volume, "adj_volume", r, nb);
    if (r.volB == 0) nb = 1;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    std::lock_guard<std::mutex> lock(w->m);
    if (!w->accumulating) fail(DVR_ERR_INVALID_ARGUMENT, "no volume-gradient scope is open on this workspace");
    if (w->adjPtr != adj_volume->ptr || w->adjVoxels != (long long)r.X * r.Y * r.Z || w->adjNb != nb)
      fail(DVR_ERR_INVALID_ARGUMENT, "adj_volume differs from the tensor the scope was opened with");
    AdjointArgs aa;
    std::memset(&aa, 0, sizeof(aa));
    aa.adjVolume = static_cast<float*>(adj_volume->ptr);
    aa.avB = r.volB; aa.avx = r.sx; aa.avy = r.sy; aa.avz = r.sz;
    aa.gradQuads = w->grad;
    aa.gradQuadB = nb == 1 ? 0 : quad_count(r.X, r.Y, r.Z);
    w->accumulating = false;
    after_launch(launch_fold_quads(r, aa, nb, s), "dvr_fold_quads", s);
    g_folds.fetch_add(1, std::memory_order_relaxed);
  });
}

int dvr_trim_scratch(void) {
  return guarded([&] {
    int dev = 0;
    cuda_check(cudaGetDevice(&dev), "dvr_trim_scratch");
    std::lock_guard<std::mutex> lock(g_pool_mutex);
    if (dev >= 0 && dev < kMaxDevices && g_pools[dev]) cuda_check(cudaMemPoolTrimTo(g_pools[dev], 0), "dvr_trim_scratch");
  });
}

int dvr_device_synchronize(void) {
  return guarded([&] { cuda_check(cudaDeviceSynchronize(), "dvr_device_synchronize"); });
}

struct DvrTimer {
  cudaEvent_t start, stop;
};
DvrTimer* dvr_timer_create(void) {
  DvrTimer* t = new DvrTimer{nullptr, nullptr};
  if (cudaEventCreate(&t->start) != cudaSuccess || cudaEventCreate(&t->stop) != cudaSuccess) {
    g_error = "cudaEventCreate failed";
    dvr_timer_destroy(t);
    return nullptr;
  }
  return t;
}
void dvr_timer_destroy(DvrTimer* t) {
  if (!t) return;
  if (t->start) cudaEventDestroy(t->start);
  if (t->stop) cudaEventDestroy(t->stop);
  delete t;
}
int dvr_timer_start(DvrTimer* t, void* stream) {
  return guarded([&] {
    if (!t) fail(DVR_ERR_INVALID_ARGUMENT, "NULL timer");
    cuda_check(cudaEventRecord(t->start, static_cast<cudaStream_t>(stream)), "dvr_timer_start");
  });
}
int dvr_timer_stop(DvrTimer* t, void* stream) {
  return guarded([&] {
    if (!t) fail(DVR_ERR_INVALID_ARGUMENT, "NULL timer");
    cuda_check(cudaEventRecord(t->stop, static_cast<cudaStream_t>(stream)), "dvr_timer_stop");
  });
}
int dvr_timer_elapsed_ms(DvrTimer* t, float* ms) {
  return guarded([&] {
    if (!t || !ms) fail(DVR_ERR_INVALID_ARGUMENT, "NULL argument");
    cuda_check(cudaEventSynchronize(t->stop), "dvr_timer_elapsed_ms");
    cuda_check(cudaEventElapsedTime(ms, t->start, t->stop), "dvr_timer_elapsed_ms");
  });
}

int64_t dvr_launch_count(void) { return g_launches.load(); }
const char* dvr_last_error(void) { return g_error.c_str(); }
int32_t dvr_abi_version(void) { return DVR_ABI_VERSION; }

}  // extern "C"
