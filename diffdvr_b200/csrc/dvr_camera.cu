// Callers either side of the marcher ("next" rows f1/f2 of SURVEY.md section 8):
//   * Camera::generateRays forward (renderer/camera_impl.cu:17-94) and adjoint (:96-356, a
//     cub::DeviceSegmentedReduce over a transform iterator in the reference)
//   * Renderer::compareToImage (renderer/renderer_compareToImage.cuh:9-65,
//     renderer/renderer_cuda_impl.cu:9-172; the reference's CUDA path only accepts D == 1)
// Both are streaming kernels bounded by HBM: 24 B written (forward) / read (adjoint) per pixel for
// the rays, (32 + 16 D) B read and (4 + 4 D) B written per pixel for the comparison.
#include "dvr_device.cuh"
#include "dvr_launch.h"

namespace dvr {

namespace {

struct Frame {
  f3 eye, right, up, front;
};

DVR_DEV Frame load_frame(const GenRaysArgs& a, int b) {
  auto m = [&](int r, int c) { return __ldg(a.viewport.p + b * a.viewport.b + r * a.viewport.r + c * a.viewport.c); };
  Frame f;
  f.eye = {m(0, 0), m(0, 1), m(0, 2)};
  f.right = {m(1, 0), m(1, 1), m(1, 2)};
  f.up = {m(2, 0), m(2, 1), m(2, 2)};
  // cross(up, right), helper_math.cuh cross: a.y*b.z - a.z*b.y, ... without contraction
  f.front = {__fsub_rn(__fmul_rn(f.up.y, f.right.z), __fmul_rn(f.up.z, f.right.y)),
             __fsub_rn(__fmul_rn(f.up.z, f.right.x), __fmul_rn(f.up.x, f.right.z)),
             __fsub_rn(__fmul_rn(f.up.x, f.right.y), __fmul_rn(f.up.y, f.right.x))};
  return f;
}

// un-normalised direction of pixel (x, y): front + fx*tanFovX*right + fy*tanFovY*up
// (camera_impl.cu:28-33), same operation order as the marcher's ReferenceFrame camera (make_ray)
DVR_DEV f3 pixel_dir(const GenRaysArgs& a, const Frame& f, int x, int y, float& ax, float& ay) {
  const float fx = 2 * (x + 0.5f) / float(a.W) - 1;
  const float fy = 2 * (y + 0.5f) / float(a.H) - 1;
  ax = __fmul_rn(fx, a.tanFovX);
  ay = __fmul_rn(fy, a.tanFovY);
  return {__fadd_rn(__fadd_rn(f.front.x, __fmul_rn(ax, f.right.x)), __fmul_rn(ay, f.up.x)),
          __fadd_rn(__fadd_rn(f.front.y, __fmul_rn(ax, f.right.y)), __fmul_rn(ay, f.up.y)),
          __fadd_rn(__fadd_rn(f.front.z, __fmul_rn(ax, f.right.z)), __fmul_rn(ay, f.up.z))};
}

DVR_DEV float dot3(const f3& a, const f3& b) {
  return __fadd_rn(__fadd_rn(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y)), __fmul_rn(a.z, b.z));
}

// ---- generateRays forward ----------------------------------------------------------------------
// packed: one thread writes 4 consecutive pixels = 12 floats = 3 x STG.128 per tensor
__global__ void __launch_bounds__(256) k_generate_rays(const GenRaysArgs a) {
  const int b = blockIdx.z;
  const Frame f = load_frame(a, b);
  const int per = a.packed ? 4 : 1;
  const unsigned items = unsigned(a.W) * unsigned(a.H) / per;  // W*H < 2^31 (checked by the C-ABI)
  const unsigned it = blockIdx.x * blockDim.x + threadIdx.x;
  if (it >= items) return;
  const unsigned pix = it * per;
  const int y = int(pix / unsigned(a.W)), x0 = int(pix - unsigned(y) * unsigned(a.W));
  float d[12];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if (k < per) {
      float ax, ay;
      const f3 v = pixel_dir(a, f, x0 + k, y, ax, ay);
      const float inv = 1.0f / sqrtf(dot3(v, v));
      d[3 * k] = v.x * inv; d[3 * k + 1] = v.y * inv; d[3 * k + 2] = v.z * inv;
    }
  }
  if (a.packed) {
    float4* ps = reinterpret_cast<float4*>(a.rayStart + ((long long)b * a.H * a.W + pix) * 3);
    float4* pd = reinterpret_cast<float4*>(a.rayDir + ((long long)b * a.H * a.W + pix) * 3);
    const f3 e = f.eye;
    ps[0] = make_float4(e.x, e.y, e.z, e.x);
    ps[1] = make_float4(e.y, e.z, e.x, e.y);
    ps[2] = make_float4(e.z, e.x, e.y, e.z);
    pd[0] = make_float4(d[0], d[1], d[2], d[3]);
    pd[1] = make_float4(d[4], d[5], d[6], d[7]);
    pd[2] = make_float4(d[8], d[9], d[10], d[11]);
  } else {
    float* ps = a.rayStart + b * a.sB + y * a.sY + x0 * a.sX;
    float* pd = a.rayDir + b * a.dB + y * a.dY + x0 * a.dX;
    ps[0] = f.eye.x; ps[a.sC] = f.eye.y; ps[2 * a.sC] = f.eye.z;
    pd[0] = d[0]; pd[a.dC] = d[1]; pd[2 * a.dC] = d[2];
  }
}

// ---- generateRays adjoint ----------------------------------------------------------------------
// adj_viewport[b] = sum over the pixels of view b. The chain rule through
// dir = front + ax*right + ay*up is linear in adj_dir, so each thread sums adj_eye, adj_dir,
// ax*adj_dir and ay*adj_dir; the cross-product adjoint (renderer_adjoint.cuh:46-64) is applied once
// per CTA to the CTA's partial sums, followed by 9 atomics per CTA.
__global__ void __launch_bounds__(256) k_zero_viewport(const GenRaysArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < a.B * 9) {
    const int b = i / 9, r = (i % 9) / 3, c = i % 3;
    a.adjViewport[b * a.avB + r * a.avR + c * a.avC] = 0.f;
  }
}

__global__ void __launch_bounds__(256) k_generate_rays_adjoint(const GenRaysArgs a) {
  __shared__ float s_part[8][12];
  const int b = blockIdx.z;
  const Frame f = load_frame(a, b);
  const int per = a.packed ? 4 : 1;
  const unsigned items = unsigned(a.W) * unsigned(a.H) / per;
  float acc[12];
#pragma unroll
  for (int i = 0; i < 12; ++i) acc[i] = 0.f;
  for (unsigned it = blockIdx.x * blockDim.x + threadIdx.x; it < items; it += gridDim.x * blockDim.x) {
    const unsigned pix = it * per;
    const int y = int(pix / unsigned(a.W)), x0 = int(pix - unsigned(y) * unsigned(a.W));
    float gs[12], gd[12];
    if (a.packed) {
      const float4* ps = reinterpret_cast<const float4*>(a.rayStart + ((long long)b * a.H * a.W + pix) * 3);
      const float4* pd = reinterpret_cast<const float4*>(a.rayDir + ((long long)b * a.H * a.W + pix) * 3);
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const float4 s = __ldg(ps + k), d = __ldg(pd + k);
        gs[4 * k] = s.x; gs[4 * k + 1] = s.y; gs[4 * k + 2] = s.z; gs[4 * k + 3] = s.w;
        gd[4 * k] = d.x; gd[4 * k + 1] = d.y; gd[4 * k + 2] = d.z; gd[4 * k + 3] = d.w;
      }
    } else {
      const float* ps = a.rayStart + b * a.sB + y * a.sY + x0 * a.sX;
      const float* pd = a.rayDir + b * a.dB + y * a.dY + x0 * a.dX;
#pragma unroll
      for (int c = 0; c < 3; ++c) { gs[c] = __ldg(ps + c * a.sC); gd[c] = __ldg(pd + c * a.dC); }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (k < per) {
        float ax, ay;
        const f3 v = pixel_dir(a, f, x0 + k, y, ax, ay);
        const f3 g = {gd[3 * k], gd[3 * k + 1], gd[3 * k + 2]};
        // adjNormalize (renderer_adjoint.cuh:28-44): adj_in = g/len - <g, v>/len^3 * v
        const float len = sqrtf(dot3(v, v));
        const float s = -dot3(g, v) / (len * len) / len;
        const f3 ad = {g.x / len + s * v.x, g.y / len + s * v.y, g.z / len + s * v.z};
        acc[0] += gs[3 * k]; acc[1] += gs[3 * k + 1]; acc[2] += gs[3 * k + 2];
        acc[3] += ad.x; acc[4] += ad.y; acc[5] += ad.z;
        acc[6] = fmaf(ad.x, ax, acc[6]); acc[7] = fmaf(ad.y, ax, acc[7]); acc[8] = fmaf(ad.z, ax, acc[8]);
        acc[9] = fmaf(ad.x, ay, acc[9]); acc[10] = fmaf(ad.y, ay, acc[10]); acc[11] = fmaf(ad.z, ay, acc[11]);
      }
    }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int i = 0; i < 12; ++i) {
    const float v = warp_sum(acc[i]);
    if (lane == 0) s_part[warp][i] = v;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float t[12];
    for (int i = 0; i < 12; ++i) {
      float v = 0.f;
      for (int w = 0; w < 8; ++w) v += s_part[w][i];
      t[i] = v;
    }
    const f3 c = {t[3], t[4], t[5]};            // adj_front
    f3 ar = {t[6], t[7], t[8]};                 // adj_right (b of cross(a = up, b = right))
    f3 au = {t[9], t[10], t[11]};               // adj_up
    const f3 A = f.up, Bv = f.right;
    au.y += Bv.z * c.x; ar.z += A.y * c.x; au.z -= Bv.y * c.x; ar.y -= A.z * c.x;
    au.z += Bv.x * c.y; ar.x += A.z * c.y; au.x -= Bv.z * c.y; ar.z -= A.x * c.y;
    au.x += Bv.y * c.z; ar.y += A.x * c.z; au.y -= Bv.x * c.z; ar.x -= A.y * c.z;
    float* o = a.adjViewport + b * a.avB;
    atomicAdd(o, t[0]); atomicAdd(o + a.avC, t[1]); atomicAdd(o + 2 * a.avC, t[2]);
    atomicAdd(o + a.avR, ar.x); atomicAdd(o + a.avR + a.avC, ar.y); atomicAdd(o + a.avR + 2 * a.avC, ar.z);
    atomicAdd(o + 2 * a.avR, au.x); atomicAdd(o + 2 * a.avR + a.avC, au.y); atomicAdd(o + 2 * a.avR + 2 * a.avC, au.z);
  }
}

// ---- compareToImage ------------------------------------------------------------------------------
// value = |input - reference|_2 over RGBA, d value / d p = <input - reference, d input / d p> / value
// (cudAD::length). reduce: mean over all B*H*W pixels (each term scaled by 1/(B*H*W) before the sum,
// renderer_cuda_impl.cu:59), pre-reduced per warp and per CTA, D+1 atomics per CTA.
DVR_DEV f4 load4(const float* p, long long c) {
  if (c == 1 && (reinterpret_cast<unsigned long long>(p) & 15ull) == 0) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(p));
    return {v.x, v.y, v.z, v.w};
  }
  return {__ldg(p), __ldg(p + c), __ldg(p + 2 * c), __ldg(p + 3 * c)};
}

__global__ void __launch_bounds__(256) k_zero_compare(const CompareArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) a.diffOut[0] = 0.f;
  else if (i <= a.D) a.gradOut[(i - 1) * a.qD] = 0.f;
}

__global__ void __launch_bounds__(256) k_compare_to_image(const CompareArgs a) {
  extern __shared__ float s_sum[];  // reduce: D + 1
  if (a.reduce) {
    for (int i = threadIdx.x; i <= a.D; i += blockDim.x) s_sum[i] = 0.f;
    __syncthreads();
  }
  const unsigned plane = unsigned(a.W) * unsigned(a.H);  // < 2^31 (checked by the C-ABI)
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.z;
  // uniform trip count: every lane takes part in the shuffles of the reduced variant
  const unsigned span = gridDim.x * blockDim.x;
  const unsigned rounds = (plane + span - 1) / span;
  for (unsigned r = 0; r < rounds; ++r) {
    const unsigned i = r * span + blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = i < plane;
    int y = 0, x = 0;
    if (active) {
      y = int(i / unsigned(a.W));
      x = int(i - unsigned(y) * unsigned(a.W));
    }
    f4 diff = {0.f, 0.f, 0.f, 0.f};
    float len = 0.f;
    if (active) {
      const f4 c = load4(a.color.p + b * a.color.b + y * a.color.y + x * a.color.x, a.color.c);
      const f4 q = load4(a.reference.p + b * a.reference.b + y * a.reference.y + x * a.reference.x, a.reference.c);
      diff = {c.x - q.x, c.y - q.y, c.z - q.z, c.w - q.w};
      len = sqrtf(dot4(diff, diff));
    }
    if (a.reduce) {
      const float v = warp_sum(active ? len * a.normalization : 0.f);
      if (lane == 0) atomicAdd(s_sum, v);
    } else if (active) {
      a.diffOut[b * a.oB + y * a.oY + x * a.oX] = len;
    }
    for (int d = 0; d < a.D; ++d) {
      float g = 0.f;
      if (active) {
        const f4 j = load4(a.grad + b * a.gB + y * a.gY + x * a.gX + d * a.gD, a.gC);
        g = dot4(diff, j) / len;
      }
      if (a.reduce) {
        const float v = warp_sum(active ? g * a.normalization : 0.f);
        if (lane == 0) atomicAdd(s_sum + 1 + d, v);
      } else if (active) {
        a.gradOut[b * a.qB + y * a.qY + x * a.qX + d * a.qD] = g;
      }
    }
  }
  if (a.reduce) {
    __syncthreads();
    for (int i = threadIdx.x; i <= a.D; i += blockDim.x) {
      if (i == 0) atomicAdd(a.diffOut, s_sum[0]);
      else atomicAdd(a.gradOut + (i - 1) * a.qD, s_sum[i]);
    }
  }
}

}  // namespace

cudaError_t launch_generate_rays(const GenRaysArgs& a, cudaStream_t s) {
  const long long items = (long long)a.W * a.H / (a.packed ? 4 : 1);
  const dim3 grid((unsigned)((items + 255) / 256), 1, a.B);
  return launch_kernel(k_generate_rays, grid, 256, 0, s, a);
}

cudaError_t launch_generate_rays_adjoint(const GenRaysArgs& a, cudaStream_t s) {
  cudaError_t e = launch_kernel(k_zero_viewport, dim3((a.B * 9 + 255) / 256), 256, 0, s, a);
  if (e != cudaSuccess) return e;
  const long long items = (long long)a.W * a.H / (a.packed ? 4 : 1);
  const dim3 grid((unsigned)((items + 1023) / 1024), 1, a.B);  // up to 4 items per thread
  return launch_kernel(k_generate_rays_adjoint, grid, 256, 0, s, a);
}

cudaError_t launch_compare_to_image(const CompareArgs& a, cudaStream_t s) {
  const long long plane = (long long)a.W * a.H;
  long long blocks = (plane + 255) / 256;
  if (a.reduce) {
    cudaError_t e = launch_kernel(k_zero_compare, dim3((a.D + 256) / 256), 256, 0, s, a);
    if (e != cudaSuccess) return e;
    // about one wave of 8 CTAs per SM over all views, grid-stride inside a view
    const long long cap = (148 * 8 + a.B - 1) / a.B;
    if (blocks > cap) blocks = cap;
  }
  const size_t smem = a.reduce ? size_t(a.D + 1) * sizeof(float) : 0;
  return launch_kernel(k_compare_to_image, dim3((unsigned)blocks, 1, a.B), 256, smem, s, a);
}

}  // namespace dvr
