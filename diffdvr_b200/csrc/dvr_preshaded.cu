// sm_100a kernels for PRE-SHADED volumes (FilterPreshaded + TFPreshaded): the volume tensor holds
// (r, g, b, absorption) in its batch dimension, 4*X*Y*Z, the transfer function is the identity on
// those four channels (renderer/renderer_tf.cuh:819-854) and the sampler interpolates all four
// (renderer/renderer_interpolation.cuh:393-579). Same tiling as the scalar kernels: one thread per
// ray, 8x4-pixel warp tiles, per-cell register cache (4 x 8 corners), RED flush on a cell change.
#include "dvr_device.cuh"
#include "dvr_launch.h"

namespace dvr {

template <bool ADJ_VOLUME>
struct Sampler4 {
  const float* vol;  // channel 0
  float* adjVol;
  long long cs, acs;  // channel strides (the batch stride of the 4*X*Y*Z tensors)
  float ox, oy, oz;   // origin of the cached cell (NaN: none)
  float d[4][8];      // [channel][corner], corner bit2 = x, bit1 = y, bit0 = z
  float a[4][8];
  unsigned aoff[8];   // element offsets of the 8 corners in adj_volume (clamped)
  bool have;
  float fx, fy, fz;

  DVR_DEV void init(const RayInputs& in, float* adjVolume, long long avB) {
    vol = in.volume;
    cs = in.volB;
    adjVol = adjVolume;
    acs = avB;
    ox = oy = oz = __int_as_float(0x7fc00000);
    have = false;
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
      for (int i = 0; i < 8; ++i) { d[c][i] = 0.f; a[c][i] = 0.f; }
#pragma unroll
    for (int i = 0; i < 8; ++i) aoff[i] = 0u;
    fx = fy = fz = 0.f;
  }
  DVR_DEV static void red(float* p, float v) {
    asm volatile("red.global.add.f32 [%0], %1;" ::"l"(__cvta_generic_to_global(p)), "f"(v));
  }
  DVR_DEV void flush() {
    if constexpr (ADJ_VOLUME) {
      if (!have) return;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        float* base = adjVol + c * acs;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          if (a[c][i] != 0.f) red(base + aoff[i], a[c][i]);
          a[c][i] = 0.f;
        }
      }
    }
  }
  DVR_DEV void load_cell(const RayInputs& in, const AdjointArgs* adj, float px, float py, float pz) {
    flush();
    have = true;
    const float bx = biased_trunc(px), by = biased_trunc(py), bz = biased_trunc(pz);
    ox = bx - kTwo23; oy = by - kTwo23; oz = bz - kTwo23;
    const int cx = __float_as_int(bx) - kTwo23Bits, cy = __float_as_int(by) - kTwo23Bits, cz = __float_as_int(bz) - kTwo23Bits;
    // corners clamped into the volume (renderer_interpolation.cuh:403-404, 459-460)
    const unsigned lx = min(cx, in.X - 1), hx = min(cx + 1, in.X - 1);
    const unsigned ly = min(cy, in.Y - 1), hy = min(cy + 1, in.Y - 1);
    const unsigned lz = min(cz, in.Z - 1), hz = min(cz + 1, in.Z - 1);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const unsigned x = (i & 4) ? hx : lx, y = (i & 2) ? hy : ly, z = (i & 1) ? hz : lz;
      const unsigned off = x * unsigned(in.sx) + y * unsigned(in.sy) + z * unsigned(in.sz);
#pragma unroll
      for (int c = 0; c < 4; ++c) d[c][i] = __ldg(vol + c * cs + off);
      if constexpr (ADJ_VOLUME) aoff[i] = x * unsigned(adj->avx) + y * unsigned(adj->avy) + z * unsigned(adj->avz);
    }
  }
  DVR_DEV f4 fetch(const RayInputs& in, const AdjointArgs* adj, float px, float py, float pz) {
    fx = px - ox; fy = py - oy; fz = pz - oz;
    constexpr unsigned kOne = 0x3f800000u;
    if (!(__float_as_uint(fx) < kOne && __float_as_uint(fy) < kOne && __float_as_uint(fz) < kOne)) {
      load_cell(in, adj, px, py, pz);
      fx = px - ox; fy = py - oy; fz = pz - oz;
    }
    float ch[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const float c00 = lerpf(d[c][0], d[c][4], fx), c10 = lerpf(d[c][2], d[c][6], fx);
      const float c01 = lerpf(d[c][1], d[c][5], fx), c11 = lerpf(d[c][3], d[c][7], fx);
      ch[c] = lerpf(lerpf(c00, c10, fy), lerpf(c01, c11, fy), fz);
    }
    return {ch[0], ch[1], ch[2], ch[3]};
  }
  // adjoint of the last fetch (renderer_interpolation.cuh:548-578). WITH_POS reproduces the
  // reference's position adjoint INCLUDING its macro typo (:561): the z-derivative of a corner is
  // credited to the vector selected by the corner's Y weight (adj_<Y>.z), not by its Z weight.
  template <bool WITH_POS>
  DVR_DEV f3 adjoint(const f4& adjC) {
    const float cw[3] = {fx, fy, fz};
    const float fw[3] = {1.f - fx, 1.f - fy, 1.f - fz};
    const float av[4] = {adjC.x, adjC.y, adjC.z, adjC.w};
    f3 g = {0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < 4; ++c) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const bool hx = i & 4, hy = i & 2, hz = i & 1;
        const float wx = hx ? cw[0] : fw[0], wy = hy ? cw[1] : fw[1], wz = hz ? cw[2] : fw[2];
        if constexpr (ADJ_VOLUME) a[c][i] = fmaf(av[c], wx * wy * wz, a[c][i]);
        if constexpr (WITH_POS) {
          const float ad = av[c] * d[c][i];
          g.x += (hx ? ad : -ad) * (wy * wz);
          g.y += (hy ? ad : -ad) * (wx * wz);
          g.z += (hy ? ad : -ad) * (wx * wy);
        }
      }
    }
    return g;
  }
};

// DvrKernelForwardImpl with FilterPreshaded / TFPreshaded (renderer_kernels.cuh:22-90)
template <int BLEND>
__global__ void __launch_bounds__(kThreads) k_forward_preshaded(const RayInputs in, const ForwardOutputs out) {
  const int b = blockIdx.z;
  int x, y;
  thread_pixel(x, y);
  if (x >= in.W || y >= in.H) return;
  const Ray ray = make_ray(in, x, y, b);
  Sampler4<false> smp;
  smp.init(in, nullptr, 0);
  f4 acc = {0.f, 0.f, 0.f, 0.f};
  float fi = 0.f;
  for (;; fi += 1.0f) {
    const float t = sample_t(ray, fi);
    if (t > ray.tmax) break;
    if (acc.w > in.earlyOut) break;
    const f3 p = sample_pos<false>(ray, t);
    const f4 cur = smp.fetch(in, nullptr, p.x, p.y, p.z);
    acc = blend_over(acc, cur, blend_alpha<BLEND>(cur.w, ray.step));
  }
  float* c = out.color + b * out.cB + y * out.cY + x * out.cX;
  c[0] = acc.x; c[out.cC] = acc.y; c[2 * out.cC] = acc.z; c[3 * out.cC] = acc.w;
  out.term[b * out.tB + y * out.tY + x * out.tX] = int(fi);
}

// DvrKernelAdjointImpl with FilterPreshaded / TFPreshaded (renderer_kernels.cuh:315-444); the TF
// adjoint passes the colour adjoint through and has no parameters (renderer_tf.cuh:827-834)
template <int BLEND, bool HAS_VOL, bool HAS_RAY>
__global__ void __launch_bounds__(kThreads) k_adjoint_preshaded(const RayInputs in, const AdjointArgs aa) {
  const int b = blockIdx.z;
  int x, y;
  thread_pixel(x, y);
  if (x >= in.W || y >= in.H) return;
  const Ray ray = make_ray(in, x, y, b);
  const float* cp = aa.color.p + b * aa.color.b + y * aa.color.y + x * aa.color.x;
  const float* ap = aa.adjColor.p + b * aa.adjColor.b + y * aa.adjColor.y + x * aa.adjColor.x;
  f4 acc = {__ldg(cp), __ldg(cp + aa.color.c), __ldg(cp + 2 * aa.color.c), __ldg(cp + 3 * aa.color.c)};
  f4 adjAcc = {__ldg(ap), __ldg(ap + aa.adjColor.c), __ldg(ap + 2 * aa.adjColor.c), __ldg(ap + 3 * aa.adjColor.c)};
  const int n = __ldg(aa.term + b * aa.tB + y * aa.tY + x * aa.tX);
  float adjStep = 0.f;
  f3 adjStart = {0.f, 0.f, 0.f}, adjDir = {0.f, 0.f, 0.f};
  Sampler4<HAS_VOL> smp;
  smp.init(in, aa.adjVolume, aa.avB);
  float fi = float(n - 1);
  for (int i = n - 1; i >= 0; --i, fi -= 1.0f) {
    const float t = sample_t(ray, fi);
    const f3 p = sample_pos<HAS_RAY>(ray, t);
    const f4 cur = smp.fetch(in, &aa, p.x, p.y, p.z);
    f4 adjCur;
    float adjStepBlend;
    blend_adjoint<BLEND>(acc, adjAcc, cur, ray.step, adjCur, adjStepBlend);
    const f3 g = smp.template adjoint<HAS_RAY>(adjCur);
    if constexpr (HAS_RAY) {
      adjStep = aa.compatStepOverwrite ? adjStepBlend : adjStep + adjStepBlend;
      const f3 w = {g.x * ray.invVoxel.x, g.y * ray.invVoxel.y, g.z * ray.invVoxel.z};
      adjStart.x += w.x; adjStart.y += w.y; adjStart.z += w.z;
      adjDir.x = fmaf(w.x, t, adjDir.x); adjDir.y = fmaf(w.y, t, adjDir.y); adjDir.z = fmaf(w.z, t, adjDir.z);
      const float adjT = w.x * ray.dir.x + w.y * ray.dir.y + w.z * ray.dir.z;
      adjStep = fmaf(adjT, fi, adjStep);
    }
  }
  smp.flush();
  if constexpr (HAS_RAY) {
    if (aa.hasStep) {
      float* p = aa.adjStep + b * aa.asB + y * aa.asY + x * aa.asX;
      if (aa.stepAtomic) atomicAdd(p, adjStep);
      else *p = adjStep;
    }
    if (aa.hasCam) {
      float* ps = aa.adjRayStart + b * aa.rsB + y * aa.rsY + x * aa.rsX;
      float* pd = aa.adjRayDir + b * aa.rdB + y * aa.rdY + x * aa.rdX;
      if (aa.camAtomic) {
        atomicAdd(ps, adjStart.x); atomicAdd(ps + aa.rsC, adjStart.y); atomicAdd(ps + 2 * aa.rsC, adjStart.z);
        atomicAdd(pd, adjDir.x); atomicAdd(pd + aa.rdC, adjDir.y); atomicAdd(pd + 2 * aa.rdC, adjDir.z);
      } else {
        ps[0] = adjStart.x; ps[aa.rsC] = adjStart.y; ps[2 * aa.rsC] = adjStart.z;
        pd[0] = adjDir.x; pd[aa.rdC] = adjDir.y; pd[2 * aa.rdC] = adjDir.z;
      }
    }
  }
}

cudaError_t launch_forward_preshaded(int blend, const RayInputs& in, const ForwardOutputs& out, dim3 grid,
                                     cudaStream_t s) {
  if (blend == kBlendBeer) return launch_kernel(k_forward_preshaded<kBlendBeer>, grid, kThreads, 0, s, in, out);
  if (blend == kBlendAlpha) return launch_kernel(k_forward_preshaded<kBlendAlpha>, grid, kThreads, 0, s, in, out);
  return kErrNoInstance;
}

template <int BLEND>
static cudaError_t adj_by_flags(bool hasVol, bool hasRay, const RayInputs& in, const AdjointArgs& aa, dim3 grid,
                                cudaStream_t s) {
  if (hasVol && hasRay) return launch_kernel(k_adjoint_preshaded<BLEND, true, true>, grid, kThreads, 0, s, in, aa);
  if (hasVol) return launch_kernel(k_adjoint_preshaded<BLEND, true, false>, grid, kThreads, 0, s, in, aa);
  if (hasRay) return launch_kernel(k_adjoint_preshaded<BLEND, false, true>, grid, kThreads, 0, s, in, aa);
  return kErrNoInstance;
}

cudaError_t launch_adjoint_preshaded(int blend, bool hasVol, bool hasRay, const RayInputs& in, const AdjointArgs& aa,
                                     dim3 grid, cudaStream_t s) {
  if (blend == kBlendBeer) return adj_by_flags<kBlendBeer>(hasVol, hasRay, in, aa, grid, s);
  if (blend == kBlendAlpha) return adj_by_flags<kBlendAlpha>(hasVol, hasRay, in, aa, grid, s);
  return kErrNoInstance;
}

}  // namespace dvr
