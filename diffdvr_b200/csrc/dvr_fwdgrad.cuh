// sm_100a kernels: forward-mode differentiation (DvrKernelForwardGradientsImpl,
// renderer/renderer_kernels.cuh:161-255) and the forward-variables -> gradients contraction
// (ForwardVariablesToGradientsKernelImpl, :504-561).
//
// The reference instantiates cudAD::fvar<T,D> for the full D at run time through NVRTC; here
// the dual number carries a compile-time CHUNK of N tangent slots and the host runs
// ceil(D/N) passes, each re-marching the ray for the slots [d0, d0+N) and writing a disjoint
// slice of the B*H*W*D*4 Jacobian (the value part is identical in every pass).
#pragma once
#include "dvr_device.cuh"

namespace dvr {

template <int N>
struct Du {
  float v;
  float d[N];
};

template <int N> DVR_DEV Du<N> du_const(float v) {
  Du<N> r;
  r.v = v;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = 0.f;
  return r;
}
// fvar::input: unit tangent in slot (idx - d0) if that slot belongs to this pass
template <int N> DVR_DEV Du<N> du_input(float v, int idx, int d0) {
  Du<N> r;
  r.v = v;
  const int s = idx - d0;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = (idx >= 0 && i == s) ? 1.f : 0.f;
  return r;
}
template <int N> DVR_DEV float val(const Du<N>& a) { return a.v; }
DVR_DEV float val(float a) { return a; }

#define DVR_DU_LOOP _Pragma("unroll") for (int i = 0; i < N; ++i)
template <int N> DVR_DEV Du<N> operator+(const Du<N>& a, const Du<N>& b) { Du<N> r; r.v = a.v + b.v; DVR_DU_LOOP r.d[i] = a.d[i] + b.d[i]; return r; }
template <int N> DVR_DEV Du<N> operator+(const Du<N>& a, float b) { Du<N> r = a; r.v = a.v + b; return r; }
template <int N> DVR_DEV Du<N> operator+(float a, const Du<N>& b) { Du<N> r = b; r.v = a + b.v; return r; }
template <int N> DVR_DEV Du<N> operator-(const Du<N>& a, const Du<N>& b) { Du<N> r; r.v = a.v - b.v; DVR_DU_LOOP r.d[i] = a.d[i] - b.d[i]; return r; }
template <int N> DVR_DEV Du<N> operator-(const Du<N>& a, float b) { Du<N> r = a; r.v = a.v - b; return r; }
template <int N> DVR_DEV Du<N> operator-(float a, const Du<N>& b) { Du<N> r; r.v = a - b.v; DVR_DU_LOOP r.d[i] = -b.d[i]; return r; }
template <int N> DVR_DEV Du<N> operator-(const Du<N>& a) { Du<N> r; r.v = -a.v; DVR_DU_LOOP r.d[i] = -a.d[i]; return r; }
template <int N> DVR_DEV Du<N> operator*(const Du<N>& a, const Du<N>& b) { Du<N> r; r.v = a.v * b.v; DVR_DU_LOOP r.d[i] = fmaf(a.d[i], b.v, a.v * b.d[i]); return r; }
template <int N> DVR_DEV Du<N> operator*(const Du<N>& a, float b) { Du<N> r; r.v = a.v * b; DVR_DU_LOOP r.d[i] = a.d[i] * b; return r; }
template <int N> DVR_DEV Du<N> operator*(float a, const Du<N>& b) { return b * a; }
template <int N> DVR_DEV Du<N> operator/(const Du<N>& a, const Du<N>& b) {
  Du<N> r;
  const float inv = 1.0f / b.v;
  r.v = a.v * inv;
  DVR_DU_LOOP r.d[i] = (a.d[i] - r.v * b.d[i]) * inv;  // (a'b - ab')/b^2
  return r;
}
template <int N> DVR_DEV Du<N> operator/(const Du<N>& a, float b) { return a * (1.0f / b); }
template <int N> DVR_DEV Du<N> du_exp(const Du<N>& a) { Du<N> r; r.v = expf(a.v); DVR_DU_LOOP r.d[i] = a.d[i] * r.v; return r; }
DVR_DEV float du_exp(float a) { return expf(a); }
// exp of the Gaussian TF kernel: the same gauss_exp as the other kernels
template <int N> DVR_DEV Du<N> du_gauss_exp(const Du<N>& a) { Du<N> r; r.v = gauss_exp(a.v); DVR_DU_LOOP r.d[i] = a.d[i] * r.v; return r; }
// exp of the Beer-Lambert blend: the same blend_exp as the plain forward kernel, so both write the same colour
template <int N> DVR_DEV Du<N> du_blend_exp(const Du<N>& a) { Du<N> r; r.v = blend_exp(a.v); DVR_DU_LOOP r.d[i] = a.d[i] * r.v; return r; }
DVR_DEV float du_blend_exp(float a) { return blend_exp(a); }
// selection rules of the restated cudAD (oracle/cudad_restated/forward.h): a<b ? a : b
template <int N> DVR_DEV Du<N> du_min(const Du<N>& a, const Du<N>& b) { return a.v < b.v ? a : b; }
template <int N> DVR_DEV Du<N> du_max(const Du<N>& a, const Du<N>& b) { return a.v > b.v ? a : b; }
template <int N> DVR_DEV Du<N> du_clamp(const Du<N>& x, const Du<N>& lo, const Du<N>& hi) { return du_max(lo, du_min(x, hi)); }
template <int N> DVR_DEV Du<N> du_lerp(const Du<N>& a, const Du<N>& b, const Du<N>& t) { return a + t * (b - a); }
template <int N> DVR_DEV Du<N> du_lerp(float a, float b, const Du<N>& t) { return a + t * (b - a); }
template <int N> DVR_DEV Du<N> du_lerp(const Du<N>& a, const Du<N>& b, float t) { return a + (b - a) * t; }
template <int N> DVR_DEV Du<N> du_lerp(float a, float b, float t) { return du_const<N>(lerpf(a, b, t)); }
#undef DVR_DU_LOOP

template <int N> struct Du4 { Du<N> x, y, z, w; };

// four channels of a pixel / Jacobian slot: one STG.128 when the layout allows it (vec4: channel stride 1, 16-byte
// aligned), else four strided stores
DVR_DEV void store4(float* p, long long cs, int vec4, float a, float b, float c, float d) {
  if (vec4) {
    *reinterpret_cast<float4*>(p) = make_float4(a, b, c, d);
  } else {
    p[0] = a; p[cs] = b; p[2 * cs] = c; p[3 * cs] = d;
  }
}

// TF evaluation in dual numbers (renderer_tf.cuh:121-154, 255-288, 436-485, 521-560).
// DENS is float (no positional tangents) or Du<N>. s_dtf mirrors s_tf's layout with the
// tangent slot of every TF entry (or is unused when !HAS_TF).
template <int TF, int N, bool HAS_TF, typename DENS>
DVR_DEV Du4<N> tf_eval_dual(const float* s_tf, const int* s_dtf, int R, const DENS& density, int d0) {
  constexpr int ROW = TfLayout<TF>::kRow;
  // TF entries are dual numbers only when TF derivatives are traced; otherwise plain floats, so that the
  // products below are Du x float (N multiplies) instead of Du x Du with N zero tangents (2N FMAs that the
  // compiler may not fold: x * 0 is not 0 for NaN / inf)
  auto seed = [&](int r, int c) {
    if constexpr (HAS_TF) return du_input<N>(s_tf[r * ROW + c], s_dtf[r * ROW + c], d0);
    else return s_tf[r * ROW + c];
  };
  auto to_du = [&](auto v) {
    if constexpr (sizeof(v) == sizeof(float)) return du_const<N>(v);
    else return v;
  };
  auto dens = [&]() {
    if constexpr (sizeof(DENS) == sizeof(float)) return du_const<N>(density);
    else return density;
  };
  if constexpr (TF == kTfIdentity) {
    const Du<N> dc = du_clamp(dens(), du_const<N>(0.f), du_const<N>(1.f));
    const auto sa = seed(0, 0), sc = seed(0, 1);
    const Du<N> col = dc * sc;
    return {col, col, col, dc * sa};
  } else if constexpr (TF == kTfTexture) {
    const Du<N> d = dens() * float(R) - 0.5f;
    const int di = int(floorf(d.v));
    const Du<N> df = d - float(di);
    const int i0 = min(max(di, 0), R - 1), i1 = min(max(di + 1, 0), R - 1);
    return {du_lerp(seed(i0, 0), seed(i1, 0), df), du_lerp(seed(i0, 1), seed(i1, 1), df),
            du_lerp(seed(i0, 2), seed(i1, 2), df), du_lerp(seed(i0, 3), seed(i1, 3), df)};
  } else if constexpr (TF == kTfLinear) {
    const Du<N> dd = dens();
    const int i = linear_interval(s_tf, R, dd.v);
    const Du<N> p0 = to_du(seed(i, 4)), p1 = to_du(seed(i + 1, 4));
    const Du<N> d2 = du_clamp(dd, p0, p1);
    const Du<N> frac = (d2 - p0) / (p1 - p0);
    return {du_lerp(seed(i, 0), seed(i + 1, 0), frac), du_lerp(seed(i, 1), seed(i + 1, 1), frac),
            du_lerp(seed(i, 2), seed(i + 1, 2), frac), du_lerp(seed(i, 3), seed(i + 1, 3), frac)};
  } else {  // kTfGaussian
    const Du<N> dd = dens();
    Du4<N> c = {du_const<N>(0.f), du_const<N>(0.f), du_const<N>(0.f), du_const<N>(0.f)};
    for (int i = 0; i < R; ++i) {
      const auto mu = seed(i, 4), sigma = seed(i, 5);
      const Du<N> diff = dd - mu;
      const Du<N> ni = du_gauss_exp(-(diff * diff) / (2.0f * sigma * sigma));
      c.x = c.x + seed(i, 0) * ni;
      c.y = c.y + seed(i, 1) * ni;
      c.z = c.z + seed(i, 2) * ni;
      c.w = c.w + seed(i, 3) * ni;
    }
    return c;
  }
}

// Slot-major kernel for TF tangents (HAS_TF; instantiated for HAS_TF only). HAS_POS: stepsize and / or ray tangents
// are traced as well (position, t and stepsize are dual).
template <int FILTER, int TF, int BLEND, int N, bool HAS_POS, bool HAS_TF>
__global__ void __launch_bounds__(kThreads, N <= 6 ? 4 : 3) k_forward_gradients(const RayInputs in, const ForwardOutputs out,
                                                                const ForwardGradArgs ga) {
  extern __shared__ __align__(16) float smem[];
  constexpr int ROW = TfLayout<TF>::kRow;
  const int b = blockIdx.z;
  float* s_tf = smem;
  int* s_dtf = reinterpret_cast<int*>(smem + ((in.R * ROW + 3) & ~3));
  stage_tf<TF>(in, b, s_tf);
  if constexpr (HAS_TF) {
    const int* src = ga.dTf + b * ga.dtB;
    for (int i = threadIdx.x; i < in.R * ROW; i += blockDim.x) {
      const int r = i / ROW, c = i - r * ROW;
      s_dtf[i] = c < in.C ? __ldg(src + r * ga.dtR + c * ga.dtC) : -1;
    }
  }
  __syncthreads();
  int x, y;
  thread_pixel(x, y);
  if (x >= in.W || y >= in.H) return;

  const Ray ray = make_ray(in, x, y, b);  // box intersection on VALUES only (:201-208)
  const int d0 = ga.d0;
  // volume-space affine ray with tangents: ps = (start-boxMin)/voxel, pd = dir/voxel
  Du<N> psx, psy, psz, pdx, pdy, pdz, step;
  if constexpr (HAS_POS) {
    psx = du_input<N>(ray.ps.x, ga.dRayStart[0], d0) ; psy = du_input<N>(ray.ps.y, ga.dRayStart[1], d0); psz = du_input<N>(ray.ps.z, ga.dRayStart[2], d0);
    pdx = du_input<N>(ray.pd.x, ga.dRayDir[0], d0); pdy = du_input<N>(ray.pd.y, ga.dRayDir[1], d0); pdz = du_input<N>(ray.pd.z, ga.dRayDir[2], d0);
#pragma unroll
    for (int i = 0; i < N; ++i) {  // chain rule of the division by voxelSize
      psx.d[i] *= ray.invVoxel.x; psy.d[i] *= ray.invVoxel.y; psz.d[i] *= ray.invVoxel.z;
      pdx.d[i] *= ray.invVoxel.x; pdy.d[i] *= ray.invVoxel.y; pdz.d[i] *= ray.invVoxel.z;
    }
    step = du_input<N>(ray.step, ga.dStep, d0);
  }
  Sampler<FILTER, false> smp;
  smp.init(in, b, nullptr, 0);
  Du4<N> acc = {du_const<N>(0.f), du_const<N>(0.f), du_const<N>(0.f), du_const<N>(0.f)};
  int i = 0;
  float fi = 0.f;
  for (;; ++i, fi += 1.0f) {
    const float t = sample_t(ray, fi);
    if (t > ray.tmax) break;
    if (acc.w.v > in.earlyOut) break;
    const f3 p = sample_pos<HAS_POS>(ray, t);
    const float density = smp.fetch(in, nullptr, p.x, p.y, p.z);
    Du4<N> cur;
    if constexpr (HAS_POS) {
      Du<N> dd;
      if constexpr (FILTER == kFilterNearest) {
        dd = du_const<N>(density);  // no positional tangent (interpolation.cuh:52-65)
      } else {
        // t = tmin + i*step: dt/dstep = i; positions keep the value computed above
        Du<N> td = step * fi;
        td.v = t;
        Du<N> fxd = psx + td * pdx, fyd = psy + td * pdy, fzd = psz + td * pdz;
        fxd.v = smp.fx; fyd.v = smp.fy; fzd.v = smp.fz;  // fractional parts share the tangents
        dd = du_lerp(du_lerp(du_lerp(smp.d(0), smp.d(4), fxd), du_lerp(smp.d(2), smp.d(6), fxd), fyd),
                     du_lerp(du_lerp(smp.d(1), smp.d(5), fxd), du_lerp(smp.d(3), smp.d(7), fxd), fyd), fzd);
      }
      cur = tf_eval_dual<TF, N, HAS_TF>(s_tf, s_dtf, in.R, dd, d0);
    } else {
      cur = tf_eval_dual<TF, N, HAS_TF>(s_tf, s_dtf, in.R, density, d0);
    }
    // Blending::blend generic (renderer_blending.cuh:40-51, 105-116): alpha = 1 - exp(-w s) | min(1, w s),
    // acc += (1 - acc.w) * cur * alpha. The VALUES follow blend_alpha / blend_over of the plain forward kernel
    // operation for operation (both kernels write the same colour); the tangents are the product rule written out
    // once per slot with the shared factors hoisted (12 operations per slot instead of the ~26 of chained dual
    // products):  d alpha = e (w' s + w s'),  d(om alpha) = -acc.w' alpha + om d alpha,
    //             acc.c' += cur.c' om alpha + cur.c d(om alpha),  acc.w' += d(om alpha).
    {
      const float wv = cur.w.v;
      float sv;
      if constexpr (HAS_POS) sv = step.v;
      else sv = ray.step;
      float e = 0.f, av;
      bool active = true;  // Alpha blend: du_min(1, w s) keeps the tangent of w s unless 1 < w s
      if constexpr (BLEND == kBlendBeer) {
        e = blend_exp(-wv * sv);
        av = 1.f - e;
      } else {
        active = !(1.f < wv * sv);
        av = active ? wv * sv : 1.f;
      }
      const float om = 1.f - acc.w.v;
      const float oav = om * av;
#pragma unroll
      for (int k = 0; k < N; ++k) {
        float ws = cur.w.d[k] * sv;  // (w s)'
        if constexpr (HAS_POS) ws = fmaf(wv, step.d[k], ws);
        float ad;
        if constexpr (BLEND == kBlendBeer) ad = e * ws;
        else ad = active ? ws : 0.f;
        const float oad = fmaf(-acc.w.d[k], av, om * ad);
        acc.x.d[k] += fmaf(cur.x.d[k], oav, cur.x.v * oad);
        acc.y.d[k] += fmaf(cur.y.d[k], oav, cur.y.v * oad);
        acc.z.d[k] += fmaf(cur.z.d[k], oav, cur.z.v * oad);
        acc.w.d[k] += oad;
      }
      acc.x.v = fmaf(om * cur.x.v, av, acc.x.v);
      acc.y.v = fmaf(om * cur.y.v, av, acc.y.v);
      acc.z.v = fmaf(om * cur.z.v, av, acc.z.v);
      acc.w.v = fmaf(om, av, acc.w.v);
    }
  }
  if (ga.writeColor) {
    store4(out.color + b * out.cB + y * out.cY + x * out.cX, out.cC, out.colorVec4, acc.x.v, acc.y.v, acc.z.v, acc.w.v);
    out.term[b * out.tB + y * out.tY + x * out.tX] = i;
  }
  float* g = ga.grad + b * ga.gB + y * ga.gY + x * ga.gX;
#pragma unroll
  for (int k = 0; k < N; ++k) {
    const int d = d0 + k;
    if (d < ga.D) store4(g + d * ga.gD, ga.gC, ga.gradVec4, acc.x.d[k], acc.y.d[k], acc.z.d[k], acc.w.d[k]);
  }
}

// Forward-mode when ONLY the ray (start, direction) and / or the stepsize are differentiated -- the camera
// optimisation of the reference (C3; pytests/diffdvr/renderer.py:67-69). The D tangent slots are then a linear map
// of at most seven INPUT tangents (start xyz, direction xyz, stepsize), each of which has a closed form per sample:
//   d pos / d start_a = 1/voxel_a (axis a only),  d pos / d dir_a = t / voxel_a,  d pos / d step = i * pd,
// so the kernel carries the accumulator's tangents per input instead of per slot: no position duals in registers
// (42 fewer than the slot-major kernel), density tangents = cell gradient x those closed forms (10 operations per
// sample for all seven), colour tangents = dTF/d density x density tangent (rank one), blend tangents written out per
// input with the shared factors hoisted. ONE marching pass for any D; the slots are assembled at the end (several
// inputs may share a slot: their tangents add, as unit seeds in the same slot would).
#ifndef DVR_FWD_RANK1
#define DVR_FWD_RANK1 1
#endif
template <int FILTER, int TF, int BLEND>
__global__ void __launch_bounds__(kThreads, 4) k_forward_gradients_inputs(const RayInputs in, const ForwardOutputs out,
                                                                          const ForwardGradArgs ga) {
  extern __shared__ __align__(16) float smem[];
  constexpr int NI = 7;
  const int b = blockIdx.z;
  float* s_tf = smem;
  stage_tf<TF>(in, b, s_tf);
  __syncthreads();
  int x, y;
  thread_pixel(x, y);
  if (x >= in.W || y >= in.H) return;

  const Ray ray = make_ray(in, x, y, b);  // box intersection on VALUES only (:201-208)
  // d acc / d input, the seven inputs as four register pairs (0,1) (2,3) (4,5) (6,-): the per-input tangent update is
  // the same arithmetic for every input, so it runs on FFMA2 / FMUL2 / FADD2 (half the issue slots; per-lane results
  // identical to the scalar form)
  constexpr int NP = 4;
  float2 ax[NP], ay[NP], az[NP], aw[NP];
#pragma unroll
  for (int j = 0; j < NP; ++j) ax[j] = ay[j] = az[j] = aw[j] = make_float2(0.f, 0.f);
  f4 acc = {0.f, 0.f, 0.f, 0.f};
  Sampler<FILTER, false> smp;
  smp.init(in, b, nullptr, 0);
  int i = 0;
  float fi = 0.f;
  constexpr bool LOOKAHEAD = FILTER == kFilterTrilinearQuad && DVR_LOOKAHEAD_RAY;
  // look-ahead state (Sampler::lookahead): parameter and position of the NEXT sample, its cell already requested
  [[maybe_unused]] float tNext = sample_t(ray, 0.f);
  if constexpr (LOOKAHEAD) {
    if (!(tNext > ray.tmax)) {
      const f3 pn = sample_pos<true>(ray, tNext);
      smp.lookahead_f(in, pn.x, pn.y, pn.z);
    }
  }
  for (;; ++i, fi += 1.0f) {
    float t;
    [[maybe_unused]] f3 p;
    if constexpr (LOOKAHEAD) t = tNext;
    else t = sample_t(ray, fi);
    if (t > ray.tmax) break;
    if (acc.w > in.earlyOut) break;
    float density;
    [[maybe_unused]] f3 gCell = {0.f, 0.f, 0.f};
    if constexpr (LOOKAHEAD) {
#if DVR_GRAD_SHARED
      density = smp.interp_grad(gCell);  // value and cell gradient from one lerp chain, before the look-ahead overwrites either
#else
      density = smp.interp_f();
      gCell = smp.unit_gradient();   // from the same corner values and fractions, before either is overwritten
#endif
      tNext = sample_t(ray, fi + 1.0f);
      if (!(tNext > ray.tmax)) {
        const f3 pn = sample_pos<true>(ray, tNext);
        smp.lookahead_f(in, pn.x, pn.y, pn.z);
      }
    } else {
      p = sample_pos<true>(ray, t);
      density = smp.fetch(in, nullptr, p.x, p.y, p.z);
    }
    if constexpr (DVR_PREFETCH_RAY && !LOOKAHEAD) {
      const f3 pn = sample_pos<false>(ray, sample_t(ray, fi + 1.0f));
      smp.prefetch_cell(in, pn.x, pn.y, pn.z);
    }
    // d density / d input
    float2 dd[NP];
    if constexpr (FILTER == kFilterNearest) {
#pragma unroll
      for (int j = 0; j < NP; ++j) dd[j] = make_float2(0.f, 0.f);  // no positional tangent (interpolation.cuh:52-65)
    } else {
      f3 g;  // gradient of the trilinear cell at the sample
      if constexpr (LOOKAHEAD) g = gCell;
      else g = smp.template adjoint<true>(1.0f);
      const float d0 = g.x * ray.invVoxel.x, d1 = g.y * ray.invVoxel.y, d2 = g.z * ray.invVoxel.z;
      dd[0] = make_float2(d0, d1);
      dd[1] = make_float2(d2, t * d0);
      dd[2] = pk_mul(t, make_float2(d1, d2));
      dd[3] = make_float2(fi * fmaf(g.z, ray.pd.z, fmaf(g.y, ray.pd.y, g.x * ray.pd.x)), 0.f);
    }
    // colour cv and its slope cs = d colour / d density
    f4 cv, cs;
    if constexpr (TF == kTfGaussian) {
      // the plain kernel's evaluation (tf_eval: staged reciprocals, one expf per row) plus, per row,
      // d/dd exp(-(d - mu)^2 / 2 sigma^2) = n (mu - d) / sigma^2 as in tf_adjoint
      cv = {0.f, 0.f, 0.f, 0.f};
      cs = {0.f, 0.f, 0.f, 0.f};
      for (int r = 0; r < in.R; ++r) {
        const float* row = s_tf + 8 * r;
        const f4 ci = lds4(row);
        const float n = gauss_exp(gauss_arg<false>(density, row));
        const float k = n * gauss_slope(density, row);
        cv.x = fmaf(ci.x, n, cv.x); cv.y = fmaf(ci.y, n, cv.y); cv.z = fmaf(ci.z, n, cv.z); cv.w = fmaf(ci.w, n, cv.w);
        cs.x = fmaf(ci.x, k, cs.x); cs.y = fmaf(ci.y, k, cs.y); cs.z = fmaf(ci.z, k, cs.z); cs.w = fmaf(ci.w, k, cs.w);
      }
    } else {  // the dual TF code with one tangent
      Du<1> d1;
      d1.v = density;
      d1.d[0] = 1.0f;
      const Du4<1> c1 = tf_eval_dual<TF, 1, false>(s_tf, nullptr, in.R, d1, 0);
      cv = {c1.x.v, c1.y.v, c1.z.v, c1.w.v};
      cs = {c1.x.d[0], c1.y.d[0], c1.z.d[0], c1.w.d[0]};
    }
    // blend: values exactly as blend_alpha / blend_over of the plain forward kernel
    const float wv = cv.w, sv = ray.step;
    float e = 0.f, av;
    bool active = true;
    if constexpr (BLEND == kBlendBeer) {
      e = blend_exp(-wv * sv);
      av = 1.f - e;
    } else {
      active = !(1.f < wv * sv);
      av = active ? wv * sv : 1.f;
    }
    const float om = 1.f - acc.w;
    const float oav = om * av;
#if DVR_FWD_RANK1
    // Every input's colour tangent is (d colour / d density) x (that input's density tangent), so the product rule of
    // the blend collapses to TWO sample-wide vectors: with E = d alpha / d (w s), B = om E cs.w s and the tangent of
    // om alpha being -alpha aw_k + B dd_k (+ om E w for the stepsize input, whose s' = 1),
    //   a.c_k += Q_c dd_k - P_c aw_k,  Q_c = cs.c om alpha + cv.c B,  P_c = cv.c alpha;   aw_k = aw_k (1 - alpha) + B dd_k
    // -- 16 FMA-pipe operations per input pair instead of 28 (same sums, re-associated).
    float E;
    if constexpr (BLEND == kBlendBeer) E = e;
    else E = active ? 1.f : 0.f;
    const float omE = om * E;
    const float Bq = omE * (cs.w * sv), Cw = omE * wv, keep = 1.f - av;
    const float Qx = fmaf(cv.x, Bq, cs.x * oav), Qy = fmaf(cv.y, Bq, cs.y * oav), Qz = fmaf(cv.z, Bq, cs.z * oav);
    const float Px = -(cv.x * av), Py = -(cv.y * av), Pz = -(cv.z * av);
#pragma unroll
    for (int j = 0; j < NP; ++j) {
      ax[j] = pk_fma(Px, aw[j], pk_fma(Qx, dd[j], ax[j]));
      ay[j] = pk_fma(Py, aw[j], pk_fma(Qy, dd[j], ay[j]));
      az[j] = pk_fma(Pz, aw[j], pk_fma(Qz, dd[j], az[j]));
      aw[j] = pk_fma(Bq, dd[j], pk_mul(keep, aw[j]));
    }
    // the stepsize input (pair 3, first half): alpha depends on it directly as well
    ax[3].x = fmaf(cv.x, Cw, ax[3].x);
    ay[3].x = fmaf(cv.y, Cw, ay[3].x);
    az[3].x = fmaf(cv.z, Cw, az[3].x);
    aw[3].x += Cw;
#else
    const float kx = cs.x * oav, ky = cs.y * oav, kz = cs.z * oav, kw = cs.w * sv;
#pragma unroll
    for (int j = 0; j < NP; ++j) {
      float2 ws = pk_mul(kw, dd[j]);     // (w s)' = w' s + w s', s' = 1 for the stepsize input (pair 3, first half) only
      if (j == 3) ws.x += wv;
      float2 ad;
      if constexpr (BLEND == kBlendBeer) ad = pk_mul(e, ws);
      else ad = active ? ws : make_float2(0.f, 0.f);
      const float2 oad = pk_fma(-av, aw[j], pk_mul(om, ad));  // d((1 - acc.w) alpha) = -acc.w' alpha + om alpha'
      ax[j] = pk_add(ax[j], pk_fma(kx, dd[j], pk_mul(cv.x, oad)));
      ay[j] = pk_add(ay[j], pk_fma(ky, dd[j], pk_mul(cv.y, oad)));
      az[j] = pk_add(az[j], pk_fma(kz, dd[j], pk_mul(cv.z, oad)));
      aw[j] = pk_add(aw[j], oad);
    }
#endif
    acc.x = fmaf(om * cv.x, av, acc.x);
    acc.y = fmaf(om * cv.y, av, acc.y);
    acc.z = fmaf(om * cv.z, av, acc.z);
    acc.w = fmaf(om, av, acc.w);
  }
  store4(out.color + b * out.cB + y * out.cY + x * out.cX, out.cC, out.colorVec4, acc.x, acc.y, acc.z, acc.w);
  out.term[b * out.tB + y * out.tY + x * out.tX] = i;
  // slots <- inputs
  float* g = ga.grad + b * ga.gB + y * ga.gY + x * ga.gX;
  for (int d = 0; d < ga.D; ++d) {
    float gx = 0.f, gy = 0.f, gz = 0.f, gw = 0.f;
#pragma unroll
    for (int j = 0; j < NI; ++j) {
      const int slot = j < 3 ? ga.dRayStart[j] : (j < 6 ? ga.dRayDir[j - 3] : ga.dStep);
      if (slot == d) {
        gx += (j & 1) ? ax[j >> 1].y : ax[j >> 1].x; gy += (j & 1) ? ay[j >> 1].y : ay[j >> 1].x;
        gz += (j & 1) ? az[j >> 1].y : az[j >> 1].x; gw += (j & 1) ? aw[j >> 1].y : aw[j >> 1].x;
      }
    }
    store4(g + d * ga.gD, ga.gC, ga.gradVec4, gx, gy, gz, gw);
  }
}

}  // namespace dvr
