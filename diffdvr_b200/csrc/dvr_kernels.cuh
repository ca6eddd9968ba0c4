// sm_100a kernels: plain forward render and adjoint pass.
// One thread marches one ray; a warp owns an 8x4 pixel tile so that its 32 rays stay in
// neighbouring cells (coherent L1 sectors, coherent trip counts); a CTA owns 16x8 pixels of one
// view (blockIdx.z = batch entry), stages that view's transfer function in shared memory once
// and, in the adjoint, owns a shared-memory accumulator for the TF gradient that is flushed to
// global memory once per CTA.
#pragma once
#include "dvr_device.cuh"

namespace dvr {

// ===============================================================================================
// forward: DvrKernelForwardImpl (renderer/renderer_kernels.cuh:22-90)
// ===============================================================================================
template <int FILTER, int TF, int BLEND>
__global__ void __launch_bounds__(kThreads) k_forward(const RayInputs in, const ForwardOutputs out) {
  extern __shared__ __align__(16) float smem[];
  const int b = blockIdx.z;
  stage_tf<TF>(in, b, smem);
  __syncthreads();
  int x, y;
  thread_pixel(x, y);
  if (x >= in.W || y >= in.H) return;

  const Ray ray = make_ray(in, x, y, b);
  Sampler<FILTER, false> smp;
  smp.init(in, b, nullptr, 0);
  f4 acc = {0.f, 0.f, 0.f, 0.f};
  float fi = 0.f;  // sample counter; exact in fp32 (terminationIndex < 2^24)
  if constexpr (FILTER == kFilterTrilinearQuad && DVR_LOOKAHEAD) {
    // look-ahead loop (Sampler::lookahead): same samples, same arithmetic, the next cell's loads issued one sample early
    [[maybe_unused]] const float sa = smem[0], sc = smem[1];
    // With predicated loads the look-ahead is unconditional: past the last sample it requests a clamped, valid quad
    // (Sampler::quad_index) that nobody interpolates, and the loop body has no reconvergence point.
    constexpr bool ALWAYS = DVR_PRED_LOADS && DVR_BRANCHFREE_CELL && (TF == kTfIdentity || TF == kTfTexture);
    float t = sample_t(ray, 0.f);
    if (ALWAYS || !(t > ray.tmax)) {
      const f3 p = sample_pos<false>(ray, t);
      smp.template lookahead_f<ALWAYS>(in, p.x, p.y, p.z);
    }
    for (;;) {
      if (t > ray.tmax) break;
      if (acc.w > in.earlyOut) break;
      const float density = smp.interp_f();
      fi += 1.0f;
      t = sample_t(ray, fi);
      if (ALWAYS || !(t > ray.tmax)) {
        const f3 p = sample_pos<false>(ray, t);
        smp.template lookahead_f<ALWAYS>(in, p.x, p.y, p.z);
      }
      if constexpr (TF == kTfIdentity) {
        const float dc = fmaxf(0.f, fminf(density, 1.f));
        const float alpha = blend_alpha<BLEND, ALWAYS>(dc * sa, ray.step);
        const float om = 1.f - acc.w;
        acc.x = fmaf(om * (dc * sc), alpha, acc.x);
        acc.w = fmaf(om, alpha, acc.w);
      } else {
        const f4 cur = tf_eval<TF>(smem, in.R, density);
        acc = blend_over(acc, cur, blend_alpha<BLEND, ALWAYS>(cur.w, ray.step));
      }
    }
    if constexpr (TF == kTfIdentity) acc.y = acc.z = acc.x;
  } else if constexpr (TF == kTfIdentity) {
    // r = g = b for this TF (renderer_tf.cuh:62-73): march one colour channel; the table's two
    // scales stay in registers
    const float sa = smem[0], sc = smem[1];
    for (;; fi += 1.0f) {
      const float t = sample_t(ray, fi);
      if (t > ray.tmax) break;
      if (acc.w > in.earlyOut) break;
      const f3 p = sample_pos<false>(ray, t);
      const float dc = fmaxf(0.f, fminf(smp.fetch(in, nullptr, p.x, p.y, p.z), 1.f));
      if constexpr (DVR_PREFETCH_FWD) {
        const f3 pn = sample_pos<false>(ray, sample_t(ray, fi + 1.0f));
        smp.prefetch_cell(in, pn.x, pn.y, pn.z);
      }
      const float alpha = blend_alpha<BLEND>(dc * sa, ray.step);
      const float om = 1.f - acc.w;
      acc.x = fmaf(om * (dc * sc), alpha, acc.x);
      acc.w = fmaf(om, alpha, acc.w);
    }
    acc.y = acc.z = acc.x;
  } else {
    for (;; fi += 1.0f) {
      const float t = sample_t(ray, fi);
      if (t > ray.tmax) break;
      if (acc.w > in.earlyOut) break;
      const f3 p = sample_pos<false>(ray, t);
      const float density = smp.fetch(in, nullptr, p.x, p.y, p.z);
      if constexpr (DVR_PREFETCH_FWD) {
        const f3 pn = sample_pos<false>(ray, sample_t(ray, fi + 1.0f));
        smp.prefetch_cell(in, pn.x, pn.y, pn.z);
      }
      const f4 cur = tf_eval<TF>(smem, in.R, density);
      acc = blend_over(acc, cur, blend_alpha<BLEND>(cur.w, ray.step));
    }
  }
  const int i = int(fi);
  float* c = out.color + b * out.cB + y * out.cY + x * out.cX;
  if (out.colorVec4) {
    *reinterpret_cast<float4*>(c) = make_float4(acc.x, acc.y, acc.z, acc.w);
  } else {
    c[0] = acc.x; c[out.cC] = acc.y; c[2 * out.cC] = acc.z; c[3 * out.cC] = acc.w;
  }
  out.term[b * out.tB + y * out.tY + x * out.tX] = i;
}

// ===============================================================================================
// adjoint: DvrKernelAdjointImpl (renderer/renderer_kernels.cuh:315-444)
// ===============================================================================================
// TF-gradient accumulation strategies (replace both the reference's immediate mode -- 8..12
// global atomics per SAMPLE, renderer_tf.cuh:208-216 -- and its delayed mode -- R*C floats of
// shared memory per THREAD, renderer_cuda.cpp:273-295):
//   kTfAccReg : R <= kRegRows: per-thread register accumulators, one warp-shuffle reduction and
//               one shared-memory atomic per table entry per WARP at the end of the ray
//   kTfAccSmem: larger tables: per-warp private table in shared memory; Texture/Linear keep the
//               two rows they are currently touching in registers and flush on a row change
// Either way global memory sees R*C atomics per CTA.
enum : int { kTfAccNone = 0, kTfAccReg = 1, kTfAccSmem = 2 };
constexpr int kRegRows = 6;

template <int TF> struct TfChannels { static constexpr int v = 6; };
template <> struct TfChannels<kTfIdentity> { static constexpr int v = 2; };
template <> struct TfChannels<kTfTexture> { static constexpr int v = 4; };
template <> struct TfChannels<kTfLinear> { static constexpr int v = 5; };

// FACTORED (Texture table, kTfAccSmem, Beer-Lambert on the transmittance; texture_adjoint_factored below): the three
// colour adjoints of a ray never change along it, so the colour channels of a row's gradient are adjColour.c x (sum of
// kc w over the samples): the window keeps ONE colour sum and the alpha sum per row (kq[k] = (K, A)) and multiplies by
// the ray's colour adjoints when a row leaves the window.
template <int TF, int TFACC, bool FACTORED = false>
struct TfGrad {
  static constexpr int CH = TfChannels<TF>::v;
  static constexpr int NREG = TFACC == kTfAccReg ? kRegRows * CH : 1;
  static constexpr int kNoBase = -0x40000000;
  float g[NREG];     // kTfAccReg: whole table; indexing is fully unrolled (stays in registers)
  float row[2][CH];  // kTfAccSmem (Texture/Linear): the two rows the ray is currently between
  // kTfAccSmem window key. Texture: floor of the texture coordinate, -1 .. R (the rows are min(max(base, 0), R-1) and
  // min(max(base + 1, 0), R-1): below the first / above the last texel centre both are the same clamped row);
  // Linear: the interval index (rows base, base + 1)
  int base;
  int rows;          // R
  float* table;      // kTfAccSmem: this warp's table
  float2 kq[2];      // FACTORED: (sum kc w, sum adjW w) of the window's two rows
  float cx, cy, cz;  // FACTORED: the ray's colour adjoints

  DVR_DEV void init(float* warpTable, int R) {
#pragma unroll
    for (int i = 0; i < NREG; ++i) g[i] = 0.f;
#pragma unroll
    for (int c = 0; c < CH; ++c) row[0][c] = row[1][c] = 0.f;
    base = kNoBase;
    rows = R;
    table = warpTable;
    kq[0] = kq[1] = make_float2(0.f, 0.f);
    cx = cy = cz = 0.f;
  }
  DVR_DEV void set_colour_adjoint(float x, float y, float z) { cx = x; cy = y; cz = z; }
  DVR_DEV void flush_rows() {
    if constexpr (FACTORED) {
      if (base != kNoBase) {
        flush_one(0);
        flush_one(1);
        kq[0] = kq[1] = make_float2(0.f, 0.f);
      }
    } else if (base != kNoBase) {
      int r0 = base, r1 = base + 1;
      if constexpr (TF == kTfTexture) {
        r0 = min(max(r0, 0), rows - 1);
        r1 = min(max(r1, 0), rows - 1);
      }
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        if (row[0][c] != 0.f) atomicAdd(table + r0 * CH + c, row[0][c]);
        if (row[1][c] != 0.f) atomicAdd(table + r1 * CH + c, row[1][c]);
        row[0][c] = row[1][c] = 0.f;
      }
    }
  }
  // kTfAccSmem, Texture / Linear: the sample lies between rows (key, key + 1) with weights (1 - w, w); v = adjoint of
  // the interpolated value. One compare per sample, the products fused into the accumulation (the reference: 8
  // global atomics per sample, renderer_tf.cuh:208-216).
  // one row of the window into the shared table (k = 0: row `base`, k = 1: row `base + 1`; Texture rows clamp)
  DVR_DEV void flush_one(int k) {
    int r = base + k;
    if constexpr (TF == kTfTexture) r = min(max(r, 0), rows - 1);
    if constexpr (FACTORED) {
      const float2 v = k ? kq[1] : kq[0];
      if (v.x != 0.f) {
        atomicAdd(table + r * CH, cx * v.x);
        atomicAdd(table + r * CH + 1, cy * v.x);
        atomicAdd(table + r * CH + 2, cz * v.x);
      }
      if (v.y != 0.f) atomicAdd(table + r * CH + 3, v.y);
    } else {
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        const float v = k ? row[1][c] : row[0][c];
        if (v != 0.f) atomicAdd(table + r * CH + c, v);
      }
    }
  }
  DVR_DEV void add_pair(int key, const float* v, float w) {
    if (key != base) {
      // the density usually drifts into the NEIGHBOURING row pair: one row of the window stays, only the other goes out
      if (key == base + 1) {
        flush_one(0);
#pragma unroll
        for (int c = 0; c < CH; ++c) { row[0][c] = row[1][c]; row[1][c] = 0.f; }
      } else if (key == base - 1) {
        flush_one(1);
#pragma unroll
        for (int c = 0; c < CH; ++c) { row[1][c] = row[0][c]; row[0][c] = 0.f; }
      } else {
        flush_rows();
      }
      base = key;
    }
    const float w0 = 1.f - w;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      row[0][c] = fmaf(v[c], w0, row[0][c]);
      row[1][c] = fmaf(v[c], w, row[1][c]);
    }
  }
  // FACTORED: the sample lies between rows (key, key + 1) with weights (1 - w, w); kc = the factor of the colour adjoints
  // in the adjoint of the interpolated colour, adjW = the adjoint of the interpolated absorption
  DVR_DEV void add_pair_k(int key, float kc, float adjW, float w) {
    if (key != base) {
      if (key == base + 1) {
        flush_one(0);
        kq[0] = kq[1]; kq[1] = make_float2(0.f, 0.f);
      } else if (key == base - 1) {
        flush_one(1);
        kq[1] = kq[0]; kq[0] = make_float2(0.f, 0.f);
      } else {
        flush_rows();
      }
      base = key;
    }
    const float2 ka = make_float2(kc, adjW);
    kq[0] = pk_fma(1.f - w, ka, kq[0]);
    kq[1] = pk_fma(w, ka, kq[1]);
  }
  // contribution v[0..CH) to row r (register accumulators, or straight into the shared table)
  DVR_DEV void add_row(int r, const float* v) {
    if constexpr (TFACC == kTfAccReg) {
#pragma unroll
      for (int rr = 0; rr < kRegRows; ++rr)
        if (rr == r) {
#pragma unroll
          for (int c = 0; c < CH; ++c) g[rr * CH + c] += v[c];
        }
    } else {
#pragma unroll
      for (int c = 0; c < CH; ++c) atomicAdd(table + r * CH + c, v[c]);
    }
  }
};

// What the evaluation of a sample leaves behind for its adjoint (evaluated back to back in the adjoint kernel):
//   Gaussian: exp(-(d-mu)^2 / 2 sigma^2) of the first kRegRows rows; Texture / Linear: the row pair, the interpolation
//   weight and v1 - v0 (the table is not read twice and the coordinates are not computed twice).
template <int TF> struct TfCache { float ni[kRegRows]; };
template <> struct TfCache<kTfTexture> { int fl; float df; f4 dv; };
template <> struct TfCache<kTfLinear> { int i; float frac, d2, p0, p1, rcp; f4 dv; };

template <int TF>
DVR_DEV f4 tf_eval_cached(const float* s_tf, int R, float density, TfCache<TF>& cache) {
  if constexpr (TF == kTfGaussian) {
    float (&ni)[kRegRows] = cache.ni;
    f4 c = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < kRegRows; ++i) {
      ni[i] = 0.f;
      if (i < R) {
        const f4 ci = lds4(s_tf + 8 * i);
        ni[i] = gauss_exp(gauss_arg<false>(density, s_tf + 8 * i));
        c.x = fmaf(ci.x, ni[i], c.x); c.y = fmaf(ci.y, ni[i], c.y); c.z = fmaf(ci.z, ni[i], c.z); c.w = fmaf(ci.w, ni[i], c.w);
      }
    }
    for (int i = kRegRows; i < R; ++i) {
      const f4 ci = lds4(s_tf + 8 * i);
      const float n = gauss_exp(gauss_arg<false>(density, s_tf + 8 * i));
      c.x = fmaf(ci.x, n, c.x); c.y = fmaf(ci.y, n, c.y); c.z = fmaf(ci.z, n, c.z); c.w = fmaf(ci.w, n, c.w);
    }
    return c;
  } else if constexpr (TF == kTfTexture) {  // tf_eval, keeping the coordinates and v1 - v0
    int i0, i1;
    texture_coords(density, R, i0, i1, cache.df, cache.fl);
    const f4 v0 = lds4(s_tf + 4 * i0), v1 = lds4(s_tf + 4 * i1);
    const float2 dlo = pk_sub(make_float2(v1.x, v1.y), make_float2(v0.x, v0.y));
    const float2 dhi = pk_sub(make_float2(v1.z, v1.w), make_float2(v0.z, v0.w));
    cache.dv = {dlo.x, dlo.y, dhi.x, dhi.y};
    const float2 lo = pk_fma(cache.df, dlo, make_float2(v0.x, v0.y)), hi = pk_fma(cache.df, dhi, make_float2(v0.z, v0.w));
    return {lo.x, lo.y, hi.x, hi.y};
  } else if constexpr (TF == kTfLinear) {
    const int i = linear_interval(s_tf, R, density);
    const f4 v0 = lds4(s_tf + 8 * i), v1 = lds4(s_tf + 8 * (i + 1));
    cache.i = i;
    cache.p0 = s_tf[8 * i + 4]; cache.p1 = s_tf[8 * (i + 1) + 4]; cache.rcp = s_tf[8 * i + 5];
    cache.d2 = fmaxf(cache.p0, fminf(density, cache.p1));
    cache.frac = div_rcp(cache.d2 - cache.p0, cache.p1 - cache.p0, cache.rcp);
    cache.dv = {v1.x - v0.x, v1.y - v0.y, v1.z - v0.z, v1.w - v0.w};
    const float f = cache.frac;
    return {fmaf(f, cache.dv.x, v0.x), fmaf(f, cache.dv.y, v0.y), fmaf(f, cache.dv.z, v0.z), fmaf(f, cache.dv.w, v0.w)};
  } else {
    return tf_eval<TF>(s_tf, R, density);
  }
}

// adjoint of tf_eval; returns adj_density and adds the parameter adjoints to `acc`
template <int TF, int TFACC>
DVR_DEV float tf_adjoint(const float* s_tf, int R, float density, const f4& adjC, TfGrad<TF, TFACC>& acc,
                         const TfCache<TF>& cache) {
  constexpr bool HAS_TF = TFACC != kTfAccNone;
  if constexpr (TF == kTfIdentity) {  // renderer_tf.cuh:75-104
    const float dc = fmaxf(0.f, fminf(density, 1.f));
    const float sa = s_tf[0], sc = s_tf[1];
    const float adjDc = (adjC.x * sc + adjC.y * sc + adjC.z * sc) + adjC.w * sa;
    if constexpr (TFACC == kTfAccReg) {
      acc.g[0] += adjC.w * dc;                               // d/d scaleAbsorption
      acc.g[1] += adjC.x * dc + adjC.y * dc + adjC.z * dc;   // d/d scaleColor
    } else if constexpr (HAS_TF) {  // never chosen by the host dispatch (two scalars fit registers); kept correct
      const float v[2] = {adjC.w * dc, adjC.x * dc + adjC.y * dc + adjC.z * dc};
      acc.add_row(0, v);
    }
    return (density > 0.f && density < 1.f) ? adjDc : 0.f;
  } else if constexpr (TF == kTfTexture) {  // :173-220
    const int fl = cache.fl;
    const float df = cache.df;
    const float adjDf = dot4(adjC, cache.dv);
    if constexpr (TFACC == kTfAccSmem) {
      const float a[4] = {adjC.x, adjC.y, adjC.z, adjC.w};
      acc.add_pair(fl, a, df);
    } else if constexpr (HAS_TF) {
      const int i0 = min(max(fl, 0), R - 1), i1 = min(max(fl + 1, 0), R - 1);
      const float w0 = 1.f - df;
      const float a0[4] = {adjC.x * w0, adjC.y * w0, adjC.z * w0, adjC.w * w0};
      const float a1[4] = {adjC.x * df, adjC.y * df, adjC.z * df, adjC.w * df};
      acc.add_row(i0, a0);
      acc.add_row(i1, a1);
    }
    return adjDf * float(R);
  } else if constexpr (TF == kTfLinear) {  // :315-400
    const int i = cache.i;
    const float p0 = cache.p0, p1 = cache.p1, d2 = cache.d2;
    const float rcp = cache.rcp;                                 // 1 / (p1 - p0), staged per row
    const float frac = cache.frac;                               // the IEEE quotient (d2 - p0) / (p1 - p0)
    const float adjFrac = dot4(adjC, cache.dv);
    const float adjD = div_rcp(adjFrac, p1 - p0, rcp);           // adjFrac / (p1 - p0)
    if constexpr (HAS_TF) {
      // adjFrac (d2 - p1) / (p1 - p0)^2 and adjFrac (p0 - d2) / (p1 - p0)^2 (:377-384) as two exact quotients by
      // (p1 - p0): the value of the reference's single division up to one rounding
      const float g0 = div_rcp(adjD * (d2 - p1), p1 - p0, rcp), g1 = div_rcp(adjD * (p0 - d2), p1 - p0, rcp);
      if constexpr (TFACC == kTfAccSmem) {
        const float a[4] = {adjC.x, adjC.y, adjC.z, adjC.w};
        acc.add_pair(i, a, frac);
        acc.row[0][4] += g0;
        acc.row[1][4] += g1;
      } else {
        const float w0 = 1.f - frac;
        const float a0[5] = {adjC.x * w0, adjC.y * w0, adjC.z * w0, adjC.w * w0, g0};
        const float a1[5] = {adjC.x * frac, adjC.y * frac, adjC.z * frac, adjC.w * frac, g1};
        acc.add_row(i, a0);
        acc.add_row(i + 1, a1);
      }
    }
    return adjD;  // passes through the clamp unconditionally ("#if 1", :350-352)
  } else if constexpr (TF == kTfGaussian) {  // :562-614
    float adjDensity = 0.f;
    auto body = [&](int i, float* g, float ni) {
      const f4 ci = lds4(s_tf + 8 * i);
      const float mu = s_tf[8 * i + 4], sigma = s_tf[8 * i + 5];
      const float f1 = gauss_slope(density, s_tf + 8 * i) * ni;
      const float dc = dot4(adjC, ci);
      adjDensity = fmaf(dc, f1, adjDensity);
      if constexpr (HAS_TF) {
#if DVR_FAST_GAUSS_ARG
        const float f2 = (f1 * (mu - density)) * (sigma * s_tf[8 * i + 7]);  // 1 / sigma = sigma / sigma^2
#else
        const float f2 = f1 * (mu - density) / sigma;
#endif
        g[0] = adjC.x * ni; g[1] = adjC.y * ni; g[2] = adjC.z * ni; g[3] = adjC.w * ni;
        g[4] = -dc * f1;
        g[5] = dc * f2;
      }
    };
    if constexpr (TFACC == kTfAccReg) {
#pragma unroll
      for (int i = 0; i < kRegRows; ++i)
        if (i < R) {
          float g[6];
          body(i, g, cache.ni[i]);
#pragma unroll
          for (int c = 0; c < 6; ++c) acc.g[i * 6 + c] += g[c];
        }
    } else {
#pragma unroll
      for (int i = 0; i < kRegRows; ++i)
        if (i < R) {
          float g[6];
          body(i, g, cache.ni[i]);
          if constexpr (HAS_TF) acc.add_row(i, g);
        }
      for (int i = kRegRows; i < R; ++i) {
        float g[6];
        body(i, g, gauss_exp(gauss_arg<false>(density, s_tf + 8 * i)));
        if constexpr (HAS_TF) acc.add_row(i, g);
      }
    }
    return adjDensity;
  } else {  // kTfGaussianLog, :705-780
    float adjDensity = 0.f;
    auto tval = [&](int i) {
      const f4 ci = lds4(s_tf + 8 * i);
      const float ln = gauss_arg(density, s_tf + 8 * i);
      return f4{logf(ci.x) + ln, logf(ci.y) + ln, logf(ci.z) + ln, logf(ci.w) + ln};
    };
    auto body = [&](int i, float* g) {
      const f4 ci = lds4(s_tf + 8 * i);
      const float mu = s_tf[8 * i + 4], sigma = s_tf[8 * i + 5];
      const f4 ti = tval(i);
      f4 den = {1.f, 1.f, 1.f, 1.f};
      for (int j = 0; j < R; ++j)
        if (j != i) {
          const f4 tj = tval(j);
          den.x += expf(tj.x - ti.x); den.y += expf(tj.y - ti.y); den.z += expf(tj.z - ti.z); den.w += expf(tj.w - ti.w);
        }
      const f4 adjTi = {adjC.x / den.x, adjC.y / den.y, adjC.z / den.z, adjC.w / den.w};
      const float f1 = div_rcp(mu - density, sigma * sigma, s_tf[8 * i + 7]);
      const float adjXmu = adjTi.x * f1 + adjTi.y * f1 + adjTi.z * f1 + adjTi.w * f1;
      adjDensity += adjXmu;
      if constexpr (HAS_TF) {
        const float f2 = f1 * (mu - density) / sigma;
        g[0] = adjTi.x / ci.x; g[1] = adjTi.y / ci.y; g[2] = adjTi.z / ci.z; g[3] = adjTi.w / ci.w;
        g[4] = -adjXmu;
        g[5] = adjTi.x * f2 + adjTi.y * f2 + adjTi.z * f2 + adjTi.w * f2;
      }
    };
    if constexpr (TFACC == kTfAccReg) {
#pragma unroll
      for (int i = 0; i < kRegRows; ++i)
        if (i < R) {
          float g[6];
          body(i, g);
#pragma unroll
          for (int c = 0; c < 6; ++c) acc.g[i * 6 + c] += g[c];
        }
    } else {
      for (int i = 0; i < R; ++i) {
        float g[6];
        body(i, g);
        if constexpr (HAS_TF) acc.add_row(i, g);
      }
    }
    return adjDensity;
  }
}

// Texture TF + Beer-Lambert on the transmittance, evaluation / blend adjoint / TF adjoint of one sample in one piece
// (DVR_TF_FACTORED). The adjoint never needs the interpolated COLOUR, only S = adjColour . colour, and the colour
// adjoints are constants of the ray: with q_i = adjColour . row_i (3 FMAs per row), S = q0 + df (q1 - q0), the colour
// part of d loss / d df is kc (q1 - q0), and the TF gradient of the colour channels is adjColour.c x sum(kc w) (the
// factored window of TfGrad). 41 FMA-pipe operations per sample instead of 51 (tf_eval_cached + blend_adjoint_T +
// tf_adjoint); same algebra, sums re-associated.
#ifndef DVR_TF_FACTORED
#define DVR_TF_FACTORED 1
#endif
template <int TFACC, bool FACTORED>
DVR_DEV float texture_adjoint_factored(const float* s_tf, int R, float density, float& T, f4& adjAcc, float step,
                                       TfGrad<kTfTexture, TFACC, FACTORED>& tfg, float& adjStep) {
  int i0, i1, fl;
  float df;
  texture_coords(density, R, i0, i1, df, fl);
  const f4 v0 = lds4(s_tf + 4 * i0), v1 = lds4(s_tf + 4 * i1);
  const float q0 = fmaf(adjAcc.z, v0.z, fmaf(adjAcc.y, v0.y, adjAcc.x * v0.x));
  const float q1 = fmaf(adjAcc.z, v1.z, fmaf(adjAcc.y, v1.y, adjAcc.x * v1.x));
  const float d3 = q1 - q0;
  const float S = fmaf(df, d3, q0);
  const float dvw = v1.w - v0.w;
  const float curw = fmaf(df, dvw, v0.w);              // the absorption the forward pass blended (tf_eval)
  const float e = blend_exp(-curw * step);
  const float alpha = 1.f - e;
  const float Tin = T * rcp_refined(e);
  const float kc = alpha * Tin;
  const float adjAlphaE = (adjAcc.w + S) * T;
  adjAcc.w = fmaf(-S, alpha, adjAcc.w * e);
  adjStep = adjAlphaE * curw;
  const float adjW = adjAlphaE * step;
  T = Tin;
  if constexpr (FACTORED) tfg.add_pair_k(fl, kc, adjW, df);
  return fmaf(kc, d3, adjW * dvw) * float(R);
}

// Which per-ray gradients the instance carries: none, the stepsize only (one accumulator; C5), or the camera's ray
// start / direction (+ the stepsize when AdjointArgs::hasStep)
enum : int { kRayNone = 0, kRayStep = 1, kRayCam = 2 };

// Resident CTAs the register allocation aims for. Instances that differentiate the ray: 6 CTAs/SM (<= 80 registers;
// C3 adjoint(camera) 26.2 -> 24.8 ms). The full-flag instance of C5 (volume + TF + stepsize), measured: before the
// look-ahead loads 4 CTAs / 108 registers 39.1 ms, 5 / 96 33.0 ms, 6 / 80 30.9 ms (8 bytes of stack; the first version
// of this instance spilled 32 bytes under that cap and took 36.3 ms); with the look-ahead loads (which carry the next
// sample's position) 6 / 80 spills 16 bytes: 28.1 ms, 5 / 96 without spills: 27.7 ms -> 5. Register accumulators of a
// Gaussian table and the GaussianLog adjoint need 4 CTAs (<= 128 registers) to stay out of local memory. Forcing more
// CTAs on the instances without ray gradients (forward <= 40 registers, adjoint(volume) <= 48) was measured and is slower.
#ifndef DVR_ADJ_CTAS_2
#define DVR_ADJ_CTAS_2 5   // ray gradients + volume + TF state: <= 96 registers
#endif
#ifndef DVR_ADJ_CTAS_1
#define DVR_ADJ_CTAS_1 6   // ray gradients + one of the two: <= 80 registers
#endif
template <int FILTER, int TF, bool HAS_VOL, int TFACC, int RAY>
constexpr int adjoint_min_ctas() {
  if (RAY == kRayNone) return 1;
  // register accumulators of a Gaussian table (up to kRegRows x 6 floats of state): <= 128 registers
  if ((TFACC == kTfAccReg && TF == kTfGaussian) || TF == kTfGaussianLog) return 4;
  const int extra = (HAS_VOL ? 1 : 0) + (TFACC != kTfAccNone ? 1 : 0);
  // the Linear TF's gradient window carries two more registers per row (the control-point positions): one CTA fewer
  const int linear = (TF == kTfLinear && TFACC != kTfAccNone) ? 1 : 0;
  return (extra == 2 ? DVR_ADJ_CTAS_2 : (extra == 1 ? DVR_ADJ_CTAS_1 : 6)) - linear;
}

template <int FILTER, int TF, int BLEND, bool HAS_VOL, int TFACC, int RAY>
__global__ void __launch_bounds__(kThreads, adjoint_min_ctas<FILTER, TF, HAS_VOL, TFACC, RAY>()) k_adjoint(const RayInputs in,
                                                                                             const AdjointArgs aa) {
  constexpr bool HAS_RAY = RAY != kRayNone;
  extern __shared__ __align__(16) float smem[];
  constexpr int ROW = TfLayout<TF>::kRow;
  constexpr int CH = TfChannels<TF>::v;
  const int b = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* s_tf = smem;
  float* s_acc = smem + ((in.R * ROW + 3) & ~3);  // TF-gradient tables
  const int entries = in.R * CH;
  constexpr int NTAB = TFACC == kTfAccSmem ? kThreads / 32 : 1;
  stage_tf<TF>(in, b, s_tf);
  if constexpr (TFACC != kTfAccNone)
    for (int i = threadIdx.x; i < entries * NTAB; i += blockDim.x) s_acc[i] = 0.f;
  __syncthreads();

  int x, y;
  thread_pixel(x, y);
  const bool active = x < in.W && y < in.H;
  // Beer-Lambert: the loop carries the transmittance in acc.w (blend_adjoint_T). Not in the instances that differentiate
  // the ray position: their config-size gradients sit within 1e-4 of the fp32 reference build only because the
  // inversion rounds like the reference's (with T they move to 8e-4 from it -- and from 8.6e-4 to 3.3e-4 from the
  // reference's fp64 build, profiles/r02_xinner.md; the parity gate is the fp32 build)
  constexpr bool TMODE = BLEND == kBlendBeer && DVR_BLEND_T && RAY != kRayCam;
  // Texture tables: the factored sample adjoint (texture_adjoint_factored), with the factored window when the instance
  // accumulates the TF gradient in shared memory
  constexpr bool TEXF = TMODE && TF == kTfTexture && TFACC != kTfAccReg && DVR_TF_FACTORED;
  TfGrad<TF, TFACC, TEXF && TFACC == kTfAccSmem> tfg;
  tfg.init(TFACC == kTfAccSmem ? s_acc + warp * entries : s_acc, in.R);

  if (active) {
    const Ray ray = make_ray(in, x, y, b);
    const float* cp = aa.color.p + b * aa.color.b + y * aa.color.y + x * aa.color.x;
    const float* ap = aa.adjColor.p + b * aa.adjColor.b + y * aa.adjColor.y + x * aa.adjColor.x;
    f4 acc = {__ldg(cp), __ldg(cp + aa.color.c), __ldg(cp + 2 * aa.color.c), __ldg(cp + 3 * aa.color.c)};
    f4 adjAcc = {__ldg(ap), __ldg(ap + aa.adjColor.c), __ldg(ap + 2 * aa.adjColor.c), __ldg(ap + 3 * aa.adjColor.c)};
    int n = __ldg(aa.term + b * aa.tB + y * aa.tY + x * aa.tX);
    // A ray whose adjoint colour is zero in all four channels contributes exact zeros to every output (converged or
    // masked pixels of a reconstruction): skip its march. The per-ray outputs below are still written (as zeros).
    if ((((__float_as_uint(adjAcc.x) | __float_as_uint(adjAcc.y) | __float_as_uint(adjAcc.z) | __float_as_uint(adjAcc.w)) &
          0x7fffffffu) | unsigned(aa.marchZeroRays)) == 0u)
      n = 0;
    float adjStep = 0.f;
    f3 adjStart = {0.f, 0.f, 0.f}, adjDir = {0.f, 0.f, 0.f};
    Sampler<FILTER, HAS_VOL, DVR_BRANCHFREE_ADJ || (DVR_BRANCHFREE_NORAY && !HAS_RAY)> smp;
    if constexpr (FILTER == kFilterTrilinearQuad && HAS_VOL)
      smp.init(in, b, reinterpret_cast<float*>(aa.gradQuads), aa.gradQuadB * 4);
    else
      smp.init(in, b, aa.adjVolume, aa.avB);
    // terminationIndex-1 was the last sample that contributed (renderer_kernels.cuh:374-376)
    float fi = float(n - 1);
    if constexpr (TMODE) acc.w = 1.f - acc.w;
    if constexpr (TEXF) tfg.set_colour_adjoint(adjAcc.x, adjAcc.y, adjAcc.z);
    // Identity TF: ray constants of blend_adjoint_grey
    [[maybe_unused]] const float idSa = s_tf[0], idSc = s_tf[1];
    [[maybe_unused]] const float idKS = (adjAcc.x + adjAcc.y + adjAcc.z) * idSc;
    [[maybe_unused]] const float idStepSa = ray.step * idSa;
    // Without position derivatives the sample position needs no t: p_i = (ps + tmin pd) + i (step pd), two hoisted
    // vectors and one FMA per axis (the forward pass rounds tmin + i step first; the positions agree to an ulp, which
    // moves the reconstructed densities by ~1e-6 relative, far inside the gradient tolerance).
    [[maybe_unused]] const f3 p0 = {fmaf(ray.tmin, ray.pd.x, ray.ps.x), fmaf(ray.tmin, ray.pd.y, ray.ps.y),
                                    fmaf(ray.tmin, ray.pd.z, ray.ps.z)};
    [[maybe_unused]] const f3 dp = {ray.step * ray.pd.x, ray.step * ray.pd.y, ray.step * ray.pd.z};
    // Look-ahead cell loads (Sampler::lookahead): measured per config (profiles/r02_lookahead.md) -- C2 adjoint(tf)
    // 3.52 -> 3.35 ms, C3 adjoint(camera) 20.1 -> 18.9, C5 adjoint(volume + tf + stepsize) 30.2 -> 28.3; the lean
    // Identity-TF adjoint(volume) of C4 is the exception (7.04 -> 7.27 ms: the gradient flush becomes a second divergent
    // region at the bottom of an iteration that has little arithmetic to hide a load under), so it keeps fetch().
    constexpr bool LOOKAHEAD = FILTER == kFilterTrilinearQuad && (HAS_RAY ? DVR_LOOKAHEAD_RAY : DVR_LOOKAHEAD_ADJ) &&
                               !(TF == kTfIdentity && HAS_VOL && !HAS_RAY);
    // position of sample number f (and its ray parameter t for the kernels that differentiate the ray)
    auto position = [&](float f, float& tOut) -> f3 {
      if constexpr (HAS_RAY) {
        tOut = sample_t(ray, f);
        return sample_pos<true>(ray, tOut);
      } else {
        tOut = 0.f;
        const float2 xy = pk_fma(f, make_float2(dp.x, dp.y), make_float2(p0.x, p0.y));
        return {xy.x, xy.y, fmaf(f, dp.z, p0.z)};
      }
    };
    // look-ahead state (Sampler::lookahead): position of the sample the loop is at, already inside the cached cell
    [[maybe_unused]] float tCur = 0.f;
    [[maybe_unused]] f3 pCur = {0.f, 0.f, 0.f};
    if constexpr (LOOKAHEAD) {
      if (fi >= 0.f) {
        pCur = position(fi, tCur);
        unsigned q0 = 0u;
        smp.lookahead(in, pCur.x, pCur.y, pCur.z, q0);  // first cell: loads unconditionally (no cell cached yet)
        smp.qoff = q0;
      }
    }
    for (; fi >= 0.f; fi -= 1.0f) {  // fi counts n-1 .. 0 exactly (n < 2^24)
      float t = 0.f;
      f3 p;
      float density;
      [[maybe_unused]] f3 gUnit = {0.f, 0.f, 0.f};
      [[maybe_unused]] bool cellChanges = false;
      [[maybe_unused]] unsigned qNext = 0u;
      if constexpr (LOOKAHEAD) {
        t = tCur;
        p = pCur;
        if constexpr (HAS_RAY && DVR_GRAD_SHARED) {
          density = smp.interp_grad(p.x, p.y, p.z, gUnit);
        } else {
          density = smp.interp(p.x, p.y, p.z);
          if constexpr (HAS_RAY) gUnit = smp.unit_gradient();
        }
        // the corner registers are dead from here on: the previous sample's cell loads run under this sample's arithmetic
        const float fn = fi - 1.0f;
        if (fn >= 0.f) {
          pCur = position(fn, tCur);
          cellChanges = smp.lookahead(in, pCur.x, pCur.y, pCur.z, qNext);
        }
      } else {
        p = position(fi, t);
        if constexpr (FILTER == kFilterTrilinearQuad && HAS_VOL && DVR_PRED_ADJ) density = smp.fetch_pred(in, p.x, p.y, p.z);
        else density = smp.fetch(in, &aa, p.x, p.y, p.z);
      }
      float adjDensity, adjStepBlend;
      if constexpr (TF == kTfIdentity) {
        // r = g = b (renderer_tf.cuh:62-104): one colour lane through the blend inversion and the TF adjoint
        const float dc = fmaxf(0.f, fminf(density, 1.f));
        float ak, adjAlphaE;
        float adjDc;
        if constexpr (TMODE) adjDc = blend_adjoint_grey_T(acc.w, adjAcc.w, dc, idKS, idSa, ray.step, idStepSa, ak, adjAlphaE);
        else adjDc = blend_adjoint_grey<BLEND>(acc.w, adjAcc.w, dc, idKS, idSa, ray.step, ak, adjAlphaE);
        adjStepBlend = adjAlphaE * (dc * idSa);
        if constexpr (TFACC == kTfAccReg) {
          tfg.g[0] = fmaf(adjAlphaE * ray.step, dc, tfg.g[0]);                       // d/d scaleAbsorption
          tfg.g[1] = fmaf((adjAcc.x + adjAcc.y + adjAcc.z) * ak, dc, tfg.g[1]);      // d/d scaleColor
        } else if constexpr (TFACC != kTfAccNone) {  // never dispatched (two scalars fit registers); kept correct
          const float v[2] = {adjAlphaE * ray.step * dc, (adjAcc.x + adjAcc.y + adjAcc.z) * ak * dc};
          tfg.add_row(0, v);
        }
        adjDensity = (density > 0.f && density < 1.f) ? adjDc : 0.f;
      } else if constexpr (TEXF) {
        adjDensity = texture_adjoint_factored(s_tf, in.R, density, acc.w, adjAcc, ray.step, tfg, adjStepBlend);
      } else {
        TfCache<TF> tfCache;
        const f4 cur = tf_eval_cached<TF>(s_tf, in.R, density, tfCache);
        f4 adjCur;
        if constexpr (TMODE) blend_adjoint_T(acc.w, adjAcc, cur, ray.step, adjCur, adjStepBlend);
        else blend_adjoint<BLEND, false>(acc, adjAcc, cur, ray.step, adjCur, adjStepBlend);
        adjDensity = tf_adjoint<TF, TFACC>(s_tf, in.R, density, adjCur, tfg, tfCache);
      }
      f3 g;
      if constexpr (LOOKAHEAD) {
        smp.accumulate(adjDensity);
        g = {adjDensity * gUnit.x, adjDensity * gUnit.y, adjDensity * gUnit.z};
      } else {
        g = smp.template adjoint<HAS_RAY>(adjDensity);
      }
      if constexpr (RAY == kRayStep) {
        // d pos / d stepsize = i dir / voxel = i pd: one dot product and one accumulator
        adjStep = aa.compatStepOverwrite ? adjStepBlend : adjStep + adjStepBlend;
        adjStep = fmaf(fmaf(g.z, ray.pd.z, fmaf(g.y, ray.pd.y, g.x * ray.pd.x)), fi, adjStep);
      } else if constexpr (RAY == kRayCam) {
        // the reference overwrites its accumulator with the blend term (see compat flag)
        adjStep = aa.compatStepOverwrite ? adjStepBlend : adjStep + adjStepBlend;
        const f3 w = {g.x * ray.invVoxel.x, g.y * ray.invVoxel.y, g.z * ray.invVoxel.z};
        adjStart.x += w.x; adjStart.y += w.y; adjStart.z += w.z;
        adjDir.x = fmaf(w.x, t, adjDir.x); adjDir.y = fmaf(w.y, t, adjDir.y); adjDir.z = fmaf(w.z, t, adjDir.z);
        const float adjT = w.x * ray.dir.x + w.y * ray.dir.y + w.z * ray.dir.z;
        adjStep = fmaf(adjT, fi, adjStep);
      }
      if constexpr (LOOKAHEAD) {
        if (cellChanges) {  // the gradient of the cell this sample was in goes out before the next cell is adopted
          smp.flush(&aa);
          smp.qoff = qNext;
        }
      }
    }
    smp.flush(&aa);
    if constexpr (HAS_RAY) {
      if (aa.hasStep) {
        float* p = aa.adjStep + b * aa.asB + y * aa.asY + x * aa.asX;
        if (aa.stepAtomic) atomicAdd(p, adjStep);
        else *p = adjStep;
      }
      if (RAY == kRayCam && aa.hasCam) {
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

  if constexpr (TFACC != kTfAccNone) {
    if constexpr (TFACC == kTfAccReg) {
      constexpr int NREG = TfGrad<TF, TFACC>::NREG;
#pragma unroll
      for (int e = 0; e < NREG; ++e) {
        if (e < entries) {  // uniform
          const float v = warp_sum(tfg.g[e]);
          if (lane == 0 && v != 0.f) atomicAdd(s_acc + e, v);
        }
      }
    } else {
      tfg.flush_rows();
    }
    __syncthreads();
    for (int e = threadIdx.x; e < entries; e += blockDim.x) {
      float v = 0.f;
#pragma unroll
      for (int w = 0; w < NTAB; ++w) v += s_acc[w * entries + e];
      const int r = e / CH, c = e - r * CH;
      if (v != 0.f) atomicAdd(aa.adjTf + b * aa.atB + r * aa.atR + c * aa.atC, v);
    }
  }
}

}  // namespace dvr
