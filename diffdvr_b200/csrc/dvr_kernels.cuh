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
  if constexpr (TF == kTfIdentity) {
    // r = g = b for this TF (renderer_tf.cuh:62-73): march one colour channel; the table's two
    // scales stay in registers
    const float sa = smem[0], sc = smem[1];
    for (;; fi += 1.0f) {
      const float t = sample_t(ray, fi);
      if (t > ray.tmax) break;
      if (acc.w > in.earlyOut) break;
      const f3 p = sample_pos<false>(ray, t);
      const float dc = fmaxf(0.f, fminf(smp.fetch(in, nullptr, p.x, p.y, p.z), 1.f));
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

template <int TF, int TFACC>
struct TfGrad {
  static constexpr int CH = TfChannels<TF>::v;
  static constexpr int NREG = TFACC == kTfAccReg ? kRegRows * CH : 1;
  float g[NREG];     // kTfAccReg: whole table; indexing is fully unrolled (stays in registers)
  float row[2][CH];  // kTfAccSmem (Texture/Linear): cached rows base, base+1
  int base;
  float* table;      // kTfAccSmem: this warp's table

  DVR_DEV void init(float* warpTable) {
#pragma unroll
    for (int i = 0; i < NREG; ++i) g[i] = 0.f;
#pragma unroll
    for (int c = 0; c < CH; ++c) row[0][c] = row[1][c] = 0.f;
    base = -2;  // never aliases a valid row or row-1
    table = warpTable;
  }
  DVR_DEV void flush_rows() {
    if (base >= 0) {
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        if (row[0][c] != 0.f) atomicAdd(table + base * CH + c, row[0][c]);
        if (row[1][c] != 0.f) atomicAdd(table + (base + 1) * CH + c, row[1][c]);
        row[0][c] = row[1][c] = 0.f;
      }
    }
  }
  // contribution v[0..CH) to row r (two-row modes call this with r = i and r = i or i+1)
  DVR_DEV void add_row(int r, const float* v) {
    if constexpr (TFACC == kTfAccReg) {
#pragma unroll
      for (int rr = 0; rr < kRegRows; ++rr)
        if (rr == r) {
#pragma unroll
          for (int c = 0; c < CH; ++c) g[rr * CH + c] += v[c];
        }
    } else if constexpr (TF == kTfTexture || TF == kTfLinear) {
      if (r != base && r != base + 1) {  // only the first of the two calls can miss
        flush_rows();
        base = r;
      }
      if (r == base) {
#pragma unroll
        for (int c = 0; c < CH; ++c) row[0][c] += v[c];
      } else {
#pragma unroll
        for (int c = 0; c < CH; ++c) row[1][c] += v[c];
      }
    } else {
#pragma unroll
      for (int c = 0; c < CH; ++c) atomicAdd(table + r * CH + c, v[c]);
    }
  }
};

// adjoint of tf_eval; returns adj_density and adds the parameter adjoints to `acc`
// Gaussian TF: the adjoint needs the same exp(-(d-mu)^2 / 2 sigma^2) per row as the evaluation; the first
// kRegRows of them are kept in registers between the two (the precise expf is ~10 instructions).
template <int TF>
DVR_DEV f4 tf_eval_cached(const float* s_tf, int R, float density, float (&ni)[kRegRows]) {
  if constexpr (TF == kTfGaussian) {
    f4 c = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < kRegRows; ++i) {
      ni[i] = 0.f;
      if (i < R) {
        const f4 ci = lds4(s_tf + 8 * i);
        ni[i] = gauss_exp(gauss_arg(density, s_tf + 8 * i));
        c.x = fmaf(ci.x, ni[i], c.x); c.y = fmaf(ci.y, ni[i], c.y); c.z = fmaf(ci.z, ni[i], c.z); c.w = fmaf(ci.w, ni[i], c.w);
      }
    }
    for (int i = kRegRows; i < R; ++i) {
      const f4 ci = lds4(s_tf + 8 * i);
      const float n = gauss_exp(gauss_arg(density, s_tf + 8 * i));
      c.x = fmaf(ci.x, n, c.x); c.y = fmaf(ci.y, n, c.y); c.z = fmaf(ci.z, n, c.z); c.w = fmaf(ci.w, n, c.w);
    }
    return c;
  } else {
    return tf_eval<TF>(s_tf, R, density);
  }
}

template <int TF, int TFACC>
DVR_DEV float tf_adjoint(const float* s_tf, int R, float density, const f4& adjC, TfGrad<TF, TFACC>& acc,
                         const float (&cache)[kRegRows]) {
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
    int i0, i1;
    float df;
    texture_coords(density, R, i0, i1, df);
    const f4 v0 = lds4(s_tf + 4 * i0), v1 = lds4(s_tf + 4 * i1);
    const float adjDf = dot4(adjC, f4{v1.x - v0.x, v1.y - v0.y, v1.z - v0.z, v1.w - v0.w});
    if constexpr (HAS_TF) {
      const float w0 = 1.f - df;
      const float a0[4] = {adjC.x * w0, adjC.y * w0, adjC.z * w0, adjC.w * w0};
      const float a1[4] = {adjC.x * df, adjC.y * df, adjC.z * df, adjC.w * df};
      acc.add_row(i0, a0);
      acc.add_row(i1, a1);
    }
    return adjDf * float(R);
  } else if constexpr (TF == kTfLinear) {  // :315-400
    const int i = linear_interval(s_tf, R, density);
    const f4 v0 = lds4(s_tf + 8 * i), v1 = lds4(s_tf + 8 * (i + 1));
    const float p0 = s_tf[8 * i + 4], p1 = s_tf[8 * (i + 1) + 4];
    const float d2 = fmaxf(p0, fminf(density, p1));
    const float frac = (d2 - p0) / (p1 - p0);
    const float adjFrac = dot4(adjC, f4{v1.x - v0.x, v1.y - v0.y, v1.z - v0.z, v1.w - v0.w});
    if constexpr (HAS_TF) {
      const float denom = (p1 - p0) * (p1 - p0);
      const float w0 = 1.f - frac;
      const float a0[5] = {adjC.x * w0, adjC.y * w0, adjC.z * w0, adjC.w * w0, adjFrac * (d2 - p1) / denom};
      const float a1[5] = {adjC.x * frac, adjC.y * frac, adjC.z * frac, adjC.w * frac, adjFrac * (p0 - d2) / denom};
      acc.add_row(i, a0);
      acc.add_row(i + 1, a1);
    }
    return adjFrac / (p1 - p0);  // passes through the clamp unconditionally ("#if 1", :350-352)
  } else if constexpr (TF == kTfGaussian) {  // :562-614
    float adjDensity = 0.f;
    auto body = [&](int i, float* g, float ni) {
      const f4 ci = lds4(s_tf + 8 * i);
      const float mu = s_tf[8 * i + 4], sigma = s_tf[8 * i + 5];
      const float f1 = div_rcp(mu - density, sigma * sigma, s_tf[8 * i + 7]) * ni;
      const float dc = dot4(adjC, ci);
      adjDensity = fmaf(dc, f1, adjDensity);
      if constexpr (HAS_TF) {
        const float f2 = f1 * (mu - density) / sigma;
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
          body(i, g, cache[i]);
#pragma unroll
          for (int c = 0; c < 6; ++c) acc.g[i * 6 + c] += g[c];
        }
    } else {
#pragma unroll
      for (int i = 0; i < kRegRows; ++i)
        if (i < R) {
          float g[6];
          body(i, g, cache[i]);
          if constexpr (HAS_TF) acc.add_row(i, g);
        }
      for (int i = kRegRows; i < R; ++i) {
        float g[6];
        body(i, g, gauss_exp(gauss_arg(density, s_tf + 8 * i)));
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

template <int FILTER, int TF, int BLEND, bool HAS_VOL, int TFACC, bool HAS_RAY>
// HAS_RAY instances: <= 80 registers for 6 CTAs/SM (C3 adjoint(camera) 26.2 -> 24.8 ms). Forcing more CTAs on the
// others (forward <= 40 registers, adjoint(volume) <= 48) was measured and is slower: they are issue bound.
__global__ void __launch_bounds__(kThreads, HAS_RAY ? 6 : 1) k_adjoint(const RayInputs in, const AdjointArgs aa) {
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
  TfGrad<TF, TFACC> tfg;
  tfg.init(TFACC == kTfAccSmem ? s_acc + warp * entries : s_acc);

  if (active) {
    const Ray ray = make_ray(in, x, y, b);
    const float* cp = aa.color.p + b * aa.color.b + y * aa.color.y + x * aa.color.x;
    const float* ap = aa.adjColor.p + b * aa.adjColor.b + y * aa.adjColor.y + x * aa.adjColor.x;
    f4 acc = {__ldg(cp), __ldg(cp + aa.color.c), __ldg(cp + 2 * aa.color.c), __ldg(cp + 3 * aa.color.c)};
    f4 adjAcc = {__ldg(ap), __ldg(ap + aa.adjColor.c), __ldg(ap + 2 * aa.adjColor.c), __ldg(ap + 3 * aa.adjColor.c)};
    const int n = __ldg(aa.term + b * aa.tB + y * aa.tY + x * aa.tX);
    float adjStep = 0.f;
    f3 adjStart = {0.f, 0.f, 0.f}, adjDir = {0.f, 0.f, 0.f};
    Sampler<FILTER, HAS_VOL> smp;
    if constexpr (FILTER == kFilterTrilinearQuad && HAS_VOL)
      smp.init(in, b, reinterpret_cast<float*>(aa.gradQuads), aa.gradQuadB * 4);
    else
      smp.init(in, b, aa.adjVolume, aa.avB);
    // terminationIndex-1 was the last sample that contributed (renderer_kernels.cuh:374-376)
    float fi = float(n - 1);
    // Identity TF: ray constants of blend_adjoint_grey
    [[maybe_unused]] const float idSa = s_tf[0], idSc = s_tf[1];
    [[maybe_unused]] const float idKS = (adjAcc.x + adjAcc.y + adjAcc.z) * idSc;
    // Without position derivatives the sample position needs no t: p_i = (ps + tmin pd) + i (step pd), two hoisted
    // vectors and one FMA per axis (the forward pass rounds tmin + i step first; the positions agree to an ulp, which
    // moves the reconstructed densities by ~1e-6 relative, far inside the gradient tolerance).
    [[maybe_unused]] const f3 p0 = {fmaf(ray.tmin, ray.pd.x, ray.ps.x), fmaf(ray.tmin, ray.pd.y, ray.ps.y),
                                    fmaf(ray.tmin, ray.pd.z, ray.ps.z)};
    [[maybe_unused]] const f3 dp = {ray.step * ray.pd.x, ray.step * ray.pd.y, ray.step * ray.pd.z};
    for (; fi >= 0.f; fi -= 1.0f) {  // fi counts n-1 .. 0 exactly (n < 2^24)
      float t = 0.f;
      f3 p;
      if constexpr (HAS_RAY) {
        t = sample_t(ray, fi);
        p = sample_pos<true>(ray, t);
      } else {
        const float2 xy = pk_fma(fi, make_float2(dp.x, dp.y), make_float2(p0.x, p0.y));
        p = {xy.x, xy.y, fmaf(fi, dp.z, p0.z)};
      }
      const float density = smp.fetch(in, &aa, p.x, p.y, p.z);
      float adjDensity, adjStepBlend;
      if constexpr (TF == kTfIdentity) {
        // r = g = b (renderer_tf.cuh:62-104): one colour lane through the blend inversion and the TF adjoint
        const float dc = fmaxf(0.f, fminf(density, 1.f));
        float ak, adjAlphaE;
        const float adjDc = blend_adjoint_grey<BLEND>(acc.w, adjAcc.w, dc, idKS, idSa, ray.step, ak, adjAlphaE);
        adjStepBlend = adjAlphaE * (dc * idSa);
        if constexpr (TFACC == kTfAccReg) {
          tfg.g[0] = fmaf(adjAlphaE * ray.step, dc, tfg.g[0]);                       // d/d scaleAbsorption
          tfg.g[1] = fmaf((adjAcc.x + adjAcc.y + adjAcc.z) * ak, dc, tfg.g[1]);      // d/d scaleColor
        } else if constexpr (TFACC != kTfAccNone) {  // never dispatched (two scalars fit registers); kept correct
          const float v[2] = {adjAlphaE * ray.step * dc, (adjAcc.x + adjAcc.y + adjAcc.z) * ak * dc};
          tfg.add_row(0, v);
        }
        adjDensity = (density > 0.f && density < 1.f) ? adjDc : 0.f;
      } else {
        float tfCache[kRegRows];
        const f4 cur = tf_eval_cached<TF>(s_tf, in.R, density, tfCache);
        f4 adjCur;
        blend_adjoint<BLEND>(acc, adjAcc, cur, ray.step, adjCur, adjStepBlend);
        adjDensity = tf_adjoint<TF, TFACC>(s_tf, in.R, density, adjCur, tfg, tfCache);
      }
      const f3 g = smp.template adjoint<HAS_RAY>(adjDensity);
      if constexpr (HAS_RAY) {
        // the reference overwrites its accumulator with the blend term (see compat flag)
        adjStep = aa.compatStepOverwrite ? adjStepBlend : adjStep + adjStepBlend;
        const f3 w = {g.x * ray.invVoxel.x, g.y * ray.invVoxel.y, g.z * ray.invVoxel.z};
        adjStart.x += w.x; adjStart.y += w.y; adjStart.z += w.z;
        adjDir.x = fmaf(w.x, t, adjDir.x); adjDir.y = fmaf(w.y, t, adjDir.y); adjDir.z = fmaf(w.z, t, adjDir.z);
        const float adjT = w.x * ray.dir.x + w.y * ray.dir.y + w.z * ray.dir.z;
        adjStep = fmaf(adjT, fi, adjStep);
      }
    }
    smp.flush(&aa);
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
