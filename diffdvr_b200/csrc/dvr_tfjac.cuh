// sm_100a kernel: forward-mode derivatives with respect to the TRANSFER FUNCTION in one marching pass for any D
// (DvrKernelForwardGradientsImpl with hasTFDerivative, renderer/renderer_kernels.cuh:161-255; TF seeding at
// renderer/renderer_tf.cuh:121-154, 255-288, 436-485).
//
// The reference pushes D tangents through every operation of every sample (cudAD::fvar<real_t, D>). The slot-major
// kernel of dvr_fwdgrad.cuh does the same in chunks of 8 slots, i.e. ceil(D / 8) marches (32 for a 64 x 4 texture).
// But the Jacobian is sparse in the sample index: one sample touches two rows of a Texture / Linear table, and the only
// coupling between samples is the transmittance. Written per sample j with T_j = 1 - acc.w before the sample,
// alpha_j, colour c_j, and C the final pixel value:
//     dC/d c_j     = T_j alpha_j                                                    (colour channels, same channel only)
//     dC/d alpha_j = T_j c_j - (C - C_{<=j}) / (1 - alpha_j)                        (all four channels; for the alpha
//                                                                                    channel c_j = 1)
// where C - C_{<=j} is everything behind sample j. C is known from the plain forward kernel, C_{<=j} is the running
// accumulator of this march: ONE pass, whatever D. Per sample the kernel adds 2 x 5 (Texture) or 2 x 9 (Linear)
// values into a two-row register window, which is flushed into the thread's own slice of the B*H*W*D*4 output when the
// ray moves to another pair of rows (plain read-modify-writes: a pixel's slots belong to one thread).
//
// The output must hold zeros (TF tangents only: the host zero-fills it) or the ray / stepsize tangents written by
// k_forward_gradients_inputs (slots shared between a TF entry and a ray input add, as unit seeds in one slot do).
// Gaussian tables touch every row at every sample: nothing sparse, but the same per-sample sensitivities hold, and the
// per-row parameter tangents reduce to three moments of the row's kernel n_r(d) -- sum w n, sum w n u, sum w n u (d - mu) / sigma
// with u = (d - mu) / sigma^2 -- for five weights w (T alpha for the colour path, (d alpha / d absorption) G[0..3] for the
// alpha path): 15 accumulators per row, kept in a per-thread column of shared memory ([entry][thread]: conflict free).
// One pass for any R that fits (15 R floats per thread: R <= 28); the colour constants c_r are applied once at the end.
#pragma once
#include "dvr_device.cuh"
#include "dvr_kernels.cuh"  // TfCache / tf_eval_cached

namespace dvr {

template <int TF> struct JacRow { static constexpr int n = 5; };          // (colour weight, d/d absorption x 4)
template <> struct JacRow<kTfLinear> { static constexpr int n = 9; };     // + d/d position x 4
template <> struct JacRow<kTfGaussian> { static constexpr int n = 1; };   // (unused: the moments live in shared memory)
constexpr int kGaussMoments = 15;  // per row: 3 moments x (colour weight + 4 alpha-path weights)

template <int FILTER, int TF, int BLEND>
__global__ void __launch_bounds__(kThreads, 4) k_tf_jacobian(const RayInputs in, const ForwardOutputs fwd,
                                                              const ForwardGradArgs ga) {
  static_assert(TF == kTfIdentity || TF == kTfTexture || TF == kTfLinear || TF == kTfGaussian, "no GaussianLog");
  extern __shared__ __align__(16) float smem[];
  constexpr int ROW = TfLayout<TF>::kRow;
  constexpr int NR = JacRow<TF>::n;
  const int b = blockIdx.z;
  float* s_tf = smem;
  int* s_dtf = reinterpret_cast<int*>(smem + ((in.R * ROW + 3) & ~3));
  // Gaussian: this thread's column of the moment table, tab[e * kThreads] for e in [0, 15 R)
  [[maybe_unused]] float* tab = smem + 2 * ((in.R * ROW + 3) & ~3) + threadIdx.x;
  stage_tf<TF>(in, b, s_tf);
  if constexpr (TF == kTfGaussian)
    for (int e = 0; e < kGaussMoments * in.R; ++e) tab[e * kThreads] = 0.f;
  {
    const int* src = ga.dTf + b * ga.dtB;
    for (int i = threadIdx.x; i < in.R * ROW; i += blockDim.x) {
      const int r = i / ROW, c = i - r * ROW;
      const int s = c < in.C ? __ldg(src + r * ga.dtR + c * ga.dtC) : -1;
      s_dtf[i] = (s >= 0 && s < ga.D) ? s : -1;
    }
  }
  __syncthreads();
  int x, y;
  thread_pixel(x, y);
  if (x >= in.W || y >= in.H) return;

  const Ray ray = make_ray(in, x, y, b);
  // outputs of the plain forward kernel for this pixel (written by the launch before this one)
  const float* cp = fwd.color + b * fwd.cB + y * fwd.cY + x * fwd.cX;
  const f4 fin = {cp[0], cp[fwd.cC], cp[2 * fwd.cC], cp[3 * fwd.cC]};
  const int n = fwd.term[b * fwd.tB + y * fwd.tY + x * fwd.tX];
  float* g = ga.grad + b * ga.gB + y * ga.gY + x * ga.gX;

  // g[slot][0..3] += v[0..3]
  auto add4 = [&](int slot, const float* v) {
    if (slot < 0) return;
    float* p = g + slot * ga.gD;
    if (ga.gradVec4) {
      float4 o = *reinterpret_cast<float4*>(p);
      o.x += v[0]; o.y += v[1]; o.z += v[2]; o.w += v[3];
      *reinterpret_cast<float4*>(p) = o;
    } else {
      p[0] += v[0]; p[ga.gC] += v[1]; p[2 * ga.gC] += v[2]; p[3 * ga.gC] += v[3];
    }
  };
  // the colour entries (r, 0..2) of one table row: entry (r, c) moves output channel c only
  auto add_colour = [&](int r, float v) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const int slot = s_dtf[r * ROW + c];
      if (slot >= 0) g[slot * ga.gD + c * ga.gC] += v;
    }
  };

  float row[2][NR];
#pragma unroll
  for (int k = 0; k < NR; ++k) row[0][k] = row[1][k] = 0.f;
  constexpr int kNoBase = -0x40000000;
  int base = kNoBase;  // Texture: floor of the texture coordinate (-1 .. R); Linear: interval index
  // one row of the window into the pixel's Jacobian slice (k = 0: row `base`, k = 1: row `base + 1`; Texture rows clamp)
  auto flush_one = [&](int k) {
    int r = base + k;
    if constexpr (TF == kTfTexture) r = min(max(r, 0), in.R - 1);
    const float* v = k ? row[1] : row[0];
    add_colour(r, v[0]);
    add4(s_dtf[r * ROW + 3], v + 1);
    if constexpr (TF == kTfLinear) add4(s_dtf[r * ROW + 4], v + 5);
  };
  auto flush = [&]() {
    if (base == kNoBase) return;
    flush_one(0);
    flush_one(1);
#pragma unroll
    for (int e = 0; e < NR; ++e) row[0][e] = row[1][e] = 0.f;
  };
  // the ray moves to another row pair: usually the neighbouring one, then one row of the window stays
  auto move_window = [&](int key) {
    if (key == base + 1) {
      flush_one(0);
#pragma unroll
      for (int e = 0; e < NR; ++e) { row[0][e] = row[1][e]; row[1][e] = 0.f; }
    } else if (key == base - 1) {
      flush_one(1);
#pragma unroll
      for (int e = 0; e < NR; ++e) { row[1][e] = row[0][e]; row[0][e] = 0.f; }
    } else {
      flush();
    }
    base = key;
  };

  Sampler<FILTER, false> smp;
  smp.init(in, b, nullptr, 0);
  f4 acc = {0.f, 0.f, 0.f, 0.f};
  float fi = 0.f;
  constexpr bool LOOKAHEAD = FILTER == kFilterTrilinearQuad && DVR_LOOKAHEAD;
  // the positions of the kernel that wrote `fwd` (uniform branch), so that the accumulator ends at `fin`
  auto position = [&](float f) -> f3 {
    const float t = sample_t(ray, f);
    return ga.exactPos ? sample_pos<true>(ray, t) : sample_pos<false>(ray, t);
  };
  if constexpr (LOOKAHEAD) {
    if (n > 0) {
      const f3 p = position(0.f);
      smp.lookahead_f(in, p.x, p.y, p.z);
    }
  }
  for (int i = 0; i < n; ++i, fi += 1.0f) {
    float density;
    if constexpr (LOOKAHEAD) {
      density = smp.interp_f();
      if (i + 1 < n) {  // the next sample's cell loads run under this sample's arithmetic (Sampler::lookahead)
        const f3 p = position(fi + 1.0f);
        smp.lookahead_f(in, p.x, p.y, p.z);
      }
    } else {
      const f3 p = position(fi);
      density = smp.fetch(in, nullptr, p.x, p.y, p.z);
    }
    [[maybe_unused]] TfCache<TF> tfc;
    f4 cur;
    if constexpr (TF == kTfGaussian) cur = tf_eval_cached<TF>(s_tf, in.R, density, tfc);  // the values of tf_eval
    else cur = tf_eval<TF>(s_tf, in.R, density);
    // blend: the values of blend_alpha / blend_over (the plain forward kernel), plus d alpha / d absorption
    const float xs = cur.w * ray.step;
    float alpha, da;
    if constexpr (BLEND == kBlendBeer) {
      const float e = blend_exp(-xs);
      alpha = 1.f - e;
      da = e * ray.step;
    } else {
      alpha = rminf(1.f, xs);
      da = (1.f < xs) ? 0.f : ray.step;  // du_min(1, w s) keeps the tangent of w s unless 1 < w s
    }
    const float T = 1.f - acc.w;
    const float q = T * alpha;  // d C[c] / d cur[c], c < 3
    acc = blend_over(acc, cur, alpha);
    // d C / d alpha_j: T c_j - (C - C_<=j) / (1 - alpha_j); nothing lies behind a sample with alpha = 1
    const float om = 1.f - alpha;
    float inv = 0.f;
    if (om > 0.f) {
      asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(om));
      inv = fmaf(fmaf(-om, inv, 1.0f), inv, inv);
    }
    const float G[4] = {fmaf(T, cur.x, -(fin.x - acc.x) * inv), fmaf(T, cur.y, -(fin.y - acc.y) * inv),
                        fmaf(T, cur.z, -(fin.z - acc.z) * inv), T - (fin.w - acc.w) * inv};
    if constexpr (TF == kTfGaussian) {
      const float wG[4] = {da * G[0], da * G[1], da * G[2], da * G[3]};
      auto one_row = [&](int r, float nr) {
        const float* rw = s_tf + 8 * r;
        const float diff = density - rw[4];
        const float m1 = nr * (diff * rw[7]);               // d n / d mu    = n (d - mu) / sigma^2
        const float m2 = m1 * diff * (rw[5] * rw[7]);        // d n / d sigma = n (d - mu)^2 / sigma^3
        float* t = tab + r * kGaussMoments * kThreads;
        const float m[3] = {nr, m1, m2};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          t[(5 * k) * kThreads] = fmaf(q, m[k], t[(5 * k) * kThreads]);
#pragma unroll
          for (int c = 0; c < 4; ++c) t[(5 * k + 1 + c) * kThreads] = fmaf(wG[c], m[k], t[(5 * k + 1 + c) * kThreads]);
        }
      };
#pragma unroll
      for (int r = 0; r < kRegRows; ++r)
        if (r < in.R) one_row(r, tfc.ni[r]);
      for (int r = kRegRows; r < in.R; ++r) one_row(r, gauss_exp(gauss_arg<false>(density, s_tf + 8 * r)));
    } else if constexpr (TF == kTfIdentity) {
      // cur = dc (sc, sc, sc, sa): row[0] = (d/d sc [one value for the three colour outputs], d/d sa x 4)
      const float dc = fmaxf(0.f, fminf(density, 1.f));
      row[0][0] = fmaf(q, dc, row[0][0]);
      const float k = dc * da;
#pragma unroll
      for (int c = 0; c < 4; ++c) row[0][1 + c] = fmaf(k, G[c], row[0][1 + c]);
    } else {
      int key;
      float w1;
      [[maybe_unused]] float fp0 = 0.f, fp1 = 0.f;
      [[maybe_unused]] f4 dv = {0.f, 0.f, 0.f, 0.f};
      if constexpr (TF == kTfTexture) {
        int i0, i1;
        texture_coords(density, in.R, i0, i1, w1, key);
      } else {
        key = linear_interval(s_tf, in.R, density);
        const float p0 = s_tf[8 * key + 4], p1 = s_tf[8 * (key + 1) + 4], rcp = s_tf[8 * key + 5];
        // d2 = du_max(p0, du_min(density, p1)) with the selection rules of the dual numbers (a < b ? a : b, a > b ? a : b)
        const float m = density < p1 ? density : p1;
        const bool isP0 = p0 > m, isP1 = !isP0 && !(density < p1);
        const float d2 = isP0 ? p0 : m;
        w1 = div_rcp(d2 - p0, p1 - p0, rcp);
        // d frac / d p0 = ((d2 is p0) - 1 + frac) / (p1 - p0), d frac / d p1 = ((d2 is p1) - frac) / (p1 - p0)
        fp0 = ((isP0 ? 0.f : -1.f) + w1) * rcp;
        fp1 = ((isP1 ? 1.f : 0.f) - w1) * rcp;
        const f4 v0 = lds4(s_tf + 8 * key), v1 = lds4(s_tf + 8 * (key + 1));
        dv = {v1.x - v0.x, v1.y - v0.y, v1.z - v0.z, v1.w - v0.w};
      }
      if (key != base) move_window(key);
      const float w0 = 1.f - w1;
      row[0][0] = fmaf(w0, q, row[0][0]);
      row[1][0] = fmaf(w1, q, row[1][0]);
      const float k0 = w0 * da, k1 = w1 * da;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        row[0][1 + c] = fmaf(k0, G[c], row[0][1 + c]);
        row[1][1 + c] = fmaf(k1, G[c], row[1][1 + c]);
      }
      if constexpr (TF == kTfLinear) {
        // a control-point position moves the interpolation weight of all four channels of the sample
        const float ka = dv.w * da;
        const float H[4] = {fmaf(q, dv.x, ka * G[0]), fmaf(q, dv.y, ka * G[1]), fmaf(q, dv.z, ka * G[2]), ka * G[3]};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          row[0][5 + c] = fmaf(fp0, H[c], row[0][5 + c]);
          row[1][5 + c] = fmaf(fp1, H[c], row[1][5 + c]);
        }
      }
    }
  }
  if constexpr (TF == kTfGaussian) {
    for (int r = 0; r < in.R; ++r) {
      const float* t = tab + r * kGaussMoments * kThreads;
      const f4 cr = lds4(s_tf + 8 * r);
      add_colour(r, t[0]);                                   // (r, 0..2): d C[c] / d c_r[c] = sum T alpha n_r
      const float a[4] = {t[1 * kThreads], t[2 * kThreads], t[3 * kThreads], t[4 * kThreads]};
      add4(s_dtf[r * ROW + 3], a);                           // (r, 3): absorption
#pragma unroll
      for (int k = 1; k < 3; ++k) {                          // (r, 4) mu and (r, 5) sigma: through all four channels
        const float w0 = t[(5 * k) * kThreads];
        const float v[4] = {fmaf(cr.x, w0, cr.w * t[(5 * k + 1) * kThreads]), fmaf(cr.y, w0, cr.w * t[(5 * k + 2) * kThreads]),
                            fmaf(cr.z, w0, cr.w * t[(5 * k + 3) * kThreads]), cr.w * t[(5 * k + 4) * kThreads]};
        add4(s_dtf[r * ROW + 3 + k], v);
      }
    }
  } else if constexpr (TF == kTfIdentity) {
    // table = (scaleAbsorption, scaleColor): entry (0, 1) moves the three colour outputs alike
    const int sColour = s_dtf[1];
    if (sColour >= 0) {
#pragma unroll
      for (int c = 0; c < 3; ++c) g[sColour * ga.gD + c * ga.gC] += row[0][0];
    }
    add4(s_dtf[0], &row[0][1]);
  } else {
    flush();
  }
}

}  // namespace dvr
