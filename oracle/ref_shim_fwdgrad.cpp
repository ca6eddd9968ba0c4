// TEST INFRASTRUCTURE ONLY -- never linked into the product.
// The reference's DvrKernelForwardGradientsHost (renderer/renderer_kernels.cuh:281-308)
// instantiated over the RESTATED cudAD (oracle/cudad_restated/, labelled "restated cudAD"
// wherever its numbers are reported). The reference's CPU back-end only instantiates
// D in {1,2,8,16,20,32,64} (renderer/renderer_cpu.cpp:100-110); the kernel itself is generic
// in D, and unused derivative slots stay zero, so the shim runs the smallest compiled
// D' >= D into a scratch B*H*W*D'*4 buffer and copies the first D slices out.
// Compiled once per TF mode (SHIM_FG_TF = 0 Identity, 1 Texture, 2 Linear, 3 Gaussian).
#include "ref_shim_common.h"

using namespace refshim;

#ifdef SHIM_FG_TF
#define FG_CAT2(a, b) a##b
#define FG_CAT(a, b) FG_CAT2(a, b)
#define FG_FN FG_CAT(REFSYM(render_forward_gradients_tf), SHIM_FG_TF)

template <int D, kernel::TFMode T>
static void run(const DvrInputs* inputs, InputsHolder& h, kernel::ForwardDifferencesSettings& sd,
                kernel::RendererOutputs& out, kernel::ForwardDifferencesOutput& go, bool hasStep,
                bool hasCam, bool hasTf) {
  dim3 vs{unsigned(h.W), unsigned(h.H), unsigned(h.B)};
  with_filter(inputs->volume_filter_mode, [&](auto filter) {
    with_camera(inputs->camera_mode, [&](auto cam) {
      with_blend(inputs->blend_mode, [&](auto blend) {
        with_bool(hasStep, [&](auto s) {
          with_bool(hasCam, [&](auto c) {
            with_bool(hasTf, [&](auto t) {
              if constexpr (decltype(s)::value || decltype(c)::value || decltype(t)::value) {
                kernel::DvrKernelForwardGradientsHost<decltype(filter)::value,
                                                      decltype(cam)::value, T,
                                                      decltype(blend)::value, D,
                                                      decltype(s)::value, decltype(c)::value,
                                                      decltype(t)::value>(vs, h.in, sd, out, go);
              } else {
                throw std::runtime_error("at least one derivative must be activated");
              }
            });
          });
        });
      });
    });
  });
}

extern "C" int FG_FN(const DvrInputs* inputs, const DvrForwardDifferencesSettings* s,
                     const DvrOutputs* outputs, const DvrTensor* gradients_out, void*) {
  return guarded([&] {
    InputsHolder h;
    build_inputs(*inputs, h);
    if (s->D <= 0) throw std::runtime_error("number of derivatives must be positive");
    const bool hasStep = s->d_stepsize >= 0;
    const bool hasCam = s->d_ray_start[0] >= 0 || s->d_ray_start[1] >= 0 ||
                        s->d_ray_start[2] >= 0 || s->d_ray_dir[0] >= 0 || s->d_ray_dir[1] >= 0 ||
                        s->d_ray_dir[2] >= 0;
    const bool hasTf = s->has_tf_derivatives != 0;
    // dummy d_tf (all -1) when not provided, like renderer.cpp:439-445 (there: uninitialised)
    std::vector<int> dummy;
    DvrTensor d_tf = s->d_tf;
    if (!d_tf.ptr) {
      if (hasTf) throw std::runtime_error("d_tf not defined, hence hasTfDerivatives must be false");
      const int64_t n = inputs->tf.size[0] * inputs->tf.size[1] * inputs->tf.size[2];
      dummy.assign(size_t(n), -1);
      d_tf = DvrTensor{dummy.data(), 3,
                       {inputs->tf.size[0], inputs->tf.size[1], inputs->tf.size[2]},
                       {inputs->tf.size[1] * inputs->tf.size[2], inputs->tf.size[2], 1}};
    }
    kernel::ForwardDifferencesSettings sd{
        s->d_stepsize, make_int3(s->d_ray_start[0], s->d_ray_start[1], s->d_ray_start[2]),
        make_int3(s->d_ray_dir[0], s->d_ray_dir[1], s->d_ray_dir[2]),
        make_acc<kernel::ITensor3Read, 3>(d_tf), hasTf};
    kernel::RendererOutputs out{make_acc<kernel::Tensor4RW, 4>(outputs->color),
                                make_acc<kernel::ITensor3RW, 3>(outputs->termination_index)};
    const int D = s->D;
    const int Dc = D <= 1 ? 1 : D <= 8 ? 8 : D <= 32 ? 32 : 0;
    if (Dc == 0) throw std::runtime_error("reference oracle compiled for D <= 32 only");
    const int64_t B = h.B, H = h.H, W = h.W;
    std::vector<real_t> scratch(size_t(B * H * W * Dc * 4), real_t(0));
    DvrTensor st{scratch.data(), 5, {B, H, W, Dc, 4}, {H * W * Dc * 4, W * Dc * 4, Dc * 4, 4, 1}};
    kernel::ForwardDifferencesOutput go{make_acc<kernel::Tensor5RW, 5>(st)};
    constexpr kernel::TFMode T = kernel::TFMode(SHIM_FG_TF);
    if (Dc == 1) run<1, T>(inputs, h, sd, out, go, hasStep, hasCam, hasTf);
    else if (Dc == 8) run<8, T>(inputs, h, sd, out, go, hasStep, hasCam, hasTf);
    else run<32, T>(inputs, h, sd, out, go, hasStep, hasCam, hasTf);
    // copy the first D slices
    real_t* g = static_cast<real_t*>(gradients_out->ptr);
    const int64_t* gs = gradients_out->stride;
    for (int64_t b = 0; b < B; ++b)
      for (int64_t y = 0; y < H; ++y)
        for (int64_t x = 0; x < W; ++x)
          for (int64_t d = 0; d < D; ++d)
            for (int64_t c = 0; c < 4; ++c)
              g[b * gs[0] + y * gs[1] + x * gs[2] + d * gs[3] + c * gs[4]] =
                  scratch[size_t((((b * H + y) * W + x) * Dc + d) * 4 + c)];
  });
}
#endif

#ifdef SHIM_FG_DISPATCH
extern "C" int REFSYM(render_forward_gradients_tf0)(const DvrInputs*, const DvrForwardDifferencesSettings*, const DvrOutputs*, const DvrTensor*, void*);
extern "C" int REFSYM(render_forward_gradients_tf1)(const DvrInputs*, const DvrForwardDifferencesSettings*, const DvrOutputs*, const DvrTensor*, void*);
extern "C" int REFSYM(render_forward_gradients_tf2)(const DvrInputs*, const DvrForwardDifferencesSettings*, const DvrOutputs*, const DvrTensor*, void*);
extern "C" int REFSYM(render_forward_gradients_tf3)(const DvrInputs*, const DvrForwardDifferencesSettings*, const DvrOutputs*, const DvrTensor*, void*);

extern "C" int REFSYM(render_forward_gradients)(const DvrInputs* inputs,
                                                const DvrForwardDifferencesSettings* s,
                                                const DvrOutputs* outputs,
                                                const DvrTensor* gradients_out, void* st) {
  switch (inputs->tf_mode) {
    case DVR_TF_IDENTITY: return REFSYM(render_forward_gradients_tf0)(inputs, s, outputs, gradients_out, st);
    case DVR_TF_TEXTURE: return REFSYM(render_forward_gradients_tf1)(inputs, s, outputs, gradients_out, st);
    case DVR_TF_LINEAR: return REFSYM(render_forward_gradients_tf2)(inputs, s, outputs, gradients_out, st);
    case DVR_TF_GAUSSIAN: return REFSYM(render_forward_gradients_tf3)(inputs, s, outputs, gradients_out, st);
    default: refshim::g_error = "forward-mode oracle: unsupported tf mode"; return DVR_ERR_UNSUPPORTED;
  }
}
#endif

#ifdef SHIM_FG_DISPATCH
// Known-answer fixture of the reference: unittests/testNonConvexDensity.cpp:8-80 (its committed
// output is unittests/NonConvexDensity.tsv). Same driver loop, calling the reference's own
// TransferFunctionEval<TFGaussian>::evalForwardGradients<1> and Blending<BeerLambert>::blend
// over the restated fvar<.,1>. out = {red, alpha, loss, dloss/dd1}.
extern "C" int REFSYM(nonconvex_density)(double d1_in, double* out) {
  return guarded([&] {
    real_t tfv[6] = {1, 0, 0, 1, 0, real_t(0.5)};  // red, green, blue, alpha, mean, variance
    DvrTensor tft{tfv, 3, {1, 1, 6}, {6, 6, 1}};
    const kernel::Tensor3Read tf_acc = make_acc<kernel::Tensor3Read, 3>(tft);
    const kernel::ITensor3Read tf_d_acc;
    const real_t stepsize = 1 / 32.f;
    const real_t d0 = -1;
    typedef cudAD::fvar<real4, 1> v4;
    typedef cudAD::fvar<real_t, 1> v1;
    const auto eval = [&](real_t d1_) {
      v4 color = v4::constant(make_real4(0));
      v1 d1 = v1::input<0>(d1_);
      int num_steps = std::max(32, static_cast<int>(std::abs((d1_ - d0) * stepsize)));
      real_t local_stepsize = 1.0f / num_steps;
      kernel::TransferFunctionEval<kernel::TFGaussian> tf;
      for (int i = 0; i <= num_steps; ++i) {
        real_t f = i / real_t(num_steps);
        v1 d = (1 - f) * d0 + f * d1;
        v4 c = tf.evalForwardGradients<1>(tf_acc, 0, d, tf_d_acc,
                                          kernel::integral_constant<bool, false>());
        color = kernel::Blending<kernel::BlendBeerLambert>::blend(color, c, local_stepsize);
      }
      return color;
    };
    const real4 ground_truth = eval(d0).value();
    v4 c = eval(real_t(d1_in));
    v1 l = cudAD::length(ground_truth - c);
    out[0] = c.value().x;
    out[1] = c.value().w;
    out[2] = l.value();
    out[3] = l.derivative<0>();
  });
}
#endif
