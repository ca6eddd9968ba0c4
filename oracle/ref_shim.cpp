// TEST INFRASTRUCTURE ONLY -- never linked into the product.
// extern "C" entry points over the REFERENCE's own OpenMP host kernels; see ref_shim_common.h.
// The file is compiled several times (oracle/Makefile) with one SHIM_PART_* macro each so the
// template instantiations build in parallel:
//   SHIM_PART_CORE        error string, forward render, forward-variables->gradients
//   SHIM_PART_ADJOINT=tf  adjoint for one TF mode (0 Identity, 1 Texture, 2 Linear, 3 Gaussian,
//                         6 GaussianLog, 5 Preshaded)
#include "ref_shim_common.h"

using namespace refshim;

#ifdef SHIM_PART_CORE
namespace refshim {
thread_local std::string g_error;
}

extern "C" const char* REFSYM(last_error)() { return g_error.c_str(); }
extern "C" int REFSYM(real_size)() { return int(sizeof(real_t)); }

// Renderer::renderForward -> RendererCpu::renderForward (renderer/renderer_cpu.cpp:81-97)
extern "C" int REFSYM(render_forward)(const DvrInputs* inputs, const DvrOutputs* outputs, void*) {
  return guarded([&] {
    InputsHolder h;
    build_inputs(*inputs, h);
    kernel::RendererOutputs out{make_acc<kernel::Tensor4RW, 4>(outputs->color),
                                make_acc<kernel::ITensor3RW, 3>(outputs->termination_index)};
    dim3 vs{unsigned(h.W), unsigned(h.H), unsigned(h.B)};
    if (inputs->volume_filter_mode == DVR_FILTER_PRESHADED) {
      if (inputs->tf_mode != DVR_TF_PRESHADED)
        throw std::runtime_error("preshaded filter requires preshaded TF");
      with_camera(inputs->camera_mode, [&](auto cam) {
        with_blend(inputs->blend_mode, [&](auto blend) {
          kernel::DvrKernelForwardHost<kernel::FilterPreshaded, decltype(cam)::value,
                                       kernel::TFPreshaded, decltype(blend)::value>(vs, h.in, out);
        });
      });
      return;
    }
    with_filter(inputs->volume_filter_mode, [&](auto filter) {
      with_camera(inputs->camera_mode, [&](auto cam) {
        with_tf<true>(inputs->tf_mode, [&](auto tf) {
          with_blend(inputs->blend_mode, [&](auto blend) {
            kernel::DvrKernelForwardHost<decltype(filter)::value, decltype(cam)::value,
                                         decltype(tf)::value, decltype(blend)::value>(vs, h.in,
                                                                                      out);
          });
        });
      });
    });
  });
}

// Renderer::forwardVariablesToGradients -> RendererCpu (renderer/renderer_cpu.cpp:317-324)
extern "C" int REFSYM(forward_variables_to_gradients)(const DvrTensor* forward_variables,
                                                      const DvrTensor* grad_output_color,
                                                      const DvrForwardDifferencesSettings* s,
                                                      const DvrAdjointOutputs* adj, void*) {
  return guarded([&] {
    if (adj->has_volume_derivatives)
      throw std::runtime_error("volume derivatives are not supported with forward variables");
    const int B = int(forward_variables->size[0]), H = int(forward_variables->size[1]),
              W = int(forward_variables->size[2]);
    kernel::ForwardDifferencesOutput fv{make_acc<kernel::Tensor5RW, 5>(*forward_variables)};
    auto gc = make_acc<kernel::AdjointColor_t, 4>(*grad_output_color);
    DvrTensor none{};
    kernel::ForwardDifferencesSettings sd{
        s->d_stepsize, make_int3(s->d_ray_start[0], s->d_ray_start[1], s->d_ray_start[2]),
        make_int3(s->d_ray_dir[0], s->d_ray_dir[1], s->d_ray_dir[2]),
        make_acc<kernel::ITensor3Read, 3>(s->has_tf_derivatives ? s->d_tf : none),
        s->has_tf_derivatives != 0};
    kernel::AdjointOutputs ao = build_adjoint_outputs(*adj, B, H, W);
    dim3 vs{unsigned(W), unsigned(H), unsigned(B)};
    kernel::ForwardVariablesToGradientsHost(vs, fv, gc, sd, ao);
  });
}
#endif  // SHIM_PART_CORE

#ifdef SHIM_PART_ADJOINT
// Renderer::renderAdjoint -> RendererCpu::renderAdjoint (renderer/renderer_cpu.cpp:299-315);
// tfDelayedAccumulation is CUDA-only in the reference (renderer_kernels.cuh:370).
#define ADJ_CAT2(a, b) a##b
#define ADJ_CAT(a, b) ADJ_CAT2(a, b)
#define ADJ_FN ADJ_CAT(REFSYM(render_adjoint_tf), SHIM_PART_ADJOINT)

template <kernel::VolumeFilterMode F, kernel::TFMode T>
static void run_adjoint(const DvrInputs* inputs, InputsHolder& h,
                        kernel::RendererOutputsAsInput& fwd, kernel::AdjointColor_t& ac,
                        kernel::AdjointOutputs& ao, const DvrAdjointOutputs* adj) {
  dim3 vs{unsigned(h.W), unsigned(h.H), unsigned(h.B)};
  with_camera(inputs->camera_mode, [&](auto cam) {
    with_blend(inputs->blend_mode, [&](auto blend) {
      with_bool(adj->has_stepsize_derivatives, [&](auto sd) {
        with_bool(adj->has_camera_derivatives, [&](auto cd) {
          with_bool(adj->has_tf_derivatives, [&](auto tfd) {
            with_bool(adj->has_volume_derivatives, [&](auto vd) {
              kernel::DvrKernelAdjointHost<F, decltype(cam)::value, T, decltype(blend)::value,
                                           decltype(sd)::value, decltype(cd)::value,
                                           decltype(tfd)::value, decltype(vd)::value>(
                  vs, h.in, fwd, ac, ao);
            });
          });
        });
      });
    });
  });
}

extern "C" int ADJ_FN(const DvrInputs* inputs, const DvrOutputs* from_forward,
                      const DvrTensor* adj_color, const DvrAdjointOutputs* adj, void*) {
  return guarded([&] {
    if (!adj->has_stepsize_derivatives && !adj->has_camera_derivatives &&
        !adj->has_tf_derivatives && !adj->has_volume_derivatives)
      throw std::runtime_error("At least one adjoint variable must be activated!");
    InputsHolder h;
    build_inputs(*inputs, h);
    kernel::RendererOutputsAsInput fwd{make_acc<kernel::Tensor4Read, 4>(from_forward->color),
                                       make_acc<kernel::ITensor3Read, 3>(from_forward->termination_index)};
    auto ac = make_acc<kernel::AdjointColor_t, 4>(*adj_color);
    kernel::AdjointOutputs ao = build_adjoint_outputs(*adj, h.B, h.H, h.W);
    constexpr kernel::TFMode T = kernel::TFMode(
        SHIM_PART_ADJOINT == 0   ? kernel::TFIdentity
        : SHIM_PART_ADJOINT == 1 ? kernel::TFTexture
        : SHIM_PART_ADJOINT == 2 ? kernel::TFLinear
        : SHIM_PART_ADJOINT == 3 ? kernel::TFGaussian
        : SHIM_PART_ADJOINT == 5 ? kernel::TFPreshaded
                                 : kernel::TFGaussianLog);
    if constexpr (T == kernel::TFPreshaded) {
      if (inputs->volume_filter_mode != DVR_FILTER_PRESHADED)
        throw std::runtime_error("preshaded TF requires preshaded filter");
      run_adjoint<kernel::FilterPreshaded, T>(inputs, h, fwd, ac, ao, adj);
    } else {
      with_filter(inputs->volume_filter_mode, [&](auto filter) {
        run_adjoint<decltype(filter)::value, T>(inputs, h, fwd, ac, ao, adj);
      });
    }
  });
}
#endif  // SHIM_PART_ADJOINT

#ifdef SHIM_PART_ADJOINT_DISPATCH
extern "C" int REFSYM(render_adjoint_tf0)(const DvrInputs*, const DvrOutputs*, const DvrTensor*, const DvrAdjointOutputs*, void*);
extern "C" int REFSYM(render_adjoint_tf1)(const DvrInputs*, const DvrOutputs*, const DvrTensor*, const DvrAdjointOutputs*, void*);
extern "C" int REFSYM(render_adjoint_tf2)(const DvrInputs*, const DvrOutputs*, const DvrTensor*, const DvrAdjointOutputs*, void*);
extern "C" int REFSYM(render_adjoint_tf3)(const DvrInputs*, const DvrOutputs*, const DvrTensor*, const DvrAdjointOutputs*, void*);
extern "C" int REFSYM(render_adjoint_tf5)(const DvrInputs*, const DvrOutputs*, const DvrTensor*, const DvrAdjointOutputs*, void*);
extern "C" int REFSYM(render_adjoint_tf6)(const DvrInputs*, const DvrOutputs*, const DvrTensor*, const DvrAdjointOutputs*, void*);

extern "C" int REFSYM(render_adjoint)(const DvrInputs* inputs, const DvrOutputs* from_forward,
                                      const DvrTensor* adj_color, const DvrAdjointOutputs* adj,
                                      void* s) {
  switch (inputs->tf_mode) {
    case DVR_TF_IDENTITY: return REFSYM(render_adjoint_tf0)(inputs, from_forward, adj_color, adj, s);
    case DVR_TF_TEXTURE: return REFSYM(render_adjoint_tf1)(inputs, from_forward, adj_color, adj, s);
    case DVR_TF_LINEAR: return REFSYM(render_adjoint_tf2)(inputs, from_forward, adj_color, adj, s);
    case DVR_TF_GAUSSIAN: return REFSYM(render_adjoint_tf3)(inputs, from_forward, adj_color, adj, s);
    case DVR_TF_PRESHADED: return REFSYM(render_adjoint_tf5)(inputs, from_forward, adj_color, adj, s);
    case DVR_TF_GAUSSIAN_LOG: return REFSYM(render_adjoint_tf6)(inputs, from_forward, adj_color, adj, s);
    default: refshim::g_error = "unknown tf enum value"; return DVR_ERR_INVALID_ARGUMENT;
  }
}
#endif
