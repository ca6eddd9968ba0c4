/*
 * dvr_b200.h -- C-ABI of the B200-native differentiable volume ray marcher.
 *
 * Drop-in boundary for ONE hot path of shamanDevel/DiffDVR: the three data-parallel
 * kernels of renderer/renderer_kernels.cuh (plain forward, forward-mode derivatives,
 * adjoint) plus the forward-variables -> gradients reduction.
 *
 * Every entry point below replaces one method of the reference's CUDA back-end
 * `renderer::RendererCuda` (renderer/renderer_cuda.h:34-123), which in the reference
 * assembles a template name, NVRTC-compiles it (renderer/kernel_loader.cpp:193-261) and
 * calls cuLaunchKernel (renderer/renderer_cuda.cpp:85-118,120-183,185-224,226-317).
 * Here the variants are precompiled sm_100a instances behind a static dispatch table.
 *
 * Conventions
 *  - plain C, no torch types; all pointers are DEVICE pointers (fp32 / int32);
 *  - a tensor is {ptr, ndim, size[], stride[]} with strides in ELEMENTS (like torch);
 *    a leading/any dimension of size 1 broadcasts (stride forced to 0), exactly like the
 *    reference's BroadcastingTensorAccessor (renderer/renderer_tensor.cuh:140-144);
 *  - enums are int32 with the reference's numeric values
 *    (renderer/renderer_settings.cuh:81-146);
 *  - adjoint outputs are ACCUMULATED into caller-zeroed tensors
 *    (renderer/renderer_settings.cuh:227-229);
 *  - no device allocations except the QUAD scratch described at DvrCallOptions / DvrWorkspace (below);
 *  - reentrant: no process-wide state changes the result of a call (options travel per call in
 *    DvrCallOptions; the only global switch is "sync_mode", which the reference has too);
 *  - every function returns DVR_OK (0) or a negative error code, never throws;
 *    dvr_last_error() returns a thread-local message for the last failure;
 *  - launches are asynchronous on `stream` (a cudaStream_t passed as void*), like the
 *    reference's launches on the current torch stream (renderer/renderer_cuda.cpp:105).
 */
#ifndef DVR_B200_H
#define DVR_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DVR_MAX_DIMS 5

#if defined(__GNUC__)
#define DVR_API __attribute__((visibility("default")))
#else
#define DVR_API
#endif

/* error codes */
#define DVR_OK 0
#define DVR_ERR_INVALID_ARGUMENT (-1) /* shape / enum / flag validation failed      */
#define DVR_ERR_UNSUPPORTED (-2)      /* variant not compiled (see dvr_has_variant) */
#define DVR_ERR_CUDA (-3)             /* CUDA runtime reported an error             */

/* renderer/renderer_settings.cuh:81-90 */
enum DvrVolumeFilterMode {
  DVR_FILTER_NEAREST = 0,
  DVR_FILTER_TRILINEAR = 1,
  DVR_FILTER_PRESHADED = 3, /* volume holds RGBA in the batch dimension */
  DVR_FILTER_NETWORK = 4    /* not supported (scene networks are off the path) */
};
/* renderer/renderer_settings.cuh:92-126 */
enum DvrCameraMode {
  DVR_CAMERA_RAY_START_DIR = 0,       /* per-pixel rays  (B)*H*W*3 twice   */
  DVR_CAMERA_INVERSE_VIEW_MATRIX = 1, /* (B)*4*4 inverse view-projection   */
  DVR_CAMERA_REFERENCE_FRAME = 2      /* (B)*3*3 eye/right/up + fovY       */
};
/* renderer/renderer_settings.cuh:128-139 */
enum DvrTFMode {
  DVR_TF_IDENTITY = 0,    /* (B)*1*2 */
  DVR_TF_TEXTURE = 1,     /* (B)*R*4 */
  DVR_TF_LINEAR = 2,      /* (B)*R*5 */
  DVR_TF_GAUSSIAN = 3,    /* (B)*R*6 */
  DVR_TF_PRESHADED = 5,   /* TF ignored */
  DVR_TF_GAUSSIAN_LOG = 6 /* (B)*R*6, returns log-space colours */
};
/* renderer/renderer_settings.cuh:141-146 */
enum DvrBlendMode { DVR_BLEND_BEER_LAMBERT = 0, DVR_BLEND_ALPHA = 1 };

/* Strided tensor view; replaces kernel::GenericPackedTensorAccessor
 * (renderer/renderer_tensor.cuh:122-269). ptr == NULL means "not provided". */
typedef struct DvrTensor {
  void* ptr;
  int32_t ndim;
  int64_t size[DVR_MAX_DIMS];
  int64_t stride[DVR_MAX_DIMS]; /* in elements */
} DvrTensor;

/* Replaces renderer::RendererInputsHost (renderer/renderer.h:20-64) and
 * kernel::RendererInputs (renderer/renderer_settings.cuh:150-180).
 * Scalar variants of the host struct are passed BY VALUE: when the tensor's ptr is NULL
 * the *_value field next to it is used (no temporary tensor, no H2D copy; the reference
 * uploads a 1-element tensor per call, renderer/renderer.cpp:101-102,117-118,217-218). */
typedef struct DvrInputs {
  int32_t screen_width;  /* W */
  int32_t screen_height; /* H */

  int32_t volume_filter_mode; /* DvrVolumeFilterMode */
  DvrTensor volume;           /* (B)*X*Y*Z fp32, any strides (4*X*Y*Z in preshaded mode) */

  DvrTensor box_min;  /* (B)*3 */
  float box_min_value[3];
  DvrTensor box_size; /* (B)*3 */
  float box_size_value[3];

  int32_t camera_mode;         /* DvrCameraMode */
  DvrTensor camera_ray_start;  /* (B)*H*W*3, mode RAY_START_DIR */
  DvrTensor camera_ray_dir;    /* (B)*H*W*3, mode RAY_START_DIR */
  DvrTensor camera_matrix;     /* (B)*4*4 (INVERSE_VIEW_MATRIX) or (B)*3*3 (REFERENCE_FRAME) */
  float camera_matrix_value[16]; /* row-major 4*4 used when camera_matrix.ptr==NULL in
                                    INVERSE_VIEW_MATRIX mode (the glm::mat4 variant) */
  float camera_fov_y_radians;  /* mode REFERENCE_FRAME */

  DvrTensor step_size; /* (B)*(H)*(W) */
  float step_size_value;

  int32_t tf_mode; /* DvrTFMode */
  DvrTensor tf;    /* (B)*R*C */

  int32_t blend_mode;       /* DvrBlendMode */
  float blending_early_out; /* reference default: 1 - 1e-5 */
} DvrInputs;

/* Replaces renderer::RendererOutputsHost (renderer/renderer.h:67-71). */
typedef struct DvrOutputs {
  DvrTensor color;             /* B*H*W*4 fp32 */
  DvrTensor termination_index; /* B*H*W   int32 */
} DvrOutputs;

/* Replaces renderer::ForwardDifferencesSettingsHost (renderer/renderer.h:74-93). */
typedef struct DvrForwardDifferencesSettings {
  int32_t D;              /* number of traced derivatives */
  int32_t d_stepsize;     /* derivative slot or -1 */
  int32_t d_ray_start[3]; /* derivative slots or -1 */
  int32_t d_ray_dir[3];   /* derivative slots or -1 */
  DvrTensor d_tf;         /* (B)*R*C int32 slots or -1; may be NULL if !has_tf_derivatives */
  int32_t has_tf_derivatives;
} DvrForwardDifferencesSettings;

/* Replaces renderer::AdjointOutputsHost (renderer/renderer.h:95-111). */
typedef struct DvrAdjointOutputs {
  int32_t has_volume_derivatives;
  DvrTensor adj_volume; /* (B)*X*Y*Z */
  int32_t has_stepsize_derivatives;
  DvrTensor adj_step_size; /* (B)*(H)*(W) */
  int32_t has_camera_derivatives;
  DvrTensor adj_camera_ray_start; /* (B)*H*W*3 */
  DvrTensor adj_camera_ray_dir;   /* (B)*H*W*3 */
  int32_t has_tf_derivatives;
  int32_t tf_delayed_accumulation; /* accepted for API parity; both settings give the same result */
  DvrTensor adj_tf;                /* (B)*R*C */
} DvrAdjointOutputs;

/* RendererCuda::renderForward (renderer/renderer_cuda.cpp:85-118) ->
 * DvrKernelForwardDevice (renderer/renderer_kernels.cuh:92-106). */
DVR_API int dvr_render_forward(const DvrInputs* inputs, const DvrOutputs* outputs, void* stream);

/* RendererCuda::renderForwardGradients (renderer/renderer_cuda.cpp:120-183) ->
 * DvrKernelForwardGradientsDevice (renderer/renderer_kernels.cuh:258-279).
 * gradients_out: B*H*W*D*4 fp32. */
DVR_API int dvr_render_forward_gradients(const DvrInputs* inputs,
                                 const DvrForwardDifferencesSettings* settings,
                                 const DvrOutputs* outputs, const DvrTensor* gradients_out,
                                 void* stream);

/* RendererCuda::forwardVariablesToGradients (renderer/renderer_cuda.cpp:185-224) ->
 * ForwardVariablesToGradientsDevice (renderer/renderer_kernels.cuh:564-576). */
DVR_API int dvr_forward_variables_to_gradients(const DvrTensor* forward_variables,
                                       const DvrTensor* grad_output_color,
                                       const DvrForwardDifferencesSettings* settings,
                                       const DvrAdjointOutputs* adj_outputs, void* stream);

/* RendererCuda::renderAdjoint (renderer/renderer_cuda.cpp:226-317) ->
 * DvrKernelAdjointDevice (renderer/renderer_kernels.cuh:447-468). */
DVR_API int dvr_render_adjoint(const DvrInputs* inputs, const DvrOutputs* outputs_from_forward,
                       const DvrTensor* adj_color, const DvrAdjointOutputs* adj_outputs,
                       void* stream);

/* ---- callers either side of the marcher ------------------------ */

/* Camera::generateRays, forward (renderer/camera.cpp:182-213 -> kernel::generateRaysForward,
 * renderer/camera_impl.cu:17-94). viewport: B*3*3 rows eye/right/up; writes ray_start and
 * ray_dir, both B*H*W*3 (any strides; contiguous tensors with W % 4 == 0 take a 128-bit path). */
DVR_API int dvr_generate_rays(const DvrTensor* viewport, float fov_y_radians, int32_t screen_width,
                      int32_t screen_height, const DvrTensor* ray_start, const DvrTensor* ray_dir,
                      void* stream);

/* Camera::generateRays, backward (renderer/camera.cpp:215-246 -> kernel::generateRaysAdjoint,
 * renderer/camera_impl.cu:96-356, a cub segmented reduction in the reference).
 * adj_viewport (B*3*3) is OVERWRITTEN with the per-view sums, like the reference. */
DVR_API int dvr_generate_rays_adjoint(const DvrTensor* viewport, float fov_y_radians,
                              const DvrTensor* adj_ray_start, const DvrTensor* adj_ray_dir,
                              const DvrTensor* adj_viewport, void* stream);

/* Renderer::compareToImage (renderer/renderer.cpp:562-657 -> RendererCuda::compareToImage_*,
 * renderer/renderer_cuda_impl.cu:112-172 over renderer/renderer_compareToImage.cuh:9-65):
 * per-pixel L2 norm |color_input - color_reference| over RGBA and its D forward derivatives.
 *   reduce == 0: difference_out B*H*W*1, gradients_out B*H*W*D
 *   reduce != 0: difference_out 1, gradients_out D; the mean over all B*H*W pixels
 * Both outputs are overwritten (the reference zeroes them first). Any D >= 0 is accepted
 * (the reference's CUDA path prints an error and does nothing for D != 1,
 * renderer_cuda_impl.cu:106-110). */
DVR_API int dvr_compare_to_image(const DvrTensor* color_input, const DvrTensor* gradients_input,
                         const DvrTensor* color_reference, const DvrTensor* difference_out,
                         const DvrTensor* gradients_out, int32_t reduce, void* stream);

/* TFUtils::preshadeVolume (renderer/tf_utils.cpp:172-300 -> kernel::PreshadeVolume,
 * renderer/tf_utils_impl.cu:16-92): color_out[c][x][y][z] = TF(density_volume[0][x][y][z]), the input of
 * the DVR_FILTER_PRESHADED / DVR_TF_PRESHADED marcher. density_volume 1*X*Y*Z, tf 1*R*C of `tf_mode`
 * (Identity, Texture, Linear, Gaussian, GaussianLog), color_out 4*X*Y*Z (r, g, b, absorption). */
DVR_API int dvr_preshade_volume(const DvrTensor* density_volume, const DvrTensor* tf, int32_t tf_mode,
                        const DvrTensor* color_out, void* stream);

/* Its adjoint with respect to the density (kernel::PreshadeVolumeAdj, renderer/tf_utils_impl.cu:98-183):
 * adj_volume_out[0][x][y][z] = <adj_color[:, x, y, z], dTF/d density>; OVERWRITTEN like the reference
 * (no TF adjoint there either). */
DVR_API int dvr_preshade_volume_adjoint(const DvrTensor* density_volume, const DvrTensor* tf, int32_t tf_mode,
                                const DvrTensor* adj_color, const DvrTensor* adj_volume_out,
                                void* stream);

/* ---- renderer::Volume on the device ---------------------------------- */

/* Volume::ImplicitEquation (renderer/volume.h:163-181) */
enum DvrImplicitEquation {
  DVR_EQ_MARSCHNER_LOBB = 0, DVR_EQ_CUBE = 1, DVR_EQ_SPHERE = 2, DVR_EQ_INVERSE_SPHERE = 3, DVR_EQ_DING_DONG = 4,
  DVR_EQ_ENDRASS = 5, DVR_EQ_BARTH = 6, DVR_EQ_HEART = 7, DVR_EQ_KLEINE = 8, DVR_EQ_CASSINI = 9, DVR_EQ_STEINER = 10,
  DVR_EQ_CROSS_CAP = 11, DVR_EQ_KUMMER = 12, DVR_EQ_BLOBBY = 13, DVR_EQ_TUBE = 14
};

/* Volume::createImplicitDataset (renderer/volume_implicit.cpp:224-256 over Volume::createSyntheticDataset,
 * renderer/volume.cpp:369-387): volume_out[0][x][y][z] = f(boxMin + (x, y, z) * (boxMax - boxMin) / (n - 1)),
 * volume_out 1*n*n*n, any strides. params: HOST pointer to 4 floats {"fM", "alpha" (MarschnerLobb), "scale" (Cube),
 * "a" (Cassini)}; NULL = the reference's defaults {6, 0.25, 0.5, 0.25}. */
DVR_API int dvr_create_implicit_volume(int32_t equation, const float* params, const DvrTensor* volume_out,
                                       void* stream);

/* The conversion step of Volume::Volume(filename) (renderer/volume.cpp:232-276): `raw` = DEVICE pointer to the
 * decompressed .cvol body, X*Y*Z elements with X FASTEST, dtype 0 = u8 (/255), 1 = u16 (/65535), 2 = f32;
 * volume_out 1*X*Y*Z fp32, any strides. */
DVR_API int dvr_volume_from_raw(const void* raw, int32_t dtype, const DvrTensor* volume_out, void* stream);

/* The inverse, for Volume::save (renderer/volume.cpp:86-151): raw_out = DEVICE pointer to X*Y*Z floats, X fastest. */
DVR_API int dvr_volume_to_raw(const DvrTensor* volume, float* raw_out, void* stream);

/* Volume::createMipmapLevel (renderer/volume.cpp:359-365, at::adaptive_avg_pool3d): volume 1*X*Y*Z ->
 * volume_out 1*oX*oY*oZ, cell o averaging the input range [floor(o*I/O), ceil((o+1)*I/O)) per axis. */
DVR_API int dvr_downsample_volume(const DvrTensor* volume, const DvrTensor* volume_out, void* stream);

/* Measurement utility (bench accounting): re-walks every ray up to its
 * terminationIndex and counts samples and CELL VISITS (a sample whose integer cell differs
 * from the ray's previous one -- the unit the reference's cell cache reloads 8 corners for,
 * renderer/renderer_interpolation.cuh:221,337). counts = device pointer to 2 uint64
 * {N_samples, N_cells}, accumulated (caller zeroes). */
DVR_API int dvr_count_cell_visits(const DvrInputs* inputs, const DvrTensor* termination_index,
                                  uint64_t* counts, void* stream);

/* Computes the broadcast batch size and volume resolution exactly like
 * Renderer::checkInputs (renderer/renderer.cpp:49-263); validates shapes.
 * Any out pointer may be NULL. */
DVR_API int dvr_check_inputs(const DvrInputs* inputs, int32_t* B, int32_t* X, int32_t* Y, int32_t* Z);

/* ---- per-call options and the volume-layout workspace (ABI version 2) ------------------------------
 *
 * Trilinear sampling has two layouts:
 *   in place : the caller's strided tensor, 8 x 4-byte loads and 8 scalar REDs per cell;
 *   QUAD     : a 16-byte-per-voxel copy holding the (y,z) 2x2 footprint of every voxel: 2 x 16-byte loads per
 *              cell; the adjoint accumulates into a scratch of the same layout with 2 x red.v4 per cell, which
 *              is folded into adj_volume afterwards.
 * Where the QUAD buffers live:
 *   - options == NULL or options->workspace == NULL: stream-ordered scratch per call from a PRIVATE memory pool of
 *     the library (cudaMallocFromPoolAsync on `stream`; the device's default pool is not touched). The copy is
 *     rebuilt, the gradient scratch zeroed and folded on every call. dvr_trim_scratch() returns the pool's pages.
 *   - options->workspace != NULL: the buffers belong to the workspace and persist. The copy is rebuilt only when
 *     the volume tensor's (ptr, sizes, strides) differ from the cached key or after dvr_workspace_set_volume /
 *     dvr_workspace_invalidate (the caller's "version": call one of them when the CONTENTS changed).
 *     Between dvr_workspace_begin_volume_gradient and dvr_workspace_end_volume_gradient every adjoint call
 *     accumulates into ONE zeroed gradient scratch that is folded into adj_volume once at the end (the fold is
 *     linear), instead of memset + fold per call.
 * A workspace must be used from one stream at a time (its buffers are not stream-ordered). */
typedef struct DvrWorkspace DvrWorkspace;

typedef struct DvrCallOptions {
  uint32_t struct_size;              /* sizeof(DvrCallOptions), set by dvr_call_options_init */
  int32_t volume_layout;             /* -1 automatic (QUAD when rays * 64 >= voxels, or a valid workspace copy
                                        exists), 0 in place, 1 QUAD */
  int32_t compat_stepsize_overwrite; /* adjoint: 1 reproduces the reference's stepsize recurrence bit for bit (the
                                        blend term of the stepsize adjoint is OVERWRITTEN every sample,
                                        renderer/renderer_kernels.cuh:393-395 with renderer_blending.cuh:93,158);
                                        0 (default) accumulates it (the mathematically complete gradient) */
  int32_t sync;                      /* 1: cudaStreamSynchronize after the launches of this call */
  int64_t max_scratch_bytes;         /* cap on the per-call / workspace QUAD buffers of this call; above it the call
                                        takes the in-place path (counted, see dvr_get_stat). <= 0: 16 GiB */
  DvrWorkspace* workspace;           /* may be NULL */
  int32_t forward_mode_tf;           /* forward-mode with TF tangents: 0 automatic (ONE sample-major marching pass for any
                                        D, csrc/dvr_tfjac.cuh; Gaussian tables of more than 28 rows: ceil(D / 8)
                                        slot-major passes), 1 always the slot-major passes (the reference's scheme in
                                        chunks; kept for A/B measurements). Read only when struct_size covers it */
  int32_t adjoint_march_zero_rays;   /* adjoint: 0 (default) a ray whose adjoint colour is zero in all four channels is
                                        not marched (its contributions are exact zeros: converged / masked pixels of
                                        a reconstruction); 1 marches every ray like the reference does (same results;
                                        for like-for-like timings) */
} DvrCallOptions;

DVR_API void dvr_call_options_init(DvrCallOptions* options); /* the defaults: -1, 0, 0, 0, NULL, 0, 0 */

DVR_API DvrWorkspace* dvr_workspace_create(void);            /* on the current device; NULL on failure */
DVR_API void dvr_workspace_destroy(DvrWorkspace* ws);        /* frees its device buffers (synchronises) */
/* Builds the QUAD copy of `volume` ((B)*X*Y*Z) now and caches it under the tensor's (ptr, sizes, strides). */
DVR_API int dvr_workspace_set_volume(DvrWorkspace* ws, const DvrTensor* volume, void* stream);
/* Marks the cached copy stale (the next call that wants it rebuilds it). */
DVR_API int dvr_workspace_invalidate(DvrWorkspace* ws);
/* Zeroes the workspace's QUAD gradient scratch for adj_volume ((B)*X*Y*Z) and opens an accumulation scope. */
DVR_API int dvr_workspace_begin_volume_gradient(DvrWorkspace* ws, const DvrTensor* adj_volume, void* stream);
/* Folds the scratch into adj_volume (accumulating) and closes the scope. */
DVR_API int dvr_workspace_end_volume_gradient(DvrWorkspace* ws, const DvrTensor* adj_volume, void* stream);
DVR_API int64_t dvr_workspace_bytes(const DvrWorkspace* ws); /* device bytes currently owned */

/* The render entry points with per-call options (options == NULL: the defaults). The plain entry points above
 * are these with options == NULL. */
DVR_API int dvr_render_forward_ex(const DvrInputs* inputs, const DvrOutputs* outputs, const DvrCallOptions* options,
                                  void* stream);
DVR_API int dvr_render_forward_gradients_ex(const DvrInputs* inputs, const DvrForwardDifferencesSettings* settings,
                                            const DvrOutputs* outputs, const DvrTensor* gradients_out,
                                            const DvrCallOptions* options, void* stream);
DVR_API int dvr_render_adjoint_ex(const DvrInputs* inputs, const DvrOutputs* outputs_from_forward,
                                  const DvrTensor* adj_color, const DvrAdjointOutputs* adj_outputs,
                                  const DvrCallOptions* options, void* stream);

/* Forward render that ALLOCATES NOTHING: dvr_render_forward with the QUAD copy of the volume (the scratch the trilinear
 * marcher reads, see above) in a static device array of the library instead of pool / workspace memory. No counterpart
 * in the reference; added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it). No cudaMalloc, cudaFree, memory-pool or other
 * allocating call is made. Same inputs, outputs, checks, kernels and results as dvr_render_forward_ex with
 * volume_layout = 1.
 *   Limit : B_vol * (X + 1) * Y * Z <= DVR_NOALLOC_MAX_QUADS (B_vol = 1 for a volume shared by the batch), i.e. one
 *           volume of up to 128^3; larger volumes return DVR_ERR_UNSUPPORTED with a message (no fallback). The array
 *           (33.8 MB) is held by every GPU that loads the library, until the process exits.
 *   Shared: every call on a device uses the same array. Calls that may overlap on the GPU (different streams without
 *           an ordering between them) RACE on it; order them. The array keeps the last call's contents; the build
 *           kernel of each call overwrites every entry its march reads, so nothing has to start at zero.
 *   Nearest / preshaded filters and volumes thinner than 2 voxels run their usual kernels, which use no scratch. */
#define DVR_NOALLOC_MAX_QUADS 2113536 /* 129 * 128 * 128 float4 entries */
DVR_API int dvr_render_forward_noalloc(const DvrInputs* inputs, const DvrOutputs* outputs, void* stream);

/* Releases the pages the library's private scratch pool holds on the current device. */
DVR_API int dvr_trim_scratch(void);

/* Which precompiled instances exist. Replaces "does it compile under NVRTC"
 * (unittests/testCudaCompilation.cpp:109-287). Generated from the same list that declares and dispatches the
 * per-(tf, blend) instance translation units (DVR_FOR_EACH_*_INSTANCE in csrc/dvr_launch.h), so it cannot drift
 * from what links. pass: one of DvrPass. */
enum DvrPass { DVR_PASS_FORWARD = 0, DVR_PASS_FORWARD_GRADIENTS = 1, DVR_PASS_ADJOINT = 2 };
DVR_API int dvr_has_variant_pass(int32_t pass, int32_t volume_filter_mode, int32_t camera_mode, int32_t tf_mode,
                                 int32_t blend_mode);
/* forward AND adjoint available (the pair every optimisation needs) */
DVR_API int dvr_has_variant(int32_t volume_filter_mode, int32_t camera_mode, int32_t tf_mode,
                    int32_t blend_mode);

/* Process-wide switch, like Renderer::setCudaSyncMode (renderer/renderer.h:128-150):
 *   "sync_mode" 0/1  cudaStreamSynchronize after every launch (default 0)
 * Everything that changes RESULTS or memory use is per call (DvrCallOptions).
 * Returns DVR_OK or DVR_ERR_INVALID_ARGUMENT for an unknown name. */
DVR_API int dvr_set_option(const char* name, int32_t value);
DVR_API int dvr_get_option(const char* name, int32_t* value);

/* Counters since load: "launches", "quad_builds" (QUAD copies built), "quad_reuses" (calls served by a cached
 * workspace copy), "quad_fallbacks" (calls that wanted the QUAD layout and took the in-place path: scratch above
 * the cap, allocation failure, aliased adj_volume strides), "folds". */
DVR_API int dvr_get_stat(const char* name, int64_t* value);

/* Number of kernel launches issued by this library since load (for bench accounting). */
DVR_API int64_t dvr_launch_count(void);

/* pyrenderer.sync (bindings/bindings.cpp:306): cudaDeviceSynchronize on the current device. */
DVR_API int dvr_device_synchronize(void);

/* GPUTimer of the reference's bindings (bindings/bindings.cpp:99-129): a cudaEvent pair. */
typedef struct DvrTimer DvrTimer;
DVR_API DvrTimer* dvr_timer_create(void);
DVR_API void dvr_timer_destroy(DvrTimer* timer);
DVR_API int dvr_timer_start(DvrTimer* timer, void* stream);
DVR_API int dvr_timer_stop(DvrTimer* timer, void* stream);
DVR_API int dvr_timer_elapsed_ms(DvrTimer* timer, float* ms); /* waits for the stop event */

/* Thread-local description of the last error returned on this thread ("" if none). */
DVR_API const char* dvr_last_error(void);

/* ABI version of this header. */
DVR_API int32_t dvr_abi_version(void);
#define DVR_ABI_VERSION 3 /* 3: DvrCallOptions grew (forward_mode_tf); dvr_call_options_init writes sizeof(DvrCallOptions) of THIS header */

#ifdef __cplusplus
}
#endif
#endif /* DVR_B200_H */
