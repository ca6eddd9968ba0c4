"""renderer::RendererCuda over the C-ABI (diffdvr_b200/csrc/host/renderer_cuda_facade.{h,cpp}), CPU side: the
reference's accessor layout is converted faithfully, the kernel names are the reference's, and the variant
enumeration agrees with the library's. No compute call is made."""
import ctypes

import numpy as np
import pytest

import facade_loader as fl
import scenes
from diffdvr_b200 import capi


@pytest.fixture(scope="module")
def shim():
    return fl.load()


def test_accessor_layout(shim):
    out = (ctypes.c_int32 * 6)()
    shim.shim_sizes(out)
    # kernel::RendererInputs (renderer_settings.cuh:150-180): int2, Tensor4Read(40), int3, 2 x Tensor2Read(24), 2 x
    # Tensor4Read, Tensor3Read(32), real_t, 2 x Tensor3Read, real_t -- with natural alignment
    assert out[3] == 8            # volume follows the int2
    assert out[0] == 8 + 40 + 16 + 2 * 24 + 2 * 40 + 32 + 8 + 2 * 32 + 8
    assert out[1] == 40 + 32      # Tensor4RW + ITensor3RW
    assert out[2] == 40 + 8 + 32 + 8 + 2 * 40 + 32
    assert shim.shim_init() == 1  # ABI versions of the header and of the loaded library agree


@pytest.mark.parametrize("camera_mode", [capi.CAMERA_RAY_START_DIR, capi.CAMERA_INVERSE_VIEW_MATRIX,
                                         capi.CAMERA_REFERENCE_FRAME])
def test_inputs_conversion(shim, camera_mode):
    sc = scenes.make_scene(n=12, W=10, H=6, B=2, camera_mode=camera_mode, tf_mode=capi.TF_LINEAR, R=5, tensor_scalars=True)
    # an x-fastest volume (the reference's Volume class) and a batch-1 transfer function: strides must survive, the
    # stride of a size-1 dimension becomes 0 (the reference's broadcasting accessors, renderer_tensor.cuh:242-268)
    vol = np.ascontiguousarray(sc.volume.transpose(0, 3, 2, 1)).transpose(0, 3, 2, 1)
    sc.volume = vol
    c_in = sc.inputs().to_c()
    c_out = capi.DvrInputs()
    shim.shim_inputs_roundtrip(ctypes.byref(c_in), ctypes.byref(c_out))
    for name in ("screen_width", "screen_height", "volume_filter_mode", "camera_mode", "tf_mode", "blend_mode"):
        assert getattr(c_in, name) == getattr(c_out, name), name
    assert c_in.blending_early_out == c_out.blending_early_out
    names = ["volume", "box_min", "box_size", "step_size", "tf"]
    names += ["camera_ray_start", "camera_ray_dir"] if camera_mode == capi.CAMERA_RAY_START_DIR else ["camera_matrix"]
    for name in names:
        ptr, nd, size, stride = fl.tensor_fields(getattr(c_in, name))
        ptr2, nd2, size2, stride2 = fl.tensor_fields(getattr(c_out, name))
        assert ptr and ptr == ptr2 and nd == nd2 and size == size2, name
        assert stride2 == tuple(0 if n == 1 else s for n, s in zip(size, stride)), name
    assert fl.tensor_fields(c_out.volume)[3] == (0, 1, 12, 144)
    if camera_mode == capi.CAMERA_RAY_START_DIR:
        assert not c_out.camera_matrix.ptr
    else:
        assert not c_out.camera_ray_start.ptr and not c_out.camera_ray_dir.ptr
        assert c_out.camera_fov_y_radians == c_in.camera_fov_y_radians


def test_settings_and_adjoint_outputs_conversion(shim):
    d_tf = np.arange(12, dtype=np.int32).reshape(1, 3, 4)
    s_in = capi.make_fd_settings(12, 3, (0, 1, 2), (-1, 4, -1), d_tf, True)
    s_out = capi.DvrForwardDifferencesSettings()
    shim.shim_settings_roundtrip(ctypes.byref(s_in), ctypes.byref(s_out))
    assert (s_out.D, s_out.d_stepsize, tuple(s_out.d_ray_start), tuple(s_out.d_ray_dir), s_out.has_tf_derivatives) == \
        (12, 3, (0, 1, 2), (-1, 4, -1), 1)
    assert fl.tensor_fields(s_out.d_tf) == (d_tf.ctypes.data, 3, (1, 3, 4), (0, 4, 1))

    adj_vol = np.zeros((1, 4, 5, 6), np.float32)
    adj_step = np.zeros((2, 1, 1), np.float32)
    adj_tf = np.zeros((1, 3, 4), np.float32)
    a_in = capi.make_adjoint_outputs(adj_volume=adj_vol, adj_step_size=adj_step, adj_tf=adj_tf, tf_delayed_accumulation=True)
    a_out = capi.DvrAdjointOutputs()
    shim.shim_adjoint_outputs_roundtrip(ctypes.byref(a_in), ctypes.byref(a_out))
    assert (a_out.has_volume_derivatives, a_out.has_stepsize_derivatives, a_out.has_camera_derivatives,
            a_out.has_tf_derivatives, a_out.tf_delayed_accumulation) == (1, 1, 0, 1, 1)
    assert fl.tensor_fields(a_out.adj_volume) == (adj_vol.ctypes.data, 4, (1, 4, 5, 6), (0, 30, 6, 1))
    assert fl.tensor_fields(a_out.adj_step_size) == (adj_step.ctypes.data, 3, (2, 1, 1), (1, 0, 0))
    assert fl.tensor_fields(a_out.adj_tf) == (adj_tf.ctypes.data, 3, (1, 3, 4), (0, 4, 1))
    assert not a_out.adj_camera_ray_start.ptr and not a_out.adj_camera_ray_dir.ptr


def test_kernel_names_are_the_references(shim):
    # renderer_cuda.cpp:64-83, 120-147, 226-258: what the reference hands to its NVRTC loader
    f, c, t, b = capi.FILTER_TRILINEAR, capi.CAMERA_REFERENCE_FRAME, capi.TF_TEXTURE, capi.BLEND_BEER_LAMBERT
    modes = ("kernel::VolumeFilterMode::FilterTrilinear, kernel::CameraMode::CameraReferenceFrame, "
             "kernel::TFMode::TFTexture, kernel::BlendMode::BlendBeerLambert")
    assert fl.kernel_name(shim, 0, f, c, t, b) == f"kernel::DvrKernelForwardDevice<{modes}>"
    assert fl.kernel_name(shim, 1, f, c, t, b, D=6, has_cam=True) == \
        f"kernel::DvrKernelForwardGradientsDevice<{modes}, 6, false, true, false>"
    assert fl.kernel_name(shim, 2, f, c, t, b, has_tf=True, delayed=True, has_vol=True) == \
        f"kernel::DvrKernelAdjointDevice<{modes}, false, false, true, true, true>"
    assert fl.kernel_name(shim, 3, f, c, t, b) == "kernel::ForwardVariablesToGradientsDevice"
    assert "TFGaussianLog" in fl.kernel_name(shim, 2, f, c, capi.TF_GAUSSIAN_LOG, capi.BLEND_ALPHA)
    assert "BlendAlpha" in fl.kernel_name(shim, 2, f, c, capi.TF_GAUSSIAN_LOG, capi.BLEND_ALPHA)


def test_variant_enumeration_matches_the_library(shim):
    lib = capi.load()
    n = 0
    for p in (capi.PASS_FORWARD, capi.PASS_FORWARD_GRADIENTS, capi.PASS_ADJOINT):
        for f in range(4):
            for c in range(3):
                for t in range(7):
                    for b in range(2):
                        got = bool(shim.shim_has_variant(p, f, c, t, b))
                        assert got == lib.has_variant_pass(p, f, c, t, b)
                        n += got
    assert n > 100
    # what unittests/testCudaCompilation.cpp:109-287 compiles: trilinear x 3 cameras x 4 TFs x 2 blends, every pass
    for p in (capi.PASS_FORWARD, capi.PASS_FORWARD_GRADIENTS, capi.PASS_ADJOINT):
        for c in range(3):
            for t in (capi.TF_IDENTITY, capi.TF_TEXTURE, capi.TF_LINEAR, capi.TF_GAUSSIAN):
                for b in range(2):
                    assert shim.shim_has_variant(p, capi.FILTER_TRILINEAR, c, t, b)


def test_sync_mode_round_trip(shim):
    lib = capi.load()
    shim.shim_set_sync_mode(1)
    assert shim.shim_get_sync_mode() == 1 and lib.get_option("sync_mode") == 1
    shim.shim_set_sync_mode(0)
    assert shim.shim_get_sync_mode() == 0 and lib.get_option("sync_mode") == 0
