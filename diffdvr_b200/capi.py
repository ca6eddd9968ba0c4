"""ctypes mirror of ``include/dvr_b200.h`` (the C-ABI of the B200 ray marcher).

Host-side plumbing only: builds the plain-C structs from torch tensors (device memory,
strides) and calls ``libdvr_b200.so``.  There is NO CPU fallback: if the CUDA library is not
built, loading raises.  The struct classes are also what the tests use to drive the CPU
oracles (which export the same signatures with host pointers) -- the oracle itself is never
imported from here.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional, Sequence

DVR_MAX_DIMS = 5
DVR_OK = 0
DVR_ERR_INVALID_ARGUMENT = -1
DVR_ERR_UNSUPPORTED = -2
DVR_ERR_CUDA = -3

# enum values of renderer/renderer_settings.cuh:81-146
FILTER_NEAREST, FILTER_TRILINEAR, FILTER_PRESHADED, FILTER_NETWORK = 0, 1, 3, 4
CAMERA_RAY_START_DIR, CAMERA_INVERSE_VIEW_MATRIX, CAMERA_REFERENCE_FRAME = 0, 1, 2
TF_IDENTITY, TF_TEXTURE, TF_LINEAR, TF_GAUSSIAN, TF_PRESHADED, TF_GAUSSIAN_LOG = 0, 1, 2, 3, 5, 6
BLEND_BEER_LAMBERT, BLEND_ALPHA = 0, 1


class DvrTensor(ctypes.Structure):
    _fields_ = [
        ("ptr", ctypes.c_void_p),
        ("ndim", ctypes.c_int32),
        ("size", ctypes.c_int64 * DVR_MAX_DIMS),
        ("stride", ctypes.c_int64 * DVR_MAX_DIMS),
    ]


class DvrInputs(ctypes.Structure):
    _fields_ = [
        ("screen_width", ctypes.c_int32),
        ("screen_height", ctypes.c_int32),
        ("volume_filter_mode", ctypes.c_int32),
        ("volume", DvrTensor),
        ("box_min", DvrTensor),
        ("box_min_value", ctypes.c_float * 3),
        ("box_size", DvrTensor),
        ("box_size_value", ctypes.c_float * 3),
        ("camera_mode", ctypes.c_int32),
        ("camera_ray_start", DvrTensor),
        ("camera_ray_dir", DvrTensor),
        ("camera_matrix", DvrTensor),
        ("camera_matrix_value", ctypes.c_float * 16),
        ("camera_fov_y_radians", ctypes.c_float),
        ("step_size", DvrTensor),
        ("step_size_value", ctypes.c_float),
        ("tf_mode", ctypes.c_int32),
        ("tf", DvrTensor),
        ("blend_mode", ctypes.c_int32),
        ("blending_early_out", ctypes.c_float),
    ]


class DvrOutputs(ctypes.Structure):
    _fields_ = [("color", DvrTensor), ("termination_index", DvrTensor)]


class DvrForwardDifferencesSettings(ctypes.Structure):
    _fields_ = [
        ("D", ctypes.c_int32),
        ("d_stepsize", ctypes.c_int32),
        ("d_ray_start", ctypes.c_int32 * 3),
        ("d_ray_dir", ctypes.c_int32 * 3),
        ("d_tf", DvrTensor),
        ("has_tf_derivatives", ctypes.c_int32),
    ]


class DvrAdjointOutputs(ctypes.Structure):
    _fields_ = [
        ("has_volume_derivatives", ctypes.c_int32),
        ("adj_volume", DvrTensor),
        ("has_stepsize_derivatives", ctypes.c_int32),
        ("adj_step_size", DvrTensor),
        ("has_camera_derivatives", ctypes.c_int32),
        ("adj_camera_ray_start", DvrTensor),
        ("adj_camera_ray_dir", DvrTensor),
        ("has_tf_derivatives", ctypes.c_int32),
        ("tf_delayed_accumulation", ctypes.c_int32),
        ("adj_tf", DvrTensor),
    ]


ABI_VERSION = 3  # include/dvr_b200.h DVR_ABI_VERSION
NOALLOC_MAX_QUADS = 129 * 128 * 128  # include/dvr_b200.h DVR_NOALLOC_MAX_QUADS


class DvrCallOptions(ctypes.Structure):
    """include/dvr_b200.h DvrCallOptions (ABI version 3)."""
    _fields_ = [
        ("struct_size", ctypes.c_uint32),
        ("volume_layout", ctypes.c_int32),
        ("compat_stepsize_overwrite", ctypes.c_int32),
        ("sync", ctypes.c_int32),
        ("max_scratch_bytes", ctypes.c_int64),
        ("workspace", ctypes.c_void_p),
        ("forward_mode_tf", ctypes.c_int32),
        ("adjoint_march_zero_rays", ctypes.c_int32),
    ]


def make_call_options(volume_layout: int = -1, compat_stepsize_overwrite: bool = False, sync: bool = False,
                      max_scratch_bytes: int = 0, workspace=None, forward_mode_tf: int = 0,
                      adjoint_march_zero_rays: bool = False) -> DvrCallOptions:
    o = DvrCallOptions()
    o.struct_size = ctypes.sizeof(DvrCallOptions)
    o.volume_layout = int(volume_layout)
    o.compat_stepsize_overwrite = int(bool(compat_stepsize_overwrite))
    o.sync = int(bool(sync))
    o.max_scratch_bytes = int(max_scratch_bytes)
    o.workspace = workspace.handle if workspace is not None else None
    o.forward_mode_tf = int(forward_mode_tf)   # 1: the slot-major ceil(D/8)-pass kernel (A/B measurements)
    o.adjoint_march_zero_rays = int(bool(adjoint_march_zero_rays))  # 1: no skipping of rays with zero adjoint colour
    o._keep = workspace
    return o


def _check_tensor(t, what: str):
    """The reference's CHECK_CUDA / CHECK_DTYPE (renderer/pytorch_utils.h:5-19) for the ctypes path: the C-ABI sees
    raw pointers, so dtype and device are validated here. numpy arrays (the CPU oracles of the tests) only get the
    dtype check."""
    name = str(t.dtype).replace("torch.", "")
    if name not in ("float32", "float64", "int32"):
        raise TypeError(f"{what}: dtype {name} is not supported by the C-ABI (float32 / int32; float64 for the fp64 test oracle)")


def tensor_view(t, device=None, dtype=None, what: str = "tensor") -> DvrTensor:
    """DvrTensor over a torch tensor or numpy array (no copy; keep ``t`` alive). ``device`` (a torch device) and
    ``dtype`` (a name such as "float32") are enforced when given."""
    v = DvrTensor()
    if t is None:
        return v
    _check_tensor(t, what)
    if dtype is not None and str(t.dtype).replace("torch.", "") != dtype:
        raise TypeError(f"{what} must be {dtype}, but is {str(t.dtype).replace('torch.', '')}")
    if hasattr(t, "data_ptr"):  # torch
        if device is not None and t.device != device:
            raise ValueError(f"{what} must live on {device}, but is on {t.device}")
        v.ptr = t.data_ptr()
        shape, strides = tuple(t.shape), tuple(t.stride())
    else:  # numpy
        if device is not None:
            raise ValueError(f"{what} must be a CUDA tensor on {device}, but is a host array")
        v.ptr = t.ctypes.data
        shape = tuple(t.shape)
        strides = tuple(s // t.itemsize for s in t.strides)
    if len(shape) > DVR_MAX_DIMS:
        raise ValueError("too many dimensions")
    v.ndim = len(shape)
    for i, (n, s) in enumerate(zip(shape, strides)):
        v.size[i] = n
        v.stride[i] = s
    return v


def _is_tensor(x) -> bool:
    return hasattr(x, "shape") and not isinstance(x, (tuple, list))


class Inputs:
    """Python-side description of ``RendererInputsHost`` (renderer/renderer.h:20-64).

    Scalar-or-tensor fields (box_min, box_size, step_size, camera matrix) accept either kind,
    like the std::variant members of the reference struct.
    """

    def __init__(self, *, screen_size, volume, box_min, box_size, camera_mode, step_size, tf_mode,
                 tf, blend_mode=BLEND_BEER_LAMBERT, volume_filter_mode=FILTER_TRILINEAR,
                 camera_ray_start=None, camera_ray_dir=None, camera_matrix=None,
                 camera_fov_y_radians=0.0, blending_early_out=1.0 - 1e-5):
        self.screen_size = tuple(screen_size)  # (W, H)
        self.volume = volume
        self.box_min = box_min
        self.box_size = box_size
        self.camera_mode = camera_mode
        self.camera_ray_start = camera_ray_start
        self.camera_ray_dir = camera_ray_dir
        self.camera_matrix = camera_matrix
        self.camera_fov_y_radians = camera_fov_y_radians
        self.step_size = step_size
        self.tf_mode = tf_mode
        self.tf = tf
        self.blend_mode = blend_mode
        self.volume_filter_mode = volume_filter_mode
        self.blending_early_out = blending_early_out

    def to_c(self) -> DvrInputs:
        c = DvrInputs()
        c.screen_width, c.screen_height = self.screen_size
        c.volume_filter_mode = int(self.volume_filter_mode)
        c.volume = tensor_view(self.volume)
        if _is_tensor(self.box_min):
            c.box_min = tensor_view(self.box_min)
        else:
            c.box_min_value[:] = [float(v) for v in self.box_min]
        if _is_tensor(self.box_size):
            c.box_size = tensor_view(self.box_size)
        else:
            c.box_size_value[:] = [float(v) for v in self.box_size]
        c.camera_mode = int(self.camera_mode)
        if self.camera_mode == CAMERA_RAY_START_DIR:
            c.camera_ray_start = tensor_view(self.camera_ray_start)
            c.camera_ray_dir = tensor_view(self.camera_ray_dir)
        elif _is_tensor(self.camera_matrix) and len(self.camera_matrix.shape) == 3:
            c.camera_matrix = tensor_view(self.camera_matrix)
        else:  # 4x4 scalar variant (glm::mat4 in the reference), row-major nested sequence
            flat = [float(v) for row in self.camera_matrix for v in row]
            c.camera_matrix_value[:] = flat
        c.camera_fov_y_radians = float(self.camera_fov_y_radians)
        if _is_tensor(self.step_size):
            c.step_size = tensor_view(self.step_size)
        else:
            c.step_size_value = float(self.step_size)
        c.tf_mode = int(self.tf_mode)
        c.tf = tensor_view(self.tf)
        c.blend_mode = int(self.blend_mode)
        c.blending_early_out = float(self.blending_early_out)
        return c


def make_outputs(color, termination_index) -> DvrOutputs:
    o = DvrOutputs()
    o.color = tensor_view(color)
    o.termination_index = tensor_view(termination_index)
    return o


def make_fd_settings(D, d_stepsize=-1, d_ray_start=(-1, -1, -1), d_ray_dir=(-1, -1, -1), d_tf=None,
                     has_tf_derivatives=False) -> DvrForwardDifferencesSettings:
    s = DvrForwardDifferencesSettings()
    s.D = int(D)
    s.d_stepsize = int(d_stepsize)
    s.d_ray_start[:] = [int(v) for v in d_ray_start]
    s.d_ray_dir[:] = [int(v) for v in d_ray_dir]
    s.d_tf = tensor_view(d_tf)
    s.has_tf_derivatives = int(bool(has_tf_derivatives))
    return s


def make_adjoint_outputs(adj_volume=None, adj_step_size=None, adj_camera_ray_start=None,
                         adj_camera_ray_dir=None, adj_tf=None,
                         tf_delayed_accumulation=False) -> DvrAdjointOutputs:
    a = DvrAdjointOutputs()
    a.has_volume_derivatives = int(adj_volume is not None)
    a.adj_volume = tensor_view(adj_volume)
    a.has_stepsize_derivatives = int(adj_step_size is not None)
    a.adj_step_size = tensor_view(adj_step_size)
    a.has_camera_derivatives = int(adj_camera_ray_start is not None)
    a.adj_camera_ray_start = tensor_view(adj_camera_ray_start)
    a.adj_camera_ray_dir = tensor_view(adj_camera_ray_dir)
    a.has_tf_derivatives = int(adj_tf is not None)
    a.tf_delayed_accumulation = int(bool(tf_delayed_accumulation))
    a.adj_tf = tensor_view(adj_tf)
    return a


class DvrError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"dvr error {code}: {message}")
        self.code = code


class Library:
    """A loaded library exporting the dvr C-ABI under ``prefix`` (``dvr_`` for the product)."""

    def __init__(self, path: str, prefix: str = "dvr_"):
        self.path = path
        self.prefix = prefix
        self.lib = ctypes.CDLL(path)
        P = ctypes.POINTER
        self._sig("render_forward", [P(DvrInputs), P(DvrOutputs), ctypes.c_void_p])
        self._sig("render_forward_gradients",
                  [P(DvrInputs), P(DvrForwardDifferencesSettings), P(DvrOutputs), P(DvrTensor),
                   ctypes.c_void_p])
        self._sig("forward_variables_to_gradients",
                  [P(DvrTensor), P(DvrTensor), P(DvrForwardDifferencesSettings),
                   P(DvrAdjointOutputs), ctypes.c_void_p])
        self._sig("render_adjoint",
                  [P(DvrInputs), P(DvrOutputs), P(DvrTensor), P(DvrAdjointOutputs), ctypes.c_void_p])
        # callers either side of the marcher
        self._sig("generate_rays", [P(DvrTensor), ctypes.c_float, ctypes.c_int32, ctypes.c_int32,
                                    P(DvrTensor), P(DvrTensor), ctypes.c_void_p])
        self._sig("generate_rays_adjoint", [P(DvrTensor), ctypes.c_float, P(DvrTensor), P(DvrTensor),
                                            P(DvrTensor), ctypes.c_void_p])
        self._sig("compare_to_image", [P(DvrTensor)] * 5 + [ctypes.c_int32, ctypes.c_void_p])
        self._sig("preshade_volume", [P(DvrTensor), P(DvrTensor), ctypes.c_int32, P(DvrTensor), ctypes.c_void_p])
        self._sig("preshade_volume_adjoint", [P(DvrTensor), P(DvrTensor), ctypes.c_int32, P(DvrTensor), P(DvrTensor),
                                              ctypes.c_void_p])
        # renderer::Volume producers (row f3); the reference-built shim of the marcher does not export them
        self._sig("create_implicit_volume", [ctypes.c_int32, P(ctypes.c_float), P(DvrTensor), ctypes.c_void_p], True)
        self._sig("volume_from_raw", [ctypes.c_void_p, ctypes.c_int32, P(DvrTensor), ctypes.c_void_p], True)
        self._sig("volume_to_raw", [P(DvrTensor), ctypes.c_void_p, ctypes.c_void_p], True)
        self._sig("downsample_volume", [P(DvrTensor), P(DvrTensor), ctypes.c_void_p], True)
        f = getattr(self.lib, prefix + "last_error")
        f.restype = ctypes.c_char_p
        f.argtypes = []

    def _sig(self, name, argtypes, optional=False):
        f = getattr(self.lib, self.prefix + name, None) if optional else getattr(self.lib, self.prefix + name)
        if f is None:
            return
        f.restype = ctypes.c_int
        f.argtypes = argtypes

    @staticmethod
    def _stream(stream):
        return 0 if stream is None else stream

    def last_error(self) -> str:
        return getattr(self.lib, self.prefix + "last_error")().decode()

    def _check(self, rc):
        if rc != DVR_OK:
            raise DvrError(rc, self.last_error())

    # -- the four entry points -------------------------------------------------------------
    def render_forward(self, inputs: Inputs, color, termination_index, stream=None):
        ci, co = inputs.to_c(), make_outputs(color, termination_index)
        self._check(getattr(self.lib, self.prefix + "render_forward")(ci, co, self._stream(stream)))

    def render_forward_gradients(self, inputs: Inputs, settings: DvrForwardDifferencesSettings,
                                 color, termination_index, gradients_out, stream=None):
        ci, co, g = inputs.to_c(), make_outputs(color, termination_index), tensor_view(gradients_out)
        self._check(getattr(self.lib, self.prefix + "render_forward_gradients")(ci, settings, co, g,
                                                                               self._stream(stream)))

    def forward_variables_to_gradients(self, forward_variables, grad_output_color,
                                       settings: DvrForwardDifferencesSettings,
                                       adj: DvrAdjointOutputs, stream=None):
        fv, gc = tensor_view(forward_variables), tensor_view(grad_output_color)
        self._check(getattr(self.lib, self.prefix + "forward_variables_to_gradients")(
            fv, gc, settings, adj, self._stream(stream)))

    def render_adjoint(self, inputs: Inputs, color, termination_index, adj_color,
                       adj: DvrAdjointOutputs, stream=None):
        ci, co, ac = inputs.to_c(), make_outputs(color, termination_index), tensor_view(adj_color)
        self._check(getattr(self.lib, self.prefix + "render_adjoint")(ci, co, ac, adj, self._stream(stream)))

    # -- Camera::generateRays / Renderer::compareToImage --------------------------------------
    def generate_rays(self, viewport, fov_y_radians, ray_start, ray_dir, stream=None):
        """viewport B*3*3 -> ray_start, ray_dir B*H*W*3 (written)."""
        H, W = int(ray_start.shape[1]), int(ray_start.shape[2])
        self._check(getattr(self.lib, self.prefix + "generate_rays")(
            tensor_view(viewport), float(fov_y_radians), W, H, tensor_view(ray_start), tensor_view(ray_dir),
            self._stream(stream)))

    def generate_rays_adjoint(self, viewport, fov_y_radians, adj_ray_start, adj_ray_dir, adj_viewport,
                              stream=None):
        """adj_viewport (B*3*3) is overwritten with the per-view sums."""
        self._check(getattr(self.lib, self.prefix + "generate_rays_adjoint")(
            tensor_view(viewport), float(fov_y_radians), tensor_view(adj_ray_start), tensor_view(adj_ray_dir),
            tensor_view(adj_viewport), self._stream(stream)))

    def compare_to_image(self, color_input, gradients_input, color_reference, difference_out, gradients_out,
                         reduce: bool = False, stream=None):
        self._check(getattr(self.lib, self.prefix + "compare_to_image")(
            tensor_view(color_input), tensor_view(gradients_input), tensor_view(color_reference),
            tensor_view(difference_out), tensor_view(gradients_out), int(bool(reduce)), self._stream(stream)))

    # -- TFUtils::preshadeVolume --------------------------------------------------------------------
    def preshade_volume(self, density_volume, tf, tf_mode, color_out, stream=None):
        """density_volume 1*X*Y*Z, tf 1*R*C -> color_out 4*X*Y*Z (written)."""
        self._check(getattr(self.lib, self.prefix + "preshade_volume")(
            tensor_view(density_volume), tensor_view(tf), int(tf_mode), tensor_view(color_out), self._stream(stream)))

    def preshade_volume_adjoint(self, density_volume, tf, tf_mode, adj_color, adj_volume_out, stream=None):
        """adj_volume_out (1*X*Y*Z) is overwritten."""
        self._check(getattr(self.lib, self.prefix + "preshade_volume_adjoint")(
            tensor_view(density_volume), tensor_view(tf), int(tf_mode), tensor_view(adj_color),
            tensor_view(adj_volume_out), self._stream(stream)))

    # -- renderer::Volume producers -----------------------------------------------------------------
    def create_implicit_volume(self, equation, volume_out, params=None, stream=None):
        """volume_out 1*n*n*n (written); params = (fM, alpha, scale, a) or None for the reference's defaults."""
        p = (ctypes.c_float * 4)(*[float(v) for v in params]) if params is not None else None
        self._check(getattr(self.lib, self.prefix + "create_implicit_volume")(
            int(equation), p, tensor_view(volume_out), self._stream(stream)))

    def volume_from_raw(self, raw, dtype, volume_out, stream=None):
        """raw: flat u8 / u16 / f32 array or tensor of X*Y*Z elements, X fastest (.cvol body); dtype 0 / 1 / 2."""
        self._check(getattr(self.lib, self.prefix + "volume_from_raw")(
            _data_ptr(raw), int(dtype), tensor_view(volume_out), self._stream(stream)))

    def volume_to_raw(self, volume, raw_out, stream=None):
        self._check(getattr(self.lib, self.prefix + "volume_to_raw")(tensor_view(volume), _data_ptr(raw_out), self._stream(stream)))

    def downsample_volume(self, volume, volume_out, stream=None):
        self._check(getattr(self.lib, self.prefix + "downsample_volume")(
            tensor_view(volume), tensor_view(volume_out), self._stream(stream)))


def _data_ptr(a):
    return a.data_ptr() if hasattr(a, "data_ptr") else a.ctypes.data


_PRODUCT: Optional["ProductLibrary"] = None


class ProductLibrary(Library):
    """``libdvr_b200.so`` -- the CUDA product library (device pointers, CUDA streams)."""

    def __init__(self, path: str):
        super().__init__(path, "dvr_")
        L = self.lib
        L.dvr_set_option.restype = ctypes.c_int
        L.dvr_set_option.argtypes = [ctypes.c_char_p, ctypes.c_int32]
        L.dvr_get_option.restype = ctypes.c_int
        L.dvr_get_option.argtypes = [ctypes.c_char_p, ctypes.POINTER(ctypes.c_int32)]
        L.dvr_launch_count.restype = ctypes.c_int64
        L.dvr_launch_count.argtypes = []
        L.dvr_has_variant.restype = ctypes.c_int
        L.dvr_has_variant.argtypes = [ctypes.c_int32] * 4
        L.dvr_abi_version.restype = ctypes.c_int32
        L.dvr_abi_version.argtypes = []
        if L.dvr_abi_version() != ABI_VERSION:   # the struct mirrors above are those of ONE header version
            raise RuntimeError(f"{path}: ABI version {L.dvr_abi_version()}, this module mirrors version {ABI_VERSION} "
                               "(rebuild with __graft_entry__.build())")
        L.dvr_count_cell_visits.restype = ctypes.c_int
        L.dvr_count_cell_visits.argtypes = [ctypes.POINTER(DvrInputs), ctypes.POINTER(DvrTensor), ctypes.c_void_p,
                                            ctypes.c_void_p]
        L.dvr_check_inputs.restype = ctypes.c_int
        L.dvr_check_inputs.argtypes = [ctypes.POINTER(DvrInputs)] + [ctypes.POINTER(ctypes.c_int32)] * 4

        P = ctypes.POINTER
        L.dvr_render_forward_ex.restype = ctypes.c_int
        L.dvr_render_forward_ex.argtypes = [P(DvrInputs), P(DvrOutputs), P(DvrCallOptions), ctypes.c_void_p]
        L.dvr_render_forward_gradients_ex.restype = ctypes.c_int
        L.dvr_render_forward_gradients_ex.argtypes = [P(DvrInputs), P(DvrForwardDifferencesSettings), P(DvrOutputs),
                                                      P(DvrTensor), P(DvrCallOptions), ctypes.c_void_p]
        L.dvr_render_adjoint_ex.restype = ctypes.c_int
        L.dvr_render_adjoint_ex.argtypes = [P(DvrInputs), P(DvrOutputs), P(DvrTensor), P(DvrAdjointOutputs),
                                            P(DvrCallOptions), ctypes.c_void_p]
        L.dvr_call_options_init.restype = None
        L.dvr_call_options_init.argtypes = [P(DvrCallOptions)]
        L.dvr_workspace_create.restype = ctypes.c_void_p
        L.dvr_workspace_create.argtypes = []
        L.dvr_workspace_destroy.restype = None
        L.dvr_workspace_destroy.argtypes = [ctypes.c_void_p]
        L.dvr_workspace_bytes.restype = ctypes.c_int64
        L.dvr_workspace_bytes.argtypes = [ctypes.c_void_p]
        L.dvr_workspace_invalidate.restype = ctypes.c_int
        L.dvr_workspace_invalidate.argtypes = [ctypes.c_void_p]
        for f in (L.dvr_workspace_set_volume, L.dvr_workspace_begin_volume_gradient,
                  L.dvr_workspace_end_volume_gradient):
            f.restype = ctypes.c_int
            f.argtypes = [ctypes.c_void_p, P(DvrTensor), ctypes.c_void_p]
        L.dvr_render_forward_noalloc.restype = ctypes.c_int
        L.dvr_render_forward_noalloc.argtypes = [P(DvrInputs), P(DvrOutputs), ctypes.c_void_p]
        L.dvr_trim_scratch.restype = ctypes.c_int
        L.dvr_trim_scratch.argtypes = []
        L.dvr_get_stat.restype = ctypes.c_int
        L.dvr_get_stat.argtypes = [ctypes.c_char_p, P(ctypes.c_int64)]
        L.dvr_has_variant_pass.restype = ctypes.c_int
        L.dvr_has_variant_pass.argtypes = [ctypes.c_int32] * 5

    # -- validation the C-ABI cannot do (it sees raw pointers): device, dtype; default stream ------------------
    @staticmethod
    def _stream(stream):
        if stream is not None:
            return stream
        import torch
        return torch.cuda.current_stream().cuda_stream

    @staticmethod
    def _validate(inputs: Inputs, extra):
        """Every tensor must be a CUDA tensor on the volume's device with the dtype the kernels read (the
        reference's CHECK_CUDA / CHECK_DTYPE, renderer/renderer.cpp:49-263)."""
        vol = inputs.volume
        if not hasattr(vol, "is_cuda") or not vol.is_cuda:
            raise ValueError("volume must be a CUDA tensor: the product has no CPU path")
        dev = vol.device
        named = [("volume", vol, "float32"), ("tf", inputs.tf, "float32"),
                 ("boxMin", inputs.box_min, "float32"), ("boxSize", inputs.box_size, "float32"),
                 ("cameraRayStart", inputs.camera_ray_start, "float32"), ("cameraRayDir", inputs.camera_ray_dir, "float32"),
                 ("cameraMatrix", inputs.camera_matrix, "float32"), ("stepSize", inputs.step_size, "float32")] + list(extra)
        for name, t, dtype in named:
            if t is None or not _is_tensor(t):
                continue
            tensor_view(t, device=dev, dtype=dtype, what=name)

    def render_forward(self, inputs: Inputs, color, termination_index, stream=None, options: DvrCallOptions = None):
        self._validate(inputs, [("color", color, "float32"), ("terminationIndex", termination_index, "int32")])
        ci, co = inputs.to_c(), make_outputs(color, termination_index)
        self._check(self.lib.dvr_render_forward_ex(ci, co, options, self._stream(stream)))

    def render_forward_noalloc(self, inputs: Inputs, color, termination_index, stream=None):
        """dvr_render_forward_noalloc: the forward pass with its QUAD copy in the library's static device array
        (NOALLOC_MAX_QUADS entries); allocates nothing; calls that overlap on the GPU race on that array."""
        self._validate(inputs, [("color", color, "float32"), ("terminationIndex", termination_index, "int32")])
        ci, co = inputs.to_c(), make_outputs(color, termination_index)
        self._check(self.lib.dvr_render_forward_noalloc(ci, co, self._stream(stream)))

    def render_forward_gradients(self, inputs: Inputs, settings: DvrForwardDifferencesSettings, color,
                                 termination_index, gradients_out, stream=None, options: DvrCallOptions = None):
        self._validate(inputs, [("color", color, "float32"), ("terminationIndex", termination_index, "int32"),
                                ("gradientsOut", gradients_out, "float32")])
        ci, co, g = inputs.to_c(), make_outputs(color, termination_index), tensor_view(gradients_out)
        self._check(self.lib.dvr_render_forward_gradients_ex(ci, settings, co, g, options, self._stream(stream)))

    def render_adjoint(self, inputs: Inputs, color, termination_index, adj_color, adj: DvrAdjointOutputs,
                       stream=None, options: DvrCallOptions = None):
        self._validate(inputs, [("output_color", color, "float32"), ("terminationIndex", termination_index, "int32"),
                                ("adj_color", adj_color, "float32")])
        ci, co, ac = inputs.to_c(), make_outputs(color, termination_index), tensor_view(adj_color)
        self._check(self.lib.dvr_render_adjoint_ex(ci, co, ac, adj, options, self._stream(stream)))

    def workspace(self) -> "Workspace":
        return Workspace(self)

    def trim_scratch(self):
        self._check(self.lib.dvr_trim_scratch())

    def stat(self, name: str) -> int:
        v = ctypes.c_int64()
        self._check(self.lib.dvr_get_stat(name.encode(), ctypes.byref(v)))
        return int(v.value)

    def has_variant_pass(self, pass_id, filter_mode, camera_mode, tf_mode, blend_mode) -> bool:
        return bool(self.lib.dvr_has_variant_pass(pass_id, filter_mode, camera_mode, tf_mode, blend_mode))

    def set_option(self, name: str, value: int):
        self._check(self.lib.dvr_set_option(name.encode(), int(value)))

    def get_option(self, name: str) -> int:
        v = ctypes.c_int32()
        self._check(self.lib.dvr_get_option(name.encode(), ctypes.byref(v)))
        return v.value

    def launch_count(self) -> int:
        return int(self.lib.dvr_launch_count())

    def has_variant(self, filter_mode, camera_mode, tf_mode, blend_mode) -> bool:
        return bool(self.lib.dvr_has_variant(filter_mode, camera_mode, tf_mode, blend_mode))

    def count_cell_visits(self, inputs: Inputs, termination_index, counts, stream=None):
        """counts: device tensor of 2 int64 {N_samples, N_cells}, accumulated."""
        self._validate(inputs, [("terminationIndex", termination_index, "int32")])
        ci, t = inputs.to_c(), tensor_view(termination_index)
        self._check(self.lib.dvr_count_cell_visits(ci, t, counts.data_ptr(), self._stream(stream)))

    def check_inputs(self, inputs: Inputs):
        B, X, Y, Z = (ctypes.c_int32() for _ in range(4))
        ci = inputs.to_c()
        self._check(self.lib.dvr_check_inputs(ci, ctypes.byref(B), ctypes.byref(X), ctypes.byref(Y),
                                              ctypes.byref(Z)))
        return B.value, X.value, Y.value, Z.value


PASS_FORWARD, PASS_FORWARD_GRADIENTS, PASS_ADJOINT = 0, 1, 2


class Workspace:
    """Persistent QUAD buffers of the library (include/dvr_b200.h DvrWorkspace): the volume copy is built once per
    volume version, the gradient scratch is zeroed / folded once per accumulation scope."""

    def __init__(self, lib: ProductLibrary):
        self.lib = lib
        self.handle = lib.lib.dvr_workspace_create()
        if not self.handle:
            raise DvrError(DVR_ERR_CUDA, lib.last_error())

    def close(self):
        if self.handle:
            self.lib.lib.dvr_workspace_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_volume(self, volume, stream=None):
        self.lib._check(self.lib.lib.dvr_workspace_set_volume(self.handle, tensor_view(volume, dtype="float32", what="volume"),
                                                              self.lib._stream(stream)))

    def invalidate(self):
        self.lib._check(self.lib.lib.dvr_workspace_invalidate(self.handle))

    def begin_volume_gradient(self, adj_volume, stream=None):
        self.lib._check(self.lib.lib.dvr_workspace_begin_volume_gradient(
            self.handle, tensor_view(adj_volume, dtype="float32", what="adj_volume"), self.lib._stream(stream)))

    def end_volume_gradient(self, adj_volume, stream=None):
        self.lib._check(self.lib.lib.dvr_workspace_end_volume_gradient(
            self.handle, tensor_view(adj_volume, dtype="float32", what="adj_volume"), self.lib._stream(stream)))

    def bytes(self) -> int:
        return int(self.lib.lib.dvr_workspace_bytes(self.handle))


def library_path() -> str:
    """The product library; ``DVR_LIB_PATH`` selects an A/B build of the same sources (csrc/Makefile, measurements only)."""
    return os.environ.get("DVR_LIB_PATH") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libdvr_b200.so")


def load() -> ProductLibrary:
    """Loads the CUDA library; raises (loudly) if it has not been built -- no fallback."""
    global _PRODUCT
    if _PRODUCT is None:
        path = library_path()
        if not os.path.exists(path):
            raise RuntimeError(
                f"{path} not found: the CUDA extension is not built (run "
                "`python -c 'import __graft_entry__ as g; g.build()'`); there is no CPU fallback")
        _PRODUCT = ProductLibrary(path)
    return _PRODUCT
