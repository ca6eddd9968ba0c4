"""ctypes access to tests/_build/libdvr_facade_shim.so: the stand-in for the reference's host layer that drives
renderer::RendererCuda (diffdvr_b200/csrc/host/renderer_cuda_facade.h) with the reference's packed accessors."""
from __future__ import annotations

import ctypes
import os
import subprocess

from diffdvr_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_P = ctypes.POINTER


def load():
    path = os.path.join(ROOT, "tests", "_build", "libdvr_facade_shim.so")
    if not os.path.exists(path):  # normally built by __graft_entry__.build()
        subprocess.run(["make", "-C", os.path.join(ROOT, "diffdvr_b200", "csrc")], check=True, capture_output=True)
        subprocess.run(["make", "-C", os.path.join(ROOT, "tests", "csrc"), "all"], check=True, capture_output=True)
    lib = ctypes.CDLL(path)
    lib.shim_last_error.restype = ctypes.c_char_p
    lib.shim_kernel_name.argtypes = [ctypes.c_int] * 11 + [ctypes.c_char_p, ctypes.c_int]
    lib.shim_inputs_roundtrip.argtypes = [_P(capi.DvrInputs), _P(capi.DvrInputs)]
    lib.shim_adjoint_outputs_roundtrip.argtypes = [_P(capi.DvrAdjointOutputs), _P(capi.DvrAdjointOutputs)]
    lib.shim_settings_roundtrip.argtypes = [_P(capi.DvrForwardDifferencesSettings), _P(capi.DvrForwardDifferencesSettings)]
    lib.shim_render_forward.argtypes = [_P(capi.DvrInputs), _P(capi.DvrOutputs), ctypes.c_void_p]
    lib.shim_render_forward_gradients.argtypes = [_P(capi.DvrInputs), _P(capi.DvrForwardDifferencesSettings),
                                                  _P(capi.DvrOutputs), _P(capi.DvrTensor), ctypes.c_void_p]
    lib.shim_forward_variables_to_gradients.argtypes = [_P(capi.DvrTensor), _P(capi.DvrTensor),
                                                        _P(capi.DvrForwardDifferencesSettings),
                                                        _P(capi.DvrAdjointOutputs), ctypes.c_void_p]
    lib.shim_render_adjoint.argtypes = [_P(capi.DvrInputs), _P(capi.DvrOutputs), _P(capi.DvrTensor),
                                        _P(capi.DvrAdjointOutputs), ctypes.c_void_p]
    return lib


def kernel_name(lib, pass_id, f, c, t, b, D=0, has_step=False, has_cam=False, has_tf=False, delayed=False, has_vol=False):
    buf = ctypes.create_string_buffer(512)
    lib.shim_kernel_name(pass_id, f, c, t, b, D, int(has_step), int(has_cam), int(has_tf), int(delayed), int(has_vol), buf, 512)
    return buf.value.decode()


def tensor_fields(t: capi.DvrTensor):
    return (t.ptr or 0, t.ndim, tuple(t.size[: t.ndim]), tuple(t.stride[: t.ndim]))
