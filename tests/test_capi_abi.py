"""CPU tests of the drop-in boundary: libdvr_b200.so loads without a GPU, exports every symbol that
include/dvr_b200.h declares, its struct layout matches the ctypes mirror, and the host-side
validation (the part of Renderer::checkInputs that needs no device) reports the reference's errors.
No compute entry point is called with valid inputs here (there is no GPU)."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import scenes
from diffdvr_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "dvr_b200.h")


def declared_symbols():
    text = open(HEADER).read()
    return sorted(set(re.findall(r"DVR_API\s+[\w\s\*]+?\b(dvr_\w+)\s*\(", text)))


def test_header_declares_the_expected_entry_points():
    syms = declared_symbols()
    for s in ("dvr_render_forward", "dvr_render_forward_gradients", "dvr_forward_variables_to_gradients",
              "dvr_render_adjoint", "dvr_last_error", "dvr_set_option", "dvr_has_variant"):
        assert s in syms


def test_library_loads_without_gpu_and_exports_every_declared_symbol():
    lib = capi.load()   # raises if the .so is missing: there is no fallback
    for s in declared_symbols():
        assert hasattr(lib.lib, s), f"{s} declared in include/dvr_b200.h but not exported"
    assert lib.lib.dvr_abi_version() == 3
    # nothing else with default visibility leaks out of the library
    out = subprocess.run(["nm", "-D", "--defined-only", capi.library_path()], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    assert exported == set(declared_symbols()), exported ^ set(declared_symbols())


def test_struct_layout_matches_the_c_header(tmp_path):
    src = tmp_path / "layout.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "dvr_b200.h"\n'
                   "int main(void){\n"
                   'printf("%zu %zu %zu %zu %zu %zu %zu\\n", sizeof(DvrTensor), sizeof(DvrInputs), sizeof(DvrOutputs),'
                   " sizeof(DvrForwardDifferencesSettings), sizeof(DvrAdjointOutputs), sizeof(DvrCallOptions),"
                   " offsetof(DvrCallOptions, workspace));\n"
                   'printf("%zu %zu %zu %zu %zu\\n", offsetof(DvrInputs, tf), offsetof(DvrInputs, step_size_value),'
                   " offsetof(DvrInputs, blending_early_out), offsetof(DvrAdjointOutputs, adj_tf),"
                   " offsetof(DvrForwardDifferencesSettings, d_tf));\nreturn 0;}\n")
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    a, b = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines()
    assert [int(v) for v in a.split()] == [ctypes.sizeof(c) for c in (
        capi.DvrTensor, capi.DvrInputs, capi.DvrOutputs, capi.DvrForwardDifferencesSettings, capi.DvrAdjointOutputs,
        capi.DvrCallOptions)] + [capi.DvrCallOptions.workspace.offset]
    # dvr_call_options_init writes the documented defaults, and the ctypes helper builds the same struct
    o = capi.DvrCallOptions()
    capi.load().lib.dvr_call_options_init(ctypes.byref(o))
    d = capi.make_call_options()
    assert (o.struct_size, o.volume_layout, o.compat_stepsize_overwrite, o.sync, o.max_scratch_bytes, o.workspace) == \
           (d.struct_size, d.volume_layout, d.compat_stepsize_overwrite, d.sync, d.max_scratch_bytes, d.workspace) == \
           (ctypes.sizeof(capi.DvrCallOptions), -1, 0, 0, 0, None)
    assert [int(v) for v in b.split()] == [
        capi.DvrInputs.tf.offset, capi.DvrInputs.step_size_value.offset, capi.DvrInputs.blending_early_out.offset,
        capi.DvrAdjointOutputs.adj_tf.offset, capi.DvrForwardDifferencesSettings.d_tf.offset]


def test_has_variant_covers_the_reference_instantiation_set():
    """unittests/testCudaCompilation.cpp:109-287 enumerates filter x camera x tf x blend."""
    lib = capi.load()
    for f in (capi.FILTER_NEAREST, capi.FILTER_TRILINEAR):
        for c in (0, 1, 2):
            for t in (capi.TF_IDENTITY, capi.TF_TEXTURE, capi.TF_LINEAR, capi.TF_GAUSSIAN, capi.TF_GAUSSIAN_LOG):
                for b in (0, 1):
                    assert lib.has_variant(f, c, t, b)
    for c in (0, 1, 2):   # pre-shaded volumes: filter and TF only together
        for b in (0, 1):
            assert lib.has_variant(capi.FILTER_PRESHADED, c, capi.TF_PRESHADED, b)
    assert not lib.has_variant(capi.FILTER_PRESHADED, 0, capi.TF_TEXTURE, 0)
    assert not lib.has_variant(capi.FILTER_TRILINEAR, 0, capi.TF_PRESHADED, 0)
    assert not lib.has_variant(capi.FILTER_NETWORK, 0, capi.TF_TEXTURE, 0)
    assert not lib.has_variant(capi.FILTER_TRILINEAR, 7, capi.TF_TEXTURE, 0)


def test_options_round_trip_and_unknown_option():
    lib = capi.load()
    for name in ("sync_mode",):
        old = lib.get_option(name)
        lib.set_option(name, 1)
        assert lib.get_option(name) == 1
        lib.set_option(name, old)
    with pytest.raises(capi.DvrError):
        lib.set_option("no_such_option", 1)
    # options that change results are per call (DvrCallOptions), never process state
    for name in ("compat_stepsize_overwrite", "volume_layout"):
        with pytest.raises(capi.DvrError, match="per call"):
            lib.set_option(name, 1)


def test_check_inputs_shapes_and_reference_error_messages():
    """dvr_check_inputs == Renderer::checkInputs (renderer/renderer.cpp:49-263): host pointers are
    fine here, nothing is dereferenced."""
    lib = capi.load()
    sc = scenes.make_scene(n=10, W=7, H=5, B=3, tf_mode=capi.TF_LINEAR, R=4)
    assert lib.check_inputs(sc.inputs()) == (3, 10, 10, 10)
    bad = scenes.make_scene(n=10, W=7, H=5, B=1, tf_mode=capi.TF_LINEAR, R=4)
    bad.tf = bad.tf[:, :, :4].copy()
    with pytest.raises(capi.DvrError) as e:
        lib.check_inputs(bad.inputs())
    assert e.value.code == capi.DVR_ERR_INVALID_ARGUMENT and "tf must have 5 entries at dimension 2" in str(e.value)
    bad = scenes.make_scene(n=10, W=7, H=5, B=2, camera_mode=capi.CAMERA_RAY_START_DIR)
    bad.ray_dir = np.concatenate([bad.ray_dir] * 2)[:3]  # batch 3 vs 2
    with pytest.raises(capi.DvrError) as e:
        lib.check_inputs(bad.inputs())
    assert "same batch size" in str(e.value)
    pre = scenes.make_scene(n=10, W=7, H=5, B=1)
    pre.filter_mode = capi.FILTER_PRESHADED   # without TFPreshaded
    with pytest.raises(capi.DvrError) as e:
        lib.check_inputs(pre.inputs())
    assert "pre-shading" in str(e.value)


def test_null_arguments_are_rejected_not_dereferenced():
    lib = capi.load()
    assert lib.lib.dvr_render_forward(None, None, None) == capi.DVR_ERR_INVALID_ARGUMENT
    assert lib.lib.dvr_render_adjoint(None, None, None, None, None) == capi.DVR_ERR_INVALID_ARGUMENT
    assert b"NULL" in lib.lib.dvr_last_error()
