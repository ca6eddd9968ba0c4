"""B200-native differentiable volume ray marcher -- drop-in for the hot path of DiffDVR.

Only what the path needs lives here:
  csrc/          hand-written CUDA (sm_100a) + the C-ABI (include/dvr_b200.h) -> libdvr_b200.so
  csrc/host/     C++ host layer: renderer::Renderer (the reference's renderer.h API) + the pybind11
                 module ``pyrenderer`` (the reference's Python API) -> pyrenderer*.so
  capi.py        ctypes mirror of the C-ABI (host plumbing; no CPU fallback)
  synthetic.py   deterministic synthetic inputs (volumes, cameras, TFs) for tests and bench
  multigpu.py    view sharding + NCCL all-reduce of the volume / TF gradients
"""
import importlib
import os
import sys

from . import capi  # noqa: F401

__all__ = ["capi", "load_pyrenderer"]


def load_pyrenderer():
    """Imports the in-tree ``pyrenderer`` extension module (same name as the reference's module, so
    ``import pyrenderer`` in the reference's Python code resolves to it afterwards). Raises if it
    has not been built -- there is no fallback."""
    if "pyrenderer" in sys.modules:
        return sys.modules["pyrenderer"]
    import torch  # noqa: F401  (libtorch must be loaded before the extension)
    here = os.path.dirname(os.path.abspath(__file__))
    if here not in sys.path:
        sys.path.insert(0, here)
    try:
        return importlib.import_module("pyrenderer")
    except ImportError as e:
        raise RuntimeError(
            "diffdvr_b200/pyrenderer*.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
            f"there is no fallback: {e}") from e
