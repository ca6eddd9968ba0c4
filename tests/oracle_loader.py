"""Test-only access to the CPU oracles (oracle/).  The product never imports this.

* ``port(real)``  -> oracle/_build/libdvr_oracle.so, the CPU restatement (built on demand with
  the oracle Makefile; needs only g++).
* ``ref(real)``   -> oracle/_ref/libdvr_ref{32,64}.so, the REFERENCE's own host kernels compiled
  from /root/reference (prebuilt in the build container; travels to the GPU box as a binary).
  Returns None when the binary is absent and cannot be built (no /root/reference).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from diffdvr_b200 import capi  # noqa: E402  (struct definitions only)

ORACLE_DIR = os.path.join(ROOT, "oracle")
_cache = {}


def _make(target: str) -> bool:
    try:
        subprocess.run(["make", "-C", ORACLE_DIR, "-j8", target], check=True,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
        return True
    except (subprocess.CalledProcessError, FileNotFoundError):
        return False


class OracleLibrary(capi.Library):
    def __init__(self, path, prefix, real_bits):
        super().__init__(path, prefix)
        self.real_bits = real_bits
        f = getattr(self.lib, prefix + "nonconvex_density", None)
        if f is not None:
            f.restype = ctypes.c_int
            f.argtypes = [ctypes.c_double, ctypes.POINTER(ctypes.c_double)]

        f = getattr(self.lib, prefix + "face_distance", None)
        if f is not None:
            f.restype = ctypes.c_int
            f.argtypes = [ctypes.POINTER(capi.DvrInputs), ctypes.POINTER(capi.DvrTensor),
                          ctypes.POINTER(capi.DvrTensor)]

    def face_distance(self, inputs, termination_index, out):
        """Per-ray minimum distance (voxels) of the samples to a trilinear cell face (port only)."""
        ci, t, o = inputs.to_c(), capi.tensor_view(termination_index), capi.tensor_view(out)
        self._check(getattr(self.lib, self.prefix + "face_distance")(ci, t, o))

    def nonconvex_density(self, d1: float):
        out = (ctypes.c_double * 4)()
        self._check(getattr(self.lib, self.prefix + "nonconvex_density")(d1, out))
        return list(out)


def port(real_bits: int = 32) -> OracleLibrary:
    key = ("port", real_bits)
    if key not in _cache:
        path = os.path.join(ORACLE_DIR, "_build", "libdvr_oracle.so")
        src = os.path.join(ORACLE_DIR, "dvr_oracle.cpp")
        if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(src):
            if not _make("port"):
                raise RuntimeError("cannot build oracle/_build/libdvr_oracle.so (g++ missing?)")
        _cache[key] = OracleLibrary(path, f"dvro{real_bits}_", real_bits)
    return _cache[key]


def port_set_option(name: str, value: int):
    lib = port(32).lib
    lib.dvro_set_option.restype = ctypes.c_int
    lib.dvro_set_option.argtypes = [ctypes.c_char_p, ctypes.c_int32]
    assert lib.dvro_set_option(name.encode(), value) == 0


def ref(real_bits: int = 32):
    key = ("ref", real_bits)
    if key not in _cache:
        path = os.path.join(ORACLE_DIR, "_ref", f"libdvr_ref{real_bits}.so")
        if not os.path.exists(path) and os.path.isdir("/root/reference/renderer"):
            _make(f"_ref/libdvr_ref{real_bits}.so")
        _cache[key] = OracleLibrary(path, f"dvrref{real_bits}_", real_bits) if os.path.exists(path) else None
    return _cache[key]
