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
            # the reference is here: a missing build is an error, not a reason to fall back to the port silently
            if not _make(f"_ref/libdvr_ref{real_bits}.so"):
                raise RuntimeError(f"/root/reference is present but oracle/_ref/libdvr_ref{real_bits}.so does not build "
                                   "(run `make -C oracle ref` to see why)")
        _cache[key] = OracleLibrary(path, f"dvrref{real_bits}_", real_bits) if os.path.exists(path) else None
    return _cache[key]


class RefVolume:
    """oracle/_ref/libdvr_refvol.so: the reference's own renderer::Volume (volume.cpp, volume_implicit.cpp and its
    vendored LZ4, compiled unmodified; oracle/ref_shim_volume.cpp). Arrays are float32 [X][Y][Z]."""

    def __init__(self, path):
        self.lib = ctypes.CDLL(path)
        self.lib.refvol_last_error.restype = ctypes.c_char_p

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError(self.lib.refvol_last_error().decode())

    def create_implicit(self, equation, resolution, **params):
        import numpy as np
        out = np.zeros((resolution,) * 3, np.float32)
        keys = (ctypes.c_char_p * len(params))(*[k.encode() for k in params])
        vals = (ctypes.c_float * len(params))(*[float(v) for v in params.values()])
        self._check(self.lib.refvol_create_implicit(int(equation), int(resolution), len(params), keys, vals,
                                                    out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def save(self, filename, data, world=(1.0, 1.0, 1.0), compression=0):
        import numpy as np
        data = np.ascontiguousarray(data, np.float32)
        X, Y, Z = data.shape
        self._check(self.lib.refvol_save(filename.encode(), data.ctypes.data_as(ctypes.c_void_p), X, Y, Z,
                                         ctypes.c_double(world[0]), ctypes.c_double(world[1]),
                                         ctypes.c_double(world[2]), int(compression)))

    def load(self, filename):
        import numpy as np
        X, Y, Z = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        world = (ctypes.c_double * 3)()
        self._check(self.lib.refvol_load(filename.encode(), ctypes.byref(X), ctypes.byref(Y), ctypes.byref(Z), world,
                                         None))
        out = np.zeros((X.value, Y.value, Z.value), np.float32)
        self._check(self.lib.refvol_load(filename.encode(), ctypes.byref(X), ctypes.byref(Y), ctypes.byref(Z), world,
                                         out.ctypes.data_as(ctypes.c_void_p)))
        return out, tuple(world)


def refvol():
    """None when the binary is absent and cannot be built (no /root/reference, or no libtorch to link against)."""
    if "refvol" not in _cache:
        path = os.path.join(ORACLE_DIR, "_ref", "libdvr_refvol.so")
        if not os.path.exists(path) and os.path.isdir("/root/reference/renderer"):
            _make("_ref/libdvr_refvol.so")
        try:
            _cache["refvol"] = RefVolume(path) if os.path.exists(path) else None
        except OSError:
            _cache["refvol"] = None
    return _cache["refvol"]
