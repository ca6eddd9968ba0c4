"""The reference's device-function unit tests (unittests/testAdjointBlending.cpp, testAdjointTF.cpp,
testAdjointInterpolation.cpp over unittests/check_adjoint.h) restated over a uniform function-level interface that
three libraries export: the reference's own functions (oracle/ref_shim_fn.cpp), the CPU port (oracle/dvr_oracle_fn.h)
and the product's CUDA device functions (tests/csrc/devfn_harness.cu).

Inputs follow the unit tests: seed 42, the same distributions and counts (std::default_random_engine's stream itself
is not reproduced: numpy.random.RandomState(42))."""
from __future__ import annotations

import ctypes
import os

import numpy as np

from diffdvr_b200 import capi

SEED = 42
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TF_CHANNELS = {capi.TF_IDENTITY: 2, capi.TF_TEXTURE: 4, capi.TF_LINEAR: 5, capi.TF_GAUSSIAN: 6, capi.TF_GAUSSIAN_LOG: 6}


class HostFns:
    """fn_* of an oracle library (host pointers); real = float32 or float64 by the library's prefix."""

    def __init__(self, oracle_library):
        self.lib, self.prefix = oracle_library.lib, oracle_library.prefix
        self.dt = np.float32 if oracle_library.real_bits == 32 else np.float64

    def _call(self, name, *args):
        conv = []
        for a in args:
            if isinstance(a, np.ndarray):
                assert a.dtype == self.dt and a.flags.c_contiguous
                conv.append(a.ctypes.data_as(ctypes.c_void_p))
            else:
                conv.append(ctypes.c_int32(int(a)))
        rc = getattr(self.lib, self.prefix + name)(*conv)
        assert rc == 0, getattr(self.lib, self.prefix + "last_error")()

    def arr(self, a):
        return np.ascontiguousarray(a, self.dt)

    def blend(self, mode, x):
        x = self.arr(x)
        out = np.zeros((len(x), 4), self.dt)
        self._call("fn_blend", mode, len(x), x, out)
        return out

    def blend_adjoint(self, mode, x, e, g):
        x, e, g = self.arr(x), self.arr(e), self.arr(g)
        z, acc_in = np.zeros((len(x), 9), self.dt), np.zeros((len(x), 4), self.dt)
        self._call("fn_blend_adjoint", mode, len(x), x, e, g, z, acc_in)
        return z, acc_in

    def tf(self, mode, table, density):
        table, density = self.arr(table), self.arr(density)
        out = np.zeros((len(density), 4), self.dt)
        self._call("fn_tf", mode, table.shape[0], table.shape[1], table, len(density), density, out)
        return out

    def tf_adjoint(self, mode, table, density, g, acc=None):
        table, density, g = self.arr(table), self.arr(density), self.arr(g)
        adj_d = np.zeros(len(density), self.dt)
        adj_tf = np.zeros((len(density),) + table.shape, self.dt)
        self._call("fn_tf_adjoint", mode, table.shape[0], table.shape[1], table, len(density), density, g, adj_d, adj_tf)
        return adj_d, adj_tf

    def interp(self, volume, pos, quad=False):
        volume, pos = self.arr(volume), self.arr(pos)
        out = np.zeros(len(pos), self.dt)
        self._call("fn_interp", *volume.shape, volume, len(pos), pos, out)
        return out

    def interp_adjoint(self, volume, pos, g, quad=False):
        volume, pos, g = self.arr(volume), self.arr(pos), self.arr(g)
        adj_pos = np.zeros((len(pos), 3), self.dt)
        adj_vol = np.zeros((len(pos),) + volume.shape, self.dt)
        self._call("fn_interp_adjoint", *volume.shape, volume, len(pos), pos, g, adj_pos, adj_vol)
        return adj_pos, adj_vol


def build_quads(volume):
    """numpy statement of the QUAD copy: Q[x][y][z] = (v[y][z], v[y][z'], v[y'][z], v[y'][z']) of
    slice min(x, X - 1) for x in [0, X] (X + 1 slices), with y' = min(y + 1, Y - 1), z' = min(z + 1, Z - 1)."""
    X, Y, Z = volume.shape
    yp, zp = np.minimum(np.arange(Y) + 1, Y - 1), np.minimum(np.arange(Z) + 1, Z - 1)
    q = np.stack([volume, volume[:, :, zp], volume[:, yp, :], volume[:, yp][:, :, zp]], axis=-1)
    q = np.concatenate([q, q[-1:]], axis=0)
    return np.ascontiguousarray(q, np.float32)


def fold_quads(g):
    """inverse for gradients: scatter the four components of every quad back onto the voxels they alias (slice X of the
    X + 1 slices lands on slice X - 1)"""
    g = np.concatenate([g[:, :-2], g[:, -2:-1] + g[:, -1:]], axis=1)
    n, X, Y, Z, _ = g.shape
    yp, zp = np.minimum(np.arange(Y) + 1, Y - 1), np.minimum(np.arange(Z) + 1, Z - 1)
    out = np.zeros((n, X, Y, Z), np.float64)
    out += g[..., 0]
    np.add.at(out, (slice(None), slice(None), slice(None), zp), g[..., 1])
    np.add.at(out, (slice(None), slice(None), yp, slice(None)), g[..., 2])
    tmp = np.zeros_like(out)
    np.add.at(tmp, (slice(None), slice(None), slice(None), zp), g[..., 3])
    np.add.at(out, (slice(None), slice(None), yp, slice(None)), tmp)
    return out


class DeviceFns:
    """devfn_* of tests/_build/libdvr_devfn.so: the product's CUDA device functions (float32, device pointers)."""

    dt = np.float32

    def __init__(self):
        import torch
        self.torch = torch
        path = os.path.join(ROOT, "tests", "_build", "libdvr_devfn.so")
        if not os.path.exists(path):  # normally built by __graft_entry__.build(); nvcc is on the GPU box too
            import subprocess
            subprocess.run(["make", "-C", os.path.join(ROOT, "tests", "csrc")], check=True, capture_output=True)
        self.lib = ctypes.CDLL(path)

    def _call(self, name, *args):
        conv = []
        for a in args:
            if a is None:
                conv.append(ctypes.c_void_p(0))
            elif hasattr(a, "data_ptr"):
                conv.append(ctypes.c_void_p(a.data_ptr()))
            else:
                conv.append(ctypes.c_int32(int(a)))
        rc = getattr(self.lib, name)(*conv)
        assert rc == 0, f"{name} failed with {rc}"

    def dev(self, a):
        return self.torch.from_numpy(np.ascontiguousarray(a, np.float32)).cuda()

    def zeros(self, *shape):
        return self.torch.zeros(*shape, device="cuda")

    def blend(self, mode, x):
        x = self.dev(x)
        out = self.zeros(len(x), 4)
        self._call("devfn_blend", mode, len(x), x, out)
        return out.cpu().numpy()

    def blend_adjoint(self, mode, x, e, g):
        x, e, g = self.dev(x), self.dev(e), self.dev(g)
        z, acc_in = self.zeros(len(x), 9), self.zeros(len(x), 4)
        self._call("devfn_blend_adjoint", mode, len(x), x, e, g, z, acc_in)
        return z.cpu().numpy(), acc_in.cpu().numpy()

    def tf(self, mode, table, density):
        t, d = self.dev(table), self.dev(density)
        out = self.zeros(len(d), 4)
        self._call("devfn_tf", mode, t.shape[0], t.shape[1], t, len(d), d, out)
        return out.cpu().numpy()

    def tf_adjoint(self, mode, table, density, g, acc=2):
        t, d, g = self.dev(table), self.dev(density), self.dev(g)
        adj_d, adj_tf = self.zeros(len(d)), self.zeros(len(d), *t.shape)
        self._call("devfn_tf_adjoint", mode, t.shape[0], t.shape[1], t, len(d), d, g, adj_d, adj_tf, acc)
        return adj_d.cpu().numpy(), adj_tf.cpu().numpy()

    def quads(self, volume):
        """device QUAD copy in the layout the harness build expects ([x][y][z] quads, or [y][z][x] with DVR_QUAD_XINNER)"""
        q = build_quads(np.asarray(volume, np.float32))
        return self.dev(q.transpose(1, 2, 0, 3) if self.lib.devfn_quad_xinner() else q)

    def interp(self, volume, pos, quad=False):
        v, p = self.dev(volume), self.dev(pos)
        q = self.quads(volume) if quad else None
        out = self.zeros(len(p))
        self._call("devfn_interp", *v.shape, v, q, len(p), p, out)
        return out.cpu().numpy()

    def interp_adjoint(self, volume, pos, g, quad=False):
        v, p, g = self.dev(volume), self.dev(pos), self.dev(g)
        q = self.quads(volume) if quad else None
        adj_pos = self.zeros(len(p), 3)
        xinner = bool(self.lib.devfn_quad_xinner())
        if not quad:
            adj_vol = self.zeros(len(p), *v.shape)
        elif xinner:
            adj_vol = self.zeros(len(p), v.shape[1], v.shape[2], v.shape[0] + 1, 4)
        else:
            adj_vol = self.zeros(len(p), v.shape[0] + 1, *v.shape[1:], 4)
        self._call("devfn_interp_adjoint", *v.shape, v, q, len(p), p, g, adj_pos, adj_vol)
        av = adj_vol.cpu().numpy()
        if quad and xinner:
            av = np.ascontiguousarray(av.transpose(0, 3, 1, 2, 4))
        return adj_pos.cpu().numpy(), (fold_quads(av).astype(np.float32) if quad else av)


# ---- the unit tests' inputs ---------------------------------------------------------------------------
def blend_inputs(n=20):
    """testAdjointBlending.cpp:57-66: 20 vectors of 9 values (acc, current, stepsize) uniform in [0.01, 0.99]"""
    return np.random.RandomState(SEED).uniform(0.01, 0.99, (n, 9))


def create_tf(mode, rng):
    """createTF<mode> of testAdjointTF.cpp:10-74 (row-major R x C here)"""
    if mode == capi.TF_IDENTITY:
        return rng.uniform(0, 1, (1, 2))
    if mode == capi.TF_TEXTURE:
        return rng.uniform(0, 1, (8, 4))
    if mode == capi.TF_LINEAR:
        R = 5
        pos = np.concatenate([[0.0], np.sort(rng.uniform(0.1, 0.9, R - 2)), [1.0]])
        return np.concatenate([rng.uniform(0.1, 0.9, (R, 4)), pos[:, None]], axis=1)
    tf = rng.uniform(0, 1, (4, 6))
    tf[:, 5] += 0.1  # sigma
    return tf


def tf_inputs(mode, n_tf=10, n_density=10):
    """testAdjointTF.cpp:99-110: 10 TFs x 10 densities uniform in [0.01, 0.99]"""
    rng = np.random.RandomState(SEED)
    return [(create_tf(mode, rng), rng.uniform(0.01, 0.99, n_density)) for _ in range(n_tf)]


def interp_inputs(n=20):
    """testAdjointInterpolation.cpp:29-36, 91-103: a 2 x 3 x 4 volume uniform in [0.01, 0.99], 20 positions inside it
    (voxel coordinates, i.e. uniform / voxelSize)"""
    rng = np.random.RandomState(SEED)
    res = (2, 3, 4)
    volume = rng.uniform(0.01, 0.99, res)
    pos = rng.uniform(0.01, 0.99, (n, 3)) * np.array(res)
    return volume, pos


def approx(a, b, epsilon, margin):
    """Catch's `a == Approx(b).epsilon(epsilon).margin(margin)`"""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return (np.abs(a - b) <= margin) | (np.abs(a - b) <= epsilon * np.abs(b))


def check_adjoint(x, forward, adjoint, epsilon=1e-5, precision=1e-6, margin=1e-10, rng=None):
    """check_adjoint.h:36-88: x (n), forward(x) -> (m), adjoint(x, e, g) -> z (n); the trial directions are the m unit
    vectors and 2n random ones; z[i] must equal the numerical derivative of the output along g within the unit test's
    precision / margin. The unit tests take a one-sided difference, whose O(epsilon f'') error sits right at their
    margins (a Gaussian TF with sigma 0.1 has f'' ~ 400: 1e-5 * 200 * |g| against a margin of 1e-4) and passes on
    the particular stream std::default_random_engine(42) gives; the central difference used here has the same
    meaning and an O(epsilon^2) error, so the check does not depend on the random stream."""
    rng = rng or np.random.RandomState(SEED)
    n = len(x)
    ref = forward(x)
    m = len(ref)
    shifted = []
    for i in range(n):
        hi, lo = x.copy(), x.copy()
        hi[i] += epsilon
        lo[i] -= epsilon
        shifted.append((forward(hi) - forward(lo)) / (2 * epsilon))
    worst = 0.0
    for trial in range(m + 2 * n):
        g = np.eye(m)[trial] if trial < m else rng.uniform(-1, 1, m)
        z = adjoint(x, ref, g)
        want = np.array([s @ g for s in shifted])
        ok = approx(z, want, precision, margin)
        assert ok.all(), (trial, z, want)
        worst = max(worst, float(np.abs(z - want).max()))
    return worst
