"""Seeded inputs for renderer::Volume's producers of the density tensor and the routines that
run ANY library exporting the C-ABI entry points on them: the CPU port when pinning it, the CUDA product when
checking it. The reference-built Volume (oracle_loader.refvol) generates the fixtures
tests/golden/ref_volume_*.npz and tests/golden/ref_*.cvol (tests/golden/make_golden.py)."""
from __future__ import annotations

import struct

import numpy as np

SEED = 91
EQUATIONS = ["MarschnerLobb", "Cube", "Sphere", "InverseSphere", "DingDong", "Endrass", "Barth", "Heart", "Kleine",
             "Cassini", "Steiner", "CrossCap", "Kummer", "Blobbly", "Tube"]
# equations whose value goes through sin / cos of the platform's libm (glibc vs CUDA): not bit-comparable
TRANSCENDENTAL = {0: 2.5e-7, 13: 1e-6, 14: 2e-5}   # absolute tolerance (Tube is scaled by 10, |v| <= ~25)
PARAM_ORDER = ("fM", "alpha", "scale", "a")
PARAM_DEFAULTS = (6.0, 0.25, 0.5, 0.25)


def implicit_cases():
    """name -> (equation, resolution, params dict)"""
    cases = {f"implicit_{i:02d}_{EQUATIONS[i].lower()}": (i, 19, {}) for i in range(15)}
    cases["implicit_00_marschnerlobb_params"] = (0, 17, dict(fM=4.0, alpha=0.5))
    cases["implicit_01_cube_params"] = (1, 16, dict(scale=0.3))
    cases["implicit_09_cassini_params"] = (9, 18, dict(a=0.4))
    return cases


def param_array(params):
    if not params:
        return None
    return [params.get(k, d) for k, d in zip(PARAM_ORDER, PARAM_DEFAULTS)]


def run_implicit(lib, case, dtype=np.float32, wrap=None, unwrap=None, sync=None):
    wrap = wrap or (lambda a: a)
    unwrap = unwrap or (lambda t: t)
    eq, n, params = case
    out = wrap(np.full((1, n, n, n), 7.0, dtype))
    lib.create_implicit_volume(eq, out, param_array(params))
    if sync:
        sync()
    return unwrap(out)[0]


def raw_cases():
    """name -> (cvol dtype code, (X, Y, Z)); the sizes straddle the 32 x 32 transposition tiles of the CUDA kernel"""
    return {"raw_u8": (0, (33, 5, 70)), "raw_u16": (1, (64, 3, 32)), "raw_f32": (2, (7, 9, 11)),
            "raw_f32_line": (2, (1, 1, 5))}


def make_raw(case):
    dtype, (X, Y, Z) = case
    rng = np.random.RandomState(SEED + dtype)
    n = X * Y * Z
    if dtype == 0:
        return rng.randint(0, 256, n).astype(np.uint8)
    if dtype == 1:
        return rng.randint(0, 65536, n).astype(np.uint16)
    return rng.standard_normal(n).astype(np.float32)


def expected_from_raw(case, raw):
    """numpy statement of renderer/volume.cpp:232-256: X fastest body -> [X][Y][Z], u8 / 255, u16 / 65535"""
    dtype, (X, Y, Z) = case
    v = raw.reshape(Z, Y, X).transpose(2, 1, 0)
    if dtype == 0:
        return v.astype(np.float32) / np.float32(255.0)
    if dtype == 1:
        return v.astype(np.float32) / np.float32(65535.0)
    return v.copy()


def run_raw(lib, case, dtype=np.float32, wrap=None, unwrap=None, sync=None):
    wrap = wrap or (lambda a: a)
    unwrap = unwrap or (lambda t: t)
    code, (X, Y, Z) = case
    raw = make_raw(case)
    vol = wrap(np.full((1, X, Y, Z), 7.0, dtype))
    lib.volume_from_raw(wrap(raw), code, vol)
    back = wrap(np.full(X * Y * Z, 7.0, np.float32))
    lib.volume_to_raw(vol, back)
    if sync:
        sync()
    return raw, unwrap(vol)[0], unwrap(back)


def mipmap_cases():
    """name -> ((X, Y, Z), level): divisible and ragged pooling windows"""
    return {"mip_even_l1": ((8, 6, 10), 1), "mip_ragged_l2": ((11, 7, 13), 2), "mip_small_l3": ((3, 9, 5), 3)}


def make_volume(shape, seed=SEED):
    return np.random.RandomState(seed).uniform(0, 1, shape).astype(np.float32)


def run_mipmap(lib, case, dtype=np.float32, wrap=None, unwrap=None, sync=None):
    wrap = wrap or (lambda a: a)
    unwrap = unwrap or (lambda t: t)
    shape, level = case
    f = level + 1
    vol = make_volume(shape).astype(dtype)
    out = wrap(np.full((1,) + tuple(max(1, s // f) for s in shape), 7.0, dtype))
    lib.downsample_volume(wrap(vol[None]), out)
    if sync:
        sync()
    return vol, unwrap(out)[0]


GOLDEN_CVOL_WORLD = (2.0, 1.3, 0.9)


def golden_cvol_volume():
    """20 x 13 x 9: smooth ramp + quantised noise + an empty slab, so that the LZ4 stream has long matches, short
    matches and literal runs"""
    rng = np.random.RandomState(SEED + 7)
    x, y, z = np.meshgrid(np.arange(20), np.arange(13), np.arange(9), indexing="ij")
    v = 0.02 * x + 0.03 * y + rng.randint(0, 3, (20, 13, 9)) / 8.0
    v[:, :, 6:] = 0.0
    return v.astype(np.float32)


# ---- .cvol in pure Python: the independent statement of the file format (renderer/volume.cpp:64-279) --------------
def lz4_block_decode(src: bytes, out: bytearray):
    """Appends the decoded LZ4 block to `out` (which holds the previously decoded chunks = the match window)."""
    i, n = 0, len(src)
    while i < n:
        token = src[i]; i += 1
        lit = token >> 4
        if lit == 15:
            while True:
                b = src[i]; i += 1; lit += b
                if b != 255:
                    break
        out += src[i:i + lit]; i += lit
        if i >= n:
            break
        off = src[i] | src[i + 1] << 8; i += 2
        ml = token & 15
        if ml == 15:
            while True:
                b = src[i]; i += 1; ml += b
                if b != 255:
                    break
        ml += 4
        assert 0 < off <= len(out)
        start = len(out) - off
        for k in range(ml):  # byte-wise: matches may overlap their own output
            out.append(out[start + k])


def cvol_read(data: bytes):
    """-> (float32 [X][Y][Z], world size, compressed flag, dtype code)"""
    assert data[:4] == b"cvol"
    X, Y, Z, vx, vy, vz, dtype, comp = struct.unpack_from("<QQQdddiB", data, 4)
    size = X * Y * Z * (1, 2, 4)[dtype]
    body = data[64:]
    if comp:
        out, i = bytearray(), 0
        while len(out) < size:
            (nb,) = struct.unpack_from("<i", body, i)
            lz4_block_decode(body[i + 4:i + 4 + nb], out)
            i += 4 + nb
        body = bytes(out)
    raw = np.frombuffer(body[:size], (np.uint8, np.uint16, np.float32)[dtype])
    return expected_from_raw((dtype, (X, Y, Z)), raw), (vx * X, vy * Y, vz * Z), comp, dtype


def cvol_header(shape, world=(1.0, 1.0, 1.0), dtype=2, compressed=0):
    X, Y, Z = shape
    return b"cvol" + struct.pack("<QQQdddiB7x", X, Y, Z, world[0] / X, world[1] / Y, world[2] / Z, dtype, compressed)


def cvol_write_raw(vol, world=(1.0, 1.0, 1.0), dtype=2):
    """uncompressed .cvol bytes of a float32 [X][Y][Z] array (u8 / u16: quantised), for reader tests"""
    head = cvol_header(vol.shape, world, dtype)
    scale = (255.0, 65535.0, 1.0)[dtype]
    body = np.ascontiguousarray((vol * scale).astype((np.uint8, np.uint16, np.float32)[dtype]).transpose(2, 1, 0))
    return head + body.tobytes()
