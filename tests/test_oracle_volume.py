"""CPU: pins the oracle's restatement of renderer::Volume's producers (oracle/dvr_oracle_volume.h) and the product's
HOST code for the .cvol format (own LZ4 block codec, diffdvr_b200/csrc/host/lz4_block.cpp) against
 * the reference's own volume.cpp / volume_implicit.cpp / liblz4 compiled unmodified (oracle/_ref/libdvr_refvol.so),
 * the committed fixtures tests/golden/ref_volume_implicit.npz and ref_volume_c{0,1,9}.cvol written by it,
 * a pure-Python statement of the file format (volume_cases.cvol_read),
 * torch.nn.functional.adaptive_avg_pool3d, the operator Volume::createMipmapLevel calls.
Bit-exact everywhere except the pooled mipmaps (summation order: 1e-6)."""
import os

import numpy as np
import pytest
import torch

import oracle_loader as ol
import volume_cases as vc
from diffdvr_b200 import load_pyrenderer

pyr = load_pyrenderer()
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
IMPLICIT = vc.implicit_cases()


@pytest.mark.parametrize("name", sorted(IMPLICIT))
def test_implicit_port_matches_reference_build(name):
    rv = ol.refvol()
    if rv is None:
        pytest.skip("oracle/_ref/libdvr_refvol.so not built (needs /root/reference)")
    eq, n, params = IMPLICIT[name]
    assert np.array_equal(vc.run_implicit(ol.port(32), IMPLICIT[name]), rv.create_implicit(eq, n, **params))


@pytest.mark.parametrize("bits", [32, 64])
def test_implicit_port_matches_golden(bits):
    gold = np.load(os.path.join(GOLDEN, "ref_volume_implicit.npz"))
    dt = np.float32 if bits == 32 else np.float64
    for name, case in IMPLICIT.items():
        got = vc.run_implicit(ol.port(bits), case, dt)
        assert got.dtype == dt
        assert np.array_equal(got.astype(np.float32), gold[name]), name  # surfaces are fp32 whatever `real` is


def test_implicit_known_answers():
    # values that follow from the formulas alone (renderer/volume_implicit.cpp): sphere centre and corner,
    # Marschner-Lobb on the axis r = 0 (cos(2 pi fM cos 0) = 1 for integer fM) at z = 0
    port = ol.port(32)
    n = 19
    sphere = vc.run_implicit(port, (2, n, {}))
    assert sphere[9, 9, 9] == 1.0 and abs(sphere[0, 0, 0] - (1 - np.sqrt(3))) < 1e-6
    inv = vc.run_implicit(port, (3, n, {}))
    assert np.array_equal(inv, 1 - sphere) or np.abs(inv - (1 - sphere)).max() < 1e-6
    ml = vc.run_implicit(port, (0, n, {}))
    assert abs(ml[9, 9, 9] - (1 + 0.25 * 2) / (2 * 1.25)) < 1e-6
    with pytest.raises(Exception):
        vc.run_implicit(port, (15, n, {}))


@pytest.mark.parametrize("name", sorted(vc.raw_cases()))
def test_raw_conversion_port(name):
    case = vc.raw_cases()[name]
    raw, vol, back = vc.run_raw(ol.port(32), case)
    want = vc.expected_from_raw(case, raw)
    assert np.array_equal(vol, want)
    assert np.array_equal(back, want.transpose(2, 1, 0).ravel())


@pytest.mark.parametrize("name", sorted(vc.mipmap_cases()))
def test_mipmap_port_matches_adaptive_avg_pool3d(name):
    vol, out = vc.run_mipmap(ol.port(32), vc.mipmap_cases()[name])
    want = torch.nn.functional.adaptive_avg_pool3d(torch.from_numpy(vol)[None], out.shape)[0].numpy()
    assert out.shape == want.shape
    assert np.abs(out - want).max() <= 1e-6


# ---- the .cvol format -------------------------------------------------------------------------------
@pytest.mark.parametrize("comp", [0, 1, 9])
def test_python_reader_on_reference_written_files(comp):
    data = open(os.path.join(GOLDEN, f"ref_volume_c{comp}.cvol"), "rb").read()
    vol, world, flag, dtype = vc.cvol_read(data)
    assert np.array_equal(vol, vc.golden_cvol_volume())
    assert np.allclose(world, vc.GOLDEN_CVOL_WORLD, rtol=1e-15)
    assert flag == (1 if comp else 0) and dtype == 2


@pytest.mark.parametrize("comp", [1, 9])
def test_product_decoder_on_reference_written_files(comp):
    data = open(os.path.join(GOLDEN, f"ref_volume_c{comp}.cvol"), "rb").read()
    want = np.ascontiguousarray(vc.golden_cvol_volume().transpose(2, 1, 0)).tobytes()
    assert pyr.cvol_decompress_body(data[64:], len(want)) == want


def _bodies():
    rng = np.random.RandomState(vc.SEED)
    smooth = np.fromfunction(lambda x, y, z: np.sin(x / 9) + y / 50 + (z > 20), (48, 40, 36)).astype(np.float32)
    return {
        "empty_slab": np.zeros(70000, np.uint8).tobytes(),                    # > 1 chunk of one repeated byte
        "noise": rng.randint(0, 256, 150000).astype(np.uint8).tobytes(),      # incompressible, 3 chunks
        "quantised": (rng.randint(0, 4, 100000) / 4).astype(np.float32).tobytes(),
        "smooth_planes": smooth.transpose(2, 1, 0).tobytes(),                 # planes repeat across chunk borders
        "chunk_minus_1": rng.randint(0, 3, 65535).astype(np.uint8).tobytes(),
        "chunk_exact": rng.randint(0, 3, 65536).astype(np.uint8).tobytes(),
        "chunk_plus_1": rng.randint(0, 3, 65537).astype(np.uint8).tobytes(),
        "twelve": bytes(range(12)), "thirteen": bytes([7] * 13), "one": b"\x05",
    }


@pytest.mark.parametrize("level", [1, 3, 9])
@pytest.mark.parametrize("name", sorted(_bodies()))
def test_product_encoder_round_trips(name, level):
    raw = _bodies()[name]
    body = pyr.cvol_compress_body(raw, level)
    assert pyr.cvol_decompress_body(body, len(raw)) == raw
    # every chunk is [int32 n][block] with at most 64 KiB of input; count them
    i, chunks = 0, 0
    while i < len(body):
        i += 4 + int.from_bytes(body[i:i + 4], "little", signed=True)
        chunks += 1
    assert i == len(body) and chunks == -(-len(raw) // 65536)
    if len(raw) <= 70000:  # the pure-Python decoder is slow
        out = bytearray()
        i = 0
        while i < len(body):
            nb = int.from_bytes(body[i:i + 4], "little", signed=True)
            vc.lz4_block_decode(body[i + 4:i + 4 + nb], out)
            i += 4 + nb
        assert bytes(out) == raw
    if name in ("empty_slab", "quantised", "smooth_planes"):
        assert len(body) < 0.6 * len(raw)  # it does compress (fast mode on 2-bit noise: ~0.5)


def test_product_decoder_rejects_malformed_streams():
    raw = _bodies()["quantised"]
    body = pyr.cvol_compress_body(raw, 1)
    with pytest.raises(Exception):
        pyr.cvol_decompress_body(body[:len(body) // 2], len(raw))           # truncated
    with pytest.raises(Exception):
        pyr.cvol_decompress_body(body, len(raw) // 2)                       # output larger than announced
    bad = bytearray(pyr.cvol_compress_body(bytes(64), 1))
    bad[5:7] = b"\xff\xff"                                                  # a match offset before the buffer
    with pytest.raises(Exception):
        pyr.cvol_decompress_body(bytes(bad), 64)
    with pytest.raises(Exception):
        pyr.cvol_decompress_body((0).to_bytes(4, "little"), 8)              # empty chunk
    with pytest.raises(Exception):
        pyr.cvol_decompress_body((1 << 30).to_bytes(4, "little") + bytes(100), 8)  # absurd chunk size: no 1 GiB alloc


@pytest.mark.parametrize("level", [1, 3, 9])
def test_reference_reads_files_written_by_the_product_encoder(level, tmp_path):
    rv = ol.refvol()
    if rv is None:
        pytest.skip("oracle/_ref/libdvr_refvol.so not built (needs /root/reference)")
    rng = np.random.RandomState(vc.SEED + level)
    vol = (rng.randint(0, 5, (47, 41, 39)) / 4).astype(np.float32)          # 300 KB: 5 chunks
    vol[:, :, 20:30] = vol[:, :, 5:15]                                      # repeats 10 planes = 77 KB back: out of window
    vol[:, :, 31:33] = vol[:, :, 29:31]                                     # and 15 KB back: inside it, across a chunk border
    raw = np.ascontiguousarray(vol.transpose(2, 1, 0)).tobytes()
    path = str(tmp_path / "product.cvol")
    open(path, "wb").write(vc.cvol_header(vol.shape, (1.0, 2.0, 3.0), 2, 1) + pyr.cvol_compress_body(raw, level))
    got, world = rv.load(path)
    assert np.array_equal(got, vol) and np.allclose(world, (1.0, 2.0, 3.0))
    # and the other direction on the same volume: the reference's LZ4 stream (streaming dictionary across chunks)
    rv.save(path, vol, (1.0, 2.0, 3.0), level)
    data = open(path, "rb").read()
    assert pyr.cvol_decompress_body(data[64:], len(raw)) == raw


def test_volume_class_surface_and_loud_failure_without_a_gpu(tmp_path):
    for a in ["create_mipmap_level", "world_size", "set_world_size", "resolution", "create_implicit", "from_numpy",
              "copy_to_gpu", "save", "getDataCpu", "getDataGpu"]:
        assert hasattr(pyr.Volume, a), a
    names = ["MarschnerLobb", "Cube", "Sphere", "InverseSphere", "DingDong", "Endrass", "Barth", "Heart", "Kleine",
             "Cassini", "Steiner", "CrossCap", "Kummer", "Blobbly", "Tube"]
    assert [int(getattr(pyr.ImplicitEquation, n)) for n in names] == list(range(15))
    if not torch.cuda.is_available():  # no CPU back-end, and no silent one either
        with pytest.raises(Exception, match="CUDA"):
            pyr.Volume.create_implicit(pyr.ImplicitEquation.Sphere, 8)
        with pytest.raises(Exception, match="CUDA"):
            pyr.Volume(os.path.join(GOLDEN, "ref_volume_c0.cvol"))
    with pytest.raises(Exception, match="magic"):
        p = tmp_path / "bad.cvol"
        p.write_bytes(b"nope" + bytes(60))
        pyr.Volume(str(p))
