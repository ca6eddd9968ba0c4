"""GPU: renderer::Volume's producers of the density tensor on sm_100a through the C-ABI and the
host layer, against the CPU oracle, the reference-built fixtures (tests/golden/ref_volume_*) and, when the binary
travelled, the reference-built Volume itself (oracle/_ref/libdvr_refvol.so).
Tolerances: BIT-EXACT for the 12 polynomial / sqrt surfaces (the kernel is compiled without FMA contraction and keeps
the reference's operation order), for the raw <-> tensor conversion (u8 / 255, u16 / 65535 are IEEE divisions) and for
the .cvol round trips; absolute 2.5e-7 / 1e-6 / 2e-5 for MarschnerLobb / Blobby / Tube, whose sin and cos come from the
device's libm instead of glibc's (volume_cases.TRANSCENDENTAL); 1e-6 for pooled mipmaps (summation order)."""
import os

import numpy as np
import pytest
import torch

import oracle_loader as ol
import volume_cases as vc
from diffdvr_b200 import capi, load_pyrenderer

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
IMPLICIT = vc.implicit_cases()


def _wrap(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _run(fn, lib, case):
    return fn(lib, case, np.float32, _wrap, lambda t: t.cpu().numpy(), torch.cuda.synchronize)


@pytest.mark.parametrize("name", sorted(IMPLICIT))
def test_implicit_matches_oracle_and_golden(name):
    case = IMPLICIT[name]
    got = _run(vc.run_implicit, capi.load(), case)
    want = vc.run_implicit(ol.port(32), case)
    gold = np.load(os.path.join(GOLDEN, "ref_volume_implicit.npz"))[name]
    tol = vc.TRANSCENDENTAL.get(case[0])
    for ref in (want, gold):
        if tol is None:
            assert np.array_equal(got, ref), np.abs(got - ref).max()
        else:
            assert np.abs(got - ref).max() <= tol, np.abs(got - ref).max()


def test_implicit_strided_output_and_errors():
    lib = capi.load()
    n = 21
    a = torch.empty(1, n, n, n, device="cuda")
    lib.create_implicit_volume(5, a)
    b = torch.empty(1, n, n, n, device="cuda").permute(0, 3, 2, 1)  # X fastest, like the reference's tensors
    lib.create_implicit_volume(5, b)
    torch.cuda.synchronize()
    assert not b.is_contiguous() and torch.equal(a, b)
    with pytest.raises(capi.DvrError):
        lib.create_implicit_volume(15, a)
    with pytest.raises(capi.DvrError):
        lib.create_implicit_volume(2, torch.empty(1, 4, 5, 6, device="cuda"))


def test_implicit_full_size_properties():
    """256^3 (BASELINE configs' volume size): symmetries that need no oracle."""
    lib = capi.load()
    n = 256
    s = torch.empty(1, n, n, n, device="cuda")
    lib.create_implicit_volume(2, s)  # sphere: 1 - |p|, invariant under axis permutations and flips
    assert torch.equal(s, s.permute(0, 2, 1, 3))  # x <-> y: the same fp32 operations (x*x + y*y commutes)
    # other permutations re-associate the sum, flips meet -1 + 255 * step != 1 exactly: rounding only
    for t in (s.permute(0, 2, 3, 1), s.flip(1), s.flip(3)):
        assert float((s - t).abs().max()) <= 1e-6
    inv = torch.empty_like(s)
    lib.create_implicit_volume(3, inv)
    assert torch.equal(1 - inv, s)
    ml = torch.empty_like(s)
    lib.create_implicit_volume(0, ml)  # Marschner-Lobb: rotationally symmetric about z, within [0, 1]
    assert float(ml.min()) >= 0.0 and float(ml.max()) <= 1.0
    assert float((ml - ml.permute(0, 2, 1, 3)).abs().max()) <= 2.5e-7
    for eq in (4, 7, 9, 10, 11, 12, 13):  # clamped densities
        v = torch.empty_like(s)
        lib.create_implicit_volume(eq, v)
        assert float(v.min()) >= 0.0 and float(v.max()) <= 1.0 and float(v.max()) > float(v.min())


@pytest.mark.parametrize("name", sorted(vc.raw_cases()))
def test_raw_conversion_both_ways(name):
    case = vc.raw_cases()[name]
    raw, vol, back = _run(vc.run_raw, capi.load(), case)
    want = vc.expected_from_raw(case, raw)
    assert np.array_equal(vol, want)
    assert np.array_equal(back, want.transpose(2, 1, 0).ravel())
    _, vol_port, _ = vc.run_raw(ol.port(32), case)
    assert np.array_equal(vol, vol_port)


def test_raw_conversion_full_size_round_trip():
    lib = capi.load()
    X, Y, Z = 256, 200, 300
    vol = torch.rand(1, X, Y, Z, device="cuda")
    raw = torch.empty(X * Y * Z, device="cuda")
    lib.volume_to_raw(vol, raw)
    assert torch.equal(raw.view(Z, Y, X), vol[0].permute(2, 1, 0))
    back = torch.empty(1, Z, Y, X, device="cuda").permute(0, 3, 2, 1)  # an X-fastest destination
    lib.volume_from_raw(raw, 2, back)
    assert torch.equal(back, vol)


@pytest.mark.parametrize("name", sorted(vc.mipmap_cases()))
def test_mipmap_matches_oracle_and_adaptive_avg_pool3d(name):
    case = vc.mipmap_cases()[name]
    vol, out = _run(vc.run_mipmap, capi.load(), case)
    _, want = vc.run_mipmap(ol.port(32), case)
    pool = torch.nn.functional.adaptive_avg_pool3d(torch.from_numpy(vol)[None], out.shape)[0].numpy()
    assert np.abs(out - want).max() <= 1e-6 and np.abs(out - pool).max() <= 1e-6


# ---- the host layer: pyrenderer.Volume --------------------------------------------------------------
def test_volume_loads_reference_written_files():
    pyr = load_pyrenderer()
    want = vc.golden_cvol_volume()
    for comp in (0, 1, 9):
        v = pyr.Volume(os.path.join(GOLDEN, f"ref_volume_c{comp}.cvol"))
        assert (v.resolution.x, v.resolution.y, v.resolution.z) == want.shape
        ws = v.world_size
        assert np.allclose((ws.x, ws.y, ws.z), vc.GOLDEN_CVOL_WORLD, rtol=1e-6)
        gpu = v.getDataGpu(0)
        assert gpu.is_cuda and tuple(gpu.shape) == (1,) + want.shape
        assert np.array_equal(gpu.cpu().numpy()[0], want)
        assert np.array_equal(v.getDataCpu(0).numpy()[0], want)


@pytest.mark.parametrize("dtype", [0, 1])
def test_volume_loads_quantised_files(dtype, tmp_path):
    pyr = load_pyrenderer()
    vol = vc.make_volume((33, 6, 40))
    data = vc.cvol_write_raw(vol, (1.0, 1.0, 1.0), dtype)
    want, _, _, _ = vc.cvol_read(data)
    path = tmp_path / "q.cvol"
    path.write_bytes(data)
    got = pyr.Volume(str(path)).getDataGpu(0).cpu().numpy()[0]
    assert np.array_equal(got, want)
    assert np.abs(got - vol).max() <= 1.0 / (255, 65535)[dtype]


@pytest.mark.parametrize("comp", [0, 1, 5, 9])
def test_volume_save_round_trip_and_reference_reads_it(comp, tmp_path):
    pyr = load_pyrenderer()
    rng = np.random.RandomState(vc.SEED + comp)
    vol = (rng.randint(0, 6, (45, 40, 50)) / 5).astype(np.float32)  # 360 KB: 6 LZ4 chunks
    vol[:, :, 30:] = vol[:, :, :20]
    v = pyr.Volume.from_numpy(vol)
    v.set_world_size(2.0, 3.0, 4.0)
    path = str(tmp_path / "out.cvol")
    v.save(path, comp)
    data = open(path, "rb").read()
    got, world, flag, dtype = vc.cvol_read(data) if comp == 0 or len(data) < 100000 else (None, None, None, None)
    if got is not None:
        assert np.array_equal(got, vol) and np.allclose(world, (2.0, 3.0, 4.0)) and flag == (comp > 0) and dtype == 2
    if comp:
        assert len(data) < 0.7 * vol.nbytes
    again = pyr.Volume(path)
    assert np.array_equal(again.getDataGpu(0).cpu().numpy()[0], vol)
    assert np.allclose((again.world_size.x, again.world_size.y, again.world_size.z), (2.0, 3.0, 4.0))
    rv = ol.refvol()
    if rv is not None:  # the reference's own reader (its liblz4) on the product's file
        ref_got, ref_world = rv.load(path)
        assert np.array_equal(ref_got, vol) and np.allclose(ref_world, (2.0, 3.0, 4.0))


def test_volume_create_implicit_from_numpy_and_mipmaps():
    pyr = load_pyrenderer()
    v = pyr.Volume.create_implicit(pyr.ImplicitEquation.MarschnerLobb, 32, fM=4.0, alpha=0.5)
    want = vc.run_implicit(ol.port(32), (0, 32, dict(fM=4.0, alpha=0.5)))
    assert np.abs(v.getDataGpu(0).cpu().numpy()[0] - want).max() <= 2.5e-7
    assert (v.resolution.x, v.resolution.y, v.resolution.z) == (32, 32, 32)
    v.copy_to_gpu()  # kept for source compatibility
    with pytest.raises(Exception, match="was not created"):
        v.getDataGpu(2)
    v.create_mipmap_level(2)
    m = v.getDataGpu(2)
    assert tuple(m.shape) == (1, 10, 10, 10)
    pool = torch.nn.functional.adaptive_avg_pool3d(v.getDataGpu(0), (10, 10, 10))
    assert float((m - pool).abs().max()) <= 1e-6

    # from_numpy takes strided arrays (a transposed view) like the reference's buffer protocol path
    a = vc.make_volume((5, 6, 7))
    assert np.array_equal(pyr.Volume.from_numpy(a.transpose(2, 0, 1)).getDataGpu(0).cpu().numpy()[0], a.transpose(2, 0, 1))
    assert np.array_equal(pyr.Volume.from_numpy(a[::-1]).getDataGpu(0).cpu().numpy()[0], a[::-1])
    with pytest.raises(Exception):
        pyr.Volume.from_numpy(a.astype(np.float64))



def test_implicit_volume_renders_like_the_oracle():
    """Volume.create_implicit -> Renderer.render_forward (the reference's train_volume / example pipeline):
    the image matches the CPU oracle rendering the oracle's own implicit volume; rel-L2 <= 1e-5."""
    import scenes
    pyr = load_pyrenderer()
    n, W, H = 48, 40, 32
    v = pyr.Volume.create_implicit(pyr.ImplicitEquation.Tube, n)
    i = pyr.RendererInputs()
    i.screen_size = pyr.int2(W, H)
    i.volume_filter_mode = pyr.VolumeFilterMode.Trilinear
    i.volume = v.getDataGpu(0)
    i.box_min = pyr.real3(-0.5, -0.5, -0.5)
    i.box_size = pyr.real3(1, 1, 1)
    i.camera_mode = pyr.CameraMode.ReferenceFrame
    vp = pyr.Camera.viewport_from_sphere(torch.zeros(1, 3, device="cuda"), torch.tensor([[0.5]], device="cuda"),
                                         torch.tensor([[0.3]], device="cuda"), torch.tensor([[1.8]], device="cuda"),
                                         pyr.Orientation.Ym)
    i.camera = pyr.CameraReferenceFrame(vp, 0.6)
    i.step_size = 0.01
    i.tf_mode = pyr.TFMode.Identity
    i.tf = torch.tensor([[[4.0, 1.0]]], device="cuda")
    i.blend_mode = pyr.BlendMode.BeerLambert
    out = pyr.RendererOutputs(torch.empty(1, H, W, 4, device="cuda"),
                              torch.empty(1, H, W, dtype=torch.int32, device="cuda"))
    pyr.Renderer.render_forward(i, out)
    torch.cuda.synchronize()

    port = ol.port(32)
    ci = capi.Inputs(screen_size=(W, H), volume=np.ascontiguousarray(vc.run_implicit(port, (14, n, {}))[None]),
                     volume_filter_mode=capi.FILTER_TRILINEAR, box_min=(-0.5, -0.5, -0.5), box_size=(1.0, 1.0, 1.0),
                     camera_mode=capi.CAMERA_REFERENCE_FRAME, camera_matrix=np.ascontiguousarray(vp.cpu().numpy()),
                     camera_fov_y_radians=0.6, step_size=0.01, tf_mode=capi.TF_IDENTITY,
                     tf=np.array([[[4.0, 1.0]]], np.float32), blend_mode=capi.BLEND_BEER_LAMBERT)
    color, term = np.zeros((1, H, W, 4), np.float32), np.zeros((1, H, W), np.int32)
    port.render_forward(ci, color, term)
    got = out.color.cpu().numpy()
    assert got[..., 3].max() > 0.1
    assert scenes.rel_l2(got, color) <= 1e-5
