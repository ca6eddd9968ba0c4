#!/usr/bin/env python
"""Generates the golden fixtures under tests/golden/ from the REFERENCE ITSELF.

Run in the build container (needs /root/reference and `make -C oracle ref`):

    python tests/golden/make_golden.py

* ``ref_*.npz``: outputs of the reference's own OpenMP host kernels (oracle/_ref/libdvr_ref32.so =
  DvrKernelForwardHost / DvrKernelAdjointHost / DvrKernelForwardGradientsHost of
  renderer/renderer_kernels.cuh compiled unmodified from /root/reference; the forward-mode kernel is
  instantiated over the restated cudAD of oracle/cudad_restated) on the seeded scenes of
  tests/scenes.py. The scene INPUTS are regenerated from the seed by the tests; only the outputs are
  stored.
* ``ref_camera_*.npz``: Camera::generateRays (forward + backward) and Renderer::compareToImage from the
  same library (oracle/ref_shim_camera.cpp: the reference's CameraEval / adjNormalize / adjCross /
  CompareToImage headers) on the seeded inputs of tests/camera_cases.py.
* ``ref_volume_implicit.npz``: Volume::createImplicitDataset of the reference's own volume_implicit.cpp / volume.cpp
  (oracle/_ref/libdvr_refvol.so, compiled unmodified; oracle/ref_shim_volume.cpp) on the cases of
  tests/volume_cases.py; ``ref_volume_c{0,1,9}.cvol``: one seeded 20 x 13 x 9 volume written by the reference's
  Volume::save uncompressed, with LZ4 fast and with LZ4 HC level 9 (its vendored liblz4).
* ``NonConvexDensity.tsv``: copied verbatim from /root/reference/unittests/NonConvexDensity.tsv
  (the reference's committed known-answer table, generator unittests/testNonConvexDensity.cpp:8-80).
"""
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import camera_cases  # noqa: E402
import golden_cases  # noqa: E402
import oracle_loader as ol  # noqa: E402
import scenes  # noqa: E402
import volume_cases  # noqa: E402


def main():
    ref = ol.ref(32)
    assert ref is not None, "oracle/_ref/libdvr_ref32.so missing: run `make -C oracle ref`"
    for name, sc, flags, fg in golden_cases.cases():
        if os.path.exists(os.path.join(HERE, f"ref_{name}.npz")) and "--all" not in sys.argv:
            continue  # fixtures are immutable once committed; --all regenerates every one
        out = golden_cases.run(ref, sc, flags, fg)
        np.savez_compressed(os.path.join(HERE, f"ref_{name}.npz"), **out)
        print(name, {k: v.shape for k, v in out.items()})
    # Camera::generateRays / compareToImage built on the reference's own headers (ref_shim_camera.cpp)
    for name, case in camera_cases.ray_cases().items():
        path = os.path.join(HERE, f"ref_camera_{name}.npz")
        if not os.path.exists(path) or "--all" in sys.argv:
            out = camera_cases.run_rays(ref, case)
            np.savez_compressed(path, **{k: out[k] for k in ("viewport", "ray_start", "ray_dir", "adj_viewport")})
            print(name)
    for name, case in camera_cases.compare_cases().items():
        path = os.path.join(HERE, f"ref_camera_{name}.npz")
        if not os.path.exists(path) or "--all" in sys.argv:
            np.savez_compressed(path, **camera_cases.run_compare(ref, case))
            print(name)
    for name, case in camera_cases.preshade_cases().items():
        path = os.path.join(HERE, f"ref_{name}.npz")
        if not os.path.exists(path) or "--all" in sys.argv:
            out = camera_cases.run_preshade(ref, case)
            np.savez_compressed(path, **{k: out[k] for k in ("tf", "color", "adj_volume")})
            print(name)
    rv = ol.refvol()
    assert rv is not None, "oracle/_ref/libdvr_refvol.so missing: run `make -C oracle refvol`"
    path = os.path.join(HERE, "ref_volume_implicit.npz")
    if not os.path.exists(path) or "--all" in sys.argv:
        np.savez_compressed(path, **{name: rv.create_implicit(eq, n, **params)
                                     for name, (eq, n, params) in volume_cases.implicit_cases().items()})
        print("ref_volume_implicit")
    for comp in (0, 1, 9):
        path = os.path.join(HERE, f"ref_volume_c{comp}.cvol")
        if not os.path.exists(path) or "--all" in sys.argv:
            rv.save(path, volume_cases.golden_cvol_volume(), volume_cases.GOLDEN_CVOL_WORLD, comp)
            print(os.path.basename(path), os.path.getsize(path))
    src = "/root/reference/unittests/NonConvexDensity.tsv"
    if os.path.exists(src):
        shutil.copyfile(src, os.path.join(HERE, "NonConvexDensity.tsv"))


if __name__ == "__main__":
    main()
