#!/usr/bin/env python
"""Builds diffdvr_b200/pyrenderer*.so: the C++ host layer (renderer.cpp) + pybind11 bindings
(bindings.cpp) over the C-ABI library diffdvr_b200/libdvr_b200.so (linked with rpath $ORIGIN).

Host code only (no device code here): compiled with the system g++ against the installed libtorch
and pybind11 headers; the include / library paths come from torch.utils.cpp_extension, the compile
itself is two explicit g++ calls so that the artefact lands IN-TREE (it travels to the GPU box).
"""
import os
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(os.path.dirname(HERE))          # diffdvr_b200/
BUILD = os.path.join(PKG, "_build", "host")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")


def main():
    import torch
    from torch.utils import cpp_extension as ce
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    out = os.path.join(PKG, "pyrenderer" + sysconfig.get_config_var("EXT_SUFFIX"))
    lib = os.path.join(PKG, "libdvr_b200.so")
    if not os.path.exists(lib):
        raise SystemExit(f"{lib} missing: build the CUDA library first (make -C diffdvr_b200/csrc)")
    srcs = [os.path.join(HERE, "renderer.cpp"), os.path.join(HERE, "camera.cpp"), os.path.join(HERE, "tf_utils.cpp"),
            os.path.join(HERE, "volume.cpp"), os.path.join(HERE, "lz4_block.cpp"), os.path.join(HERE, "bindings.cpp")]
    deps = srcs + [os.path.join(HERE, "renderer.h"), os.path.join(HERE, "kernel_types.h"), os.path.join(HERE, "camera.h"), os.path.join(HERE, "tf_utils.h"),
                   os.path.join(HERE, "volume.h"), os.path.join(HERE, "lz4_block.h"), os.path.join(PKG, "..", "include", "dvr_b200.h"), lib, __file__]
    if os.path.exists(out) and all(os.path.getmtime(out) >= os.path.getmtime(d) for d in deps):
        print("pyrenderer is up to date")
        return
    os.makedirs(BUILD, exist_ok=True)
    inc = [f"-I{p}" for p in ce.include_paths()] + [f"-I{sysconfig.get_paths()['include']}", f"-I{CUDA}/include"]
    cflags = ["-std=c++17", "-O2", "-fPIC", "-fvisibility=hidden", "-w", "-DTORCH_API_INCLUDE_EXTENSION_H",
              f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}", "-DTORCH_EXTENSION_NAME=pyrenderer"]
    objs = [os.path.join(BUILD, os.path.basename(s) + ".o") for s in srcs]

    def compile_one(so):
        cmd = [cxx] + cflags + inc + ["-c", so[0], "-o", so[1]]
        print("+", " ".join(cmd), flush=True)
        subprocess.run(cmd, check=True)

    with ThreadPoolExecutor(6) as ex:
        list(ex.map(compile_one, zip(srcs, objs)))
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    link = [cxx, "-shared", "-o", out] + objs + [f"-L{tlib}", "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch",
                                                  "-ltorch_python", f"-L{PKG}", "-ldvr_b200",
                                                  "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{tlib}"]
    print("+", " ".join(link), flush=True)
    subprocess.run(link, check=True)
    print("built", out)


if __name__ == "__main__":
    sys.exit(main())
