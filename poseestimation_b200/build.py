"""Build ``libpe_rotation.so`` in-tree for sm_100a:  ``python poseestimation_b200/build.py [-v]``.

nvcc cross-compiles without a GPU.  The library is git-ignored (source-only history) but travels to the GPU box with
the working-tree snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SOURCES = ["pe_api.cu", "pe_api_svd.cu", "pe_api_step.cu", "pe_api_metric.cu", "pe_kabsch.cu", "pe_cloud.cu", "pe_addl1.cu", "pe_static.cu"]
HEADERS = ["pe_common.h", "pe_math.cuh", "pe_stage.cuh", "pe_stream.cuh", "pe_lanes.cuh", "pe_ops.cuh", "pe_api_run.cuh",
           os.path.join("..", "..", "include", "pe_rotation.h")]
OUT = os.path.join(HERE, "libpe_rotation.so")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found")
    return exe


def up_to_date() -> bool:
    if not os.path.exists(OUT):
        return False
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return all(os.path.getmtime(d) <= t for d in deps)


def build(force: bool = False, verbose: bool = False, out: str = OUT, defines: tuple = ()) -> str:
    """`out` / `defines`: A/B variants of the library (e.g. -DPE_ANGLE_LIBM=1 into scratch_bin/), selected at run time with
    the environment variable PE_ROTATION_LIB; the product build takes neither."""
    if not force and out == OUT and up_to_date():
        return OUT
    # -cudart shared: the static runtime would embed the names of every runtime entry point (e.g. the batched memcpy
    # calls, which this library never makes) in the shipped artefact; the process already has torch's libcudart loaded
    cmd = [nvcc(), "-O3", "-std=c++17", "--threads", "0", *ARCH, "-lineinfo", "-cudart", "shared", "-Xcompiler", "-fPIC", "-shared",
           *defines, *[os.path.join(CSRC, s) for s in SOURCES], "-o", out]
    if verbose:
        cmd[1:1] = ["-Xptxas", "-v"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed building libpe_rotation.so")
    return out


TORCH_SRC = os.path.join(CSRC, "pe_torch.cpp")
TORCH_OUT = os.path.join(HERE, "_pe_torch.so")


def torch_ext_up_to_date() -> bool:
    if not os.path.exists(TORCH_OUT) or not os.path.exists(OUT):
        return False
    t = os.path.getmtime(TORCH_OUT)
    deps = [TORCH_SRC, os.path.join(CSRC, "..", "..", "include", "pe_rotation.h"), os.path.abspath(__file__)]
    return all(os.path.getmtime(d) <= t for d in deps)


def build_torch_ext(force: bool = False, verbose: bool = False) -> str:
    """`_pe_torch.so`: the C++ autograd node of the training step (csrc/pe_torch.cpp; host code only - no kernels),
    compiled with g++ against the installed torch and linked to libpe_rotation.so next to it."""
    if not force and torch_ext_up_to_date():
        return TORCH_OUT
    import sysconfig
    import torch
    from torch.utils import cpp_extension as ce
    cuda_home = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    inc = ce.include_paths() + [sysconfig.get_paths()["include"], os.path.join(cuda_home, "include")]
    torch_lib = os.path.join(os.path.dirname(torch.__file__), "lib")
    cxx = shutil.which("g++") or "g++"
    cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden",
           f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}", "-DTORCH_EXTENSION_NAME=_pe_torch",
           "-DTORCH_API_INCLUDE_EXTENSION_H", *[f"-I{i}" for i in inc], TORCH_SRC, "-o", TORCH_OUT,
           f"-L{HERE}", "-lpe_rotation", f"-L{torch_lib}", "-ltorch", "-ltorch_cpu", "-ltorch_python", "-lc10", "-lc10_cuda",
           "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{torch_lib}"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("g++ failed building _pe_torch.so")
    return TORCH_OUT


if __name__ == "__main__":
    dest = sys.argv[sys.argv.index("-o") + 1] if "-o" in sys.argv else OUT
    print(build(force="-f" in sys.argv or "--force" in sys.argv, verbose="-v" in sys.argv, out=dest,
                defines=tuple(a for a in sys.argv[1:] if a.startswith("-D"))))
    if dest == OUT:
        print(build_torch_ext(force="-f" in sys.argv or "--force" in sys.argv, verbose="-v" in sys.argv))
