"""TEST INFRASTRUCTURE: builds (g++) and loads the host instantiation of csrc/pe_math.cuh."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "hostmath.cpp")
HDRS = [os.path.join(HERE, "..", "..", "poseestimation_b200", "csrc", h) for h in ("pe_math.cuh", "pe_lanes.cuh")]
OUT = os.path.join(HERE, "_hostmath.so")
CUDA_INC = os.environ.get("CUDA_HOME", "/usr/local/cuda") + "/include"


def load():
    if not os.path.exists(OUT) or os.path.getmtime(OUT) < max(os.path.getmtime(f) for f in [SRC] + HDRS):
        # -ffp-contract=off: only the explicit fmaf() calls fuse, as in the device code's explicit FMAs
        subprocess.run(["g++", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-I", CUDA_INC, "-x", "c++", SRC,
                        "-o", OUT], check=True)
    return ctypes.CDLL(OUT)


_FP = ctypes.POINTER(ctypes.c_float)


def call(lib, name, *arrays_and_outshapes):
    """call(lib, 'hm_quat', x, (n, 9)) -> out ; inputs are float32 C-contiguous arrays, tuples are output shapes."""
    args, outs, n, keep = [], [], None, []
    for a in arrays_and_outshapes:
        if isinstance(a, tuple):
            o = np.empty(a, np.float32)
            outs.append(o)
            args.append(o.ctypes.data_as(_FP))
        else:
            a = np.ascontiguousarray(a, np.float32)
            n = a.shape[0] if n is None else n
            outs.append(None)
            args.append(a.ctypes.data_as(_FP))
            keep.append(a)
    getattr(lib, name)(*args, ctypes.c_long(n))
    res = [o for o in outs if o is not None]
    return res[0] if len(res) == 1 else res
