"""`torch.ops.funasr_b200.*` -- the torch custom-op layer over the C-ABI (csrc/torch_shim.cpp, TORCH_LIBRARY).

    from funasr_b200 import torch_ops
    ops = torch_ops.load()                       # torch.ops.funasr_b200
    ops.layernorm(x, gamma, beta, 1e-12, None, xs, 2, 0, None)

Same kernels, same caller-owned buffers and the same error behaviour as the ctypes binding (`_lib.py`): the shim only
unwraps tensors into pointers / pitches, takes the current CUDA stream and raises on a non-zero return code.  The ops are
visible to the torch dispatcher (schemas with mutable-argument annotations), which is what a FunASR module written against
`torch.ops` would call; `funasr_b200/ops.py` keeps the ctypes route (no torch types across the boundary)."""
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfunasr_b200_torch.so")
_loaded = False


def load():
    global _loaded
    if not _loaded:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("funasr_b200: %s not found -- run `python -m funasr_b200.build` (there is no fallback path)" % LIB_PATH)
        torch.ops.load_library(LIB_PATH)
        _loaded = True
    return torch.ops.funasr_b200
