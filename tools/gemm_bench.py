#!/usr/bin/env python
"""Per-shape timing of fb200_gemm on the GPU (CUDA events, L2 flushed between shapes by size).
usage: python tools/gemm_bench.py [nparts]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from funasr_b200 import ops  # noqa: E402

DEV = "cuda:0"
NP = int(sys.argv[1]) if len(sys.argv) > 1 else 2
TERMS = {1: 1, 2: 3, 3: 6}[NP]
M = int(sys.argv[2]) if len(sys.argv) > 2 else 18944
# (name, n, k, f32 out, split out, residual, relu, f32_col0)
SHAPES = [
    ("qkv   ", 1536, 512, True, True, False, False, 1024),
    ("out   ", 512, 512, True, False, True, False, 0),
    ("ffn_w1", 2048, 512, False, True, False, True, 0),
    ("ffn_w2", 512, 2048, True, False, True, False, 0),
    ("plain ", 2048, 2048, True, False, False, False, 0),
    ("vocab ", 8404, 512, True, False, False, False, 0),
]


# Conformer (d = 256) shapes and epilogue variants: `python tools/gemm_bench.py 2 18944 conformer`
CONFORMER = [
    ("c_w1 swish split", 2048, 256, False, True, False, "swish", 0),
    ("c_w1 none  split", 2048, 256, False, True, False, False, 0),
    ("c_w1 none  f32  ", 2048, 256, True, False, False, False, 0),
    ("c_w2 res   f32  ", 256, 2048, True, False, True, False, 0),
    ("c_qkv      split", 1024, 256, False, True, False, False, 0),
    ("c_out res  f32  ", 256, 256, True, False, True, False, 0),
    ("c_pw1      f32  ", 512, 256, True, False, False, False, 0),
    ("c_bd  k64  f32  ", 752, 64, True, False, False, False, 0),
    ("c_ctc      f32  ", 4236, 256, True, False, False, False, 0),
]


def main():
    g = torch.Generator().manual_seed(0)
    shapes = CONFORMER if len(sys.argv) > 3 and sys.argv[3] == "conformer" else SHAPES
    for name, n, k, f32, split, res, relu, c0 in shapes:
        m = M if name != "vocab " else 5000
        a = ops.split_rows(torch.randn(m, k, generator=g).to(DEV), NP)
        w = ops.split_rows((torch.randn(n, k, generator=g) / k ** 0.5).to(DEV), NP)
        bias = torch.randn(n, generator=g).to(DEV)
        out = torch.empty(m, n - c0, device=DEV) if f32 else None
        sp = torch.empty(m, ops.split_width(n, NP), dtype=torch.bfloat16, device=DEV) if split else None
        r = torch.randn(m, n, generator=g).to(DEV) if res else None

        def run():
            if relu == "swish":
                ops.gemm_ex(a, NP, w, NP, m, n, k, bias=bias, act="swish", res1=r, out_f32=out, out_split=sp, out_parts=NP)
            else:
                ops.gemm(a, NP, w, NP, m, n, k, bias=bias, relu=relu, res1=r, out_f32=out, out_split=sp, out_parts=NP,
                         f32_col0=c0)

        for _ in range(3):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 20
        e0.record()
        for _ in range(iters):
            run()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / iters
        fl = 2.0 * m * n * k
        byts = a.numel() * 2 + w.numel() * 2 + (out.numel() * 4 if f32 else 0) + (sp.numel() * 2 if split else 0) + \
            (r.numel() * 4 if res else 0)
        print("%s m=%5d n=%4d k=%4d  %8.1f us  algo %6.1f TF/s  executed %6.1f TF/s  min-bytes %6.1f MB -> %5.0f GB/s"
              % (name, m, n, k, us, fl / us / 1e6, TERMS * fl / us / 1e6, byts / 1e6, byts / us / 1e3))


if __name__ == "__main__":
    main()
