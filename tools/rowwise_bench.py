"""HBM-bound row kernels at the full-batch shape (18944 x 512 fp32): device time (CUDA events, L2 flushed by
rotating over buffers larger than L2) and achieved GB/s against the algorithmic bytes.

    python tools/rowwise_bench.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from funasr_b200 import ops  # noqa: E402

DEV = torch.device("cuda:0")


def timeit(fn, nbuf, iters=40):
    for i in range(nbuf):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i % nbuf)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3  # us


def main():
    r, d, nbuf = 18944, 512, 6  # 6 x (3 x 38.8 MB) > 126 MB L2
    lens = [r // 56] * 55
    lens.append(r - sum(lens))
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=DEV)
    v = [torch.randn(r, d, device=DEV) for _ in range(nbuf)]
    x = [torch.randn(r, d, device=DEV) for _ in range(nbuf)]
    o = [torch.empty(r, d, device=DEV) for _ in range(nbuf)]
    w = torch.randn(d, 11, device=DEV)
    us = timeit(lambda i: ops.fsmn(v[i], cu, w, o[i], residual=x[i]), nbuf)
    print("fsmn k=11 + residual   %7.1f us  %6.0f GB/s (algorithmic 3 x R x 512 x 4 B)" % (us, 3 * r * d * 4 / us / 1e3))
    us = timeit(lambda i: ops.fsmn(v[i], cu, w, o[i]), nbuf)
    print("fsmn k=11              %7.1f us  %6.0f GB/s (2 x R x 512 x 4 B)" % (us, 2 * r * d * 4 / us / 1e3))
    g, b = torch.ones(d, device=DEV), torch.zeros(d, device=DEV)
    sp = [torch.empty(r, ops.split_width(d, 2), dtype=torch.bfloat16, device=DEV) for _ in range(nbuf)]
    us = timeit(lambda i: ops.layernorm(x[i], g, b, 1e-12, out_split=sp[i], nparts=2), nbuf)
    print("layernorm -> split(2)  %7.1f us  %6.0f GB/s (R x 512 x (4 + 4) B)" % (us, 2 * r * d * 4 / us / 1e3))
    us = timeit(lambda i: ops.layernorm(x[i], g, b, 1e-12, out_f32=o[i]), nbuf)
    print("layernorm -> fp32      %7.1f us  %6.0f GB/s (R x 512 x (4 + 4) B)" % (us, 2 * r * d * 4 / us / 1e3))
    h = [torch.randn(r // 2, 2048, device=DEV) for _ in range(nbuf)]
    hs = [torch.empty(r // 2, ops.split_width(2048, 2), dtype=torch.bfloat16, device=DEV) for _ in range(nbuf)]
    g2, b2 = torch.ones(2048, device=DEV), torch.zeros(2048, device=DEV)
    us = timeit(lambda i: ops.layernorm(h[i], g2, b2, 1e-12, out_split=hs[i], nparts=2), nbuf)
    print("layernorm 2048 split   %7.1f us  %6.0f GB/s (R/2 x 2048 x 8 B)" % (us, r // 2 * 2048 * 8 / us / 1e3))
    us = timeit(lambda i: o[i].copy_(x[i]), nbuf)
    print("torch copy (reference) %7.1f us  %6.0f GB/s" % (us, 2 * r * d * 4 / us / 1e3))


if __name__ == "__main__":
    main()
