#!/usr/bin/env python
"""Per-shape timing of fb200_attention (self-attention over a packed qkv split) with CUDA events.
usage: python tools/attn_bench.py [nparts] [iters]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from funasr_b200 import _lib, ops  # noqa: E402

if os.environ.get("FB200_LIB"):  # an experimental build of the library next to the shipped one (tools only)
    _lib.LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), os.environ["FB200_LIB"])

DEV = "cuda:0"
NP = int(sys.argv[1]) if len(sys.argv) > 1 else 2
ITERS = int(sys.argv[2]) if len(sys.argv) > 2 else 20
ROWS = 18944
TERMS = {1: 1, 2: 3}[NP]


def lens_for(kind):
    g = torch.Generator().manual_seed(1)
    if isinstance(kind, int):
        n = ROWS // kind
        return [kind] * n
    lens, tot = [], 0
    while True:
        t = int(torch.randint(33, 501, (1,), generator=g))
        if tot + t > ROWS:
            break
        lens.append(t)
        tot += t
    return sorted(lens, reverse=True)


def main():
    g = torch.Generator().manual_seed(0)
    for kind in (490, 256, 128, 64, 33, "mix"):
        lens = lens_for(kind)
        rows = sum(lens)
        cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=DEV)
        kp = 1536
        qkv = ops.split_rows(torch.randn(rows, kp, generator=g).to(DEV), NP)
        out = torch.empty(rows, ops.split_width(512, NP), dtype=torch.bfloat16, device=DEV)

        def run():
            ops.attention(qkv, 0, kp, qkv, 512, qkv, 1024, kp, NP, cu, cu, max(lens), 4, 128 ** -0.5, out_split=out,
                          out_parts=NP)

        for _ in range(3):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(ITERS):
            run()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / ITERS
        tiles = sum(4 * ((t + 127) // 128) * ((t + 63) // 64) for t in lens)
        ctas = sum(4 * ((t + 127) // 128) for t in lens)
        fl = sum(4.0 * t * t * 512 for t in lens)
        print("T=%-4s utts=%4d rows=%5d ctas=%5d tiles=%6d  %8.1f us  %6.3f us/tile/SM  algo %6.1f TF/s  executed(padded) %6.1f TF/s"
              % (kind, len(lens), rows, ctas, tiles, us, us * 148 / tiles, fl / us / 1e6,
                 TERMS * tiles * 4.0 * 128 * 64 * 128 / us / 1e6))


if __name__ == "__main__":
    main()
