#!/usr/bin/env python
"""Timeline of CTA 0 of gemm_tcgen05_kernel (library built by tools/build_trace.sh; debugging aid).
usage: python tools/gemm_trace.py n k [res] [nparts] [f32|split|both] [relu]   (add "summary" for per-tag means only)"""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from funasr_b200 import _lib, ops  # noqa: E402

_lib.LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), "libfunasr_b200_trace.so")
DEV = "cuda:0"
N, K = int(sys.argv[1]), int(sys.argv[2])
RES = len(sys.argv) > 3 and sys.argv[3] == "1"
NP = int(sys.argv[4]) if len(sys.argv) > 4 else 2
M = 18944
MODE = sys.argv[5] if len(sys.argv) > 5 else "f32"
RELU = len(sys.argv) > 6 and sys.argv[6] == "1"
SUMMARY = "summary" in sys.argv
NAMES = {1: "stage free -> loads", 10: "acc buffer free", 11: "stage full", 12: "k-block issued", 20: "acc ready", 21: "chunk stored",
         22: "chunk in registers", 23: "bias / act / residual done", 24: "boxes free (wait_group.read)", 25: "boxes written + fenced", 26: "acc handed back (last chunk)", 27: "bias added"}
ROLES = ["producer", "mma", "epilogue"]


def main():
    g = torch.Generator().manual_seed(0)
    a = ops.split_rows(torch.randn(M, K, generator=g).to(DEV), NP)
    w = ops.split_rows((torch.randn(N, K, generator=g) / K ** 0.5).to(DEV), NP)
    out = torch.empty(M, N, device=DEV) if MODE in ("f32", "both") else None
    sp = torch.empty(M, ops.split_width(N, NP), dtype=torch.bfloat16, device=DEV) if MODE in ("split", "both") else None
    bias = torch.randn(N, generator=g).to(DEV)
    r = torch.randn(M, N, generator=g).to(DEV) if RES else None
    for _ in range(3):
        ops.gemm(a, NP, w, NP, M, N, K, bias=bias, relu=RELU, res1=r, out_f32=out, out_split=sp, out_parts=NP)
    torch.cuda.synchronize()
    buf = (ctypes.c_longlong * 3072)()
    lib = _lib.load()
    lib.fb200_debug_gemm_trace.restype = ctypes.c_int
    assert lib.fb200_debug_gemm_trace(buf) == 0
    ev = []
    for w_ in range(3):
        for k in range(512):
            tag, t = buf[w_ * 1024 + 2 * k], buf[w_ * 1024 + 2 * k + 1]
            if tag:
                ev.append((t, w_, tag))
    ev.sort()
    t0 = ev[0][0]
    last = {w_: t0 for w_ in range(3)}
    if SUMMARY:
        import collections

        acc = collections.defaultdict(list)
        for t, w_, tag in ev:
            acc[(w_, tag)].append(t - last[w_])
            last[w_] = t
        for (w_, tag), v in sorted(acc.items()):
            v = v[len(v) // 4:]
            print("%-9s %-30s n=%4d  mean %7.0f  median %7.0f" % (ROLES[w_], NAMES.get(tag, tag), len(v), sum(v) / len(v), sorted(v)[len(v) // 2]))
        return
    for t, w_, tag in ev:
        print("%8d  (+%6d)  %s%-8s %s" % (t - t0, t - last[w_], " " * (26 * w_), ROLES[w_], NAMES.get(tag, tag)))
        last[w_] = t


if __name__ == "__main__":
    main()
