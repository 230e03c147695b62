#!/usr/bin/env python
"""Timeline of one CTA of attention_tcgen05_kernel (library built with -DFB200_TRACE; debugging aid).
usage: python tools/attn_trace.py [T] [nparts]"""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from funasr_b200 import _lib, ops  # noqa: E402

# the instrumented build lives next to the product library (tools/build_trace.sh)
_lib.LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), "libfunasr_b200_trace.so")

DEV = "cuda:0"
T = int(sys.argv[1]) if len(sys.argv) > 1 else 490
NP = int(sys.argv[2]) if len(sys.argv) > 2 else 2
NAMES = {1: "start", 3: "Q ready", 4: "first S issued", 11: "P_G arrived", 12: "V_G / O free", 13: "PV issued",
         14: "S issued", 20: "tile top", 21: "S_G ready", 22: "S in regs", 23: "max/rescale done", 24: "P packed",
         25: "bar_o(G-2) waited", 26: "arrived", 40: "K,V loads issued", 50: "Q_0 stored", 51: "Q_n+1 stored",
         52: "O final + l", 53: "O in regs", 54: "output issued", 60: "P_G seen", 61: "V_G seen", 62: "O free seen", 63: "K_G+2 seen", 55: "box packed", 56: "box free", 99: "end"}
ROLES = ["softmax", "output", "producer", "mma", "pv.l0", "pv.l1", "pv.l2", "s-issue"]


def main():
    g = torch.Generator().manual_seed(0)
    lens = [T] * (18944 // T)
    rows = sum(lens)
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=DEV)
    kp = 1536
    qkv = ops.split_rows(torch.randn(rows, kp, generator=g).to(DEV), NP)
    out = torch.empty(rows, ops.split_width(512, NP), dtype=torch.bfloat16, device=DEV)
    for _ in range(3):
        ops.attention(qkv, 0, kp, qkv, 512, qkv, 1024, kp, NP, cu, cu, max(lens), 4, 128 ** -0.5, out_split=out,
                      out_parts=NP)
    torch.cuda.synchronize()
    buf = (ctypes.c_longlong * 4096)()
    lib = _lib.load()
    lib.fb200_debug_attn_trace.restype = ctypes.c_int
    assert lib.fb200_debug_attn_trace(buf) == 0
    ev = []
    for w in range(8):
        for k in range(256):
            tag, t = buf[w * 512 + 2 * k], buf[w * 512 + 2 * k + 1]
            if tag:
                ev.append((t, w, tag))
    ev.sort()
    t0 = ev[0][0]
    last = {w: t0 for w in range(8)}
    for t, w, tag in ev:
        print("%8d  (+%6d)  %s%-8s %s" % (t - t0, t - last[w], " " * (22 * min(w, 4)), ROLES[w], NAMES.get(tag, tag)))
        last[w] = t


if __name__ == "__main__":
    main()
