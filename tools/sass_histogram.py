#!/usr/bin/env python
"""Opcode histogram per kernel of the shipped library (`cuobjdump -sass funasr_b200/libfunasr_b200.so`): shows which kernels
are Blackwell-native -- UTCHMMA / UTCHMMA.2CTA (tcgen05.mma), LDTM / STTM (tcgen05.ld / st), UTMALDG / UTMASTG (TMA), UBLKCP
(bulk copy), SYNCS (mbarrier) -- and which run on the legacy pipes (HMMA = mma.sync).  Runs without a GPU.

    python tools/sass_histogram.py > profiles/r02_sass_histogram.md"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "funasr_b200", "libfunasr_b200.so")
KEY = ["UTCHMMA", "UTCHMMA.2CTA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "SYNCS", "HMMA", "MUFU", "FFMA",
       "LDG", "STG", "LDS", "STS", "LDGSTS", "BAR", "ACQBULK"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            cur = re.sub(r"\(.*", "", cur).replace("fb200::", "")
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
        if m and cur:
            op = m.group(1)
            base = op.split(".")[0]
            kernels[cur][base] += 1
            if op.startswith("UTCHMMA.2CTA"):
                kernels[cur]["UTCHMMA.2CTA"] += 1
    print("# SASS opcode histogram per kernel (`cuobjdump -sass funasr_b200/libfunasr_b200.so`, sm_100a)\n")
    print("Static instruction counts.  UTCHMMA = tcgen05.mma (of which .2CTA = cta_group::2), LDTM / STTM = tcgen05.ld / st, "
          "UTMALDG / UTMASTG = TMA tensor loads / stores, UBLKCP = cp.async.bulk, SYNCS = mbarrier ops, HMMA = legacy mma.sync.\n")
    print("| kernel | total | " + " | ".join(KEY) + " |")
    print("|---|---|" + "---|" * len(KEY))
    for name, c in kernels.items():
        print("| `%s` | %d | " % (name[:90], sum(v for k, v in c.items() if k != "UTCHMMA.2CTA")) + " | ".join(str(c.get(k, 0) or "") for k in KEY) + " |")


if __name__ == "__main__":
    sys.exit(main())
