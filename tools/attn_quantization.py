#!/usr/bin/env python
"""How much of the attention's tile time the round structure of its items costs on the headline set (CPU only, no GPU):
every batch's items (utterance x head x 128-row query block, cost = its 64-key tiles, at least the ~2.3 tile-times the
output warps need per item) dealt round-robin to 148 CTAs as the kernel does, against a longest-first dynamic deal and the
ideal (sum / 148).   usage: python tools/attn_quantization.py [max_rows]"""
import heapq
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from funasr_b200 import schedule, synth  # noqa: E402

G = 148
MAX_ROWS = int(sys.argv[1]) if len(sys.argv) > 1 else 56832


def main():
    durs = [int(x) for x in synth.synth_durations(1024)]
    tot_rr = tot_lpt = tot_ideal = 0.0
    for b in schedule.batches(durs, list(range(len(durs))), MAX_ROWS):
        items = []
        for i in sorted(b, key=lambda i: -durs[i]):
            t = schedule.encoder_rows(durs[i])
            items += [max((t + 63) // 64, 2.3)] * (4 * ((t + 127) // 128))
        loads = [0.0] * G
        for k, c in enumerate(items):
            loads[k % G] += c
        heap = [0.0] * G
        for c in sorted(items, reverse=True):
            heapq.heapreplace(heap, heap[0] + c)
        rr, lpt, ideal = max(loads), max(heap), sum(items) / G
        tot_rr, tot_lpt, tot_ideal = tot_rr + rr, tot_lpt + lpt, tot_ideal + ideal
        print("batch of %3d utterances, %4d items: round-robin %.1f  longest-first %.1f  ideal %.1f tile-times" % (len(b), len(items), rr, lpt, ideal))
    print("set: round-robin %.0f, longest-first %.0f, ideal %.0f  ->  %.3f / %.3f of the ideal" % (tot_rr, tot_lpt, tot_ideal, tot_rr / tot_ideal, tot_lpt / tot_ideal))


if __name__ == "__main__":
    main()
