"""Per-stage / per-kernel-kind device time of one packed batch (CUDA events, no profiler).

    python tools/stage_profile.py [--utts 80] [--reps 5]

Prints, for the frontend, encoder, predictor and decoder of ParaformerEngine.forward_packed, the wall device
time of the stage (events around the stage, steady state) and the per-kind kernel times measured by the
library's own profile hooks (fb200_profile_*), plus the row / token counts -- the basis for the "which kernel
is off its roofline" table in DESIGN.md.
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from funasr_b200 import _lib, engine, params, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--utts", type=int, default=80)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--max-rows", type=int, default=18944)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.paraformer_large_shapes(), seed=0).items()}
    shift, scale = synth.synth_cmvn()
    eng = engine.ParaformerEngine(sd, dev, precision="fp32", cmvn=(torch.from_numpy(shift), torch.from_numpy(scale)),
                                  max_rows=args.max_rows)
    durs = synth.synth_durations(args.utts)
    waves = [synth.synth_waveform(1000 + i, int(n)) for i, n in enumerate(durs)]
    batch = eng.plan_batches([len(w) for w in waves])[0]
    wave, lens = eng.upload([waves[i] for i in batch])
    torch.cuda.synchronize()

    def stages(profile):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
        prof = {}

        def mark(i, name=None):
            if profile and name:
                torch.cuda.synchronize()
                prof[name] = _lib.profile_read(reset=True)
            ev[i].record()

        mark(0)
        feats, cu_rows, rows = eng.frontend.run(wave, lens, pe_scale=float(eng.encoder.d) ** 0.5)
        mark(1, "frontend")
        b, r = len(lens), feats.shape[0]
        halo = eng.predictor.halo_buffer(r, b)
        enc = eng.encoder.run(feats, cu_rows, max(rows), out_halo=halo)
        mark(2, "encoder")
        pred = eng.predictor.run(enc, halo, cu_rows, b)
        mark(3, "predictor")
        tok = pred["token_num"]
        eng.decoder.run(enc, cu_rows, pred["embeds"], pred["cu_tok"], max(tok))
        mark(4, "decoder")
        torch.cuda.synchronize()
        return [ev[i].elapsed_time(ev[i + 1]) for i in range(4)], prof, r, sum(tok), b

    for _ in range(3):
        stages(False)
    acc = [0.0] * 4
    for _ in range(args.reps):
        ms, _, r, l, b = stages(False)
        acc = [a + m for a, m in zip(acc, ms)]
    _lib.profile_read(reset=True)
    _lib.profile_enable(True)
    _, prof, r, l, b = stages(True)
    _lib.profile_enable(False)
    print("batch: %d utterances, %d encoder rows, %d tokens" % (b, r, l))
    for name, a in zip(("frontend", "encoder", "predictor", "decoder"), acc):
        kinds = prof[name]
        ksum = sum(v["ms"] for v in kinds.values())
        print("%-10s %8.3f ms/stage   (sum of kernels %.3f ms)" % (name, a / args.reps, ksum))
        for k, v in sorted(kinds.items(), key=lambda kv: -kv[1]["ms"]):
            print("    %-12s %8.3f ms  %5d calls  %7.1f us/call" % (k, v["ms"], v["calls"], 1e3 * v["ms"] / v["calls"]))


if __name__ == "__main__":
    main()
