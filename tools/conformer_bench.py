"""BASELINE config 4 on one B200: Conformer (aishell recipe, 12 blocks, d = 256) encoder forward + CTC greedy over
synthetic batches -- throughput (audio s / wall s) with waveforms resident in HBM, end to end from pinned host
memory, the per-kernel-kind device times of one step (fb200_profile_*), and the oracle on the host cores.

    python tools/conformer_bench.py [--utts 256] [--steps 3] [--warmup 3] [--precision fp32|bf16]

Prints ONE JSON line (same keys as bench.py's; config 4 is a parity-test configuration, not the headline bench).
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from funasr_b200 import _lib, conformer, params, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--utts", type=int, default=256)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--precision", default="fp32", choices=["fp32", "bf16"])
    ap.add_argument("--max-rows", type=int, default=18944)
    ap.add_argument("--cpu-utts", type=int, default=6)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.conformer_ctc_shapes(), seed=0).items()}
    eng = conformer.ConformerCTCEngine(sd, dev, args.precision, max_rows=args.max_rows)
    durs = synth.synth_durations(args.utts)
    waves = [synth.synth_waveform(1000 + i, int(n)) for i, n in enumerate(durs)]
    audio_s = float(sum(len(w) for w in waves)) / synth.SAMPLE_RATE
    batches = eng.plan_batches([len(w) for w in waves])
    resident = [eng.upload([waves[i] for i in b]) for b in batches]
    pinned = [torch.from_numpy(w).pin_memory() for w in waves]
    torch.cuda.synchronize()

    def step_resident():
        for wave, lens in resident:
            eng.forward_packed(wave, lens)

    def timed(fn, steps):
        beg, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        beg.record()
        for _ in range(steps):
            fn()
        end.record()
        torch.cuda.synchronize()
        return beg.elapsed_time(end) / steps

    for _ in range(args.warmup):
        step_resident()
    _lib.launch_count(reset=True)
    ms = timed(step_resident, args.steps)
    launches = _lib.launch_count() // args.steps
    for _ in range(1):
        eng.recognize(pinned)
    ms_e2e = timed(lambda: eng.recognize(pinned), args.steps)
    _lib.profile_read(reset=True)
    _lib.profile_enable(True)
    step_resident()
    torch.cuda.synchronize()
    prof = _lib.profile_read(reset=True)
    _lib.profile_enable(False)
    gemm = prof.get("gemm", {"ms": 0.0, "work": 0.0, "calls": 0})
    terms = 3 if args.precision == "fp32" else 1
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("bf16_tflops_sustained", 1357.3))
    achieved = gemm["work"] / (gemm["ms"] * 1e-3) / 1e12 if gemm["ms"] else 0.0
    # oracle on the host cores, a bounded sample
    from oracle import conformer_oracle as C

    sample = waves[: args.cpu_utts]
    t0 = time.perf_counter()
    with torch.no_grad():
        for w in sample:
            C.recognize(torch.from_numpy(w), sd)
    cpu_s = time.perf_counter() - t0
    cpu_audio = float(sum(len(w) for w in sample)) / synth.SAMPLE_RATE
    line = {
        "metric": "Conformer-CTC RTFx (audio s / wall s)", "value": audio_s / (ms * 1e-3), "unit": "audio_s/s", "n_gpus": 1,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "dtype": "f32" if terms == 3 else "bf16",
        "data": "synthetic",
        "config": {"workload": "configs[3]: Conformer 12 blocks d=256 (aishell recipe) encoder forward + CTC greedy, %d synthetic "
                               "utterances U(2,30) s (%.0f s audio), packed var-len batches <= %d rows" % (args.utts, audio_s, args.max_rows),
                   "precision": args.precision, "batches_per_step": len(batches)},
        "e2e": {"value": audio_s / (ms_e2e * 1e-3), "unit": "audio_s/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": int(sum(w.nbytes for w in waves))},
        "gpu_launches": int(launches),
        "roofline": {"bound": "tensor", "kernel": "gemm_tcgen05_kernel", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                     "frac": achieved / peak, "tensor_terms_per_flop": terms, "executed_tflops": achieved * terms,
                     "launches": gemm["calls"]},
        "breakdown_ms": {k: round(v["ms"], 3) for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])},
        "cpu_baseline": {"value": cpu_audio / cpu_s, "unit": "audio_s/s", "cores": torch.get_num_threads(), "kind": "port",
                         "sample": "first %d utterances (%.1f s audio, %.1f s CPU), batch 1, oracle" % (len(sample), cpu_audio, cpu_s)},
    }
    print(json.dumps(line))


if __name__ == "__main__":
    main()
