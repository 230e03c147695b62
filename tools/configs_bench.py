"""BASELINE configs 3 and 5 on one B200 (parity-test configurations; the headline bench is bench.py = configs[1]).

    python tools/configs_bench.py --config 3     # SenseVoiceSmall (50 + 20 SAN-M blocks, V = 25055), CTC greedy
    python tools/configs_bench.py --config 5     # Paraformer-large long-audio sweep: 60 / 120 / 300 / 600 s x 8 each
    python tools/configs_bench.py --config 6     # BiCifParaformer: ids + time stamps
    python tools/configs_bench.py --config 7     # SeacoParaformer (what `paraformer-zh` resolves to) with 20 hot words

Prints ONE JSON line: throughput with waveforms resident in HBM, end to end from pinned host memory, per-kernel-kind
device times of one step (fb200_profile_*), and the oracle timed on the host cores over a bounded sample.
"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from funasr_b200 import _lib, engine, params, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, required=True, choices=[3, 5, 6, 7])
    ap.add_argument("--utts", type=int, default=256)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--max-rows", type=int, default=18944)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    shift, scale = synth.synth_cmvn()
    cmvn = (torch.from_numpy(shift), torch.from_numpy(scale))
    from oracle import paraformer_oracle as O

    if args.config == 3:
        sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.sense_voice_small_shapes(), seed=0).items()}
        eng = engine.SenseVoiceEngine(sd, dev, "fp32", cmvn=cmvn, max_rows=args.max_rows)
        durs = synth.synth_durations(args.utts)
        name = "configs[2]: SenseVoiceSmall (50 + 20 SAN-M blocks, V = 25055) CTC greedy, %d synthetic utterances U(2,30) s" % args.utts
        metric = "SenseVoiceSmall RTFx (audio s / wall s)"
        cpu_fn, cpu_n = (lambda w: O.sense_voice_recognize(torch.from_numpy(w), sd, cmvn)), 4
    elif args.config == 6:
        from funasr_b200 import bicif
        from oracle import bicif_oracle as BO

        sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.bicif_paraformer_shapes(), seed=0).items()}
        eng = bicif.BiCifParaformerEngine(sd, dev, "fp32", cmvn=cmvn, max_rows=args.max_rows)
        durs = synth.synth_durations(args.utts)
        name = "BiCifParaformer (SURVEY 8f rank 3): Paraformer-large + CifPredictorV3 time-stamp head, %d synthetic utterances U(2,30) s" % args.utts
        metric = "BiCifParaformer RTFx (audio s / wall s), token ids + time stamps"
        cpu_fn, cpu_n = (lambda w: BO.recognize(torch.from_numpy(w), sd, cmvn)), 2
    elif args.config == 7:
        import numpy as np

        from funasr_b200 import seaco
        from oracle import seaco_oracle as SO

        sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.seaco_paraformer_shapes(), seed=0).items()}
        eng = seaco.SeacoParaformerEngine(sd, dev, "fp32", cmvn=cmvn, max_rows=args.max_rows)
        rng = np.random.default_rng(5)
        hot = [rng.integers(3, 8000, size=int(rng.integers(2, 5))).tolist() for _ in range(20)]
        plain = eng.recognize
        eng.recognize = lambda src: plain(src, hotwords=hot)
        durs = synth.synth_durations(args.utts)
        name = "SeacoParaformer (SURVEY 8f rank 3) with 20 hot words of 2-4 tokens: BiCifParaformer + bias-encoder LSTM + SeACo decoder, %d synthetic utterances U(2,30) s" % args.utts
        metric = "SeacoParaformer RTFx (audio s / wall s), 20 hot words, token ids + time stamps"
        cpu_fn, cpu_n = (lambda w: SO.recognize(torch.from_numpy(w), sd, cmvn, hotwords=hot + [[1]])), 2
    else:
        sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.paraformer_large_shapes(), seed=0).items()}
        eng = engine.ParaformerEngine(sd, dev, "fp32", cmvn=cmvn, max_rows=args.max_rows)
        durs = [int(s * synth.SAMPLE_RATE) for s in (60, 120, 300, 600) for _ in range(8)]
        name = "configs[4]: Paraformer-large long-audio sweep, 8 segments each of 60 / 120 / 300 / 600 s"
        metric = "Paraformer-large long-audio RTFx (audio s / wall s)"
        cpu_fn, cpu_n = (lambda w: O.recognize(torch.from_numpy(w), sd, cmvn)), 1
    waves = [synth.synth_waveform(2000 + i, int(n)) for i, n in enumerate(durs)]
    audio_s = float(sum(len(w) for w in waves)) / synth.SAMPLE_RATE
    pinned = [torch.from_numpy(w).pin_memory() for w in waves]
    resident = [w.to(dev) for w in pinned]
    torch.cuda.synchronize()

    def step(src):
        eng.recognize(src)

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
        step(resident)
    _lib.launch_count(reset=True)
    ms = timed(lambda: step(resident), args.steps)
    launches = _lib.launch_count() // args.steps
    step(pinned)
    ms_e2e = timed(lambda: step(pinned), args.steps)
    _lib.profile_read(reset=True)
    _lib.profile_enable(True)
    step(resident)
    torch.cuda.synchronize()
    prof = _lib.profile_read(reset=True)
    _lib.profile_enable(False)
    gemm = prof.get("gemm", {"ms": 0.0, "work": 0.0, "calls": 0})
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("bf16_tflops_sustained", 1357.3))
    achieved = gemm["work"] / (gemm["ms"] * 1e-3) / 1e12 if gemm["ms"] else 0.0
    sample = waves[:cpu_n]
    t0 = time.perf_counter()
    with torch.no_grad():
        for w in sample:
            cpu_fn(w)
    cpu_s = time.perf_counter() - t0
    cpu_audio = float(sum(len(w) for w in sample)) / synth.SAMPLE_RATE
    line = {
        "metric": metric, "value": audio_s / (ms * 1e-3), "unit": "audio_s/s", "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "%s (%.0f s audio), packed var-len batches <= %d rows" % (name, audio_s, args.max_rows),
                   "precision": "fp32-parity (hi + mid bf16 operands, 3 tcgen05 terms)"},
        "e2e": {"value": audio_s / (ms_e2e * 1e-3), "unit": "audio_s/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": int(sum(w.nbytes for w in waves))},
        "gpu_launches": int(launches),
        "roofline": {"bound": "tensor", "kernel": "gemm_tcgen05_kernel", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                     "frac": achieved / peak, "tensor_terms_per_flop": 3, "executed_tflops": achieved * 3, "launches": gemm["calls"]},
        "breakdown_ms": {k: round(v["ms"], 3) for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])},
        "cpu_baseline": {"value": cpu_audio / cpu_s, "unit": "audio_s/s", "cores": torch.get_num_threads(), "kind": "port",
                         "sample": "first %d utterance(s) (%.1f s audio, %.1f s CPU), batch 1, oracle" % (len(sample), cpu_audio, cpu_s)},
    }
    print(json.dumps(line))


if __name__ == "__main__":
    main()
