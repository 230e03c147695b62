"""Throughput of the native C++ offline runtime (fb200_offline_infer, host waveforms in, token ids out) on the
headline workload (BASELINE configs[1]: 1024 synthetic utterances U(2,30) s), next to the Python engine's e2e figure.

    python tools/runtime_bench.py [--utts 1024] [--steps 3]
"""
import argparse
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from funasr_b200 import params, runtime, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--utts", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--precision", default="fp32")
    args = ap.parse_args()
    sd = synth.synth_state_dict(params.paraformer_large_shapes(), seed=0)
    d = tempfile.mkdtemp()
    runtime.export_model(sd, d, cmvn=synth.synth_cmvn())
    rec = runtime.OfflineRecognizer(d, precision=args.precision)
    durs = synth.synth_durations(args.utts)
    waves = [synth.synth_waveform(1000 + i, int(n)) for i, n in enumerate(durs)]
    audio_s = float(sum(len(w) for w in waves)) / synth.SAMPLE_RATE
    for _ in range(2):
        ids = rec.recognize(waves)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ids = rec.recognize(waves)
    dt = (time.perf_counter() - t0) / args.steps
    print(json.dumps({"metric": "Paraformer-large RTFx (audio s / wall s), native C++ runtime, host waveforms in / ids out",
                      "value": audio_s / dt, "unit": "audio_s/s", "ms_per_step": dt * 1e3, "utterances": args.utts,
                      "audio_s": audio_s, "precision": args.precision, "tokens": sum(len(x) for x in ids),
                      "timing": "host wall clock around fb200_offline_infer (it returns after the last device->host read)"}))


if __name__ == "__main__":
    main()
