"""The "existing GPU path" bar: the UNMODIFIED reference's PyTorch-eager path on the same B200 --
`AutoModel(..., device="cuda").generate(input=[...], batch_size=64)` (funasr/auto/auto_model.py:184-188,301-376;
funasr/models/paraformer/model.py:443-596), fp32 and bf16, padded batches -- over BASELINE configs[1]'s 1024 utterances,
next to this repo's engine on the same box in the same process.

    python tools/reference_gpu_bench.py [--utts 1024] [--batch-size 64] [--steps 2]

The reference is given its best case: the utterances are sorted by length before batching (least padding; its own VAD
pipeline does the same, auto_model.py:423-459) and TF32 is left at torch's default for the fp32 run (off) and switched on
for a second fp32 run.  Not a driver arm (the driver's reference arm is `bench.py --impl reference`, the CPU path); the
line is recorded in BASELINE.md section 3 / profiles/.  Prints ONE JSON line."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from funasr_b200 import engine, params, schedule, synth  # noqa: E402
from oracle import reference_runner as R  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--utts", type=int, default=1024)
    ap.add_argument("--batch-size", type=int, default=64)
    ap.add_argument("--steps", type=int, default=2)
    args = ap.parse_args()
    durs = [int(n) for n in synth.synth_durations(args.utts)]
    order = sorted(range(len(durs)), key=lambda i: -durs[i])
    waves = [synth.synth_waveform(i, durs[i]) for i in order]
    secs = sum(schedule.encoder_rows(n) for n in durs) * 0.06
    out = {"workload": "configs[1]: %d synthetic utterances U(2,30) s (%.0f s audio), sorted by length, padded batches of %d"
                       % (args.utts, secs, args.batch_size), "unit": "audio_s/s"}

    def timed(fn):
        fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            res = fn()
        torch.cuda.synchronize()
        return secs * args.steps / (time.perf_counter() - t0), res

    ref = R.build_paraformer("cuda")
    assert type(ref.model).__module__.startswith("funasr.models.paraformer") and next(ref.model.parameters()).is_cuda
    v, ids_ref = timed(lambda: R.token_ids(ref, waves, args.batch_size))
    out["reference_eager_fp32"] = v
    torch.backends.cuda.matmul.allow_tf32 = True
    torch.backends.cudnn.allow_tf32 = True
    out["reference_eager_fp32_tf32"], _ = timed(lambda: R.token_ids(ref, waves, args.batch_size))
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    del ref
    torch.cuda.empty_cache()
    try:
        ref16 = R.build_paraformer("cuda", bf16=True)
        out["reference_eager_bf16"], _ = timed(lambda: R.token_ids(ref16, waves, args.batch_size))
        del ref16
    except Exception as e:  # the reference's bf16 switch casts the model only (auto_model.py:281-282); inputs stay fp32
        out["reference_eager_bf16"] = None
        out["reference_eager_bf16_error"] = "%s: %s" % (type(e).__name__, str(e)[:200])
    torch.cuda.empty_cache()
    sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.paraformer_large_shapes(), seed=0).items()}
    cm = tuple(torch.from_numpy(x) for x in synth.synth_cmvn())
    pinned = [torch.from_numpy(w).pin_memory() for w in waves]
    for prec in ("fp32", "bf16"):
        eng = engine.ParaformerEngine(sd, "cuda:0", prec, cmvn=cm, max_rows=18944)
        v, ids = timed(lambda: eng.recognize(pinned))
        out["funasr_b200_%s" % prec] = v
        if prec == "fp32":
            # the reference's PADDED batches leak padding into the CIF conv, so its batched ids are not
            # its own batch-1 ids; agreement is reported, the parity gate is tests/test_gpu_parity_gate.py (batch 1)
            out["ids_identical_to_reference_eager_batched"] = int(sum(1 for a, b in zip(ids, ids_ref) if a == b))
        del eng
        torch.cuda.empty_cache()
    out["speedup_vs_reference_eager_fp32"] = out["funasr_b200_fp32"] / out["reference_eager_fp32"]
    out["gpu"] = torch.cuda.get_device_name(0)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
