"""Write the model directory of the native offline runtime (model.fb200, am.mvn, tokens.json) from a checkpoint.

    python tools/export_model.py --model-pt model.pt --am-mvn am.mvn --tokens tokens.json --out model_dir
    python tools/export_model.py --synthetic --out model_dir        # seeded random-init Paraformer-large

`model.pt` is a torch state_dict (or a dict with a "state_dict" / "model" entry) under the reference's parameter names.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from funasr_b200 import params, runtime, synth, tables  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model-pt")
    ap.add_argument("--am-mvn")
    ap.add_argument("--tokens")
    ap.add_argument("--synthetic", action="store_true")
    ap.add_argument("--window", default="hamming")
    ap.add_argument("--out", required=True)
    args = ap.parse_args()
    if args.synthetic:
        sd = synth.synth_state_dict(params.paraformer_large_shapes(), seed=0)
        cmvn = synth.synth_cmvn()
        tokens = synth.synth_token_list()
    else:
        sd = torch.load(args.model_pt, map_location="cpu")
        for key in ("state_dict", "model"):
            if isinstance(sd, dict) and key in sd and isinstance(sd[key], dict):
                sd = sd[key]
        cmvn = tables.load_cmvn(args.am_mvn) if args.am_mvn else None
        tokens = json.load(open(args.tokens, encoding="utf-8")) if args.tokens else None
    print(runtime.export_model(sd, args.out, cmvn=cmvn, token_list=tokens, window=args.window))


if __name__ == "__main__":
    main()
