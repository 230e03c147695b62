"""Numerics probe for a 2-term GEMM scheme: activations hi + mid, WEIGHTS ROUNDED TO ONE 11-bit-mantissa
part (fp16) -- two tensor passes per product instead of three.  Emulated on the shipped kernels by rounding every weight
matrix to fp16 before the hi + mid bf16 split (hi + mid then carries the 11 bits exactly), so the engine computes
(a_hi + a_mid) . fp16(w) up to the dropped mid.mid term.  Gate: token ids of all 1024 utterances of configs[1] against
the reference's golden ids (tests/golden/configs1_ids_1024.npz).  Prints one JSON line."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from funasr_b200 import engine, params, synth  # noqa: E402


def main():
    mant = int(sys.argv[1]) if len(sys.argv) > 1 else 11
    g = np.load(os.path.join(ROOT, "tests", "golden", "configs1_ids_1024.npz"))
    n = len(g["token_num"])
    sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.paraformer_large_shapes(), seed=0).items()}

    def round_mantissa(w, bits):  # keep `bits` significant bits (round to nearest even), exponent range of fp32
        if bits == 11:
            return w.half().float()
        i = w.view(torch.int32)
        drop = 24 - bits
        half = (1 << (drop - 1)) - 1 + ((i >> drop) & 1)
        return ((i + half) >> drop << drop).view(torch.float32)

    for k in list(sd):
        if sd[k].dim() >= 2 and k.endswith("weight") and "norm" not in k and "fsmn" not in k and "embed" not in k:
            sd[k] = round_mantissa(sd[k].clone(), mant)
    cm = tuple(torch.from_numpy(x) for x in synth.synth_cmvn())
    eng = engine.ParaformerEngine(sd, "cuda:0", "fp32", cmvn=cm, max_rows=18944)
    waves = [synth.synth_waveform(i, int(m)) for i, m in enumerate(g["num_samples"])]
    got = eng.recognize(waves)
    same = sum(1 for i in range(n) if got[i] == g["ids"][g["offsets"][i]: g["offsets"][i + 1]].tolist())
    cnt = sum(1 for i in range(n) if eng.last_token_num[i] == int(g["token_num"][i]))
    tok_same = sum(sum(1 for a, b in zip(got[i], g["ids"][g["offsets"][i]: g["offsets"][i + 1]].tolist()) if a == b) for i in range(n))
    print(json.dumps({"weight_mantissa_bits": mant, "utterances": n, "ids_identical": same, "counts_identical": cnt,
                      "ids_identical_pct": 100.0 * same / n, "tokens_identical_pct": 100.0 * tok_same / int(g["offsets"][-1])}))


if __name__ == "__main__":
    main()
