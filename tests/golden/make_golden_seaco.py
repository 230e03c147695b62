"""Generate `tests/golden/seaco.npz`: the UNMODIFIED reference SeacoParaformer (what the hub alias `paraformer-zh` resolves to,
funasr/download/name_maps_from_hub.py:3) run WITHOUT hot words on the BiCif golden utterances.

    python tests/golden/make_golden_seaco.py      # needs /root/reference; ~1 min on 8 threads

Without a `hotword` argument `_seaco_decode_with_ASF` (funasr/models/seaco_paraformer/model.py:233-327) returns the plain
decoder's log-softmax and the rest of `inference` is BiCifParaformer's: the fixture pins that claim (token ids and time stamps
for seeded weights; the tensors shared with BiCifParaformer get the same seeded values because `synth` keys values by
parameter name).  The hot-word branch (bias encoder, SeACo decoder, ASF) is not on the accelerated path.
"""
import os
import sys
import tempfile

import numpy as np
import torch
import yaml

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

from ref_loader import REFERENCE_ROOT, import_reference  # noqa: E402

from funasr_b200 import synth  # noqa: E402

CASES = ((200, 32000), (201, 16000 * 5 + 321), (202, 16000 * 11))


def main():
    torch.manual_seed(0)
    shift, scale = synth.synth_cmvn(560, 3)
    tmp = tempfile.mkdtemp()
    cmvn_path = os.path.join(tmp, "am.mvn")
    synth.write_cmvn_file(cmvn_path, shift, scale)
    import_reference()
    from funasr import AutoModel

    cfg = yaml.safe_load(open(os.path.join(REFERENCE_ROOT, "funasr/models/seaco_paraformer/template.yaml")))
    cfg["tokenizer_conf"]["token_list"] = synth.synth_token_list(8404)
    cfg["frontend_conf"]["dither"] = 0.0
    cfg["frontend_conf"]["cmvn_file"] = cmvn_path
    cfg["specaug"] = None
    m = AutoModel(**cfg, device="cpu", ncpu=8, disable_update=True, disable_pbar=True)
    synth.fill_module_(m.model, 0)
    m.model.eval()
    model, frontend, tokenizer = m.model, m.kwargs["frontend"], m.kwargs["tokenizer"]
    import funasr.models.seaco_paraformer.model as seaco_module

    seen = []
    original = seaco_module.ts_prediction_lfr6_standard

    def recording(*a, **k):
        r = original(*a, **k)
        seen.append(r)
        return r

    seaco_module.ts_prediction_lfr6_standard = recording
    extra = sorted(k for k in model.state_dict() if k.startswith(("bias_encoder.", "seaco_decoder.", "hotword_output_layer.")))
    out = {"num_samples": np.asarray([n for _, n in CASES], dtype=np.int64), "seeds": np.asarray([s for s, _ in CASES], dtype=np.int64),
           "extra_keys": np.asarray(extra)}
    with torch.no_grad():
        for ci, (seed, n) in enumerate(CASES):
            wave = synth.synth_waveform(seed, n)
            del seen[:]
            model.inference(data_in=[wave], key=["utt"], tokenizer=tokenizer, frontend=frontend, device="cpu")
            stamps = seen[0][1]
            ids = model.inference(data_in=[wave], key=["utt"], tokenizer=None, frontend=frontend, device="cpu")[0][0]["token_int"]
            out["ids_%d" % ci] = np.asarray(ids, dtype=np.int32)
            out["timestamp_%d" % ci] = np.asarray(stamps, dtype=np.int64).reshape(-1, 2)
            print(ci, n, "ids", len(ids), "stamps", len(stamps), stamps[:2])
    bicif = np.load(os.path.join(HERE, "bicif.npz"))
    for ci in range(len(CASES)):
        assert np.array_equal(out["ids_%d" % ci], bicif["ids_%d" % ci]), "SeACo without hot words differs from BiCif (ids)"
        assert np.array_equal(out["timestamp_%d" % ci], bicif["timestamp_%d" % ci]), "SeACo without hot words differs from BiCif (stamps)"
    np.savez_compressed(os.path.join(HERE, "seaco.npz"), **out)
    print("seaco.npz: identical to bicif.npz on", len(CASES), "utterances;", len(extra), "hot-word tensors unused")


if __name__ == "__main__":
    main()
