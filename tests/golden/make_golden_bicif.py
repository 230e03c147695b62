"""Generate `tests/golden/bicif.npz` by running the UNMODIFIED reference BiCifParaformer in the build container.

    python tests/golden/make_golden_bicif.py      # needs /root/reference; ~1 min on 8 threads

Inputs (waveforms, CMVN, weights) are regenerated from integer seeds by `funasr_b200/synth.py`; the fixture holds the
reference's own outputs: `model.calc_predictor` (CifPredictorV3), `model.calc_predictor_timestamp`, and the token ids
/ time stamps of `model.inference` through the reference's CharTokenizer.  The time stamps are the value
`ts_prediction_lfr6_standard` returns INSIDE `model.inference` (observed through a recording wrapper): the synthetic
vocabulary ("t17", ...) is neither Chinese nor alphabetic, so `sentence_postprocess` -- text post-processing outside
the path -- drops every stamp from the final result dict.
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

from funasr_b200 import params, synth  # noqa: E402

WEIGHT_SEED = 0
CASES = ((200, 32000), (201, 16000 * 5 + 321), (202, 16000 * 11))


def main():
    torch.manual_seed(0)
    shift, scale = synth.synth_cmvn(560, 3)
    tmp = tempfile.mkdtemp()
    cmvn_path = os.path.join(tmp, "am.mvn")
    synth.write_cmvn_file(cmvn_path, shift, scale)
    import_reference()
    from funasr import AutoModel
    from funasr.utils.load_utils import extract_fbank

    cfg = yaml.safe_load(open(os.path.join(REFERENCE_ROOT, "funasr/models/bicif_paraformer/template.yaml")))
    cfg["tokenizer_conf"]["token_list"] = synth.synth_token_list(8404)
    cfg["frontend_conf"]["dither"] = 0.0
    cfg["frontend_conf"]["cmvn_file"] = cmvn_path
    cfg["specaug"] = None
    m = AutoModel(**cfg, device="cpu", ncpu=8, disable_update=True, disable_pbar=True)
    synth.fill_module_(m.model, WEIGHT_SEED)
    m.model.eval()
    model, frontend, tokenizer = m.model, m.kwargs["frontend"], m.kwargs["tokenizer"]
    import funasr.models.bicif_paraformer.model as bicif_module

    seen = []
    original = bicif_module.ts_prediction_lfr6_standard

    def recording(*a, **k):
        r = original(*a, **k)
        seen.append(r)
        return r

    bicif_module.ts_prediction_lfr6_standard = recording
    want = params.bicif_paraformer_shapes()
    have = {k: tuple(v.shape) for k, v in model.state_dict().items()}
    assert have == {k: tuple(v) for k, v in want.items()}, sorted(set(have) ^ set(want))[:8]
    out = {"num_samples": np.asarray([n for _, n in CASES], dtype=np.int64),
           "seeds": np.asarray([s for s, _ in CASES], dtype=np.int64)}
    with torch.no_grad():
        for ci, (seed, n) in enumerate(CASES):
            wave = synth.synth_waveform(seed, n)
            speech, lens = extract_fbank([torch.from_numpy(wave)], data_type="sound", frontend=frontend)
            enc, enc_lens = model.encode(speech, lens)
            embeds, tok_len, alphas, fires = model.calc_predictor(enc, enc_lens)
            tok = tok_len.round().long()
            _, _, us_alphas, us_peaks = model.calc_predictor_timestamp(enc, enc_lens, tok)
            del seen[:]
            res = model.inference(data_in=[wave], key=["utt"], tokenizer=tokenizer, frontend=frontend, device="cpu")
            r = res[0][0]
            stamps = seen[0][1]
            ids = model.inference(data_in=[wave], key=["utt"], tokenizer=None, frontend=frontend, device="cpu")[0][0]["token_int"]
            out.update({
                "enc_%d" % ci: enc[0].numpy(), "embeds_%d" % ci: embeds[0].numpy(), "alphas_%d" % ci: alphas[0].numpy(),
                "fires_%d" % ci: fires[0].numpy(), "token_num_%d" % ci: np.int64(tok[0].item()),
                "us_alphas_%d" % ci: us_alphas[0].numpy(), "us_peaks_%d" % ci: us_peaks[0].numpy(),
                "ids_%d" % ci: np.asarray(ids, dtype=np.int32),
                "timestamp_%d" % ci: np.asarray(stamps, dtype=np.int64).reshape(-1, 2),
                "text_%d" % ci: np.asarray(r["text"]),
            })
            print(ci, n, "tokens", int(tok[0]), "ids", len(ids), "stamps", len(stamps), stamps[:3], stamps[-2:])
    np.savez_compressed(os.path.join(HERE, "bicif.npz"), **out)
    print("bicif.npz", {k: getattr(v, "shape", v) for k, v in out.items()})


if __name__ == "__main__":
    main()
