"""Generate `tests/golden/*.npz` by running the UNMODIFIED reference in the build container.

    python tests/golden/make_golden.py            # needs /root/reference; ~2 min on 8 threads

The reference cannot travel to the GPU box, so the fixtures are committed.  Inputs are NOT
stored: waveforms, CMVN and weights are regenerated from integer seeds by
`funasr_b200/synth.py` (numpy PCG64, platform independent).  Outputs are the reference's
own tensors at every stage of the path, obtained through the
reference's public objects: `AutoModel(**template.yaml)`, `frontend(...)`, `model.encode`,
`model.calc_predictor`, `model.cal_decoder_with_predictor`, `model.inference`.
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

WEIGHT_SEED = 0


def build_reference_model(cmvn_path, window="hamming"):
    import_reference()
    from funasr import AutoModel

    cfg = yaml.safe_load(open(os.path.join(REFERENCE_ROOT, "funasr/models/paraformer/template.yaml")))
    cfg["tokenizer_conf"]["token_list"] = synth.synth_token_list(8404)
    cfg["frontend_conf"]["dither"] = 0.0
    cfg["frontend_conf"]["window"] = window
    cfg["frontend_conf"]["cmvn_file"] = cmvn_path
    cfg["specaug"] = None
    m = AutoModel(**cfg, device="cpu", ncpu=8, disable_update=True, disable_pbar=True)
    synth.fill_module_(m.model, WEIGHT_SEED)
    m.model.eval()
    return m


def stage_taps(m, wave):
    """Run the reference stage by stage on ONE utterance (batch 1 = CPU semantics)."""
    from funasr.utils.load_utils import extract_fbank

    model, frontend = m.model, m.kwargs["frontend"]
    with torch.no_grad():
        speech, lens = extract_fbank([torch.from_numpy(wave)], data_type="sound", frontend=frontend)
        taps = []
        hooks = []
        layers = list(model.encoder.encoders0) + list(model.encoder.encoders)
        for idx in (0, 1, len(layers) - 1):
            hooks.append(layers[idx].register_forward_hook(lambda mod, i, o, idx=idx: taps.append((idx, o[0][0].clone()))))
        enc, enc_lens = model.encode(speech, lens)
        for h in hooks:
            h.remove()
        embeds, tok_len, alphas, fires = model.calc_predictor(enc, enc_lens)
        tok_len = tok_len.round().long()
        out = dict(feats=speech[0].numpy(), enc=enc[0].numpy(), alphas=alphas[0].numpy(), fires=fires[0].numpy(),
                   token_num=np.int64(tok_len[0].item()), embeds=embeds[0].numpy())
        for idx, t in taps:
            out["enc_layer%d" % idx] = t.numpy()
        if tok_len.max() >= 1:
            dec, _ = model.cal_decoder_with_predictor(enc, enc_lens, embeds, tok_len)
            lp = dec[0, : tok_len[0]]
            top = lp.topk(2, dim=-1)
            out.update(logp_top2=top.values.numpy(), ids_top2=top.indices.numpy().astype(np.int32))
        res = model.inference(data_in=[wave], key=["utt"], tokenizer=None, frontend=frontend, device="cpu")
        ids = res[0][0]["token_int"] if len(res) == 2 and len(res[0]) else []
        out["token_int"] = np.asarray(ids, dtype=np.int32)
    return out


def main():
    torch.manual_seed(0)
    shift, scale = synth.synth_cmvn(560, 3)
    tmp = tempfile.mkdtemp()
    cmvn_path = os.path.join(tmp, "am.mvn")
    synth.write_cmvn_file(cmvn_path, shift, scale)

    # ---- 1. frontend only: windows x lengths, with and without CMVN ----------------------
    import_reference()
    from funasr.frontends.wav_frontend import WavFrontend

    fe = {}
    for window in ("hamming", "povey"):
        for cm in (None, cmvn_path):
            f = WavFrontend(cmvn_file=cm, window=window, lfr_m=7, lfr_n=6, dither=0.0)
            for ui, n in ((0, 8000), (1, 32491), (2, 80000)):
                w = torch.from_numpy(synth.synth_waveform(ui, n)).unsqueeze(0)
                feats, lens = f(w, torch.tensor([n]))
                fe["%s_%s_%d" % (window, "cmvn" if cm else "raw", n)] = feats[0].numpy()
    # plain fbank (no LFR) for one length, to pin the fbank stage by itself
    f = WavFrontend(cmvn_file=None, window="hamming", lfr_m=1, lfr_n=1, dither=0.0)
    fe["hamming_fbank_32491"] = f(torch.from_numpy(synth.synth_waveform(1, 32491)).unsqueeze(0), torch.tensor([32491]))[0][0].numpy()
    np.savez_compressed(os.path.join(HERE, "frontend.npz"), **fe)
    print("frontend.npz", {k: v.shape for k, v in fe.items()})

    # ---- 2. full Paraformer-large, stage taps ----------------------------------------------
    m = build_reference_model(cmvn_path)
    nparam = sum(p.numel() for p in m.model.parameters())
    print("reference params: %.3f M" % (nparam / 1e6))
    for ui, n, name in ((10, 32000, "stages_2s"), (11, 160000, "stages_10s")):
        out = stage_taps(m, synth.synth_waveform(ui, n))
        if name == "stages_10s":  # keep the fixture small: drop the big per-layer taps
            out = {k: v for k, v in out.items() if not k.startswith("enc_layer")}
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, {k: getattr(v, "shape", v) for k, v in out.items()})

    # ---- 3. transcripts for a ragged set (batch-1 reference semantics) ------------------------
    lens = [16000 * 2, 16000 * 3 + 77, 47999, 64000, 16000 * 5 + 1234, 100000, 123457, 16000 * 9]
    tr = {"num_samples": np.asarray(lens, dtype=np.int64)}
    for i, n in enumerate(lens):
        wave = synth.synth_waveform(100 + i, n)
        res = m.model.inference(data_in=[wave], key=["u%d" % i], tokenizer=None, frontend=m.kwargs["frontend"], device="cpu")
        ids = res[0][0]["token_int"] if len(res) == 2 and len(res[0]) else []
        tr["ids_%d" % i] = np.asarray(ids, dtype=np.int32)
    # text through the public API, one utterance
    txt = m.generate(input=synth.synth_waveform(100, lens[0]))[0]["text"]
    tr["text_0"] = np.asarray(txt)
    np.savez_compressed(os.path.join(HERE, "transcripts.npz"), **tr)
    print("transcripts", {k: (v.shape if v.ndim else v) for k, v in tr.items()})


if __name__ == "__main__":
    main()
