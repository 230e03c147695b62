"""Generate `tests/golden/sensevoice.npz` by running the UNMODIFIED reference SenseVoiceSmall
(funasr/models/sense_voice/model.py) in the build container -- BASELINE config 3.

    python tests/golden/make_golden_sensevoice.py      # needs /root/reference; ~1 min

Random-init weights from integer seeds (`funasr_b200/synth.py`), V = 25055, 50 + 20 blocks; the
frontend, the model object and `model.inference` are the reference's own.  Stored: encoder output,
per-frame argmax and the CTC-greedy token ids of two utterances (one per language/text-norm query).
"""
import os
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

from ref_loader import import_reference  # noqa: E402

from funasr_b200 import synth  # noqa: E402

WEIGHT_SEED = 0
VOCAB = 25055


class RecordingTokenizer:
    """`decode(list) -> str` is all SenseVoiceSmall.inference needs; it keeps what it was given."""

    def __init__(self):
        self.seen = []

    def decode(self, ids):
        self.seen.append(list(ids))
        return " ".join(str(i) for i in ids)


def main():
    torch.manual_seed(0)
    import_reference()
    from funasr.frontends.wav_frontend import WavFrontend
    from funasr.models.sense_voice.model import SenseVoiceSmall

    shift, scale = synth.synth_cmvn(560, 3)
    tmp = tempfile.mkdtemp()
    cmvn_path = os.path.join(tmp, "am.mvn")
    synth.write_cmvn_file(cmvn_path, shift, scale)
    frontend = WavFrontend(cmvn_file=cmvn_path, window="hamming", lfr_m=7, lfr_n=6, dither=0.0)
    enc_conf = dict(output_size=512, attention_heads=4, linear_units=2048, num_blocks=50, tp_blocks=20, dropout_rate=0.1,
                    positional_dropout_rate=0.1, attention_dropout_rate=0.1, input_layer="pe", normalize_before=True,
                    kernel_size=11, sanm_shfit=0, selfattention_layer_type="sanm")
    model = SenseVoiceSmall(specaug=None, normalize=None, encoder="SenseVoiceEncoderSmall", encoder_conf=enc_conf,
                            ctc_conf={}, input_size=560, vocab_size=VOCAB, blank_id=0)
    synth.fill_module_(model, WEIGHT_SEED)
    model.eval()
    print("reference SenseVoiceSmall params: %.3f M" % (sum(p.numel() for p in model.parameters()) / 1e6))
    out = {}
    cases = ((20, 48000, "auto", False), (21, 100000, "en", True))
    for ci, (ui, n, lang, itn) in enumerate(cases):
        wave = synth.synth_waveform(ui, n)
        tok = RecordingTokenizer()
        got = {}
        h = model.encoder.register_forward_hook(lambda mod, i, o: got.update(enc=o[0][0].clone()))
        with torch.no_grad():
            model.inference(data_in=[wave], key=["utt%d" % ci], tokenizer=tok, frontend=frontend, device="cpu",
                            language=lang, use_itn=itn)
        h.remove()
        with torch.no_grad():
            logp = model.ctc.log_softmax(got["enc"].unsqueeze(0))[0]
        out["enc_%d" % ci] = got["enc"].numpy()
        out["argmax_%d" % ci] = logp.argmax(-1).numpy().astype(np.int32)
        out["ids_%d" % ci] = np.asarray(tok.seen[0], dtype=np.int32)
        out["case_%d" % ci] = np.asarray([ui, n, {"auto": 0, "en": 4}[lang], int(itn)], dtype=np.int64)
        print(ci, n, lang, itn, out["enc_%d" % ci].shape, len(tok.seen[0]))
    np.savez_compressed(os.path.join(HERE, "sensevoice.npz"), **out)


if __name__ == "__main__":
    main()
