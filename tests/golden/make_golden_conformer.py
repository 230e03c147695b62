"""Generate `tests/golden/conformer.npz` by running the UNMODIFIED reference in the build container --
BASELINE config 4: `model_classes["CTC"]` (funasr/models/ctc/model.py) over `ConformerEncoder`
(funasr/models/conformer/encoder.py) with the aishell recipe's encoder options
(examples/aishell/conformer/conf/conformer_12e_6d_2048_256.yaml:16-31), CTC greedy decoding.

    python tests/golden/make_golden_conformer.py      # needs /root/reference; < 1 min

Random-init weights from integer seeds (`funasr_b200/synth.py`), V = 4234, 12 blocks; the frontend, the model
object and `model.inference` are the reference's own.  Stored per utterance: the embed output (input of block
0), the outputs of blocks 0 and 11, the encoder output, the per-frame argmax and the token ids `inference` kept.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

from ref_loader import import_reference  # noqa: E402

from funasr_b200 import synth  # noqa: E402

WEIGHT_SEED = 0
VOCAB = 4234
ENCODER_CONF = dict(output_size=256, attention_heads=4, linear_units=2048, num_blocks=12, dropout_rate=0.1,
                    positional_dropout_rate=0.1, attention_dropout_rate=0.0, input_layer="conv2d",
                    normalize_before=True, pos_enc_layer_type="rel_pos", selfattention_layer_type="rel_selfattn",
                    activation_type="swish", macaron_style=True, use_cnn_module=True, cnn_module_kernel=15)
CASES = ((30, 40000), (31, 112000))  # (utterance index, samples): 2.5 s and 7 s


class RecordingTokenizer:
    """Keeps the ids `inference` hands to the tokenizer."""

    def __init__(self):
        self.seen = []

    def ids2tokens(self, ids):
        self.seen.append(list(ids))
        return [str(i) for i in ids]

    def tokens2text(self, tokens):
        return " ".join(tokens)


class FirstTwo(torch.nn.Module):
    def __init__(self, inner):
        super().__init__()
        self.inner = inner

    def forward(self, *args, **kwargs):
        return self.inner(*args, **kwargs)[:2]


def build_reference_model():
    import_reference()
    from funasr.register import tables

    model = tables.model_classes["CTC"](encoder="ConformerEncoder", encoder_conf=dict(ENCODER_CONF), input_size=80,
                                        vocab_size=VOCAB)
    synth.fill_module_(model, WEIGHT_SEED)
    model.eval()
    # The CTC model's `encode` unpacks two values (funasr/models/ctc/model.py:139) while ConformerEncoder.forward
    # returns three (encoder.py:607, as the Transformer/Conformer model's encode expects, transformer/model.py:254):
    # a two-value view of the unmodified encoder lets the unmodified `inference` (CTC greedy) run over it.
    model.encoder = FirstTwo(model.encoder)
    return model


def main():
    torch.manual_seed(0)
    model = build_reference_model()
    from funasr.frontends.wav_frontend import WavFrontend

    frontend = WavFrontend(cmvn_file=None, window="hamming", n_mels=80, lfr_m=1, lfr_n=1, dither=0.0)
    print("reference Conformer-CTC params: %.3f M" % (sum(p.numel() for p in model.parameters()) / 1e6))
    out = {}
    for ci, (ui, n) in enumerate(CASES):
        wave = synth.synth_waveform(ui, n)
        tok = RecordingTokenizer()
        got = {}
        hooks = [
            model.encoder.inner.register_forward_hook(lambda mod, i, o: got.update(enc=o[0][0].clone())),
            model.encoder.inner.embed.register_forward_hook(lambda mod, i, o: got.update(embed=o[0][0][0].clone())),
            model.encoder.inner.encoders[0].register_forward_hook(lambda mod, i, o: got.update(block0=o[0][0][0].clone())),
            model.encoder.inner.encoders[11].register_forward_hook(lambda mod, i, o: got.update(block11=o[0][0][0].clone())),
        ]
        with torch.no_grad():
            model.inference(data_in=[wave], key=["utt%d" % ci], tokenizer=tok, frontend=frontend, device="cpu")
        for h in hooks:
            h.remove()
        with torch.no_grad():
            logp = model.ctc.log_softmax(got["enc"].unsqueeze(0))[0]
        for k in ("embed", "block0", "block11", "enc"):
            out["%s_%d" % (k, ci)] = got[k].numpy()
        out["argmax_%d" % ci] = logp.argmax(-1).numpy().astype(np.int32)
        out["ids_%d" % ci] = np.asarray(tok.seen[0], dtype=np.int32)
        out["case_%d" % ci] = np.asarray([ui, n], dtype=np.int64)
        print(ci, n, out["enc_%d" % ci].shape, len(tok.seen[0]))
    np.savez_compressed(os.path.join(HERE, "conformer.npz"), **out)


if __name__ == "__main__":
    main()
