"""Pin oracle/conformer_oracle.py to the UNMODIFIED reference's outputs (tests/golden/conformer.npz, made by
tests/golden/make_golden_conformer.py) -- BASELINE config 4.  CPU only."""
import os

import numpy as np
import pytest
import torch

from funasr_b200 import params, synth
from oracle import conformer_oracle as C
from oracle import paraformer_oracle as P


def _rel(a, b):
    a = torch.as_tensor(a, dtype=torch.float64)
    b = torch.as_tensor(b, dtype=torch.float64)
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def test_param_inventory_matches_reference_count():
    # 34 552 970 parameters (the reference prints 34.553 M) + BatchNorm running statistics (2 x 256 x 12 buffers)
    assert params.count_params(params.conformer_ctc_shapes()) == 34552970 + 2 * 256 * 12


def test_rel_shift_closed_form():
    """The closed form the CUDA kernel indexes by (docstring of rel_shift_legacy) against the
    pad / view / slice transcription."""
    for t in (1, 2, 3, 7, 20):
        x = torch.arange(2 * t * t, dtype=torch.float32).reshape(2, t, t) + 1
        y = C.rel_shift_legacy(x)
        for i in range(t):
            for j in range(t):
                want = x[:, i, t - 1 - i + j] if j <= i else (torch.zeros(2) if j == i + 1 else x[:, i + 1, j - i - 2])
                assert torch.equal(y[:, i, j], want)


@pytest.mark.parametrize("ci", [0, 1])
def test_stages_match_reference(golden_dir, conformer_sd, ci):
    g = np.load(os.path.join(golden_dir, "conformer.npz"))
    ui, n = (int(v) for v in g["case_%d" % ci])
    wave = torch.from_numpy(synth.synth_waveform(ui, n))
    taps = []
    with torch.no_grad():
        feats = P.wav_frontend(wave, None, window="hamming", lfr_m=1, lfr_n=1)
        enc = C.conformer_encoder(feats, conformer_sd, taps=taps)
        logits = C.ctc_logits(enc, conformer_sd)
    assert enc.shape == g["enc_%d" % ci].shape
    assert _rel(taps[0], g["embed_%d" % ci]) < 2e-6
    assert _rel(taps[1], g["block0_%d" % ci]) < 1e-5
    assert _rel(taps[12], g["block11_%d" % ci]) < 2e-5
    assert _rel(enc, g["enc_%d" % ci]) < 2e-5
    assert np.array_equal(logits.argmax(-1).numpy().astype(np.int32), g["argmax_%d" % ci])
    assert P.ctc_greedy(logits, blank=0, drop=(0, 1, 2)) == g["ids_%d" % ci].tolist()
    assert C.recognize(wave, conformer_sd) == g["ids_%d" % ci].tolist()


def test_too_short_utterance_raises(conformer_sd):
    with pytest.raises(ValueError):
        C.conformer_encoder(torch.zeros(6, 80), conformer_sd)
