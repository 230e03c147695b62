"""Pin the oracle (oracle/paraformer_oracle.py) to the UNMODIFIED reference's outputs
(tests/golden/*.npz, made by tests/golden/make_golden.py).  CPU only."""
import os

import numpy as np
import pytest
import torch

from funasr_b200 import params, synth
from oracle import paraformer_oracle as O


def _rel(a, b):
    a = torch.as_tensor(a, dtype=torch.float64)
    b = torch.as_tensor(b, dtype=torch.float64)
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def test_param_inventory_matches_reference_count():
    assert params.count_params(params.paraformer_large_shapes()) == 220084533  # 220.08 M


@pytest.mark.parametrize("window", ["hamming", "povey"])
@pytest.mark.parametrize("n,ui", [(8000, 0), (32491, 1), (80000, 2)])
def test_frontend_matches_reference(golden_dir, cmvn, window, n, ui):
    g = np.load(os.path.join(golden_dir, "frontend.npz"))
    wave = torch.from_numpy(synth.synth_waveform(ui, n))
    raw = O.wav_frontend(wave, None, window)
    assert raw.shape == g["%s_raw_%d" % (window, n)].shape
    assert _rel(raw, g["%s_raw_%d" % (window, n)]) < 2e-6
    cm = O.wav_frontend(wave, cmvn, window)
    assert _rel(cm, g["%s_cmvn_%d" % (window, n)]) < 2e-6


def test_fbank_stage_matches_reference(golden_dir):
    g = np.load(os.path.join(golden_dir, "frontend.npz"))
    fb = O.kaldi_fbank(torch.from_numpy(synth.synth_waveform(1, 32491)) * 32768.0)
    assert _rel(fb, g["hamming_fbank_32491"]) < 2e-6


def test_lfr_closed_form_matches_reference_algorithm():
    """apply_lfr closed form vs a literal transcription-free check: every output row is a
    window of 7 consecutive clamped frames starting at 6i-3."""
    for f in list(range(1, 40)) + [167, 998, 1000, 2998]:
        x = torch.arange(f * 2, dtype=torch.float32).reshape(f, 2)
        y = O.apply_lfr(x, 7, 6)
        assert y.shape == ((f + 5) // 6, 14)
        for i in (0, y.shape[0] // 2, y.shape[0] - 1):
            for j in range(7):
                src = min(max(6 * i - 3 + j, 0), f - 1)
                assert torch.equal(y[i, 2 * j : 2 * j + 2], x[src])


@pytest.mark.parametrize("name,ui,n", [("stages_2s", 10, 32000), ("stages_10s", 11, 160000)])
def test_stages_match_reference(golden_dir, paraformer_sd, cmvn, name, ui, n):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    sd = paraformer_sd
    wave = torch.from_numpy(synth.synth_waveform(ui, n))
    with torch.no_grad():
        feats = O.wav_frontend(wave, cmvn)
        assert _rel(feats, g["feats"]) < 2e-6
        taps = []
        enc = O.sanm_encoder(torch.from_numpy(g["feats"]), sd, taps=taps)
        for idx in (0, 1, 49):
            key = "enc_layer%d" % idx
            if key in g.files:
                assert _rel(taps[idx], g[key]) < 2e-5, key
        assert _rel(enc, g["enc"]) < 2e-5
        embeds, token_num, alphas, fires = O.cif_predictor(torch.from_numpy(g["enc"]), sd)
        assert token_num == int(g["token_num"])
        assert np.abs(alphas.numpy() - g["alphas"]).max() < 1e-6
        assert np.abs(fires.numpy() - g["fires"]).max() < 1e-5
        assert embeds.shape == g["embeds"].shape
        assert _rel(embeds, g["embeds"]) < 1e-5
        logits = O.paraformer_decoder(torch.from_numpy(g["enc"]), torch.from_numpy(g["embeds"]), sd)
        lp = torch.log_softmax(logits, dim=-1)
        top = lp.topk(2, dim=-1)
        assert np.abs(top.values.numpy() - g["logp_top2"]).max() < 1e-4
        assert np.array_equal(top.indices.numpy()[:, 0], g["ids_top2"][:, 0])
        ids = O.recognize(wave, sd, cmvn)
        assert ids == g["token_int"].tolist()


def test_transcripts_match_reference(golden_dir, paraformer_sd, cmvn):
    g = np.load(os.path.join(golden_dir, "transcripts.npz"))
    for i, n in enumerate(g["num_samples"].tolist()):
        ids = O.recognize(torch.from_numpy(synth.synth_waveform(100 + i, n)), paraformer_sd, cmvn)
        assert ids == g["ids_%d" % i].tolist(), "utterance %d" % i


def test_fbank_matches_torchaudio_when_present():
    ta = pytest.importorskip("torchaudio.compliance.kaldi")
    w = torch.from_numpy(synth.synth_waveform(5, 24000)) * 32768.0
    ref = ta.fbank(w.unsqueeze(0), num_mel_bins=80, frame_length=25, frame_shift=10, dither=0.0,
                   energy_floor=0.0, window_type="hamming", sample_frequency=16000, snip_edges=True)
    assert _rel(O.kaldi_fbank(w), ref) < 2e-6


def test_ctc_greedy_collapse():
    logits = torch.full((7, 5), -5.0)
    for t, k in enumerate([0, 3, 3, 0, 3, 2, 2]):
        logits[t, k] = 1.0
    assert O.ctc_greedy(logits) == [3, 3, 2]
    assert O.ctc_greedy(logits, drop=(0, 1, 2)) == [3, 3]


# ---------------------------------------------------------------- SenseVoiceSmall (config 3)
SV_LANG = {0: "auto", 4: "en"}


def test_sensevoice_param_inventory():
    assert params.count_params(params.sense_voice_small_shapes()) == 233999167  # the reference build reports 233.999 M


@pytest.mark.parametrize("ci", [0, 1])
def test_sensevoice_oracle_matches_reference(golden_dir, sensevoice_sd, cmvn, ci):
    """Encoder output, per-frame argmax and CTC-greedy ids of the reference SenseVoiceSmall.inference."""
    g = np.load(os.path.join(golden_dir, "sensevoice.npz"))
    ui, n, lang, itn = (int(v) for v in g["case_%d" % ci])
    wave = torch.from_numpy(synth.synth_waveform(ui, n))
    taps = {}
    ids = O.sense_voice_recognize(wave, sensevoice_sd, cmvn, language=SV_LANG[lang],
                                  textnorm="withitn" if itn else "woitn", taps=taps)
    assert taps["enc"].shape == g["enc_%d" % ci].shape
    assert _rel(taps["enc"], g["enc_%d" % ci]) < 2e-5
    assert taps["logits"].argmax(-1).tolist() == g["argmax_%d" % ci].tolist()
    assert ids == g["ids_%d" % ci].tolist()
