"""Pin oracle/bicif_oracle.py to the UNMODIFIED reference's BiCifParaformer outputs (tests/golden/bicif.npz, made by
tests/golden/make_golden_bicif.py) -- SURVEY 8(f) rank 3.  CPU only."""
import os

import numpy as np
import pytest
import torch

from funasr_b200 import params, synth
from oracle import bicif_oracle as B


def _rel(a, b):
    a = torch.as_tensor(a, dtype=torch.float64)
    b = torch.as_tensor(b, dtype=torch.float64)
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.fixture(scope="module")
def golden(golden_dir):
    return np.load(os.path.join(golden_dir, "bicif.npz"))


def test_param_inventory():
    # Paraformer-large 220 084 533 + upsample conv + 2 x LSTM direction + second alpha head
    extra = 512 * 512 * 3 + 512 + 2 * (2 * 2048 * 512 + 2 * 2048) + 1024 + 1
    assert params.count_params(params.bicif_paraformer_shapes()) == 220084533 + extra


@pytest.mark.parametrize("ci", [0, 1, 2])
def test_predictor_v3_matches_reference(golden, bicif_sd, ci):
    enc = torch.from_numpy(golden["enc_%d" % ci])
    with torch.no_grad():
        embeds, token_num, alphas, fires = B.cif_predictor_v3(enc, bicif_sd)
    assert token_num == int(golden["token_num_%d" % ci])
    assert _rel(alphas, golden["alphas_%d" % ci]) < 1e-6
    assert _rel(fires, golden["fires_%d" % ci]) < 1e-5
    assert embeds.shape == golden["embeds_%d" % ci].shape
    assert _rel(embeds, golden["embeds_%d" % ci]) < 1e-5


@pytest.mark.parametrize("ci", [0, 1, 2])
def test_upsampled_alphas_and_peaks_match_reference(golden, bicif_sd, ci):
    enc = torch.from_numpy(golden["enc_%d" % ci])
    with torch.no_grad():
        us_alphas, us_peaks = B.upsample_timestamp(enc, bicif_sd, int(golden["token_num_%d" % ci]))
    assert us_alphas.shape == golden["us_alphas_%d" % ci].shape
    assert _rel(us_alphas, golden["us_alphas_%d" % ci]) < 1e-5
    want = torch.from_numpy(golden["us_peaks_%d" % ci])
    assert _rel(us_peaks, want) < 1e-4
    thr = 1.0 - 1e-4
    assert torch.equal(torch.where(us_peaks >= thr)[0], torch.where(want >= thr)[0])


@pytest.mark.parametrize("ci", [0, 1, 2])
def test_timestamps_from_reference_peaks_are_exact(golden, ci):
    stamps = B.ts_prediction(torch.from_numpy(golden["us_alphas_%d" % ci]), torch.from_numpy(golden["us_peaks_%d" % ci]),
                             len(golden["ids_%d" % ci]))
    assert stamps == golden["timestamp_%d" % ci].tolist()


def test_recognize_matches_reference_inference(golden, bicif_sd):
    shift, scale = synth.synth_cmvn()
    cmvn = (torch.from_numpy(shift), torch.from_numpy(scale))
    wave = torch.from_numpy(synth.synth_waveform(int(golden["seeds"][0]), int(golden["num_samples"][0])))
    ids, stamps = B.recognize(wave, bicif_sd, cmvn)
    assert ids == golden["ids_0"].tolist()
    assert stamps == golden["timestamp_0"].tolist()


def test_timestamp_edge_cases():
    assert B.ts_prediction(torch.zeros(6), torch.zeros(6), 0) == []
    # a long gap between two fires is cut at 12 upsampled frames (the rest is silence and is not reported)
    a = torch.zeros(60)
    a[[2, 40, 50]] = 1.0
    peaks = B.fires_wo_hidden(a, 1.0 - 1e-4)
    assert B.ts_prediction(a, peaks, 2) == [[10, 250], [770, 1085]]
