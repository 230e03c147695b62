"""Host time-stamp arithmetic (funasr_b200/timestamp.py) against the oracle's restatement of
ts_prediction_lfr6_standard and the reference's golden stamps.  CPU only."""
import os

import numpy as np
import pytest
import torch

from funasr_b200 import timestamp
from oracle import bicif_oracle as B

THR = 1.0 - 1e-4


def _fires(alphas):
    peaks = B.fires_wo_hidden(alphas, THR)
    return peaks, torch.where(peaks >= THR)[0].numpy()


@pytest.mark.parametrize("seed", range(8))
def test_matches_oracle_on_random_alphas(seed):
    g = torch.Generator().manual_seed(seed)
    n = int(torch.randint(30, 400, (1,), generator=g))
    alphas = torch.rand(n, generator=g) * 0.3
    if seed % 2:  # long silences: the 12-frame cut and the begin / end silence rules
        alphas[: n // 5] = 0
        alphas[n // 2: n // 2 + 40] = 0
        alphas[-(10 + seed):] = 0
    peaks, idx = _fires(alphas)
    chars = len(idx) - 1
    for off in (0.0, 1234.0):
        assert timestamp.stamps_from_fires(idx, n, chars, vad_offset=off) == B.ts_prediction(alphas, peaks, chars, vad_offset=off)


def test_reference_golden_stamps(golden_dir):
    g = np.load(os.path.join(golden_dir, "bicif.npz"))
    for ci in range(3):
        alphas, peaks = torch.from_numpy(g["us_alphas_%d" % ci]), torch.from_numpy(g["us_peaks_%d" % ci])
        chars = len(g["ids_%d" % ci])
        idx = torch.where(peaks >= THR)[0].numpy()
        if len(idx) != chars + 1:  # the reference's second pass (timestamp_tools.py:46-53)
            alphas = alphas / (alphas.sum() / (chars + 1))
            _, idx = _fires(alphas)
        assert timestamp.stamps_from_fires(idx, len(peaks), chars) == g["timestamp_%d" % ci].tolist()


def test_edges():
    assert timestamp.stamps_from_fires(np.array([3, 9]), 20, 0) == []
    assert timestamp.stamps_from_fires(np.array([], dtype=np.int64), 20, 2) == []
    assert timestamp.stamps_from_fires(np.array([7]), 20, 1) == []
    # more fire intervals than characters: cut to the characters
    assert len(timestamp.stamps_from_fires(np.array([2, 8, 14, 19]), 21, 2)) == 2
