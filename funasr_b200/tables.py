"""Constant tables of the frontend and encoder prologue, computed once on the host at load.

These are parameters of the kernels (window, mel filterbank, positional inverse timescales,
CMVN vectors), not part of the per-utterance path.  Formulas:
  * window / mel bank: Kaldi's fbank as published (HTK mel scale 1127 ln(1 + f/700), triangular
    filters in the mel domain between 20 Hz and Nyquist, window 0.54 - 0.46 cos(2 pi n/(N-1)) for
    Hamming, hann^0.85 for Povey) -- the algorithm torchaudio.compliance.kaldi implements and the
    reference calls at funasr/frontends/wav_frontend.py:133-143;
  * positional timescales: funasr/models/transformer/embedding.py:396-406;
  * CMVN file: Kaldi-nnet text format parsed as in funasr/frontends/wav_frontend.py:15-38.
"""
import math
from typing import Optional, Tuple

import numpy as np
import torch


def feature_window(kind: str, size: int) -> torch.Tensor:
    n = torch.arange(size, dtype=torch.float32)
    if kind == "hamming":
        return torch.hamming_window(size, periodic=False, alpha=0.54, beta=0.46, dtype=torch.float32)
    if kind == "hanning":
        return torch.hann_window(size, periodic=False, dtype=torch.float32)
    if kind == "povey":
        return torch.hann_window(size, periodic=False, dtype=torch.float32).pow(0.85)
    if kind == "rectangular":
        return torch.ones_like(n)
    raise ValueError("funasr_b200: unsupported window type %r" % kind)


def _mel(f: float) -> float:
    return 1127.0 * math.log(1.0 + f / 700.0)


def mel_filterbank(n_mels: int, nfft: int, fs: float, low_hz: float = 20.0, high_hz: float = 0.0) -> torch.Tensor:
    """[n_mels, nfft/2 + 1] fp32; the Nyquist column is zero (Kaldi drops that bin)."""
    nbins = nfft // 2
    if high_hz <= 0.0:
        high_hz += 0.5 * fs
    lo, hi = _mel(low_hz), _mel(high_hz)
    delta = (hi - lo) / (n_mels + 1)
    b = torch.arange(n_mels, dtype=torch.float32).unsqueeze(1)
    left, center, right = lo + b * delta, lo + (b + 1.0) * delta, lo + (b + 2.0) * delta
    freqs = (fs / nfft) * torch.arange(nbins, dtype=torch.float32)
    mel = (1127.0 * (1.0 + freqs / 700.0).log()).unsqueeze(0)
    up = (mel - left) / (center - left)
    down = (right - mel) / (right - center)
    bank = torch.clamp(torch.minimum(up, down), min=0.0)
    return torch.cat([bank, torch.zeros(n_mels, 1)], dim=1).contiguous()


def pe_inverse_timescales(depth: int) -> torch.Tensor:
    """[depth/2] fp32: exp(-i ln(1e4) / (depth/2 - 1))."""
    inc = torch.log(torch.tensor([10000.0], dtype=torch.float32)) / (depth / 2 - 1)
    return torch.exp(torch.arange(depth // 2, dtype=torch.float32) * (-inc)).contiguous()


def load_cmvn(path: str) -> Tuple[torch.Tensor, torch.Tensor]:
    """Kaldi-nnet `am.mvn`: the `<LearnRateCoef>` line after `<AddShift>` is the shift, the one
    after `<Rescale>` the scale; values are fields [3:-1] of the line."""
    with open(path, "r", encoding="utf-8") as f:
        lines = f.readlines()
    shift = scale = None
    for i, line in enumerate(lines):
        item = line.split()
        if not item:
            continue
        if item[0] in ("<AddShift>", "<Rescale>"):
            nxt = lines[i + 1].split()
            if nxt and nxt[0] == "<LearnRateCoef>":
                vals = np.asarray(nxt[3:-1], dtype=np.float64)
                if item[0] == "<AddShift>":
                    shift = vals
                else:
                    scale = vals
    if shift is None or scale is None:
        raise ValueError("funasr_b200: %s holds no <AddShift>/<Rescale> statistics" % path)
    return torch.from_numpy(shift.astype(np.float32)), torch.from_numpy(scale.astype(np.float32))


def num_frames(num_samples: int, frame_len: int = 400, frame_shift: int = 160) -> int:
    """snip_edges=True frame count."""
    return 0 if num_samples < frame_len else 1 + (num_samples - frame_len) // frame_shift


def num_lfr_rows(frames: int, lfr_n: int = 6) -> int:
    return (frames + lfr_n - 1) // lfr_n


def cu_from_lengths(lengths, dtype=torch.int32, device: Optional[torch.device] = None) -> torch.Tensor:
    cu = np.zeros(len(lengths) + 1, dtype=np.int64)
    np.cumsum(np.asarray(lengths, dtype=np.int64), out=cu[1:])
    return torch.from_numpy(cu).to(dtype).to(device) if device is not None else torch.from_numpy(cu).to(dtype)


def legacy_rel_position_table(rows: int, d: int, max_len: int = 5000) -> torch.Tensor:
    """[rows, d] fp32: the first `rows` rows of LegacyRelPositionalEncoding's table
    (funasr/models/transformer/embedding.py:59-78 with reverse=True, :215-252): row i encodes position
    max_len - 1 - i, sin in even columns, cos in odd ones.  The reference builds the table once for
    max_len = 5000 and hands its first T rows to every layer, so row i does not depend on T (T <= max_len)."""
    if rows > max_len:
        raise ValueError("funasr_b200: %d rows exceed the reference's %d-row position table" % (rows, max_len))
    position = torch.arange(max_len - 1, -1, -1.0, dtype=torch.float32).unsqueeze(1)[:rows]
    div_term = torch.exp(torch.arange(0, d, 2, dtype=torch.float32) * -(math.log(10000.0) / d))
    pe = torch.zeros(rows, d)
    pe[:, 0::2] = torch.sin(position * div_term)
    pe[:, 1::2] = torch.cos(position * div_term)
    return pe
