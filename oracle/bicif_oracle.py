"""CPU restatement of BiCifParaformer's additions to the Paraformer path -- TEST INFRASTRUCTURE ONLY.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU legs may import this module; the product path
(`funasr_b200/`) never does.  Pinned against the reference itself: `tests/golden/bicif.npz` holds the outputs of the
unmodified `BiCifParaformer` (`tests/golden/make_golden_bicif.py`), `tests/test_oracle_bicif.py` checks this file
against them.

What BiCifParaformer (funasr/models/bicif_paraformer/model.py:37) adds to Paraformer at inference:
  * CifPredictorV3 (funasr/models/bicif_paraformer/cif_predictor.py:96-330): the same alpha head as V2, the
    integrate-and-fire done as a sequential fp32 loop (`cif`, :27-69), and a second, 3x upsampled alpha head
    (ConvTranspose1d k=3 s=3 -> BLSTM -> Linear -> sigmoid) whose fires give token time stamps;
  * `ts_prediction_lfr6_standard` (funasr/utils/timestamp_tools.py:31-106): fires -> [begin ms, end ms] per token.
One unpadded utterance at a time (batch 1 = the reference's CPU semantics; in a padded batch the reference's backward
LSTM starts in the padding, so its batched time stamps depend on the batch composition).
"""
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

from . import paraformer_oracle as P

SD = Dict[str, torch.Tensor]

UPSAMPLE = 3


def cif_loop(hidden: torch.Tensor, alphas: torch.Tensor, threshold: float = 1.0) -> Tuple[torch.Tensor, torch.Tensor]:
    """`cif` (bicif_paraformer/cif_predictor.py:27-69) for one utterance: hidden [T, D], alphas [T].
    Returns (frames at the fires [n, D], fires [T]) -- fp32, strictly sequential like the reference loop."""
    t_len, d = hidden.shape
    integrate = torch.zeros((), dtype=torch.float32)
    frame = torch.zeros(d)
    fires, frames = [], []
    one = torch.ones((), dtype=torch.float32)
    for t in range(t_len):
        alpha = alphas[t]
        completion = one - integrate
        integrate = integrate + alpha
        fires.append(integrate)
        fire = bool(integrate >= threshold)
        if fire:
            integrate = integrate - one
        cur = completion if fire else alpha
        remainds = alpha - cur
        frame = frame + cur * hidden[t]
        frames.append(frame)
        if fire:
            frame = remainds * hidden[t]
    fires = torch.stack(fires)
    frames = torch.stack(frames)
    return frames[fires >= threshold], fires


def cif_predictor_v3(enc: torch.Tensor, sd: SD, prefix: str = "predictor.", tail_threshold: float = 0.45):
    """CifPredictorV3.forward at inference (cif_predictor.py:171-246) + tail_process_fn (:295-316).
    Returns (acoustic_embeds [L, D], token_num L, alphas [T+1], fires [T+1])."""
    alphas = P.cif_alphas(enc, sd, prefix)
    t, d = enc.shape
    tail = torch.zeros(1)
    tail[0] = tail_threshold
    alphas_tail = torch.cat([alphas, tail])
    hidden_tail = torch.cat([enc, torch.zeros(1, d)], dim=0)
    token_num = int(torch.floor(alphas_tail.sum()).item())
    frames, fires = cif_loop(hidden_tail, alphas_tail)
    max_label_len = int(torch.round(alphas_tail.sum()).int().item())
    embeds = torch.zeros(max_label_len, d)
    n = min(frames.shape[0], max_label_len)
    embeds[:n] = frames[:n]
    return embeds[:token_num], token_num, alphas_tail, fires


def lstm_direction(x: torch.Tensor, w_ih, w_hh, b_ih, b_hh, reverse: bool) -> torch.Tensor:
    """One direction of torch.nn.LSTM (gate order i, f, g, o) over x [S, D] from zero state -> h [S, H]."""
    s_len, hid = x.shape[0], w_hh.shape[1]
    xg = F.linear(x, w_ih, b_ih)
    h = torch.zeros(hid)
    c = torch.zeros(hid)
    out = torch.zeros(s_len, hid)
    steps = range(s_len - 1, -1, -1) if reverse else range(s_len)
    for s in steps:
        g = xg[s] + F.linear(h, w_hh, b_hh)
        i_, f_, g_, o_ = g.split(hid)
        c = torch.sigmoid(f_) * c + torch.sigmoid(i_) * torch.tanh(g_)
        h = torch.sigmoid(o_) * torch.tanh(c)
        out[s] = h
    return out


def upsample_hidden(enc: torch.Tensor, sd: SD, prefix: str = "predictor.") -> torch.Tensor:
    """`upsample_cnn` then `blstm` (cif_predictor.py:262-265, use_cif1_cnn = false): enc [T, D] -> [3T, 2D]."""
    w = sd[prefix + "upsample_cnn.weight"]                      # ConvTranspose1d weight [in, out, k], stride = k
    k = w.shape[-1]
    # out[3t + j, o] = sum_c enc[t, c] * w[c, o, j] + b[o]
    up = torch.einsum("tc,coj->tjo", enc, w).reshape(enc.shape[0] * k, w.shape[1]) + sd[prefix + "upsample_cnn.bias"]
    b = prefix + "blstm."
    fwd = lstm_direction(up, sd[b + "weight_ih_l0"], sd[b + "weight_hh_l0"], sd[b + "bias_ih_l0"], sd[b + "bias_hh_l0"], False)
    bwd = lstm_direction(up, sd[b + "weight_ih_l0_reverse"], sd[b + "weight_hh_l0_reverse"], sd[b + "bias_ih_l0_reverse"],
                         sd[b + "bias_hh_l0_reverse"], True)
    return torch.cat([fwd, bwd], dim=-1)


def fires_wo_hidden(alphas: torch.Tensor, threshold: float) -> torch.Tensor:
    """`cif_wo_hidden` (cif_predictor.py:71-93; timestamp_tools.py:11-28) for one utterance: the integrate value
    after each frame, fp32, sequential; a fire subtracts `threshold`."""
    thr = torch.tensor(threshold, dtype=torch.float32)
    integrate = torch.zeros((), dtype=torch.float32)
    out = []
    for t in range(alphas.shape[0]):
        integrate = integrate + alphas[t]
        out.append(integrate)
        if bool(integrate >= thr):
            integrate = integrate - thr
    return torch.stack(out) if out else torch.zeros(0)


def upsample_timestamp(enc: torch.Tensor, sd: SD, token_num: int, prefix: str = "predictor.", smooth2: float = 0.25,
                       noise2: float = 0.01, threshold: float = 1.0) -> Tuple[torch.Tensor, torch.Tensor]:
    """CifPredictorV3.get_upsample_timestamp (cif_predictor.py:248-293) -> (us_alphas [3T], us_peaks [3T])."""
    h = upsample_hidden(enc, sd, prefix)
    a = torch.sigmoid(F.linear(h, sd[prefix + "cif_output2.weight"], sd[prefix + "cif_output2.bias"])).squeeze(-1)
    a = torch.relu(a * smooth2 - noise2)
    a = a * (torch.tensor(float(token_num)) / a.sum())
    return a, fires_wo_hidden(a, threshold - 1e-4)


def ts_prediction(us_alphas: torch.Tensor, us_peaks: torch.Tensor, num_tokens: int, vad_offset: float = 0.0,
                  force_time_shift: float = -1.5, upsample_rate: int = UPSAMPLE) -> List[List[int]]:
    """`ts_prediction_lfr6_standard` (funasr/utils/timestamp_tools.py:31-106), the `res` list: [begin ms, end ms]
    of each of the `num_tokens` tokens ("</s>" already removed).  The character strings only label the entries, so
    they are not needed here."""
    if num_tokens == 0:
        return []
    start_end_threshold, max_token_duration = 5, 12
    time_rate = 10.0 * 6 / 1000 / upsample_rate
    alphas, peaks = us_alphas.clone(), us_peaks
    thr = 1.0 - 1e-4
    fire_place = torch.where(peaks >= thr)[0].numpy() + force_time_shift
    if len(fire_place) != num_tokens + 1:
        alphas = alphas / (alphas.sum() / (num_tokens + 1))
        peaks = fires_wo_hidden(alphas, thr)
        fire_place = torch.where(peaks >= thr)[0].numpy() + force_time_shift
    num_frames = peaks.shape[0]
    stamps, is_sil = [], []
    if fire_place[0] > start_end_threshold:
        stamps.append([0.0, fire_place[0] * time_rate])
        is_sil.append(True)
    for i in range(len(fire_place) - 1):
        is_sil.append(False)
        if fire_place[i + 1] - fire_place[i] <= max_token_duration:
            stamps.append([fire_place[i] * time_rate, fire_place[i + 1] * time_rate])
        else:
            split = fire_place[i] + max_token_duration
            stamps.append([fire_place[i] * time_rate, split * time_rate])
            stamps.append([split * time_rate, fire_place[i + 1] * time_rate])
            is_sil.append(True)
    if num_frames - fire_place[-1] > start_end_threshold:
        end = (num_frames + fire_place[-1]) * 0.5
        stamps[-1][1] = end * time_rate
        stamps.append([end * time_rate, num_frames * time_rate])
        is_sil.append(True)
    elif stamps:
        stamps[-1][1] = num_frames * time_rate
    if vad_offset:
        for s in stamps:
            s[0] += vad_offset / 1000.0
            s[1] += vad_offset / 1000.0
    return [[int(s[0] * 1000), int(s[1] * 1000)] for s, sil in zip(stamps, is_sil) if not sil]


def recognize(wave: torch.Tensor, sd: SD, cmvn=None, window: str = "hamming", heads: int = 4,
              taps: Optional[dict] = None) -> Tuple[List[int], List[List[int]]]:
    """BiCifParaformer.inference for one utterance (bicif_paraformer/model.py:222-369): (token ids, time stamps)."""
    with torch.no_grad():
        feats = P.wav_frontend(wave, cmvn, window)
        enc = P.sanm_encoder(feats, sd, heads=heads)
        embeds, token_num, alphas, fires = cif_predictor_v3(enc, sd)
        if taps is not None:
            taps.update(feats=feats, enc=enc, embeds=embeds, token_num=token_num, alphas=alphas, fires=fires)
        if token_num < 1:
            return [], []
        logits = P.paraformer_decoder(enc, embeds, sd, heads=heads)
        ids, _ = P.greedy_pick(logits)
        us_alphas, us_peaks = upsample_timestamp(enc, sd, token_num)
        if taps is not None:
            taps.update(logits=logits, us_alphas=us_alphas, us_peaks=us_peaks)
        # the token list the reference passes is the filtered one; a "</s>" string at its end cannot occur (id 2 is
        # filtered), so every kept id is one character
        return ids, ts_prediction(us_alphas, us_peaks, len(ids))
