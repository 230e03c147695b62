"""ORACLE -- test infrastructure only, never the product path.

CPU (torch fp32) restatement of the reference's batched offline non-autoregressive
recognition path, written functionally over a flat state_dict so that it needs nothing
from `/root/reference` at run time (the reference tree does not exist on the GPU box).
Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
leg may import this module.  `funasr_b200/` must never import it.

Parity pin: the reference's own tests hold no golden vector for this path (SURVEY.md
section 4), so the oracle is pinned against outputs of the UNMODIFIED reference imported in
the build container: `tests/golden/make_golden.py` writes `tests/golden/*.npz`, and
`tests/test_oracle_golden.py` checks every stage of this file against them.

Semantics restated = reference CPU path = batch_size 1 (funasr/auto/auto_model.py:185-187):
one utterance at a time, no padding anywhere.

Each function cites the reference lines it follows (paths relative to /root/reference).
fbank follows torchaudio.compliance.kaldi (third party, un-pinned by the reference; the
installed 2.11 is what the goldens were made with).
"""
import math
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F

SD = Dict[str, torch.Tensor]

# --------------------------------------------------------------------------------------
# Frontend: kaldi fbank -> LFR -> CMVN
# --------------------------------------------------------------------------------------


def mel_scale(freq):
    return 1127.0 * (1.0 + freq / 700.0).log()


def mel_scale_scalar(freq: float) -> float:
    return 1127.0 * math.log(1.0 + freq / 700.0)


def get_mel_banks(num_bins: int, padded: int, fs: float, low: float = 20.0, high: float = 0.0) -> torch.Tensor:
    """torchaudio/compliance/kaldi.py `get_mel_banks` (vtln_warp 1.0): [num_bins, padded/2]."""
    num_fft_bins = padded / 2
    nyquist = 0.5 * fs
    if high <= 0.0:
        high += nyquist
    fft_bin_width = fs / padded
    mel_low = mel_scale_scalar(low)
    mel_high = mel_scale_scalar(high)
    delta = (mel_high - mel_low) / (num_bins + 1)
    b = torch.arange(num_bins).unsqueeze(1)
    left = mel_low + b * delta
    center = mel_low + (b + 1.0) * delta
    right = mel_low + (b + 2.0) * delta
    mel = mel_scale(fft_bin_width * torch.arange(num_fft_bins)).unsqueeze(0)
    up = (mel - left) / (center - left)
    down = (right - mel) / (right - center)
    return torch.max(torch.zeros(1), torch.min(up, down))


def window_function(kind: str, size: int) -> torch.Tensor:
    """torchaudio `_feature_window_function`."""
    if kind == "hamming":
        return torch.hamming_window(size, periodic=False, alpha=0.54, beta=0.46)
    if kind == "hanning":
        return torch.hann_window(size, periodic=False)
    if kind == "povey":
        return torch.hann_window(size, periodic=False).pow(0.85)
    if kind == "rectangular":
        return torch.ones(size)
    raise ValueError("window " + kind)


def kaldi_fbank(
    wave: torch.Tensor,
    fs: int = 16000,
    n_mels: int = 80,
    frame_length_ms: float = 25.0,
    frame_shift_ms: float = 10.0,
    window: str = "hamming",
    preemph: float = 0.97,
) -> torch.Tensor:
    """`kaldi.fbank(dither=0, energy_floor=0, snip_edges=True)` as called at
    funasr/frontends/wav_frontend.py:133-143.  wave: [N] already scaled by 32768."""
    shift = int(fs * frame_shift_ms * 0.001)
    size = int(fs * frame_length_ms * 0.001)
    padded = 1 << (size - 1).bit_length()
    n = wave.numel()
    if n < size:
        return torch.zeros(0, n_mels)
    m = 1 + (n - size) // shift
    frames = wave.as_strided((m, size), (shift, 1))
    frames = frames - frames.mean(dim=1, keepdim=True)  # remove_dc_offset
    prev = F.pad(frames.unsqueeze(0), (1, 0), mode="replicate").squeeze(0)[:, :-1]
    frames = frames - preemph * prev
    frames = frames * window_function(window, size).unsqueeze(0)
    frames = F.pad(frames, (0, padded - size))
    spec = torch.fft.rfft(frames).abs().pow(2.0)  # [m, padded/2+1]
    mel = F.pad(get_mel_banks(n_mels, padded, float(fs)), (0, 1))  # [n_mels, padded/2+1]
    e = torch.mm(spec, mel.T)
    eps = torch.tensor(torch.finfo(torch.float32).eps)
    return torch.max(e, eps).log()


def apply_lfr(feat: torch.Tensor, m: int = 7, n: int = 6) -> torch.Tensor:
    """funasr/frontends/wav_frontend.py:58-74 in closed form (verified for F=1..3000 in
    tests): out[i, 80j:80j+80] = feat[clamp(n*i - (m-1)//2 + j, 0, F-1)]."""
    f = feat.shape[0]
    t = (f + n - 1) // n
    idx = torch.arange(t).unsqueeze(1) * n - (m - 1) // 2 + torch.arange(m).unsqueeze(0)
    idx = idx.clamp(0, f - 1)
    return feat[idx].reshape(t, m * feat.shape[1]).contiguous()


def apply_cmvn(feat: torch.Tensor, shift: torch.Tensor, scale: torch.Tensor) -> torch.Tensor:
    """funasr/frontends/wav_frontend.py:41-55: (x + shift) * scale."""
    d = feat.shape[1]
    return (feat + shift[:d]) * scale[:d]


def wav_frontend(
    wave: torch.Tensor,
    cmvn: Optional[Tuple[torch.Tensor, torch.Tensor]] = None,
    window: str = "hamming",
    lfr_m: int = 7,
    lfr_n: int = 6,
    n_mels: int = 80,
    fs: int = 16000,
) -> torch.Tensor:
    """WavFrontend.forward for one utterance (funasr/frontends/wav_frontend.py:118-158);
    wave in [-1, 1], upscaled by 1 << 15 (:130-131).  Returns [T, n_mels * lfr_m]."""
    w = wave.to(torch.float32) * (1 << 15)
    frame_len = min(25.0, w.numel() / fs * 1000)
    fb = kaldi_fbank(w, fs=fs, n_mels=n_mels, frame_length_ms=frame_len, window=window)
    out = apply_lfr(fb, lfr_m, lfr_n) if (lfr_m != 1 or lfr_n != 1) else fb
    if cmvn is not None:
        out = apply_cmvn(out, cmvn[0], cmvn[1])
    return out


# --------------------------------------------------------------------------------------
# SAN-M encoder
# --------------------------------------------------------------------------------------


def sinusoidal_pe(t: int, depth: int) -> torch.Tensor:
    """funasr/models/transformer/embedding.py:383-413: positions 1..T, [sin | cos]."""
    pos = torch.arange(1, t + 1, dtype=torch.float32)
    inc = torch.log(torch.tensor([10000.0])) / (depth / 2 - 1)
    inv = torch.exp(torch.arange(depth / 2).to(torch.float32) * (-inc))
    st = pos.reshape(-1, 1) * inv.reshape(1, -1)
    return torch.cat([torch.sin(st), torch.cos(st)], dim=1)


def layer_norm(x: torch.Tensor, sd: SD, prefix: str, eps: float) -> torch.Tensor:
    return F.layer_norm(x, (x.shape[-1],), sd[prefix + "weight"], sd[prefix + "bias"], eps)


def fsmn_memory(v: torch.Tensor, weight: torch.Tensor) -> torch.Tensor:
    """funasr/models/sanm/attention.py:207-223 for one unpadded utterance (mask all ones):
    zero pad ((k-1)//2, k-1-(k-1)//2), depthwise conv, + v.   v: [T, D]; weight: [D, 1, k]."""
    k = weight.shape[-1]
    left = (k - 1) // 2
    x = F.pad(v.T.unsqueeze(0), (left, k - 1 - left))
    x = F.conv1d(x, weight, groups=weight.shape[0]).squeeze(0).T
    return x + v


def mha_core(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int) -> torch.Tensor:
    """q: [L, D], k, v: [T, D] -> [L, D]; scale d_k^-0.5 applied to q first
    (funasr/models/sanm/attention.py:306-311, :254-290)."""
    l, d = q.shape
    t = k.shape[0]
    dk = d // heads
    qh = q.reshape(l, heads, dk).transpose(0, 1) * dk ** (-0.5)
    kh = k.reshape(t, heads, dk).transpose(0, 1)
    vh = v.reshape(t, heads, dk).transpose(0, 1)
    att = torch.softmax(torch.matmul(qh, kh.transpose(-2, -1)), dim=-1)
    return torch.matmul(att, vh).transpose(0, 1).reshape(l, d)


def sanm_attention(x: torch.Tensor, sd: SD, p: str, heads: int) -> torch.Tensor:
    """MultiHeadedAttentionSANM.forward (funasr/models/sanm/attention.py:292-311)."""
    qkv = F.linear(x, sd[p + "linear_q_k_v.weight"], sd[p + "linear_q_k_v.bias"])
    q, k, v = torch.split(qkv, qkv.shape[-1] // 3, dim=-1)
    mem = fsmn_memory(v, sd[p + "fsmn_block.weight"])
    ctx = mha_core(q, k, v, heads)
    return F.linear(ctx, sd[p + "linear_out.weight"], sd[p + "linear_out.bias"]) + mem


def encoder_layer(x: torch.Tensor, sd: SD, p: str, heads: int, eps: float) -> torch.Tensor:
    """EncoderLayerSANM.forward (funasr/models/sanm/encoder.py:72-148), eval, pre-norm."""
    size = sd[p + "norm2.weight"].shape[0]
    a = sanm_attention(layer_norm(x, sd, p + "norm1.", eps), sd, p + "self_attn.", heads)
    x = x + a if x.shape[-1] == size else a  # no residual when in_size != size (:120-137)
    h = layer_norm(x, sd, p + "norm2.", eps)
    h = F.linear(torch.relu(F.linear(h, sd[p + "feed_forward.w_1.weight"], sd[p + "feed_forward.w_1.bias"])),
                 sd[p + "feed_forward.w_2.weight"], sd[p + "feed_forward.w_2.bias"])
    return x + h


def count_layers(sd: SD, prefix: str) -> int:
    n = 0
    while (prefix + "%d.norm1.weight" % n) in sd:
        n += 1
    return n


def sanm_encoder(feats: torch.Tensor, sd: SD, prefix: str = "encoder.", heads: int = 4,
                 eps: float = 1e-12, taps: Optional[List[torch.Tensor]] = None) -> torch.Tensor:
    """SANMEncoder.forward (funasr/models/sanm/encoder.py:361-430), input_layer 'pe'."""
    size = sd[prefix + "after_norm.weight"].shape[0]
    x = feats * size ** 0.5
    x = x + sinusoidal_pe(x.shape[0], x.shape[1])
    for group in ("encoders0.", "encoders."):
        for i in range(count_layers(sd, prefix + group)):
            x = encoder_layer(x, sd, prefix + group + "%d." % i, heads, eps)
            if taps is not None:
                taps.append(x)
    return layer_norm(x, sd, prefix + "after_norm.", eps)


# --------------------------------------------------------------------------------------
# CIF predictor
# --------------------------------------------------------------------------------------


def cif_alphas(enc: torch.Tensor, sd: SD, prefix: str = "predictor.", smooth: float = 1.0,
               noise: float = 0.0) -> torch.Tensor:
    """alpha estimation, funasr/models/paraformer/cif_predictor.py:210-222 for one
    unpadded utterance: zero pad (l_order, r_order) -> conv1d k=3 -> relu -> linear -> sigmoid."""
    w = sd[prefix + "cif_conv1d.weight"]
    k = w.shape[-1]
    lo = (k - 1) // 2
    x = F.pad(enc.T.unsqueeze(0), (lo, k - 1 - lo))
    y = torch.relu(F.conv1d(x, w, sd[prefix + "cif_conv1d.bias"])).squeeze(0).T
    a = torch.sigmoid(F.linear(y, sd[prefix + "cif_output.weight"], sd[prefix + "cif_output.bias"]))
    return torch.relu(a * smooth - noise).squeeze(-1)


def cif_fire(alphas_tail: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """cif_wo_hidden_v1 (funasr/models/paraformer/cif_predictor.py:668-693), threshold is
    hard-wired to 1.0 there.  Returns fires [T+1] f32, fire mask [T+1] bool, prefix sum f32."""
    cs = torch.cumsum(alphas_tail, dim=0, dtype=torch.float64).to(torch.float32)
    fl = torch.floor(cs)
    prev = torch.roll(fl, 1)
    prev[0] = 0
    fire = (fl - prev) > 0
    fires = torch.zeros_like(cs)
    fires[fire] = 1
    fires = fires + cs - fl
    return fires, fire, cs


def cif_integrate(hidden_tail: torch.Tensor, alphas_tail: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """cif_v1 (funasr/models/paraformer/cif_predictor.py:696-735), batch 1.
    Returns frames [n_fires, D] and fires [T+1]."""
    fires, fire, _ = cif_fire(alphas_tail)
    psh = torch.cumsum(alphas_tail.unsqueeze(-1) * hidden_tail, dim=0)
    frames = psh[fire]
    shift_frames = torch.roll(frames, 1, dims=0)
    if frames.shape[0] > 0:
        shift_frames[0] = 0
    remains = fires - torch.floor(fires)
    remain_frames = remains[fire].unsqueeze(-1) * hidden_tail[fire]
    shift_remain = torch.roll(remain_frames, 1, dims=0)
    if frames.shape[0] > 0:
        shift_remain[0] = 0
    frames = frames - shift_frames + shift_remain - remain_frames
    return frames, fires


def cif_predictor(enc: torch.Tensor, sd: SD, prefix: str = "predictor.", tail_threshold: float = 0.45):
    """CifPredictorV2.forward at inference (cif_predictor.py:202-253) + tail_process_fn
    (:346-370).  Returns (acoustic_embeds [L, D], token_num L (int), alphas [T+1], fires [T+1])."""
    alphas = cif_alphas(enc, sd, prefix)
    t, d = enc.shape
    tail = torch.zeros(1)
    tail[0] = tail_threshold
    alphas_tail = torch.cat([alphas, tail])  # alpha[T] = 0 + 0.45
    hidden_tail = torch.cat([enc, torch.zeros(1, d)], dim=0)
    token_num = int(torch.floor(alphas_tail.sum()).item())
    frames, fires = cif_integrate(hidden_tail, alphas_tail)
    max_label_len = int(torch.round(alphas_tail.sum()).int().item())
    embeds = torch.zeros(max_label_len, d)
    embeds[: frames.shape[0]] = frames[:max_label_len]
    return embeds[:token_num], token_num, alphas_tail, fires


# --------------------------------------------------------------------------------------
# Paraformer SAN-M decoder
# --------------------------------------------------------------------------------------


def decoder_ffn(x: torch.Tensor, sd: SD, p: str, eps: float) -> torch.Tensor:
    """PositionwiseFeedForwardDecoderSANM (funasr/models/sanm/positionwise_feed_forward.py:12-33)."""
    h = torch.relu(F.linear(x, sd[p + "w_1.weight"], sd[p + "w_1.bias"]))
    h = layer_norm(h, sd, p + "norm.", eps)
    return F.linear(h, sd[p + "w_2.weight"])


def cross_attention(x: torch.Tensor, memory: torch.Tensor, sd: SD, p: str, heads: int) -> torch.Tensor:
    """MultiHeadedAttentionCrossAtt.forward (funasr/models/sanm/attention.py:700-717)."""
    q = F.linear(x, sd[p + "linear_q.weight"], sd[p + "linear_q.bias"])
    kv = F.linear(memory, sd[p + "linear_k_v.weight"], sd[p + "linear_k_v.bias"])
    k, v = torch.split(kv, kv.shape[-1] // 2, dim=-1)
    ctx = mha_core(q, k, v, heads)
    return F.linear(ctx, sd[p + "linear_out.weight"], sd[p + "linear_out.bias"])


def decoder_layer(x: torch.Tensor, memory: torch.Tensor, sd: SD, p: str, heads: int, eps: float) -> torch.Tensor:
    """DecoderLayerSANM.forward (funasr/models/paraformer/decoder.py:78-121)."""
    t = decoder_ffn(layer_norm(x, sd, p + "norm1.", eps), sd, p + "feed_forward.", eps)
    if (p + "self_attn.fsmn_block.weight") not in sd:
        return t  # decoders3: FFN only, no residual (:102-108)
    y = x + fsmn_memory(layer_norm(t, sd, p + "norm2.", eps), sd[p + "self_attn.fsmn_block.weight"])
    if (p + "src_attn.linear_q.weight") in sd:
        y = y + cross_attention(layer_norm(y, sd, p + "norm3.", eps), memory, sd, p + "src_attn.", heads)
    return y


def paraformer_decoder(enc: torch.Tensor, embeds: torch.Tensor, sd: SD, prefix: str = "decoder.",
                       heads: int = 4, eps: float = 1e-12) -> torch.Tensor:
    """ParaformerSANMDecoder.forward (funasr/models/paraformer/decoder.py:359-411) -> logits [L, V]."""
    x = embeds
    for group in ("decoders.", "decoders2.", "decoders3."):
        for i in range(count_layers(sd, prefix + group)):
            x = decoder_layer(x, enc, sd, prefix + group + "%d." % i, heads, eps)
    x = layer_norm(x, sd, prefix + "after_norm.", eps)
    return F.linear(x, sd[prefix + "output_layer.weight"], sd[prefix + "output_layer.bias"])


def greedy_pick(logits: torch.Tensor, drop=(0, 1, 2)) -> Tuple[List[int], float]:
    """Paraformer.inference greedy branch (funasr/models/paraformer/model.py:527-565):
    argmax of log_softmax per token, score = sum of max log-probs, drop blank/sos/eos."""
    lp = torch.log_softmax(logits, dim=-1)
    ids = lp.argmax(dim=-1).tolist()
    score = float(lp.max(dim=-1)[0].sum().item()) if logits.shape[0] else 0.0
    return [i for i in ids if i not in drop], score


def ctc_greedy(logits: torch.Tensor, blank: int = 0, drop=(0,)) -> List[int]:
    """argmax -> unique_consecutive -> drop blank (funasr/models/sense_voice/model.py:895-906;
    funasr/models/ctc/model.py:222-247 drops {0, 1, 2})."""
    ids = torch.unique_consecutive(logits.argmax(dim=-1), dim=-1).tolist()
    return [i for i in ids if i not in drop]


# --------------------------------------------------------------------------------------
# Whole path
# --------------------------------------------------------------------------------------


def recognize(wave: torch.Tensor, sd: SD, cmvn=None, window: str = "hamming", heads: int = 4,
              taps: Optional[dict] = None) -> List[int]:
    """Paraformer.inference for one utterance (funasr/models/paraformer/model.py:443-596),
    token ids only (tokenizer/text post-processing is host code outside the path)."""
    with torch.no_grad():
        feats = wav_frontend(wave, cmvn, window)
        enc = sanm_encoder(feats, sd, heads=heads)
        embeds, token_num, alphas, fires = cif_predictor(enc, sd)
        if taps is not None:
            taps.update(feats=feats, enc=enc, embeds=embeds, token_num=token_num, alphas=alphas, fires=fires)
        if token_num < 1:
            return []  # model.py:514-515
        logits = paraformer_decoder(enc, embeds, sd, heads=heads)
        ids, score = greedy_pick(logits)
        if taps is not None:
            taps.update(logits=logits, score=score)
        return ids

# --------------------------------------------------------------------------------------
# SenseVoiceSmall (CTC) path -- BASELINE config 3, SURVEY.md 8a rows a22/a23
# --------------------------------------------------------------------------------------
SENSE_VOICE_LID = {"auto": 0, "zh": 3, "en": 4, "yue": 7, "ja": 11, "ko": 12, "nospeech": 13}  # sense_voice/model.py:642
SENSE_VOICE_TEXTNORM = {"withitn": 14, "woitn": 15}                                             # :644


def sense_voice_prompt(sd: SD, language: str = "auto", textnorm: str = "woitn") -> torch.Tensor:
    """The 4 query rows prepended to the features (funasr/models/sense_voice/model.py:852-876):
    [language, event (id 1), emotion (id 2), text-norm] rows of `embed.weight` [16, 560]."""
    e = sd["embed.weight"]
    ids = [SENSE_VOICE_LID.get(language, 0), 1, 2, SENSE_VOICE_TEXTNORM[textnorm]]
    return e[torch.tensor(ids)]


def sense_voice_encoder(x: torch.Tensor, sd: SD, prefix: str = "encoder.", heads: int = 4,
                        eps: float = 1e-5) -> torch.Tensor:
    """SenseVoiceEncoderSmall.forward (funasr/models/sense_voice/model.py:553-585) for one utterance:
    x [T + 4, 560] (prompt rows included) -> [T + 4, 512].  Same SAN-M blocks as the Paraformer
    encoder, LayerNorm eps 1e-5 (:275-287), then the 20 `tp_encoders` and `tp_norm`."""
    x = sanm_encoder(x, sd, prefix=prefix, heads=heads, eps=eps)
    for i in range(count_layers(sd, prefix + "tp_encoders.")):
        x = encoder_layer(x, sd, prefix + "tp_encoders.%d." % i, heads, eps)
    return layer_norm(x, sd, prefix + "tp_norm.", eps)


def sense_voice_recognize(wave: torch.Tensor, sd: SD, cmvn=None, window: str = "hamming", heads: int = 4,
                          language: str = "auto", textnorm: str = "woitn", taps: Optional[dict] = None) -> List[int]:
    """SenseVoiceSmall.inference for one utterance (funasr/models/sense_voice/model.py:836-906):
    frontend -> prompt rows -> encoder -> ctc_lo -> log_softmax -> argmax -> unique_consecutive ->
    drop blank 0 (funasr/models/ctc/ctc.py:173-197)."""
    with torch.no_grad():
        feats = wav_frontend(wave, cmvn, window)
        x = torch.cat([sense_voice_prompt(sd, language, textnorm), feats], dim=0)
        enc = sense_voice_encoder(x, sd, heads=heads)
        logits = torch.log_softmax(F.linear(enc, sd["ctc.ctc_lo.weight"], sd["ctc.ctc_lo.bias"]), dim=-1)
        if taps is not None:
            taps.update(feats=feats, enc=enc, logits=logits)
        return ctc_greedy(logits, blank=0, drop=(0,))
