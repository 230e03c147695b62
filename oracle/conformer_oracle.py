"""CPU restatement (torch fp32, batch 1) of the reference's Conformer encoder + CTC greedy path --
BASELINE config 4.  TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's CPU legs
may import this module; the product path (funasr_b200/) never does.

Pinned: tests/golden/make_golden_conformer.py ran the UNMODIFIED reference (`model_classes["CTC"]` with
`ConformerEncoder`, aishell recipe options) in the build container and committed encoder outputs, per-frame
argmax and collapsed ids (tests/golden/conformer.npz); tests/test_oracle_conformer.py checks every function
here against them.  All file:line citations are into /root/reference.

Shapes: feats [F, 80] (WavFrontend with lfr_m = lfr_n = 1, no CMVN), T1 = (F - 1) // 2, T2 = (T1 - 1) // 2,
d = 256, 4 heads of 64, FFN 2048, depthwise kernel 15, 12 blocks.
"""
import math
from typing import Dict, List

import torch
import torch.nn.functional as F

from . import paraformer_oracle as P

SD = Dict[str, torch.Tensor]
EPS = 1e-12  # funasr/models/transformer/layer_norm.py:13-39


def conv2d_subsampling(feats: torch.Tensor, sd: SD, prefix: str = "encoder.embed.") -> torch.Tensor:
    """Conv2dSubsampling.forward without the position module
    (funasr/models/transformer/utils/subsampling.py:63-72,88-91): two 3x3 stride-2 convolutions with ReLU,
    then Linear over the (channel, frequency) pairs in channel-major order."""
    x = feats[None, None]  # (b, c, t, f)
    x = F.relu(F.conv2d(x, sd[prefix + "conv.0.weight"], sd[prefix + "conv.0.bias"], stride=2))
    x = F.relu(F.conv2d(x, sd[prefix + "conv.2.weight"], sd[prefix + "conv.2.bias"], stride=2))
    b, c, t, f = x.shape
    x = x.transpose(1, 2).contiguous().view(t, c * f)
    return F.linear(x, sd[prefix + "out.0.weight"], sd[prefix + "out.0.bias"])


def legacy_rel_positions(t: int, d: int, max_len: int = 5000) -> torch.Tensor:
    """LegacyRelPositionalEncoding's `pos_emb` (funasr/models/transformer/embedding.py:59-78,215-252):
    a table of REVERSED positions max_len-1 ... 0 built once for max_len = 5000 (regrown only for longer
    inputs), of which the first t rows are handed to every layer -- row i encodes position L - 1 - i,
    L = max(max_len, t), not t - 1 - i."""
    length = max(max_len, t)
    position = torch.arange(length - 1, -1, -1.0, dtype=torch.float32).unsqueeze(1)
    div_term = torch.exp(torch.arange(0, d, 2, dtype=torch.float32) * -(math.log(10000.0) / d))
    pe = torch.zeros(length, d)
    pe[:, 0::2] = torch.sin(position * div_term)
    pe[:, 1::2] = torch.cos(position * div_term)
    return pe[:t]


def rel_shift_legacy(x: torch.Tensor) -> torch.Tensor:
    """LegacyRelPositionMultiHeadedAttention.rel_shift (funasr/models/transformer/attention.py:249-269) for
    x [h, t, t]: out[i, j] = x[i, t-1-i+j] for j <= i, 0 for j = i+1, x[i+1, j-i-2] for j >= i+2."""
    h, t1, t2 = x.shape
    padded = torch.cat([torch.zeros(h, t1, 1, dtype=x.dtype), x], dim=-1)
    padded = padded.view(h, t2 + 1, t1)
    return padded[:, 1:].reshape(h, t1, t2)


def rel_attention(x: torch.Tensor, pos_emb: torch.Tensor, sd: SD, p: str, heads: int) -> torch.Tensor:
    """LegacyRelPositionMultiHeadedAttention.forward (attention.py:271-312) + forward_attention (:69-102)."""
    t, d = x.shape
    dk = d // heads
    q = F.linear(x, sd[p + "linear_q.weight"], sd[p + "linear_q.bias"]).view(t, heads, dk)
    k = F.linear(x, sd[p + "linear_k.weight"], sd[p + "linear_k.bias"]).view(t, heads, dk).transpose(0, 1)
    v = F.linear(x, sd[p + "linear_v.weight"], sd[p + "linear_v.bias"]).view(t, heads, dk).transpose(0, 1)
    pp = F.linear(pos_emb, sd[p + "linear_pos.weight"]).view(t, heads, dk).transpose(0, 1)
    q_u = (q + sd[p + "pos_bias_u"]).transpose(0, 1)
    q_v = (q + sd[p + "pos_bias_v"]).transpose(0, 1)
    ac = torch.matmul(q_u, k.transpose(-2, -1))
    bd = rel_shift_legacy(torch.matmul(q_v, pp.transpose(-2, -1)))
    attn = torch.softmax((ac + bd) / math.sqrt(dk), dim=-1)
    ctx = torch.matmul(attn, v).transpose(0, 1).contiguous().view(t, d)
    return F.linear(ctx, sd[p + "linear_out.weight"], sd[p + "linear_out.bias"])


def swish(x: torch.Tensor) -> torch.Tensor:
    """Swish with beta = 1, use_builtin = False (funasr/models/transformer/utils/nets_utils.py:496-528)."""
    return x * torch.sigmoid(x)


def feed_forward(x: torch.Tensor, sd: SD, p: str) -> torch.Tensor:
    """PositionwiseFeedForward with swish (funasr/models/transformer/positionwise_feed_forward.py:14-34)."""
    return F.linear(swish(F.linear(x, sd[p + "w_1.weight"], sd[p + "w_1.bias"])), sd[p + "w_2.weight"], sd[p + "w_2.bias"])


def conv_module(x: torch.Tensor, sd: SD, p: str, bn_eps: float = 1e-5) -> torch.Tensor:
    """ConvolutionModule.forward (funasr/models/conformer/encoder.py:98-120): pointwise -> GLU -> depthwise
    k = 15 ('same' zero padding) -> BatchNorm1d (eval: running statistics) -> swish -> pointwise."""
    y = x.t().unsqueeze(0)  # (1, C, T)
    y = F.conv1d(y, sd[p + "pointwise_conv1.weight"], sd[p + "pointwise_conv1.bias"])
    y = F.glu(y, dim=1)
    w = sd[p + "depthwise_conv.weight"]
    y = F.conv1d(y, w, sd[p + "depthwise_conv.bias"], padding=(w.shape[-1] - 1) // 2, groups=w.shape[0])
    y = F.batch_norm(y, sd[p + "norm.running_mean"], sd[p + "norm.running_var"], sd[p + "norm.weight"],
                     sd[p + "norm.bias"], training=False, eps=bn_eps)
    y = swish(y)
    y = F.conv1d(y, sd[p + "pointwise_conv2.weight"], sd[p + "pointwise_conv2.bias"])
    return y.squeeze(0).t()


def conformer_layer(x: torch.Tensor, pos_emb: torch.Tensor, sd: SD, p: str, heads: int) -> torch.Tensor:
    """EncoderLayer.forward, eval, normalize_before, macaron, cnn module (encoder.py:183-283)."""
    x = x + 0.5 * feed_forward(P.layer_norm(x, sd, p + "norm_ff_macaron.", EPS), sd, p + "feed_forward_macaron.")
    x = x + rel_attention(P.layer_norm(x, sd, p + "norm_mha.", EPS), pos_emb, sd, p + "self_attn.", heads)
    x = x + conv_module(P.layer_norm(x, sd, p + "norm_conv.", EPS), sd, p + "conv_module.")
    x = x + 0.5 * feed_forward(P.layer_norm(x, sd, p + "norm_ff.", EPS), sd, p + "feed_forward.")
    return P.layer_norm(x, sd, p + "norm_final.", EPS)


def count_blocks(sd: SD, prefix: str) -> int:
    n = 0
    while (prefix + "encoders.%d.norm_ff.weight" % n) in sd:
        n += 1
    return n


def conformer_encoder(feats: torch.Tensor, sd: SD, prefix: str = "encoder.", heads: int = 4, taps=None) -> torch.Tensor:
    """ConformerEncoder.forward (encoder.py:530-607) for one utterance: feats [F, 80] -> [T2, 256]."""
    if feats.shape[0] < 7:  # check_short_utt (subsampling.py:36-46): the reference raises TooShortUttError
        raise ValueError("has %d frames and is too short for subsampling (it needs more than 7 frames)" % feats.shape[0])
    x = conv2d_subsampling(feats, sd, prefix + "embed.")
    d = x.shape[1]
    x = x * math.sqrt(d)  # LegacyRelPositionalEncoding.forward (embedding.py:238-252): x * xscale, pos_emb apart
    pos_emb = legacy_rel_positions(x.shape[0], d)
    if taps is not None:
        taps.append(x.clone())
    for i in range(count_blocks(sd, prefix)):
        x = conformer_layer(x, pos_emb, sd, prefix + "encoders.%d." % i, heads)
        if taps is not None:
            taps.append(x.clone())
    return P.layer_norm(x, sd, prefix + "after_norm.", EPS)


def ctc_logits(enc: torch.Tensor, sd: SD, prefix: str = "ctc.") -> torch.Tensor:
    """CTC.log_softmax's projection (funasr/models/ctc/ctc.py:173-184); argmax is taken on the logits
    (log_softmax is monotone per row)."""
    return F.linear(enc, sd[prefix + "ctc_lo.weight"], sd[prefix + "ctc_lo.bias"])


def recognize(wave: torch.Tensor, sd: SD, heads: int = 4) -> List[int]:
    """`model_classes["CTC"].inference` (funasr/models/ctc/model.py:162-247) for one waveform: WavFrontend
    (hamming, 80 mel, lfr 1/1, no CMVN, dither 0) -> ConformerEncoder -> ctc_lo -> argmax -> unique_consecutive
    -> drop blank / sos / eos (0, 1, 2)."""
    feats = P.wav_frontend(wave, None, window="hamming", lfr_m=1, lfr_n=1)
    enc = conformer_encoder(feats, sd, heads=heads)
    return P.ctc_greedy(ctc_logits(enc, sd), blank=0, drop=(0, 1, 2))
