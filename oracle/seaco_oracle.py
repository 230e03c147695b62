"""CPU restatement of SeacoParaformer's hot-word branch -- TEST INFRASTRUCTURE ONLY.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU legs may import this module.  Pinned against the reference
itself: `tests/golden/seaco.npz` (made by `tests/golden/make_golden_seaco.py`) holds the unmodified
`SeacoParaformer`'s hot-word embeddings, decoder hidden states and token ids WITH hot words;
`tests/test_oracle_seaco.py` checks this file against them.

What the branch adds to BiCifParaformer (funasr/models/seaco_paraformer/model.py:233-359), for one utterance:
  * `_hotword_representation` (:329-359): token embedding (`decoder.embed`) of every hot word -> 2-layer unidirectional
    LSTM (`bias_encoder`) -> the top layer's state at the word's last token;
  * the SeACo decoder (`seaco_decoder`, a ParaformerSANMDecoder without input / output layer) run twice with the hot-word
    embeddings as memory: over the CIF embeddings and over the decoder's hidden states; the two results are added (:178-179);
  * `hotword_output_layer` -> log-softmax; where its argmax is NO_BIAS the plain decoder's log-softmax is kept, elsewhere
    the hot-word distribution replaces it (`_merge_res`, :312-322, with seaco_weight = 1).
  * with more than `nfilter` = 50 hot words the attention-score filter (ASF, :268-295) first keeps the 50 hot words the
    decoder's hidden states attend to most, plus the trailing [sos].
"""
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

from . import bicif_oracle as B
from . import paraformer_oracle as P

SD = Dict[str, torch.Tensor]
NO_BIAS = 8377
NFILTER = 50


def lstm_layer(x: torch.Tensor, sd: SD, prefix: str, layer: int) -> torch.Tensor:
    s = "_l%d" % layer
    return B.lstm_direction(x, sd[prefix + "weight_ih" + s], sd[prefix + "weight_hh" + s], sd[prefix + "bias_ih" + s],
                            sd[prefix + "bias_hh" + s], False)


def hotword_embeddings(hotwords: Sequence[Sequence[int]], sd: SD) -> torch.Tensor:
    """`_hotword_representation` (model.py:329-359): [N, D], one row per hot word (the list ends with [sos])."""
    rows = []
    for ids in hotwords:
        x = sd["decoder.embed.0.weight"][torch.tensor(list(ids), dtype=torch.long)]
        x = lstm_layer(x, sd, "bias_encoder.", 0)
        x = lstm_layer(x, sd, "bias_encoder.", 1)
        rows.append(x[-1])
    return torch.stack(rows)


def decoder_hidden(memory: torch.Tensor, x: torch.Tensor, sd: SD, prefix: str, heads: int = 4, eps: float = 1e-12) -> torch.Tensor:
    """ParaformerSANMDecoder.forward up to `after_norm` (funasr/models/paraformer/decoder.py:395-404)."""
    for group in ("decoders.", "decoders2.", "decoders3."):
        for i in range(P.count_layers(sd, prefix + group)):
            x = P.decoder_layer(x, memory, sd, prefix + group + "%d." % i, heads, eps)
    return P.layer_norm(x, sd, prefix + "after_norm.", eps)


def asf_filter(ctx: torch.Tensor, hidden: torch.Tensor, sd: SD, heads: int = 4, eps: float = 1e-12, nfilter: int = NFILTER,
               prefix: str = "seaco_decoder.") -> List[int]:
    """The attention-score filter (model.py:268-295; `forward_asf6`, paraformer/decoder.py:439-458; `get_attn_mat`, :123-137):
    the decoder's hidden states go through SeACo decoder layers 0-4 with ALL hot words as memory; the attention
    probabilities of layer 5's cross-attention are summed over heads and tokens; the `nfilter` best hot words (never
    the trailing [sos]... which is always appended) are kept.  Returns the kept row indices: top-k order, then N - 1.
    (In a padded batch the reference scores utterance 0 only and applies its choice to every utterance; this is the
    batch-1 form.)"""
    x = hidden
    for i in range(5):
        x = P.decoder_layer(x, ctx, sd, prefix + "decoders.%d." % i, heads, eps)
    p = prefix + "decoders.5."
    t = P.decoder_ffn(P.layer_norm(x, sd, p + "norm1.", eps), sd, p + "feed_forward.", eps)
    y = x + P.fsmn_memory(P.layer_norm(t, sd, p + "norm2.", eps), sd[p + "self_attn.fsmn_block.weight"])
    q = F.linear(P.layer_norm(y, sd, p + "norm3.", eps), sd[p + "src_attn.linear_q.weight"], sd[p + "src_attn.linear_q.bias"])
    kv = F.linear(ctx, sd[p + "src_attn.linear_k_v.weight"], sd[p + "src_attn.linear_k_v.bias"])
    k = kv[:, : kv.shape[-1] // 2]
    dk = q.shape[-1] // heads
    scores = torch.zeros(ctx.shape[0])
    for h in range(heads):
        att = torch.softmax((q[:, h * dk:(h + 1) * dk] * dk ** -0.5) @ k[:, h * dk:(h + 1) * dk].T, dim=-1)
        scores = scores + att.sum(0)
    n = ctx.shape[0]
    keep = torch.topk(scores, min(nfilter, n - 1))[1].tolist()
    keep.append(n - 1)
    return keep


def seaco_decode(enc: torch.Tensor, embeds: torch.Tensor, sd: SD, hotwords: Optional[Sequence[Sequence[int]]], heads: int = 4,
                 taps: Optional[dict] = None) -> torch.Tensor:
    """`_seaco_decode_with_ASF` (model.py:233-327) -> merged log-probabilities [L, V]."""
    hidden = decoder_hidden(enc, embeds, sd, "decoder.", heads)
    dec_pred = torch.log_softmax(F.linear(hidden, sd["decoder.output_layer.weight"], sd["decoder.output_layer.bias"]), dim=-1)
    if hotwords is None:
        return dec_pred
    ctx = hotword_embeddings(hotwords, sd)
    if len(hotwords) > NFILTER:
        keep_rows = asf_filter(ctx, hidden, sd, heads)
        if taps is not None:
            taps.update(asf_keep=keep_rows)
        ctx = ctx[keep_rows]
    cif_att = decoder_hidden(ctx, embeds, sd, "seaco_decoder.", heads)
    dec_att = decoder_hidden(ctx, hidden, sd, "seaco_decoder.", heads)
    merged = cif_att + dec_att
    dha_pred = torch.log_softmax(F.linear(merged, sd["hotword_output_layer.weight"], sd["hotword_output_layer.bias"]), dim=-1)
    keep = (dha_pred.argmax(dim=-1) == NO_BIAS).to(dec_pred.dtype).unsqueeze(-1)
    if taps is not None:
        taps.update(hw_embed=ctx, hidden=hidden, merged=merged, dha_ids=dha_pred.argmax(dim=-1))
    return dec_pred * keep + dha_pred * (1 - keep)


def recognize(wave: torch.Tensor, sd: SD, cmvn=None, hotwords: Optional[Sequence[Sequence[int]]] = None, window: str = "hamming",
              heads: int = 4, taps: Optional[dict] = None) -> Tuple[List[int], List[List[int]]]:
    """SeacoParaformer.inference for one utterance (model.py:361-510): (token ids, time stamps)."""
    with torch.no_grad():
        feats = P.wav_frontend(wave, cmvn, window)
        enc = P.sanm_encoder(feats, sd, heads=heads)
        embeds, token_num, _, _ = B.cif_predictor_v3(enc, sd)
        if token_num < 1:
            return [], []
        pred = seaco_decode(enc, embeds, sd, hotwords, heads, taps)
        ids = [i for i in pred.argmax(dim=-1).tolist() if i not in (0, 1, 2)]
        us_alphas, us_peaks = B.upsample_timestamp(enc, sd, token_num)
        if taps is not None:
            taps.update(enc=enc, embeds=embeds, pred=pred)
        return ids, B.ts_prediction(us_alphas, us_peaks, len(ids))
