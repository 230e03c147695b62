"""SeacoParaformer at inference (funasr/models/seaco_paraformer/model.py:361-510) on one B200: BiCifParaformer's path
(`bicif.BiCifParaformerEngine`) plus the hot-word branch of `_seaco_decode_with_ASF` (:233-327).

Per packed batch, when hot words are given:
  hot-word embeddings (`_hotword_representation`, :329-359)   embedding rows of `decoder.embed` -> two unidirectional LSTM layers
      (`bias_encoder`) through `fb200_lstm_bidir` (its forward direction; input projections = one GEMM per layer) -> the
      state at each word's last token.  Computed once per hot-word list and cached.
  plain decoder up to `after_norm`, its hidden states kept in fp32        `DecoderPack.hidden`
  SeACo decoder twice, memory = the hot-word embeddings of every utterance `DecoderPack.hidden` (prefix `seaco_decoder.`)
  hotword_output_layer(cif_attended + dec_attended)                        two GEMMs, the second adding the first (linearity)
  `_merge_res`                                                             `fb200_merge_hotword_ids` on the two greedy picks
With more than `nfilter` = 50 hot words the attention-score filter (ASF, :268-295) first keeps, per utterance, the 50 hot
words its decoder states attend to most (`fb200_asf_scores` over SeACo layer 5's projected queries / keys, top-k on the host).
"""
from typing import Dict, List, Optional, Sequence

import torch

from . import ops, tables
from .bicif import BiCifParaformerEngine
from .engine import DecoderPack, _Linear, _dev

NFILTER = 50


class HotwordPack:
    def __init__(self, sd: Dict[str, torch.Tensor], device, nparts: int = 2, prefix: str = "bias_encoder."):
        self.device, self.nparts = torch.device(device), nparts
        self.embed = _dev(sd["decoder.embed.0.weight"], self.device)
        self.h = sd[prefix + "weight_hh_l0"].shape[1]
        if self.h != 512 or (prefix + "weight_ih_l0_reverse") in sd:
            raise RuntimeError("funasr_b200: bias encoder with hidden size %d / bidirectional is unsupported" % self.h)
        self.layers = []
        layer = 0
        while prefix + "weight_ih_l%d" % layer in sd:
            s = "_l%d" % layer
            # fb200_lstm_bidir runs two directions: both get this layer's weights, the forward one is used
            w_ih = torch.cat([sd[prefix + "weight_ih" + s]] * 2, dim=0)
            b = torch.cat([sd[prefix + "bias_ih" + s] + sd[prefix + "bias_hh" + s]] * 2)
            w_hh = _dev(torch.stack([sd[prefix + "weight_hh" + s]] * 2).contiguous(), self.device)
            self.layers.append((_Linear(w_ih, b, nparts, self.device), w_hh))
            layer += 1
        self.w_out = torch.zeros(2, self.h, dtype=torch.float32, device=self.device)
        self.parts = ops.lstm_parts(self.h)

    def run(self, hotwords: Sequence[Sequence[int]]) -> torch.Tensor:
        """-> fp32 [N, H]: one embedding per hot word."""
        dev, np_, h = self.device, self.nparts, self.h
        lens = [len(w) for w in hotwords]
        if min(lens) < 1:
            raise ValueError("funasr_b200: empty hot word")
        flat = torch.tensor([i for w in hotwords for i in w], dtype=torch.long, device=dev)
        cu = tables.cu_from_lengths(lens, torch.int32, dev)
        x = self.embed.index_select(0, flat)  # token embeddings of every hot word, packed (row gather: data movement only)
        rows = x.shape[0]
        for proj, w_hh in self.layers:
            xs = ops.split_rows(x, np_)
            xg = torch.empty(rows, 8 * h, dtype=torch.float32, device=dev)
            ops.gemm(xs, np_, proj.w, np_, rows, 8 * h, proj.k, bias=proj.bias, out_f32=xg)
            part = torch.empty(self.parts, rows, dtype=torch.float32, device=dev)
            out = torch.empty(rows, 2 * h, dtype=torch.float32, device=dev)
            ops.lstm_bidir(xg, cu, 1, max(lens), w_hh, self.w_out, part, out_h=out, hidden=h, impl=1)  # fp32 recurrence
            x = out[:, :h]
        last = (cu[1:] - 1).long()
        return x.index_select(0, last).contiguous()


class SeacoParaformerEngine(BiCifParaformerEngine):
    """SeacoParaformer.inference for packed batches: token ids and time stamps, with or without hot words."""

    NO_BIAS = 8377  # model_conf NO_BIAS (seaco_paraformer/template.yaml)

    def __init__(self, state_dict, device="cuda:0", precision: str = "fp32", no_bias: int = 8377, **kw):
        super().__init__(state_dict, device, precision, **kw)
        self.NO_BIAS = no_bias
        with torch.cuda.device(self.device):
            self.hot = HotwordPack(state_dict, self.device, nparts=self.nparts)
            self.seaco_decoder = DecoderPack(state_dict, self.device, prefix="seaco_decoder.", heads=self.decoder.heads,
                                             nparts=self.nparts)
            self.hot_output = _Linear(state_dict["hotword_output_layer.weight"], state_dict["hotword_output_layer.bias"],
                                      self.nparts, self.device)
        self._hot_key, self._hot_embed = None, None
        self.hotwords: Optional[List[List[int]]] = None

    def set_hotwords(self, hotwords: Optional[Sequence[Sequence[int]]]) -> None:
        """Hot words as token-id lists; the reference's list ends with [sos] (model.py:570) -- appended here when missing."""
        if hotwords is None:
            self.hotwords = None
            return
        hw = [list(map(int, w)) for w in hotwords]
        if not hw or hw[-1] != [1]:
            hw.append([1])
        if len(hw) > 2048:
            raise NotImplementedError("funasr_b200: %d hot words (at most 2048)" % len(hw))
        self.hotwords = hw

    def _hot_embeddings(self) -> torch.Tensor:
        key = tuple(tuple(w) for w in self.hotwords)
        if key != self._hot_key:
            self._hot_embed, self._hot_key = self.hot.run(self.hotwords), key
        return self._hot_embed

    def decode(self, enc, cu_rows, embeds, cu_tok, max_l, taps=None):
        """`_seaco_decode_with_ASF` on the greedy path -> (ids, logp, logits or None)."""
        dec = self.decoder
        if self.hotwords is None:
            return dec.run(enc, cu_rows, embeds, cu_tok, max_l, want_logits=taps is not None)
        dev, np_, d = self.device, self.nparts, dec.d
        hidden, hidden_split = dec.hidden(enc, cu_rows, embeds, cu_tok, max_l, want_f32=True)
        dec_ids, dec_logp, logits = dec.head(hidden_split, want_logits=taps is not None)
        ctx = self._hot_embeddings()
        b, nh, l = cu_tok.numel() - 1, ctx.shape[0], embeds.shape[0]
        memory = ctx.repeat(b, 1)  # every utterance attends to the same hot-word embeddings
        cu_mem = tables.cu_from_lengths([nh] * b, torch.int32, dev)
        mem_split = ops.split_rows(memory, np_)
        if nh > NFILTER:
            # attention-score filter (model.py:268-295): per utterance, the NFILTER hot words its decoder states attend to most
            # in SeACo layer 5, plus the trailing [sos].  (The reference scores utterance 0 of a padded batch for all of them;
            # this is its batch-1 behaviour.)
            heads = self.seaco_decoder.heads
            q32, k32 = self.seaco_decoder.hidden(memory, cu_mem, hidden, cu_tok, max_l, enc_split=mem_split, query_layer=5)
            scores = torch.empty(b * nh, dtype=torch.float32, device=dev)
            ops.asf_scores(q32, k32, cu_tok, cu_mem, nh, heads, float(d // heads) ** -0.5, scores)
            top = torch.topk(scores.cpu().view(b, nh), min(NFILTER, nh - 1), dim=1)[1]  # the reference's own CPU top-k (:274)
            keep = torch.cat([top, torch.full((b, 1), nh - 1, dtype=top.dtype)], dim=1)  # + the trailing [sos] (:276)
            memory = ctx.index_select(0, keep.reshape(-1).to(dev))
            nh = keep.shape[1]
            cu_mem = tables.cu_from_lengths([nh] * b, torch.int32, dev)
            mem_split = ops.split_rows(memory, np_)
            if taps is not None:
                taps.update(asf_scores=scores.view(b, -1), asf_keep=keep)
        _, cif_att = self.seaco_decoder.hidden(memory, cu_mem, embeds, cu_tok, max_l, enc_split=mem_split)
        _, dec_att = self.seaco_decoder.hidden(memory, cu_mem, hidden, cu_tok, max_l, enc_split=mem_split)
        v = self.hot_output.n
        part = torch.empty(l, v, dtype=torch.float32, device=dev)
        ops.gemm(cif_att, np_, self.hot_output.w, np_, l, v, d, bias=self.hot_output.bias, out_f32=part)
        hw_logits = torch.empty(l, v, dtype=torch.float32, device=dev)
        ops.gemm(dec_att, np_, self.hot_output.w, np_, l, v, d, res1=part, out_f32=hw_logits)  # W (a + b) = W a + W b
        hw_ids = torch.empty(l, dtype=torch.int32, device=dev)
        hw_logp = torch.empty(l, dtype=torch.float32, device=dev)
        ops.argmax_logprob(hw_logits, v, hw_ids, hw_logp)
        ids = torch.empty(l, dtype=torch.int32, device=dev)
        logp = torch.empty(l, dtype=torch.float32, device=dev)
        ops.merge_hotword_ids(dec_ids, dec_logp, hw_ids, hw_logp, self.NO_BIAS, ids, logp)
        if taps is not None:
            taps.update(hw_embed=ctx, hidden=hidden, hw_ids=hw_ids, hw_logits=hw_logits, dec_ids=dec_ids)
        return ids, logp, logits

    def recognize(self, waves, with_scores: bool = False, begin_time: float = 0.0, hotwords=None):
        """As `BiCifParaformerEngine.recognize`; `hotwords` (token-id lists, or None) apply to this call."""
        self.set_hotwords(hotwords)
        try:
            return super().recognize(waves, with_scores=with_scores, begin_time=begin_time)
        finally:
            self.hotwords = None
