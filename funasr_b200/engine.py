"""Packed (var-len) execution of the offline Paraformer path on one B200.

Host-side orchestration only: every numeric step is a hand-written CUDA kernel reached through
the C-ABI (`funasr_b200.ops` -> libfunasr_b200.so).  torch supplies device memory and the stream.

Layout: a batch of B utterances is one packed matrix of R = sum(T_b) rows with int32 prefix
sums `cu_rows` -- the reference's batch_size=1 semantics (funasr/auto/auto_model.py:185-187) for
a whole batch, with none of the padded path's leaks.

Stages (reference call stack):
  frontend   funasr/frontends/wav_frontend.py:118-158  + encoder prologue sanm/encoder.py:378,397
  encode     funasr/models/sanm/encoder.py:361-430
  predict    funasr/models/paraformer/cif_predictor.py:202-253, :346-370, :668-735
  decode     funasr/models/paraformer/decoder.py:359-411 + greedy pick paraformer/model.py:527-565

precision = "fp32": every dense contraction runs on a 2-way bf16 split of both operands (hi + mid,
16 mantissa bits; 3 tcgen05 cross terms, fp32 accumulate) -- parity mode.  "fp32x6": 3-way split,
6 terms (24 mantissa bits).  "bf16": single bf16 operands -- throughput mode.
"""
from typing import Dict, List, Optional, Sequence, Tuple

import os

import numpy as np
import torch

from . import ops, tables

# vocabulary / CTC projections fused with the greedy pick (fb200_gemm_argmax); FB200_FUSED_ARGMAX=0 keeps logits + argmax kernels
FUSED_ARGMAX = os.environ.get("FB200_FUSED_ARGMAX", "1") != "0"
FRAME_LEN = 400
FRAME_SHIFT = 160
NFFT = 512


def _dev(t: torch.Tensor, device) -> torch.Tensor:
    return t.detach().to(device=device, dtype=torch.float32).contiguous()



# Packed encoder rows per batch: three waves of 148 SMs x 128-row GEMM tiles.  Measured on the 1024-utterance set
# (`tools/max_rows_sweep.sh`): 18 944 rows (one wave) 397 ms per step, 37 888 373.6, 56 832 356.7-365.2, 75 776 353.9,
# 136 192 348.3 -- every launch has fixed costs (a GEMM's last epilogue, the attention's pipeline fill and the tail round of
# its items, launch tails of the row-wise kernels) that larger batches spread over more rows; 56 832 keeps element offsets of
# the widest matrix (the decoder's [rows, 2 x 16384] K / V projection) below 2^31 and still leaves ~5 batches per 1024
# utterances for the upload of batch k + 1 to hide under batch k.
DEFAULT_MAX_ROWS = 3 * 148 * 128
MAX_ROWS_LIMIT = 65536  # larger budgets ran in the sweep but their results were never checked: refused

class _Linear:
    """A Linear (or k-tap Conv1d) as a bf16 split operand + fp32 bias."""

    __slots__ = ("w", "bias", "n", "k", "taps")

    def __init__(self, weight: torch.Tensor, bias: Optional[torch.Tensor], nparts: int, device):
        if weight.dim() == 3:  # Conv1d [out, in, taps] -> [taps][out, in]
            w2 = weight.permute(2, 0, 1).contiguous()
            self.taps, self.n, self.k = w2.shape
            w2 = w2.reshape(self.taps * self.n, self.k)
        else:
            self.taps = 1
            self.n, self.k = weight.shape
            w2 = weight
        wd = _dev(w2, device)
        self.w = torch.empty(wd.shape[0], ops.split_width(self.k, nparts), dtype=torch.bfloat16, device=device)
        ops.split_rows(wd, nparts, out=self.w)
        self.bias = None if bias is None else _dev(bias, device)


class _Norm:
    __slots__ = ("g", "b")

    def __init__(self, sd, prefix, device):
        self.g = _dev(sd[prefix + "weight"], device)
        self.b = _dev(sd[prefix + "bias"], device)


class _EncLayer:
    def __init__(self, sd, p, nparts, device):
        self.norm1 = _Norm(sd, p + "norm1.", device)
        self.norm2 = _Norm(sd, p + "norm2.", device)
        self.qkv = _Linear(sd[p + "self_attn.linear_q_k_v.weight"], sd[p + "self_attn.linear_q_k_v.bias"], nparts, device)
        self.out = _Linear(sd[p + "self_attn.linear_out.weight"], sd[p + "self_attn.linear_out.bias"], nparts, device)
        fw = sd[p + "self_attn.fsmn_block.weight"]
        self.fsmn = _dev(fw.reshape(fw.shape[0], fw.shape[-1]), device)
        self.w1 = _Linear(sd[p + "feed_forward.w_1.weight"], sd[p + "feed_forward.w_1.bias"], nparts, device)
        self.w2 = _Linear(sd[p + "feed_forward.w_2.weight"], sd[p + "feed_forward.w_2.bias"], nparts, device)
        self.din = self.qkv.k
        self.d = self.out.n


class _DecLayer:
    def __init__(self, sd, p, nparts, device):
        self.norm1 = _Norm(sd, p + "norm1.", device)
        self.w1 = _Linear(sd[p + "feed_forward.w_1.weight"], sd[p + "feed_forward.w_1.bias"], nparts, device)
        self.ffn_norm = _Norm(sd, p + "feed_forward.norm.", device)
        self.w2 = _Linear(sd[p + "feed_forward.w_2.weight"], sd.get(p + "feed_forward.w_2.bias"), nparts, device)
        self.has_fsmn = (p + "self_attn.fsmn_block.weight") in sd
        self.has_src = (p + "src_attn.linear_q.weight") in sd
        if self.has_fsmn:
            self.norm2 = _Norm(sd, p + "norm2.", device)
            fw = sd[p + "self_attn.fsmn_block.weight"]
            self.fsmn = _dev(fw.reshape(fw.shape[0], fw.shape[-1]), device)
        if self.has_src:
            self.norm3 = _Norm(sd, p + "norm3.", device)
            self.q = _Linear(sd[p + "src_attn.linear_q.weight"], sd[p + "src_attn.linear_q.bias"], nparts, device)
            self.out = _Linear(sd[p + "src_attn.linear_out.weight"], sd[p + "src_attn.linear_out.bias"], nparts, device)


def _count(sd, prefix: str) -> int:
    n = 0
    while (prefix + "%d.norm1.weight" % n) in sd:
        n += 1
    return n


class FrontendPack:
    """WavFrontend (funasr/frontends/wav_frontend.py:78-158) as two kernels over a packed batch:
    fbank, then LFR + CMVN (+ the encoder's x*sqrt(d) + sinusoidal position prologue when asked)."""

    def __init__(self, device, cmvn: Optional[Tuple[torch.Tensor, torch.Tensor]] = None, window: str = "hamming",
                 n_mels: int = 80, lfr_m: int = 7, lfr_n: int = 6, fs: int = 16000, frame_length: float = 25.0,
                 frame_shift: float = 10.0, upscale: bool = True):
        self.device = torch.device(device)
        self.n_mels, self.lfr_m, self.lfr_n, self.fs = n_mels, lfr_m, lfr_n, fs
        self.frame_len = int(fs * frame_length * 0.001)
        self.frame_shift = int(fs * frame_shift * 0.001)
        self.nfft = 1 << (self.frame_len - 1).bit_length()
        self.upscale = float(1 << 15) if upscale else 1.0
        self.window = tables.feature_window(window, self.frame_len).to(self.device)
        self.mel = tables.mel_filterbank(n_mels, self.nfft, float(fs)).to(self.device)
        self.dim = n_mels * lfr_m
        self.shift = self.scale = None
        if cmvn is not None:
            self.shift = _dev(cmvn[0][: self.dim], self.device)
            self.scale = _dev(cmvn[1][: self.dim], self.device)
        self.pe_inv = tables.pe_inverse_timescales(self.dim).to(self.device)

    def plan(self, num_samples: Sequence[int]):
        frames = [tables.num_frames(int(n), self.frame_len, self.frame_shift) for n in num_samples]
        rows = [tables.num_lfr_rows(f, self.lfr_n) for f in frames]
        return frames, rows

    def run(self, wave: torch.Tensor, num_samples: Sequence[int], pe_scale: Optional[float] = None,
            fbank_out: Optional[list] = None, row_skip: int = 0, feats_out: Optional[torch.Tensor] = None):
        """wave: packed fp32 [sum N] on the device.  Returns (feats [R, dim], cu_rows i32 dev, rows list).
        pe_scale given: feats = cmvn(lfr(fbank)) * pe_scale + PE(pos)  (the encoder input).
        row_skip > 0: every utterance gets that many extra rows at its head (counted in `rows`, positions
        included, left for the caller to fill -- the SenseVoice prompt frames)."""
        frames, rows = self.plan(num_samples)
        rows = [r + row_skip for r in rows]
        if min(frames) < 1:
            raise RuntimeError("funasr_b200: an utterance is shorter than one %d-sample frame" % self.frame_len)
        dev = self.device
        cu_samples = tables.cu_from_lengths(num_samples, torch.int64, dev)
        cu_frames = tables.cu_from_lengths(frames, torch.int32, dev)
        cu_rows = tables.cu_from_lengths(rows, torch.int32, dev)
        fb = ops.fbank(wave, cu_samples, cu_frames, int(sum(frames)), self.window, self.mel, self.frame_shift,
                       self.nfft, self.upscale, 0.97)
        if fbank_out is not None:
            fbank_out.append(fb)
        feats = ops.lfr_cmvn(fb, cu_frames, cu_rows, int(sum(rows)), self.lfr_m, self.lfr_n, self.shift, self.scale,
                             self.pe_inv if pe_scale is not None else None, 1.0 if pe_scale is None else pe_scale,
                             row_skip=row_skip,
                             out=None if feats_out is None or feats_out.shape[0] < sum(rows) else feats_out[: int(sum(rows))])
        return feats, cu_rows, rows


class EncoderPack:
    """SANMEncoder (funasr/models/sanm/encoder.py:361-430) over packed rows."""

    def __init__(self, sd: Dict[str, torch.Tensor], device, prefix: str = "encoder.", heads: int = 4,
                 eps: float = 1e-12, nparts: int = 3):
        self.device, self.heads, self.eps, self.nparts = torch.device(device), heads, eps, nparts
        self.layers: List[_EncLayer] = []
        for group in ("encoders0.", "encoders."):
            for i in range(_count(sd, prefix + group)):
                self.layers.append(_EncLayer(sd, prefix + group + "%d." % i, nparts, self.device))
        self.after_norm = _Norm(sd, prefix + "after_norm.", self.device)
        self.tp_layers: List[_EncLayer] = []
        for i in range(_count(sd, prefix + "tp_encoders.")):
            self.tp_layers.append(_EncLayer(sd, prefix + "tp_encoders.%d." % i, nparts, self.device))
        self.tp_norm = _Norm(sd, prefix + "tp_norm.", self.device) if self.tp_layers else None
        self.d = self.layers[0].d
        self.din = self.layers[0].din
        self.ff = self.layers[0].w1.n
        self.pe_inv = tables.pe_inverse_timescales(self.din).to(self.device)

    def prologue(self, feats: torch.Tensor, cu_rows: torch.Tensor) -> torch.Tensor:
        """x * sqrt(d) + PE (sanm/encoder.py:378,397) for callers that hand over plain features."""
        return ops.scale_add_pe(feats, cu_rows, self.pe_inv, float(self.d) ** 0.5)

    def _layer(self, ly: _EncLayer, x: torch.Tensor, cu_rows, max_t: int, ws) -> torch.Tensor:
        r, np_, d = x.shape[0], self.nparts, self.d
        xs = ws["xs"] if ly.din == d else ws["xs0"]
        ops.layernorm(x, ly.norm1.g, ly.norm1.b, self.eps, out_split=xs, nparts=np_)
        qkv, v = ws["qkv"], ws["v"]  # q | k | v as a bf16 split for the attention UMMAs; v also in fp32 for the FSMN
        ops.gemm(xs, np_, ly.qkv.w, np_, r, 3 * d, ly.din, bias=ly.qkv.bias, out_f32=v, f32_col0=2 * d, out_split=qkv,
                 out_parts=np_)
        mem = ws["mem"]
        ops.fsmn(v, cu_rows, ly.fsmn, mem, residual=x if ly.din == d else None)  # conv(v) + v (+ x)
        ctx = ws["xs"]
        ops.attention(qkv, 0, 3 * d, qkv, d, qkv, 2 * d, 3 * d, np_, cu_rows, cu_rows, max_t, self.heads,
                      float(d // self.heads) ** -0.5, out_split=ctx, plan=ws["plan"])
        x1 = ws["x1"]
        ops.gemm(ctx, np_, ly.out.w, np_, r, d, d, bias=ly.out.bias, res1=mem, out_f32=x1)
        ops.layernorm(x1, ly.norm2.g, ly.norm2.b, self.eps, out_split=ws["xs"], nparts=np_)
        ops.gemm(ws["xs"], np_, ly.w1.w, np_, r, self.ff, d, bias=ly.w1.bias, relu=True, out_split=ws["h"],
                 out_parts=np_)
        xo = ws["x0"]
        ops.gemm(ws["h"], np_, ly.w2.w, np_, r, d, self.ff, bias=ly.w2.bias, res1=x1, out_f32=xo)
        return xo

    def record(self, x: torch.Tensor, cu_rows: torch.Tensor, enc: torch.Tensor, halo: Optional[torch.Tensor],
               split: Optional[torch.Tensor]):
        """The layer walk of `run` over capacity-sized buffers, for an open `ops.StackProgram` recording: x [R, din] in,
        after_norm(x) -> enc fp32 [R, d], its split in the CIF halo layout and / or as a plain split (the decoder's memory)."""
        ws = self.workspace(x.shape[0])
        ws["plan"] = None
        for ly in self.layers:
            x = self._layer(ly, x, cu_rows, x.shape[0], ws)
        an = self.after_norm
        ops.layernorm(x, an.g, an.b, self.eps, out_f32=enc, out_split=halo, nparts=self.nparts, halo=halo is not None,
                      cu_rows=cu_rows if halo is not None else None)
        if split is not None:
            ops.layernorm(x, an.g, an.b, self.eps, out_split=split, nparts=self.nparts)

    def workspace(self, r: int):
        dev, np_, d = self.device, self.nparts, self.d
        bf = torch.bfloat16
        return {
            "xs0": torch.empty(r, ops.split_width(self.din, np_), dtype=bf, device=dev),
            "xs": torch.empty(r, ops.split_width(d, np_), dtype=bf, device=dev),
            "h": torch.empty(r, ops.split_width(self.ff, np_), dtype=bf, device=dev),
            "qkv": torch.empty(r, ops.split_width(3 * d, np_), dtype=bf, device=dev),
            "v": torch.empty(r, d, dtype=torch.float32, device=dev),
            "mem": torch.empty(r, d, dtype=torch.float32, device=dev),
            "x0": torch.empty(r, d, dtype=torch.float32, device=dev),
            "x1": torch.empty(r, d, dtype=torch.float32, device=dev),
        }

    def run(self, x: torch.Tensor, cu_rows: torch.Tensor, max_t: int, taps: Optional[list] = None,
            out_halo: Optional[torch.Tensor] = None, tp: bool = True) -> torch.Tensor:
        """x: fp32 [R, din] AFTER the prologue.  Returns after_norm(x) fp32 [R, d]; `out_halo` (bf16
        split, R + B + 1 rows) additionally receives it in the CIF conv's halo layout."""
        r = x.shape[0]
        ws = self.workspace(r)
        # attention work list of this batch, shared by every layer (the SIMT 3-part mode needs none)
        ws["plan"] = ops.attention_plan(cu_rows, cu_rows, r) if self.nparts < 3 else None
        for ly in self.layers:
            x = self._layer(ly, x, cu_rows, max_t, ws)
            if taps is not None:
                taps.append(x.clone())
        enc = torch.empty(r, self.d, dtype=torch.float32, device=self.device)
        use_halo = out_halo is not None and not (tp and self.tp_layers)
        ops.layernorm(x, self.after_norm.g, self.after_norm.b, self.eps, out_f32=enc,
                      out_split=out_halo if use_halo else None, nparts=self.nparts, halo=use_halo,
                      cu_rows=cu_rows if use_halo else None)
        if tp and self.tp_layers:
            x = enc
            for ly in self.tp_layers:
                x = self._layer(ly, x, cu_rows, max_t, ws)
            enc = torch.empty(r, self.d, dtype=torch.float32, device=self.device)
            ops.layernorm(x, self.tp_norm.g, self.tp_norm.b, self.eps, out_f32=enc)
        return enc


class PredictorPack:
    """CifPredictorV2 at inference (funasr/models/paraformer/cif_predictor.py:202-253) with
    tail_process_fn (:346-370) and cif_v1 (:668-735): conv k=3 as a 3-tap GEMM over the halo
    layout, alpha head, per-utterance integrate-and-fire scan, frame integration."""

    def __init__(self, sd, device, prefix: str = "predictor.", nparts: int = 3, threshold: float = 1.0,
                 smooth_factor: float = 1.0, noise_threshold: float = 0.0, tail_threshold: float = 0.45):
        self.device, self.nparts = torch.device(device), nparts
        w = sd[prefix + "cif_conv1d.weight"]
        if w.shape[-1] != 3:
            raise RuntimeError("funasr_b200: cif_conv1d kernel %d unsupported (l_order = r_order = 1 only)" % w.shape[-1])
        self.conv = _Linear(w, sd[prefix + "cif_conv1d.bias"], nparts, self.device)
        self.out_w = _dev(sd[prefix + "cif_output.weight"].reshape(-1), self.device)
        self.out_b = _dev(sd[prefix + "cif_output.bias"].reshape(-1), self.device)
        self.d = self.conv.k
        self.smooth, self.noise, self.tail = smooth_factor, noise_threshold, tail_threshold

    def halo_buffer(self, r: int, b: int) -> torch.Tensor:
        return torch.empty(r + b + 1, ops.split_width(self.d, self.nparts), dtype=torch.bfloat16, device=self.device)

    def run(self, enc: torch.Tensor, enc_halo: torch.Tensor, cu_rows: torch.Tensor, batch: int,
            want_fires: bool = False, embeds_out: Optional[torch.Tensor] = None):
        """enc fp32 [R, d]; enc_halo its bf16 split in halo layout.  Returns dict with alphas [R],
        token_num (host list), cu_tok (device i32), embeds [L, d], fires [R + B] or None."""
        r, d, dev = enc.shape[0], self.d, self.device
        m = r + batch - 1
        conv = torch.empty(r + batch, d, dtype=torch.float32, device=dev)
        ops.gemm(enc_halo, self.nparts, self.conv.w, self.nparts, m, d, d, bias=self.conv.bias, relu=True,
                 out_f32=conv, taps=3, a_rows=r + batch + 1)
        alphas = torch.empty(r, dtype=torch.float32, device=dev)
        ops.cif_alpha(conv, cu_rows, r, self.out_w, self.out_b, alphas, self.smooth, self.noise)
        token_num = torch.empty(batch, dtype=torch.int32, device=dev)
        fire_pos = torch.empty(r + batch, dtype=torch.int32, device=dev)
        fire_rem = torch.empty(r + batch, dtype=torch.float32, device=dev)
        fires = torch.empty(r + batch, dtype=torch.float32, device=dev) if want_fires else None
        ops.cif_scan(alphas, cu_rows, self.tail, token_num, fire_pos, fire_rem, fires)
        tok = token_num.cpu().tolist()  # the one D2H of the batch: [B] token counts
        cu_tok = tables.cu_from_lengths(tok, torch.int32, dev)
        total = int(sum(tok))
        if embeds_out is not None and embeds_out.shape[0] >= total:
            embeds = embeds_out[:total]
        else:
            embeds = torch.empty(total, d, dtype=torch.float32, device=dev)
        ops.cif_integrate(enc, alphas, cu_rows, cu_tok, total, self.tail, fire_pos, fire_rem, embeds)
        return dict(alphas=alphas, token_num=tok, token_num_dev=token_num, cu_tok=cu_tok, embeds=embeds, fires=fires)


class DecoderPack:
    """ParaformerSANMDecoder.forward (funasr/models/paraformer/decoder.py:359-411) over packed
    token rows, cross-attending to the packed encoder rows of the same utterance."""

    def __init__(self, sd, device, prefix: str = "decoder.", heads: int = 4, eps: float = 1e-12, nparts: int = 3):
        self.device, self.heads, self.eps, self.nparts = torch.device(device), heads, eps, nparts
        self.layers: List[_DecLayer] = []
        for group in ("decoders.", "decoders2.", "decoders3."):
            for i in range(_count(sd, prefix + group)):
                self.layers.append(_DecLayer(sd, prefix + group + "%d." % i, nparts, self.device))
        self.after_norm = _Norm(sd, prefix + "after_norm.", self.device)
        self.output = None  # the SeACo decoder has no output layer (`use_output_layer: false`)
        if prefix + "output_layer.weight" in sd:
            self.output = _Linear(sd[prefix + "output_layer.weight"], sd[prefix + "output_layer.bias"], nparts, self.device)
        self.d = self.layers[0].w1.k
        self.vocab = self.output.n if self.output is not None else 0
        self.ff = self.layers[0].w1.n
        # K/V projections of the memory for every cross-attention layer as ONE contraction
        src = [i for i, ly in enumerate(self.layers) if ly.has_src]
        self.src_index = {li: j for j, li in enumerate(src)}
        kvw = torch.cat([sd[prefix + self._name(li) + "src_attn.linear_k_v.weight"] for li in src], dim=0)
        kvb = torch.cat([sd[prefix + self._name(li) + "src_attn.linear_k_v.bias"] for li in src], dim=0)
        self.kv = _Linear(kvw, kvb, nparts, self.device)

    def _name(self, li: int) -> str:
        return "decoders.%d." % li  # cross-attention layers are the first `att_layer_num` of `decoders`

    def run(self, enc: torch.Tensor, cu_rows: torch.Tensor, embeds: torch.Tensor, cu_tok: torch.Tensor,
            max_l: int, enc_split: Optional[torch.Tensor] = None, want_logits: bool = False):
        """Returns (ids i32 [L], logp f32 [L], logits or None)."""
        _, xs = self.hidden(enc, cu_rows, embeds, cu_tok, max_l, enc_split=enc_split)
        return self.head(xs, want_logits=want_logits)

    def head(self, xs: torch.Tensor, want_logits: bool = False):
        """`output_layer` + log-softmax argmax over the split hidden states -> (ids, logp, logits or None)."""
        l, dev = xs.shape[0], self.device
        ids = torch.empty(l, dtype=torch.int32, device=dev)
        logp = torch.empty(l, dtype=torch.float32, device=dev)
        if not want_logits and self.nparts < 3 and FUSED_ARGMAX:
            # the [L, V] logits never exist: the projection's epilogue keeps (max, first argmax, sum of exponentials) per row
            ops.gemm_argmax(xs, self.nparts, self.output.w, self.nparts, l, self.vocab, self.d, ids, logp, bias=self.output.bias)
            return ids, logp, None
        logits = torch.empty(l, self.vocab, dtype=torch.float32, device=dev)
        ops.gemm(xs, self.nparts, self.output.w, self.nparts, l, self.vocab, self.d, bias=self.output.bias, out_f32=logits)
        ops.argmax_logprob(logits, self.vocab, ids, logp)
        return ids, logp, (logits if want_logits else None)

    def hidden(self, enc: torch.Tensor, cu_rows: torch.Tensor, embeds: torch.Tensor, cu_tok: torch.Tensor,
               max_l: int, enc_split: Optional[torch.Tensor] = None, want_f32: bool = False, query_layer: Optional[int] = None):
        """The decoder up to `after_norm` (decoder.py:395-404) -> (hidden fp32 [L, d] or None, its bf16 split).
        With `query_layer` the walk stops at that layer's cross-attention and returns its projected queries and keys in
        fp32, (q [L, d], k [R, d]) -- `get_attn_mat` (decoder.py:123-137), for the SeACo attention-score filter."""
        dev, np_, d, eps = self.device, self.nparts, self.d, self.eps
        r, l = enc.shape[0], embeds.shape[0]
        bf = torch.bfloat16
        if enc_split is None:
            enc_split = ops.split_rows(enc, np_)
        kv = torch.empty(r, ops.split_width(self.kv.n, np_), dtype=bf, device=dev)
        ops.gemm(enc_split, np_, self.kv.w, np_, r, self.kv.n, d, bias=self.kv.bias, out_split=kv, out_parts=np_)
        qs = torch.empty(l, ops.split_width(d, np_), dtype=bf, device=dev)
        xs = torch.empty(l, ops.split_width(d, np_), dtype=bf, device=dev)
        hs = torch.empty(l, ops.split_width(self.ff, np_), dtype=bf, device=dev)
        h = torch.empty(l, self.ff, dtype=torch.float32, device=dev)
        t = torch.empty(l, d, dtype=torch.float32, device=dev)
        n2 = torch.empty(l, d, dtype=torch.float32, device=dev)
        bufs = [torch.empty(l, d, dtype=torch.float32, device=dev) for _ in range(2)]
        # one work list for all cross-attentions (a stack-program recording needs none)
        plan = ops.attention_plan(cu_tok, cu_rows, l) if np_ < 3 and not ops.recording() else None
        x = embeds
        for li, ly in enumerate(self.layers):
            ops.layernorm(x, ly.norm1.g, ly.norm1.b, eps, out_split=xs, nparts=np_)
            ops.gemm(xs, np_, ly.w1.w, np_, l, self.ff, d, bias=ly.w1.bias, relu=True, out_f32=h)
            ops.layernorm(h, ly.ffn_norm.g, ly.ffn_norm.b, eps, out_split=hs, nparts=np_)
            ops.gemm(hs, np_, ly.w2.w, np_, l, d, self.ff, bias=ly.w2.bias, out_f32=t)
            if not ly.has_fsmn:  # decoders3: FFN only, no residual (decoder.py:102-108)
                x = t
                t = torch.empty(l, d, dtype=torch.float32, device=dev)
                continue
            ops.layernorm(t, ly.norm2.g, ly.norm2.b, eps, out_f32=n2)
            y = bufs[0] if x is not bufs[0] else bufs[1]
            ops.fsmn(n2, cu_tok, ly.fsmn, y, residual=x)  # x + conv(n2) + n2
            if ly.has_src:
                ops.layernorm(y, ly.norm3.g, ly.norm3.b, eps, out_split=xs, nparts=np_)
                j = self.src_index[li]
                if li == query_layer:
                    q32 = torch.empty(l, d, dtype=torch.float32, device=dev)
                    ops.gemm(xs, np_, ly.q.w, np_, l, d, d, bias=ly.q.bias, out_f32=q32)
                    k32 = torch.empty(r, d, dtype=torch.float32, device=dev)
                    ops.gemm(enc_split, np_, self.kv.w[j * 2 * d: j * 2 * d + d], np_, r, d, d, bias=self.kv.bias[j * 2 * d: j * 2 * d + d],
                             out_f32=k32)
                    return q32, k32
                ops.gemm(xs, np_, ly.q.w, np_, l, d, d, bias=ly.q.bias, out_split=qs, out_parts=np_)
                ops.attention(qs, 0, d, kv, j * 2 * d, kv, j * 2 * d + d, self.kv.n, np_, cu_tok, cu_rows, max_l,
                              self.heads, float(d // self.heads) ** -0.5, out_split=xs, plan=plan)
                xo = bufs[0] if y is not bufs[0] else bufs[1]
                ops.gemm(xs, np_, ly.out.w, np_, l, d, d, bias=ly.out.bias, res1=y, out_f32=xo)
                x = xo
            else:
                x = y
        hid = torch.empty(l, d, dtype=torch.float32, device=dev) if want_f32 else None
        out = torch.empty(l, ops.split_width(d, np_), dtype=bf, device=dev)  # `xs` is scratch of the layers above
        ops.layernorm(x, self.after_norm.g, self.after_norm.b, eps, out_f32=hid, out_split=out, nparts=np_)
        return hid, out


class _SmallPath:
    """Small requests (BASELINE configs[0]: one 10 s utterance = 167 encoder rows): the 50 encoder layers + after_norm and
    the 16 decoder layers + vocabulary projection run as TWO persistent cooperative kernels (`ops.StackProgram`) instead of
    ~570 launches.  Recorded once for capacities (rows, tokens, utterances); each request runs them with its own counts."""

    def __init__(self, eng: "ParaformerEngine"):
        dev, np_ = eng.device, eng.nparts
        enc_, dec_ = eng.encoder, eng.decoder
        rc, lc, bc = eng.STACK_ROWS, eng.STACK_TOKENS, eng.STACK_BATCH
        self.eng = eng
        d = enc_.d
        f32, bf, i32 = torch.float32, torch.bfloat16, torch.int32
        self.x = torch.zeros(rc, enc_.din, dtype=f32, device=dev)         # encoder input (the frontend writes it)
        self.cu_rows = torch.zeros(bc + 1, dtype=i32, device=dev)
        self.cu_tok = torch.zeros(bc + 1, dtype=i32, device=dev)
        self.enc = torch.zeros(rc, d, dtype=f32, device=dev)
        self.halo = torch.zeros(rc + bc + 1, ops.split_width(d, np_), dtype=bf, device=dev)
        self.split = torch.zeros(rc, ops.split_width(d, np_), dtype=bf, device=dev)
        self.embeds = torch.zeros(lc, d, dtype=f32, device=dev)           # CIF output (the predictor writes it)
        self.logits = torch.zeros(lc, dec_.vocab, dtype=f32, device=dev)
        self.enc_prog = ops.StackProgram([rc, bc])
        with self.enc_prog:
            enc_.record(self.x, self.cu_rows, self.enc, self.halo, self.split)
        self.dec_prog = ops.StackProgram([lc, rc, bc])
        with self.dec_prog:
            _, xs = dec_.hidden(self.enc, self.cu_rows, self.embeds, self.cu_tok, lc, enc_split=self.split)
            ops.gemm(xs, np_, dec_.output.w, np_, lc, dec_.vocab, dec_.d, bias=dec_.output.bias, out_f32=self.logits)

    def encode(self, cu_rows: torch.Tensor, r: int, b: int):
        """`self.x[:r]` holds the encoder input.  Returns (enc [r, d], halo split [r + b + 1, ...]) views of the buffers."""
        self.cu_rows[: b + 1].copy_(cu_rows)
        self.enc_prog.run([r, b])
        return self.enc[:r], self.halo[: r + b + 1]

    def decode(self, cu_tok: torch.Tensor, r: int, l: int, b: int):
        """`self.embeds[:l]` holds the CIF output.  Returns (ids i32 [l], logp f32 [l])."""
        self.cu_tok[: b + 1].copy_(cu_tok)
        self.dec_prog.run([l, r, b])
        dev = self.eng.device
        ids = torch.empty(l, dtype=torch.int32, device=dev)
        logp = torch.empty(l, dtype=torch.float32, device=dev)
        ops.argmax_logprob(self.logits[:l], self.eng.decoder.vocab, ids, logp)
        return ids, logp


class ParaformerEngine:
    """Paraformer.inference (funasr/models/paraformer/model.py:443-596) for a packed batch:
    host float32 waveforms in, token-id lists out.  One instance per process / GPU."""

    DROP_IDS = (0, 1, 2)  # blank, sos, eos (model.py:561-565)
    # bf16 parts per dense-contraction operand: fp32 = hi + mid (3 tcgen05 terms, "bf16x3"),
    # fp32x6 = hi + mid + lo (6 terms), bf16 = plain bf16 operands
    PRECISIONS = {"fp32": 2, "fp32x6": 3, "bf16": 1}

    def __init__(self, state_dict: Dict[str, torch.Tensor], device="cuda:0", precision: str = "fp32",
                 cmvn: Optional[Tuple[torch.Tensor, torch.Tensor]] = None, window: str = "hamming", heads: int = 4,
                 max_rows: int = DEFAULT_MAX_ROWS):
        if precision not in self.PRECISIONS:
            raise ValueError("precision must be one of %s" % sorted(self.PRECISIONS))
        if not torch.cuda.is_available():
            raise RuntimeError("funasr_b200: no CUDA device -- this path has no CPU fallback")
        if not 1 <= max_rows <= MAX_ROWS_LIMIT:
            raise ValueError("max_rows must be in [1, %d] (element offsets of the [rows, 2 x 16384] K / V projection stay below 2^31)"
                             % MAX_ROWS_LIMIT)
        self.device = torch.device(device)
        self.precision = precision
        self.nparts = self.PRECISIONS[precision]
        self.max_rows = max_rows
        self._copy_stream = None
        self._small = None
        self.last_token_num: List[int] = []
        self._drop_ids = np.array(sorted(self.DROP_IDS), dtype=np.int32)
        with torch.cuda.device(self.device):
            self.frontend = FrontendPack(self.device, cmvn=cmvn, window=window)
            self.encoder = EncoderPack(state_dict, self.device, heads=heads, nparts=self.nparts)
            self.predictor = PredictorPack(state_dict, self.device, nparts=self.nparts)
            self.decoder = DecoderPack(state_dict, self.device, heads=heads, nparts=self.nparts)

    # ---- small requests: encoder and decoder stretches as ONE persistent kernel each (csrc/stack_sm100.cu) -----------
    # capacities the two programs are recorded with (encoder rows, tokens, utterances of a request): up to one 30 s utterance
    # the path beats the per-op launches (tools/stack_capacity_sweep.py: 3.6 / 4.0 / 5.3 ms at 2 / 10 / 30 s against 6.5)
    STACK_ROWS = int(os.environ.get("FB200_STACK_ROWS", "512"))
    STACK_TOKENS = int(os.environ.get("FB200_STACK_TOKENS", "384"))
    STACK_BATCH = 8

    def small_path(self):
        """Buffers and programs of the request-stack path, recorded on first use (FB200_STACK=0 switches it off)."""
        if self._small is None:
            if os.environ.get("FB200_STACK", "1") == "0" or self.nparts > 2 or self.encoder.tp_layers:
                self._small = False
            else:
                self._small = _SmallPath(self)
        return self._small or None

    # ---- stage-level entry points (also used by the parity tests) ------------------------------
    def upload(self, waves: Sequence) -> Tuple[torch.Tensor, List[int]]:
        """Copy host waveforms (numpy arrays or torch CPU tensors, ideally pinned) into one packed
        device buffer -- one async H2D per utterance on the current stream."""
        lens = [int(w.shape[0]) for w in waves]
        dev = torch.empty(sum(lens), dtype=torch.float32, device=self.device)
        o = 0
        for w, n in zip(waves, lens):
            src = torch.from_numpy(w) if isinstance(w, np.ndarray) else w
            dev[o: o + n].copy_(src, non_blocking=True)
            o += n
        return dev, lens

    def forward_packed(self, wave: torch.Tensor, lens: Sequence[int], taps: Optional[dict] = None):
        """Device-resident packed waveforms -> (ids i32 [L] device, logp, token_num list)."""
        with torch.cuda.device(self.device):
            b = len(lens)
            small = None
            if taps is None and b <= self.STACK_BATCH and sum(self.frontend.plan(lens)[1]) <= self.STACK_ROWS:
                small = self.small_path()
            feats, cu_rows, rows = self.frontend.run(wave, lens, pe_scale=float(self.encoder.d) ** 0.5,
                                                     feats_out=small.x if small else None)
            r = feats.shape[0]
            if small:
                enc, halo = small.encode(cu_rows, r, b)
            else:
                halo = self.predictor.halo_buffer(r, b)
                enc = self.encoder.run(feats, cu_rows, max(rows), out_halo=halo)
            pred = self.predictor.run(enc, halo, cu_rows, b, want_fires=taps is not None,
                                      embeds_out=small.embeds if small else None)
            if taps is not None:
                taps.update(feats=feats, enc=enc, cu_rows=cu_rows, rows=rows, **pred)
            tok = pred["token_num"]
            if max(tok) < 1:
                empty = torch.empty(0, dtype=torch.int32, device=self.device)
                return empty, torch.empty(0, dtype=torch.float32, device=self.device), tok
            if small and sum(tok) <= self.STACK_TOKENS:
                ids, logp = small.decode(pred["cu_tok"], r, int(sum(tok)), b)
                return ids, logp, tok
            ids, logp, logits = self.decoder.run(enc, cu_rows, pred["embeds"], pred["cu_tok"], max(tok),
                                                 want_logits=taps is not None)
            if taps is not None:
                taps.update(logits=logits, ids=ids, logp=logp)
            return ids, logp, tok

    def plan_batches(self, lens: Sequence[int]) -> List[List[int]]:
        """Indices grouped longest-first under the packed-row budget (ref. analogue: the
        duration-sorted dynamic batching of funasr/auto/auto_model.py:423-459)."""
        _, rows = self.frontend.plan(lens)
        order = sorted(range(len(lens)), key=lambda i: -rows[i])
        batches, cur, cur_rows = [], [], 0
        for i in order:
            if rows[i] < 1:
                continue
            if cur and (cur_rows + rows[i] > self.max_rows or (self.max_utts and len(cur) >= self.max_utts)):
                batches.append(cur)
                cur, cur_rows = [], 0
            cur.append(i)
            cur_rows += rows[i]
        if cur:
            batches.append(cur)
        return self._ramp(batches, rows)

    def _ramp(self, batches: List[List[int]], rows: Sequence[int]) -> List[List[int]]:
        """With several batches the FIRST one is cut at a third of the row budget: nothing can overlap its upload (218 MB of
        waveforms for a full batch), so the kernels start after a third of that and the rest arrives under them."""
        if len(batches) < 2:
            return batches
        head, acc = [], 0
        for i in batches[0]:
            if head and acc + rows[i] > self.max_rows // 3:
                break
            head.append(i)
            acc += rows[i]
        if len(head) == len(batches[0]):
            return batches
        return [head, batches[0][len(head):]] + batches[1:]

    # utterances per batch (None: the row budget alone); engines with a time-stamp head set it to the LSTM launch's limit
    max_utts = None

    def uploaded_batches(self, waves: Sequence[np.ndarray]):
        """Yields (utterance indices, packed device waveforms, sample counts) per planned batch, inside this engine's
        device context; the host -> device copies of batch k + 1 run on their own stream under the kernels of batch k."""
        batches = self.plan_batches([int(w.shape[0]) for w in waves])
        if not batches:
            return
        with torch.cuda.device(self.device):
            compute = torch.cuda.current_stream()
            if self._copy_stream is None:
                self._copy_stream = torch.cuda.Stream(device=self.device)

            def start_upload(batch):
                with torch.cuda.stream(self._copy_stream):
                    wave, lens = self.upload([waves[i] for i in batch])
                    done = torch.cuda.Event()
                    done.record(self._copy_stream)
                return wave, lens, done

            nxt = start_upload(batches[0])
            for bi, batch in enumerate(batches):
                wave, lens, done = nxt
                compute.wait_event(done)
                wave.record_stream(compute)  # allocated under the copy stream, consumed on the compute stream
                if bi + 1 < len(batches):
                    nxt = start_upload(batches[bi + 1])
                yield batch, wave, lens

    RETURNS_ARRAYS = True  # recognize(as_arrays=True): int32 arrays instead of lists (schedule.recognize_sharded's payload)

    def recognize(self, waves: Sequence[np.ndarray], with_scores: bool = False, as_arrays: bool = False):
        """Token ids per utterance (sos/eos/blank dropped).  Utterances shorter than one frame or
        with no predicted token give [] (model.py:514-515)."""
        results: List[List[int]] = [[] for _ in waves]
        scores = [0.0 for _ in waves]
        self.last_token_num = [0 for _ in waves]  # predicted token counts of the last call (before id filtering)
        # Device -> host reads go to pinned buffers behind an event and become Python lists one batch LATE: the host-side
        # work of batch k (id filtering) runs while the GPU is busy with batch k + 1.
        def finish(item):
            batch, tok, ids_p, logp_p, done = item
            done.synchronize()
            ids_h = ids_p.numpy()
            logp_h = logp_p.numpy() if with_scores else None
            keep = ~np.isin(ids_h, self._drop_ids)
            o = 0
            for i, n in zip(batch, tok):
                self.last_token_num[i] = int(n)
                row = ids_h[o: o + n][keep[o: o + n]]
                results[i] = row if as_arrays else row.tolist()
                if with_scores:
                    scores[i] = float(logp_h[o: o + n].sum())
                o += n

        pending = None
        for batch, wave, lens in self.uploaded_batches(waves):
            ids, logp, tok = self.forward_packed(wave, lens)
            cur = (batch, tok, self.read_later(ids), self.read_later(logp) if with_scores else None, torch.cuda.Event())
            cur[-1].record()
            if pending is not None:
                finish(pending)
            pending = cur
        if pending is not None:
            finish(pending)
        return (results, scores) if with_scores else results

    @staticmethod
    def read_later(t: torch.Tensor) -> torch.Tensor:
        """Asynchronous device -> host copy into pinned memory on the current stream (valid after a later event / sync)."""
        h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
        h.copy_(t, non_blocking=True)
        return h


class SenseVoiceEngine:
    """SenseVoiceSmall.inference (funasr/models/sense_voice/model.py:836-906) for one GPU -- BASELINE
    config 3: host float32 waveforms in, CTC-greedy token-id lists out.  Same frontend and SAN-M kernels
    as the Paraformer path; differences: 4 query rows from `embed.weight` at the head of every utterance
    (:852-876), LayerNorm eps 1e-5 (:275-287), 20 extra `tp_encoders` + `tp_norm` (:553-585), then the
    CTC projection, per-frame argmax, unique_consecutive and blank removal (:884-906,
    funasr/models/ctc/ctc.py:173-197)."""

    LID = {"auto": 0, "zh": 3, "en": 4, "yue": 7, "ja": 11, "ko": 12, "nospeech": 13}  # model.py:642
    TEXTNORM = {"withitn": 14, "woitn": 15}                                             # model.py:644
    N_PROMPT = 4
    LOGIT_ROWS = 4096  # rows of the [rows, V] logits buffer per projection call (V = 25055: 410 MB)

    def __init__(self, state_dict: Dict[str, torch.Tensor], device="cuda:0", precision: str = "fp32",
                 cmvn: Optional[Tuple[torch.Tensor, torch.Tensor]] = None, window: str = "hamming", heads: int = 4,
                 max_rows: int = DEFAULT_MAX_ROWS, blank_id: int = 0):
        if precision not in ParaformerEngine.PRECISIONS:
            raise ValueError("precision must be one of %s" % sorted(ParaformerEngine.PRECISIONS))
        if not torch.cuda.is_available():
            raise RuntimeError("funasr_b200: no CUDA device -- this path has no CPU fallback")
        if not 0 <= blank_id < 32:
            raise ValueError("blank_id must be below 32")
        self.device = torch.device(device)
        self.precision = precision
        self.nparts = ParaformerEngine.PRECISIONS[precision]
        self.max_rows = max_rows
        self.blank_id = blank_id
        with torch.cuda.device(self.device):
            self.frontend = FrontendPack(self.device, cmvn=cmvn, window=window)
            self.encoder = EncoderPack(state_dict, self.device, heads=heads, eps=1e-5, nparts=self.nparts)
            self.ctc = _Linear(state_dict["ctc.ctc_lo.weight"], state_dict["ctc.ctc_lo.bias"], self.nparts, self.device)
            self.embed = _dev(state_dict["embed.weight"], self.device)
        self._prompts: Dict[Tuple[str, str], torch.Tensor] = {}
        self._ban_bias: Dict[Tuple[int, ...], torch.Tensor] = {}

    def _bias(self, ban_ids: Sequence[int]) -> torch.Tensor:
        """CTC bias with -inf at the banned ids (`ban_emo_unk`, model.py:886-887: the id can never win the argmax)."""
        key = tuple(sorted(int(i) for i in ban_ids))
        if not key:
            return self.ctc.bias
        if key not in self._ban_bias:
            b = self.ctc.bias.clone()
            b[torch.tensor(key, device=self.device)] = float("-inf")
            self._ban_bias[key] = b
        return self._ban_bias[key]

    def prompt(self, language: str = "auto", textnorm: str = "woitn") -> torch.Tensor:
        """[language, event, emotion, text-norm] rows of the query embedding, [4, 560] on the device."""
        key = (language, textnorm)
        if key not in self._prompts:
            ids = [self.LID.get(language, 0), 1, 2, self.TEXTNORM[textnorm]]
            self._prompts[key] = self.embed[torch.tensor(ids, device=self.device)].contiguous()
        return self._prompts[key]

    def forward_packed(self, wave: torch.Tensor, lens: Sequence[int], language: str = "auto", textnorm: str = "woitn",
                       taps: Optional[dict] = None, ban_ids: Sequence[int] = ()):
        """Device-resident packed waveforms -> (token ids i32 packed at cu_rows, counts i32 [B], cu_rows, rows)."""
        with torch.cuda.device(self.device):
            scale = float(self.encoder.d) ** 0.5
            feats, cu_rows, rows = self.frontend.run(wave, lens, pe_scale=scale, row_skip=self.N_PROMPT)
            ops.fill_head_rows(self.prompt(language, textnorm), cu_rows, feats, pe_inv=self.frontend.pe_inv, scale=scale)
            enc = self.encoder.run(feats, cu_rows, max(rows), tp=True)
            r, v = enc.shape[0], self.ctc.n
            enc_split = ops.split_rows(enc, self.nparts)
            frame_ids = torch.empty(r, dtype=torch.int32, device=self.device)
            bias = self._bias(ban_ids)
            if self.nparts < 3 and FUSED_ARGMAX:
                # CTC projection fused with the per-frame argmax: the [rows, 25055] logits are never written
                ops.gemm_argmax(enc_split, self.nparts, self.ctc.w, self.nparts, r, v, self.ctc.k, frame_ids, bias=bias)
            else:
                ld = (v + 3) // 4 * 4
                logits = torch.empty(min(r, self.LOGIT_ROWS), ld, dtype=torch.float32, device=self.device)
                for r0 in range(0, r, self.LOGIT_ROWS):
                    n = min(self.LOGIT_ROWS, r - r0)
                    ops.gemm(enc_split[r0: r0 + n], self.nparts, self.ctc.w, self.nparts, n, v, self.ctc.k, bias=bias,
                             out_f32=logits[:n])
                    ops.argmax_logprob(logits[:n], v, frame_ids[r0: r0 + n], rows=n)
            out_ids = torch.empty(r, dtype=torch.int32, device=self.device)
            out_len = torch.empty(len(lens), dtype=torch.int32, device=self.device)
            ops.ctc_collapse(frame_ids, cu_rows, 1 << self.blank_id, True, out_ids, out_len)
            if taps is not None:
                taps.update(feats=feats, enc=enc, frame_ids=frame_ids, cu_rows=cu_rows, rows=rows)
            return out_ids, out_len, cu_rows, rows

    def recognize(self, waves: Sequence[np.ndarray], language: str = "auto", textnorm: str = "woitn",
                  ban_ids: Sequence[int] = ()) -> List[List[int]]:
        """CTC-greedy token ids per utterance (blank dropped).  Utterances shorter than one frame give []."""
        results: List[List[int]] = [[] for _ in waves]

        def finish(item):
            batch, ids_p, cnt_p, cu_p, done = item
            done.synchronize()
            ids_h, cnt_h, cu_h = ids_p.numpy(), cnt_p.numpy(), cu_p.numpy()
            for j, i in enumerate(batch):
                results[i] = ids_h[cu_h[j]: cu_h[j] + cnt_h[j]].tolist()

        pending = None
        for batch, wave, lens in self.uploaded_batches(waves):  # uploads under the previous batch, results one batch late
            ids, counts, cu_rows, _ = self.forward_packed(wave, lens, language, textnorm, ban_ids=ban_ids)
            cur = (batch, self.read_later(ids), self.read_later(counts), self.read_later(cu_rows), torch.cuda.Event())
            cur[-1].record()
            if pending is not None:
                finish(pending)
            pending = cur
        if pending is not None:
            finish(pending)
        return results

    def plan_batches(self, lens: Sequence[int]) -> List[List[int]]:
        """Indices grouped longest-first under the packed-row budget (every utterance also carries its 4 prompt rows)."""
        _, rows = self.frontend.plan(lens)
        order = sorted(range(len(lens)), key=lambda i: -rows[i])
        batches, cur, cur_rows = [], [], 0
        for i in order:
            if rows[i] < 1:
                continue
            need = rows[i] + self.N_PROMPT
            if cur and cur_rows + need > self.max_rows:
                batches.append(cur)
                cur, cur_rows = [], 0
            cur.append(i)
            cur_rows += need
        if cur:
            batches.append(cur)
        return batches

    _copy_stream = None
    upload = ParaformerEngine.upload
    uploaded_batches = ParaformerEngine.uploaded_batches
    read_later = staticmethod(ParaformerEngine.read_later)
