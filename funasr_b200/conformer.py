"""Packed (var-len) execution of the reference's Conformer encoder + CTC greedy path on one B200 --
BASELINE config 4 (aishell recipe: examples/aishell/conformer/conf/conformer_12e_6d_2048_256.yaml).

Host-side orchestration only: every numeric step is a hand-written CUDA kernel reached through the C-ABI
(`funasr_b200.ops` -> libfunasr_b200.so).  Reference call stack:
  frontend     funasr/frontends/wav_frontend.py:118-158 (hamming, 80 mel, lfr 1/1, no CMVN)
  input layer  funasr/models/transformer/utils/subsampling.py:49-94 (two 3x3 stride-2 convolutions, Linear) and
               LegacyRelPositionalEncoding (funasr/models/transformer/embedding.py:215-252)
  blocks       funasr/models/conformer/encoder.py:183-283 (macaron FFN, rel-pos attention, conv module, FFN, norm)
  attention    funasr/models/transformer/attention.py:221-312 (legacy rel_shift)
  decoding     funasr/models/ctc/model.py:162-247 (ctc_lo, argmax, unique_consecutive, drop blank/sos/eos)

How the dense work reaches the tensor cores (fb200_gemm_ex):
  * conv 2 (256 -> 256 channels, 3x3 stride 2; 27 % of the encoder's FLOPs) = 9-tap contraction over four parity
    planes of the conv-1 output: tap (dt, df) is a fixed row offset, no im2col matrix exists;
  * the Linear over (channel, frequency) = 19-tap contraction with column offsets over the conv-2 output;
  * q+u | q+v | k | v = ONE projection (N = 1024, the position biases folded into the bias vector);
  * (q+v) . p^T per head against a table p = linear_pos(pe) built once at load, rel_shift is indexing only;
  * 0.5 * FFN: the factor is folded into w_2 (power of two: exact).
"""
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from .engine import FUSED_ARGMAX
from . import ops, tables
from .engine import FrontendPack, ParaformerEngine, _dev, _Linear, _Norm


def subsampled_lengths(frames: int):
    """(T1, T2) of Conv2dSubsampling for `frames` input frames (two kernel-3 stride-2 convolutions)."""
    t1 = (frames - 3) // 2 + 1 if frames >= 3 else 0
    t2 = (t1 - 3) // 2 + 1 if t1 >= 3 else 0
    return t1, t2


class _ConformerLayer:
    def __init__(self, sd, p, nparts, device, bn_eps=1e-5):
        def lin(name, scale=1.0, bias=True):
            b = sd[p + name + ".bias"] * scale if bias else None
            return _Linear(sd[p + name + ".weight"] * scale, b, nparts, device)

        self.norm_ff_macaron = _Norm(sd, p + "norm_ff_macaron.", device)
        self.norm_mha = _Norm(sd, p + "norm_mha.", device)
        self.norm_conv = _Norm(sd, p + "norm_conv.", device)
        self.norm_ff = _Norm(sd, p + "norm_ff.", device)
        self.norm_final = _Norm(sd, p + "norm_final.", device)
        # x + 0.5 * FFN(x): ff_scale (encoder.py:178,229) folded into w_2
        self.mac_w1, self.mac_w2 = lin("feed_forward_macaron.w_1"), lin("feed_forward_macaron.w_2", 0.5)
        self.ff_w1, self.ff_w2 = lin("feed_forward.w_1"), lin("feed_forward.w_2", 0.5)
        a = p + "self_attn."
        wq, bq = sd[a + "linear_q.weight"], sd[a + "linear_q.bias"]
        u, v = sd[a + "pos_bias_u"].reshape(-1), sd[a + "pos_bias_v"].reshape(-1)
        self.heads, self.dk = sd[a + "pos_bias_u"].shape
        w = torch.cat([wq, wq, sd[a + "linear_k.weight"], sd[a + "linear_v.weight"]], dim=0)
        b = torch.cat([bq + u, bq + v, sd[a + "linear_k.bias"], sd[a + "linear_v.bias"]], dim=0)
        self.quqvkv = _Linear(w, b, nparts, device)
        self.out = lin("self_attn.linear_out")
        self.pos_w = _Linear(sd[a + "linear_pos.weight"], None, nparts, device)
        self.pos: Optional[torch.Tensor] = None  # linear_pos(pe)[:, head] as split operands, heads stacked along the rows
        c = p + "conv_module."
        w1 = sd[c + "pointwise_conv1.weight"]
        self.pw1 = _Linear(w1.reshape(w1.shape[0], w1.shape[1]), sd[c + "pointwise_conv1.bias"], nparts, device)
        w2 = sd[c + "pointwise_conv2.weight"]
        self.pw2 = _Linear(w2.reshape(w2.shape[0], w2.shape[1]), sd[c + "pointwise_conv2.bias"], nparts, device)
        dw = sd[c + "depthwise_conv.weight"]
        self.dw = _dev(dw.reshape(dw.shape[0], dw.shape[-1]), device)
        self.dw_bias = _dev(sd[c + "depthwise_conv.bias"], device)
        # BatchNorm1d in eval mode (running statistics) as one multiply-add per channel
        scale = sd[c + "norm.weight"].float() / torch.sqrt(sd[c + "norm.running_var"].float() + bn_eps)
        self.bn_scale = _dev(scale, device)
        self.bn_shift = _dev(sd[c + "norm.bias"].float() - sd[c + "norm.running_mean"].float() * scale, device)


class ConformerEncoderPack:
    """ConformerEncoder.forward (funasr/models/conformer/encoder.py:530-607) over packed fbank frames."""

    MAX_POS = 5000  # rows of the reference's position table (embedding.py:48); longer inputs are not supported

    def __init__(self, sd: Dict[str, torch.Tensor], device, prefix: str = "encoder.", nparts: int = 2,
                 eps: float = 1e-12, max_len: int = 2048):
        self.device, self.nparts, self.eps = torch.device(device), nparts, eps
        dev = self.device
        e = prefix + "embed."
        w1 = sd[e + "conv.0.weight"]
        self.c = w1.shape[0]
        self.conv1_w = _dev(w1.reshape(self.c, 9), dev)
        self.conv1_b = _dev(sd[e + "conv.0.bias"], dev)
        w2 = sd[e + "conv.2.weight"]  # [out, in, 3, 3] -> taps (dt, df) of [out, in]
        self.conv2 = _Linear(w2.permute(2, 3, 0, 1).reshape(9 * self.c, self.c), None, nparts, dev)
        self.conv2.taps, self.conv2.n = 9, self.c
        self.conv2_b = _dev(sd[e + "conv.2.bias"], dev)
        wo = sd[e + "out.0.weight"]
        self.d = wo.shape[0]
        self.f2 = wo.shape[1] // self.c
        xscale = float(self.d) ** 0.5  # LegacyRelPositionalEncoding.forward: x * xscale (16 for d = 256: exact)
        wo = wo.reshape(self.d, self.c, self.f2).permute(2, 0, 1).reshape(self.f2 * self.d, self.c) * xscale
        self.out = _Linear(wo, None, nparts, dev)
        self.out.taps, self.out.n = self.f2, self.d
        self.out_b = _dev(sd[e + "out.0.bias"] * xscale, dev)
        self.layers: List[_ConformerLayer] = []
        i = 0
        while (prefix + "encoders.%d.norm_ff.weight" % i) in sd:
            self.layers.append(_ConformerLayer(sd, prefix + "encoders.%d." % i, nparts, dev))
            i += 1
        self.after_norm = _Norm(sd, prefix + "after_norm.", dev)
        self.heads, self.dk = self.layers[0].heads, self.layers[0].dk
        self.ff = self.layers[0].ff_w1.n
        self.n_mels = None
        self._pos_rows = 0
        self.ensure_positions(max_len)

    def ensure_positions(self, rows: int) -> None:
        """p = linear_pos(pos_emb) for the first `rows` table rows, per layer and head, as split operands
        (attention.py:294-295; the table does not depend on the utterance)."""
        if rows <= self._pos_rows:
            return
        rows = min(max((rows + 255) // 256 * 256, 1024), self.MAX_POS)
        np_, d, dk = self.nparts, self.d, self.dk
        pe = ops.split_rows(tables.legacy_rel_position_table(rows, d, self.MAX_POS).to(self.device), np_)
        p = torch.empty(rows, d, dtype=torch.float32, device=self.device)
        for ly in self.layers:
            ops.gemm(pe, np_, ly.pos_w.w, np_, rows, d, d, out_f32=p)
            # heads stacked along the rows: ONE grouped launch computes every head's (q+v).p^T (ops.gemm_heads)
            ly.pos = torch.cat([ops.split_rows(p[:, h * dk:(h + 1) * dk], np_) for h in range(self.heads)], dim=0)
        self._pos_rows = rows

    def plan(self, frames: Sequence[int]):
        t2 = [subsampled_lengths(int(f))[1] for f in frames]
        return t2

    def embed(self, feats: torch.Tensor, cu_frames: torch.Tensor, frames: Sequence[int]):
        """Conv2dSubsampling + x * sqrt(d) over packed frames [sum F, n_mels] -> (x fp32 [R, d], cu_rows, rows)."""
        dev, np_, c, d = self.device, self.nparts, self.c, self.d
        rows = self.plan(frames)
        if min(rows) < 1:  # check_short_utt (subsampling.py:36-46): the reference raises TooShortUttError below 7 frames
            raise RuntimeError("funasr_b200: an utterance has fewer than 7 frames, too short for the conv2d input layer")
        n_mels = feats.shape[1]
        f1 = (n_mels - 3) // 2 + 1
        fp = (f1 + 1) // 2
        f2 = (f1 - 3) // 2 + 1
        if f2 != self.f2:
            raise RuntimeError("funasr_b200: embed.out expects %d frequency bins after subsampling, input gives %d" % (self.f2, f2))
        seg = [r + 1 for r in rows]
        total_seg = int(sum(seg))
        cu_seg = tables.cu_from_lengths(seg, torch.int32, dev)
        cu_rows = tables.cu_from_lengths(rows, torch.int32, dev)
        pr = total_seg * fp
        planes = torch.empty(4 * pr, np_ * c, dtype=torch.bfloat16, device=dev)
        ops.conv2d_sub1(feats, cu_frames, cu_seg, total_seg, self.conv1_w, self.conv1_b, planes, np_)
        tap_row = [((dt & 1) * 2 + (df & 1)) * pr + (dt >> 1) * fp + (df >> 1) for dt in range(3) for df in range(3)]
        y2 = torch.empty(pr, np_ * c, dtype=torch.bfloat16, device=dev)
        ops.gemm_ex(planes, np_, self.conv2.w, np_, pr, c, c, taps=9, tap_row=tap_row, bias=self.conv2_b, act="relu",
                    out_split=y2, out_parts=np_)
        xs = torch.empty(total_seg, d, dtype=torch.float32, device=dev)
        ops.gemm_ex(y2, np_, self.out.w, np_, total_seg, d, c, lda=fp * np_ * c, a_part_stride=c, taps=f2,
                    tap_row=[0] * f2, tap_col=[f * np_ * c for f in range(f2)], bias=self.out_b, out_f32=xs,
                    a_rows=total_seg)
        x = torch.empty(int(sum(rows)), d, dtype=torch.float32, device=dev)
        ops.gather_rows(xs, cu_seg, x, cu_rows)
        return x, cu_rows, rows

    def workspace(self, r: int, max_t: int):
        dev, np_, d = self.device, self.nparts, self.d
        bf, f32 = torch.bfloat16, torch.float32
        bd_pitch = (max_t + 255) // 256 * 256  # columns per head of the grouped (q+v).p^T product
        return {
            "xs": torch.empty(r, ops.split_width(d, np_), dtype=bf, device=dev),
            "h": torch.empty(r, ops.split_width(self.ff, np_), dtype=bf, device=dev),
            "qkv": torch.empty(r, ops.split_width(4 * d, np_), dtype=bf, device=dev),
            "bd": torch.empty(r, self.heads, bd_pitch, dtype=f32, device=dev),
            "g": torch.empty(r, 2 * d, dtype=f32, device=dev),
            "xa": torch.empty(r, d, dtype=f32, device=dev),
            "xb": torch.empty(r, d, dtype=f32, device=dev),
        }

    def _ffn(self, x, norm, w1, w2, out, ws):
        r, np_, d = x.shape[0], self.nparts, self.d
        ops.layernorm(x, norm.g, norm.b, self.eps, out_split=ws["xs"], nparts=np_)
        ops.gemm_ex(ws["xs"], np_, w1.w, np_, r, self.ff, d, bias=w1.bias, act="swish", out_split=ws["h"], out_parts=np_)
        ops.gemm(ws["h"], np_, w2.w, np_, r, d, self.ff, bias=w2.bias, res1=x, out_f32=out)
        return out

    def _layer(self, ly: _ConformerLayer, x, cu_rows, max_t, ws, out):
        r, np_, d, dk, heads = x.shape[0], self.nparts, self.d, self.dk, self.heads
        xa, xb = ws["xa"], ws["xb"]
        x1 = self._ffn(x, ly.norm_ff_macaron, ly.mac_w1, ly.mac_w2, xa, ws)
        # rel-pos self-attention
        ops.layernorm(x1, ly.norm_mha.g, ly.norm_mha.b, self.eps, out_split=ws["xs"], nparts=np_)
        qkv = ws["qkv"]
        ops.gemm(ws["xs"], np_, ly.quqvkv.w, np_, r, 4 * d, d, bias=ly.quqvkv.bias, out_split=qkv, out_parts=np_)
        bd = ws["bd"]  # bd[r, h, c] = (q_r + v)_h . p_c for c < max_t, every head in one launch
        ops.gemm_heads(qkv, np_, ly.pos, self._pos_rows, r, max_t, heads, dk, bd.view(r, -1), lda=np_ * 4 * d,
                       a_part_stride=4 * d, a_col0=d, a_col_step=dk)
        ops.attention_relpos(qkv, 0, 4 * d, qkv, 2 * d, qkv, 3 * d, 4 * d, np_, bd.permute(1, 0, 2), cu_rows, max_t, heads,
                             float(dk) ** -0.5, out_split=ws["xs"], plan=ws["plan"])
        ops.gemm(ws["xs"], np_, ly.out.w, np_, r, d, d, bias=ly.out.bias, res1=x1, out_f32=xb)
        # convolution module
        ops.layernorm(xb, ly.norm_conv.g, ly.norm_conv.b, self.eps, out_split=ws["xs"], nparts=np_)
        ops.gemm(ws["xs"], np_, ly.pw1.w, np_, r, 2 * d, d, bias=ly.pw1.bias, out_f32=ws["g"])
        ops.glu_dwconv(ws["g"], cu_rows, max_t, d, ly.dw, ly.dw_bias, ly.bn_scale, ly.bn_shift, ws["xs"], np_)
        ops.gemm(ws["xs"], np_, ly.pw2.w, np_, r, d, d, bias=ly.pw2.bias, res1=xb, out_f32=xa)
        x4 = self._ffn(xa, ly.norm_ff, ly.ff_w1, ly.ff_w2, xb, ws)
        ops.layernorm(x4, ly.norm_final.g, ly.norm_final.b, self.eps, out_f32=out)
        return out

    def run(self, x: torch.Tensor, cu_rows: torch.Tensor, max_t: int, taps: Optional[list] = None,
            out_split: Optional[torch.Tensor] = None) -> torch.Tensor:
        """x fp32 [R, d] from `embed`.  Returns after_norm(x) fp32 [R, d] (and its split when asked)."""
        if max_t > self.MAX_POS:
            raise RuntimeError("funasr_b200: %d encoder frames exceed the reference's %d-row position table" % (max_t, self.MAX_POS))
        self.ensure_positions(max_t)
        r = x.shape[0]
        ws = self.workspace(r, max_t)
        ws["plan"] = ops.attention_plan(cu_rows, cu_rows, r)  # 128-row query blocks, shared by every layer
        bufs = [torch.empty_like(x), torch.empty_like(x)]
        for i, ly in enumerate(self.layers):
            x = self._layer(ly, x, cu_rows, max_t, ws, bufs[i & 1])
            if taps is not None:
                taps.append(x.clone())
        enc = torch.empty(r, self.d, dtype=torch.float32, device=self.device)
        ops.layernorm(x, self.after_norm.g, self.after_norm.b, self.eps, out_f32=enc, out_split=out_split, nparts=self.nparts)
        return enc


class ConformerCTCEngine:
    """`model_classes["CTC"].inference` (funasr/models/ctc/model.py:162-247) over a ConformerEncoder for one GPU:
    host float32 waveforms in, CTC-greedy token-id lists out (blank / sos / eos dropped)."""

    DROP_MASK = 0b111  # ids 0, 1, 2 (model.py:243-247)

    def __init__(self, state_dict: Dict[str, torch.Tensor], device="cuda:0", precision: str = "fp32",
                 window: str = "hamming", max_rows: int = 56832):
        if precision not in ParaformerEngine.PRECISIONS:
            raise ValueError("precision must be one of %s" % sorted(ParaformerEngine.PRECISIONS))
        if not torch.cuda.is_available():
            raise RuntimeError("funasr_b200: no CUDA device -- this path has no CPU fallback")
        self.device = torch.device(device)
        self.precision = precision
        self.nparts = ParaformerEngine.PRECISIONS[precision]
        if self.nparts == 3:
            raise ValueError("the Conformer path runs in fp32 (hi + mid) or bf16 precision")
        self.max_rows = max_rows
        with torch.cuda.device(self.device):
            self.frontend = FrontendPack(self.device, cmvn=None, window=window, lfr_m=1, lfr_n=1)
            self.encoder = ConformerEncoderPack(state_dict, self.device, nparts=self.nparts)
            self.ctc = _Linear(state_dict["ctc.ctc_lo.weight"], state_dict["ctc.ctc_lo.bias"], self.nparts, self.device)

    def encoder_rows(self, num_samples: int) -> int:
        return subsampled_lengths(tables.num_frames(int(num_samples), self.frontend.frame_len, self.frontend.frame_shift))[1]

    def forward_packed(self, wave: torch.Tensor, lens: Sequence[int], taps: Optional[dict] = None):
        """Device-resident packed waveforms -> (token ids i32 packed at cu_rows, counts i32 [B], cu_rows, rows)."""
        with torch.cuda.device(self.device):
            dev, np_ = self.device, self.nparts
            frames = [tables.num_frames(int(n), self.frontend.frame_len, self.frontend.frame_shift) for n in lens]
            cu_samples = tables.cu_from_lengths(lens, torch.int64, dev)
            cu_frames = tables.cu_from_lengths(frames, torch.int32, dev)
            fe = self.frontend
            fb = ops.fbank(wave, cu_samples, cu_frames, int(sum(frames)), fe.window, fe.mel, fe.frame_shift, fe.nfft,
                           fe.upscale, 0.97)
            x, cu_rows, rows = self.encoder.embed(fb, cu_frames, frames)
            r, v = x.shape[0], self.ctc.n
            layer_taps = [] if taps is not None else None
            if taps is not None:
                taps.update(embed=x.clone())
            enc_split = torch.empty(r, ops.split_width(self.encoder.d, np_), dtype=torch.bfloat16, device=dev)
            enc = self.encoder.run(x, cu_rows, max(rows), taps=layer_taps, out_split=enc_split)
            frame_ids = torch.empty(r, dtype=torch.int32, device=dev)
            if np_ < 3 and FUSED_ARGMAX:
                ops.gemm_argmax(enc_split, np_, self.ctc.w, np_, r, v, self.ctc.k, frame_ids, bias=self.ctc.bias)
            else:
                ld = (v + 3) // 4 * 4
                logits = torch.empty(r, ld, dtype=torch.float32, device=dev)
                ops.gemm(enc_split, np_, self.ctc.w, np_, r, v, self.ctc.k, bias=self.ctc.bias, out_f32=logits)
                ops.argmax_logprob(logits, v, frame_ids, rows=r)
            out_ids = torch.empty(r, dtype=torch.int32, device=dev)
            out_len = torch.empty(len(lens), dtype=torch.int32, device=dev)
            ops.ctc_collapse(frame_ids, cu_rows, self.DROP_MASK, True, out_ids, out_len)
            if taps is not None:
                taps.update(fbank=fb, enc=enc, layers=layer_taps, frame_ids=frame_ids, cu_rows=cu_rows, rows=rows)
            return out_ids, out_len, cu_rows, rows

    def plan_batches(self, lens: Sequence[int]) -> List[List[int]]:
        rows = [self.encoder_rows(n) for n in lens]
        order = sorted(range(len(lens)), key=lambda i: -rows[i])
        batches, cur, cur_rows = [], [], 0
        for i in order:
            if rows[i] < 1:
                continue
            if cur and cur_rows + rows[i] > self.max_rows:
                batches.append(cur)
                cur, cur_rows = [], 0
            cur.append(i)
            cur_rows += rows[i]
        if cur:
            batches.append(cur)
        return batches

    def upload(self, waves: Sequence):
        lens = [int(w.shape[0]) for w in waves]
        dev = torch.empty(sum(lens), dtype=torch.float32, device=self.device)
        o = 0
        for w, n in zip(waves, lens):
            src = torch.from_numpy(w) if isinstance(w, np.ndarray) else w
            dev[o: o + n].copy_(src, non_blocking=True)
            o += n
        return dev, lens

    def recognize(self, waves: Sequence[np.ndarray]) -> List[List[int]]:
        """CTC-greedy token ids per utterance.  Utterances too short for the input layer (fewer than 7 frames: the
        reference raises TooShortUttError and AutoModel returns no result for them) give []."""
        results: List[List[int]] = [[] for _ in waves]

        # uploads of batch k + 1 on a copy stream under the kernels of batch k; results read into pinned memory behind an
        # event and unpacked one batch late (engine.ParaformerEngine.uploaded_batches / read_later)
        def finish(item):
            batch, ids_p, cnt_p, cu_p, done = item
            done.synchronize()
            ids_h, cnt_h, cu_h = ids_p.numpy(), cnt_p.numpy(), cu_p.numpy()
            for j, i in enumerate(batch):
                results[i] = ids_h[cu_h[j]: cu_h[j] + cnt_h[j]].tolist()

        pending = None
        for batch, wave, lens in self.uploaded_batches(waves):
            ids, counts, cu_rows, _ = self.forward_packed(wave, lens)
            cur = (batch, self.read_later(ids), self.read_later(counts), self.read_later(cu_rows), torch.cuda.Event())
            cur[-1].record()
            if pending is not None:
                finish(pending)
            pending = cur
        if pending is not None:
            finish(pending)
        return results

    _copy_stream = None
    uploaded_batches = ParaformerEngine.uploaded_batches
    read_later = staticmethod(ParaformerEngine.read_later)
