"""BiCifParaformer at inference (funasr/models/bicif_paraformer/model.py:222-369) on one B200: the Paraformer path of
`engine.ParaformerEngine` plus the time-stamp head of CifPredictorV3 (funasr/models/bicif_paraformer/cif_predictor.py:248-293).

Per packed batch, after the decoder's greedy pick:
  upsample_cnn (ConvTranspose1d k = s = 3)   one GEMM [R, 512] x [1536, 512]^T, read back as [3R, 512] rows
  blstm input projections, both directions   three GEMMs (one per sub-frame) into xg fp32 [3R, 2 x 2048]
  blstm recurrence + cif_output2 dot         fb200_lstm_bidir (cooperative persistent kernel)
  alphas2, length scaling, fires, rescan     fb200_us_fires (one warp per utterance)
  [begin ms, end ms] per token               timestamp.stamps_from_fires on the host (float64, as the reference)

CifPredictorV3's first head is the recurrence of CifPredictorV2 written as a frame loop (`cif`, cif_predictor.py:27-69, fp32
sequential) instead of the prefix-sum form (`cif_v1`); the device keeps the prefix-sum kernels of the Paraformer path (fp64
prefix sums) -- the acoustic embeddings agree to fp32 rounding, checked against the reference's V3 outputs in
tests/test_gpu_bicif.py.
"""
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import ops, timestamp
from .engine import ParaformerEngine, _Linear, _dev


class TimestampPack:
    def __init__(self, sd: Dict[str, torch.Tensor], device, prefix: str = "predictor.", nparts: int = 2,
                 smooth_factor2: float = 0.25, noise_threshold2: float = 0.01, threshold: float = 1.0):
        self.device, self.nparts = torch.device(device), nparts
        w = sd[prefix + "upsample_cnn.weight"]  # ConvTranspose1d [in, out, k], stride = k (cif_predictor.py:139-141)
        d, dout, k = w.shape
        self.d, self.up = d, k
        # out[3t + j, o] = sum_c enc[t, c] w[c, o, j] + b[o]: a GEMM whose weight rows are (j, o)
        self.up_w = _Linear(w.permute(2, 1, 0).reshape(k * dout, d).contiguous(), sd[prefix + "upsample_cnn.bias"].repeat(k),
                            nparts, self.device)
        b = prefix + "blstm."
        if sd[b + "weight_hh_l0"].shape != (4 * dout, dout) or dout != 512:
            raise RuntimeError("funasr_b200: blstm hidden size %d unsupported (512 only)" % sd[b + "weight_hh_l0"].shape[1])
        self.h = dout
        w_ih = torch.cat([sd[b + "weight_ih_l0"], sd[b + "weight_ih_l0_reverse"]], dim=0)
        b_ih = torch.cat([sd[b + "bias_ih_l0"] + sd[b + "bias_hh_l0"], sd[b + "bias_ih_l0_reverse"] + sd[b + "bias_hh_l0_reverse"]])
        self.proj = _Linear(w_ih, b_ih, nparts, self.device)
        self.w_hh = _dev(torch.stack([sd[b + "weight_hh_l0"], sd[b + "weight_hh_l0_reverse"]]).contiguous(), self.device)
        self.w_out = _dev(sd[prefix + "cif_output2.weight"].reshape(2, dout).contiguous(), self.device)
        self.b_out = float(sd[prefix + "cif_output2.bias"].reshape(-1)[0])
        self.smooth, self.noise, self.threshold = smooth_factor2, noise_threshold2, threshold
        self.parts = ops.lstm_parts(dout)

    def run(self, enc: torch.Tensor, cu_rows: torch.Tensor, max_rows: int, token_num: torch.Tensor, ids: torch.Tensor,
            cu_tok: torch.Tensor, drop_ids: Sequence[int], want_taps: bool = False):
        """enc fp32 [R, d] packed; token_num i32 [B] (device).  Returns dict(us_alphas [3R], fire_idx [3R] i32,
        fire_count [B], num_chars [B], and with want_taps us_peaks [3R] and lstm_out [3R, 2H])."""
        dev, np_, d, k, h = self.device, self.nparts, self.d, self.up, self.h
        r = enc.shape[0]
        xs = ops.split_rows(enc, np_)
        up = torch.empty(r, ops.split_width(k * h, np_), dtype=torch.bfloat16, device=dev)
        ops.gemm(xs, np_, self.up_w.w, np_, r, k * h, d, bias=self.up_w.bias, out_split=up, out_parts=np_)
        xg = torch.empty(r * k, 8 * h, dtype=torch.float32, device=dev)
        xg_rows = xg.view(r, k * 8 * h)
        for j in range(k):  # sub-frame j of every encoder frame: columns [j h, (j + 1) h) of the upsampling GEMM's rows
            ops.gemm_ex(up, np_, self.proj.w, np_, r, 8 * h, h, lda=np_ * k * h, a_part_stride=k * h, a_col0=j * h,
                        bias=self.proj.bias, out_f32=xg_rows[:, j * 8 * h:(j + 1) * 8 * h])
        part = torch.empty(self.parts, r * k, dtype=torch.float32, device=dev)
        lstm_out = torch.empty(r * k, 2 * h, dtype=torch.float32, device=dev) if want_taps else None
        ops.lstm_bidir(xg, cu_rows, k, k * max_rows, self.w_hh, self.w_out, part, out_h=lstm_out, hidden=h)
        b = cu_rows.numel() - 1
        us_alphas = torch.empty(r * k, dtype=torch.float32, device=dev)
        us_peaks = torch.empty(r * k, dtype=torch.float32, device=dev) if want_taps else None
        fire_idx = torch.empty(r * k, dtype=torch.int32, device=dev)
        fire_count = torch.empty(b, dtype=torch.int32, device=dev)
        num_chars = torch.empty(b, dtype=torch.int32, device=dev)
        ops.us_fires(part, self.b_out, self.smooth, self.noise, self.threshold - 1e-4, cu_rows, k, token_num, ids, cu_tok,
                     drop_ids, us_alphas, fire_idx, fire_count, num_chars, us_peaks=us_peaks)
        return dict(us_alphas=us_alphas, us_peaks=us_peaks, fire_idx=fire_idx, fire_count=fire_count, num_chars=num_chars,
                    lstm_out=lstm_out)


class BiCifParaformerEngine(ParaformerEngine):
    """BiCifParaformer.inference for packed batches: token ids and [begin ms, end ms] per token."""

    RETURNS_ARRAYS = False  # recognize() here returns (id lists, time-stamp lists)
    max_utts = 1024  # fb200_lstm_bidir takes at most 1024 utterances per launch (many short VAD segments fit the row budget)

    def __init__(self, state_dict, device="cuda:0", precision: str = "fp32", **kw):
        super().__init__(state_dict, device, precision, **kw)
        with torch.cuda.device(self.device):
            self.stamper = TimestampPack(state_dict, self.device, nparts=self.nparts)

    def decode(self, enc, cu_rows, embeds, cu_tok, max_l, taps=None):
        """Decoder + greedy pick -> (ids, logp, logits or None); `seaco.SeacoParaformerEngine` adds the hot-word branch."""
        return self.decoder.run(enc, cu_rows, embeds, cu_tok, max_l, want_logits=taps is not None)

    def forward_packed_ts(self, wave: torch.Tensor, lens: Sequence[int], taps: Optional[dict] = None):
        """-> (ids i32 [L] device, logp, token_num list, rows list, time-stamp dict or None)."""
        inner = {} if taps is None else taps
        with torch.cuda.device(self.device):
            b = len(lens)
            small = None  # small requests: encoder / decoder stretches as persistent kernels (engine._SmallPath)
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
            inner.update(feats=feats, enc=enc, cu_rows=cu_rows, rows=rows, **pred)
            tok = pred["token_num"]
            if max(tok) < 1:
                empty = torch.empty(0, dtype=torch.int32, device=self.device)
                return empty, torch.empty(0, dtype=torch.float32, device=self.device), tok, rows, None
            if small and sum(tok) <= self.STACK_TOKENS and type(self).decode is BiCifParaformerEngine.decode:
                ids, logp = small.decode(pred["cu_tok"], r, int(sum(tok)), b)
                logits = None
            else:
                ids, logp, logits = self.decode(enc, cu_rows, pred["embeds"], pred["cu_tok"], max(tok), taps)
            ts = self.stamper.run(enc, cu_rows, max(rows), pred["token_num_dev"], ids, pred["cu_tok"], self.DROP_IDS,
                                  want_taps=taps is not None)
            inner.update(logits=logits, ids=ids, logp=logp, **ts)
            return ids, logp, tok, rows, ts

    def recognize(self, waves: Sequence[np.ndarray], with_scores: bool = False, begin_time: float = 0.0):
        """-> (token-id lists, time-stamp lists [[begin ms, end ms], ...]) per utterance (sos/eos/blank dropped; one stamp
        per kept token).  `begin_time` (ms) is the reference's `begin_time` kwarg (vad offset, model.py:349)."""
        results: List[List[int]] = [[] for _ in waves]
        stamps: List[List[List[int]]] = [[] for _ in waves]
        scores = [0.0 for _ in waves]
        self.last_token_num = [0 for _ in waves]
        # Device -> host reads go to pinned buffers behind an event and are turned into Python lists one batch LATE: the
        # host-side work of batch k (id filtering, time stamps) runs while the GPU is busy with batch k + 1.
        def stage(batch, ids, logp, tok, rows, ts):
            pinned = self.read_later
            item = dict(batch=batch, tok=tok, rows=rows, ids=pinned(ids), logp=pinned(logp) if with_scores else None, ts=None)
            if ts is not None:
                item["ts"] = {k: pinned(ts[k]) for k in ("fire_idx", "fire_count", "num_chars")}
            item["done"] = torch.cuda.Event()
            item["done"].record()
            return item

        def finish(item):
            item["done"].synchronize()
            ids_h = item["ids"].numpy()
            keep = ~np.isin(ids_h, self._drop_ids)
            ts = item["ts"]
            if ts is not None:
                fire_idx, fire_count, num_chars = ts["fire_idx"].numpy(), ts["fire_count"].numpy(), ts["num_chars"].numpy()
            o = r0 = 0
            for j, (i, n) in enumerate(zip(item["batch"], item["tok"])):
                self.last_token_num[i] = int(n)
                results[i] = ids_h[o: o + n][keep[o: o + n]].tolist()
                if with_scores:
                    scores[i] = float(item["logp"].numpy()[o: o + n].sum())
                if ts is not None and n > 0:
                    f0 = 3 * r0
                    stamps[i] = timestamp.stamps_from_fires(fire_idx[f0: f0 + int(fire_count[j])], 3 * item["rows"][j],
                                                            int(num_chars[j]), vad_offset=begin_time)
                o += n
                r0 += item["rows"][j]

        pending = None
        for batch, wave, lens in self.uploaded_batches(waves):
            ids, logp, tok, rows, ts = self.forward_packed_ts(wave, lens)
            cur = stage(batch, ids, logp, tok, rows, ts)
            if pending is not None:
                finish(pending)
            pending = cur
        if pending is not None:
            finish(pending)
        return (results, stamps, scores) if with_scores else (results, stamps)
