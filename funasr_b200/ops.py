"""Thin torch-facing wrappers over the C-ABI (include/funasr_b200.h).

torch is used for device memory and the current stream only; every function here enqueues
hand-written CUDA kernels from libfunasr_b200.so and raises if the library is unavailable.
Naming: `cu_*` are int32 prefix-sum tensors on the device ([batch + 1]).
"""
import ctypes
from typing import Optional

import torch

from . import _lib

BK = 64  # K padding of split operands


def _p(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _chk(t: torch.Tensor, dtype, name: str):
    if not t.is_cuda:
        raise RuntimeError("funasr_b200.%s: tensor must live on a CUDA device (no CPU fallback)" % name)
    if t.dtype != dtype:
        raise RuntimeError("funasr_b200.%s: expected %s, got %s" % (name, dtype, t.dtype))
    return t


class StackProgram:
    """A stretch of the path recorded as ONE persistent cooperative kernel (csrc/stack_sm100.cu; fb200_stack_* in the
    C-ABI) for small requests.  Inside `with prog:` the wrappers `gemm`, `layernorm`, `fsmn` and `attention` of this module
    launch nothing: the calls are appended to the program (every tensor they name is kept alive by it).  `dim_values` are
    the row / batch CAPACITIES the recording runs with (pairwise distinct); `run(dims)` launches the program for a request
    with the actual counts."""

    def __init__(self, dim_values):
        self.dim_values = [int(v) for v in dim_values]
        self.handle = ctypes.c_void_p()
        self.keep = []
        self.n_ops = self.n_barriers = 0
        self._open = False

    def __enter__(self):
        global _recording
        if _recording is not None:
            raise RuntimeError("funasr_b200: a stack program is already being recorded")
        arr = (ctypes.c_int32 * len(self.dim_values))(*self.dim_values)
        _lib.call("fb200_stack_begin", ctypes.byref(self.handle), arr, len(self.dim_values))
        _recording, self._open = self, True
        return self

    def __exit__(self, exc_type, exc, tb):
        global _recording
        _recording, self._open = None, False
        if exc_type is not None:
            _lib.call("fb200_stack_destroy", self.handle)
            self.handle = ctypes.c_void_p()
            return False
        _lib.call("fb200_stack_end", self.handle)
        a, b = ctypes.c_int32(), ctypes.c_int32()
        _lib.call("fb200_stack_info", self.handle, ctypes.byref(a), ctypes.byref(b))
        self.n_ops, self.n_barriers = a.value, b.value
        return False

    def run(self, dims):
        arr = (ctypes.c_int32 * len(dims))(*[int(v) for v in dims])
        _lib.call("fb200_stack_run", self.handle, arr, len(dims), _stream())

    def profile(self, on: bool = True):
        """Per-op time stamps of every following run (fb200_stack_profile)."""
        _lib.call("fb200_stack_profile", self.handle, 1 if on else 0)

    def profile_read(self):
        """[(kind, barrier_after, [8 stamps in us since the program's start, -1 = not taken])] per op of the last run; kind
        0 GEMM, 1 LayerNorm, 2 FSMN, 3 attention; stamps: 0 start, 1 producer done, 2 first stage landed, 3 last MMA
        committed, 4 the last CTA's body done, 5 stores landed, 6 body done, 7 barrier passed."""
        n = max(self.n_ops, 1)
        kinds, bars, us = (ctypes.c_int32 * n)(), (ctypes.c_int32 * n)(), (ctypes.c_double * (8 * n))()
        mhz = ctypes.c_double()
        got = _lib.load().fb200_stack_profile_read(self.handle, kinds, bars, us, n, ctypes.byref(mhz))
        self.sm_mhz = mhz.value  # SM clock of the profiled run
        if got < 0:
            raise RuntimeError("fb200_stack_profile_read failed (%d): %s" % (got, _lib.load().fb200_last_error().decode()))
        return [(int(kinds[i]), int(bars[i]), [float(us[8 * i + k]) for k in range(8)]) for i in range(got)]

    def close(self):
        if self.handle:
            _lib.call("fb200_stack_destroy", self.handle)
            self.handle = ctypes.c_void_p()
        self.keep = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_recording: Optional[StackProgram] = None


def recording() -> bool:
    return _recording is not None


def _keep(*tensors):
    if _recording is not None:
        _recording.keep.extend(t for t in tensors if t is not None)


def round_up(x: int, m: int = BK) -> int:
    return (x + m - 1) // m * m


def split_width(k: int, nparts: int) -> int:
    return nparts * round_up(k)


def fbank(wave, cu_samples, cu_frames, total_frames, window, mel_bank, frame_shift=160, nfft=512,
          upscale=32768.0, preemph=0.97, out=None):
    _chk(wave, torch.float32, "fbank")
    _chk(cu_samples, torch.int64, "fbank")
    _chk(cu_frames, torch.int32, "fbank")
    n_mels = mel_bank.shape[0]
    if out is None:
        out = torch.empty(total_frames, n_mels, dtype=torch.float32, device=wave.device)
    _lib.call("fb200_fbank", _p(wave), _p(cu_samples), _p(cu_frames), cu_frames.numel() - 1, total_frames,
              _p(window), window.numel(), frame_shift, nfft, _p(mel_bank), n_mels, upscale, preemph, _p(out),
              _stream())
    return out


def lfr_cmvn(fb, cu_frames, cu_rows, total_rows, lfr_m=7, lfr_n=6, shift=None, scale=None, pe_inv=None,
             pe_scale=1.0, row_skip=0, out=None):
    n_mels = fb.shape[1]
    if out is None:
        out = torch.empty(total_rows, n_mels * lfr_m, dtype=torch.float32, device=fb.device)
    _lib.call("fb200_lfr_cmvn", _p(fb), _p(cu_frames), _p(cu_rows), cu_rows.numel() - 1, total_rows, n_mels,
              lfr_m, lfr_n, _p(shift), _p(scale), _p(pe_inv), pe_scale, row_skip, _p(out), out.stride(0), _stream())
    return out


def fill_head_rows(head, cu_rows, out, pe_inv=None, scale=1.0):
    """out[cu_rows[b] + j] = head[j] * scale + PE(j + 1) for the first head.shape[0] rows of every sequence."""
    _chk(head, torch.float32, "fill_head_rows")
    _lib.call("fb200_fill_head_rows", _p(head), head.shape[0], head.shape[1], _p(cu_rows), cu_rows.numel() - 1, _p(pe_inv),
              scale, _p(out), out.stride(0), _stream())
    return out


def scale_add_pe(x, cu_rows, pe_inv, scale, out=None):
    _chk(x, torch.float32, "scale_add_pe")
    if out is None:
        out = torch.empty_like(x)
    _lib.call("fb200_scale_add_pe", _p(x), x.stride(0), _p(cu_rows), cu_rows.numel() - 1, x.shape[0], x.shape[1],
              _p(pe_inv), scale, _p(out), out.stride(0), _stream())
    return out


def layernorm(x, gamma, beta, eps, out_f32=None, out_split=None, nparts=1, halo=False, cu_rows=None, rows=None):
    """x fp32 [rows, dim] (row stride allowed).  Writes out_f32 and/or out_split (preallocated)."""
    _chk(x, torch.float32, "layernorm")
    rows = x.shape[0] if rows is None else rows
    batch = 0 if cu_rows is None else cu_rows.numel() - 1
    _keep(x, gamma, beta, out_f32, out_split, cu_rows)
    _lib.call("fb200_layernorm", _p(x), x.stride(0), rows, x.shape[1], _p(gamma), _p(beta), eps, _p(out_f32),
              0 if out_f32 is None else out_f32.stride(0), _p(out_split), nparts, 1 if halo else 0, _p(cu_rows),
              batch, _stream())


def split_rows(x, nparts, out=None):
    _chk(x, torch.float32, "split_rows")
    rows, dim = x.shape
    if out is None:
        out = torch.empty(rows, split_width(dim, nparts), dtype=torch.bfloat16, device=x.device)
    _lib.call("fb200_split_rows", _p(x), x.stride(0), rows, dim, _p(out), nparts, _stream())
    return out


def pack_rows(padded, cu_rows, total_rows, out=None):
    """padded fp32 [B, Tmax, D] (any batch/row strides, unit column stride) -> packed [sum T, D]."""
    _chk(padded, torch.float32, "pack_rows")
    if padded.stride(2) != 1:
        padded = padded.contiguous()
    dim = padded.shape[2]
    if out is None:
        out = torch.empty(total_rows, dim, dtype=torch.float32, device=padded.device)
    _lib.call("fb200_pack_rows", _p(padded), padded.stride(0), padded.stride(1), _p(cu_rows), cu_rows.numel() - 1,
              total_rows, dim, _p(out), out.stride(0), _stream())
    return out


def unpack_rows(packed, cu_rows, max_len, pad_value=0.0, dim=None):
    """packed fp32 [sum T, >= D] -> dense padded [B, max_len, D]."""
    _chk(packed, torch.float32, "unpack_rows")
    dim = packed.shape[1] if dim is None else dim
    batch = cu_rows.numel() - 1
    out = torch.empty(batch, max_len, dim, dtype=torch.float32, device=packed.device)
    _lib.call("fb200_unpack_rows", _p(packed), packed.stride(0), _p(cu_rows), batch, max_len, dim, pad_value, _p(out),
              _stream())
    return out


def split_rows_halo(x, nparts, cu_rows):
    """fp32 [R, D] -> bf16 split in the CIF conv's halo layout [R + B + 1, nparts * Kp]."""
    _chk(x, torch.float32, "split_rows_halo")
    rows, dim = x.shape
    batch = cu_rows.numel() - 1
    out = torch.empty(rows + batch + 1, split_width(dim, nparts), dtype=torch.bfloat16, device=x.device)
    _lib.call("fb200_split_rows_halo", _p(x), x.stride(0), rows, dim, _p(out), nparts, _p(cu_rows), batch, _stream())
    return out


def fsmn(v, cu_rows, weight, out, residual=None, rows=None):
    """v fp32 [rows, dim] view (row stride allowed); weight fp32 [dim, k]; out fp32 [rows, dim]."""
    rows = v.shape[0] if rows is None else rows
    _keep(v, cu_rows, weight, out, residual)
    _lib.call("fb200_fsmn", _p(v), v.stride(0), _p(cu_rows), cu_rows.numel() - 1, rows, v.shape[1], _p(weight),
              weight.shape[-1], _p(residual), 0 if residual is None else residual.stride(0), _p(out), out.stride(0),
              _stream())
    return out


def gemm(a_split, a_parts, w_split, w_parts, m, n, k, bias=None, relu=False, res1=None, out_f32=None,
         out_split=None, out_parts=1, taps=1, impl=0, a_rows=None, f32_col0=0):
    """C[m, n] = A . W^T (+bias)(relu)(+res1) over bf16 split operands (tcgen05).
    `out_f32` receives columns [f32_col0, n) only (column c at c - f32_col0)."""
    _chk(a_split, torch.bfloat16, "gemm")
    _chk(w_split, torch.bfloat16, "gemm")
    a_rows = a_split.shape[0] if a_rows is None else a_rows
    _keep(a_split, w_split, bias, res1, out_f32, out_split)
    _lib.call("fb200_gemm", _p(a_split), a_rows, a_parts, _p(w_split), w_parts, m, n, k, taps, _p(bias),
              1 if relu else 0, _p(res1), 0 if res1 is None else res1.stride(0), _p(out_f32),
              0 if out_f32 is None else out_f32.stride(0), f32_col0, _p(out_split), out_parts, impl, _stream())


def attention_plan(cu_q, cu_kv, q_rows):
    """Work list of `attention` for one (cu_q, cu_kv) pair; reuse it for every call over the same batch."""
    batch = cu_q.numel() - 1
    cap = q_rows // 128 + batch
    plan = torch.empty(4 + 4 * cap, dtype=torch.int32, device=cu_q.device)
    _lib.call("fb200_attention_plan", _p(cu_q), _p(cu_kv), batch, _p(plan), cap, _stream())
    return plan


def attention(q, q_col0, q_ps, k, k_col0, v, v_col0, kv_ps, nparts, cu_q, cu_kv, max_q, heads, scale, out_f32=None,
              out_split=None, out_parts=None, impl=0, q_rows=None, kv_rows=None, head_dim=128, plan=None):
    """q / k / v: bf16 split matrices; head h of part p starts at column col0 + p * part_stride + h * head_dim."""
    _chk(q, torch.bfloat16, "attention")
    _chk(k, torch.bfloat16, "attention")
    _chk(v, torch.bfloat16, "attention")
    q_rows = q.shape[0] if q_rows is None else q_rows
    kv_rows = k.shape[0] if kv_rows is None else kv_rows
    _keep(q, k, v, cu_q, cu_kv, out_f32, out_split)
    if plan is None and impl == 0 and nparts < 3 and _recording is None:  # a recorded attention op needs no work list
        plan = attention_plan(cu_q, cu_kv, q_rows)
    _lib.call("fb200_attention", _p(q), q.stride(0), q_col0, q_ps, _p(k), k.stride(0), k_col0, _p(v), v.stride(0),
              v_col0, kv_ps, nparts, _p(cu_q), _p(cu_kv), _p(plan), cu_q.numel() - 1, q_rows, kv_rows, max_q, heads,
              head_dim, scale, _p(out_f32), 0 if out_f32 is None else out_f32.stride(0), _p(out_split),
              nparts if out_parts is None else out_parts, impl, _stream())


def cif_alpha(conv_relu, cu_rows, total_rows, w, b, alphas, smooth=1.0, noise=0.0):
    _lib.call("fb200_cif_alpha", _p(conv_relu), conv_relu.stride(0), _p(cu_rows), cu_rows.numel() - 1, total_rows,
              conv_relu.shape[1], _p(w), _p(b), smooth, noise, _p(alphas), _stream())


def cif_scan(alphas, cu_rows, tail, token_num, fire_pos, fire_rem, fires=None):
    _lib.call("fb200_cif_scan", _p(alphas), _p(cu_rows), cu_rows.numel() - 1, tail, _p(token_num), _p(fire_pos),
              _p(fire_rem), _p(fires), _stream())


def cif_integrate(hidden, alphas, cu_rows, cu_tok, total_tok, tail, fire_pos, fire_rem, embeds):
    _lib.call("fb200_cif_integrate", _p(hidden), hidden.stride(0), _p(alphas), _p(cu_rows), _p(cu_tok),
              cu_rows.numel() - 1, total_tok, hidden.shape[1], tail, _p(fire_pos), _p(fire_rem), _p(embeds),
              embeds.stride(0), _stream())


def lstm_parts(hidden: int = 512) -> int:
    import ctypes

    n = ctypes.c_int()
    _lib.call("fb200_lstm_parts", hidden, ctypes.byref(n))
    return int(n.value)


def lstm_bidir(xg, cu_rows, upsample, max_len3, w_hh, w_out, part, out_h=None, hidden=512, impl=0):
    """Bidirectional LSTM over packed rows fused with the Linear(2H -> 1) dot product: xg fp32 [rows3, >= 8H], w_hh fp32
    [2, 4H, H], w_out fp32 [2, H], part fp32 [lstm_parts(H), rows3] (out), out_h fp32 [rows3, 2H] or None."""
    import ctypes

    _chk(xg, torch.float32, "lstm_bidir")
    _chk(part, torch.float32, "lstm_bidir")
    batch, rows3 = cu_rows.numel() - 1, xg.shape[0]
    nbytes = ctypes.c_longlong()
    _lib.call("fb200_lstm_workspace_bytes", batch, hidden, ctypes.byref(nbytes))
    ws = torch.empty((int(nbytes.value) + 3) // 4, dtype=torch.float32, device=xg.device)
    _lib.call("fb200_lstm_bidir", _p(xg), xg.stride(0), _p(cu_rows), upsample, batch, rows3, max_len3, hidden, _p(w_hh),
              _p(w_out), _p(ws), int(nbytes.value), _p(part), _p(out_h), impl, _stream())


def us_fires(part, bias, smooth, noise, threshold, cu_rows, upsample, token_num, ids, cu_tok, drop_ids, us_alphas, fire_idx,
             fire_count, num_chars, us_peaks=None):
    """Upsampled alphas, their fires and the per-utterance character counts (see include/funasr_b200.h)."""
    import ctypes

    _chk(part, torch.float32, "us_fires")
    drop = (ctypes.c_int32 * 3)(*[int(d) for d in drop_ids])
    _lib.call("fb200_us_fires", _p(part), part.shape[0], part.shape[1], float(bias), float(smooth), float(noise),
              float(threshold), _p(cu_rows), upsample, cu_rows.numel() - 1, _p(token_num), _p(ids), _p(cu_tok),
              ctypes.cast(drop, ctypes.c_void_p), _p(us_alphas), _p(us_peaks), _p(fire_idx), _p(fire_count), _p(num_chars),
              _stream())


def asf_scores(q, k, cu_tok, cu_mem, max_words, heads, scale, scores):
    """Per hot-word row: attention probabilities summed over heads and tokens (fb200_asf_scores); q / k fp32."""
    _chk(q, torch.float32, "asf_scores")
    _chk(k, torch.float32, "asf_scores")
    _lib.call("fb200_asf_scores", _p(q), q.stride(0), _p(k), k.stride(0), _p(cu_tok), _p(cu_mem), cu_tok.numel() - 1, max_words,
              heads, q.shape[1] // heads, float(scale), _p(scores), _stream())


def merge_hotword_ids(dec_ids, dec_logp, hw_ids, hw_logp, no_bias, out_ids, out_logp=None):
    _lib.call("fb200_merge_hotword_ids", _p(dec_ids), _p(dec_logp), _p(hw_ids), _p(hw_logp), int(no_bias), dec_ids.numel(),
              _p(out_ids), _p(out_logp), _stream())


def argmax_logprob(logits, vocab, ids, score=None, rows=None):
    rows = logits.shape[0] if rows is None else rows
    _lib.call("fb200_argmax_logprob", _p(logits), logits.stride(0), rows, vocab, _p(ids), _p(score), _stream())


def gemm_argmax(a_split, a_parts, w_split, w_parts, m, n, k, ids, score=None, bias=None, workspace=None, a_rows=None):
    """ids[r] = argmax over the n columns of A . W^T + bias (first index on ties), score[r] = the log-softmax there; the
    [m, n] matrix is never written (fused epilogue + a small second pass).  Returns the workspace for reuse."""
    _chk(a_split, torch.bfloat16, "gemm_argmax")
    _chk(w_split, torch.bfloat16, "gemm_argmax")
    need = ctypes.c_longlong()
    _lib.call("fb200_gemm_argmax_workspace", m, n, ctypes.byref(need))
    if workspace is None or workspace.numel() * workspace.element_size() < need.value:
        workspace = torch.empty(max(need.value // 4, 4), dtype=torch.float32, device=a_split.device)
    a_rows = a_split.shape[0] if a_rows is None else a_rows
    _lib.call("fb200_gemm_argmax", _p(a_split), a_rows, a_parts, _p(w_split), w_parts, m, n, k, _p(bias), _p(workspace),
              workspace.numel() * workspace.element_size(), _p(ids), _p(score), _stream())
    return workspace


def ctc_collapse(ids, cu_rows, drop_mask, collapse, out_ids, out_len):
    _lib.call("fb200_ctc_collapse", _p(ids), _p(cu_rows), cu_rows.numel() - 1, drop_mask, 1 if collapse else 0,
              _p(out_ids), _p(out_len), _stream())


# ---- Conformer path (BASELINE config 4) -------------------------------------------------------------------

_ACT = {None: 0, "none": 0, "relu": 1, "swish": 2}


def _host_i32(values):
    import ctypes

    arr = (ctypes.c_int32 * len(values))(*[int(v) for v in values])
    return arr, ctypes.cast(arr, ctypes.c_void_p)


def gemm_ex(a, a_parts, w_split, w_parts, m, n, k, lda=0, a_part_stride=0, a_col0=0, tap_row=None, tap_col=None,
            taps=1, bias=None, act=None, res1=None, out_f32=None, out_split=None, out_parts=1, impl=0, a_rows=None,
            f32_col0=0):
    """fb200_gemm with a general A operand: row pitch, part stride, first column, per-tap row / column offsets
    (host lists), activation "relu" / "swish"."""
    _chk(a, torch.bfloat16, "gemm_ex")
    _chk(w_split, torch.bfloat16, "gemm_ex")
    a_rows = a.shape[0] if a_rows is None else a_rows
    keep_r, p_r = _host_i32(tap_row) if tap_row is not None else (None, None)
    keep_c, p_c = _host_i32(tap_col) if tap_col is not None else (None, None)
    _lib.call("fb200_gemm_ex", _p(a), a_rows, lda, a_part_stride, a_col0, a_parts, _p(w_split), w_parts, m, n, k, taps,
              p_r, p_c, _p(bias), _ACT[act], _p(res1), 0 if res1 is None else res1.stride(0), _p(out_f32),
              0 if out_f32 is None else out_f32.stride(0), f32_col0, _p(out_split), out_parts, impl, _stream())
    del keep_r, keep_c


def gemm_heads(a, a_parts, w_all, w_group_rows, m, n_group, groups, k, out_f32, lda=0, a_part_stride=0, a_col0=0,
               a_col_step=0, impl=0, a_rows=None):
    """out[r, g * P + c] = A[r, a_col0 + g * a_col_step : +k] . W[g * w_group_rows + c, :k], P = n_group rounded up to
    256, for every group g in ONE launch; out fp32 [m, >= groups * P]."""
    _chk(a, torch.bfloat16, "gemm_heads")
    _chk(w_all, torch.bfloat16, "gemm_heads")
    a_rows = a.shape[0] if a_rows is None else a_rows
    _lib.call("fb200_gemm_heads", _p(a), a_rows, lda, a_part_stride, a_col0, a_col_step, a_parts, _p(w_all), w_group_rows,
              a_parts, m, n_group, groups, k, _p(out_f32), out_f32.stride(0), impl, _stream())


def conv2d_sub1(feats, cu_frames, cu_seg, total_seg, weight, bias, planes, nparts):
    """feats fp32 [frames, n_mels]; planes bf16 [4 * total_seg * FP, nparts * C] (every slot written)."""
    _chk(feats, torch.float32, "conv2d_sub1")
    _lib.call("fb200_conv2d_sub1", _p(feats), feats.stride(0), _p(cu_frames), _p(cu_seg), cu_frames.numel() - 1,
              total_seg, feats.shape[1], _p(weight), _p(bias), weight.shape[0], _p(planes), nparts, _stream())


def gather_rows(src, cu_src, dst, cu_dst, dim=None):
    _chk(src, torch.float32, "gather_rows")
    dim = dst.shape[1] if dim is None else dim
    _lib.call("fb200_gather_rows", _p(src), src.stride(0), _p(cu_src), _p(dst), dst.stride(0), _p(cu_dst),
              cu_dst.numel() - 1, dst.shape[0], dim, _stream())
    return dst


def glu_dwconv(g, cu_rows, max_len, channels, weight, bias, bn_scale, bn_shift, out_split, nparts, rows=None):
    _chk(g, torch.float32, "glu_dwconv")
    rows = g.shape[0] if rows is None else rows
    _lib.call("fb200_glu_dwconv", _p(g), g.stride(0), _p(cu_rows), cu_rows.numel() - 1, rows, max_len, channels,
              _p(weight), weight.shape[-1], _p(bias), _p(bn_scale), _p(bn_shift), _p(out_split), nparts, _stream())


def attention_relpos(q, q_col0, q_ps, k, k_col0, v, v_col0, kv_ps, nparts, bd, cu_rows, max_len, heads, scale,
                     out_f32=None, out_split=None, out_parts=None, rows=None, head_dim=64, plan=None, impl=0):
    """bd fp32 [heads, rows, ld_bd] (contiguous heads); plan = attention_plan(cu_rows, cu_rows, rows) (impl 0)."""
    _chk(q, torch.bfloat16, "attention_relpos")
    _chk(bd, torch.float32, "attention_relpos")
    rows = q.shape[0] if rows is None else rows
    if plan is None and impl == 0 and nparts < 3:
        plan = attention_plan(cu_rows, cu_rows, rows)
    _lib.call("fb200_attention_relpos", _p(q), q.stride(0), q_col0, q_ps, _p(k), k.stride(0), k_col0, _p(v), v.stride(0),
              v_col0, kv_ps, nparts, _p(bd), bd.stride(1), bd.stride(0), _p(cu_rows), _p(plan), cu_rows.numel() - 1, rows,
              max_len, heads, head_dim, scale, _p(out_f32), 0 if out_f32 is None else out_f32.stride(0), _p(out_split),
              nparts if out_parts is None else out_parts, impl, _stream())
