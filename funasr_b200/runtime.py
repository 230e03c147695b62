"""Python view of the native offline runtime (csrc/runtime.cu, `fb200_offline_*` in include/funasr_b200.h) and the
writer of its model directory.  The runtime itself is C++ over the library's kernels -- this module only exports
weights and calls it through ctypes (what a C++ server would do directly, cf. the reference's
runtime/websocket/bin/funasr-wss-server.cpp over funasrruntime.h)."""
import ctypes
import json
import os
import struct
from typing import Dict, List, Optional, Sequence

import numpy as np

from . import _lib, tables

MAGIC = b"FB200W1\x00"


def export_model(state_dict: Dict[str, "np.ndarray"], model_dir: str, cmvn=None, token_list: Optional[Sequence[str]] = None,
                 window: str = "hamming", n_mels: int = 80, lfr_m: int = 7, lfr_n: int = 6, fs: int = 16000,
                 frame_length_ms: float = 25.0, frame_shift_ms: float = 10.0, heads: int = 4,
                 tail_threshold: float = 0.45) -> str:
    """Write `model_dir/model.fb200` (+ `am.mvn`, `tokens.json`): every floating-point tensor of the state_dict under
    its reference name, fp32, plus the frontend tables the Python engine uses (so both front ends are bit-identical).
    Layout: magic, u32 count, then per tensor u16 name length, name, u8 ndim, i64 dims, fp32 data (little endian)."""
    os.makedirs(model_dir, exist_ok=True)
    frame_len = int(fs * frame_length_ms * 0.001)
    frame_shift = int(fs * frame_shift_ms * 0.001)
    nfft = 1 << (frame_len - 1).bit_length()
    items = {}
    for k, v in state_dict.items():
        a = v.detach().cpu().numpy() if hasattr(v, "detach") else np.asarray(v)
        if a.dtype.kind == "f":
            items[k] = np.ascontiguousarray(a, dtype=np.float32)
    items["frontend.window"] = tables.feature_window(window, frame_len).numpy()
    items["frontend.mel"] = tables.mel_filterbank(n_mels, nfft, float(fs)).numpy()
    items["frontend.pe_inv"] = tables.pe_inverse_timescales(n_mels * lfr_m).numpy()
    items["frontend.meta"] = np.asarray([lfr_m, lfr_n, frame_shift, nfft, heads, tail_threshold], dtype=np.float32)
    path = os.path.join(model_dir, "model.fb200")
    with open(path, "wb") as f:
        f.write(MAGIC)
        f.write(struct.pack("<I", len(items)))
        for name, a in items.items():
            nb = name.encode("utf-8")
            f.write(struct.pack("<H", len(nb)))
            f.write(nb)
            f.write(struct.pack("<B", a.ndim))
            f.write(struct.pack("<%dq" % a.ndim, *a.shape))
            f.write(np.ascontiguousarray(a, dtype="<f4").tobytes())
    if cmvn is not None:
        from . import synth

        synth.write_cmvn_file(os.path.join(model_dir, "am.mvn"), np.asarray(cmvn[0]), np.asarray(cmvn[1]))
    if token_list is not None:
        with open(os.path.join(model_dir, "tokens.json"), "w", encoding="utf-8") as f:
            json.dump(list(token_list), f, ensure_ascii=False)
    return path


class OfflineRecognizer:
    """`fb200_offline_init` / `_infer` / `_uninit` behind a Python object (tests and examples)."""

    def __init__(self, model_dir: str, device: int = 0, precision: str = "fp32", max_rows: int = 56832):
        self._lib = _lib.load()
        self._h = self._lib.fb200_offline_init(model_dir.encode("utf-8"), device, {"fp32": 0, "bf16": 1}[precision], max_rows)
        if not self._h:
            raise RuntimeError("fb200_offline_init failed: %s" % self._lib.fb200_last_error().decode())

    def _collect(self, res):
        if not res:
            raise RuntimeError("fb200_offline_infer failed: %s" % self._lib.fb200_last_error().decode())
        try:
            ids, text, stamps = [], [], []
            for i in range(self._lib.fb200_result_count(res)):
                n = ctypes.c_int(0)
                p = self._lib.fb200_result_ids(res, i, ctypes.byref(n))
                ids.append(np.ctypeslib.as_array(p, shape=(n.value,)).tolist() if n.value else [])
                text.append(self._lib.fb200_result_text(res, i).decode("utf-8"))
                p = self._lib.fb200_result_stamps(res, i, ctypes.byref(n))
                stamps.append(np.ctypeslib.as_array(p, shape=(n.value, 2)).tolist() if n.value else [])
            self.last_stamps = stamps  # [[begin ms, end ms], ...] per utterance; empty lists without a time-stamp head
            return ids, text
        finally:
            self._lib.fb200_free_result(res)

    def recognize(self, waves: Sequence[np.ndarray], with_text: bool = False, with_stamps: bool = False):
        """Token ids per utterance; `with_stamps` adds the [begin ms, end ms] lists of a BiCifParaformer model directory."""
        arrs = [np.ascontiguousarray(w, dtype=np.float32) for w in waves]
        ptrs = (ctypes.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])
        lens = (ctypes.c_int32 * len(arrs))(*[a.shape[0] for a in arrs])
        ids, text = self._collect(self._lib.fb200_offline_infer(self._h, ptrs, lens, len(arrs)))
        if with_stamps:
            return (ids, text, self.last_stamps) if with_text else (ids, self.last_stamps)
        return (ids, text) if with_text else ids

    def recognize_pcm16(self, pcm: bytes, sampling_rate: int = 16000):
        ids, text = self._collect(self._lib.fb200_offline_infer_buffer(self._h, pcm, len(pcm), sampling_rate))
        return ids[0], text[0]

    def close(self):
        if self._h:
            self._lib.fb200_offline_uninit(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # interpreter shutdown
            pass
