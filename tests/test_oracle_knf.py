"""Second pin of the fbank stage: the reference's OWN vendored C++ fbank (kaldi-native-fbank, used by its C++ runtime,
runtime/onnxruntime/src/paraformer.cpp:298-311), compiled from /root/reference by oracle/Makefile into oracle/_ref/.
The Python reference calls torchaudio.compliance.kaldi (third party, un-pinned by the reference's tests):
this checks the oracle's restatement -- and the CUDA kernel -- against independent reference code as well."""
import ctypes
import os

import numpy as np
import pytest
import torch

from funasr_b200 import synth, tables
from oracle import paraformer_oracle as O

LIB = os.path.join(os.path.dirname(os.path.abspath(O.__file__)), "_ref", "libknf_fbank.so")
needs_knf = pytest.mark.skipif(not os.path.exists(LIB), reason="oracle/_ref/libknf_fbank.so not built (make -C oracle)")

CASES = [(60, 400), (61, 16000), (62, 32491), (63, 160000)]


def knf_fbank(wave: np.ndarray, window: str = "hamming", n_mels: int = 80) -> np.ndarray:
    lib = ctypes.CDLL(LIB)
    lib.knf_fbank.restype = ctypes.c_int
    lib.knf_fbank.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_int, ctypes.c_float, ctypes.c_float,
                              ctypes.c_char_p, ctypes.c_void_p, ctypes.c_int]
    wave = np.ascontiguousarray(wave, dtype=np.float32)
    frames = lib.knf_fbank(wave.ctypes.data, wave.size, 16000.0, n_mels, 25.0, 10.0, window.encode(), None, 0)
    out = np.empty((frames, n_mels), dtype=np.float32)
    lib.knf_fbank(wave.ctypes.data, wave.size, 16000.0, n_mels, 25.0, 10.0, window.encode(), out.ctypes.data, frames)
    return out


def rel(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return float((a - b).abs().max() / b.abs().max())


@needs_knf
@pytest.mark.parametrize("window", ["hamming", "povey"])
def test_oracle_fbank_matches_the_references_cpp_fbank(window):
    for ui, n in CASES:
        wave = synth.synth_waveform(ui, n)
        ref = knf_fbank(wave, window)
        got = O.kaldi_fbank(torch.from_numpy(wave) * 32768.0, window=window)
        assert tuple(got.shape) == ref.shape == (tables.num_frames(n), 80)
        assert rel(got, ref) < 1e-4  # north_star tolerance; measured ~1e-5 between the two CPU implementations


@needs_knf
@pytest.mark.gpu
def test_cuda_fbank_matches_the_references_cpp_fbank():
    from funasr_b200 import ops

    dev = torch.device("cuda:0")
    waves = [synth.synth_waveform(ui, n) for ui, n in CASES]
    lens = [len(w) for w in waves]
    frames = [tables.num_frames(n) for n in lens]
    cu = lambda v, dt: torch.tensor([0] + list(np.cumsum(v)), dtype=dt, device=dev)  # noqa: E731
    out = ops.fbank(torch.from_numpy(np.concatenate(waves)).to(dev), cu(lens, torch.int64), cu(frames, torch.int32), sum(frames),
                    tables.feature_window("hamming", 400).to(dev), tables.mel_filterbank(80, 512, 16000.0).to(dev))
    # The two reference implementations differ from each other by 7e-5 on these inputs (one chance spectral null,
    # 95 dB below the peak, where fp32 rounding of the SAMPLES already moves the log-energy by 1e-3); the kernel is
    # held to the north-star 1e-4 against the Python reference's fbank (tests/test_gpu_kernels.py) and to 2e-4 here.
    o = 0
    for w, f in zip(waves, frames):
        assert rel(out[o: o + f], knf_fbank(w)) < 2e-4
        o += f
