"""The reference's C++ offline-stream API over the native runtime (csrc/funasr_compat.cpp -> libfunasr_b200_compat.so):
a caller written against the reference's OWN header (runtime/onnxruntime/include/funasrruntime.h: /root/reference here, the
unmodified copy under oracle/_ref/include on the GPU box) compiles and LINKS against the adapter -- the exported names and
C++ signatures are the reference's -- and on the GPU returns the text and time stamps of the native runtime."""
import os
import struct
import subprocess

import numpy as np
import pytest

from funasr_b200 import build, params, runtime, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER_DIRS = ["/root/reference/runtime/onnxruntime/include", os.path.join(ROOT, "oracle", "_ref", "include")]


def _header_dir():
    for d in HEADER_DIRS:
        if os.path.exists(os.path.join(d, "funasrruntime.h")):
            return d
    return None


needs_header = pytest.mark.skipif(_header_dir() is None, reason="reference header not present (make -C oracle)")


def _caller():
    build.build()
    libdir = os.path.join(ROOT, "funasr_b200")
    src, exe = os.path.join(ROOT, "examples", "funasr_api_caller.cpp"), os.path.join(ROOT, "examples", "funasr_api_caller")
    subprocess.run(["g++", "-O2", "-std=c++17", "-I" + _header_dir(), src, "-L" + libdir, "-lfunasr_b200_compat", "-lfunasr_b200",
                    "-Wl,-rpath," + libdir, "-o", exe], check=True)
    return exe


@needs_header
def test_caller_against_reference_header_links_and_fails_loudly_without_a_model(tmp_path):
    exe = _caller()
    r = subprocess.run([exe, str(tmp_path / "nothing"), str(tmp_path / "a.pcm")], capture_output=True, text=True)
    assert r.returncode == 1 and "FunOfflineInit failed" in r.stderr


@needs_header
def test_adapter_exports_the_reference_offline_api():
    build.build()
    out = subprocess.run(["nm", "-DC", "--defined-only", os.path.join(ROOT, "funasr_b200", "libfunasr_b200_compat.so")],
                         capture_output=True, text=True, check=True).stdout
    for name in ("FunOfflineInit(", "FunOfflineInferBuffer(", "FunOfflineInfer(", "FunOfflineUninit(", "FunOfflineReset(",
                 "FunASRGetResult(", "FunASRGetStamp(", "FunASRGetStampSents(", "FunASRGetRetNumber(", "FunASRFreeResult(",
                 "FunASRGetRetSnippetTime(", "CompileHotwordEmbedding("):
        assert name in out, name


def _write_wav(path, pcm):
    data = pcm.tobytes()
    with open(path, "wb") as f:
        f.write(b"RIFF" + struct.pack("<I", 36 + len(data)) + b"WAVEfmt " + struct.pack("<IHHIIHH", 16, 1, 1, 16000, 32000, 2, 16))
        f.write(b"data" + struct.pack("<I", len(data)) + data)


@needs_header
@pytest.mark.gpu
def test_reference_api_returns_the_native_runtime_results(tmp_path):
    exe = _caller()
    shapes = params.bicif_paraformer_shapes(enc_blocks=3, dec_blocks=2)
    sd = synth.synth_state_dict(shapes, seed=0)
    runtime.export_model(sd, str(tmp_path), cmvn=synth.synth_cmvn(), token_list=synth.synth_token_list())
    pcm = (np.clip(synth.synth_waveform(90, 48000), -1, 1) * 32768.0).astype(np.int16)
    pcm.tofile(tmp_path / "a.pcm")
    _write_wav(tmp_path / "a.wav", pcm)
    rec = runtime.OfflineRecognizer(str(tmp_path))
    ids, text = rec.recognize([pcm.astype(np.float32) / 32768.0], with_text=True)
    _, stamps = rec.recognize([pcm.astype(np.float32) / 32768.0], with_stamps=True)
    rec.close()
    want_stamp = "[" + ",".join("[%d,%d]" % (b, e) for b, e in stamps[0]) + "]" if stamps[0] else ""
    for name in ("a.pcm", "a.wav"):
        r = subprocess.run([exe, str(tmp_path), str(tmp_path / name)], capture_output=True, text=True, check=True)
        out = dict(line.split("\t", 1) for line in r.stdout.strip().splitlines())
        assert out["text"] == text[0] and out["text"] != ""
        assert out["stamp"] == want_stamp and want_stamp != ""
        assert out["seconds"] == "3.000" and out["number"] == "1"
