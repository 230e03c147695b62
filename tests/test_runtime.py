"""Native offline runtime (csrc/runtime.cu): model directory writer and failure behaviour on CPU; on the GPU the
C++ orchestration must give the token ids of the Python engine (and so of the oracle / reference goldens)."""
import os
import struct

import numpy as np
import pytest
import torch

from funasr_b200 import _lib, params, runtime, synth


def _small_sd():
    shapes = params.paraformer_large_shapes(enc_blocks=3, dec_blocks=2)
    return synth.synth_state_dict(shapes, seed=0)


def test_model_file_layout(tmp_path):
    sd = _small_sd()
    path = runtime.export_model(sd, str(tmp_path), cmvn=synth.synth_cmvn(), token_list=synth.synth_token_list())
    raw = open(path, "rb").read()
    assert raw[:8] == runtime.MAGIC
    (count,) = struct.unpack("<I", raw[8:12])
    assert count == len(sd) + 4  # + frontend.window / mel / pe_inv / meta
    (nlen,) = struct.unpack("<H", raw[12:14])
    name = raw[14:14 + nlen].decode()
    assert name in sd
    assert os.path.exists(tmp_path / "am.mvn") and os.path.exists(tmp_path / "tokens.json")


def test_init_fails_loudly_without_model_or_gpu(tmp_path):
    lib = _lib.load()
    h = lib.fb200_offline_init(str(tmp_path / "missing").encode(), 0, 0, 0)
    assert not h
    assert lib.fb200_last_error().decode() != ""
    assert lib.fb200_result_count(None) == 0
    with pytest.raises(RuntimeError):
        runtime.OfflineRecognizer(str(tmp_path / "missing"))


@pytest.mark.gpu
def test_native_runtime_matches_python_engine(tmp_path, cmvn):
    from funasr_b200 import engine

    sd = _small_sd()
    runtime.export_model(sd, str(tmp_path), cmvn=synth.synth_cmvn(), token_list=synth.synth_token_list())
    waves = [synth.synth_waveform(80 + i, n) for i, n in enumerate([32000, 300, 57001, 16000, 96000])]
    rec = runtime.OfflineRecognizer(str(tmp_path), precision="fp32", max_rows=600)  # several packed batches
    ids, text = rec.recognize(waves, with_text=True)
    eng = engine.ParaformerEngine({k: torch.from_numpy(v) for k, v in sd.items()}, "cuda:0", "fp32", cmvn=cmvn)
    want = eng.recognize(waves)
    assert ids == want
    assert ids[1] == [] and text[1] == ""
    toks = synth.synth_token_list()
    assert text[0] == "".join(toks[i] for i in ids[0])
    pcm = (np.clip(waves[3], -1, 1) * 32768.0).astype(np.int16)
    ids_pcm, _ = rec.recognize_pcm16(pcm.tobytes())
    assert ids_pcm == eng.recognize([pcm.astype(np.float32) / 32768.0])[0]
    rec.close()


@pytest.mark.gpu
def test_native_runtime_full_depth_matches_reference_golden(tmp_path, paraformer_sd, golden_dir):
    """Paraformer-large at full depth through the C++ runtime == the unmodified reference's transcripts."""
    g = np.load(os.path.join(golden_dir, "stages_2s.npz"))
    runtime.export_model(paraformer_sd, str(tmp_path), cmvn=synth.synth_cmvn())
    rec = runtime.OfflineRecognizer(str(tmp_path))
    ids, stamps = rec.recognize([synth.synth_waveform(10, 32000)], with_stamps=True)
    assert ids[0] == g["token_int"].tolist()
    assert stamps == [[]]  # no CifPredictorV3 head in this model directory
    rec.close()


@pytest.mark.gpu
def test_native_runtime_time_stamps_match_reference_golden(tmp_path, bicif_sd, golden_dir):
    """A BiCifParaformer model directory (predictor.blstm.* present) through the C++ runtime: ids and
    [begin ms, end ms] per token of a ragged batch == the unmodified reference's batch-1 results; a Paraformer
    directory gives no stamps."""
    g = np.load(os.path.join(golden_dir, "bicif.npz"))
    runtime.export_model(bicif_sd, str(tmp_path), cmvn=synth.synth_cmvn())
    rec = runtime.OfflineRecognizer(str(tmp_path), max_rows=2048)
    waves = [synth.synth_waveform(int(g["seeds"][ci]), int(g["num_samples"][ci])) for ci in range(3)]
    ids, stamps = rec.recognize(waves, with_stamps=True)
    for ci in range(3):
        assert ids[ci] == g["ids_%d" % ci].tolist()
        assert stamps[ci] == g["timestamp_%d" % ci].tolist()
    rec.close()


@pytest.mark.gpu
def test_native_runtime_graph_replay_matches_eager(tmp_path, cmvn, monkeypatch):
    """Requests whose shape comes back are replayed from a CUDA graph (third call on): same ids as the eager calls, for
    the same waveform and for ANOTHER waveform of the same length (the graph holds no data of the request), and as a
    runtime with graphs switched off."""
    sd = _small_sd()
    runtime.export_model(sd, str(tmp_path), cmvn=synth.synth_cmvn())
    a, b = synth.synth_waveform(70, 40000), synth.synth_waveform(71, 40000)
    monkeypatch.setenv("FB200_RT_GRAPHS", "0")
    eager = runtime.OfflineRecognizer(str(tmp_path))
    want_a, want_b = eager.recognize([a])[0], eager.recognize([b])[0]
    want_pair = eager.recognize([a, b])
    eager.close()
    monkeypatch.delenv("FB200_RT_GRAPHS")
    rec = runtime.OfflineRecognizer(str(tmp_path))
    for _ in range(4):
        assert rec.recognize([a])[0] == want_a
        assert rec.recognize([b])[0] == want_b
    for _ in range(3):
        assert rec.recognize([a, b]) == want_pair
    big = synth.synth_waveform(72, 16000 * 20)  # grows the waveform buffers: the graphs are rebuilt
    first = rec.recognize([big])[0]
    for _ in range(3):
        assert rec.recognize([big])[0] == first
        assert rec.recognize([a])[0] == want_a
    assert want_a != want_b
    rec.close()


@pytest.mark.gpu
def test_native_runtime_graph_replay_same_rows_other_frame_count(tmp_path, cmvn, monkeypatch):
    """Two request shapes with the SAME encoder rows, token counts apart or not, but different fbank frame counts
    (40000 vs 40160 samples: F = 248 / 249, R = 42 both): the activation layout behind the fbank buffer depends on F, so a
    decoder graph captured for one must never be replayed for the other (both graphs key on F)."""
    sd = _small_sd()
    runtime.export_model(sd, str(tmp_path), cmvn=synth.synth_cmvn())
    a, b = synth.synth_waveform(73, 40000), synth.synth_waveform(73, 40160)
    b[:40000] = a  # same audio plus 10 ms: most likely the same token count as well
    monkeypatch.setenv("FB200_RT_GRAPHS", "0")
    eager = runtime.OfflineRecognizer(str(tmp_path))
    want_a, want_b = eager.recognize([a])[0], eager.recognize([b])[0]
    eager.close()
    monkeypatch.delenv("FB200_RT_GRAPHS")
    rec = runtime.OfflineRecognizer(str(tmp_path))
    for _ in range(4):  # graphs captured on the second sighting, replayed afterwards, the two shapes alternating
        assert rec.recognize([a])[0] == want_a
        assert rec.recognize([b])[0] == want_b
    rec.close()


@pytest.mark.gpu
def test_native_runtime_more_than_1024_short_utterances_with_stamps(tmp_path):
    """A time-stamp model directory and 1100 half-second utterances: they fit one packed batch by rows (9 rows each), but the
    LSTM launch takes at most 1024 utterances -- the batcher must cut there instead of failing after the ids were computed."""
    shapes = params.bicif_paraformer_shapes(enc_blocks=2, dec_blocks=1)
    sd = synth.synth_state_dict(shapes, seed=0)
    runtime.export_model(sd, str(tmp_path), cmvn=synth.synth_cmvn())
    rec = runtime.OfflineRecognizer(str(tmp_path))
    base = [synth.synth_waveform(200 + i, 8000) for i in range(4)]
    waves = [base[i % 4] for i in range(1100)]
    ids, stamps = rec.recognize(waves, with_stamps=True)
    want_ids, want_stamps = rec.recognize(base, with_stamps=True)
    rec.close()
    assert len(ids) == 1100
    for i in range(1100):
        assert ids[i] == want_ids[i % 4] and stamps[i] == want_stamps[i % 4]
    from funasr_b200 import bicif

    eng = bicif.BiCifParaformerEngine({k: torch.from_numpy(v) for k, v in sd.items()}, "cuda:0", "fp32",
                                      cmvn=tuple(torch.from_numpy(x) for x in synth.synth_cmvn()))
    assert max(len(b) for b in eng.plan_batches([8000] * 1100)) <= 1024
    e_ids, e_stamps = eng.recognize(waves)
    assert e_ids == ids and e_stamps == stamps


def test_corrupt_model_file_is_rejected_not_fatal(tmp_path):
    """Header fields of model.fb200 are validated: a negative or absurd extent gives an error string, not an exception
    through the C boundary (std::terminate)."""
    sd = _small_sd()
    path = runtime.export_model(sd, str(tmp_path), cmvn=synth.synth_cmvn())
    raw = bytearray(open(path, "rb").read())
    (nlen,) = struct.unpack("<H", raw[12:14])
    at = 14 + nlen  # ndim byte of the first tensor, then int64 extents
    for bad in (-5, 1 << 40):
        broken = bytearray(raw)
        broken[at + 1: at + 9] = struct.pack("<q", bad)
        d = tmp_path / ("bad%d" % (bad > 0))
        d.mkdir()
        open(d / "model.fb200", "wb").write(bytes(broken))
        lib = _lib.load()
        assert not lib.fb200_offline_init(str(d).encode(), 0, 0, 0)
        assert "bad shape" in lib.fb200_last_error().decode()


def _example_binary():
    """examples/offline_recognize, built with g++ against the in-tree library when missing or stale."""
    import subprocess

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src, exe = os.path.join(root, "examples", "offline_recognize.cpp"), os.path.join(root, "examples", "offline_recognize")
    libdir = os.path.join(root, "funasr_b200")
    if not os.path.exists(exe) or os.path.getmtime(exe) < os.path.getmtime(src):
        subprocess.run(["g++", "-O2", "-std=c++17", "-I" + os.path.join(root, "include"), src, "-L" + libdir, "-lfunasr_b200",
                        "-Wl,-rpath," + libdir, "-o", exe], check=True)
    return exe, libdir


def test_example_program_builds_and_fails_loudly_without_a_model(tmp_path):
    import subprocess

    exe, libdir = _example_binary()
    r = subprocess.run([exe, str(tmp_path / "missing"), str(tmp_path / "a.pcm")], capture_output=True, text=True,
                       env=dict(os.environ, LD_LIBRARY_PATH=libdir))
    assert r.returncode == 1 and "fb200_offline_init" in r.stderr


@pytest.mark.gpu
def test_example_program_without_python_matches_engine(tmp_path, cmvn):
    """PCM16 files -> stand-alone C++ program (no Python in the process) -> ids == the Python engine on the same samples."""
    import subprocess

    from funasr_b200 import engine

    sd = _small_sd()
    runtime.export_model(sd, str(tmp_path), cmvn=synth.synth_cmvn(), token_list=synth.synth_token_list())
    exe, libdir = _example_binary()
    paths, quant = [], []
    for i, n in enumerate([24000, 40000]):
        pcm = (np.clip(synth.synth_waveform(90 + i, n), -1, 1) * 32767.0).astype(np.int16)
        path = str(tmp_path / ("u%d.pcm" % i))
        pcm.tofile(path)
        paths.append(path)
        quant.append(pcm.astype(np.float32) / 32768.0)
    r = subprocess.run([exe, str(tmp_path)] + paths, capture_output=True, text=True, env=dict(os.environ, LD_LIBRARY_PATH=libdir))
    assert r.returncode == 0, r.stderr
    got = [[int(t) for t in line.split("\t")[1].split()] for line in r.stdout.strip().splitlines()]
    eng = engine.ParaformerEngine({k: torch.from_numpy(v) for k, v in sd.items()}, "cuda:0", "fp32", cmvn=cmvn)
    assert got == eng.recognize(quant)
