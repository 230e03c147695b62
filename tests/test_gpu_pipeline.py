"""End-to-end parity of the CUDA path on the B200 against (a) the committed outputs of the
UNMODIFIED reference (tests/golden/*.npz) and (b) the oracle on fresh seeded inputs.

Tolerances are BASELINE.json's: fbank/features within 1e-4 relative, encoder hidden within
1e-3 max relative (fp32 mode), CIF token counts and token ids identical.  bf16 mode:
encoder within 5e-2 max relative, token counts equal on the fixtures."""
import os

import numpy as np
import pytest
import torch

from funasr_b200 import engine, synth
from oracle import paraformer_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def rel(a, b):
    a = torch.as_tensor(a).double().cpu()
    b = torch.as_tensor(b).double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.fixture(scope="module")
def eng_fp32(paraformer_sd, cmvn):
    return engine.ParaformerEngine(paraformer_sd, DEV, "fp32", cmvn=cmvn)


@pytest.fixture(scope="module")
def eng_fp32x6(paraformer_sd, cmvn):
    return engine.ParaformerEngine(paraformer_sd, DEV, "fp32x6", cmvn=cmvn)


@pytest.fixture(scope="module")
def eng_bf16(paraformer_sd, cmvn):
    return engine.ParaformerEngine(paraformer_sd, DEV, "bf16", cmvn=cmvn)


@pytest.mark.parametrize("name,ui,n", [("stages_2s", 10, 32000), ("stages_10s", 11, 160000)])
def test_stages_match_reference_golden_fp32(golden_dir, eng_fp32, name, ui, n):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    eng = eng_fp32
    wave, lens = eng.upload([synth.synth_waveform(ui, n)])
    # frontend alone (no prologue) against the reference's extract_fbank output
    feats, cu_rows, rows = eng.frontend.run(wave, lens)
    assert feats.shape == g["feats"].shape
    assert rel(feats, g["feats"]) < 1e-4
    taps = {}
    ids, logp, tok = eng.forward_packed(wave, lens, taps=taps)
    torch.cuda.synchronize()
    assert rel(taps["enc"], g["enc"]) < 1e-3
    assert tok == [int(g["token_num"])]
    t = rows[0]
    assert float((taps["alphas"].cpu() - torch.from_numpy(g["alphas"][:t])).abs().max()) < 1e-4
    assert float((taps["fires"].cpu() - torch.from_numpy(g["fires"])).abs().max()) < 1e-3
    assert rel(taps["embeds"], g["embeds"]) < 1e-3
    lp = torch.log_softmax(taps["logits"].cpu(), dim=-1)
    top = lp.topk(2, dim=-1)
    assert float((top.values - torch.from_numpy(g["logp_top2"])).abs().max()) < 5e-3
    assert ids.cpu().tolist() == g["ids_top2"][:, 0].tolist()
    assert float((logp.cpu() - torch.from_numpy(g["logp_top2"][:, 0])).abs().max()) < 5e-3
    kept = [i for i in ids.cpu().tolist() if i not in (0, 1, 2)]
    assert kept == g["token_int"].tolist()


def test_encoder_layer_taps_match_reference_golden(golden_dir, eng_fp32):
    g = np.load(os.path.join(golden_dir, "stages_2s.npz"))
    eng = eng_fp32
    wave, lens = eng.upload([synth.synth_waveform(10, 32000)])
    feats, cu_rows, rows = eng.frontend.run(wave, lens, pe_scale=512 ** 0.5)
    taps = []
    eng.encoder.run(feats, cu_rows, max(rows), taps=taps)
    for idx in (0, 1, 49):
        assert rel(taps[idx], g["enc_layer%d" % idx]) < 1e-3, idx


def test_ragged_batch_transcripts_match_reference_golden(golden_dir, eng_fp32):
    """8 utterances of different lengths in ONE packed batch == the reference's batch-1 runs."""
    g = np.load(os.path.join(golden_dir, "transcripts.npz"))
    lens = g["num_samples"].tolist()
    waves = [synth.synth_waveform(100 + i, n) for i, n in enumerate(lens)]
    out = eng_fp32.recognize(waves)
    for i in range(len(lens)):
        assert out[i] == g["ids_%d" % i].tolist(), "utterance %d" % i
    # batch composition must not matter (no padding leak): one at a time gives the same ids
    assert eng_fp32.recognize(waves[2:3])[0] == out[2]


def test_multiple_batches_and_short_inputs(eng_fp32, paraformer_sd, cmvn):
    lens = [399, 16000, 24000, 400 + 160 * 5, 48000]
    waves = [synth.synth_waveform(300 + i, n) for i, n in enumerate(lens)]
    old = eng_fp32.max_rows
    eng_fp32.max_rows = 40  # force several batches
    try:
        out = eng_fp32.recognize(waves)
    finally:
        eng_fp32.max_rows = old
    assert out[0] == []  # shorter than one frame: no features, no tokens
    for i in (1, 2, 3, 4):
        ref = O.recognize(torch.from_numpy(waves[i]), paraformer_sd, cmvn)
        assert out[i] == ref, i


def test_long_audio_config5(eng_fp32, paraformer_sd, cmvn):
    """BASELINE configs[4]: long segments.  60 s and 120 s (T = 1000 / 2000 rows, 16 / 32 query tiles x 16 / 32 key tiles
    per head) against the oracle; 600 s (T = 10 000 rows, more than half the row budget of a batch) through
    batch-composition invariance: alone, and packed behind another utterance, it gives the same ids."""
    w60, w120 = synth.synth_waveform(700, 16000 * 60), synth.synth_waveform(701, 16000 * 120)
    taps = {}
    wave, lens = eng_fp32.upload([w60, w120])
    ids, _, tok = eng_fp32.forward_packed(wave, lens, taps=taps)
    assert taps["rows"] == [1000, 2000]
    o_enc, o_tok = 0, 0
    for w, t, n in zip((w60, w120), taps["rows"], tok):
        ref = {}
        want = O.recognize(torch.from_numpy(w), paraformer_sd, cmvn, taps=ref)
        assert rel(taps["enc"][o_enc: o_enc + t], ref["enc"]) < 1e-3
        assert n == ref["token_num"]
        got = ids[o_tok: o_tok + n].cpu().tolist()
        assert [v for v in got if v not in (0, 1, 2)] == want
        o_enc, o_tok = o_enc + t, o_tok + n
    w600 = synth.synth_waveform(702, 16000 * 600)
    alone = eng_fp32.recognize([w600])[0]
    packed = eng_fp32.recognize([w60, w600])
    assert len(alone) > 1000 and packed[1] == alone
    assert packed[0] == eng_fp32.recognize([w60])[0]


def test_bf16_mode_within_stated_tolerance(golden_dir, eng_bf16):
    g = np.load(os.path.join(golden_dir, "stages_10s.npz"))
    wave, lens = eng_bf16.upload([synth.synth_waveform(11, 160000)])
    taps = {}
    ids, logp, tok = eng_bf16.forward_packed(wave, lens, taps=taps)
    assert rel(taps["enc"], g["enc"]) < 5e-2
    assert abs(tok[0] - int(g["token_num"])) <= 1
    agree = np.mean(np.asarray(ids.cpu().tolist()[: len(g["ids_top2"])]) == g["ids_top2"][: len(ids), 0]) if tok[0] == int(g["token_num"]) else 0.0
    print("bf16 token agreement on random-init weights: %.3f" % agree)


def test_launch_counter_counts_kernels(eng_fp32):
    from funasr_b200 import _lib

    # a 2 s request takes the request-stack path (two persistent kernels + frontend / predictor / greedy pick) ...
    wave, lens = eng_fp32.upload([synth.synth_waveform(10, 32000)])
    _lib.launch_count(reset=True)
    eng_fp32.forward_packed(wave, lens)
    n = _lib.launch_count()
    assert 0 < n < 40 if eng_fp32.small_path() is not None else n > 400
    # ... two 30 s ones (1000 encoder rows) the per-op path: 50 encoder layers x 8 kernels + predictor + 16 decoder layers
    wave, lens = eng_fp32.upload([synth.synth_waveform(10, 480000), synth.synth_waveform(12, 480000)])
    _lib.launch_count(reset=True)
    eng_fp32.forward_packed(wave, lens)
    assert _lib.launch_count() > 400


def test_fp32x6_mode_matches_reference_golden(golden_dir, eng_fp32x6):
    g = np.load(os.path.join(golden_dir, "transcripts.npz"))
    lens = g["num_samples"].tolist()
    waves = [synth.synth_waveform(100 + i, n) for i, n in enumerate(lens)]
    out = eng_fp32x6.recognize(waves)
    for i in range(len(lens)):
        assert out[i] == g["ids_%d" % i].tolist(), "utterance %d" % i


def test_fp32_mode_error_budget_is_reported(golden_dir, eng_fp32, eng_fp32x6):
    """Encoder max-relative error of both parity modes against the reference (budget 1e-3)."""
    g = np.load(os.path.join(golden_dir, "stages_10s.npz"))
    for name, eng in (("fp32 (3 terms)", eng_fp32), ("fp32x6 (6 terms)", eng_fp32x6)):
        wave, lens = eng.upload([synth.synth_waveform(11, 160000)])
        taps = {}
        eng.forward_packed(wave, lens, taps=taps)
        err = rel(taps["enc"], g["enc"])
        print("%s: encoder max-rel error %.3e, alpha max abs %.3e" % (
            name, err, float((taps["alphas"].cpu() - torch.from_numpy(g["alphas"][:-1])).abs().max())))
        assert err < 1e-3


# ------------------------------------------------------------------- SenseVoiceSmall (config 3)
@pytest.fixture(scope="module")
def sv_eng(sensevoice_sd, cmvn):
    from funasr_b200.engine import SenseVoiceEngine

    return SenseVoiceEngine(sensevoice_sd, "cuda:0", precision="fp32", cmvn=cmvn)


SV_LANG = {0: "auto", 4: "en"}


@pytest.mark.gpu
@pytest.mark.parametrize("ci", [0, 1])
def test_sensevoice_stages_match_reference_golden(golden_dir, sv_eng, ci):
    """Encoder output within 1e-3 max-rel, per-frame argmax and CTC-greedy ids identical to the reference's
    SenseVoiceSmall.inference (tests/golden/sensevoice.npz)."""
    g = np.load(os.path.join(golden_dir, "sensevoice.npz"))
    ui, n, lang, itn = (int(v) for v in g["case_%d" % ci])
    wave = torch.from_numpy(synth.synth_waveform(ui, n)).to("cuda:0")
    taps = {}
    ids, counts, cu_rows, rows = sv_eng.forward_packed(wave, [n], SV_LANG[lang], "withitn" if itn else "woitn", taps=taps)
    assert rows == [g["enc_%d" % ci].shape[0]]
    assert rel(taps["enc"], g["enc_%d" % ci]) < 1e-3
    assert taps["frame_ids"].cpu().tolist() == g["argmax_%d" % ci].tolist()
    assert ids[: int(counts[0])].cpu().tolist() == g["ids_%d" % ci].tolist()


@pytest.mark.gpu
def test_sensevoice_ragged_batch_matches_oracle(sv_eng, sensevoice_sd, cmvn):
    """A ragged batch through recognize(): every utterance equals the oracle's batch-1 result."""
    from oracle import paraformer_oracle as O

    lens = [16000 * 2 + 11, 52000, 16000 * 4, 399]  # the last one is shorter than a frame -> []
    waves = [synth.synth_waveform(300 + i, n) for i, n in enumerate(lens)]
    got = sv_eng.recognize(waves, language="zh", textnorm="woitn")
    for w, ids in zip(waves, got):
        if len(w) < 400:
            assert ids == []
            continue
        assert ids == O.sense_voice_recognize(torch.from_numpy(w), sensevoice_sd, cmvn, language="zh", textnorm="woitn")


def test_vad_segments_second_stage_matches_oracle(eng_fp32, paraformer_sd, cmvn):
    """The long-audio pipeline's second stage (AutoModel.inference_with_vad, funasr/auto/auto_model.py:400-470):
    segments of one recording, out of duration order, as one packed call == the oracle on each slice."""
    from funasr_b200 import schedule

    speech = synth.synth_waveform(70, 16000 * 9)
    segs = [[300, 2400], [2500, 7000], [7100, 8100], [8200, 8999]]
    got = schedule.recognize_segments(eng_fp32, speech, segs)
    for ids, w in zip(got, schedule.slice_segments(speech, segs)):
        with torch.no_grad():
            assert ids == O.recognize(torch.from_numpy(w.copy()), paraformer_sd, cmvn)
