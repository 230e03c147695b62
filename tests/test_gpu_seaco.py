"""SeacoParaformer's hot-word branch on the GPU against the reference's golden outputs (tests/golden/seaco.npz) and the
oracle: hot-word embeddings, the merged greedy picks, ids and time stamps with and without hot words."""
import os

import numpy as np
import pytest
import torch

from funasr_b200 import seaco, synth
from oracle import seaco_oracle as S

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


def rel(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.fixture(scope="module")
def golden(golden_dir):
    return np.load(os.path.join(golden_dir, "seaco.npz"))


def hotwords_of(g):
    flat, lens = g["hotwords_flat"].tolist(), g["hotwords_len"].tolist()
    out, o = [], 0
    for n in lens:
        out.append(flat[o: o + n])
        o += n
    return out


def _engine(sd):
    shift, scale = synth.synth_cmvn()
    return seaco.SeacoParaformerEngine(sd, DEV, "fp32", cmvn=(torch.from_numpy(shift), torch.from_numpy(scale)), max_rows=2048)


@pytest.fixture(scope="module")
def eng(seaco_sd):
    return _engine(seaco_sd)


def test_hotword_embeddings_match_reference(eng, golden):
    emb = eng.hot.run(hotwords_of(golden))
    torch.cuda.synchronize()
    assert emb.shape == golden["hw_embed"].shape
    assert rel(emb, golden["hw_embed"]) < 1e-4


def test_ids_and_stamps_with_hot_words_match_reference(eng, golden):
    """Two utterances in ONE packed batch, the hot-word list without its trailing [sos] (appended by the engine)."""
    waves = [synth.synth_waveform(int(golden["seeds"][ci]), int(golden["num_samples"][ci])) for ci in range(2)]
    ids, stamps = eng.recognize(waves, hotwords=hotwords_of(golden)[:-1])
    for ci in range(2):
        assert ids[ci] == golden["hot_ids_%d" % ci].tolist()
        assert stamps[ci] == golden["hot_timestamp_%d" % ci].tolist()
    # the hot words are per call: the next call without them gives BiCifParaformer's results
    ids, stamps = eng.recognize(waves)
    for ci in range(2):
        assert ids[ci] == golden["ids_%d" % ci].tolist()
        assert stamps[ci] == golden["timestamp_%d" % ci].tolist()


def test_merged_picks_match_reference_top1(eng, golden):
    wave = synth.synth_waveform(int(golden["seeds"][0]), int(golden["num_samples"][0]))
    eng.set_hotwords(hotwords_of(golden))
    taps = {}
    dev_wave, lens = eng.upload([wave])
    ids, logp, tok, rows, ts = eng.forward_packed_ts(dev_wave, lens, taps=taps)
    torch.cuda.synchronize()
    eng.set_hotwords(None)
    assert ids.cpu().tolist() == golden["hot_top2_ids_0"][:, 0].tolist()
    assert rel(logp, golden["hot_top2_logp_0"][:, 0]) < 1e-3


def test_keep_branch_matches_reference(seaco_sd, golden):
    """NO_BIAS bias raised as in the golden generator: about half of the tokens keep the plain decoder's pick."""
    sd = dict(seaco_sd)
    bias = sd["hotword_output_layer.bias"].clone()
    bias[S.NO_BIAS] += float(golden["mixed_bias_shift"])
    sd["hotword_output_layer.bias"] = bias
    e = _engine(sd)
    wave = synth.synth_waveform(int(golden["seeds"][0]), int(golden["num_samples"][0]))
    e.set_hotwords(hotwords_of(golden))
    taps = {}
    dev_wave, lens = e.upload([wave])
    ids, logp, tok, rows, ts = e.forward_packed_ts(dev_wave, lens, taps=taps)
    torch.cuda.synchronize()
    keep = (taps["hw_ids"] == S.NO_BIAS).cpu().numpy()
    assert np.array_equal(keep, golden["mixed_keep_0"])
    assert ids.cpu().tolist() == golden["mixed_top2_ids_0"][:, 0].tolist()
    assert rel(logp, golden["mixed_top2_logp_0"][:, 0]) < 1e-3


def test_matches_oracle_on_another_utterance(eng, seaco_sd):
    shift, scale = synth.synth_cmvn()
    cmvn = (torch.from_numpy(shift), torch.from_numpy(scale))
    hot = [[40, 41], [900], [17, 18, 19, 20, 21, 22, 23]]
    wave = synth.synth_waveform(320, 16000 * 3 + 9)
    ids, stamps = eng.recognize([wave], hotwords=hot)
    want_ids, want_stamps = S.recognize(torch.from_numpy(wave), seaco_sd, cmvn, hotwords=hot + [[1]])
    assert ids[0] == want_ids and stamps[0] == want_stamps


def test_attention_score_filter_matches_reference(eng, golden, seaco_sd):
    """65 hot words: scores, the kept 50 + [sos] and the merged picks of the reference; in a ragged batch every utterance
    gets its OWN filter (= the reference's batch-1 results; the oracle is the check for the second utterance)."""
    flat, lens, hot, o = golden["many_flat"].tolist(), golden["many_len"].tolist(), [], 0
    for n in lens:
        hot.append(flat[o: o + n])
        o += n
    waves = [synth.synth_waveform(int(golden["seeds"][0]), int(golden["num_samples"][0])), synth.synth_waveform(321, 16000 * 4)]
    eng.set_hotwords(hot)
    taps = {}
    dev_wave, wl = eng.upload(waves)
    ids, logp, tok, rows, ts = eng.forward_packed_ts(dev_wave, wl, taps=taps)
    torch.cuda.synchronize()
    eng.set_hotwords(None)
    assert rel(taps["asf_scores"][0], golden["many_scores"]) < 1e-4
    assert taps["asf_keep"][0].tolist() == golden["many_keep"].tolist()
    n0 = tok[0]
    assert ids[:n0].cpu().tolist() == golden["many_top2_ids_0"][:, 0].tolist()
    assert rel(logp[:n0], golden["many_top2_logp_0"][:, 0]) < 1e-3
    shift, scale = synth.synth_cmvn()
    want, _ = S.recognize(torch.from_numpy(waves[1]), seaco_sd, (torch.from_numpy(shift), torch.from_numpy(scale)), hotwords=hot)
    assert [i for i in ids[n0:].cpu().tolist() if i not in (0, 1, 2)] == want
