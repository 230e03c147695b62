"""Pin oracle/seaco_oracle.py to the UNMODIFIED reference's SeacoParaformer outputs WITH hot words (tests/golden/seaco.npz,
made by tests/golden/make_golden_seaco.py).  CPU only."""
import os

import numpy as np
import pytest
import torch

from funasr_b200 import params, synth
from oracle import bicif_oracle as B
from oracle import seaco_oracle as S


def _rel(a, b):
    a = torch.as_tensor(a, dtype=torch.float64)
    b = torch.as_tensor(b, dtype=torch.float64)
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


def test_param_inventory_has_the_reference_hot_word_tensors(golden):
    extra = set(params.seaco_paraformer_shapes()) - set(params.bicif_paraformer_shapes())
    assert extra == set(golden["extra_keys"].tolist()) and len(extra) == 127


def test_hotword_list_and_embeddings(golden, seaco_sd):
    hw = hotwords_of(golden)
    assert hw[-1] == [1] and [len(h) for h in hw] == [3, 1, 2, 5, 1]  # the list ends with [sos] (model.py:570)
    with torch.no_grad():
        emb = S.hotword_embeddings(hw, seaco_sd)
    assert _rel(emb, golden["hw_embed"]) < 1e-5


@pytest.mark.parametrize("ci", [0, 1])
def test_decode_with_hot_words_matches_reference(golden_dir, golden, seaco_sd, ci):
    b = np.load(os.path.join(golden_dir, "bicif.npz"))
    enc, embeds = torch.from_numpy(b["enc_%d" % ci]), torch.from_numpy(b["embeds_%d" % ci])
    with torch.no_grad():
        pred = S.seaco_decode(enc, embeds, seaco_sd, hotwords_of(golden))
    top = pred.topk(2, dim=-1)
    assert np.array_equal(top.indices.numpy().astype(np.int32), golden["hot_top2_ids_%d" % ci])
    assert _rel(top.values, golden["hot_top2_logp_%d" % ci]) < 1e-4
    ids = [i for i in top.indices[:, 0].tolist() if i not in (0, 1, 2)]
    assert ids == golden["hot_ids_%d" % ci].tolist()
    assert ids != b["ids_%d" % ci].tolist()  # the hot words do change the result


def test_keep_branch_matches_reference(golden_dir, golden, seaco_sd):
    """`_merge_res` (model.py:312-322): tokens whose hot-word argmax is NO_BIAS keep the plain decoder's distribution."""
    b = np.load(os.path.join(golden_dir, "bicif.npz"))
    sd = dict(seaco_sd)
    bias = sd["hotword_output_layer.bias"].clone()
    bias[S.NO_BIAS] += float(golden["mixed_bias_shift"])
    sd["hotword_output_layer.bias"] = bias
    taps = {}
    with torch.no_grad():
        pred = S.seaco_decode(torch.from_numpy(b["enc_0"]), torch.from_numpy(b["embeds_0"]), sd, hotwords_of(golden), taps=taps)
    keep = (taps["dha_ids"] == S.NO_BIAS).numpy()
    assert np.array_equal(keep, golden["mixed_keep_0"]) and 0 < keep.sum() < len(keep)
    top = pred.topk(2, dim=-1)
    assert np.array_equal(top.indices.numpy().astype(np.int32), golden["mixed_top2_ids_0"])
    assert _rel(top.values, golden["mixed_top2_logp_0"]) < 1e-4


def test_without_hot_words_is_bicif(golden, seaco_sd, golden_dir):
    b = np.load(os.path.join(golden_dir, "bicif.npz"))
    with torch.no_grad():
        pred = S.seaco_decode(torch.from_numpy(b["enc_0"]), torch.from_numpy(b["embeds_0"]), seaco_sd, None)
    assert [i for i in pred.argmax(-1).tolist() if i not in (0, 1, 2)] == golden["ids_0"].tolist()


def test_recognize_with_hot_words_matches_reference_inference(golden, seaco_sd):
    shift, scale = synth.synth_cmvn()
    cmvn = (torch.from_numpy(shift), torch.from_numpy(scale))
    wave = torch.from_numpy(synth.synth_waveform(int(golden["seeds"][0]), int(golden["num_samples"][0])))
    ids, stamps = S.recognize(wave, seaco_sd, cmvn, hotwords=hotwords_of(golden))
    assert ids == golden["hot_ids_0"].tolist()
    assert stamps == golden["hot_timestamp_0"].tolist()


def many_hotwords_of(g):
    flat, lens = g["many_flat"].tolist(), g["many_len"].tolist()
    out, o = [], 0
    for n in lens:
        out.append(flat[o: o + n])
        o += n
    return out


def test_attention_score_filter_matches_reference(golden_dir, golden, seaco_sd):
    """65 hot words: the 50 the decoder's hidden states attend to most are kept (+ the trailing [sos]), then the
    decode goes on with those (model.py:268-295)."""
    b = np.load(os.path.join(golden_dir, "bicif.npz"))
    hw = many_hotwords_of(golden)
    assert len(hw) == 65 and hw[-1] == [1]
    taps = {}
    with torch.no_grad():
        pred = S.seaco_decode(torch.from_numpy(b["enc_0"]), torch.from_numpy(b["embeds_0"]), seaco_sd, hw, taps=taps)
    assert taps["asf_keep"] == golden["many_keep"].tolist()
    top = pred.topk(2, dim=-1)
    assert np.array_equal(top.indices.numpy().astype(np.int32), golden["many_top2_ids_0"])
    assert _rel(top.values, golden["many_top2_logp_0"]) < 1e-4
