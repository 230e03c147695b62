"""Drop-in boundary: the classes registered under the reference's keys keep its constructor
keywords, state_dict keys and forward contracts.  CPU part: registry + checkpoint compatibility
(against the UNMODIFIED reference when /root/reference is present).  GPU part: the reference's call
sequence of Paraformer.inference (funasr/models/paraformer/model.py:486-546) driven through the
drop-in modules with PADDED batches, compared with the oracle's batch-1 results."""
import os
import sys

import numpy as np
import pytest
import torch

from funasr_b200 import dropin, params, synth

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from ref_loader import reference_available  # noqa: E402

ENC = dict(output_size=512, attention_heads=4, linear_units=2048, num_blocks=50, dropout_rate=0.1,
           positional_dropout_rate=0.1, attention_dropout_rate=0.1, input_layer="pe", pos_enc_class="SinusoidalPositionEncoder",
           normalize_before=True, kernel_size=11, sanm_shfit=0, selfattention_layer_type="sanm")
DEC = dict(attention_heads=4, linear_units=2048, num_blocks=16, dropout_rate=0.1, positional_dropout_rate=0.1,
           self_attention_dropout_rate=0.1, src_attention_dropout_rate=0.1, att_layer_num=16, kernel_size=11, sanm_shfit=0)
PRED = dict(idim=512, threshold=1.0, l_order=1, r_order=1, tail_threshold=0.45)


def test_state_dict_keys_are_the_reference_inventory():
    enc = dropin.SANMEncoder(560, **ENC)
    dec = dropin.ParaformerSANMDecoder(8404, 512, **DEC)
    pred = dropin.CifPredictorV2(**PRED)
    got = {"encoder." + k: tuple(v.shape) for k, v in enc.state_dict().items()}
    got.update({"decoder." + k: tuple(v.shape) for k, v in dec.state_dict().items()})
    got.update({"predictor." + k: tuple(v.shape) for k, v in pred.state_dict().items()})
    assert got == dict(params.paraformer_large_shapes())
    assert enc.output_size() == 512


def test_unsupported_configurations_raise():
    with pytest.raises(NotImplementedError):
        dropin.SANMEncoder(560, **dict(ENC, input_layer="conv2d"))
    with pytest.raises(NotImplementedError):
        dropin.CifPredictorV2(512, 2, 1, tail_threshold=0.45)
    with pytest.raises(NotImplementedError):
        dropin.ParaformerSANMDecoder(8404, 512, **dict(DEC, sanm_shfit=5))


def test_forward_on_cpu_fails_loudly():
    enc = dropin.SANMEncoder(560, **dict(ENC, num_blocks=2))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        enc(torch.zeros(1, 8, 560), torch.tensor([8]))


@pytest.mark.skipif(not reference_available(), reason="reference tree not present (GPU box)")
def test_registers_into_reference_tables_and_loads_reference_checkpoint(tmp_path):
    import yaml
    from ref_loader import REFERENCE_ROOT, import_reference

    import_reference()
    from funasr import AutoModel
    from funasr.register import tables

    cfg = yaml.safe_load(open(os.path.join(REFERENCE_ROOT, "funasr/models/paraformer/template.yaml")))
    cfg["tokenizer_conf"]["token_list"] = synth.synth_token_list(8404)
    cfg["frontend_conf"]["dither"] = 0.0
    cfg["specaug"] = None
    for k in ("encoder_conf", "decoder_conf"):
        cfg[k]["num_blocks"] = 2  # keep the CPU test small; shapes per layer are the full ones
    cfg["decoder_conf"]["att_layer_num"] = 2
    ref = AutoModel(**cfg, device="cpu", disable_update=True, disable_pbar=True)
    ref_sd = ref.model.state_dict()
    saved = {k: dict(getattr(tables, k)) for k in ("frontend_classes", "encoder_classes", "predictor_classes", "decoder_classes",
                                                   "model_classes")}
    try:
        done = dropin.install()
        assert "encoder_classes[SANMEncoder]" in done
        ours = AutoModel(**cfg, device="cpu", disable_update=True, disable_pbar=True)
        assert type(ours.model.encoder) is dropin.SANMEncoder
        assert type(ours.model.decoder) is dropin.ParaformerSANMDecoder
        assert type(ours.model.predictor) is dropin.CifPredictorV2
        assert type(ours.kwargs["frontend"]) is dropin.WavFrontend
        assert ours.kwargs["frontend"].output_size() == 560
        # model_classes["Paraformer"]: the reference's class with `inference` replaced by the packed engine call
        from funasr.models.paraformer.model import Paraformer as RefParaformer
        assert isinstance(ours.model, RefParaformer) and isinstance(ours.model, dropin.ParaformerInference)
        assert type(ours.model).inference is dropin.ParaformerInference.inference
        mine = ours.model.state_dict()
        assert set(mine.keys()) == set(ref_sd.keys())
        assert all(mine[k].shape == ref_sd[k].shape for k in ref_sd)
        ours.model.load_state_dict(ref_sd, strict=True)  # a reference checkpoint loads unchanged
        with pytest.raises(RuntimeError, match="no CUDA device|no CPU fallback"):
            ours.generate(input=synth.synth_waveform(0, 16000))
    finally:
        for k, v in saved.items():
            getattr(tables, k).clear()
            getattr(tables, k).update(v)


MODEL = dict(encoder="SANMEncoder", encoder_conf=ENC, decoder="ParaformerSANMDecoder", decoder_conf=DEC,
             predictor="CifPredictorV2", predictor_conf=PRED, ctc_weight=0.0, input_size=560, vocab_size=8404)


class _Tokenizer:
    """What `inference` needs of funasr/tokenizer/char_tokenizer.py: ids2tokens, tokens2text."""

    def __init__(self):
        self.token_list = synth.synth_token_list(8404)
        self.bpemodel = None  # skips the reference's text post-processing (host side, not on the path)

    def ids2tokens(self, ids):
        return [self.token_list[i] for i in ids]

    def tokens2text(self, tokens):
        return " ".join(tokens)


def test_standalone_model_has_the_reference_state_dict_and_fails_loudly_on_cpu():
    model = dropin.Paraformer(**dict(MODEL, encoder_conf=dict(ENC, num_blocks=2), decoder_conf=dict(DEC, num_blocks=2, att_layer_num=2)))
    want = params.paraformer_large_shapes(enc_blocks=2, dec_blocks=2)
    assert {k: tuple(v.shape) for k, v in model.state_dict().items()} == dict(want)
    with_ctc = dropin.Paraformer(**dict(MODEL, ctc_weight=0.3, encoder_conf=dict(ENC, num_blocks=1),
                                        decoder_conf=dict(DEC, num_blocks=1, att_layer_num=1)))
    assert tuple(with_ctc.state_dict()["ctc.ctc_lo.weight"].shape) == (8404, 512)
    assert ("model_classes", "Paraformer") in dropin.DROPIN_CLASSES
    frontend = dropin.WavFrontend(lfr_m=7, lfr_n=6, dither=0.0)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        model.inference([synth.synth_waveform(0, 16000)], key=["a"], tokenizer=None, frontend=frontend, device="cpu")
    with pytest.raises(NotImplementedError):
        model.inference([synth.synth_waveform(0, 16000)], key=["a"], frontend=frontend, pred_timestamp=True)


# ---------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
def test_model_inference_override_returns_the_reference_results(paraformer_sd, cmvn, tmp_path):
    """`Paraformer.inference` (model.py:443-596) through the packed engine: results list + meta_data as the
    reference returns them, token ids equal to the oracle's batch-1 recognition."""
    from oracle import paraformer_oracle as O

    dev = torch.device("cuda:0")
    shift, scale = synth.synth_cmvn()
    cmvn_file = str(tmp_path / "am.mvn")
    synth.write_cmvn_file(cmvn_file, shift, scale)
    frontend = dropin.WavFrontend(cmvn_file=cmvn_file, fs=16000, window="hamming", n_mels=80, frame_length=25,
                                  frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0)
    model = dropin.Paraformer(**MODEL).to(dev)
    model.load_state_dict(paraformer_sd, strict=True)
    lens = [48000, 16000 * 2 + 77, 300, 70000]  # the third is shorter than one frame
    waves = [synth.synth_waveform(500 + i, n) for i, n in enumerate(lens)]
    tok = _Tokenizer()
    want = [O.recognize(torch.from_numpy(w), paraformer_sd, cmvn) if len(w) >= 400 else [] for w in waves]
    # 1. a list of arrays with keys, no tokenizer -> token_int
    results, meta = model.inference(waves, key=["a", "b", "c", "d"], tokenizer=None, frontend=frontend, device="cuda:0")
    assert [r["key"] for r in results] == ["a", "b", "c", "d"]
    assert [r["token_int"] for r in results] == want
    rows = sum((1 + (n - 400) // 160 + 5) // 6 for n in lens if n >= 400)
    assert meta["batch_data_time"] == pytest.approx(rows * 10 * 6 / 1000)
    assert set(meta) == {"load_data", "extract_feat", "batch_data_time"}
    # 2. a padded tensor + lengths (what extract_fbank's callers hold), tokenizer -> text
    data = torch.nn.utils.rnn.pad_sequence([torch.from_numpy(w) for w in waves], batch_first=True)
    results, _ = model.inference(data, data_lengths=torch.tensor(lens), key=[["a", "b", "c", "d"]], tokenizer=tok,
                                 frontend=frontend, device="cuda:0")
    assert [r["text"] for r in results] == [" ".join(tok.ids2tokens(w)) for w in want]
    # 3. nothing recognisable -> [] (model.py:514-515)
    assert model.inference([waves[2]], key=["c"], tokenizer=tok, frontend=frontend, device="cuda:0") == []
    # 4. new weights are picked up (the engine's weight pack follows the parameters)
    with torch.no_grad():
        model.decoder.get_parameter("output_layer.bias").add_(torch.arange(8404, device=dev).float())
    results, _ = model.inference([waves[0]], key=["a"], tokenizer=None, frontend=frontend, device="cuda:0")
    assert results[0]["token_int"] and set(results[0]["token_int"]) == {8403}


@pytest.mark.gpu
def test_reference_call_sequence_through_dropin_modules_matches_oracle(paraformer_sd, cmvn, tmp_path):
    from oracle import paraformer_oracle as O

    dev = torch.device("cuda:0")
    shift, scale = synth.synth_cmvn()
    cmvn_file = str(tmp_path / "am.mvn")
    synth.write_cmvn_file(cmvn_file, shift, scale)
    frontend = dropin.WavFrontend(cmvn_file=cmvn_file, fs=16000, window="hamming", n_mels=80, frame_length=25,
                                  frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0)
    enc = dropin.SANMEncoder(560, **ENC).to(dev)
    pred = dropin.CifPredictorV2(**PRED).to(dev)
    dec = dropin.ParaformerSANMDecoder(8404, 512, **DEC).to(dev)
    enc.load_state_dict({k[len("encoder."):]: v for k, v in paraformer_sd.items() if k.startswith("encoder.")}, strict=True)
    pred.load_state_dict({k[len("predictor."):]: v for k, v in paraformer_sd.items() if k.startswith("predictor.")}, strict=True)
    dec.load_state_dict({k[len("decoder."):]: v for k, v in paraformer_sd.items() if k.startswith("decoder.")}, strict=True)

    lens = [32000, 51234, 16000 * 4]
    waves = [torch.from_numpy(synth.synth_waveform(200 + i, n)) for i, n in enumerate(lens)]
    # extract_fbank (funasr/utils/load_utils.py:207-216): pad_sequence -> frontend(data, lens)
    data = torch.nn.utils.rnn.pad_sequence(waves, batch_first=True)
    speech, speech_lengths = frontend(data, torch.tensor(lens, dtype=torch.int32))
    speech_lengths = speech_lengths.to(dev)
    # Paraformer.encode (model.py:239-266)
    encoder_out, encoder_out_lens, _ = enc(speech, speech_lengths)
    # calc_predictor (model.py:268-276)
    t = encoder_out.shape[1]
    mask = (torch.arange(t, device=dev)[None, :] < encoder_out_lens[:, None]).float()[:, None, :]
    embeds, token_num, alphas, peaks = pred(encoder_out, None, mask, ignore_id=-1)
    pre_token_length = token_num.round().long()
    # cal_decoder_with_predictor (model.py:278-285)
    logits, _ = dec(encoder_out, encoder_out_lens, embeds, pre_token_length)
    decoder_out = torch.log_softmax(logits, dim=-1)
    for i, w in enumerate(waves):
        taps = {}
        want = O.recognize(w, paraformer_sd, cmvn, taps=taps)
        n_tok = int(pre_token_length[i])
        assert n_tok == taps["token_num"]
        am_scores = decoder_out[i, :n_tok]
        yseq = am_scores.argmax(dim=-1).tolist()
        assert [v for v in yseq if v not in (0, 1, 2)] == want
        ti = int(encoder_out_lens[i])
        ref_enc = taps["enc"]
        err = float((encoder_out[i, :ti].cpu() - ref_enc).abs().max() / ref_enc.abs().max())
        assert err < 1e-3
        assert float(encoder_out[i, ti:].abs().max()) == 0.0 if ti < t else True
        assert float((alphas[i, : ti + 1].cpu() - taps["alphas"]).abs().max()) < 1e-4
        assert float((peaks[i, : ti + 1].cpu() - taps["fires"]).abs().max()) < 1e-3


# ------------------------------------------------------------------- SenseVoiceSmall (config 3)
SV_ENC = dict(output_size=512, attention_heads=4, linear_units=2048, num_blocks=50, tp_blocks=20, dropout_rate=0.1,
              positional_dropout_rate=0.1, attention_dropout_rate=0.1, input_layer="pe", normalize_before=True,
              kernel_size=11, sanm_shfit=0, selfattention_layer_type="sanm")


def test_sensevoice_encoder_state_dict_keys_are_the_reference_inventory():
    enc = dropin.SenseVoiceEncoderSmall(560, **SV_ENC)
    got = {"encoder." + k: tuple(v.shape) for k, v in enc.state_dict().items()}
    want = {k: v for k, v in params.sense_voice_small_shapes().items() if k.startswith("encoder.")}
    assert got == want
    assert ("encoder_classes", "SenseVoiceEncoderSmall") in dropin.DROPIN_CLASSES
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        enc(torch.zeros(1, 8, 560), torch.tensor([8]))


@pytest.mark.gpu
def test_sensevoice_call_sequence_through_dropin_encoder_matches_golden(sensevoice_sd, cmvn, golden_dir, tmp_path):
    """SenseVoiceSmall.inference's own steps (funasr/models/sense_voice/model.py:852-906) around the drop-in
    frontend and encoder, on a PADDED batch of the two golden utterances."""
    dev = torch.device("cuda:0")
    g = np.load(os.path.join(golden_dir, "sensevoice.npz"))
    shift, scale = synth.synth_cmvn()
    cmvn_file = str(tmp_path / "am.mvn")
    synth.write_cmvn_file(cmvn_file, shift, scale)
    frontend = dropin.WavFrontend(cmvn_file=cmvn_file, fs=16000, window="hamming", n_mels=80, frame_length=25,
                                  frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0)
    enc = dropin.SenseVoiceEncoderSmall(560, **SV_ENC).to(dev)
    enc.load_state_dict({k[len("encoder."):]: v for k, v in sensevoice_sd.items() if k.startswith("encoder.")}, strict=True)
    embed = sensevoice_sd["embed.weight"].to(dev)
    w_ctc, b_ctc = sensevoice_sd["ctc.ctc_lo.weight"].to(dev), sensevoice_sd["ctc.ctc_lo.bias"].to(dev)
    for ci in (0, 1):  # one language / text-norm query per call, like the reference
        ui, n, lang, itn = (int(v) for v in g["case_%d" % ci])
        wave = torch.from_numpy(synth.synth_waveform(ui, n))
        speech, speech_lengths = frontend(wave[None, :], torch.tensor([n], dtype=torch.int32))
        speech_lengths = speech_lengths.to(dev)
        q = embed[torch.tensor([lang, 1, 2, 14 if itn else 15], device=dev)][None]
        speech = torch.cat([q, speech], dim=1)
        speech_lengths = speech_lengths + 4
        encoder_out, encoder_out_lens = enc(speech, speech_lengths)
        ref = torch.from_numpy(g["enc_%d" % ci]).to(dev)
        assert encoder_out.shape[1:] == ref.shape and int(encoder_out_lens[0]) == ref.shape[0]
        assert float((encoder_out[0] - ref).abs().max() / ref.abs().max()) < 1e-3
        logp = torch.log_softmax(torch.nn.functional.linear(encoder_out[0], w_ctc, b_ctc), dim=-1)
        yseq = torch.unique_consecutive(logp.argmax(-1))
        assert yseq[yseq != 0].tolist() == g["ids_%d" % ci].tolist()


SV_MODEL = dict(encoder="SenseVoiceEncoderSmall", encoder_conf=SV_ENC, input_size=560, vocab_size=25055)


class _SvTokenizer:
    """What SenseVoiceSmall.inference needs of the sentencepiece tokenizer: decode(list) -> str."""

    def decode(self, ids):
        return " ".join("t%d" % i for i in ids)


def test_standalone_sensevoice_model_has_the_reference_state_dict():
    model = dropin.SenseVoiceSmall(**dict(SV_MODEL, encoder_conf=dict(SV_ENC, num_blocks=2, tp_blocks=1)))
    want = params.sense_voice_small_shapes(enc_blocks=2, tp_blocks=1)
    assert {k: tuple(v.shape) for k, v in model.state_dict().items()} == dict(want)
    assert ("model_classes", "SenseVoiceSmall") in dropin.DROPIN_CLASSES
    frontend = dropin.WavFrontend(lfr_m=7, lfr_n=6, dither=0.0)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        model.inference([synth.synth_waveform(0, 16000)], key=["a"], tokenizer=_SvTokenizer(), frontend=frontend)
    with pytest.raises(NotImplementedError):
        model.inference([synth.synth_waveform(0, 16000)], key=["a"], frontend=frontend, output_timestamp=True)


@pytest.mark.skipif(not reference_available(), reason="reference tree not present (GPU box)")
def test_install_wraps_the_reference_model_classes():
    from ref_loader import import_reference

    import_reference()
    from funasr.models.sense_voice.model import SenseVoiceSmall as RefSenseVoice
    from funasr.register import tables

    saved = {k: dict(getattr(tables, k)) for k in ("frontend_classes", "encoder_classes", "predictor_classes", "decoder_classes",
                                                   "model_classes")}
    try:
        dropin.install()
        cls = tables.model_classes["SenseVoiceSmall"]
        assert issubclass(cls, RefSenseVoice) and issubclass(cls, dropin.SenseVoiceInference)
        assert cls.inference is dropin.SenseVoiceInference.inference
        model = cls(**dict(SV_MODEL, encoder_conf=dict(SV_ENC, num_blocks=2, tp_blocks=1)))
        assert type(model.encoder) is dropin.SenseVoiceEncoderSmall  # built through the (re-registered) tables
        want = params.sense_voice_small_shapes(enc_blocks=2, tp_blocks=1)
        assert {k: tuple(v.shape) for k, v in model.state_dict().items()} == dict(want)
    finally:
        for k, v in saved.items():
            getattr(tables, k).clear()
            getattr(tables, k).update(v)


@pytest.mark.gpu
def test_sensevoice_model_inference_override_matches_golden(sensevoice_sd, golden_dir, tmp_path):
    """`SenseVoiceSmall.inference` (sense_voice/model.py:809-906) through the packed engine: token ids of the
    reference's own run (tests/golden/sensevoice.npz), decoded by the tokenizer it is handed."""
    dev = torch.device("cuda:0")
    g = np.load(os.path.join(golden_dir, "sensevoice.npz"))
    shift, scale = synth.synth_cmvn()
    cmvn_file = str(tmp_path / "am.mvn")
    synth.write_cmvn_file(cmvn_file, shift, scale)
    frontend = dropin.WavFrontend(cmvn_file=cmvn_file, fs=16000, window="hamming", n_mels=80, frame_length=25,
                                  frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0)
    model = dropin.SenseVoiceSmall(**SV_MODEL).to(dev)
    model.load_state_dict(sensevoice_sd, strict=True)
    tok = _SvTokenizer()
    names = {0: "auto", 4: "en"}
    for ci in (0, 1):
        ui, n, lang, itn = (int(v) for v in g["case_%d" % ci])
        wave = synth.synth_waveform(ui, n)
        results, meta = model.inference([wave], key=["k%d" % ci], tokenizer=tok, frontend=frontend, device="cuda:0",
                                        language=names[lang], use_itn=bool(itn))
        assert results == [{"key": "k%d" % ci, "text": tok.decode(g["ids_%d" % ci].tolist())}]
        assert meta["batch_data_time"] == pytest.approx(((1 + (n - 400) // 160 + 5) // 6) * 0.06)
    # ban_emo_unk (model.py:886-887): the banned id never appears; banning the most frequent id changes the result
    ui, n, lang, itn = (int(v) for v in g["case_0"])
    wave = synth.synth_waveform(ui, n)
    base = model._engine_for(frontend).recognize([wave])[0]
    top = max(set(base), key=base.count)
    banned = model._engine_for(frontend).recognize([wave], ban_ids=(top,))[0]
    assert top not in banned and banned != base


# ---- ConformerEncoder (BASELINE config 4) -------------------------------------------------------------------

CONF_ENC = dict(output_size=256, attention_heads=4, linear_units=2048, num_blocks=12, dropout_rate=0.1,
                positional_dropout_rate=0.1, attention_dropout_rate=0.0, input_layer="conv2d", normalize_before=True,
                pos_enc_layer_type="rel_pos", selfattention_layer_type="rel_selfattn", activation_type="swish",
                macaron_style=True, use_cnn_module=True, cnn_module_kernel=15)


def test_conformer_encoder_state_dict_keys_are_the_reference_inventory():
    enc = dropin.ConformerEncoder(80, **CONF_ENC)
    got = {"encoder." + k: tuple(v.shape) for k, v in enc.state_dict().items() if not k.endswith("num_batches_tracked")}
    assert got == dict(params.conformer_encoder_shapes())
    assert enc.output_size() == 256
    with pytest.raises(NotImplementedError):
        dropin.ConformerEncoder(80, **dict(CONF_ENC, rel_pos_type="latest"))
    with pytest.raises(NotImplementedError):
        dropin.ConformerEncoder(80, **dict(CONF_ENC, input_layer="linear"))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        enc(torch.zeros(1, 40, 80), torch.tensor([40]))


@pytest.mark.skipif(not reference_available(), reason="reference tree not present (GPU box)")
def test_conformer_encoder_loads_the_reference_state_dict_strictly():
    from ref_loader import import_reference

    import_reference()
    from funasr.register import tables

    ref_cls = tables.encoder_classes["ConformerEncoder"]
    ref = ref_cls(input_size=80, **dict(CONF_ENC, num_blocks=2))
    mine = dropin.ConformerEncoder(80, **dict(CONF_ENC, num_blocks=2))
    assert {k: tuple(v.shape) for k, v in ref.state_dict().items()} == {k: tuple(v.shape) for k, v in mine.state_dict().items()}
    mine.load_state_dict(ref.state_dict(), strict=True)
    try:
        assert "encoder_classes[ConformerEncoder]" in dropin.install(tables)
        assert tables.encoder_classes["ConformerEncoder"] is dropin.ConformerEncoder
    finally:
        tables.encoder_classes["ConformerEncoder"] = ref_cls


@pytest.mark.gpu
def test_conformer_dropin_padded_forward_matches_oracle(conformer_sd):
    """The reference's encoder call `self.encoder(speech, speech_lengths)` (funasr/models/ctc/model.py:139) with a
    PADDED batch through the drop-in module == the oracle's batch-1 runs."""
    from oracle import conformer_oracle as C
    from oracle import paraformer_oracle as P

    enc = dropin.ConformerEncoder(80, **CONF_ENC)
    enc.load_state_dict({k[len("encoder."):]: v for k, v in conformer_sd.items() if k.startswith("encoder.")}, strict=False)
    enc = enc.cuda().eval()
    waves = [synth.synth_waveform(60, 24000), synth.synth_waveform(61, 50000), synth.synth_waveform(62, 9000)]
    feats = [P.wav_frontend(torch.from_numpy(w), None, window="hamming", lfr_m=1, lfr_n=1) for w in waves]
    lens = [f.shape[0] for f in feats]
    pad = torch.nn.utils.rnn.pad_sequence(feats, batch_first=True).cuda()
    with torch.no_grad():
        out, olens, _ = enc(pad, torch.tensor(lens).cuda())
    assert out.shape[:2] == (3, ((max(lens) - 1) // 2 - 1) // 2)
    for b, f in enumerate(feats):
        with torch.no_grad():
            want = C.conformer_encoder(f, conformer_sd)
        assert int(olens[b]) == want.shape[0]
        got = out[b, : want.shape[0]].cpu()
        assert float((got - want).abs().max() / want.abs().max()) < 1e-3
        assert float(out[b, want.shape[0]:].abs().max()) == 0.0 if want.shape[0] < out.shape[1] else True


class _CtcTokenizer:
    def ids2tokens(self, ids):
        return ["t%d" % i for i in ids]

    def tokens2text(self, tokens):
        return " ".join(tokens)


def test_standalone_ctc_model_has_the_reference_state_dict_and_fails_loudly_on_cpu():
    m = dropin.CTC(encoder="ConformerEncoder", encoder_conf=dict(CONF_ENC, num_blocks=1), input_size=80, vocab_size=50)
    keys = {k for k in m.state_dict() if not k.endswith("num_batches_tracked")}
    assert keys == set(params.conformer_ctc_shapes(50, num_blocks=1))
    fe = dropin.WavFrontend(cmvn_file=None, window="hamming", lfr_m=1, lfr_n=1, dither=0.0)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m.inference([synth.synth_waveform(0, 16000)], key=["a"], tokenizer=None, frontend=fe, device="cpu")


@pytest.mark.skipif(not reference_available(), reason="reference tree not present (GPU box)")
def test_install_wraps_the_reference_ctc_model_class():
    from ref_loader import import_reference

    import_reference()
    from funasr.register import tables

    saved = (dict(tables.model_classes), dict(tables.encoder_classes), dict(tables.frontend_classes),
             dict(tables.predictor_classes), dict(tables.decoder_classes))
    try:
        dropin.install(tables)
        cls = tables.model_classes["CTC"]
        from funasr.models.ctc.model import Transformer as ref_ctc

        assert issubclass(cls, ref_ctc) and issubclass(cls, dropin.CTCInference)
        m = cls(encoder="ConformerEncoder", encoder_conf=dict(CONF_ENC, num_blocks=1), input_size=80, vocab_size=50)
        assert isinstance(m.encoder, dropin.ConformerEncoder)
    finally:
        for table, old in zip((tables.model_classes, tables.encoder_classes, tables.frontend_classes,
                               tables.predictor_classes, tables.decoder_classes), saved):
            table.clear()
            table.update(old)


@pytest.mark.gpu
def test_ctc_model_inference_override_matches_golden(conformer_sd, golden_dir):
    g = np.load(os.path.join(golden_dir, "conformer.npz"))
    m = dropin.CTC(encoder="ConformerEncoder", encoder_conf=CONF_ENC, input_size=80, vocab_size=4234)
    missing = m.load_state_dict(conformer_sd, strict=False)
    assert not missing.unexpected_keys and all(k.endswith("num_batches_tracked") for k in missing.missing_keys)
    m = m.cuda().eval()
    fe = dropin.WavFrontend(cmvn_file=None, window="hamming", lfr_m=1, lfr_n=1, dither=0.0)
    waves = [synth.synth_waveform(*(int(v) for v in g["case_%d" % ci])) for ci in (0, 1)]
    results, meta = m.inference(waves, key=["a", "b"], tokenizer=_CtcTokenizer(), frontend=fe, device="cuda")
    assert [r["key"] for r in results] == ["a", "b"]
    for ci, r in enumerate(results):
        assert r["token"] == ["t%d" % i for i in g["ids_%d" % ci].tolist()]
    assert meta["batch_data_time"] > 0


# ---------------------------------------------------------------------------------------- BiCifParaformer (8f rank 3)
PRED_V3 = dict(idim=512, threshold=1.0, l_order=1, r_order=1, tail_threshold=0.45, smooth_factor2=0.25, noise_threshold2=0.01,
               upsample_times=3, use_cif1_cnn=False, upsample_type="cnn_blstm")


def test_standalone_bicif_model_has_the_reference_state_dict():
    conf = dict(MODEL, predictor="CifPredictorV3", predictor_conf=PRED_V3, encoder_conf=dict(ENC, num_blocks=2),
                decoder_conf=dict(DEC, num_blocks=2, att_layer_num=2))
    model = dropin.BiCifParaformer(**conf)
    want = params.bicif_paraformer_shapes(enc_blocks=2, dec_blocks=2)
    assert {k: tuple(v.shape) for k, v in model.state_dict().items()} == dict(want)
    assert dropin.DROPIN_CLASSES[("model_classes", "BiCifParaformer")] is dropin.BiCifParaformer
    with pytest.raises(NotImplementedError):
        dropin.CifPredictorV3(**dict(PRED_V3, upsample_type="cnn"))
    frontend = dropin.WavFrontend(lfr_m=7, lfr_n=6, dither=0.0)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        model.inference([synth.synth_waveform(0, 16000)], key=["a"], tokenizer=None, frontend=frontend, device="cpu")


@pytest.mark.skipif(not reference_available(), reason="reference tree not present (GPU box)")
def test_install_wraps_the_reference_bicif_model_class():
    from ref_loader import import_reference

    import_reference()
    from funasr.models.bicif_paraformer.model import BiCifParaformer as ref_cls
    from funasr.register import tables

    saved = {t: dict(getattr(tables, t)) for t in ("model_classes", "predictor_classes", "encoder_classes", "decoder_classes",
                                                    "frontend_classes")}
    try:
        dropin.install(tables)
        cls = tables.model_classes["BiCifParaformer"]
        assert issubclass(cls, ref_cls) and issubclass(cls, dropin.BiCifParaformerInference)
        assert cls.inference is dropin.ParaformerInference.inference and cls._fb200_timestamps
        assert tables.predictor_classes["CifPredictorV3"] is dropin.CifPredictorV3
    finally:
        for t, d in saved.items():
            getattr(tables, t).clear()
            getattr(tables, t).update(d)


@pytest.mark.gpu
def test_bicif_model_inference_override_matches_golden(bicif_sd, golden_dir, tmp_path):
    """`BiCifParaformer.inference` (bicif_paraformer/model.py:222-369) through the packed engine: token ids and the
    time stamps of the reference's batch-1 runs, for a ragged batch in one call."""
    g = np.load(os.path.join(golden_dir, "bicif.npz"))
    dev = torch.device("cuda:0")
    shift, scale = synth.synth_cmvn()
    cmvn_file = str(tmp_path / "am.mvn")
    synth.write_cmvn_file(cmvn_file, shift, scale)
    frontend = dropin.WavFrontend(cmvn_file=cmvn_file, fs=16000, window="hamming", n_mels=80, frame_length=25,
                                  frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0)
    model = dropin.BiCifParaformer(**dict(MODEL, predictor="CifPredictorV3", predictor_conf=PRED_V3)).to(dev)
    model.load_state_dict(bicif_sd, strict=True)
    waves = [synth.synth_waveform(int(g["seeds"][ci]), int(g["num_samples"][ci])) for ci in range(3)]
    results, meta = model.inference(waves, key=["a", "b", "c"], tokenizer=None, frontend=frontend, device="cuda:0")
    assert [r["token_int"] for r in results] == [g["ids_%d" % ci].tolist() for ci in range(3)]
    tok = _Tokenizer()
    results, _ = model.inference(waves, key=["a", "b", "c"], tokenizer=tok, frontend=frontend, device="cuda:0")
    for ci, r in enumerate(results):
        assert set(r) == {"key", "text", "timestamp"}
        if not reference_available():  # without the reference's text post-processing the raw stamps come back
            assert r["timestamp"] == g["timestamp_%d" % ci].tolist()
    shifted, _ = model.inference([waves[0]], key=["a"], tokenizer=tok, frontend=frontend, device="cuda:0", begin_time=1000)
    if not reference_available():
        assert shifted[0]["timestamp"][0] == [g["timestamp_0"][0][0] + 1000, g["timestamp_0"][0][1] + 1000]


# ---------------------------------------------------------------------------------------- SeacoParaformer without hot words
def test_seaco_golden_equals_bicif_golden(golden_dir):
    """The pin behind the SeACo drop-in: the unmodified reference's SeacoParaformer without hot words gives BiCifParaformer's
    ids and time stamps on the same seeded weights (tests/golden/make_golden_seaco.py)."""
    s, b = np.load(os.path.join(golden_dir, "seaco.npz")), np.load(os.path.join(golden_dir, "bicif.npz"))
    for ci in range(3):
        assert np.array_equal(s["ids_%d" % ci], b["ids_%d" % ci])
        assert np.array_equal(s["timestamp_%d" % ci], b["timestamp_%d" % ci])
    assert all(k.startswith(("bias_encoder.", "seaco_decoder.", "hotword_output_layer.")) for k in s["extra_keys"].tolist())


SEACO = dict(inner_dim=512, bias_encoder_type="lstm", bias_encoder_bid=False, seaco_lsm_weight=0.1, seaco_length_normal=True,
             train_decoder=False, NO_BIAS=8377, seaco_decoder="ParaformerSANMDecoder",
             seaco_decoder_conf=dict(attention_heads=4, linear_units=1024, num_blocks=4, kernel_size=21, sanm_shfit=0,
                                     use_output_layer=False, wo_input_layer=True))


def test_standalone_seaco_model_has_the_reference_state_dict():
    conf = dict(MODEL, predictor="CifPredictorV3", predictor_conf=PRED_V3, encoder_conf=dict(ENC, num_blocks=1),
                decoder_conf=dict(DEC, num_blocks=1, att_layer_num=1), **SEACO)
    model = dropin.SeacoParaformer(**conf)
    want = params.seaco_paraformer_shapes(enc_blocks=1, dec_blocks=1)
    assert {k: tuple(v.shape) for k, v in model.state_dict().items()} == dict(want)
    frontend = dropin.WavFrontend(lfr_m=7, lfr_n=6, dither=0.0)
    with pytest.raises(RuntimeError, match="tokenizer"):  # a hot-word string cannot be looked up without the tokenizer
        model.inference([synth.synth_waveform(0, 16000)], key=["a"], frontend=frontend, hotword="some words")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        model.inference([synth.synth_waveform(0, 16000)], key=["a"], frontend=frontend, hotword=[[5 + i] for i in range(60)])


@pytest.mark.skipif(not reference_available(), reason="reference tree not present (GPU box)")
def test_install_wraps_the_reference_seaco_model_class():
    from ref_loader import import_reference

    import_reference()
    from funasr.models.seaco_paraformer.model import SeacoParaformer as ref_cls
    from funasr.register import tables

    saved = {t: dict(getattr(tables, t)) for t in ("model_classes", "predictor_classes", "encoder_classes", "decoder_classes",
                                                    "frontend_classes")}
    try:
        dropin.install(tables)
        cls = tables.model_classes["SeacoParaformer"]
        assert issubclass(cls, ref_cls) and issubclass(cls, dropin.SeacoParaformerInference)
        assert cls.inference is dropin.ParaformerInference.inference and cls._fb200_hotwords and cls._fb200_empty == ([],)
    finally:
        for t, d in saved.items():
            getattr(tables, t).clear()
            getattr(tables, t).update(d)


@pytest.mark.gpu
def test_seaco_model_inference_matches_golden(seaco_sd, golden_dir, tmp_path):
    """`SeacoParaformer.inference` through the packed engine, without and with hot words, == the unmodified reference."""
    g = np.load(os.path.join(golden_dir, "seaco.npz"))
    dev = torch.device("cuda:0")
    shift, scale = synth.synth_cmvn()
    cmvn_file = str(tmp_path / "am.mvn")
    synth.write_cmvn_file(cmvn_file, shift, scale)
    frontend = dropin.WavFrontend(cmvn_file=cmvn_file, fs=16000, window="hamming", n_mels=80, frame_length=25,
                                  frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0)
    model = dropin.SeacoParaformer(**dict(MODEL, predictor="CifPredictorV3", predictor_conf=PRED_V3, **SEACO)).to(dev)
    model.load_state_dict(seaco_sd, strict=True)
    waves = [synth.synth_waveform(int(g["seeds"][ci]), int(g["num_samples"][ci])) for ci in range(3)]
    results, _ = model.inference(waves, key=["a", "b", "c"], tokenizer=None, frontend=frontend, device="cuda:0")
    assert [r["token_int"] for r in results] == [g["ids_%d" % ci].tolist() for ci in range(3)]
    tok = _Tokenizer()
    tok.tokens2ids = lambda tokens: [tok.token_list.index(t) for t in tokens]
    results, _ = model.inference(waves, key=["a", "b", "c"], tokenizer=tok, frontend=frontend, device="cuda:0")
    if not reference_available():
        assert [r["timestamp"] for r in results] == [g["timestamp_%d" % ci].tolist() for ci in range(3)]
    # hot words as id lists (what the reference's generate_hotwords_list produces, without the trailing [sos])
    flat, lens, hot, o = g["hotwords_flat"].tolist(), g["hotwords_len"].tolist(), [], 0
    for n in lens[:-1]:
        hot.append(flat[o: o + n])
        o += n
    results, _ = model.inference(waves[:2], key=["a", "b"], tokenizer=None, frontend=frontend, device="cuda:0", hotword=hot)
    assert [r["token_int"] for r in results] == [g["hot_ids_%d" % ci].tolist() for ci in range(2)]
    # a hot-word string: one single-token hot word per blank-separated item (model.py:595-605)
    results, _ = model.inference([waves[0]], key=["a"], tokenizer=tok, frontend=frontend, device="cuda:0", hotword="t5 t77")
    assert results[0]["text"]
    assert model.inference([np.zeros(200, dtype=np.float32)], key=["z"], frontend=frontend, device="cuda:0") == ([],)
