"""Generate `tests/golden/seaco.npz`: the UNMODIFIED reference SeacoParaformer (what the hub alias `paraformer-zh` resolves to,
funasr/download/name_maps_from_hub.py:3) run WITHOUT hot words on the BiCif golden utterances.

    python tests/golden/make_golden_seaco.py      # needs /root/reference; ~2 min on 8 threads

Without a `hotword` argument `_seaco_decode_with_ASF` (funasr/models/seaco_paraformer/model.py:233-327) returns the plain
decoder's log-softmax and the rest of `inference` is BiCifParaformer's: the fixture pins that claim (token ids and time stamps
for seeded weights; the tensors shared with BiCifParaformer get the same seeded values because `synth` keys values by
parameter name).  Second part: the same utterances WITH a hot-word file -- hot-word embeddings, the two best merged
log-probabilities per token, token ids and time stamps (fewer than `nfilter` = 50 hot words, so no ASF filtering).
"""
import os
import sys
import tempfile

import numpy as np
import torch
import yaml

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

from ref_loader import REFERENCE_ROOT, import_reference  # noqa: E402

from funasr_b200 import synth  # noqa: E402

CASES = ((200, 32000), (201, 16000 * 5 + 321), (202, 16000 * 11))


def main():
    torch.manual_seed(0)
    shift, scale = synth.synth_cmvn(560, 3)
    tmp = tempfile.mkdtemp()
    cmvn_path = os.path.join(tmp, "am.mvn")
    synth.write_cmvn_file(cmvn_path, shift, scale)
    import_reference()
    from funasr import AutoModel

    cfg = yaml.safe_load(open(os.path.join(REFERENCE_ROOT, "funasr/models/seaco_paraformer/template.yaml")))
    cfg["tokenizer_conf"]["token_list"] = synth.synth_token_list(8404)
    cfg["frontend_conf"]["dither"] = 0.0
    cfg["frontend_conf"]["cmvn_file"] = cmvn_path
    cfg["specaug"] = None
    m = AutoModel(**cfg, device="cpu", ncpu=8, disable_update=True, disable_pbar=True)
    synth.fill_module_(m.model, 0)
    m.model.eval()
    model, frontend, tokenizer = m.model, m.kwargs["frontend"], m.kwargs["tokenizer"]
    import funasr.models.seaco_paraformer.model as seaco_module

    seen = []
    original = seaco_module.ts_prediction_lfr6_standard

    def recording(*a, **k):
        tokens = list(a[2])
        r = original(*a, **k)
        seen.append((r[0], r[1], tokens))
        return r

    seaco_module.ts_prediction_lfr6_standard = recording
    extra = sorted(k for k in model.state_dict() if k.startswith(("bias_encoder.", "seaco_decoder.", "hotword_output_layer.")))
    out = {"num_samples": np.asarray([n for _, n in CASES], dtype=np.int64), "seeds": np.asarray([s for s, _ in CASES], dtype=np.int64),
           "extra_keys": np.asarray(extra)}
    with torch.no_grad():
        for ci, (seed, n) in enumerate(CASES):
            wave = synth.synth_waveform(seed, n)
            del seen[:]
            model.inference(data_in=[wave], key=["utt"], tokenizer=tokenizer, frontend=frontend, device="cpu")
            stamps = seen[0][1]
            ids = model.inference(data_in=[wave], key=["utt"], tokenizer=None, frontend=frontend, device="cpu")[0][0]["token_int"]
            out["ids_%d" % ci] = np.asarray(ids, dtype=np.int32)
            out["timestamp_%d" % ci] = np.asarray(stamps, dtype=np.int64).reshape(-1, 2)
            print(ci, n, "ids", len(ids), "stamps", len(stamps), stamps[:2])
    # ---- with hot words (a .txt file: one hot word per line, tokens separated by blanks, model.py:555-571) -------------
    from funasr_b200 import params

    want = params.seaco_paraformer_shapes()
    have = {k: tuple(v.shape) for k, v in model.state_dict().items()}
    assert have == {k: tuple(v) for k, v in want.items()}, sorted(set(have) ^ set(want))[:8]
    hot_path = os.path.join(tmp, "hotwords.txt")
    with open(hot_path, "w") as f:
        f.write("t10 t20 t30\nt5\nt100 t200\nt7 t8 t9 t10 t11\n")
    from funasr.utils.load_utils import extract_fbank

    with torch.no_grad():
        hw_list = model.generate_hotwords_list(hot_path, tokenizer=tokenizer, frontend=frontend)
        out["hotwords_flat"] = np.asarray([i for hw in hw_list for i in hw], dtype=np.int32)
        out["hotwords_len"] = np.asarray([len(hw) for hw in hw_list], dtype=np.int32)
        from funasr.models.transformer.utils.nets_utils import pad_list

        hw_pad = pad_list([torch.Tensor(i).long() for i in hw_list], 0)
        out["hw_embed"] = model._hotword_representation(hw_pad, torch.Tensor([len(i) for i in hw_list]).int()).numpy()
        for ci, (seed, n) in enumerate(CASES[:2]):
            wave = synth.synth_waveform(seed, n)
            speech, lens = extract_fbank([torch.from_numpy(wave)], data_type="sound", frontend=frontend)
            enc, enc_lens = model.encode(speech, lens)
            embeds, tok_len = model.calc_predictor(enc, enc_lens)[:2]
            tok_len = tok_len.round().long()
            pred = model._seaco_decode_with_ASF(enc, enc_lens, embeds, tok_len, hw_list=hw_list)
            top = pred[0, : tok_len[0]].topk(2, dim=-1)
            out["hot_top2_ids_%d" % ci] = top.indices.numpy().astype(np.int32)
            out["hot_top2_logp_%d" % ci] = top.values.numpy()
            del seen[:]
            model.inference(data_in=[wave], key=["utt"], tokenizer=tokenizer, frontend=frontend, device="cpu", hotword=hot_path)
            out["hot_timestamp_%d" % ci] = np.asarray(seen[0][1], dtype=np.int64).reshape(-1, 2)
            ids = tokenizer.tokens2ids(seen[0][2])  # without a tokenizer the reference cannot build the hot-word list (model.py:566)
            assert ids == [i for i in top.indices[:, 0].tolist() if i not in (0, 1, 2)]
            out["hot_ids_%d" % ci] = np.asarray(ids, dtype=np.int32)
            print("hot words", ci, "ids", len(ids), "differ from plain:", int((np.asarray(ids) != out["ids_%d" % ci][: len(ids)]).sum())
                  if len(ids) == len(out["ids_%d" % ci]) else "length", "NO_BIAS picks:", int((top.indices[:, 0] == 8377).sum()))
    # ---- the keep-the-decoder branch of `_merge_res`: with random weights no token's hot-word argmax is NO_BIAS, so the
    # NO_BIAS bias is raised by the median gap (about half of the tokens then keep the plain decoder's distribution)
    with torch.no_grad():
        wave = synth.synth_waveform(*CASES[0])
        speech, lens = extract_fbank([torch.from_numpy(wave)], data_type="sound", frontend=frontend)
        enc, enc_lens = model.encode(speech, lens)
        embeds, tok_len = model.calc_predictor(enc, enc_lens)[:2]
        tok_len = tok_len.round().long()
        _, hidden, _ = model.decoder(enc, enc_lens, embeds, tok_len, return_hidden=True, return_both=True)
        ctx = torch.from_numpy(out["hw_embed"]).unsqueeze(0)
        n_hw = torch.tensor([ctx.shape[1]], dtype=torch.int32)
        merged = model.seaco_decoder(ctx, n_hw, embeds, tok_len)[0] + model.seaco_decoder(ctx, n_hw, hidden, tok_len)[0]
        dha = model.hotword_output_layer(merged)[0, : tok_len[0]]
        shift = float((dha.max(dim=-1)[0] - dha[:, 8377]).median()) + 1e-3
        model.hotword_output_layer.bias.data[8377] += shift
        pred = model._seaco_decode_with_ASF(enc, enc_lens, embeds, tok_len, hw_list=hw_list)
        top = pred[0, : tok_len[0]].topk(2, dim=-1)
        out["mixed_bias_shift"] = np.float32(shift)
        out["mixed_top2_ids_0"] = top.indices.numpy().astype(np.int32)
        out["mixed_top2_logp_0"] = top.values.numpy()
        dha2 = model.hotword_output_layer(merged)[0, : tok_len[0]]
        out["mixed_keep_0"] = (dha2.argmax(dim=-1) == 8377).numpy()
        model.hotword_output_layer.bias.data[8377] -= shift
        print("mixed: shift %.4f, %d of %d tokens keep the decoder's distribution" % (shift, int(out["mixed_keep_0"].sum()), int(tok_len[0])))
    # ---- more than nfilter = 50 hot words: the attention-score filter (model.py:268-295) -----------------------------
    rng = np.random.default_rng(11)
    many_path = os.path.join(tmp, "many.txt")
    many = [["t%d" % int(v) for v in rng.integers(0, 8000, size=int(rng.integers(1, 4)))] for _ in range(64)]
    with open(many_path, "w") as f:
        f.write("\n".join(" ".join(w) for w in many) + "\n")
    with torch.no_grad():
        hw_many = model.generate_hotwords_list(many_path, tokenizer=tokenizer, frontend=frontend)
        out["many_flat"] = np.asarray([i for hw in hw_many for i in hw], dtype=np.int32)
        out["many_len"] = np.asarray([len(hw) for hw in hw_many], dtype=np.int32)
        hw_pad = pad_list([torch.Tensor(i).long() for i in hw_many], 0)
        selected = model._hotword_representation(hw_pad, torch.Tensor([len(i) for i in hw_many]).int())
        wave = synth.synth_waveform(*CASES[0])
        speech, lens = extract_fbank([torch.from_numpy(wave)], data_type="sound", frontend=frontend)
        enc, enc_lens = model.encode(speech, lens)
        embeds, tok_len = model.calc_predictor(enc, enc_lens)[:2]
        tok_len = tok_len.round().long()
        _, hidden, _ = model.decoder(enc, enc_lens, embeds, tok_len, return_hidden=True, return_both=True)
        n_hw = torch.tensor([selected.shape[0]], dtype=torch.int32)
        att = model.seaco_decoder.forward_asf6(selected.unsqueeze(0), n_hw, hidden, tok_len)
        scores = att[0].sum(0).sum(0)
        out["many_scores"] = scores.numpy()
        out["many_keep"] = np.asarray(torch.topk(scores, 50)[1].tolist() + [selected.shape[0] - 1], dtype=np.int32)
        pred = model._seaco_decode_with_ASF(enc, enc_lens, embeds, tok_len, hw_list=hw_many)
        top = pred[0, : tok_len[0]].topk(2, dim=-1)
        out["many_top2_ids_0"] = top.indices.numpy().astype(np.int32)
        out["many_top2_logp_0"] = top.values.numpy()
        srt = np.sort(out["many_scores"][:-1])[::-1]
        print("many hot words:", len(hw_many), "kept", len(out["many_keep"]), "score gap at the cut %.3e (scores ~ %.3f)" % (srt[49] - srt[50], srt[49]))
    bicif = np.load(os.path.join(HERE, "bicif.npz"))
    for ci in range(len(CASES)):
        assert np.array_equal(out["ids_%d" % ci], bicif["ids_%d" % ci]), "SeACo without hot words differs from BiCif (ids)"
        assert np.array_equal(out["timestamp_%d" % ci], bicif["timestamp_%d" % ci]), "SeACo without hot words differs from BiCif (stamps)"
    np.savez_compressed(os.path.join(HERE, "seaco.npz"), **out)
    print("seaco.npz: identical to bicif.npz on", len(CASES), "utterances;", len(extra), "hot-word tensors unused")


if __name__ == "__main__":
    main()
