"""BiCifParaformer time-stamp head on the GPU against the oracle and the reference's golden outputs
(tests/golden/bicif.npz): the cooperative LSTM kernel, the fire scan, the whole engine."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from funasr_b200 import bicif, ops, synth, tables
from oracle import bicif_oracle as B

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")
THR = 1.0 - 1e-4


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _lstm_weights(seed, h=512):
    g = torch.Generator().manual_seed(seed)
    bound = h ** -0.5
    u = lambda *s: (torch.rand(*s, generator=g) * 2 - 1) * bound
    return dict(w_ih=[u(4 * h, h), u(4 * h, h)], w_hh=[u(4 * h, h), u(4 * h, h)], b_ih=[u(4 * h), u(4 * h)],
                b_hh=[u(4 * h), u(4 * h)], w_out=u(2, h))


@pytest.mark.parametrize("impl,tol", [(0, 1e-4), (1, 2e-5)])
@pytest.mark.parametrize("lens", [[5, 33, 1, 20], [2] * 70 + [9, 4]])
def test_lstm_bidir_matches_oracle(lens, impl, tol):
    """Ragged batch (and one with more than 64 utterances = two staging chunks): LSTM output and the fused
    Linear(2H -> 1) partial sums against the oracle's explicit recurrence.  impl 0 = tensor cores with hi + mid bf16
    operands (16-bit mantissas through a recurrence: 1e-4), impl 1 = the fp32 check kernel."""
    h, up = 512, 3
    w = _lstm_weights(1)
    g = torch.Generator().manual_seed(2)
    rows3 = up * sum(lens)
    x = torch.randn(rows3, h, generator=g)
    xg = torch.cat([F.linear(x, w["w_ih"][d], w["b_ih"][d] + w["b_hh"][d]) for d in range(2)], dim=1).to(DEV)
    cu = tables.cu_from_lengths(lens, torch.int32, DEV)
    part = torch.zeros(ops.lstm_parts(h), rows3, device=DEV)
    out_h = torch.zeros(rows3, 2 * h, device=DEV)
    ops.lstm_bidir(xg, cu, up, up * max(lens), torch.stack(w["w_hh"]).to(DEV), w["w_out"].to(DEV), part, out_h=out_h,
                   impl=impl)
    torch.cuda.synchronize()
    o = 0
    want = []
    for n in lens:
        seg = x[o: o + up * n]
        want.append(torch.cat([B.lstm_direction(seg, w["w_ih"][d], w["w_hh"][d], w["b_ih"][d], w["b_hh"][d], bool(d))
                               for d in range(2)], dim=1))
        o += up * n
    want = torch.cat(want)
    assert rel(out_h, want) < tol
    logits = want @ w["w_out"].reshape(-1)
    assert rel(part.sum(0), logits) < tol


def test_us_fires_matches_oracle():
    """alphas2 / scaling / fires / second pass, including utterances where the fire count already equals chars + 1."""
    g = torch.Generator().manual_seed(4)
    up, lens = 3, [11, 40, 7, 90, 23]
    rows3 = up * sum(lens)
    logits = torch.randn(rows3, generator=g) * 2
    nparts, bias = 5, 0.3
    part = torch.randn(nparts, rows3, generator=g)
    part[0] += logits - bias - part.sum(0)
    token_num = [4, 13, 2, 30, 8]
    ids, kept = [], []
    for i, n in enumerate(token_num):
        row = torch.randint(3, 100, (n,), generator=g)
        if i % 2:
            row[0] = 2
            row[-1] = 0
        ids.append(row)
        kept.append(int(((row != 0) & (row != 1) & (row != 2)).sum()))
    cu = tables.cu_from_lengths(lens, torch.int32, DEV)
    cu_tok = tables.cu_from_lengths(token_num, torch.int32, DEV)
    us_alphas = torch.zeros(rows3, device=DEV)
    us_peaks = torch.zeros(rows3, device=DEV)
    fire_idx = torch.full((rows3,), -1, dtype=torch.int32, device=DEV)
    fire_count = torch.zeros(len(lens), dtype=torch.int32, device=DEV)
    num_chars = torch.zeros(len(lens), dtype=torch.int32, device=DEV)
    ops.us_fires(part.to(DEV), bias, 0.25, 0.01, THR, cu, up, torch.tensor(token_num, dtype=torch.int32, device=DEV),
                 torch.cat(ids).to(torch.int32).to(DEV), cu_tok, (0, 1, 2), us_alphas, fire_idx, fire_count, num_chars,
                 us_peaks=us_peaks)
    torch.cuda.synchronize()
    assert num_chars.cpu().tolist() == kept
    o = 0
    for i, n in enumerate(lens):
        a = torch.relu(torch.sigmoid(part[:, o: o + up * n].sum(0) + bias) * 0.25 - 0.01)
        a = a * (torch.tensor(float(token_num[i])) / a.sum())
        peaks = B.fires_wo_hidden(a, THR)
        got_a = us_alphas[o: o + up * n].cpu()
        assert rel(got_a, a) < 1e-5
        assert rel(us_peaks[o: o + up * n], peaks) < 1e-4
        # the fires of the device's own alphas, by the oracle's scan (a last-ulp difference in the alphas may move a
        # fire that sits exactly on the threshold, so the oracle is run on the device values)
        peaks = B.fires_wo_hidden(got_a, THR)
        idx = torch.where(peaks >= THR)[0]
        if len(idx) != kept[i] + 1 and kept[i] > 0:
            a2 = got_a / (got_a.sum() / (kept[i] + 1))
            idx = torch.where(B.fires_wo_hidden(a2, THR) >= THR)[0]
        c = int(fire_count[i])
        assert c == len(idx)
        assert fire_idx[o: o + c].cpu().tolist() == idx.tolist()
        o += up * n


@pytest.fixture(scope="module")
def eng(bicif_sd):
    shift, scale = synth.synth_cmvn()
    return bicif.BiCifParaformerEngine(bicif_sd, DEV, "fp32", cmvn=(torch.from_numpy(shift), torch.from_numpy(scale)),
                                       max_rows=2048)


def test_stages_match_reference_golden(golden_dir, eng):
    g = np.load(os.path.join(golden_dir, "bicif.npz"))
    for ci in range(3):
        wave = synth.synth_waveform(int(g["seeds"][ci]), int(g["num_samples"][ci]))
        taps = {}
        dev_wave, lens = eng.upload([wave])
        ids, logp, tok, rows, ts = eng.forward_packed_ts(dev_wave, lens, taps=taps)
        torch.cuda.synchronize()
        assert tok == [int(g["token_num_%d" % ci])]
        assert rel(taps["enc"], torch.from_numpy(g["enc_%d" % ci])) < 1e-3
        assert rel(taps["embeds"], torch.from_numpy(g["embeds_%d" % ci])) < 1e-3
        assert rel(taps["us_alphas"], torch.from_numpy(g["us_alphas_%d" % ci])) < 1e-3
        want_idx = np.where(g["us_peaks_%d" % ci] >= np.float32(THR))[0]
        got_peaks = taps["us_peaks"].cpu().numpy()
        assert np.array_equal(np.where(got_peaks >= np.float32(THR))[0], want_idx)


def test_ids_and_stamps_match_reference_golden(golden_dir, eng):
    """Ragged batch of the three golden utterances in ONE packed call: ids and time stamps equal the reference's
    batch-1 results."""
    g = np.load(os.path.join(golden_dir, "bicif.npz"))
    waves = [synth.synth_waveform(int(g["seeds"][ci]), int(g["num_samples"][ci])) for ci in range(3)]
    ids, stamps = eng.recognize(waves)
    for ci in range(3):
        assert ids[ci] == g["ids_%d" % ci].tolist()
        assert stamps[ci] == g["timestamp_%d" % ci].tolist()
    ids2, stamps2 = eng.recognize(waves, begin_time=500.0)
    assert ids2 == ids and stamps2[0][0] == [stamps[0][0][0] + 500, stamps[0][0][1] + 500]


def test_matches_oracle_on_other_utterances(eng, bicif_sd):
    shift, scale = synth.synth_cmvn()
    cmvn = (torch.from_numpy(shift), torch.from_numpy(scale))
    waves = [synth.synth_waveform(300 + i, n) for i, n in enumerate([16000 * 3 + 5, 16000 * 2])]
    ids, stamps = eng.recognize(waves)
    for w, i, s in zip(waves, ids, stamps):
        want_ids, want_stamps = B.recognize(torch.from_numpy(w), bicif_sd, cmvn)
        assert i == want_ids
        assert s == want_stamps


def test_too_short_and_ragged_inputs(eng, bicif_sd):
    """Utterances shorter than one frame give ([], []) and do not disturb their neighbours (same ids / stamps as alone)."""
    w = synth.synth_waveform(310, 16000 * 2 + 11)
    alone_ids, alone_stamps = eng.recognize([w])
    ids, stamps = eng.recognize([np.zeros(100, dtype=np.float32), w, synth.synth_waveform(311, 399)])
    assert ids[0] == [] and stamps[0] == [] and ids[2] == [] and stamps[2] == []
    assert ids[1] == alone_ids[0] and stamps[1] == alone_stamps[0]
    assert len(stamps[1]) == len(ids[1]) and all(b <= e for b, e in stamps[1])
    assert eng.recognize([]) == ([], [])


def test_kernels_stay_inside_their_buffers():
    """Guard bands around every output of the LSTM and fire kernels (compute-sanitizer is not available on the pool)."""
    h, up, lens = 512, 3, [7, 1, 12]
    w = _lstm_weights(3)
    g = torch.Generator().manual_seed(6)
    rows3, pad = up * sum(lens), 96
    xg = torch.randn(rows3, 8 * h, generator=g).to(DEV)
    cu = tables.cu_from_lengths(lens, torch.int32, DEV)
    parts = ops.lstm_parts(h)
    for impl in (0, 1):
        part = torch.full((parts + 2, rows3), 7.5, device=DEV)
        out_h = torch.full((rows3 + pad, 2 * h), 7.5, device=DEV)
        ops.lstm_bidir(xg, cu, up, up * max(lens), torch.stack(w["w_hh"]).to(DEV), w["w_out"].to(DEV), part[1:-1], out_h=out_h[:rows3],
                       impl=impl)
        torch.cuda.synchronize()
        assert bool((part[0] == 7.5).all()) and bool((part[-1] == 7.5).all())
        assert bool((out_h[rows3:] == 7.5).all()) and bool((out_h[:rows3] != 7.5).all())
        assert bool(torch.isfinite(part[1:-1]).all()) and bool((part[1:-1] != 7.5).all())
    token_num = torch.tensor([3, 1, 5], dtype=torch.int32, device=DEV)
    ids = torch.randint(3, 50, (9,), dtype=torch.int32, device=DEV)
    cu_tok = tables.cu_from_lengths([3, 1, 5], torch.int32, DEV)
    fbuf = torch.full((rows3 + 2 * pad,), 7.5, device=DEV)
    pbuf = torch.full((rows3 + 2 * pad,), 7.5, device=DEV)
    ibuf = torch.full((rows3 + 2 * pad,), -7, dtype=torch.int32, device=DEV)
    cbuf = torch.full((len(lens) + 2,), -7, dtype=torch.int32, device=DEV)
    nbuf = torch.full((len(lens) + 2,), -7, dtype=torch.int32, device=DEV)
    ops.us_fires(part[1:-1], 0.1, 0.25, 0.01, THR, cu, up, token_num, ids, cu_tok, (0, 1, 2), fbuf[pad: pad + rows3],
                 ibuf[pad: pad + rows3], cbuf[1:-1], nbuf[1:-1], us_peaks=pbuf[pad: pad + rows3])
    torch.cuda.synchronize()
    for buf, fill in ((fbuf, 7.5), (pbuf, 7.5), (ibuf, -7)):
        assert bool((buf[:pad] == fill).all()) and bool((buf[pad + rows3:] == fill).all())
    assert cbuf[0] == -7 and cbuf[-1] == -7 and nbuf[0] == -7 and nbuf[-1] == -7
    assert nbuf[1:-1].cpu().tolist() == [3, 1, 5]
    o = 0
    for i, n in enumerate(lens):  # fired frames are inside the utterance and increasing; the rest of the slot is untouched
        c = int(cbuf[1 + i])
        got = ibuf[pad + o: pad + o + up * n].cpu().tolist()
        assert 0 <= c <= up * n and all(0 <= a < b < up * n for a, b in zip(got[:c], got[1:c])) and all(v == -7 for v in got[c:])
        o += up * n
