"""Parity of the CUDA Conformer-CTC path (BASELINE config 4) on the B200 against (a) the committed outputs of the
UNMODIFIED reference (tests/golden/conformer.npz) and (b) the oracle on fresh seeded inputs, stage by stage.

Tolerances: encoder hidden within 1e-3 max relative in fp32 mode (BASELINE.json), per-frame argmax and CTC token ids
identical; bf16 mode: encoder within 5e-2."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from funasr_b200 import conformer, ops, synth, tables
from oracle import conformer_oracle as C
from oracle import paraformer_oracle as P

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def rel(a, b):
    a = torch.as_tensor(a).double().cpu()
    b = torch.as_tensor(b).double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.fixture(scope="module")
def eng(conformer_sd):
    return conformer.ConformerCTCEngine(conformer_sd, DEV, "fp32")


def test_gemm_ex_taps_offsets_and_swish_match_simt():
    """Row / column tap offsets, A column window, part stride and the swish epilogue: tcgen05 vs the SIMT check
    kernel over the same operands, and vs torch on the fp32 values the split stands for."""
    g = torch.Generator().manual_seed(5)
    rows, kfull, k, n, m, np_ = 700, 192, 64, 200, 640, 2
    a = torch.randn(rows, kfull, generator=g)
    w = torch.randn(3, n, k, generator=g) * 0.2
    bias = torch.randn(n, generator=g)
    a_s = ops.split_rows(a.to(DEV), np_)            # [rows, 2 * 192]
    w_s = ops.split_rows(w.reshape(3 * n, k).to(DEV), np_)
    tap_row, tap_col = [0, 37, -5], [64, 0, 128]
    outs = []
    for impl in (0, 1):
        o = torch.zeros(m, n, device=DEV)
        ops.gemm_ex(a_s, np_, w_s, np_, m, n, k, lda=np_ * kfull, a_part_stride=kfull, a_col0=0, tap_row=tap_row,
                    tap_col=tap_col, taps=3, bias=bias.to(DEV), act="swish", out_f32=o, impl=impl)
        outs.append(o.cpu())
    want = torch.zeros(m, n)
    for j in range(3):
        src = torch.zeros(m, k)
        for r in range(m):
            rr = r + tap_row[j]
            if 0 <= rr < rows:
                src[r] = a[rr, tap_col[j]: tap_col[j] + k]
        want += src @ w[j].T
    want = want + bias
    want = want * torch.sigmoid(want)
    assert rel(outs[0], outs[1]) < 1e-5
    assert rel(outs[0], want) < 1e-4


@pytest.mark.parametrize("m,n_group", [(640, 300), (130, 61), (1500, 770)])
def test_gemm_heads_matches_per_head_products(m, n_group):
    """One grouped launch = the per-head (q+v).p^T products: tcgen05 vs the SIMT check kernel vs torch."""
    g = torch.Generator().manual_seed(9)
    heads, dk, np_, wrows = 4, 64, 2, 1024
    a = torch.randn(m, 4 * heads * dk, generator=g)            # the fused q+u | q+v | k | v layout
    pos = torch.randn(heads, wrows, dk, generator=g) * 0.3
    a_s = ops.split_rows(a.to(DEV), np_)
    w_s = torch.cat([ops.split_rows(pos[h].to(DEV), np_) for h in range(heads)], dim=0)
    pitch = (n_group + 255) // 256 * 256
    outs = []
    for impl in (0, 1):
        o = torch.zeros(m, heads * pitch, device=DEV)
        ops.gemm_heads(a_s, np_, w_s, wrows, m, n_group, heads, dk, o, lda=np_ * a.shape[1], a_part_stride=a.shape[1],
                       a_col0=heads * dk, a_col_step=dk, impl=impl)
        outs.append(o.view(m, heads, pitch)[:, :, :n_group].cpu())
    want = torch.stack([a[:, heads * dk + h * dk: heads * dk + (h + 1) * dk] @ pos[h, :n_group].T for h in range(heads)], dim=1)
    assert rel(outs[0], outs[1]) < 1e-5
    assert rel(outs[0], want) < 1e-4


@pytest.mark.parametrize("frames", [[7, 8, 11, 250], [401, 30]])
def test_conv2d_subsampling_matches_oracle(eng, conformer_sd, frames):
    """conv 1 (parity planes) -> conv 2 (9-tap contraction) -> Linear (19-tap contraction) -> gather, ragged batch
    including the shortest legal input (7 frames -> 1 row)."""
    g = torch.Generator().manual_seed(11)
    feats = [torch.randn(f, 80, generator=g) * 3 + 10 for f in frames]
    packed = torch.cat(feats).to(DEV)
    cu_frames = tables.cu_from_lengths(frames, torch.int32, DEV)
    x, cu_rows, rows = eng.encoder.embed(packed, cu_frames, frames)
    torch.cuda.synchronize()
    o = 0
    for f, r in zip(feats, rows):
        want = C.conv2d_subsampling(f, conformer_sd) * 16.0
        assert want.shape[0] == r
        assert rel(x[o: o + r], want) < 1e-4  # operands carry 16 mantissa bits (hi + mid), K = 4864
        o += r


def test_too_short_utterance_raises(eng):
    packed = torch.zeros(6 + 40, 80, device=DEV)
    with pytest.raises(RuntimeError):
        eng.encoder.embed(packed, tables.cu_from_lengths([6, 40], torch.int32, DEV), [6, 40])


def test_glu_dwconv_matches_torch(conformer_sd):
    p = "encoder.encoders.3.conv_module."
    sd = conformer_sd
    lens = [1, 5, 33, 100]
    g = torch.Generator().manual_seed(2)
    xs = [torch.randn(n, 512, generator=g) for n in lens]
    w = sd[p + "depthwise_conv.weight"]
    scale = sd[p + "norm.weight"] / torch.sqrt(sd[p + "norm.running_var"] + 1e-5)
    shift = sd[p + "norm.bias"] - sd[p + "norm.running_mean"] * scale
    out = torch.zeros(sum(lens), 512, dtype=torch.bfloat16, device=DEV)
    ops.glu_dwconv(torch.cat(xs).to(DEV), tables.cu_from_lengths(lens, torch.int32, DEV), max(lens), 256,
                   w.reshape(256, 15).contiguous().to(DEV), sd[p + "depthwise_conv.bias"].to(DEV), scale.to(DEV), shift.to(DEV),
                   out, 2)
    got = (out[:, :256].float() + out[:, 256:].float()).cpu()
    o = 0
    for x in xs:
        y = F.glu(x.t().unsqueeze(0), dim=1)
        y = F.conv1d(y, w, sd[p + "depthwise_conv.bias"], padding=7, groups=256)
        y = F.batch_norm(y, sd[p + "norm.running_mean"], sd[p + "norm.running_var"], sd[p + "norm.weight"], sd[p + "norm.bias"],
                         training=False, eps=1e-5)
        y = C.swish(y).squeeze(0).t()
        assert rel(got[o: o + x.shape[0]], y) < 3e-5
        o += x.shape[0]


@pytest.mark.parametrize("impl", [0, 1])
def test_relpos_attention_matches_oracle(impl):
    """Streaming softmax with the closed-form rel_shift vs the pad/view/slice transcription, ragged batch
    (lengths around the 64-row tile edges, T = 1 and 2 where the shift degenerates)."""
    heads, dk, np_ = 4, 64, 2
    d = heads * dk
    lens = [1, 2, 63, 64, 65, 130, 300]
    g = torch.Generator().manual_seed(9)
    r, max_t = sum(lens), max(lens)
    qu, qv, k, v = (torch.randn(r, d, generator=g) for _ in range(4))
    pos = torch.randn(max_t, d, generator=g)
    qkv = ops.split_rows(torch.cat([qu, qv, k, v], dim=1).to(DEV), np_)
    ld_bd = (max_t + 3) // 4 * 4
    bd = torch.zeros(heads, r, ld_bd, device=DEV)
    for h in range(heads):
        ph = ops.split_rows(pos[:, h * dk:(h + 1) * dk].contiguous().to(DEV), np_)
        ops.gemm_ex(qkv, np_, ph, np_, r, max_t, dk, lda=np_ * 4 * d, a_part_stride=4 * d, a_col0=d + h * dk, out_f32=bd[h])
    out = torch.zeros(r, d, device=DEV)
    ops.attention_relpos(qkv, 0, 4 * d, qkv, 2 * d, qkv, 3 * d, 4 * d, np_, bd, tables.cu_from_lengths(lens, torch.int32, DEV),
                         max_t, heads, dk ** -0.5, out_f32=out, impl=impl)
    osp = torch.zeros(r, 2 * d, dtype=torch.bfloat16, device=DEV)
    ops.attention_relpos(qkv, 0, 4 * d, qkv, 2 * d, qkv, 3 * d, 4 * d, np_, bd, tables.cu_from_lengths(lens, torch.int32, DEV),
                         max_t, heads, dk ** -0.5, out_split=osp, impl=impl)
    assert rel(osp[:, :d].float() + osp[:, d:].float(), out) < 2e-5
    out = out.cpu()
    o = 0
    for t in lens:
        sl = slice(o, o + t)
        hv = lambda x: x[sl].view(t, heads, dk).transpose(0, 1)  # noqa: E731
        ac = hv(qu) @ hv(k).transpose(-2, -1)
        bdm = C.rel_shift_legacy(hv(qv) @ pos[:t].view(t, heads, dk).transpose(0, 1).transpose(-2, -1))
        attn = torch.softmax((ac + bdm) / dk ** 0.5, dim=-1)
        want = (attn @ hv(v)).transpose(0, 1).reshape(t, d)
        assert rel(out[sl], want) < 2e-4, t
        o += t


@pytest.mark.parametrize("ci", [0, 1])
def test_stages_match_reference_golden_fp32(golden_dir, eng, ci):
    g = np.load(os.path.join(golden_dir, "conformer.npz"))
    ui, n = (int(v) for v in g["case_%d" % ci])
    wave, lens = eng.upload([synth.synth_waveform(ui, n)])
    taps = {}
    ids, counts, cu_rows, rows = eng.forward_packed(wave, lens, taps=taps)
    torch.cuda.synchronize()
    assert taps["enc"].shape == g["enc_%d" % ci].shape
    assert rel(taps["embed"], g["embed_%d" % ci]) < 1e-4
    assert rel(taps["layers"][0], g["block0_%d" % ci]) < 1e-3
    assert rel(taps["layers"][11], g["block11_%d" % ci]) < 1e-3
    assert rel(taps["enc"], g["enc_%d" % ci]) < 1e-3
    assert np.array_equal(taps["frame_ids"].cpu().numpy(), g["argmax_%d" % ci])
    assert ids.cpu().numpy()[: int(counts[0])].tolist() == g["ids_%d" % ci].tolist()


def test_ragged_batch_matches_oracle_fp32(eng, conformer_sd):
    """Six utterances of different lengths in ONE packed batch == the oracle's batch-1 runs (ids and encoder)."""
    sizes = [(40, 5200), (41, 16000), (42, 33333), (43, 48000), (44, 70001), (45, 160000)]
    waves = [synth.synth_waveform(ui, n) for ui, n in sizes]
    got = eng.recognize(waves)
    wave, lens = eng.upload(waves)
    taps = {}
    eng.forward_packed(wave, lens, taps=taps)
    torch.cuda.synchronize()
    cu = taps["cu_rows"].cpu().tolist()
    for j, w in enumerate(waves):
        with torch.no_grad():
            feats = P.wav_frontend(torch.from_numpy(w), None, window="hamming", lfr_m=1, lfr_n=1)
            enc = C.conformer_encoder(feats, conformer_sd)
            want = P.ctc_greedy(C.ctc_logits(enc, conformer_sd), blank=0, drop=(0, 1, 2))
        assert rel(taps["enc"][cu[j]: cu[j + 1]], enc) < 1e-3, j
        assert got[j] == want, j


def test_too_short_waveforms_give_empty_results(eng):
    got = eng.recognize([synth.synth_waveform(50, 300), synth.synth_waveform(51, 1200), synth.synth_waveform(52, 16000)])
    assert got[0] == [] and got[1] == [] and isinstance(got[2], list)


def test_bf16_mode_within_stated_tolerance(golden_dir, conformer_sd):
    g = np.load(os.path.join(golden_dir, "conformer.npz"))
    e16 = conformer.ConformerCTCEngine(conformer_sd, DEV, "bf16")
    ui, n = (int(v) for v in g["case_1"])
    wave, lens = e16.upload([synth.synth_waveform(ui, n)])
    taps = {}
    e16.forward_packed(wave, lens, taps=taps)
    torch.cuda.synchronize()
    assert rel(taps["enc"], g["enc_1"]) < 5e-2


def test_gemm_heads_and_glu_dwconv_stay_inside_their_buffers(conformer_sd):
    """Guard rows around the outputs of the grouped GEMM and the depthwise kernel (ragged batch, last tile partial)."""
    g = torch.Generator().manual_seed(11)
    heads, dk, np_, wrows, m, n_group = 4, 64, 2, 1024, 333, 300
    a_s = ops.split_rows(torch.randn(m, 4 * heads * dk, generator=g).to(DEV), np_)
    w_s = torch.cat([ops.split_rows((torch.randn(wrows, dk, generator=g) * 0.3).to(DEV), np_) for _ in range(heads)], dim=0)
    pitch = (n_group + 255) // 256 * 256
    out = torch.full((m + 64, heads * pitch), 7.5, device=DEV)
    ops.gemm_heads(a_s, np_, w_s, wrows, m, n_group, heads, dk, out[32: 32 + m], lda=np_ * 4 * heads * dk,
                   a_part_stride=4 * heads * dk, a_col0=heads * dk, a_col_step=dk)
    torch.cuda.synchronize()
    assert bool((out[:32] == 7.5).all()) and bool((out[32 + m:] == 7.5).all())
    c, k, lens = 256, 15, [70, 3, 131]
    r = sum(lens)
    x = torch.randn(r, 2 * c, generator=g).to(DEV)
    cu = tables.cu_from_lengths(lens, torch.int32, DEV)
    dw = (torch.randn(c, k, generator=g) * 0.2).to(DEV)
    vec = lambda: torch.randn(c, generator=g).to(DEV)
    width = ops.split_width(c, np_)
    buf = torch.full((r + 64, width), 3.0, dtype=torch.bfloat16, device=DEV)
    ops.glu_dwconv(x, cu, max(lens), c, dw, vec(), vec().abs() + 0.5, vec(), buf[32: 32 + r], np_)
    torch.cuda.synchronize()
    assert bool((buf[:32] == 3.0).all()) and bool((buf[32 + r:] == 3.0).all())
    assert bool(torch.isfinite(buf[32: 32 + r].float()).all())
