"""Per-kernel parity on the B200: every C-ABI entry point against the oracle
(oracle/paraformer_oracle.py, pinned to the reference) or a torch fp64/fp32 restatement of
the same op, on seeded inputs.  Integer/index outputs are compared bit-exactly."""
import numpy as np
import pytest
import torch

from funasr_b200 import ops, synth, tables
from oracle import paraformer_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def rel(a, b):
    a = torch.as_tensor(a).double().cpu()
    b = torch.as_tensor(b).double().cpu()
    if b.numel() == 0:
        return 0.0 if a.shape == b.shape else float("inf")
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def cu(lengths, dtype=torch.int32):
    return tables.cu_from_lengths(lengths, dtype, torch.device(DEV))


def unsplit(s, k, nparts):
    kp = ops.round_up(k)
    s = s.float()
    return sum(s[:, p * kp: p * kp + k] for p in range(nparts))


# ------------------------------------------------------------------ frontend
@pytest.mark.parametrize("window", ["hamming", "povey"])
def test_fbank_matches_oracle(window):
    lens = [400, 8000, 32491, 80000, 559]
    waves = [synth.synth_waveform(i, n) for i, n in enumerate(lens)]
    packed = torch.from_numpy(np.concatenate(waves)).to(DEV)
    frames = [tables.num_frames(n) for n in lens]
    out = ops.fbank(packed, cu(lens, torch.int64), cu(frames), sum(frames),
                    tables.feature_window(window, 400).to(DEV), tables.mel_filterbank(80, 512, 16000.0).to(DEV))
    o = 0
    for w, f in zip(waves, frames):
        ref = O.kaldi_fbank(torch.from_numpy(w) * 32768.0, window=window)
        assert ref.shape[0] == f
        # north_star tolerance: fbank within 1e-4 relative (to the utterance's max |log-mel|)
        assert rel(out[o: o + f], ref) < 1e-4
        o += f


def test_fbank_generic_fft_size_matches_oracle():
    """8 kHz: 200-sample frames, 256-point FFT -- the generic radix-2 kernel (the 512-point one is specialised)."""
    lens = [200, 4000, 9999]
    waves = [synth.synth_waveform(40 + i, n) for i, n in enumerate(lens)]
    packed = torch.from_numpy(np.concatenate(waves)).to(DEV)
    frames = [tables.num_frames(n, 200, 80) for n in lens]
    out = ops.fbank(packed, cu(lens, torch.int64), cu(frames), sum(frames), tables.feature_window("hamming", 200).to(DEV),
                    tables.mel_filterbank(80, 256, 8000.0).to(DEV), frame_shift=80, nfft=256)
    o = 0
    for w, f in zip(waves, frames):
        ref = O.kaldi_fbank(torch.from_numpy(w) * 32768.0, fs=8000)
        assert ref.shape[0] == f
        assert rel(out[o: o + f], ref) < 1e-4
        o += f


def test_lfr_cmvn_matches_oracle_bit_exact():
    frames = [1, 5, 6, 7, 201, 998]
    g = torch.Generator().manual_seed(0)
    fbs = [torch.randn(f, 80, generator=g) for f in frames]
    rows = [tables.num_lfr_rows(f) for f in frames]
    shift, scale = (torch.from_numpy(x) for x in synth.synth_cmvn())
    out = ops.lfr_cmvn(torch.cat(fbs).to(DEV), cu(frames), cu(rows), sum(rows), shift=shift.to(DEV), scale=scale.to(DEV))
    raw = ops.lfr_cmvn(torch.cat(fbs).to(DEV), cu(frames), cu(rows), sum(rows))
    o = 0
    for fb, t in zip(fbs, rows):
        assert torch.equal(raw[o: o + t].cpu(), O.apply_lfr(fb))
        assert rel(out[o: o + t], O.apply_cmvn(O.apply_lfr(fb), shift, scale)) < 1e-6
        o += t


def test_lfr_prologue_and_scale_add_pe():
    frames = [13, 998]
    rows = [tables.num_lfr_rows(f) for f in frames]
    g = torch.Generator().manual_seed(1)
    fbs = [torch.randn(f, 80, generator=g) for f in frames]
    inv = tables.pe_inverse_timescales(560).to(DEV)
    out = ops.lfr_cmvn(torch.cat(fbs).to(DEV), cu(frames), cu(rows), sum(rows), pe_inv=inv, pe_scale=512 ** 0.5)
    plain = ops.lfr_cmvn(torch.cat(fbs).to(DEV), cu(frames), cu(rows), sum(rows))
    out2 = ops.scale_add_pe(plain, cu(rows), inv, 512 ** 0.5)
    o = 0
    for fb, t in zip(fbs, rows):
        ref = O.apply_lfr(fb) * 512 ** 0.5 + O.sinusoidal_pe(t, 560)
        assert rel(out[o: o + t], ref) < 2e-6
        assert rel(out2[o: o + t], ref) < 2e-6
        o += t


# ------------------------------------------------------------------ row-wise
@pytest.mark.parametrize("dim,eps", [(512, 1e-12), (560, 1e-12), (2048, 1e-12), (512, 1e-5)])
@pytest.mark.parametrize("nparts", [1, 2, 3])
def test_layernorm(dim, eps, nparts):
    g = torch.Generator().manual_seed(dim)
    x = torch.randn(77, dim, generator=g) * 3 + 0.5
    w, b = torch.rand(dim, generator=g) + 0.5, torch.randn(dim, generator=g)
    ref = torch.nn.functional.layer_norm(x, (dim,), w, b, eps)
    out = torch.empty(77, dim, device=DEV)
    sp = torch.full((77, ops.split_width(dim, nparts)), 7.0, dtype=torch.bfloat16, device=DEV)
    ops.layernorm(x.to(DEV), w.to(DEV), b.to(DEV), eps, out_f32=out, out_split=sp, nparts=nparts)
    assert rel(out, ref) < 2e-6
    assert rel(unsplit(sp, dim, nparts), ref) < {1: 5e-3, 2: 2e-5, 3: 1e-6}[nparts]
    kp = ops.round_up(dim)
    if kp > dim:
        assert float(sp[:, dim:kp].float().abs().max()) == 0.0


def test_layernorm_halo_layout():
    lens = [3, 1, 5]
    dim, r, b = 512, 9, 3
    x = torch.randn(r, dim, generator=torch.Generator().manual_seed(3))
    w, bb = torch.ones(dim), torch.zeros(dim)
    sp = torch.full((r + b + 1, ops.split_width(dim, 3)), 7.0, dtype=torch.bfloat16, device=DEV)
    ops.layernorm(x.to(DEV), w.to(DEV), bb.to(DEV), 1e-12, out_split=sp, nparts=3, halo=True, cu_rows=cu(lens))
    got = unsplit(sp, dim, 3).cpu()
    ref = torch.nn.functional.layer_norm(x, (dim,), eps=1e-12)
    zero_rows, o, h = [], 0, 0
    for n in lens:
        zero_rows.append(h)
        assert rel(got[h + 1: h + 1 + n], ref[o: o + n]) < 1e-6
        o, h = o + n, h + n + 1
    zero_rows.append(h)
    assert float(sp[zero_rows].float().abs().max()) == 0.0


@pytest.mark.parametrize("with_res", [False, True])
def test_fsmn_matches_oracle(with_res):
    lens = [1, 4, 11, 33, 167]
    g = torch.Generator().manual_seed(5)
    v = torch.randn(sum(lens), 1536, generator=g)
    w = torch.randn(512, 1, 11, generator=g) * 0.3
    res = torch.randn(sum(lens), 512, generator=g)
    vd = v.to(DEV)
    out = torch.empty(sum(lens), 512, device=DEV)
    ops.fsmn(vd[:, 1024:], cu(lens), w.reshape(512, 11).to(DEV), out, residual=res.to(DEV) if with_res else None)
    o = 0
    for n in lens:
        ref = O.fsmn_memory(v[o: o + n, 1024:], w)
        if with_res:
            ref = ref + res[o: o + n]
        assert rel(out[o: o + n], ref) < 2e-6
        o += n


@pytest.mark.parametrize("dim,ks", [(512, 11), (320, 11), (1024, 11), (512, 21), (512, 5)])
def test_fsmn_streams_across_cta_and_sequence_boundaries(dim, ks):
    """64-row CTAs x 512-channel column blocks: sequences that start / end anywhere inside a CTA's rows, a
    sequence spanning many CTAs, single-row sequences at CTA edges; channel counts below / above one column block.
    Kernel 11 (Paraformer) and 21 (the SeACo decoder) stream; other sizes take the generic kernel."""
    lens = [63, 1, 1, 70, 3, 64, 129, 1, 500, 5, 187]
    g = torch.Generator().manual_seed(dim + ks)
    v = torch.randn(sum(lens), dim, generator=g)
    w = torch.randn(dim, 1, ks, generator=g) * 0.3
    res = torch.randn(sum(lens), dim, generator=g)
    out = torch.full((sum(lens), dim), float("nan"), device=DEV)
    ops.fsmn(v.to(DEV), cu(lens), w.reshape(dim, ks).to(DEV), out, residual=res.to(DEV))
    o = 0
    for n in lens:
        ref = O.fsmn_memory(v[o: o + n], w) + res[o: o + n]
        assert rel(out[o: o + n], ref) < 2e-6
        o += n


# ------------------------------------------------------------------ dense contractions
GEMM_SHAPES = [(167, 1536, 560), (300, 512, 512), (129, 2048, 512), (64, 512, 2048), (71, 8404, 512), (1, 512, 512),
               (1000, 1024, 512), (18944, 512, 512), (9600, 2048, 512), (5000, 8404, 512), (10000, 512, 2048)]


@pytest.mark.parametrize("m,n,k", GEMM_SHAPES)
@pytest.mark.parametrize("nparts", [1, 2, 3])
def test_gemm_tcgen05_matches_fp64(m, n, k, nparts):
    g = torch.Generator().manual_seed(m + n + k)
    a = torch.randn(m, k, generator=g)
    w = torch.randn(n, k, generator=g) / k ** 0.5
    bias = torch.randn(n, generator=g)
    res = torch.randn(m, n, generator=g)
    a_s, w_s = ops.split_rows(a.to(DEV), nparts), ops.split_rows(w.to(DEV), nparts)
    # what the kernel is asked to compute: the contraction of the (split) operands it is given
    a_eff, w_eff = unsplit(a_s, k, nparts).double().cpu(), unsplit(w_s, k, nparts).double().cpu()
    ref = torch.relu(a_eff @ w_eff.T + bias.double()) + res.double()
    outs = {}
    for impl in (0, 1):
        out = torch.empty(m, n, device=DEV)
        sp = torch.empty(m, ops.split_width(n, nparts), dtype=torch.bfloat16, device=DEV)
        ops.gemm(a_s, nparts, w_s, nparts, m, n, k, bias=bias.to(DEV), relu=True, res1=res.to(DEV), out_f32=out,
                 out_split=sp, out_parts=nparts, impl=impl)
        torch.cuda.synchronize()
        outs[impl] = out
        # dropped cross terms: none (1 part), 2^-16-relative (2 parts), 2^-24-relative (3 parts)
        # the tensor cores' fp32 accumulation is not round-to-nearest: ~1e-5 relative over long K loops
        assert rel(out, ref) < {1: 1e-5, 2: 4e-5, 3: 1e-5}[nparts], "impl %d" % impl
        assert rel(unsplit(sp, n, nparts), ref) < {1: 5e-3, 2: 5e-5, 3: 1e-5}[nparts]
    assert rel(outs[0], outs[1]) < 1e-5  # tcgen05 vs SIMT on identical operands and term lists
    # fp32-parity modes against the unsplit fp32 operands (what the reference's nn.Linear sees)
    if nparts >= 2:
        full = torch.relu(a.double() @ w.double().T + bias.double()) + res.double()
        assert rel(outs[0], full) < (4e-5 if nparts == 2 else 1e-5)


def test_gemm_three_taps_over_halo_layout():
    """The CIF k=3 convolution: tap j reads A row m + j."""
    lens = [5, 1, 40]
    r, b, d = sum(lens), len(lens), 512
    g = torch.Generator().manual_seed(9)
    x = torch.randn(r, d, generator=g)
    w = torch.randn(d, d, 3, generator=g) / (3 * d) ** 0.5
    bias = torch.randn(d, generator=g)
    halo = torch.zeros(r + b + 1, d)
    o = 0
    for i, n in enumerate(lens):
        halo[o + i + 1: o + i + 1 + n] = x[o: o + n]
        o += n
    a_s = ops.split_rows(halo.to(DEV), 3)
    w_s = ops.split_rows(w.permute(2, 0, 1).reshape(3 * d, d).contiguous().to(DEV), 3)
    out = torch.empty(r + b, d, device=DEV)
    ops.gemm(a_s, 3, w_s, 3, r + b - 1, d, d, bias=bias.to(DEV), relu=True, out_f32=out, taps=3, a_rows=r + b + 1)
    o = 0
    for i, n in enumerate(lens):
        xi = torch.nn.functional.pad(x[o: o + n].T.unsqueeze(0), (1, 1))
        ref = torch.relu(torch.nn.functional.conv1d(xi.double(), w.double(), bias.double())).squeeze(0).T
        # tcgen05 accumulates in fp32 without round-to-nearest: ~1e-5 relative over 3 x 6 x 512 / 16 steps
        assert rel(out[o + i: o + i + n], ref) < 2e-5
        o += n


# ------------------------------------------------------------------ attention
@pytest.mark.parametrize("lens_q,lens_kv", [([33, 167, 1, 70], None), ([14, 71, 3], [33, 167, 9]), ([500], None),
                                            ([129, 64, 65], [64, 128, 257])])
@pytest.mark.parametrize("nparts", [1, 2, 3])
def test_attention_matches_oracle(lens_q, lens_kv, nparts):
    """Operands are bf16 splits (as the projection GEMMs write them); the expected value is the
    oracle's attention on the fp32 values those splits stand for."""
    self_att = lens_kv is None
    lens_kv = lens_q if self_att else lens_kv
    g = torch.Generator().manual_seed(sum(lens_q))
    q = torch.randn(sum(lens_q), 512, generator=g)
    kv = torch.randn(sum(lens_kv), 1024, generator=g)
    # keys grow along time so that later key tiles raise the running maximum by far more than 2^8
    # (exercises the O rescale in TMEM), then shrink again (exercises the no-rescale path)
    ramp = 1.0 + 12.0 * torch.sin(torch.arange(sum(lens_kv)) / 40.0).abs()
    kv[:, :512] *= ramp[:, None]
    qs, kvs = ops.split_rows(q.to(DEV), nparts), ops.split_rows(kv.to(DEV), nparts)
    q_eff, kv_eff = unsplit(qs, 512, nparts).cpu(), unsplit(kvs, 1024, nparts).cpu()
    for impl in (0, 1):
        out = torch.empty(sum(lens_q), 512, device=DEV)
        sp = torch.empty(sum(lens_q), ops.split_width(512, nparts), dtype=torch.bfloat16, device=DEV)
        ops.attention(qs, 0, 512, kvs, 0, kvs, 512, 1024, nparts, cu(lens_q), cu(lens_kv), max(lens_q), 4, 128 ** -0.5,
                      out_f32=out, out_split=sp, impl=impl)
        torch.cuda.synchronize()
        oq = ok = 0
        # tensor-core path: P is rounded to the operand precision (bf16 / 16-bit mantissa) before P.V
        tol = {1: 1e-2, 2: 5e-5, 3: 5e-6}[nparts] if impl == 0 else 5e-6
        for nq, nk in zip(lens_q, lens_kv):
            ref = O.mha_core(q_eff[oq: oq + nq], kv_eff[ok: ok + nk, :512], kv_eff[ok: ok + nk, 512:], 4)
            assert rel(out[oq: oq + nq], ref) < tol, (impl, nq, nk)
            assert rel(unsplit(sp, 512, nparts)[oq: oq + nq], ref) < max(tol, {1: 5e-3, 2: 2e-5, 3: 1e-6}[nparts])
            oq, ok = oq + nq, ok + nk


# ------------------------------------------------------------------ CIF + decoding
def test_cif_scan_and_integrate_bit_exact_fires():
    lens = [33, 167, 1, 500, 2]
    g = torch.Generator().manual_seed(11)
    alphas = torch.rand(sum(lens), generator=g) * 0.9
    alphas[-2:] = 0.5  # exact-integer prefix sum: cs = 1.0 at the last frame, tail on top
    alphas[33:37] = torch.tensor([0.5, 0.5, 1.0, 0.0])  # fires on exact integers, zero alpha
    hidden = torch.randn(sum(lens), 512, generator=g)
    r, b = sum(lens), len(lens)
    tok = torch.empty(b, dtype=torch.int32, device=DEV)
    pos = torch.full((r + b,), -1, dtype=torch.int32, device=DEV)
    rem = torch.zeros(r + b, device=DEV)
    fires = torch.empty(r + b, device=DEV)
    ad = alphas.to(DEV)
    ops.cif_scan(ad, cu(lens), 0.45, tok, pos, rem, fires)
    tok_h = tok.cpu().tolist()
    embeds = torch.empty(sum(tok_h), 512, device=DEV)
    ops.cif_integrate(hidden.to(DEV), ad, cu(lens), cu(tok_h), sum(tok_h), 0.45, pos, rem, embeds)
    o = e = 0
    for i, n in enumerate(lens):
        at = torch.cat([alphas[o: o + n], torch.tensor([0.45])])
        ht = torch.cat([hidden[o: o + n], torch.zeros(1, 512)])
        ref_fires, fire_mask, _ = O.cif_fire(at)
        frames, _ = O.cif_integrate(ht, at)
        ref_tok = int(torch.floor(at.sum()).item())
        assert tok_h[i] == ref_tok
        assert torch.equal(fires[o + i: o + i + n + 1].cpu(), ref_fires)  # bit-exact
        ref_pos = torch.nonzero(fire_mask).flatten()
        assert pos[o + i: o + i + len(ref_pos)].cpu().tolist() == ref_pos.tolist()
        assert rel(embeds[e: e + ref_tok], frames[:ref_tok]) < 5e-5
        o, e = o + n, e + ref_tok


def test_cif_alpha_matches_oracle():
    lens = [4, 9]
    r, b = sum(lens), len(lens)
    g = torch.Generator().manual_seed(2)
    conv = torch.relu(torch.randn(r + b, 512, generator=g))
    w, bias = torch.randn(512, generator=g) * 0.05, torch.tensor([0.1])
    out = torch.empty(r, device=DEV)
    ops.cif_alpha(conv.to(DEV), cu(lens), r, w.to(DEV), bias.to(DEV), out)
    rows = [i for i in range(4)] + [5 + i for i in range(9)]
    ref = torch.sigmoid(conv[rows] @ w + bias)
    assert float((out.cpu() - ref).abs().max()) < 1e-6


def test_argmax_logprob():
    g = torch.Generator().manual_seed(4)
    logits = torch.randn(71, 8404, generator=g) * 3
    logits[5, 100] = logits[5, 7000] = 50.0  # tie -> first index, like torch.argmax
    ids = torch.empty(71, dtype=torch.int32, device=DEV)
    sc = torch.empty(71, device=DEV)
    ops.argmax_logprob(logits.to(DEV), 8404, ids, sc)
    lp = torch.log_softmax(logits, dim=-1)
    assert ids.cpu().tolist() == lp.argmax(-1).tolist()
    assert float((sc.cpu() - lp.max(-1)[0]).abs().max()) < 1e-5


@pytest.mark.parametrize("drop,collapse", [((0,), True), ((0, 1, 2), True), ((0,), False)])
def test_ctc_collapse(drop, collapse):
    lens = [1, 37, 64, 100]
    g = torch.Generator().manual_seed(8)
    ids = torch.randint(0, 5, (sum(lens),), generator=g, dtype=torch.int32)
    out = torch.empty_like(ids, device=DEV)
    n_out = torch.empty(len(lens), dtype=torch.int32, device=DEV)
    mask = sum(1 << d for d in drop)
    ops.ctc_collapse(ids.to(DEV), cu(lens), mask, collapse, out, n_out)
    o = 0
    for i, n in enumerate(lens):
        seq = ids[o: o + n]
        if collapse:
            seq = torch.unique_consecutive(seq)
        ref = [int(t) for t in seq if int(t) not in drop]
        assert out[o: o + int(n_out[i])].cpu().tolist() == ref
        o += n


@pytest.mark.parametrize("m,n,k", [(5000, 8404, 512), (71, 8404, 512), (4096, 25055, 512), (300, 4234, 256), (1, 100, 512)])
def test_gemm_argmax_equals_gemm_then_argmax(m, n, k):
    """fb200_gemm_argmax (projection fused with the greedy pick, logits never written) against fb200_gemm + fb200_argmax_logprob:
    ids identical -- ties included: duplicated weight rows give equal logits and the FIRST index must win -- scores to 1e-5."""
    g = torch.Generator().manual_seed(m + n)
    a = torch.randn(m, k, generator=g).to(DEV)
    w = (torch.randn(n, k, generator=g) * k ** -0.5)
    w[n // 2] = w[3]                      # a tie inside one row of logits ...
    w[n - 1] = w[n - 2]                   # ... and one in the last, partial column tile
    w = w.to(DEV)
    bias = torch.randn(n, generator=g).to(DEV)
    bias[n // 2] = bias[3]
    bias[n - 1] = bias[n - 2]
    bias[7] = float("-inf")               # a banned id (SenseVoice's ban_emo_unk)
    a_s, w_s = ops.split_rows(a, 2), ops.split_rows(w, 2)
    logits = torch.empty(m, (n + 3) // 4 * 4, dtype=torch.float32, device=DEV)
    ops.gemm(a_s, 2, w_s, 2, m, n, k, bias=bias, out_f32=logits)
    want_ids = torch.empty(m, dtype=torch.int32, device=DEV)
    want_lp = torch.empty(m, dtype=torch.float32, device=DEV)
    ops.argmax_logprob(logits, n, want_ids, want_lp)
    got_ids = torch.full((m,), -1, dtype=torch.int32, device=DEV)
    got_lp = torch.empty(m, dtype=torch.float32, device=DEV)
    ops.gemm_argmax(a_s, 2, w_s, 2, m, n, k, got_ids, got_lp, bias=bias)
    torch.cuda.synchronize()
    assert torch.equal(want_ids, got_ids)
    assert float((want_lp - got_lp).abs().max()) < 1e-5
    # rows whose maximum sits on a tie pick the first of the equal columns
    lg = logits[:, :n]
    tied = (lg[:, 3] == lg.max(dim=1).values) | (lg[:, n - 2] == lg.max(dim=1).values)
    if bool(tied.any()):
        assert bool(((got_ids[tied] == 3) | (got_ids[tied] == n - 2)).all())
