"""Request-stack path (csrc/stack_sm100.cu): programs recorded from the per-op calls and run as ONE persistent
cooperative kernel must reproduce the per-op kernels -- bit for bit for GEMM / LayerNorm / FSMN (same sums, other tiles),
within the 16-bit-mantissa operand rounding for attention (bf16 hi / mid products on mma.sync vs the fp32 SIMT check kernel) -- and the whole engine must
return the reference's token ids for BASELINE configs[0] (tests/golden/stages_10s.npz) through it."""
import os

import numpy as np
import pytest
import torch

from funasr_b200 import engine, ops, params, synth, tables

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _rand(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(DEV)


def _split(x, nparts=2):
    return ops.split_rows(x.contiguous(), nparts)


@pytest.mark.parametrize("m_cap,m,n,k,relu,res,f32,split", [
    (256, 167, 1536, 512, False, False, True, True),    # qkv-like: split out + fp32 tail columns
    (256, 167, 512, 512, False, True, True, False),     # out-projection: residual, fp32 out
    (256, 256, 2048, 512, True, False, False, True),    # ffn w1: relu, split out
    (256, 71, 512, 2048, False, True, True, False),     # ffn w2: long K
    (192, 33, 8404, 512, False, False, True, False),    # vocabulary projection: n not a multiple of 64
    (256, 1, 512, 560, False, False, True, True),       # one row, K padded to 576
])
def test_stack_gemm_equals_per_op_kernel(m_cap, m, n, k, relu, res, f32, split):
    a = _rand(m_cap, k, seed=1)
    w = _rand(n, k, seed=2, scale=k ** -0.5)
    bias = _rand(n, seed=3)
    r1 = _rand(m_cap, n, seed=4) if res else None
    a_s, w_s = _split(a), _split(w)
    f32_col0 = 1024 if (f32 and split and n > 1024) else 0

    def outs():
        of = torch.zeros(m_cap, n - f32_col0, dtype=torch.float32, device=DEV) if f32 else None
        os_ = torch.zeros(m_cap, ops.split_width(n, 2), dtype=torch.bfloat16, device=DEV) if split else None
        return of, os_

    of0, os0 = outs()
    ops.gemm(a_s, 2, w_s, 2, m, n, k, bias=bias, relu=relu, res1=r1, out_f32=of0, out_split=os0, out_parts=2,
             f32_col0=f32_col0, a_rows=m_cap)
    of1, os1 = outs()
    prog = ops.StackProgram([m_cap])
    with prog:
        ops.gemm(a_s, 2, w_s, 2, m_cap, n, k, bias=bias, relu=relu, res1=r1, out_f32=of1, out_split=os1, out_parts=2,
                 f32_col0=f32_col0)
    assert prog.n_ops == 1 and prog.n_barriers == 0
    for _ in range(2):  # a program re-arms itself: two runs in a row
        prog.run([m])
    torch.cuda.synchronize()
    if f32:
        assert torch.equal(of0[:m], of1[:m])
    if split:
        assert torch.equal(os0[:m], os1[:m])
    prog.close()


def test_stack_layernorm_fsmn_equal_per_op_kernels():
    cap, bcap = 256, 8
    lens = [90, 1, 40, 36]
    r, b = sum(lens), len(lens)
    cu = tables.cu_from_lengths(lens, torch.int32, torch.device(DEV))
    cu_cap = torch.zeros(bcap + 1, dtype=torch.int32, device=DEV)
    cu_cap[: b + 1].copy_(cu)
    for dim in (512, 560, 2048):
        x = _rand(cap, dim, seed=dim)
        g, be = _rand(dim, seed=5), _rand(dim, seed=6)
        want_f = torch.zeros(r, dim, device=DEV)
        want_s = torch.zeros(r + b + 1, ops.split_width(dim, 2), dtype=torch.bfloat16, device=DEV)
        ops.layernorm(x[:r], g, be, 1e-12, out_f32=want_f, out_split=want_s, nparts=2, halo=True, cu_rows=cu)
        got_f = torch.zeros(cap, dim, device=DEV)
        got_s = torch.zeros(cap + bcap + 1, ops.split_width(dim, 2), dtype=torch.bfloat16, device=DEV)
        prog = ops.StackProgram([cap, bcap])
        with prog:
            ops.layernorm(x, g, be, 1e-12, out_f32=got_f, out_split=got_s, nparts=2, halo=True, cu_rows=cu_cap)
        prog.run([r, b])
        torch.cuda.synchronize()
        assert torch.equal(want_f, got_f[:r])
        assert torch.equal(want_s, got_s[: r + b + 1])
        prog.close()
    for ks in (11, 21):
        v, res = _rand(cap, 512, seed=7), _rand(cap, 512, seed=8)
        wt = _rand(512, ks, seed=9)
        want = torch.zeros(r, 512, device=DEV)
        ops.fsmn(v[:r], cu, wt, want, residual=res[:r])
        got = torch.zeros(cap, 512, device=DEV)
        prog = ops.StackProgram([cap, bcap])
        with prog:
            ops.fsmn(v, cu_cap, wt, got, residual=res)
        prog.run([r, b])
        torch.cuda.synchronize()
        assert torch.equal(want, got[:r])
        prog.close()


@pytest.mark.parametrize("q_lens,kv_lens", [
    ([167], [167]),                 # one 10 s utterance, self-attention
    ([71], [167]),                  # its decoder cross-attention
    ([5, 130, 1, 64], [40, 200, 3, 13]),  # ragged batch, more than one key tile, a one-row utterance
    ([256], [256]),                 # capacity: 7 query rows per item
    ([200, 56], [256, 1]),          # a one-key utterance; four full key tiles
])
def test_stack_attention_matches_simt_check_kernel(q_lens, kv_lens):
    qcap, kcap, bcap, d = 256, 257, 8, 512
    lq, lk, b = sum(q_lens), sum(kv_lens), len(q_lens)
    dev = torch.device(DEV)
    cu_q, cu_k = tables.cu_from_lengths(q_lens, torch.int32, dev), tables.cu_from_lengths(kv_lens, torch.int32, dev)
    cq = torch.zeros(bcap + 1, dtype=torch.int32, device=DEV)
    ck = torch.zeros(bcap + 1, dtype=torch.int32, device=DEV)
    cq[: b + 1].copy_(cu_q)
    ck[: b + 1].copy_(cu_k)
    q = _split(_rand(qcap, d, seed=11))
    kv = _split(_rand(kcap, 2 * d, seed=12))
    want = torch.zeros(lq, d, device=DEV)
    ops.attention(q[:lq], 0, d, kv[:lk], 0, kv[:lk], d, 2 * d, 2, cu_q, cu_k, max(q_lens), 4, 128 ** -0.5, out_f32=want, impl=1)
    got = torch.zeros(qcap, d, device=DEV)
    got_s = torch.zeros(qcap, ops.split_width(d, 2), dtype=torch.bfloat16, device=DEV)
    prog = ops.StackProgram([qcap, kcap, bcap])
    with prog:
        ops.attention(q, 0, d, kv, 0, kv, d, 2 * d, 2, cq, ck, qcap, 4, 128 ** -0.5, out_f32=got, out_split=got_s)
    prog.run([lq, lk, b])
    torch.cuda.synchronize()
    err = float((want - got[:lq]).abs().max() / want.abs().max())
    assert err < 3e-5, err  # hi.mid + mid.hi + hi.hi products on the tensor cores (mma.sync) against the fp32 check kernel
    hi, mid = got_s[:lq, :d].float(), got_s[:lq, d:].float()
    assert float((hi + mid - got[:lq]).abs().max() / got[:lq].abs().max()) < 2e-5
    prog.close()


def test_stack_barriers_follow_the_dependencies():
    """LN -> GEMM -> (FSMN, attention) -> GEMM: the FSMN and the attention of a layer share a phase."""
    sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.paraformer_large_shapes(enc_blocks=2, dec_blocks=1), seed=0).items()}
    enc = engine.EncoderPack(sd, DEV, nparts=2)
    cap, bcap = 256, 8
    x = torch.zeros(cap, enc.din, device=DEV)
    cu = torch.zeros(bcap + 1, dtype=torch.int32, device=DEV)
    out = torch.zeros(cap, enc.d, device=DEV)
    prog = ops.StackProgram([cap, bcap])
    with prog:
        enc.record(x, cu, out, None, None)
    # per layer: LN | QKV | FSMN + attention | out | LN | w1 | w2  = 8 ops, 7 phases; + after_norm
    assert prog.n_ops == 2 * 8 + 1
    assert prog.n_barriers == 2 * 7
    prog.close()


def test_stack_encoder_matches_per_op_path():
    sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.paraformer_large_shapes(enc_blocks=3, dec_blocks=1), seed=0).items()}
    enc = engine.EncoderPack(sd, DEV, nparts=2)
    cap, bcap = 256, 8
    lens = [100, 67, 3]
    r, b = sum(lens), len(lens)
    dev = torch.device(DEV)
    cu = tables.cu_from_lengths(lens, torch.int32, dev)
    x = _rand(cap, enc.din, seed=21)
    halo_want = torch.zeros(r + b + 1, ops.split_width(enc.d, 2), dtype=torch.bfloat16, device=DEV)
    want = enc.run(x[:r].clone(), cu, max(lens), out_halo=halo_want)
    cu_cap = torch.zeros(bcap + 1, dtype=torch.int32, device=DEV)
    cu_cap[: b + 1].copy_(cu)
    got = torch.zeros(cap, enc.d, device=DEV)
    halo = torch.zeros(cap + bcap + 1, ops.split_width(enc.d, 2), dtype=torch.bfloat16, device=DEV)
    split = torch.zeros(cap, ops.split_width(enc.d, 2), dtype=torch.bfloat16, device=DEV)
    prog = ops.StackProgram([cap, bcap])
    with prog:
        enc.record(x, cu_cap, got, halo, split)
    prog.run([r, b])
    torch.cuda.synchronize()
    err = float((want - got[:r]).abs().max() / want.abs().max())
    assert err < 1e-4, err  # the two attention kernels differ in the last bits; everything else is identical
    d = enc.d  # hi | mid parts: compare the values they stand for (a last-bit difference may move a hi part by one ulp)
    hw, hg = halo_want.float(), halo[: r + b + 1].float()
    assert float(((hw[:, :d] + hw[:, d:]) - (hg[:, :d] + hg[:, d:])).abs().max() / want.abs().max()) < 1e-4
    prog.close()


def test_engine_small_request_ids_equal_reference_golden(golden_dir, paraformer_sd, cmvn, monkeypatch):
    """BASELINE configs[0]: one 10 s utterance through ParaformerEngine.recognize -- the request-stack path (default) and
    the per-op path (FB200_STACK=0) both return the unmodified reference's token ids."""
    g = np.load(os.path.join(golden_dir, "stages_10s.npz"))
    want = [int(t) for t in g["token_int"] if int(t) not in (0, 1, 2)]
    waves = [synth.synth_waveform(11, 160000)]
    eng = engine.ParaformerEngine(paraformer_sd, DEV, "fp32", cmvn=cmvn)
    from funasr_b200 import _lib
    _lib.launch_count(reset=True)
    got = eng.recognize(waves)
    torch.cuda.synchronize()
    launches = _lib.launch_count()
    assert eng._small, "the small request did not take the request-stack path"
    assert launches < 40, launches
    monkeypatch.setenv("FB200_STACK", "0")
    eng0 = engine.ParaformerEngine(paraformer_sd, DEV, "fp32", cmvn=cmvn)
    ref = eng0.recognize(waves)
    assert eng0._small is False
    assert got == ref
    assert eng.last_token_num == [int(g["token_num"])]
    assert got[0] == want
    # a ragged small batch (3 utterances, 2 + 3 + 4 s) and a second request through the same programs
    batch = [synth.synth_waveform(20 + i, n) for i, n in enumerate((32000, 48000, 64000))]
    assert eng.recognize(batch) == eng0.recognize(batch)
    assert eng.recognize(waves) == ref
    # the capacity edge: one 30 s utterance (500 rows of the 512 recorded) and 8 utterances at once
    for req in ([synth.synth_waveform(40, 480000)], [synth.synth_waveform(50 + i, 16000 * (2 + i % 3)) for i in range(8)]):
        _lib.launch_count(reset=True)
        got = eng.recognize(req)
        assert _lib.launch_count() < 40
        assert got == eng0.recognize(req)
