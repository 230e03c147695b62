"""`torch.ops.funasr_b200.*` (csrc/torch_shim.cpp: the TORCH_LIBRARY layer over the C-ABI): loads, carries
schemas the dispatcher can see, fails loudly on CPU tensors; on the GPU one SAN-M encoder layer driven through torch.ops
(LayerNorm -> QKV GEMM -> FSMN -> attention -> out projection -> LayerNorm -> FFN) is bit-identical to the ctypes route."""
import pytest
import torch

from funasr_b200 import build, ops, tables, torch_ops


@pytest.fixture(scope="module")
def T():
    build.build()
    return torch_ops.load()


def test_ops_are_registered_with_schemas(T):
    assert T.version() == 100
    for name in ("fbank", "lfr_cmvn", "layernorm", "split_rows", "fsmn", "gemm", "attention_plan", "attention", "cif_alpha",
                 "cif_scan", "cif_integrate", "argmax_logprob", "ctc_collapse"):
        schema = str(getattr(T, name).default._schema)
        assert schema.startswith("funasr_b200::" + name + "("), schema
    assert "Tensor(a!)" in str(T.fsmn.default._schema)  # outputs are declared as mutated arguments


def test_cpu_tensors_fail_loudly(T):
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        T.split_rows(torch.zeros(4, 64), torch.zeros(4, 64, dtype=torch.bfloat16), 1)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        T.layernorm(torch.zeros(4, 512), torch.ones(512), torch.zeros(512), 1e-12, None, None, 1, 0, None)


@pytest.mark.gpu
def test_encoder_layer_through_torch_ops_equals_ctypes_route(T):
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(5)
    lens, d, ff, heads, npt = [70, 33, 200, 129], 512, 2048, 4, 2
    rows = sum(lens)
    cu = tables.cu_from_lengths(lens, torch.int32, dev)
    rnd = lambda *s: torch.randn(*s, generator=g).to(dev)
    x = rnd(rows, d)
    w = {k: rnd(*s) * 0.04 for k, s in dict(qkv=(3 * d, d), out=(d, d), w1=(ff, d), w2=(d, ff)).items()}
    b = {k: rnd(s) * 0.1 for k, s in dict(qkv=3 * d, out=d, w1=ff, w2=d).items()}
    gam, bet, fs_w = 1 + 0.1 * rnd(d), 0.1 * rnd(d), rnd(d, 11) * 0.1
    ws = {k: ops.split_rows(v, npt) for k, v in w.items()}
    kp = lambda n: ops.split_width(n, npt)

    def layer(route):
        xs = torch.empty(rows, kp(d), dtype=torch.bfloat16, device=dev)
        qkv = torch.empty(rows, kp(3 * d), dtype=torch.bfloat16, device=dev)
        v = torch.empty(rows, d, device=dev)
        mem, x1, x2 = torch.empty(rows, d, device=dev), torch.empty(rows, d, device=dev), torch.empty(rows, d, device=dev)
        ctx = torch.empty(rows, kp(d), dtype=torch.bfloat16, device=dev)
        h = torch.empty(rows, kp(ff), dtype=torch.bfloat16, device=dev)
        if route == "torch":
            plan = torch.zeros(4 + 4 * (rows // 128 + len(lens)), dtype=torch.int32, device=dev)
            T.attention_plan(cu, cu, plan)
            T.layernorm(x, gam, bet, 1e-12, None, xs, npt, 0, None)
            T.gemm(xs, npt, ws["qkv"], npt, rows, 3 * d, d, 1, b["qkv"], 0, None, v, 2 * d, qkv, npt, 0)
            T.fsmn(v, cu, fs_w, x, mem)
            T.attention(qkv, 0, kp(3 * d) // npt, qkv, d, qkv, 2 * d, kp(3 * d) // npt, npt, cu, cu, plan, max(lens), heads,
                        d // heads, (d // heads) ** -0.5, None, ctx, npt, 0)
            T.gemm(ctx, npt, ws["out"], npt, rows, d, d, 1, b["out"], 0, mem, x1, 0, None, npt, 0)
            T.layernorm(x1, gam, bet, 1e-12, None, xs, npt, 0, None)
            T.gemm(xs, npt, ws["w1"], npt, rows, ff, d, 1, b["w1"], 1, None, None, 0, h, npt, 0)
            T.gemm(h, npt, ws["w2"], npt, rows, d, ff, 1, b["w2"], 0, x1, x2, 0, None, npt, 0)
        else:
            ops.layernorm(x, gam, bet, 1e-12, out_split=xs, nparts=npt)
            ops.gemm(xs, npt, ws["qkv"], npt, rows, 3 * d, d, bias=b["qkv"], out_f32=v, f32_col0=2 * d, out_split=qkv, out_parts=npt)
            ops.fsmn(v, cu, fs_w, mem, residual=x)
            ops.attention(qkv, 0, kp(3 * d) // npt, qkv, d, qkv, 2 * d, kp(3 * d) // npt, npt, cu, cu, max(lens), heads,
                          (d // heads) ** -0.5, out_split=ctx, out_parts=npt)
            ops.gemm(ctx, npt, ws["out"], npt, rows, d, d, bias=b["out"], res1=mem, out_f32=x1)
            ops.layernorm(x1, gam, bet, 1e-12, out_split=xs, nparts=npt)
            ops.gemm(xs, npt, ws["w1"], npt, rows, ff, d, bias=b["w1"], relu=True, out_split=h, out_parts=npt)
            ops.gemm(h, npt, ws["w2"], npt, rows, d, ff, bias=b["w2"], res1=x1, out_f32=x2)
        torch.cuda.synchronize()
        return x2

    a, c = layer("torch"), layer("ctypes")
    assert torch.isfinite(a).all() and torch.equal(a, c)
