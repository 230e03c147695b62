"""Device time of each stage of the BiCifParaformer time-stamp head for one packed batch (CUDA events).

    python tools/bicif_stage_profile.py [impl]      # impl 0 = tensor-core recurrence (default), 1 = fp32 SIMT check kernel
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from funasr_b200 import bicif, ops, params, synth, tables  # noqa: E402


def main():
    impl = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    dev = torch.device("cuda:0")
    sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.bicif_predictor_shapes(), seed=0).items()}
    tp = bicif.TimestampPack(sd, dev, nparts=2)
    for lens in ([500] * 37, [250] * 75, [100] * 189, [40] * 473):
        r = sum(lens)
        enc = torch.randn(r, 512, device=dev)
        cu = tables.cu_from_lengths(lens, torch.int32, dev)
        np_, d, k, h = 2, 512, 3, 512

        def timed(fn, n=3):
            fn()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(n):
                fn()
            b.record()
            torch.cuda.synchronize()
            return a.elapsed_time(b) / n

        xs = ops.split_rows(enc, np_)
        up = torch.empty(r, ops.split_width(k * h, np_), dtype=torch.bfloat16, device=dev)
        xg = torch.empty(r * k, 8 * h, dtype=torch.float32, device=dev)
        xg_rows = xg.view(r, k * 8 * h)
        part = torch.empty(tp.parts, r * k, dtype=torch.float32, device=dev)
        t_up = timed(lambda: ops.gemm(xs, np_, tp.up_w.w, np_, r, k * h, d, bias=tp.up_w.bias, out_split=up, out_parts=np_))

        def proj():
            for j in range(k):
                ops.gemm_ex(up, np_, tp.proj.w, np_, r, 8 * h, h, lda=np_ * k * h, a_part_stride=k * h, a_col0=j * h,
                            bias=tp.proj.bias, out_f32=xg_rows[:, j * 8 * h:(j + 1) * 8 * h])

        t_proj = timed(proj)
        t_lstm = timed(lambda: ops.lstm_bidir(xg, cu, k, k * max(lens), tp.w_hh, tp.w_out, part, hidden=h, impl=impl))
        steps = k * max(lens)
        chunks = (len(lens) + 63) // 64
        print("impl %d B=%d T=%d rows=%d: upsample GEMM %.3f ms, projections %.3f ms, lstm %.3f ms (%d steps x %d chunks: %.2f us per chunk-step)"
              % (impl, len(lens), lens[0], r, t_up, t_proj, t_lstm, steps, chunks, t_lstm * 1e3 / (steps * chunks)))


if __name__ == "__main__":
    main()
