"""Request-stack path (csrc/stack_sm100.cu) on BASELINE configs[0] (one 10 s utterance: 167 encoder rows, 71 tokens):
device time of the encoder and decoder programs (CUDA events around N runs), the per-op timeline of one run, and the
wall-clock latency of ParaformerEngine.recognize with and without the path.

    python tools/stack_bench.py [seconds=10]
"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from funasr_b200 import _lib, engine, params, synth  # noqa: E402

seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 10.0
shift, scale = synth.synth_cmvn()
cmvn = (torch.from_numpy(shift), torch.from_numpy(scale))
sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.paraformer_large_shapes(), seed=0).items()}
eng = engine.ParaformerEngine(sd, "cuda:0", "fp32", cmvn=cmvn)
w = [torch.from_numpy(synth.synth_waveform(11, int(seconds * 16000))).pin_memory()]
ids = eng.recognize(w)
sp = eng._small
r, l = eng.frontend.plan([w[0].shape[0]])[1][0], eng.last_token_num[0]
print("request: %.1f s, %d encoder rows, %d tokens; encoder program %d ops / %d barriers, decoder program %d ops / %d barriers"
      % (seconds, r, l, sp.enc_prog.n_ops, sp.enc_prog.n_barriers, sp.dec_prog.n_ops, sp.dec_prog.n_barriers))


def timed(fn, n=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


enc_ms = timed(lambda: sp.enc_prog.run([r, 1]))
dec_ms = timed(lambda: sp.dec_prog.run([l, r, 1]))
print("device time per run: encoder program %.3f ms, decoder program %.3f ms" % (enc_ms, dec_ms))
weights_mb = sum(v.numel() for k, v in sd.items() if k.startswith("encoder.") and v.dim() == 2) * 4 / 1e6
print("encoder weights as hi + mid bf16: %.0f MB -> %.0f GB/s streamed" % (weights_mb, weights_mb / enc_ms))

names = {0: "gemm", 1: "layernorm", 2: "fsmn", 3: "attention"}
for tag, prog, dims, per in (("encoder", sp.enc_prog, [r, 1], 8), ("decoder", sp.dec_prog, [l, r, 1], 0)):
    prog.profile(True)
    prog.run(dims)
    prog.run(dims)
    t = prog.profile_read()
    prog.profile(False)
    dur = [(k, b, st[7] - st[0]) for k, b, st in t]
    tot = t[-1][2][7] - t[0][2][0]
    by = {}
    for k, _, u in dur:
        by[k] = by.get(k, 0.0) + u
    print("%s timeline (CTA 0, SM clock %.0f MHz): %.1f us over %d ops: %s" % (tag, prog.sm_mhz, tot, len(t), ", ".join("%s %.1f" % (names[k], v) for k, v in sorted(by.items()))))
    lo = per * 10 if per else 0
    for k, b, st in t[lo: lo + (per or 14)]:
        rel = ["%6.2f" % (x - st[0]) if x >= 0 else "     -" for x in st[1:]]
        print("   %-9s%s  producer %s  stage0 %s  mma %s  last-cta %s  stored %s  body %s  barrier %s" % (names[k], "|" if b else "+", *rel))


def wall(e, n=30):
    for _ in range(5):
        e.recognize(w)
    torch.cuda.synchronize()
    _lib.launch_count(reset=True)
    t0 = time.perf_counter()
    for _ in range(n):
        e.recognize(w)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3, _lib.launch_count() // n


ms1, n1 = wall(eng)
os.environ["FB200_STACK"] = "0"
eng0 = engine.ParaformerEngine(sd, "cuda:0", "fp32", cmvn=cmvn)
assert eng0.recognize(w) == ids
ms0, n0 = wall(eng0)
print("ParaformerEngine.recognize, host waveform -> ids: request-stack path %.2f ms (%d launches), per-op path %.2f ms (%d launches)"
      % (ms1, n1, ms0, n0))
