"""Request-stack path vs per-op path over request sizes (one utterance of 2..30 s, batches of short utterances):
    FB200_STACK_ROWS=512 FB200_STACK_TOKENS=384 python tools/stack_capacity_sweep.py"""
import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from funasr_b200 import _lib, engine, params, synth
shift, scale = synth.synth_cmvn()
cmvn = (torch.from_numpy(shift), torch.from_numpy(scale))
sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.paraformer_large_shapes(), seed=0).items()}
eng = engine.ParaformerEngine(sd, "cuda:0", "fp32", cmvn=cmvn)
assert eng.small_path() is not None  # record the programs before the switch below turns the path off for the second engine
os.environ["FB200_STACK"] = "0"
eng0 = engine.ParaformerEngine(sd, "cuda:0", "fp32", cmvn=cmvn)
def wall(e, w, n=20):
    for _ in range(4): e.recognize(w)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): e.recognize(w)
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
for secs in (2, 5, 10, 15, 20, 30):
    w = [torch.from_numpy(synth.synth_waveform(11, int(secs * 16000))).pin_memory()]
    a, b = eng.recognize(w), eng0.recognize(w)
    print("%2d s: stack %.2f ms, per-op %.2f ms, ids equal %s, rows %d tokens %d" % (secs, wall(eng, w), wall(eng0, w), a == b, eng.frontend.plan([w[0].shape[0]])[1][0], eng.last_token_num[0]))
for nb in (2, 4, 8):
    w = [torch.from_numpy(synth.synth_waveform(30 + i, 3 * 16000)).pin_memory() for i in range(nb)]
    a, b = eng.recognize(w), eng0.recognize(w)
    print("%d x 3 s: stack %.2f ms, per-op %.2f ms, ids equal %s" % (nb, wall(eng, w), wall(eng0, w), a == b))
