import sys, time, torch
sys.path.insert(0, "/root/repo")
from funasr_b200 import engine, params, synth, _lib
sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.paraformer_large_shapes(), seed=0).items()}
shift, scale = synth.synth_cmvn()
eng = engine.ParaformerEngine(sd, "cuda:0", "fp32", cmvn=(torch.from_numpy(shift), torch.from_numpy(scale)))
w = [torch.from_numpy(synth.synth_waveform(11, 160000)).pin_memory()]
for _ in range(5): eng.recognize(w)
torch.cuda.synchronize()
_lib.launch_count(reset=True)
t0 = time.perf_counter()
for _ in range(20): eng.recognize(w)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 20
print("config 1: one 10 s utterance: %.2f ms per call, %d launches, RTFx %.0f" % (dt * 1e3, _lib.launch_count() // 20, 10.0 / dt))
_lib.profile_read(reset=True); _lib.profile_enable(True); eng.recognize(w); torch.cuda.synchronize()
p = _lib.profile_read(reset=True); _lib.profile_enable(False)
print("sum of kernel times %.2f ms:" % sum(v["ms"] for v in p.values()), {k: round(v["ms"], 2) for k, v in p.items()})
