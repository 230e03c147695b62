"""BASELINE configs[0]: ONE 10 s utterance, host waveform in -> ids (and time stamps) out; wall-clock latency per call.

    python tools/latency_config1.py [paraformer|bicif|seaco]      # seaco: with 20 hot words
"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from funasr_b200 import _lib, engine, params, synth  # noqa: E402

model = sys.argv[1] if len(sys.argv) > 1 else "paraformer"
shift, scale = synth.synth_cmvn()
cmvn = (torch.from_numpy(shift), torch.from_numpy(scale))
kw = {}
if model == "paraformer":
    sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.paraformer_large_shapes(), seed=0).items()}
    eng = engine.ParaformerEngine(sd, "cuda:0", "fp32", cmvn=cmvn)
elif model == "bicif":
    from funasr_b200 import bicif

    sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.bicif_paraformer_shapes(), seed=0).items()}
    eng = bicif.BiCifParaformerEngine(sd, "cuda:0", "fp32", cmvn=cmvn)
else:
    from funasr_b200 import seaco

    sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.seaco_paraformer_shapes(), seed=0).items()}
    eng = seaco.SeacoParaformerEngine(sd, "cuda:0", "fp32", cmvn=cmvn)
    rng = np.random.default_rng(5)
    kw = dict(hotwords=[rng.integers(3, 8000, size=int(rng.integers(2, 5))).tolist() for _ in range(20)])
w = [torch.from_numpy(synth.synth_waveform(11, 160000)).pin_memory()]
for _ in range(5):
    eng.recognize(w, **kw)
torch.cuda.synchronize()
_lib.launch_count(reset=True)
t0 = time.perf_counter()
for _ in range(20):
    eng.recognize(w, **kw)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 20
print("config 1 (%s): one 10 s utterance: %.2f ms per call, %d launches, RTFx %.0f" % (model, dt * 1e3, _lib.launch_count() // 20, 10.0 / dt))
_lib.profile_read(reset=True)
_lib.profile_enable(True)
eng.recognize(w, **kw)
torch.cuda.synchronize()
p = _lib.profile_read(reset=True)
_lib.profile_enable(False)
print("sum of kernel times %.2f ms:" % sum(v["ms"] for v in p.values()), {k: round(v["ms"], 2) for k, v in p.items()})
