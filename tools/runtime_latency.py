"""Single-utterance latency of the native C++ runtime (fb200_offline_infer, no Python between the kernels):
one 10 s utterance, host waveform in -> ids (and time stamps for a BiCifParaformer directory) out.

    python tools/runtime_latency.py [paraformer|bicif]
"""
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from funasr_b200 import params, runtime, synth  # noqa: E402

model = sys.argv[1] if len(sys.argv) > 1 else "paraformer"
shapes = params.paraformer_large_shapes() if model == "paraformer" else params.bicif_paraformer_shapes()
sd = synth.synth_state_dict(shapes, seed=0)
d = tempfile.mkdtemp()
runtime.export_model(sd, d, cmvn=synth.synth_cmvn())
rec = runtime.OfflineRecognizer(d)
w = [synth.synth_waveform(11, 160000)]
for _ in range(5):
    rec.recognize(w)
t0 = time.perf_counter()
for _ in range(20):
    ids = rec.recognize(w, with_stamps=True)
dt = (time.perf_counter() - t0) / 20
print("native runtime (%s): one 10 s utterance: %.2f ms per call, %d tokens, %d stamps, RTFx %.0f"
      % (model, dt * 1e3, len(ids[0][0]), len(ids[1][0]), 10.0 / dt))
rec.close()
