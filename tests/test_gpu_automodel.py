"""The reference's own `AutoModel(...).generate(input=...)` (funasr/auto/auto_model.py:113-376) running THROUGH the drop-ins
on the B200, against the same unmodified `AutoModel` on the host CPU (batch 1, fp32) -- the two halves of the boundary
(registry + checkpoint compatibility; packed CUDA path) meeting in one call.

The reference tree is `/root/reference` in the build container and the unmodified copy under `oracle/_ref/pyref` on the GPU
box (`make -C oracle`; tests/golden/ref_loader.py picks whichever exists)."""
import numpy as np
import pytest
import torch

from funasr_b200 import dropin, synth
from oracle import reference_runner as R

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not R.available(), reason="reference tree not built (make -C oracle)")]

LENS = [32000, 52877, 80000, 123200, 160000, 14400, 240000, 47999]  # 0.9 s ... 15 s, ragged


@pytest.fixture(scope="module")
def waves():
    return [synth.synth_waveform(300 + i, n) for i, n in enumerate(LENS)]


@pytest.fixture(scope="module")
def reference_cpu(waves):
    """Texts and token ids of the unmodified reference on the CPU, one utterance per call (what AutoModel does on a CPU:
    auto_model.py:185-187)."""
    ref = R.build_paraformer("cpu")
    assert type(ref.model).__module__.startswith("funasr.models.paraformer")
    texts = [r["text"] for r in ref.generate(input=list(waves), batch_size=1)]
    ids = R.token_ids(ref, waves, 1)
    return texts, ids


def test_automodel_generate_with_dropins_on_gpu_matches_reference_cpu(waves, reference_cpu):
    want_text, want_ids = reference_cpu
    with R.registry_snapshot():
        ours = R.build_paraformer("cuda", dropins=True)
        # the registry handed out the B200 classes, inside the reference's own Paraformer / AutoModel objects
        from funasr.models.paraformer.model import Paraformer as RefParaformer

        assert isinstance(ours.model, RefParaformer) and isinstance(ours.model, dropin.ParaformerInference)
        assert type(ours.model.encoder) is dropin.SANMEncoder and type(ours.model.decoder) is dropin.ParaformerSANMDecoder
        assert type(ours.model.predictor) is dropin.CifPredictorV2 and type(ours.kwargs["frontend"]) is dropin.WavFrontend
        assert next(ours.model.parameters()).is_cuda
        res = ours.generate(input=list(waves), batch_size=64)  # ONE padded batch in the reference; one packed batch here
        assert [r["text"] for r in res] == want_text
        assert all(set(r) == {"key", "text"} for r in res)
        assert R.token_ids(ours, waves, 64) == want_ids
        # batch_size 3: the AutoModel loop cuts the list itself (auto_model.py:329-338); same results
        assert R.token_ids(ours, waves, 3) == want_ids
        # single ndarray input, the documented call
        one = ours.generate(input=waves[2], tokenizer=None)
        assert one[0]["token_int"] == want_ids[2]


def test_padded_module_contracts_inside_the_reference_inference(waves, reference_cpu):
    """The reference's OWN `Paraformer.inference` (model.py:443-596: extract_fbank -> encode -> calc_predictor ->
    cal_decoder_with_predictor -> per-utterance greedy loop) over the drop-in frontend / encoder / predictor / decoder
    MODULES -- only the model class stays the reference's, so its padded call sequence runs on the CUDA kernels."""
    _, want_ids = reference_cpu
    with R.registry_snapshot() as snap:
        ref_model_cls = snap.tables.model_classes["Paraformer"]
        dropin.install()
        snap.tables.model_classes["Paraformer"] = ref_model_cls  # modules replaced, `inference` NOT replaced
        m = R.build_paraformer("cuda")
        assert type(m.model) is ref_model_cls and type(m.model.encoder) is dropin.SANMEncoder
        for i in (0, 3, 5):  # batch 1: the reference's padded batches leak padding into the CIF conv
            assert R.token_ids(m, [waves[i]], 1)[0] == want_ids[i]
