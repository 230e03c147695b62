import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def paraformer_sd():
    """Seeded Paraformer-large weights as torch CPU tensors (220 M params, ~3 s to make)."""
    import torch

    from funasr_b200 import params, synth

    sd = synth.synth_state_dict(params.paraformer_large_shapes(), seed=0)
    return {k: torch.from_numpy(v) for k, v in sd.items()}


@pytest.fixture(scope="session")
def cmvn():
    import torch

    from funasr_b200 import synth

    shift, scale = synth.synth_cmvn(560, 3)
    return torch.from_numpy(shift), torch.from_numpy(scale)


@pytest.fixture(scope="session")
def sensevoice_sd():
    """Seeded SenseVoiceSmall weights (234 M params: 50 + 20 SAN-M blocks, V = 25055)."""
    import torch

    from funasr_b200 import params, synth

    sd = synth.synth_state_dict(params.sense_voice_small_shapes(), seed=0)
    return {k: torch.from_numpy(v) for k, v in sd.items()}


@pytest.fixture(scope="session")
def conformer_sd():
    """Seeded Conformer-CTC weights (34.6 M params: conv2d input layer, 12 blocks, V = 4234)."""
    import torch

    from funasr_b200 import params, synth

    sd = synth.synth_state_dict(params.conformer_ctc_shapes(), seed=0)
    return {k: torch.from_numpy(v) for k, v in sd.items()}


@pytest.fixture(scope="session")
def bicif_sd():
    """Seeded BiCifParaformer weights: Paraformer-large + CifPredictorV3 (upsampling conv, BLSTM, second alpha head)."""
    import torch

    from funasr_b200 import params, synth

    sd = synth.synth_state_dict(params.bicif_paraformer_shapes(), seed=0)
    return {k: torch.from_numpy(v) for k, v in sd.items()}


@pytest.fixture(scope="session")
def seaco_sd():
    """Seeded SeacoParaformer weights: BiCifParaformer + bias-encoder LSTM, SeACo decoder, hot-word output layer."""
    import torch

    from funasr_b200 import params, synth

    sd = synth.synth_state_dict(params.seaco_paraformer_shapes(), seed=0)
    return {k: torch.from_numpy(v) for k, v in sd.items()}
