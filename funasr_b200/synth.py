"""Deterministic synthetic inputs: waveforms, CMVN statistics and random-init weights.

There is no network, so no checkpoint or corpus can be fetched.  Everything that both
sides of a parity check must agree on is derived from integer seeds with numpy's PCG64
(bit-identical on every platform), so the reference run that produced
`tests/golden/*.npz` in the build container and the CUDA run on the GPU box see the
same bytes without shipping 880 MB of weights.

  * waveforms: SURVEY.md section 8(d) -- `0.1 * randn(N)` float32, utterance i seeded 1000+i
  * durations: config 2 = 1024 utterances, U(2, 30) s from seed 7
  * CMVN:      Kaldi-nnet `am.mvn` text, shift ~ N(-8, 1), scale ~ U(0.1, 0.3), seed 3
               (format parsed by the reference `load_cmvn`, funasr/frontends/wav_frontend.py:15-38)
  * weights:   every tensor of a state_dict is filled from a generator seeded by
               (seed, crc32(parameter name)); scale follows torch's default init
               (U(-1/sqrt(fan_in), 1/sqrt(fan_in)) for Linear/Conv, LayerNorm ~ 1 / 0 with a
               small perturbation so the affine terms are exercised).
"""
import zlib
from typing import Dict, Iterable, List, Tuple

import numpy as np

SAMPLE_RATE = 16000


def synth_waveform(index: int, num_samples: int, amplitude: float = 0.1) -> np.ndarray:
    """float32 white noise in [-1, 1] scale, utterance `index` (seed 1000 + index)."""
    rng = np.random.default_rng(1000 + int(index))
    return (amplitude * rng.standard_normal(int(num_samples))).astype(np.float32)


def synth_durations(count: int = 1024, lo_s: float = 2.0, hi_s: float = 30.0, seed: int = 7) -> np.ndarray:
    """Number of samples of each utterance of BASELINE config 2 (whole samples)."""
    rng = np.random.default_rng(seed)
    dur = rng.uniform(lo_s, hi_s, size=count)
    return np.round(dur * SAMPLE_RATE).astype(np.int64)


def synth_cmvn(dim: int = 560, seed: int = 3) -> Tuple[np.ndarray, np.ndarray]:
    rng = np.random.default_rng(seed)
    shift = rng.normal(-8.0, 1.0, size=dim).astype(np.float32)
    scale = rng.uniform(0.1, 0.3, size=dim).astype(np.float32)
    return shift, scale


def write_cmvn_file(path: str, shift: np.ndarray, scale: np.ndarray) -> None:
    """Kaldi-nnet text layout the reference parser expects (it takes fields [3:-1] of the
    `<LearnRateCoef>` line that follows `<AddShift>` / `<Rescale>`)."""
    dim = len(shift)

    def row(v):
        return " ".join(repr(float(x)) for x in v)

    with open(path, "w", encoding="utf-8") as f:
        f.write("<Nnet>\n")
        f.write("<Splice> %d %d\n[ 0 ]\n" % (dim, dim))
        f.write("<AddShift> %d %d\n" % (dim, dim))
        f.write("<LearnRateCoef> 0 [ %s ]\n" % row(shift))
        f.write("<Rescale> %d %d\n" % (dim, dim))
        f.write("<LearnRateCoef> 0 [ %s ]\n" % row(scale))
        f.write("</Nnet>\n")


def _tensor_rng(seed: int, name: str) -> np.random.Generator:
    return np.random.default_rng([int(seed), zlib.crc32(name.encode("utf-8"))])


def synth_tensor(name: str, shape: Iterable[int], seed: int = 0) -> np.ndarray:
    """Value of parameter `name` (reference state_dict key) with the given shape."""
    shape = tuple(int(s) for s in shape)
    rng = _tensor_rng(seed, name)
    leaf = name.rsplit(".", 1)[-1]
    is_norm = ".norm" in name or "after_norm" in name or "tp_norm" in name
    if is_norm and len(shape) == 1:
        u = rng.uniform(-1.0, 1.0, size=shape)
        if leaf == "running_var":  # BatchNorm statistics of the Conformer conv module: positive
            return (1.0 + 0.5 * u).astype(np.float32)
        return (1.0 + 0.1 * u if leaf == "weight" else 0.1 * u).astype(np.float32)
    if "embed" in name and "embed.out" not in name and len(shape) == 2:
        return rng.standard_normal(size=shape).astype(np.float32)
    if leaf == "weight":
        fan_in = int(np.prod(shape[1:])) if len(shape) > 1 else shape[0]
    else:  # bias: torch uses the fan_in of the matching weight; approximated by 512
        fan_in = 512
    bound = 1.0 / np.sqrt(max(fan_in, 1))
    return rng.uniform(-bound, bound, size=shape).astype(np.float32)


def synth_state_dict(shapes: Dict[str, Tuple[int, ...]], seed: int = 0) -> Dict[str, np.ndarray]:
    return {k: synth_tensor(k, v, seed) for k, v in shapes.items()}


def fill_module_(module, seed: int = 0) -> None:
    """Overwrite every parameter of a torch module in place with its seeded value."""
    import torch

    with torch.no_grad():
        for name, p in module.state_dict().items():
            if not p.dtype.is_floating_point:
                continue
            p.copy_(torch.from_numpy(synth_tensor(name, p.shape, seed)))


def synth_token_list(vocab_size: int = 8404) -> List[str]:
    """`<blank>`, `<s>`, `</s>`, synthetic tokens, `<unk>`."""
    return ["<blank>", "<s>", "</s>"] + ["t%d" % i for i in range(vocab_size - 4)] + ["<unk>"]
