"""TEST INFRASTRUCTURE -- drives the UNMODIFIED reference (`funasr.AutoModel`) for this path.

Where the reference comes from: `/root/reference` in the build container, `oracle/_ref/pyref` on the GPU box (an unmodified
copy made by `make -C oracle`, git-ignored; see tests/golden/ref_loader.py).  Only tests/, bench.py's reference arm /
cpu_baseline leg and tools/ import this module; nothing under funasr_b200/ does.

  build_paraformer(device="cpu")                 the reference's AutoModel for funasr/models/paraformer/template.yaml
                                                 (funasr/auto/auto_model.py:113-288) with seeded weights, dither 0, synthetic
                                                 CMVN and token list -- the model every golden fixture was made with
  build_paraformer(device="cuda", dropins=True)  the same call AFTER funasr_b200.dropin.install(): the reference's registry
                                                 (funasr/register.py:57-65) hands out the B200 classes
  token_ids(m, waves, batch_size)                AutoModel.generate(input=[...]) -> token-id lists (auto_model.py:301-376 ->
                                                 Paraformer.inference, funasr/models/paraformer/model.py:443-596)
"""
import os
import sys
import tempfile

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
_GOLDEN = os.path.join(_ROOT, "tests", "golden")
if _GOLDEN not in sys.path:
    sys.path.insert(0, _GOLDEN)
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

import ref_loader  # noqa: E402

WEIGHT_SEED = 0
_TABLE_NAMES = ("frontend_classes", "encoder_classes", "predictor_classes", "decoder_classes", "model_classes")


def available() -> bool:
    return ref_loader.reference_available()


def reference_root() -> str:
    return ref_loader.REFERENCE_ROOT


def _cmvn_file() -> str:
    from funasr_b200 import synth

    path = os.path.join(tempfile.mkdtemp(prefix="fb200_ref_"), "am.mvn")
    shift, scale = synth.synth_cmvn(560, 3)
    synth.write_cmvn_file(path, shift, scale)
    return path


def paraformer_config(enc_blocks=None, dec_blocks=None, template="funasr/models/paraformer/template.yaml"):
    import yaml

    from funasr_b200 import synth

    cfg = yaml.safe_load(open(os.path.join(reference_root(), template)))
    cfg["tokenizer_conf"]["token_list"] = synth.synth_token_list(8404)
    cfg["frontend_conf"]["dither"] = 0.0  # the default 1.0 adds noise (funasr/frontends/wav_frontend.py:94)
    cfg["frontend_conf"]["cmvn_file"] = _cmvn_file()
    cfg["specaug"] = None
    if enc_blocks is not None:
        cfg["encoder_conf"]["num_blocks"] = enc_blocks
    if dec_blocks is not None:
        cfg["decoder_conf"]["num_blocks"] = dec_blocks
        cfg["decoder_conf"]["att_layer_num"] = dec_blocks
    return cfg


class registry_snapshot:
    """Context manager: puts the reference's registry tables back as they were (dropin.install() is process-wide)."""

    def __enter__(self):
        ref_loader.import_reference()
        from funasr.register import tables

        self.tables = tables
        self.saved = {k: dict(getattr(tables, k)) for k in _TABLE_NAMES}
        return self

    def __exit__(self, *exc):
        for k, v in self.saved.items():
            getattr(self.tables, k).clear()
            getattr(self.tables, k).update(v)
        return False


def build_paraformer(device="cpu", ncpu=None, dropins=False, enc_blocks=None, dec_blocks=None, seed=WEIGHT_SEED, **extra):
    """The reference's AutoModel over the Paraformer-large template with seeded weights (synth.fill_module_ walks the
    parameters by NAME, so the reference model and the drop-in model get identical values)."""
    import torch

    ref_loader.import_reference()
    from funasr import AutoModel

    from funasr_b200 import synth

    if dropins:
        from funasr_b200 import dropin

        dropin.install()
    cfg = paraformer_config(enc_blocks, dec_blocks)
    cfg.update(extra)
    m = AutoModel(**cfg, device=device, ncpu=ncpu or (os.cpu_count() or 4), disable_update=True, disable_pbar=True)
    synth.fill_module_(m.model, seed)
    m.model.eval()
    if device == "cpu":  # build_model called torch.set_num_threads(ncpu) (auto_model.py:190)
        torch.set_num_threads(ncpu or (os.cpu_count() or 4))
    return m


def token_ids(m, waves, batch_size=1):
    """Token-id lists through the reference's public call: generate() -> inference() -> model.inference (tokenizer=None
    makes Paraformer.inference return `token_int`, model.py:589-590)."""
    res = m.generate(input=[np.asarray(w, dtype=np.float32) for w in waves], batch_size=batch_size, tokenizer=None)
    return [[int(t) for t in r.get("token_int", [])] for r in res]


# ---------------------------------------------------------------------------------------------------------------------
# Many utterances on the host cores: a pool of worker PROCESSES (spawned -- the parent may hold a CUDA context), each with
# its own copy of the model and `threads` intra-op threads.  Batch-1 semantics (what AutoModel enforces on CPU,
# auto_model.py:185-187); waveforms are regenerated from their seeds inside the workers.  This is how the reference itself
# uses a many-core host (one process per job slot, examples/aishell/paraformer/run.sh:154-193); a single process with all
# threads is slower at batch 1 because the per-utterance GEMMs are small.

_STATE = {}


def _init_worker(kind, threads):
    import torch

    torch.set_num_threads(threads)
    from funasr_b200 import params, synth

    _STATE.update(kind=kind, threads=threads)
    cmvn = tuple(torch.from_numpy(x) for x in synth.synth_cmvn())
    if kind == "reference":  # the unmodified reference's AutoModel
        m = build_paraformer("cpu", ncpu=threads)
        counts = []
        orig = m.model.calc_predictor  # observe the CIF token counts the reference computes (instance-level wrapper)

        def recording(*a, **k):
            res = orig(*a, **k)
            counts.append([int(v) for v in res[1].round().long().tolist()])
            return res

        m.model.calc_predictor = recording
        _STATE.update(model=m, counts=counts)
        return
    shapes = {"oracle": params.paraformer_large_shapes, "sensevoice": params.sense_voice_small_shapes,
              "conformer": params.conformer_ctc_shapes}[kind]()
    _STATE.update(sd={k: torch.from_numpy(v) for k, v in synth.synth_state_dict(shapes, seed=0).items()}, cmvn=cmvn)


def _work(specs):
    import torch

    from funasr_b200 import synth

    kind, out = _STATE["kind"], []
    for index, seed, n in specs:
        wave = synth.synth_waveform(seed, n)
        if kind == "reference":
            counts = _STATE["counts"]
            del counts[:]
            ids = token_ids(_STATE["model"], [wave], 1)[0]
            out.append((index, ids, counts[0][0] if counts else 0))
        elif kind == "oracle":  # the ports below are pinned to the reference by tests/test_oracle_*.py
            from oracle import paraformer_oracle as O

            taps = {}
            ids = O.recognize(torch.from_numpy(wave), _STATE["sd"], _STATE["cmvn"], taps=taps)
            out.append((index, [int(t) for t in ids], int(taps["token_num"])))
        elif kind == "sensevoice":
            from oracle import paraformer_oracle as O

            ids = O.sense_voice_recognize(torch.from_numpy(wave), _STATE["sd"], _STATE["cmvn"])
            out.append((index, [int(t) for t in ids], len(ids)))
        elif kind == "conformer":
            from oracle import conformer_oracle as C

            ids = C.recognize(torch.from_numpy(wave), _STATE["sd"])
            out.append((index, [int(t) for t in ids], len(ids)))
        else:
            raise ValueError(kind)
    return out


class CpuPool:
    """`procs` worker processes x `threads` threads (default: all cores, 4 threads per worker), models built once.
    kind: "reference" (the unmodified reference's AutoModel, batch 1), "oracle" (Paraformer port), "sensevoice",
    "conformer" (ports).  run(specs) with specs = [(index, waveform seed, samples)] -> {index: (token ids, token count)}."""

    def __init__(self, kind, procs=None, threads=None):
        import multiprocessing as mp

        cores = os.cpu_count() or 4
        self.threads = threads or min(4, cores)
        self.procs = procs or max(1, cores // self.threads)
        self.kind = kind
        self.pool = mp.get_context("spawn").Pool(self.procs, initializer=_init_worker, initargs=(kind, self.threads))
        self.pool.map(_work, [[] for _ in range(self.procs)])  # returns once every worker holds its model

    @property
    def cores(self):
        return self.procs * self.threads

    def run(self, specs):
        order = sorted(specs, key=lambda sp: -sp[2])  # longest first, dealt round-robin: equal work per worker
        chunks = [order[i::self.procs] for i in range(self.procs)]
        parts = self.pool.map(_work, [c for c in chunks if c], chunksize=1)
        return {index: (ids, count) for part in parts for index, ids, count in part}

    def close(self):
        self.pool.close()
        self.pool.join()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False


def pool_recognize(kind, specs, procs=None, threads=None):
    with CpuPool(kind, procs, threads) as pool:
        return pool.run(specs)
