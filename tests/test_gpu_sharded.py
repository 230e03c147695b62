"""The multi-GPU product path on real GPUs: `schedule.recognize_sharded` over NCCL with two ranks gives rank 0 exactly the
token ids one GPU computes for the same utterances (needs 2 visible GPUs: `gpurun --gpus 2`; skipped on a 1-GPU box, where
tests/test_schedule.py covers the same logic with gloo)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from funasr_b200 import params, schedule, synth

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")]
LENS = [32000, 80000, 47999, 123457, 16000, 64000, 100000, 52000, 24000, 300]


def _engine(rank):
    from funasr_b200 import engine

    shapes = params.paraformer_large_shapes(enc_blocks=3, dec_blocks=2)
    sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(shapes, seed=0).items()}
    cmvn = tuple(torch.from_numpy(x) for x in synth.synth_cmvn())
    return engine.ParaformerEngine(sd, "cuda:%d" % rank, "fp32", cmvn=cmvn, max_rows=700)


def _worker(rank, world, port, out_path):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    waves = [synth.synth_waveform(400 + i, n) for i, n in enumerate(LENS)]
    stats = {}
    eng = _engine(rank)
    got = schedule.recognize_sharded(eng, waves, rank, world, stats=stats)
    arr = schedule.recognize_sharded(eng, waves, rank, world, as_arrays=True)  # same ids as int32 arrays (what bench.py times)
    if rank == 0:
        assert [a.tolist() for a in arr] == got and all(a.dtype == np.int32 for a in arr)
        torch.save((got, stats), out_path)
    else:
        assert got is None and arr is None
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_nccl_equal_one_gpu(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out_path = str(tmp_path / "out.pt")
    mp.spawn(_worker, args=(2, port, out_path), nprocs=2, join=True)
    got, stats = torch.load(out_path)
    waves = [synth.synth_waveform(400 + i, n) for i, n in enumerate(LENS)]
    want = _engine(0).recognize(waves)
    assert got == want and got[-1] == [] and any(got)
    assert stats["gather_bytes"] == 2 * 4 * schedule.payload_capacity(LENS, 2)
