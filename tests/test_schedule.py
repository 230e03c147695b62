"""Host-side sharding/batching logic and the transcript gather (gloo, world_size 2, CPU)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from funasr_b200 import schedule, synth, tables


def test_encoder_rows_matches_frontend_tables():
    for n in [0, 399, 400, 559, 560, 16000, 32491, 160000, 480000]:
        f = tables.num_frames(n)
        assert schedule.encoder_rows(n) == tables.num_lfr_rows(f)


def test_shards_partition_and_balance():
    lens = synth.synth_durations(1024).tolist()
    for world in (1, 2, 4, 8):
        parts = schedule.shard(lens, world)
        flat = sorted(i for p in parts for i in p)
        assert flat == list(range(1024))
        loads = [sum(schedule.cost(schedule.encoder_rows(lens[i])) for i in p) for p in parts]
        assert max(loads) / (sum(loads) / world) < 1.01  # LPT: within 1 % of perfect balance
        assert schedule.shard(lens, world) == parts  # deterministic: every rank derives the same plan


def test_shard_plan_is_the_greedy_lpt_and_cached():
    """schedule.shard keeps a heap of (load, rank) and caches the plan per length list; the result is the plain greedy
    longest-first assignment (least-loaded rank, ties by rank), and the capacity follows from the same plan."""
    lens = synth.synth_durations(300, seed=3).tolist() + [0, 399, 400, 16000, 16000]
    for world in (1, 2, 3, 8):
        rows = [schedule.encoder_rows(n) for n in lens]
        loads, want = [0.0] * world, [[] for _ in range(world)]
        for i in sorted(range(len(rows)), key=lambda i: (-rows[i], i)):
            r = min(range(world), key=lambda k: (loads[k], k))
            want[r].append(i)
            loads[r] += schedule.cost(rows[i])
        assert schedule.shard(lens, world) == want
        for stamped, per_tok in ((False, 1), (True, 3)):
            cap = max([1] + [1 + sum(3 + per_tok * (rows[i] + 1) for i in mine) for mine in want])
            assert schedule.payload_capacity(lens, world, stamped) == cap
    a = schedule.shard(lens, 2)
    a[0].append(-1)  # callers own the lists they get: the cached plan is not touched
    assert schedule.shard(lens, 2)[0][-1] != -1
    assert schedule.shard([], 2) == [[], []] and schedule.payload_capacity([], 2) == 1


def test_batches_respect_row_budget_and_drop_empty():
    lens = [399, 16000, 480000, 32000, 100, 160000]
    b = schedule.batches(lens, range(len(lens)), max_rows=600)
    flat = [i for x in b for i in x]
    assert sorted(flat) == [1, 2, 3, 5]
    for x in b:
        rows = sum(schedule.encoder_rows(lens[i]) for i in x)
        assert rows <= 600 or len(x) == 1
    assert flat[0] == 2  # longest first


class _FakeEngine:
    """Stands in for ParaformerEngine on CPU: 'token ids' are a function of the waveform only."""

    def recognize(self, waves):
        return [[int(w.shape[0]) % 97, int(abs(float(w[0])) * 1e6) % 89] for w in waves]


class _FakeStampEngine(_FakeEngine):
    """Stands in for BiCifParaformerEngine: (ids, stamps) per utterance."""

    def recognize(self, waves):
        ids = super().recognize(waves)
        return ids, [[[i, i + 10] for i in row] for row in ids]



class _ArrayEngine(_FakeEngine):
    """An engine that can hand its ids over as int32 arrays (ParaformerEngine.RETURNS_ARRAYS)."""

    RETURNS_ARRAYS = True

    def recognize(self, waves, as_arrays=False):
        ids = super().recognize(waves)
        return [np.asarray(r, np.int32) for r in ids] if as_arrays else ids


def test_recognize_sharded_one_rank_lists_and_arrays():
    """world_size 1 (no process group): lists by default even when the engine produced arrays, arrays on request."""
    lens = [16000, 700, 52000, 399]
    waves = [synth.synth_waveform(i, n) for i, n in enumerate(lens)]
    want = _FakeEngine().recognize(waves)
    got = schedule.recognize_sharded(_ArrayEngine(), waves, 0, 1)
    assert got == want and all(isinstance(r, list) and all(isinstance(t, int) for t in r) for r in got)
    arr = schedule.recognize_sharded(_ArrayEngine(), waves, 0, 1, as_arrays=True)
    assert all(isinstance(a, np.ndarray) and a.dtype == np.int32 for a in arr) and [a.tolist() for a in arr] == want
    arr2 = schedule.recognize_sharded(_FakeStampEngine(), waves, 0, 1, as_arrays=True)
    want2 = _FakeStampEngine().recognize(waves)
    assert [a.tolist() for a in arr2[0]] == want2[0] and [a.tolist() for a in arr2[1]] == want2[1]


def _worker(rank, world, port, lens, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    waves = [synth.synth_waveform(i, n) for i, n in enumerate(lens)]
    got = schedule.recognize_sharded(_FakeEngine(), waves, rank, world)
    got2 = schedule.recognize_sharded(_FakeStampEngine(), waves, rank, world)
    arr = schedule.recognize_sharded(_FakeEngine(), waves, rank, world, as_arrays=True)
    arr2 = schedule.recognize_sharded(_FakeStampEngine(), waves, rank, world, as_arrays=True)
    if rank == 0:
        assert all(isinstance(a, np.ndarray) and a.dtype == np.int32 for a in arr)
        assert [a.tolist() for a in arr] == got  # same ids, int32 arrays instead of lists
        assert [a.tolist() for a in arr2[0]] == got2[0] and [a.tolist() for a in arr2[1]] == got2[1]
        assert all(a.ndim == 2 and a.shape[1] == 2 for a in arr2[1])
        torch.save((got, got2), out_path)
    else:
        assert got is None and got2 is None and arr is None and arr2 is None
    dist.barrier()
    dist.destroy_process_group()


def test_recognize_sharded_two_ranks_gloo(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    lens = [16000, 24000, 700, 52000, 8000, 31000, 399]
    out_path = str(tmp_path / "out.pt")
    mp.spawn(_worker, args=(2, port, lens, out_path), nprocs=2, join=True)
    got, got2 = torch.load(out_path)
    waves = [synth.synth_waveform(i, n) for i, n in enumerate(lens)]
    assert got == _FakeEngine().recognize(waves)
    assert got2 == _FakeStampEngine().recognize(waves)  # (ids, stamps) of every utterance, in order, on rank 0


def test_transcript_payload_round_trip_and_capacity():
    """The gather's int32 payload: plain id lists and (ids, stamps) pairs survive encode/decode; the capacity computed
    from the plan alone bounds every rank's payload (tokens <= rows + 1 per utterance)."""
    local = {5: [3, 9, 4], 0: [], 7: [8404 - 1]}
    assert schedule._decode_items(schedule._encode_items(local)) == local
    paired = {2: ([5, 6], [[0, 120], [130, 300]]), 9: ([], [])}
    assert schedule._decode_items(schedule._encode_items(paired)) == paired
    arr = schedule._decode_items(schedule._encode_items(paired), as_arrays=True)
    assert arr[2][0].tolist() == [5, 6] and arr[2][1].tolist() == [[0, 120], [130, 300]] and arr[9][0].size == 0 and arr[9][1].shape == (0, 2)
    assert {k: v.tolist() for k, v in schedule._decode_items(schedule._encode_items(local), as_arrays=True).items()} == local
    lens = [16000, 24000, 700, 52000, 8000, 31000, 399]
    for world in (1, 2, 3):
        for stamped in (False, True):
            cap = schedule.payload_capacity(lens, world, stamped)
            for mine in schedule.shard(lens, world):
                full = {i: [1] * (schedule.encoder_rows(lens[i]) + 1) for i in mine}
                if stamped:
                    full = {i: (v, [[0, 1]] * len(v)) for i, v in full.items()}
                assert schedule._encode_items(full).size <= cap


def test_slice_segments_matches_reference_slicing():
    speech = np.arange(16000 * 3, dtype=np.float32)
    segs = [[0, 500], [480, 1999], [2500, 3500], [100, 100]]
    got = schedule.slice_segments(speech, segs)
    assert [len(x) for x in got] == [8000, 24304, 8000, 0]
    assert got[1][0] == 480 * 16 and got[2][-1] == len(speech) - 1
    ref_root = os.environ.get("FUNASR_REFERENCE_ROOT", "/root/reference")
    path = os.path.join(ref_root, "funasr", "utils", "vad_utils.py")
    if os.path.exists(path):  # the reference's own slicer (imports torch only)
        import importlib.util

        spec = importlib.util.spec_from_file_location("ref_vad_utils", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        want, want_len = mod.slice_padding_audio_samples(speech, len(speech), [(s, i) for i, s in enumerate(segs)])
        assert [len(x) for x in want] == [len(x) for x in got]
        assert all(np.array_equal(a, b) for a, b in zip(want, got))


def test_recognize_segments_keeps_segment_order():
    speech = synth.synth_waveform(3, 16000 * 4)
    segs = [[0, 900], [1000, 3900], [3000, 3200]]
    got = schedule.recognize_segments(_FakeEngine(), speech, segs)
    assert got == _FakeEngine().recognize(schedule.slice_segments(speech, segs))


def test_recognize_segments_with_stamps_shifts_by_segment_start():
    """Result combination of the long-audio pipeline for a time-stamp engine (auto_model.py:510-518), host logic only."""
    import numpy as np

    from funasr_b200 import schedule

    class FakeEngine:
        def recognize(self, waves):
            self.lens = [len(w) for w in waves]
            return [[10 + i, 20 + i] for i in range(len(waves))], [[[5, 50], [50, 90 + i]] for i in range(len(waves))]

    eng = FakeEngine()
    speech = np.zeros(16000 * 3, dtype=np.float32)
    ids, stamps = schedule.recognize_segments_with_stamps(eng, speech, [[100, 900], [1500, 4000]])
    assert eng.lens == [800 * 16, 16000 * 3 - 1500 * 16]  # the second segment is clipped to the recording
    assert ids == [10, 20, 11, 21]
    assert stamps == [[105, 150], [150, 190], [1505, 1550], [1550, 1591]]


def test_default_row_budget_is_whole_waves_and_ramp_splits_only_the_first_batch():
    """engine.DEFAULT_MAX_ROWS = whole waves of 148 x 128-row tiles below the offset limit; ParaformerEngine._ramp (host logic,
    no GPU needed) cuts the first of several batches at a third of the budget and leaves everything else alone."""
    from funasr_b200 import engine

    assert engine.DEFAULT_MAX_ROWS % (148 * 128) == 0 and engine.DEFAULT_MAX_ROWS <= engine.MAX_ROWS_LIMIT

    class _Planner:
        max_rows = 900
        _ramp = engine.ParaformerEngine._ramp

    rows = [500, 400, 300, 200, 100, 90, 80]
    one = [[0, 1]]
    assert _Planner()._ramp(one, rows) == one  # a single batch is never split
    got = _Planner()._ramp([[0, 1], [2, 3, 4, 5, 6]], rows)
    assert got == [[0], [1], [2, 3, 4, 5, 6]]  # 500 rows already exceed a third of 900: the head is one utterance
    got = _Planner()._ramp([[4, 5, 6], [0]], rows)
    assert got == [[4, 5, 6], [0]]  # the whole first batch fits into the third: unchanged
    got = _Planner()._ramp([[3, 4, 5, 6], [0, 1]], rows)
    assert got == [[3, 4], [5, 6], [0, 1]] and sorted(sum(got, [])) == [0, 1, 3, 4, 5, 6]
