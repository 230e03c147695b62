"""Sharding and batching of utterances across the GPUs of one box (pure data parallel).

Every rank computes the SAME plan from the list of utterance lengths -- no communication -- and
works on its shard; rank 0 gathers (utterance index, token ids) at the end.  The reference's own
analogues are the `wav.scp` split into one process per GPU (examples/aishell/paraformer/run.sh:154-193)
and the duration-sorted dynamic batching of funasr/auto/auto_model.py:423-459.
"""
import functools
import heapq
from typing import List, Sequence

FRAME_LEN = 400
FRAME_SHIFT = 160
LFR_N = 6


def encoder_rows(num_samples: int) -> int:
    """T = ceil(F / 6), F = 1 + (N - 400) // 160 (snip_edges), 0 if shorter than one frame."""
    if num_samples < FRAME_LEN:
        return 0
    frames = 1 + (num_samples - FRAME_LEN) // FRAME_SHIFT
    return (frames + LFR_N - 1) // LFR_N


def cost(rows: int) -> float:
    """Algorithmic FLOPs of one utterance: encoder linear + attention + decoder
    (token count taken as 0.41 T, the random-init probe's rate)."""
    t = float(rows)
    l = 0.41 * t
    enc = 314.7e6 * t + 102.4e3 * t * t
    dec = 16 * (4 * l * 512 * 2048 + 4 * l * 512 * 512 + 2 * t * 512 * 1024 + 4 * l * t * 512) + 2 * l * 512 * 8404
    return enc + dec


@functools.lru_cache(maxsize=16)
def _plan(lens: tuple, world_size: int):
    """(shards, payload capacity without / with stamps) of one length list.  Cached: a serving loop (and `bench.py`) shards
    the same set again and again, and 1024 utterances cost ~6 ms of Python per call before this -- 10 % of an 8-GPU step."""
    rows = [encoder_rows(int(n)) for n in lens]
    order = sorted(range(len(rows)), key=lambda i: (-rows[i], i))
    heap = [(0.0, k) for k in range(world_size)]  # (load, rank): the least-loaded rank, ties by rank
    out = [[] for _ in range(world_size)]
    for i in order:
        load, r = heap[0]
        out[r].append(i)
        heapq.heapreplace(heap, (load + cost(rows[i]), r))
    caps = []
    for per_tok in (1, 3):
        caps.append(max([1] + [1 + sum(3 + per_tok * (rows[i] + 1) for i in mine) for mine in out]))
    return tuple(tuple(m) for m in out), caps[0], caps[1]


def shard(num_samples: Sequence[int], world_size: int) -> List[List[int]]:
    """Longest-processing-time-first assignment: indices per rank, deterministic (ties by index)."""
    return [list(m) for m in _plan(tuple(int(n) for n in num_samples), int(world_size))[0]]


def batches(num_samples: Sequence[int], indices: Sequence[int], max_rows: int, max_utts: int = 0) -> List[List[int]]:
    """Indices (into `num_samples`) grouped longest-first under the packed-row budget (and at most `max_utts`
    utterances per batch when given: 1024 for engines with a time-stamp head); utterances shorter than one frame are
    left out (they produce no tokens)."""
    rows = {i: encoder_rows(int(num_samples[i])) for i in indices}
    order = sorted(indices, key=lambda i: (-rows[i], i))
    out, cur, cur_rows = [], [], 0
    for i in order:
        if rows[i] < 1:
            continue
        if cur and (cur_rows + rows[i] > max_rows or (max_utts and len(cur) >= max_utts)):
            out.append(cur)
            cur, cur_rows = [], 0
        cur.append(i)
        cur_rows += rows[i]
    if cur:
        out.append(cur)
    return out


def _encode_items(local: dict) -> "np.ndarray":
    """{utterance index: ids | (ids, stamps)} -> int32 payload: [n, index[n], n_ids[n], n_stamps[n] (-1: no stamps), every
    utterance's ids back to back, every utterance's (begin, end) pairs back to back] -- 3 n + 1 header slots, then one slot
    per token and two per stamp.  Built with numpy (a Python loop over 10^5 tokens costs more than the collective)."""
    import numpy as np

    n = len(local)
    index = np.fromiter(local.keys(), dtype=np.int32, count=n)
    paired = [isinstance(v, tuple) for v in local.values()]
    ids = [np.asarray(v[0] if p else v, dtype=np.int32).reshape(-1) for v, p in zip(local.values(), paired)]
    stamps = [np.asarray(v[1], dtype=np.int32).reshape(-1) if p else np.empty(0, np.int32) for v, p in zip(local.values(), paired)]
    n_ids = np.fromiter((a.size for a in ids), dtype=np.int32, count=n)
    n_st = np.fromiter((a.size // 2 if p else -1 for a, p in zip(stamps, paired)), dtype=np.int32, count=n)
    return np.concatenate([np.asarray([n], np.int32), index, n_ids, n_st] + ids + stamps)


def _decode_items(buf, as_arrays: bool = False) -> dict:
    """Inverse of _encode_items.  as_arrays: ids as int32 array views into `buf` (stamps as [n, 2] views) instead of Python
    lists -- turning 10^5..10^6 token ids into Python ints costs 16 ns apiece, more than everything else in the gather."""
    import numpy as np

    buf = np.asarray(buf)
    n = int(buf[0])
    hdr = buf[1: 1 + 3 * n].tolist()
    index, n_ids, n_st = hdr[:n], hdr[n: 2 * n], hdr[2 * n:]
    o = 1 + 3 * n
    so = o + sum(n_ids)
    used = so + 2 * sum(k for k in n_st if k > 0)
    flat = buf if as_arrays else buf[:used].tolist()  # lists: ONE conversion of the used prefix, list slices from here on
    out = {}
    for k in range(n):
        ids = flat[o: o + n_ids[k]]
        o += n_ids[k]
        if n_st[k] < 0:
            out[index[k]] = ids
        else:
            e = so + 2 * n_st[k]
            out[index[k]] = (ids, flat[so:e].reshape(-1, 2) if as_arrays else [flat[t: t + 2] for t in range(so, e, 2)])
            so = e
    return out


def payload_capacity(num_samples: Sequence[int], world_size: int, with_stamps: bool = False) -> int:
    """int32 slots per rank of the transcript gather, from the plan alone (every rank computes the same number, nothing
    is exchanged): an utterance of T encoder rows has at most T + 1 tokens (the CIF tail frame)."""
    plan = _plan(tuple(int(n) for n in num_samples), int(world_size))
    return plan[2] if with_stamps else plan[1]


def gather_transcripts(local: dict, world_size: int, rank: int, group=None, capacity: int = 0, device=None, stats=None,
                       as_arrays: bool = False):
    """Collect {utterance index: token ids (or (ids, stamps))} from every rank on rank 0 -- the path's only collective
    (SURVEY.md 8e): ONE all_gather of a fixed-size int32 buffer per rank (ids + lengths, a few hundred KB in total) over
    NCCL (device buffers) or gloo (host buffers, the CPU tests).  `capacity` = int32 slots per rank (payload_capacity());
    0: agreed on first with one MAX all-reduce.  Returns the merged dict on rank 0, None elsewhere; `as_arrays`: ids (and
    stamps) stay int32 arrays."""
    if world_size == 1:
        return dict(local)
    import numpy as np
    import torch
    import torch.distributed as dist

    flat = _encode_items(local)
    dev = torch.device(device) if device is not None else torch.device("cpu")
    if capacity <= 0:
        need = torch.tensor([flat.size], dtype=torch.int64, device=dev)
        dist.all_reduce(need, op=dist.ReduceOp.MAX, group=group)
        capacity = int(need.item())
    if flat.size > capacity:
        raise RuntimeError("gather_transcripts: %d int32 slots needed, capacity %d" % (flat.size, capacity))
    host = torch.zeros(capacity, dtype=torch.int32, pin_memory=dev.type == "cuda")
    host[: flat.size] = torch.from_numpy(flat)
    mine = host.to(dev, non_blocking=True)
    everyone = torch.empty(world_size * capacity, dtype=torch.int32, device=dev)
    dist.all_gather_into_tensor(everyone, mine, group=group)  # one collective straight into the flat table (no per-rank copies)
    if stats is not None:
        stats["gather_bytes"] = int(world_size * capacity * 4)
    if rank != 0:
        return None
    merged = {}
    table = everyone.cpu().numpy().reshape(world_size, capacity)
    for r in range(world_size):
        merged.update(_decode_items(table[r], as_arrays))
    return merged


def recognize_sharded(engine, waves, rank: int, world_size: int, group=None, stats=None, as_arrays: bool = False):
    """Data-parallel recognition of `waves` (every rank holds or can index all of them; entries of other ranks' shards may
    be None as long as `lens` can be told -- pass objects with `.shape`): this rank runs its shard through
    `engine.recognize`, rank 0 returns the token ids of ALL utterances -- or (ids, stamps) when the engine has a
    time-stamp head (`bicif.BiCifParaformerEngine`, `seaco.SeacoParaformerEngine`).  `as_arrays`: every utterance's ids as
    an int32 numpy array (stamps as an [n, 2] array) instead of Python lists -- for callers that feed a tokenizer or a
    file writer anyway; on 8 GPUs the list conversion of 10^5 ids is 3-5 % of a step."""
    import time

    lens = [int(w.shape[0]) for w in waves]
    mine = shard(lens, world_size)[rank]
    t0 = time.perf_counter()
    # engines that can hand the ids over as int32 arrays do (no Python list per utterance on the way into the payload)
    kw = {"as_arrays": True} if getattr(engine, "RETURNS_ARRAYS", False) else {}
    out = engine.recognize([waves[i] for i in mine], **kw)
    t1 = time.perf_counter()
    paired = isinstance(out, tuple)  # engines with a time-stamp head return (ids, stamps)
    items = list(zip(*out)) if paired else out
    local = {i: ((item[0], item[1]) if paired else item) for i, item in zip(mine, items)}
    dev = getattr(engine, "device", None)
    use_dev = dev if (dev is not None and world_size > 1 and _backend(group) == "nccl") else None
    merged = gather_transcripts(local, world_size, rank, group, payload_capacity(lens, world_size, paired), use_dev, stats,
                                as_arrays)
    if stats is not None:  # host clocks around this rank's share and the gather (diagnostics for bench.py)
        stats.setdefault("local_ms", []).append((t1 - t0) * 1e3)
        stats.setdefault("gather_ms", []).append((time.perf_counter() - t1) * 1e3)
    if merged is None:
        return None
    if as_arrays:
        import numpy as np

        none_i, none_s = np.empty(0, np.int32), np.empty((0, 2), np.int32)
        if world_size == 1:  # (nothing was gathered: the engine's own objects)
            merged = {i: ((np.asarray(v[0], np.int32), np.asarray(v[1], np.int32).reshape(-1, 2)) if paired else np.asarray(v, np.int32))
                      for i, v in merged.items()}
    else:
        none_i, none_s = [], []
        if world_size == 1 and kw:  # asked the engine for arrays, the caller wants lists
            merged = {i: ((list(map(int, v[0])), [list(map(int, p)) for p in v[1]]) if paired else list(map(int, v))) for i, v in merged.items()}
    if paired:
        rows = [merged.get(i, (none_i, none_s)) for i in range(len(waves))]
        return tuple([row[k] for row in rows] for k in range(2))
    return [merged.get(i, none_i) for i in range(len(waves))]


def _backend(group=None) -> str:
    import torch.distributed as dist

    return str(dist.get_backend(group)) if dist.is_initialized() else ""


def slice_segments(speech, segments_ms):
    """Waveforms of VAD segments `[[beg_ms, end_ms], ...]` of one long recording, cut like the reference's
    `slice_padding_audio_samples` (funasr/utils/vad_utils.py:21-33: 16 samples per millisecond, end clipped to the
    recording)."""
    n = int(speech.shape[0])
    out = []
    for beg_ms, end_ms in segments_ms:
        b = int(beg_ms * 16)
        e = min(int(end_ms * 16), n)
        out.append(speech[b:e])
    return out


def recognize_segments(engine, speech, segments_ms):
    """Second stage of the reference's long-audio pipeline (`AutoModel.inference_with_vad`,
    funasr/auto/auto_model.py:400-470: VAD segments sorted by duration, batched under `batch_size_s`, one
    `inference` call per batch, results put back in time order) as ONE call into the packed engine: every segment
    of the recording goes through `engine.recognize`, which length-sorts and packs them under its row budget, and
    the token ids come back in segment order.  Because each segment is computed with batch-1 semantics the result
    does not depend on how the segments were grouped."""
    return engine.recognize(slice_segments(speech, segments_ms))


def recognize_segments_with_stamps(engine, speech, segments_ms):
    """The same for an engine with a time-stamp head (`bicif.BiCifParaformerEngine`): returns (ids, stamps) of the WHOLE
    recording in time order, every segment's `[begin ms, end ms]` pairs shifted by the segment's start as the reference's
    result combination does (funasr/auto/auto_model.py:510-518: `t[0] += vadsegments[j][0]`, integers, after the
    per-segment inference)."""
    ids, stamps = engine.recognize(slice_segments(speech, segments_ms))
    all_ids, all_stamps = [], []
    for (beg_ms, _), seg_ids, seg_stamps in zip(segments_ms, ids, stamps):
        all_ids.extend(seg_ids)
        all_stamps.extend([[b + int(beg_ms), e + int(beg_ms)] for b, e in seg_stamps])
    return all_ids, all_stamps
