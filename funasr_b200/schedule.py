"""Sharding and batching of utterances across the GPUs of one box (pure data parallel).

Every rank computes the SAME plan from the list of utterance lengths -- no communication -- and
works on its shard; rank 0 gathers (utterance index, token ids) at the end.  The reference's own
analogues are the `wav.scp` split into one process per GPU (examples/aishell/paraformer/run.sh:154-193)
and the duration-sorted dynamic batching of funasr/auto/auto_model.py:423-459.
"""
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


def shard(num_samples: Sequence[int], world_size: int) -> List[List[int]]:
    """Longest-processing-time-first assignment: indices per rank, deterministic (ties by index)."""
    rows = [encoder_rows(int(n)) for n in num_samples]
    order = sorted(range(len(rows)), key=lambda i: (-rows[i], i))
    loads = [0.0] * world_size
    out: List[List[int]] = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (loads[k], k))
        out[r].append(i)
        loads[r] += cost(rows[i])
    return out


def batches(num_samples: Sequence[int], indices: Sequence[int], max_rows: int) -> List[List[int]]:
    """Indices (into `num_samples`) grouped longest-first under the packed-row budget; utterances
    shorter than one frame are left out (they produce no tokens)."""
    rows = {i: encoder_rows(int(num_samples[i])) for i in indices}
    order = sorted(indices, key=lambda i: (-rows[i], i))
    out, cur, cur_rows = [], [], 0
    for i in order:
        if rows[i] < 1:
            continue
        if cur and cur_rows + rows[i] > max_rows:
            out.append(cur)
            cur, cur_rows = [], 0
        cur.append(i)
        cur_rows += rows[i]
    if cur:
        out.append(cur)
    return out


def gather_transcripts(local: dict, world_size: int, rank: int, group=None):
    """Collect {utterance index: token ids} from every rank on rank 0 (the path's only collective:
    a few hundred KB, latency-bound).  Returns the merged dict on rank 0, None elsewhere."""
    if world_size == 1:
        return dict(local)
    import torch.distributed as dist

    bucket = [None] * world_size if rank == 0 else None
    dist.gather_object(local, bucket, dst=0, group=group)
    if rank != 0:
        return None
    merged = {}
    for part in bucket:
        merged.update(part)
    return merged


def recognize_sharded(engine, waves, rank: int, world_size: int, group=None):
    """Data-parallel recognition of `waves` (every rank holds or can index all of them): this rank
    runs its shard through `engine.recognize`, rank 0 returns the token ids of ALL utterances -- or (ids, stamps) when
    the engine has a time-stamp head (`bicif.BiCifParaformerEngine`, `seaco.SeacoParaformerEngine`)."""
    lens = [int(w.shape[0]) for w in waves]
    mine = shard(lens, world_size)[rank]
    out = engine.recognize([waves[i] for i in mine])
    paired = isinstance(out, tuple)  # engines with a time-stamp head return (ids, stamps)
    items = list(zip(*out)) if paired else out
    merged = gather_transcripts({i: item for i, item in zip(mine, items)}, world_size, rank, group)
    if merged is None:
        return None
    if paired:
        n_out = len(out)
        empty = tuple([] for _ in range(n_out))
        rows = [merged.get(i, empty) for i in range(len(waves))]
        return tuple([row[k] for row in rows] for k in range(n_out))
    return [merged[i] for i in range(len(waves))]


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
