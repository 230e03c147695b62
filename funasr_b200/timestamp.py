"""Host side of the time-stamp path: fired frames -> [begin ms, end ms] per token.

Mirrors the arithmetic of `ts_prediction_lfr6_standard` (funasr/utils/timestamp_tools.py:31-106) in float64, the type the
reference computes in (numpy int64 indices + a Python float shift), so the truncated millisecond values are identical.
The fire search itself (and its re-normalised second pass, :46-53) runs on the device: `fb200_us_fires`.
"""
from typing import List

import numpy as np

FORCE_TIME_SHIFT = -1.5
START_END_THRESHOLD = 5
MAX_TOKEN_DURATION = 12  # in upsampled frames
UPSAMPLE_RATE = 3
TIME_RATE = 10.0 * 6 / 1000 / UPSAMPLE_RATE


def stamps_from_fires(fire_idx: np.ndarray, num_frames: int, num_chars: int, vad_offset: float = 0.0) -> List[List[int]]:
    """fire_idx: fired upsampled frames of one utterance in order (one more than tokens when the head is consistent);
    num_frames = 3 x encoder frames.  Returns one [begin_ms, end_ms] per token (silence entries are not reported,
    :101-104).  The reference indexes its character list by fire interval and raises IndexError when there are more
    intervals than characters; here the list is cut to `num_chars`."""
    n = int(len(fire_idx))
    if num_chars <= 0 or n < 1:
        return []
    fp = np.asarray(fire_idx, dtype=np.int64) + FORCE_TIME_SHIFT
    begin = fp[:-1] * TIME_RATE
    gap = fp[1:] - fp[:-1]
    split = gap > MAX_TOKEN_DURATION
    end = np.where(split, (fp[:-1] + MAX_TOKEN_DURATION) * TIME_RATE, fp[1:] * TIME_RATE)
    if n > 1 and not split[-1]:  # the last list entry is the last token (otherwise it is the silence after it)
        if num_frames - fp[-1] > START_END_THRESHOLD:
            end[-1] = ((num_frames + fp[-1]) * 0.5) * TIME_RATE
        else:
            end[-1] = num_frames * TIME_RATE
    if vad_offset:
        begin = begin + vad_offset / 1000.0
        end = end + vad_offset / 1000.0
    b = np.trunc(begin * 1000).astype(np.int64)
    e = np.trunc(end * 1000).astype(np.int64)
    return [[int(x), int(y)] for x, y in zip(b[:num_chars], e[:num_chars])]
