"""Golden token ids of ALL 1024 utterances of BASELINE configs[1], from the UNMODIFIED reference's CPU AutoModel
(batch 1, fp32, dither 0, seeded weights / CMVN / waveforms) -> tests/golden/configs1_ids_1024.npz.

    python tests/golden/make_golden_configs1.py [--utts 1024]

About 16 000 s of audio: ~100 s on the GPU box's 16 host cores (4 worker processes x 4 threads), ~10 min in the build
container.  The reference tree is /root/reference here and oracle/_ref/pyref (an unmodified copy, `make -C oracle`) on the
box.  Stored: per-utterance token ids (blank / sos / eos dropped, as `Paraformer.inference` returns them,
funasr/models/paraformer/model.py:561-565) packed back to back with their offsets, and the CIF token counts
(`pre_token_length`, model.py:513)."""
import argparse
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from funasr_b200 import synth  # noqa: E402
from oracle import reference_runner as R  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--utts", type=int, default=1024)
    ap.add_argument("--out", default=os.path.join(HERE, "configs1_ids_1024.npz"))
    args = ap.parse_args()
    assert R.available(), "reference tree not found"
    durs = synth.synth_durations(args.utts)
    specs = [(i, i, int(n)) for i, n in enumerate(durs)]
    want = R.pool_recognize("reference", specs)
    ids = [np.asarray(want[i][0], dtype=np.int32) for i in range(args.utts)]
    offsets = np.zeros(args.utts + 1, dtype=np.int64)
    offsets[1:] = np.cumsum([a.size for a in ids])
    np.savez_compressed(args.out, ids=np.concatenate(ids), offsets=offsets,
                        token_num=np.asarray([want[i][1] for i in range(args.utts)], dtype=np.int32),
                        num_samples=np.asarray(durs, dtype=np.int64))
    print("wrote %s: %d utterances, %d tokens" % (args.out, args.utts, int(offsets[-1])))


if __name__ == "__main__":
    main()
