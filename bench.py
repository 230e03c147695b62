#!/usr/bin/env python
"""Benchmark of the batched offline Paraformer-large path (BASELINE.json metric: RTFx).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--precision fp32|bf16]
    torchrun --nproc-per-node N bench.py --gpus N ...          (one rank per GPU, NCCL)

Workload (BASELINE.json configs[1]): ONE set of 1024 synthetic utterances, durations U(2,30) s (seed 7),
0.1*randn waveforms (utterance i seeded 1000+i), random-init Paraformer-large weights (seed 0), synthetic CMVN (seed 3),
dither 0.  One "step" = one pass of the whole path (fbank -> LFR/CMVN -> 50-layer SAN-M encoder -> CIF -> 16-layer
decoder -> greedy pick) over the set, packed into var-len batches under a row budget.

N = 1   value  RTFx with the waveforms already resident in HBM (CUDA events)
        e2e    RTFx through ParaformerEngine.recognize(): pinned host waveforms -> H2D -> path -> D2H ids
N > 1   STRONG scaling, as BASELINE configs[1] words it: the SAME 1024 utterances sharded over the ranks
        (funasr_b200/schedule.py: longest-processing-time-first plan, per-rank packed batches), every step ends with the
        NCCL gather of the transcripts on rank 0 (schedule.gather_transcripts: one all_gather of int32 ids + lengths);
        value = device-resident waveforms, e2e = pinned host waveforms, both through schedule.recognize_sharded and
        timed with CUDA events from a barrier, max over ranks.  Extra keys: per_rank_ms, imbalance, gather_ms,
        gather_bytes, and `large_set` = 1024 x N utterances sharded the same way (per-GPU work as at N = 1: tells fixed
        costs and imbalance apart; its value is also reported as `weak_value`).

  roofline      the tcgen05 GEMM (dominant kernel) timed live with CUDA events around every launch of one extra profiled
                step; achieved = algorithmic 2MNK FLOPs / device time
  cpu_baseline  N = 1: the UNMODIFIED reference's CPU AutoModel (oracle/_ref/pyref; the oracle port when that tree is
                missing) on a bounded sample, on all host cores as a pool of batch-1 worker processes
  parity        N = 1: token ids / CIF token counts of that sample, reference CPU vs this run's GPU results
  configs       N = 1: short sub-results for BASELINE configs[0], [2], [3], [4]; N > 1: configs[4] sharded
  --impl reference   the reference arm: the same CPU pool timed over --steps steps of the bounded sample
"""
import argparse
import json
import os
import subprocess
import sys
import time
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from funasr_b200 import params, schedule, synth  # noqa: E402

# The contract is ONE JSON line on stdout: everything else that libraries print there (the NCCL version banner,
# warnings) is sent to stderr by pointing fd 1 at fd 2 for the whole run; `emit` writes to the real stdout.
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(obj) -> None:
    os.write(_REAL_STDOUT, (json.dumps(obj) + "\n").encode())


FRAME_S = 0.06  # audio seconds per encoder row: frame_shift 10 ms * lfr_n 6 (paraformer/model.py:491-493)
METRIC = "Paraformer-large RTFx (audio s / wall s)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default="fp32", choices=["fp32", "fp32x6", "bf16"])
    ap.add_argument("--utts", type=int, default=1024)
    ap.add_argument("--max-rows", type=int, default=56832,
                    help="packed rows per batch (3 waves of 148 SMs x 128-row tiles; 18944 = one wave was the default until the end of round 2)")
    ap.add_argument("--cpu-utts", type=int, default=96, help="utterances of the workload timed on the host cores (~10 s of CPU work)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the sub-results for the other BASELINE configs")
    ap.add_argument("--no-large-set", action="store_true")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm_gbs": p["hbm_gbs"], "tf_burst": p["bf16_tflops"], "tf_sustained": p["bf16_tflops_sustained"],
                "source": "measured"}
    return {"hbm_gbs": 6650.0, "tf_burst": 1590.0, "tf_sustained": 1400.0, "source": "fallback"}


def audio_seconds(durs) -> float:
    """As the reference counts it: rows * frame_shift * lfr_n / 1000 (paraformer/model.py:491-493)."""
    return sum(schedule.encoder_rows(int(n)) for n in durs) * FRAME_S


def state_dict():
    sd = synth.synth_state_dict(params.paraformer_large_shapes(), seed=0)
    return {k: torch.from_numpy(v) for k, v in sd.items()}


def cmvn():
    return tuple(torch.from_numpy(x) for x in synth.synth_cmvn())


def cpu_pool():
    """The reference's own CPU implementation when its tree travelled (oracle/_ref/pyref), else the oracle port."""
    from oracle import reference_runner as R

    kind = "reference" if R.available() else "oracle"
    return R.CpuPool(kind), ("reference" if kind == "reference" else "port")


def cpu_sample_text(kind, n, utts, secs, pool):
    what = "the unmodified reference's AutoModel.generate (device=cpu, batch 1)" if kind == "reference" else "oracle port, batch 1"
    return "first %d of the %d utterances (%.1f s audio) per step; %s in %d worker processes x %d threads" % (
        n, utts, secs, what, pool.procs, pool.threads)


def run_reference(args):
    """Reference arm: the reference's own CPU implementation of the path on the box's host cores, bounded sample per step."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    durs = synth.synth_durations(args.utts)
    n = min(args.cpu_utts, len(durs))
    specs = [(i, i, int(durs[i])) for i in range(n)]
    secs = audio_seconds(durs[:n])
    pool, kind = cpu_pool()
    warm = min(args.warmup, 1)
    for _ in range(warm):
        pool.run(specs)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        pool.run(specs)
    dt = (time.perf_counter() - t0) / args.steps
    sample = cpu_sample_text(kind, n, args.utts, secs, pool)
    cores = pool.cores
    pool.close()
    v = secs / dt
    emit({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "audio_s/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": warm, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong" if args.gpus > 1 else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "configs[1]: 1024 synthetic utterances U(2,30) s, Paraformer-large random init", "sample": sample},
        "cpu_baseline": {"value": v, "unit": "audio_s/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "audio_s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            pass

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        out = self.proc.communicate(timeout=10)[0]
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


class Job:
    """Process-group plumbing + device timing shared by every measurement of the run."""

    def __init__(self):
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist

            dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        torch.cuda.synchronize()

    def timed(self, fn, steps):
        """Milliseconds for `steps` calls: CUDA events on the current stream, barrier + synchronize on both sides, MAX over ranks."""
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=self.dev)
        self.barrier()
        if self.dist is not None:
            self.dist.all_reduce(ms, op=self.dist.ReduceOp.MAX)
        return float(ms.item())

    def all_ranks(self, value: float):
        t = torch.tensor([value], dtype=torch.float64, device=self.dev)
        if self.dist is None:
            return [float(value)]
        out = [torch.zeros_like(t) for _ in range(self.world)]
        self.dist.all_gather(out, t)
        return [float(x.item()) for x in out]

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()


class ShardedSet:
    """One utterance set sharded over the ranks: this rank's waveforms in pinned host memory and in HBM; the other ranks'
    entries are shape-only placeholders (schedule.recognize_sharded needs every LENGTH, not every waveform)."""

    def __init__(self, job, durs, first_seed=0):
        self.durs = [int(n) for n in durs]
        self.mine = schedule.shard(self.durs, job.world)[job.rank]
        self.pinned = [types.SimpleNamespace(shape=(n,)) for n in self.durs]
        self.resident = list(self.pinned)
        for i in self.mine:
            w = torch.from_numpy(synth.synth_waveform(first_seed + i, self.durs[i])).pin_memory()
            self.pinned[i] = w
            self.resident[i] = w.to(job.dev)
        self.secs = audio_seconds(self.durs)
        self.h2d = sum(self.durs) * 4


def sharded_steps(job, eng, data, stats):
    def run(waves):
        t0 = time.perf_counter()
        out = schedule.recognize_sharded(eng, waves, job.rank, job.world, stats=stats, as_arrays=True)  # ids as int32 arrays
        stats["step_ms"].append((time.perf_counter() - t0) * 1e3)
        return out

    return (lambda: run(data.resident)), (lambda: run(data.pinned))


def other_configs(job, eng, args):
    """Short sub-results for the other BASELINE configs on rank 0's GPU (N = 1): end to end from pinned host memory
    through each engine's recognize(), 1 warm-up + 2 timed steps; parity of these paths is the job of tests/."""
    from funasr_b200 import conformer, engine

    dev, out = job.dev, {}

    def rtfx(fn, secs, steps=2):
        fn()
        return secs * steps / (job.timed(fn, steps) * 1e-3)

    # configs[0]: one 10 s utterance (the waveform of tests/golden/stages_10s.npz: ids checked against the reference's)
    wave10 = [torch.from_numpy(synth.synth_waveform(11, 160000)).pin_memory()]
    ids10 = eng.recognize(wave10)[0]
    lat = []
    for _ in range(12):
        t0 = time.perf_counter()
        eng.recognize(wave10)
        lat.append((time.perf_counter() - t0) * 1e3)
    entry = {"workload": "one synthetic 10 s utterance, host waveform -> token ids, ParaformerEngine.recognize",
             "ms_per_utterance": float(np.median(lat)), "value": 10.0 / (float(np.median(lat)) * 1e-3), "unit": "audio_s/s",
             "path": "request-stack (two persistent cooperative kernels)" if getattr(eng, "_small", None) else "per-op launches"}
    golden = os.path.join(ROOT, "tests", "golden", "stages_10s.npz")
    if os.path.exists(golden) and args.precision == "fp32":
        entry["ids_equal_reference_golden"] = bool(ids10 == np.load(golden)["token_int"].tolist())
    # the same request through the native runtime (fb200_offline_infer: no Python between the kernels), and the weight
    # stream it implies: every request reads all contraction weights once (hi + mid bf16 = 4 bytes per parameter)
    try:
        import tempfile

        from funasr_b200 import runtime

        d = tempfile.mkdtemp(prefix="fb200_bench_")
        sd_t = state_dict()
        runtime.export_model({k: v.numpy() for k, v in sd_t.items()}, d, cmvn=synth.synth_cmvn())
        rec = runtime.OfflineRecognizer(d, device=job.local, precision="bf16" if args.precision == "bf16" else "fp32")
        w10 = [wave10[0].numpy()]
        for _ in range(5):
            rec.recognize(w10)
        nat = []
        for _ in range(12):
            t0 = time.perf_counter()
            ids_n = rec.recognize(w10)
            nat.append((time.perf_counter() - t0) * 1e3)
        rec.close()
        wbytes = (2.0 if args.precision == "bf16" else 4.0) * sum(v.numel() for k, v in sd_t.items()
                                                                   if v.dim() >= 2 and not k.startswith("decoder.embed"))
        del sd_t
        ms = float(np.median(nat))
        entry["native_runtime_ms"] = ms
        entry["native_ids_equal"] = bool(list(ids_n[0]) == list(ids10))
        peak = peaks().get("hbm_gbs")
        entry["roofline"] = {"bound": "hbm", "achieved": wbytes / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                             "frac": (wbytes / (ms * 1e-3) / 1e9 / peak) if peak else None,
                             "bytes": wbytes, "note": "weights streamed once per request / host-to-ids time of the native runtime"}
        import shutil

        shutil.rmtree(d, ignore_errors=True)
    except Exception as exc:  # the sub-result is additional: never fail the bench line over it
        entry["native_runtime_error"] = str(exc)[:200]
    out["configs[0]"] = entry
    # configs[4]: long-audio sweep through the same engine
    long_durs = [int(s * synth.SAMPLE_RATE) for s in (60, 120, 300, 600) for _ in range(8)]
    long_waves = [torch.from_numpy(synth.synth_waveform(2000 + i, n)).pin_memory() for i, n in enumerate(long_durs)]
    secs = sum(long_durs) / synth.SAMPLE_RATE
    out["configs[4]"] = {"workload": "Paraformer-large long-audio sweep: 8 segments each of 60 / 120 / 300 / 600 s (%.0f s audio)" % secs,
                         "value": rtfx(lambda: eng.recognize(long_waves), secs), "unit": "audio_s/s"}
    del long_waves
    durs = synth.synth_durations(256)
    waves = [torch.from_numpy(synth.synth_waveform(2000 + i, int(n))).pin_memory() for i, n in enumerate(durs)]
    secs = float(sum(int(n) for n in durs)) / synth.SAMPLE_RATE
    # configs[2]: SenseVoiceSmall
    sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.sense_voice_small_shapes(), seed=0).items()}
    sv = engine.SenseVoiceEngine(sd, dev, args.precision, cmvn=cmvn(), max_rows=args.max_rows)
    out["configs[2]"] = {"workload": "SenseVoiceSmall (50 + 20 SAN-M blocks, V = 25055) CTC greedy, 256 synthetic utterances U(2,30) s (%.0f s audio)" % secs,
                         "value": rtfx(lambda: sv.recognize(waves), secs), "unit": "audio_s/s"}
    del sv, sd
    torch.cuda.empty_cache()
    # configs[3]: Conformer-CTC
    sd = {k: torch.from_numpy(v) for k, v in synth.synth_state_dict(params.conformer_ctc_shapes(), seed=0).items()}
    cf = conformer.ConformerCTCEngine(sd, dev, args.precision, max_rows=args.max_rows)
    out["configs[3]"] = {"workload": "Conformer-CTC (aishell recipe, 12 blocks, d = 256) encoder + CTC greedy, 256 synthetic utterances U(2,30) s (%.0f s audio)" % secs,
                         "value": rtfx(lambda: cf.recognize(waves), secs), "unit": "audio_s/s"}
    del cf, sd
    torch.cuda.empty_cache()
    return out


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
        return
    from funasr_b200 import _lib, engine

    job = Job()
    world, rank, dev = job.world, job.rank, job.dev
    durs = synth.synth_durations(args.utts)
    data = ShardedSet(job, durs)
    eng = engine.ParaformerEngine(state_dict(), dev, args.precision, cmvn=cmvn(), max_rows=args.max_rows)
    mine_pinned = [data.pinned[i] for i in data.mine]
    batches = eng.plan_batches([data.durs[i] for i in data.mine])
    stats = {"step_ms": [], "local_ms": [], "gather_ms": []}

    if world == 1:
        resident = [eng.upload([mine_pinned[i] for i in b]) for b in batches]
        torch.cuda.synchronize()

        def step_resident():
            n_tok = 0
            for wave, lens in resident:
                ids, _, tok = eng.forward_packed(wave, lens)
                n_tok += len(ids.cpu())
            return n_tok

        def step_e2e():
            return eng.recognize(mine_pinned)
    else:
        step_resident, step_e2e = sharded_steps(job, eng, data, stats)

    for _ in range(args.warmup):
        step_resident()
    sampler = ClockSampler(job.local) if rank == 0 else None
    _lib.launch_count(reset=True)
    for k in ("step_ms", "local_ms", "gather_ms"):
        del stats[k][:]
    ms_res = job.timed(step_resident, args.steps)
    launches = _lib.launch_count()
    step_e2e()
    for k in ("step_ms", "local_ms", "gather_ms"):
        del stats[k][:]
    ms_e2e = job.timed(step_e2e, args.steps)
    e2e_stats = {k: list(stats[k]) for k in ("step_ms", "local_ms", "gather_ms")}
    clocks = sampler.stop() if sampler else None
    got = step_e2e()  # rank 0: ids of every utterance of the set (the parity sample below reads them)
    counts = None
    if world == 1:  # recognize() answered in the order of `mine` (the LPT order of the plan): back to utterance order
        by_index, cnt_by_index = [None] * len(durs), [0] * len(durs)
        for k, i in enumerate(data.mine):
            by_index[i], cnt_by_index[i] = got[k], int(eng.last_token_num[k])
        got, counts = by_index, cnt_by_index

    # one extra profiled pass over this rank's share: CUDA events around every kernel entry point
    _lib.profile_read(reset=True)
    _lib.profile_enable(True)
    step_resident()
    torch.cuda.synchronize()
    _lib.profile_enable(False)
    prof = _lib.profile_read(reset=True)

    # per-rank view of the sharded steps (host clocks around the local work and the gather: diagnostics, not the headline)
    per_rank = job.all_ranks(float(np.mean(e2e_stats["local_ms"])) if e2e_stats["local_ms"] else ms_e2e / args.steps)
    gather_ms = job.all_ranks(float(np.mean(e2e_stats["gather_ms"])) if e2e_stats["gather_ms"] else 0.0)

    large = None
    if world > 1 and not args.no_large_set:
        big = ShardedSet(job, synth.synth_durations(args.utts * world, seed=11), first_seed=100000)
        bstats = {"step_ms": [], "local_ms": [], "gather_ms": []}
        _, big_e2e = sharded_steps(job, eng, big, bstats)
        big_e2e()
        ms_big = job.timed(big_e2e, 2)
        big_rank = job.all_ranks(float(np.mean(bstats["local_ms"][-2:])))
        large = {"utterances": len(big.durs), "audio_s": big.secs, "ms_per_step": ms_big / 2,
                 "value": big.secs * 2 / (ms_big * 1e-3), "unit": "audio_s/s", "timed": "e2e (pinned host -> transcripts on rank 0)",
                 "per_rank_ms": {"min": min(big_rank), "max": max(big_rank)}, "imbalance": max(big_rank) / (sum(big_rank) / world)}
        del big

    sub = None
    if not args.no_configs:
        if world == 1:
            sub = other_configs(job, eng, args)
        else:  # configs[4] at N GPUs: the 32 long segments sharded like any other set
            long_set = ShardedSet(job, [int(s * synth.SAMPLE_RATE) for s in (60, 120, 300, 600) for _ in range(8)], first_seed=2000)
            _, long_e2e = sharded_steps(job, eng, long_set, {"step_ms": [], "local_ms": [], "gather_ms": []})
            long_e2e()
            ms_long = job.timed(long_e2e, 2)
            sub = {"configs[4]": {"workload": "Paraformer-large long-audio sweep: 8 segments each of 60 / 120 / 300 / 600 s "
                                              "(%.0f s audio) sharded over %d GPUs" % (long_set.secs, world),
                                  "value": long_set.secs * 2 / (ms_long * 1e-3), "unit": "audio_s/s", "ms_per_step": ms_long / 2}}

    if rank != 0:
        job.close()
        return
    pk = peaks()
    secs = data.secs
    value = secs * args.steps / (ms_res * 1e-3)
    e2e = secs * args.steps / (ms_e2e * 1e-3)
    g = prof.get("gemm", {"ms": 0.0, "calls": 0, "work": 0.0})
    achieved = g["work"] / (g["ms"] * 1e-3) / 1e12 if g["ms"] else 0.0
    terms = {"fp32": 3, "fp32x6": 6, "bf16": 1}[args.precision]
    tot_ms = sum(v["ms"] for v in prof.values()) or 1.0
    n_batches = len(schedule.batches(data.durs, list(range(len(data.durs))), args.max_rows)) if world == 1 else None
    h2d = int(data.h2d + (sum(3 * (len(b) + 1) * 4 + (len(b) + 1) * 8 for b in batches) if world == 1 else 0))
    n_tok = sum(len(x) for x in got) if got is not None else 0
    d2h = int(n_tok * 4 + len(durs) * 4)
    # DRAM bytes per GEMM launch: one ncu pass over the same command, summarised under profiles/ (bytes, or None)
    gemm_traffic, gemm_traffic_src = None, None
    for name in ("r02_gemm_dram_56832.json", "r02_gemm_dram.json", "r01_gemm_dram.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                _t = json.load(f)
            if int(_t.get("max_rows", 18944)) == args.max_rows:  # per launch like `achieved`: same rows per launch only
                gemm_traffic = _t["dram_bytes_per_launch"]
                gemm_traffic_src = "profiles/%s (ncu dram__bytes_read+write.sum, avg over %d launches)" % (name, _t["launches"])
                break
            if gemm_traffic_src is None:
                gemm_traffic_src = "none at %d-row batches; profiles/%s holds %.0f MB per launch for %d-row batches" % (
                    args.max_rows, name, _t["dram_bytes_per_launch"] / 1e6, int(_t.get("max_rows", 18944)))
        except (OSError, KeyError, ValueError):
            continue
    if world == 1:
        work = "configs[1]: %d synthetic utterances U(2,30) s (%.0f s audio), Paraformer-large random init, packed var-len " \
               "batches <= %d rows" % (args.utts, secs, args.max_rows)
    else:
        work = "configs[1]: ONE set of %d synthetic utterances U(2,30) s (%.0f s audio) sharded over %d GPUs (LPT plan, ~%d " \
               "utterances per GPU, packed var-len batches <= %d rows), transcripts gathered on rank 0 by one NCCL all_gather " \
               "inside every step (ids returned as one int32 array per utterance)" % (args.utts, secs, world, args.utts // world, args.max_rows)
    line = {
        "metric": METRIC, "value": value, "unit": "audio_s/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_res / args.steps, "higher_is_better": True,
        "scaling": "weak" if world == 1 else "strong", "vs_baseline": None,
        "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
        "config": {"workload": work,
                   "value_is": "device-resident waveforms (e2e = the same from pinned host memory, the headline against the reference arm)",
                   "precision": {"fp32": "fp32-parity: operands split hi+mid bf16 (16-bit mantissa), 3 tcgen05 terms, "
                                         "fp32 accumulate; all other kernels fp32",
                                 "fp32x6": "fp32-parity: operands split hi+mid+lo bf16, 6 tcgen05 terms, fp32 accumulate",
                                 "bf16": "bf16 operands, fp32 accumulate"}[args.precision],
                   "batches_per_step": n_batches if world == 1 else "%d on rank 0" % len(batches), "parallelism": "dp%d" % world,
                   "l2": "inputs > L2: %.2f GB of waveforms + %.0f MB of weights per step" % (data.h2d / 1e9, 440.0 * eng.nparts)},
        "e2e": {"value": e2e, "unit": "audio_s/s", "ms_per_step": ms_e2e / args.steps, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "tensor", "kernel": "gemm_tcgen05_kernel", "achieved": achieved,
                     "peak": pk["tf_sustained"], "unit": "TFLOP/s", "frac": achieved / pk["tf_sustained"],
                     "traffic": gemm_traffic, "traffic_source": gemm_traffic_src,
                     "peak_source": pk["source"] + " (sustained bf16)",
                     "launches": g["calls"], "avg_launch_us": 1e3 * g["ms"] / max(g["calls"], 1),
                     "share_of_step": g["ms"] / tot_ms, "tensor_terms_per_flop": terms,
                     "executed_tflops": achieved * terms},
        "breakdown_ms": {k: round(v["ms"], 3) for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])},
    }
    if world > 1:
        line["per_rank_ms"] = {"min": min(per_rank), "max": max(per_rank), "all": [round(x, 2) for x in per_rank]}
        line["imbalance"] = max(per_rank) / (sum(per_rank) / world)
        line["gather_ms"] = max(gather_ms)
        line["gather_bytes"] = int(stats.get("gather_bytes", 0))
        if large is not None:
            line["large_set"] = large
            line["weak_value"] = large["value"]
    if sub is not None:
        line["configs"] = sub
    if world == 1 and not args.no_cpu_baseline:
        pool, kind = cpu_pool()
        n = min(args.cpu_utts, len(durs))
        specs = [(i, i, int(durs[i])) for i in range(n)]
        pool.run(specs[:pool.procs])  # warm-up: one utterance per worker
        t0 = time.perf_counter()
        want = pool.run(specs)
        dt = time.perf_counter() - t0
        s = audio_seconds(durs[:n])
        line["cpu_baseline"] = {"value": s / dt, "unit": "audio_s/s", "cores": pool.cores, "kind": kind,
                                "sample": cpu_sample_text(kind, n, args.utts, s, pool) + " (%.1f s CPU)" % dt}
        pool.close()
        same_ids = sum(1 for i in range(n) if got[i] == want[i][0])
        same_cnt = sum(1 for i in range(n) if counts[i] == want[i][1])
        line["parity"] = {"utterances": n, "ids_identical": same_ids, "counts_identical": same_cnt,
                          "against": "reference CPU AutoModel, batch 1" if kind == "reference" else "oracle port, batch 1",
                          "full_gate": "tests/test_gpu_parity_gate.py (all 1024 utterances; profiles/r02_parity_gate.json)"}
    emit(line)
    job.close()


if __name__ == "__main__":
    main()
