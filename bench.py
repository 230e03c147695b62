#!/usr/bin/env python
"""Benchmark of the batched offline Paraformer-large path (BASELINE.json metric: RTFx).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--precision fp32|bf16]

Workload (BASELINE.json configs[1]): 1024 synthetic utterances, durations U(2,30) s
(seed 7), 0.1*randn waveforms (utterance i seeded 1000+i), random-init Paraformer-large weights
(seed 0), synthetic CMVN (seed 3), dither 0.  One "step" = one pass of the whole path
(fbank -> LFR/CMVN -> 50-layer SAN-M encoder -> CIF -> 16-layer decoder -> greedy pick) over the
1024 utterances, packed into var-len batches under a row budget.  Weak scaling: every rank
processes its own 1024 utterances (rank r uses seeds 1000 + 1024 r + i); no data-path collective.

  value  RTFx with waveforms already resident in HBM (device-timed, CUDA events, max over ranks)
  e2e    RTFx through ParaformerEngine.recognize(): pinned host waveforms -> H2D -> path -> D2H ids
  roofline  the tcgen05 GEMM (dominant kernel) timed live with CUDA events around every launch of
         one extra profiled step; achieved = algorithmic 2MNK FLOPs / device time
  cpu_baseline  the oracle (torch CPU restatement pinned to the reference) on a bounded sample
"""
import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from funasr_b200 import params, synth  # noqa: E402

# The contract is ONE JSON line on stdout: everything else that libraries print there (the NCCL version banner,
# warnings) is sent to stderr by pointing fd 1 at fd 2 for the whole run; `emit` writes to the real stdout.
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(obj) -> None:
    os.write(_REAL_STDOUT, (json.dumps(obj) + "\n").encode())


FRAME_S = 0.06  # audio seconds per encoder row: frame_shift 10 ms * lfr_n 6 (paraformer/model.py:491-493)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default="fp32", choices=["fp32", "fp32x6", "bf16"])
    ap.add_argument("--utts", type=int, default=1024)
    ap.add_argument("--max-rows", type=int, default=18944, help="packed rows per batch (148 SMs x 128-row tiles)")
    ap.add_argument("--cpu-utts", type=int, default=96, help="utterances of the workload timed on the host cores (~12 s of CPU work)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm_gbs": p["hbm_gbs"], "tf_burst": p["bf16_tflops"], "tf_sustained": p["bf16_tflops_sustained"],
                "source": "measured"}
    return {"hbm_gbs": 6650.0, "tf_burst": 1590.0, "tf_sustained": 1400.0, "source": "fallback"}


def workload(rank: int, utts: int):
    durs = synth.synth_durations(utts)
    waves = [synth.synth_waveform(rank * utts + i, int(n)) for i, n in enumerate(durs)]
    return durs, waves


def audio_seconds(durs) -> float:
    """As the reference counts it: rows * frame_shift * lfr_n / 1000 (paraformer/model.py:491-493)."""
    tot = 0
    for n in durs:
        f = 1 + (int(n) - 400) // 160
        tot += (f + 5) // 6
    return tot * FRAME_S


def state_dict():
    sd = synth.synth_state_dict(params.paraformer_large_shapes(), seed=0)
    return {k: torch.from_numpy(v) for k, v in sd.items()}


def cmvn():
    return tuple(torch.from_numpy(x) for x in synth.synth_cmvn())


def cpu_baseline(durs, waves, count, threads):
    """Oracle (reference CPU semantics: batch 1, fp32) on the first `count` utterances."""
    from oracle import paraformer_oracle as O

    torch.set_num_threads(threads)
    sd, cm = state_dict(), cmvn()
    idx = list(range(min(count, len(waves))))
    O.recognize(torch.from_numpy(waves[idx[0]]), sd, cm)  # warm-up
    t0 = time.perf_counter()
    for i in idx:
        O.recognize(torch.from_numpy(waves[i]), sd, cm)
    dt = time.perf_counter() - t0
    secs = audio_seconds([durs[i] for i in idx])
    return secs / dt, dt, secs, len(idx)


def run_reference(args):
    """Reference arm: the reference's CPU implementation of the path (oracle port; the Python
    reference tree cannot travel to the GPU box), all host threads, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    durs, waves = workload(0, args.utts)
    from oracle import paraformer_oracle as O

    torch.set_num_threads(threads)
    sd, cm = state_dict(), cmvn()
    idx = list(range(min(args.cpu_utts, len(waves))))
    secs = audio_seconds([durs[i] for i in idx])

    def step():
        for i in idx:
            O.recognize(torch.from_numpy(waves[i]), sd, cm)

    for _ in range(min(args.warmup, 1)):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    v = secs / dt
    sample = "first %d of the %d utterances (%.1f s audio) per step, batch 1" % (len(idx), args.utts, secs)
    emit(({
        "impl": "reference", "metric": "Paraformer-large RTFx (audio s / wall s)", "value": v, "unit": "audio_s/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": min(args.warmup, 1), "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "configs[1]: 1024 synthetic utterances U(2,30) s, Paraformer-large random init",
                   "sample": sample},
        "cpu_baseline": {"value": v, "unit": "audio_s/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "audio_s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


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


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
        return
    from funasr_b200 import _lib, engine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)

    durs, waves = workload(rank, args.utts)
    secs = audio_seconds(durs)
    pinned = [torch.from_numpy(w).pin_memory() for w in waves]
    eng = engine.ParaformerEngine(state_dict(), dev, args.precision, cmvn=cmvn(), max_rows=args.max_rows)
    batches = eng.plan_batches([int(n) for n in durs])
    resident = [eng.upload([pinned[i] for i in b]) for b in batches]
    torch.cuda.synchronize()

    def step_resident():
        n_tok = 0
        for wave, lens in resident:
            ids, _, tok = eng.forward_packed(wave, lens)
            n_tok += len(ids.cpu())
        return n_tok

    def step_e2e():
        out = eng.recognize(pinned)
        return sum(len(x) for x in out)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        barrier()
        if dist is not None:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(args.warmup):
        step_resident()
    sampler = ClockSampler(local) if rank == 0 else None
    _lib.launch_count(reset=True)
    ms_res = timed(step_resident, args.steps)
    launches = _lib.launch_count()
    step_e2e()
    ms_e2e = timed(step_e2e, args.steps)
    clocks = sampler.stop() if sampler else None

    # one extra profiled pass: CUDA events around every kernel entry point
    _lib.profile_read(reset=True)
    _lib.profile_enable(True)
    step_resident()
    torch.cuda.synchronize()
    _lib.profile_enable(False)
    prof = _lib.profile_read(reset=True)

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    pk = peaks()
    value = world * secs * args.steps / (ms_res * 1e-3)
    e2e = world * secs * args.steps / (ms_e2e * 1e-3)
    g = prof.get("gemm", {"ms": 0.0, "calls": 0, "work": 0.0})
    achieved = g["work"] / (g["ms"] * 1e-3) / 1e12 if g["ms"] else 0.0
    terms = {"fp32": 3, "fp32x6": 6, "bf16": 1}[args.precision]
    tot_ms = sum(v["ms"] for v in prof.values()) or 1.0
    h2d = int(sum(int(n) for n in durs) * 4 + sum(3 * (len(b) + 1) * 4 + (len(b) + 1) * 8 for b in batches))
    n_tok = step_resident()
    d2h = int(n_tok * 4 + len(durs) * 4)
    # DRAM bytes per GEMM launch: one ncu pass over the same command, summarised under profiles/ (bytes, or None)
    gemm_traffic, gemm_traffic_src = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "r01_gemm_dram.json")) as f:
            _t = json.load(f)
        gemm_traffic, gemm_traffic_src = _t["dram_bytes_per_launch"], "profiles/r01_gemm_dram.json (ncu dram__bytes_read+write.sum, avg over %d launches)" % _t["launches"]
    except (OSError, KeyError, ValueError):
        pass
    line = {
        "metric": "Paraformer-large RTFx (audio s / wall s)", "value": value, "unit": "audio_s/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_res / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "f32",
        "data": "synthetic",
        "config": {"workload": "configs[1]: %d synthetic utterances U(2,30) s per GPU (%.0f s audio), Paraformer-large "
                               "random init, packed var-len batches <= %d rows" % (args.utts, secs, args.max_rows),
                   "precision": {"fp32": "fp32-parity: operands split hi+mid bf16 (16-bit mantissa), 3 tcgen05 terms, "
                                         "fp32 accumulate; all other kernels fp32",
                                 "fp32x6": "fp32-parity: operands split hi+mid+lo bf16, 6 tcgen05 terms, fp32 accumulate",
                                 "bf16": "bf16 operands, fp32 accumulate"}[args.precision],
                   "batches_per_step": len(batches), "parallelism": "dp%d" % world,
                   "l2": "inputs > L2: %.2f GB of waveforms + %.0f MB of weights per step" % (
                       sum(int(n) for n in durs) * 4 / 1e9, 440.0 * eng.nparts)},
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
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, dt, s, n = cpu_baseline(durs, waves, args.cpu_utts, threads)
        line["cpu_baseline"] = {"value": v, "unit": "audio_s/s", "cores": threads, "kind": "port",
                                "sample": "first %d of the %d utterances (%.1f s audio, %.1f s CPU), batch 1, oracle"
                                          % (n, args.utts, s, dt)}
    emit(line)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
