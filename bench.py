#!/usr/bin/env python
"""bench.py -- sequences folded per second on the BASELINE.json workload.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2|cfg3|cfg4] [--count n]
                  [--scaling weak|strong]
  python bench.py --impl reference ...      # the CPU implementation of the path on the host cores

A "step" is one pass of the folding hot path over one batch of synthetic sequences.
Default workload = BASELINE.json configs[1]: 10,000 synthetic RNAs, 50-500 nt (benchmark
length mix), -n 100 -ms 50 (max_branch 1000, as the reference CLI), final structures only
(the reference's benchmark driver runs `rafft --bench`, without --traj).

  value     sequences/s, inputs already resident in HBM, timed with CUDA events on the streams the
            kernels are launched on (fold graph + scan + pack kernels of every step)
  e2e       the same batch through the C-ABI call rafft_fold_batch with HOST buffers: host->device
            copy of the sequences and device->host copy of the structures inside the timed region
            (wall clock between synchronisations); median / p95 / min of the steps beside the mean
  e2e_python_api   the same batch through rafft_b200.fold_batch (list of str in, lazy per-sequence
            results out) plus reading the best structure of every sequence
  roofline  the kernel family with the largest share of the step (per-family CUDA events, measured
            in this run in the library's profile mode): warp-instructions per second (instruction
            count of the family from the committed ncu capture of the same workload,
            profiles/r02_kernel_counters.json) against the issue-slot peak measured in this run by
            the library's IADD3 microbenchmark; real DRAM bytes against the measured HBM peak beside it
  N > 1     launched by torchrun, one rank per GPU, no data-path collective.  --scaling weak (the
            default line): N x the 1-GPU batch.  The line of an N > 1 run also carries "strong":
            the SAME 10k batch of configs[1] and configs[2] partitioned over the N GPUs (BASELINE
            configs[2] as written), with the N = 1 time of the same batch measured on rank 0 in the
            same run, and a per-rank parity sample against the oracle.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "cfg2": dict(gen="cfg2", count=10000, nb_mode=100, max_stack=50,
                 desc="10k synthetic RNAs, 50-500 nt (benchmark length mix), -n 100 -ms 50"),
    "cfg3": dict(gen="cfg2", count=10000, nb_mode=200, max_stack=200,
                 desc="10k synthetic RNAs, 50-500 nt (benchmark length mix), -n 200 -ms 200"),
    "cfg4": dict(gen="cfg4", count=1000, nb_mode=200, max_stack=200,
                 desc="1k synthetic 3,000-nt RNAs, -n 200 -ms 200"),
}
COUNTERS = os.path.join(ROOT, "profiles", "r02_kernel_counters.json")
METRIC = "sequences folded/sec"
UNIT = "seq/s"


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    p.add_argument("--count", type=int, default=None, help="sequences per GPU (weak) / in total (strong)")
    p.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    p.add_argument("--cpu-sample", type=int, default=None, help="sequences of the CPU baseline sample")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-strong", action="store_true", help="N > 1: skip the strong-scaling sub-record")
    p.add_argument("--parity-sample", type=int, default=64, help="sequences per rank checked against the oracle")
    return p.parse_args()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def cpu_sample(seqs, want):
    "stratified subsample: every k-th sequence of the batch"
    want = max(1, min(want, len(seqs)))
    k = max(1, len(seqs) // want)
    return seqs[::k][:want], k


def cpu_sample_for(pyoracle, seqs, wl, cores, target_s):
    """Stratified sample of the workload that keeps the CPU restatement busy for about target_s
    seconds on `cores` threads (a 64-sequence pilot measures the rate first)."""
    pilot, _ = cpu_sample(seqs, 64)
    t0 = time.perf_counter()
    pyoracle.fold_batch(pilot, wl["nb_mode"], wl["max_stack"], 1000, nthreads=cores, want_text=False)
    rate = len(pilot) / max(time.perf_counter() - t0, 1e-3)
    return cpu_sample(seqs, int(rate * target_s))


def algorithmic_bytes(counters, lengths, which="n_eval"):
    """one int16 pair table read per evaluated structure, the position list of
    every expanded region (reference granularity)."""
    import numpy as np
    from rafft_b200._native import COUNTER_NAMES as CN
    L = np.asarray(lengths, dtype=np.int64)
    return counters[:, CN.index(which)] * 2 * (L + 1) + counters[:, CN.index("node_pos")] * 2


def algorithmic_intops(counters, lengths, which="n_eval"):
    import numpy as np
    from rafft_b200._native import COUNTER_NAMES as CN
    L = np.asarray(lengths, dtype=np.int64)
    return counters[:, CN.index("w_corr")] + counters[:, CN.index("w_scan")] + counters[:, CN.index(which)] * (L + 1)


class ClockSampler:
    "nvidia-smi clocks and throttle reasons DURING the timed region"
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.path)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


def run_reference(args, wl):
    """The reference's CPU implementation of the path on the host cores.  The reference itself
    (Python + ViennaRNA) cannot run on this box, so this arm times the
    C++ restatement (oracle/, kind "port") with all host threads, on a bounded stratified sample
    of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    from rafft_b200.synth import config_batch
    seqs = config_batch(wl["gen"], (args.count or wl["count"]))
    cores = os.cpu_count() or 1
    if args.cpu_sample:
        sample, k = cpu_sample(seqs, args.cpu_sample)
    else:  # each timed step ~12 s of host work, so that K steps + warm-up end within minutes
        sample, k = cpu_sample_for(pyoracle, seqs, wl, cores, 12.0)
    for _ in range(args.warmup):
        pyoracle.fold_batch(sample[:max(1, len(sample) // 8)], wl["nb_mode"], wl["max_stack"], 1000, nthreads=cores,
                            want_text=False)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        pyoracle.fold_batch(sample, wl["nb_mode"], wl["max_stack"], 1000, nthreads=cores, want_text=False)
    dt = time.perf_counter() - t0
    v = len(sample) * args.steps / dt
    sdesc = "every %dth sequence of the batch (%d sequences, %d nt) per step" % (k, len(sample), sum(map(len, sample)))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": {"workload": wl["desc"], "nb_mode": wl["nb_mode"], "max_stack": wl["max_stack"],
                   "max_branch": 1000, "traj": False, "sample": sdesc},
        "nt_per_s": sum(map(len, sample)) * args.steps / dt,
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sdesc},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


class Shard:
    "this rank's sequences of a batch, as host bytes and as device buffers"

    def __init__(self, seqs, dev, torch, np):
        self.seqs = seqs
        self.n = len(seqs)
        self.lengths = np.array([len(s) for s in seqs], dtype=np.int64)
        self.offs = np.zeros(self.n + 1, dtype=np.int64)
        np.cumsum(self.lengths, out=self.offs[1:])
        self.bytes = "".join(seqs).encode()
        self.d_seqs = torch.frombuffer(bytearray(self.bytes or b"\0"), dtype=torch.uint8).to(dev)
        self.d_offs = torch.from_numpy(self.offs).to(dev)


def stats_ms(xs):
    xs = sorted(xs)
    if not xs:
        return {}
    p95 = xs[min(len(xs) - 1, int(round(0.95 * (len(xs) - 1))))]
    return {"median_ms": statistics.median(xs), "p95_ms": p95, "min_ms": xs[0], "max_ms": xs[-1]}


def run_b200(args, wl):
    import numpy as np
    import torch
    from rafft_b200 import _native as N
    from rafft_b200 import rafft as R
    from rafft_b200.shard import imbalance, partition
    from rafft_b200.synth import config_batch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch N > 1 with torchrun (one rank per GPU)")
    dist = None
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"  # the version banner goes to stdout, which carries the JSON line
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    ctx = R.FoldContext(local)   # fails loudly without the CUDA library / a GPU
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier(collective=True):
        torch.cuda.synchronize()
        if dist is not None and collective:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(*xs):
        t = torch.tensor(list(xs), dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(x) for x in t.cpu()]

    def allsum(*xs):
        t = torch.tensor(list(xs), dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return [float(x) for x in t.cpu()]

    def params_of(w):
        return R.make_params(w["nb_mode"], w["max_stack"], 1000, 3, 0.0, False, 37.0, 3.0, 2.0, 1.0)

    def step_device(sh, prm):
        flush.zero_()
        torch.cuda.synchronize()
        return ctx.fold_raw_device(sh.d_seqs.data_ptr(), sh.d_offs.data_ptr(), sh.n, int(sh.offs[-1]),
                                   int(sh.lengths.max()) if sh.n else 0, prm)

    def step_host(sh, prm):
        flush.zero_()
        torch.cuda.synchronize()
        t = time.perf_counter()
        raw = ctx.fold_raw(sh.bytes, sh.offs, prm)
        return time.perf_counter() - t, raw

    def parity_sample(sh, w, want):
        "a stratified sample of this rank's shard folded again and compared with the oracle"
        if want <= 0 or sh.n == 0:
            return 0, True
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import pyoracle  # checker only, after the timed region
        k = max(1, sh.n // want)
        sample = sh.seqs[::k][:want]
        offs = np.zeros(len(sample) + 1, dtype=np.int64)
        np.cumsum([len(s) for s in sample], out=offs[1:])
        raw = ctx.fold_raw("".join(sample).encode(), offs, params_of(w))
        bad, _ = pyoracle.fold_batch_check(sample, raw, w["nb_mode"], w["max_stack"], 1000, traj=False,
                                           nthreads=max(1, (os.cpu_count() or 1) // max(1, world)))
        ok = not bad and bool((raw.status == 0).all())
        raw.close()
        return len(sample), ok

    def measure(sh, w, steps, warmup, want_e2e=True, collective=True):
        """device-resident and end-to-end legs of one shard; returns per-rank numbers (callers reduce).
        collective=False: only this rank runs it (no barrier across ranks)"""
        sync = lambda: barrier(collective)  # noqa: E731
        prm = params_of(w)
        raw = None
        for _ in range(warmup):
            if raw is not None:
                raw.close()
            raw = step_device(sh, prm)
        assert (raw.status == 0).all(), "a sequence failed: %s" % np.bincount(raw.status)
        counters = raw.counters.copy()  # views die with the result
        lanes, loop_mode = raw.lanes, raw.loop_mode
        raw.close()
        sync()
        wall0 = time.perf_counter()
        kernel_ms, launches, iters = 0.0, 0, 0
        dev_steps = []
        for _ in range(steps):
            raw = step_device(sh, prm)
            kernel_ms += raw.kernel_ms
            dev_steps.append(raw.kernel_ms)
            launches += raw.kernel_launches
            iters = max(iters, raw.iterations)
            raw.close()
        sync()
        wall_dev = time.perf_counter() - wall0
        out = dict(kernel_ms=kernel_ms, launches=launches, iterations=iters, wall_dev=wall_dev, counters=counters,
                   lanes=lanes, loop_mode=loop_mode, dev_steps=dev_steps, e2e_s=0.0, e2e_steps=[], h2d=0, d2h=0)
        if want_e2e:
            for _ in range(min(warmup, 2)):
                step_host(sh, prm)[1].close()
            sync()
            for _ in range(steps):
                dt, raw = step_host(sh, prm)
                out["e2e_s"] += dt
                out["e2e_steps"].append(dt * 1e3)
                out["h2d"], out["d2h"] = raw.h2d_bytes, raw.d2h_bytes
                raw.close()
            sync()
        return out

    per_gpu = args.count or wl["count"]
    strong = args.scaling == "strong"
    n_total = per_gpu if strong else per_gpu * world
    seqs_all = config_batch(wl["gen"], n_total)
    shards = partition([len(s) for s in seqs_all], world, wl["max_stack"])
    mine = Shard([seqs_all[int(i)] for i in shards[rank]], dev, torch, np)

    sampler = ClockSampler(local)
    sampler.start()
    m = measure(mine, wl, args.steps, args.warmup)
    clocks = sampler.stop()
    kernel_ms_max, e2e_max, wall_dev_max = allmax(m["kernel_ms"], m["e2e_s"], m["wall_dev"])
    h2d_all, d2h_all, launches_all = [int(x) for x in allsum(m["h2d"], m["d2h"], m["launches"])]
    par_n, par_ok = parity_sample(mine, wl, args.parity_sample)
    par_n_all, par_bad_all = [int(x) for x in allsum(par_n, 0 if par_ok else 1)]

    # ---- the drop-in Python API on the same batch (rank 0's shard, N = 1 only: it is a host-side figure)
    py_api = None
    if world == 1:
        prm_kw = dict(nb_mode=wl["nb_mode"], max_stack=wl["max_stack"], max_branch=1000, context=ctx)
        ts, tb = [], []
        for k in range(min(args.steps, 5) + 1):
            flush.zero_()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            res = R.fold_batch(mine.seqs, **prm_kw)
            t1 = time.perf_counter()
            dbs, es = res.best()
            t2 = time.perf_counter()
            res.raw.close()
            if k:  # the first call warms the pinned block of this shape
                ts.append((t1 - t0) * 1e3)
                tb.append((t2 - t0) * 1e3)
        py_api = {"call": "rafft_b200.fold_batch(list of str) -> lazy per-sequence results", "value": mine.n / (statistics.mean(ts) / 1e3),
                  "unit": UNIT, "ms_per_step": statistics.mean(ts),
                  "with_best_structures": {"value": mine.n / (statistics.mean(tb) / 1e3), "ms_per_step": statistics.mean(tb),
                                           "what": "plus BatchFolds.best(): dot-bracket + energy of the lowest-energy structure of every sequence"}}

    # ---- roofline of the dominant kernel family: per-family CUDA events (profile mode), this run
    roof = None
    if rank == 0:
        ctx.set_option("profile", 1)
        prm = params_of(wl)
        step_device(mine, prm).close()
        fam_ms = {k: 0.0 for k in N.FAMILY_NAMES}
        fam_n = {k: 0 for k in N.FAMILY_NAMES}
        psteps = max(1, min(args.steps, 3))
        prof_ms = 0.0
        for _ in range(psteps):
            raw = step_device(mine, prm)
            for k in N.FAMILY_NAMES:
                fam_ms[k] += raw.family_ms[k]
                fam_n[k] += raw.family_launches[k]
            prof_ms += raw.kernel_ms
            raw.close()
        ctx.set_option("profile", 0)
        issue_peak, mix_peak = ctx.int_peak()
        hbm_peak, hbm_src = peaks()
        tot = sum(fam_ms.values()) or 1e-9
        dom = max(fam_ms, key=lambda k: fam_ms[k])
        fam_step_s = fam_ms[dom] / psteps / 1e3
        launches_step = fam_n[dom] / psteps
        cnt = {}
        try:
            cnt = json.load(open(COUNTERS)).get(args.workload if per_gpu == wl["count"] else "", {}).get(dom, {})
        except Exception:
            cnt = {}
        inst = cnt.get("warp_inst_per_step")
        dram = cnt.get("dram_bytes_per_step")
        exp_bytes = int(algorithmic_bytes(m["counters"], mine.lengths, "n_eval_expand").sum())
        step_bytes = int(algorithmic_bytes(m["counters"], mine.lengths).sum())
        roof = {
            "bound": "issue", "kernel": dom + " (" + ", ".join(N.FAMILY_KERNELS[dom]) + ")",
            "achieved": inst / fam_step_s / 1e9 if inst else None,
            "peak": issue_peak / 1e9, "unit": "Gwarp-inst/s",
            "frac": inst / fam_step_s / issue_peak if inst else None,
            "peak_source": "measured in this run: rafft_int_peak (independent IADD3 chains, every scheduler issuing every cycle); "
                           "IADD3/LOP3 mix %.0f Gwarp-inst/s" % (mix_peak / 1e9),
            "traffic": dram / launches_step if dram and launches_step else None,
            "lanes_active": cnt.get("threads_per_inst"),
            "launches_per_step": launches_step, "avg_launch_ms": fam_step_s * 1e3 / max(launches_step, 1),
            "share_of_step": fam_ms[dom] / tot,
            "family_ms_per_step": {k: v / psteps for k, v in fam_ms.items()},
            "counters_source": (cnt.get("source") if cnt else "no ncu capture committed for this workload / count"),
            "hbm": {"achieved": dram / fam_step_s / 1e9 if dram else None, "peak": hbm_peak, "unit": "GB/s",
                    "frac": dram / fam_step_s / 1e9 / hbm_peak if dram else None, "peak_source": hbm_src,
                    "note": "real DRAM bytes of the family (ncu dram__bytes_read + write) over its event-timed duration"},
            "work_avoided": {"algorithmic_bytes_per_step_family": exp_bytes if dom.startswith("expand") else None,
                             "algorithmic_bytes_per_step": step_bytes,
                             "ratio_vs_dram": (exp_bytes / dram if dram and dom.startswith("expand") else None),
                             "note": "bytes the REFERENCE formulation would move (one pair table per evaluated stem + region "
                                     "position lists, SURVEY 8d) over the bytes this implementation moves"},
            "profile_mode_ms_per_step": prof_ms / psteps,
        }

    # ---- N > 1: BASELINE configs[2] as written -- the same 10k batch partitioned over the GPUs
    strong_rec = None
    if world > 1 and not args.no_strong and not strong:
        strong_rec = {}
        for name in ("cfg2", "cfg3"):
            w = WORKLOADS[name]
            batch = config_batch(w["gen"], w["count"])
            lens = [len(s) for s in batch]
            parts = partition(lens, world, w["max_stack"])
            sh = Shard([batch[int(i)] for i in parts[rank]], dev, torch, np)
            steps = max(3, min(args.steps, 10))
            r = measure(sh, w, steps, 3)
            k_max, e_max = allmax(r["kernel_ms"], r["e2e_s"])
            pn, pok = parity_sample(sh, w, args.parity_sample)
            pn_all, pbad_all = [int(x) for x in allsum(pn, 0 if pok else 1)]
            # the N = 1 figure of the same batch, on rank 0's GPU in this run (the other ranks wait)
            one_ms = 0.0
            if rank == 0:
                whole = Shard(batch, dev, torch, np)
                r1 = measure(whole, w, steps, 3, want_e2e=False, collective=False)
                one_ms = r1["kernel_ms"] / steps
                del whole
            barrier()
            one_ms = allmax(one_ms)[0]
            strong_rec[name] = {
                "workload": w["desc"] + ", the same batch partitioned over %d GPUs" % world,
                "value": w["count"] * steps / (k_max / 1e3), "unit": UNIT, "ms_per_step": k_max / steps,
                "e2e": {"value": w["count"] * steps / e_max, "ms_per_step": e_max / steps * 1e3},
                "n1_ms_per_step_same_run": one_ms, "efficiency_vs_n1": one_ms / (k_max / steps) / world,
                "shard_imbalance": imbalance(lens, parts, w["max_stack"]),
                "sequences_per_gpu": [len(p) for p in parts], "iterations": r["iterations"],
                "parity_checked": pn_all, "parity_ok": pbad_all == 0,
            }

    line = None
    if rank == 0:
        total_nt = sum(len(s) for s in seqs_all)
        e2e_ms = m["e2e_steps"]
        line = {
            "metric": METRIC, "value": n_total * args.steps / (kernel_ms_max / 1e3), "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": kernel_ms_max / args.steps, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": wl["desc"] + ((" x %d (weak scaling)" % world if not strong else
                                                  ", one batch partitioned over %d GPUs (strong scaling)" % world) if world > 1 else ""),
                       "sequences": n_total, "nt": total_nt, "nb_mode": wl["nb_mode"],
                       "max_stack": wl["max_stack"], "max_branch": 1000, "traj": False,
                       "parallelism": "sequences sharded over %d GPU(s), no collective" % world,
                       "shard_imbalance": imbalance([len(s) for s in seqs_all], shards, wl["max_stack"]),
                       "l2": "256 MiB flush buffer written between timed steps",
                       "waves": "all sequences resident (one arena each)",
                       "loop": {0: "one CUDA-graph launch per fold (device-side WHILE node)", 1: "stream launches, look-ahead 3",
                                2: "stepwise"}[m["loop_mode"]], "lanes": m["lanes"], "iterations": m["iterations"]},
            "nt_per_s": total_nt * args.steps / (kernel_ms_max / 1e3),
            "device_ms_steps_rank0": [round(x, 3) for x in m["dev_steps"]],
            "wall_ms_per_step_device_inputs": wall_dev_max / args.steps * 1e3,
            "e2e": dict({"value": n_total * args.steps / e2e_max, "unit": UNIT,
                         "h2d_bytes_per_step": h2d_all, "d2h_bytes_per_step": d2h_all,
                         "ms_per_step": e2e_max / args.steps * 1e3,
                         "ms_steps_rank0": [round(x, 3) for x in e2e_ms]}, **stats_ms(e2e_ms)),
            "gpu_launches": launches_all,
            "clocks": clocks,
            "parity_checked": par_n_all, "parity_ok": par_bad_all == 0,
            "roofline": roof,
        }
        if py_api is not None:
            line["e2e_python_api"] = py_api
        if strong_rec is not None:
            line["strong"] = strong_rec
        if not args.no_cpu_baseline and world == 1:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import pyoracle  # CPU baseline leg only
            cores = os.cpu_count() or 1
            if args.cpu_sample:
                sample, k = cpu_sample(seqs_all, args.cpu_sample)
            else:  # about 25 s of host work
                sample, k = cpu_sample_for(pyoracle, seqs_all, wl, cores, 25.0)
            t0 = time.perf_counter()
            pyoracle.fold_batch(sample, wl["nb_mode"], wl["max_stack"], 1000, nthreads=cores, want_text=False)
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": len(sample) / dt, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": "every %dth sequence of the batch (%d sequences), %.1f s"
                                              % (k, len(sample), dt)}
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line))


def main():
    args = parse_args()
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_b200(args, wl)


if __name__ == "__main__":
    main()
