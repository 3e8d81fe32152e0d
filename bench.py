#!/usr/bin/env python
"""bench.py -- sequences folded per second on the BASELINE.json workload.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2|cfg3|cfg4] [--count n]
  python bench.py --impl reference ...      # the CPU implementation of the path on the host cores

A "step" is one pass of the folding hot path over one batch of synthetic sequences.
Default workload = BASELINE.json configs[1]: 10,000 synthetic RNAs, 50-500 nt (benchmark
length mix), -n 100 -ms 50 (max_branch 1000, as the reference CLI), final structures only
(the reference's benchmark driver runs `rafft --bench`, without --traj).

  value   sequences/s, inputs already resident in HBM, timed with CUDA events on the stream
          the kernels are launched on (order + fold + scan + pack kernels of every step)
  e2e     the same batch through the C-ABI call rafft_fold_batch with HOST buffers:
          host->device copy of the sequences and device->host copy of the structures inside
          the timed region (wall clock between synchronisations)
  N > 1   launched by torchrun, one rank per GPU; the global batch is N x the 1-GPU batch
          (weak scaling), partitioned by sequence with no data-path collective; times are
          the MAX over ranks
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
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
# dram__bytes_read.sum + dram__bytes_write.sum of k_expand_fast per launch (average over the launches
# of one step) from the ncu --set full capture under profiles/ (null until one exists for the workload)
TRAFFIC = {"cfg2": 51.7e6}  # profiles/r01_dram_v13.txt: 672 MB over the 13 launches of one fold
METRIC = "sequences folded/sec"
UNIT = "seq/s"


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    p.add_argument("--count", type=int, default=None, help="sequences per GPU (default: the config's)")
    p.add_argument("--cpu-sample", type=int, default=None, help="sequences of the CPU baseline sample")
    p.add_argument("--no-cpu-baseline", action="store_true")
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
    every expanded region (reference granularity).  which = "n_eval": the whole path;
    "n_eval_expand": the share of the region expansion (find_best_consecutives, rafft.py:98)."""
    import numpy as np
    L = np.asarray(lengths, dtype=np.int64)
    c = counters
    from rafft_b200._native import COUNTER_NAMES as CN
    n_eval, node_pos = c[:, CN.index(which)], c[:, CN.index("node_pos")]
    return n_eval * 2 * (L + 1) + node_pos * 2


def algorithmic_intops(counters, lengths, which="n_eval"):
    import numpy as np
    from rafft_b200._native import COUNTER_NAMES as CN
    L = np.asarray(lengths, dtype=np.int64)
    c = counters
    return c[:, CN.index("w_corr")] + c[:, CN.index("w_scan")] + c[:, CN.index(which)] * (L + 1)


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
    else:  # each timed step ~8 s of host work, so that K steps + warm-up end within minutes
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


def run_b200(args, wl):
    import numpy as np
    import torch
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
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    ctx = R.FoldContext(local)   # fails loudly without the CUDA library / a GPU

    per_gpu = args.count or wl["count"]
    seqs_all = config_batch(wl["gen"], per_gpu * world)
    shards = partition([len(s) for s in seqs_all], world, wl["max_stack"])
    mine = [seqs_all[int(i)] for i in shards[rank]]
    n_local, n_total = len(mine), len(seqs_all)
    lengths = np.array([len(s) for s in mine], dtype=np.int64)
    offs = np.zeros(n_local + 1, dtype=np.int64)
    np.cumsum(lengths, out=offs[1:])
    seq_bytes = "".join(mine).encode()
    params = R.make_params(wl["nb_mode"], wl["max_stack"], 1000, 3, 0.0, False, 37.0, 3.0, 2.0, 1.0)

    # inputs resident in HBM (torch only provides the device memory)
    d_seqs = torch.frombuffer(bytearray(seq_bytes), dtype=torch.uint8).to(dev)
    d_offs = torch.from_numpy(offs).to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device():
        flush.zero_()
        torch.cuda.synchronize()
        return ctx.fold_raw_device(d_seqs.data_ptr(), d_offs.data_ptr(), n_local, int(offs[-1]),
                                   int(lengths.max()), params)

    def step_host():
        flush.zero_()
        torch.cuda.synchronize()
        t = time.perf_counter()
        raw = ctx.fold_raw(seq_bytes, offs, params)
        return time.perf_counter() - t, raw

    for _ in range(args.warmup):
        raw = step_device()
    assert (raw.status == 0).all(), "a sequence failed: %s" % np.bincount(raw.status)
    counters = raw.counters.copy()  # views die with the result

    # ---- device-resident leg: CUDA events inside the library, on its launch stream
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    wall0 = time.perf_counter()
    kernel_ms, launches, expand_ms, expand_launches = 0.0, 0, 0.0, 0
    for _ in range(args.steps):
        raw = step_device()
        kernel_ms += raw.kernel_ms
        launches += raw.kernel_launches
        expand_ms += raw.expand_ms
        expand_launches += raw.expand_launches
    barrier()
    wall_dev = time.perf_counter() - wall0
    clocks = sampler.stop()

    # ---- end-to-end leg: host buffers through the C-ABI, copies inside the timed region
    for _ in range(min(args.warmup, 2)):
        step_host()
    barrier()
    e2e_s, h2d, d2h, e2e_steps = 0.0, 0, 0, []
    for _ in range(args.steps):
        dt, raw = step_host()
        e2e_s += dt
        e2e_steps.append(round(dt * 1e3, 3))
        h2d, d2h = raw.h2d_bytes, raw.d2h_bytes
    barrier()

    t = torch.tensor([kernel_ms, e2e_s, wall_dev], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(h2d), float(d2h), float(launches)], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    kernel_ms_max, e2e_max, wall_dev_max = [float(x) for x in t.cpu()]
    h2d_all, d2h_all, launches_all = [int(x) for x in tot.cpu()]

    # roofline of the dominant kernel, k_expand_fast (region expansion; ~60% of the step): its
    # algorithmic bytes/int-ops (the expansion share of the reference formulation)
    # over its own launch durations, CUDA events around every launch inside the library
    peak, peak_src = peaks()
    step_bytes = int(algorithmic_bytes(counters, lengths).sum()) + int(raw.d2h_bytes)
    step_iops = int(algorithmic_intops(counters, lengths).sum())
    exp_bytes = int(algorithmic_bytes(counters, lengths, "n_eval_expand").sum())
    exp_iops = int(algorithmic_intops(counters, lengths, "n_eval_expand").sum())
    n_exp = max(1, expand_launches)
    exp_launch_s = max(expand_ms, 1e-6) / n_exp / 1e3          # average launch duration
    exp_bytes_launch = exp_bytes * args.steps / n_exp           # algorithmic bytes per launch
    exp_iops_launch = exp_iops * args.steps / n_exp
    achieved = exp_bytes_launch / exp_launch_s / 1e9
    per_step_s = kernel_ms / args.steps / 1e3

    line = None
    if rank == 0:
        total_nt = sum(len(s) for s in seqs_all)
        line = {
            "metric": METRIC, "value": n_total * args.steps / (kernel_ms_max / 1e3), "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": kernel_ms_max / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": wl["desc"] + (" x %d (weak scaling)" % world if world > 1 else ""),
                       "sequences": n_total, "nt": total_nt, "nb_mode": wl["nb_mode"],
                       "max_stack": wl["max_stack"], "max_branch": 1000, "traj": False,
                       "parallelism": "sequences sharded over %d GPU(s), no collective" % world,
                       "shard_imbalance": imbalance([len(s) for s in seqs_all], shards, wl["max_stack"]),
                       "l2": "256 MiB flush buffer written between timed steps",
                       "waves": "all sequences resident (one arena each)"},
            "nt_per_s": total_nt * args.steps / (kernel_ms_max / 1e3),
            "wall_ms_per_step_device_inputs": wall_dev_max / args.steps * 1e3,
            "e2e": {"value": n_total * args.steps / e2e_max, "unit": UNIT,
                    "h2d_bytes_per_step": h2d_all, "d2h_bytes_per_step": d2h_all,
                    "ms_per_step": e2e_max / args.steps * 1e3, "ms_steps_rank0": e2e_steps},
            "gpu_launches": launches_all,
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": TRAFFIC.get(args.workload),
                         "peak_source": peak_src, "kernel": "k_expand_fast",
                         "launches_per_step": n_exp / args.steps,
                         "avg_launch_ms": exp_launch_s * 1e3,
                         "share_of_step": expand_ms / max(kernel_ms, 1e-9),
                         "algorithmic_bytes_per_launch": exp_bytes_launch,
                         "note": "algorithmic bytes = the reference formulation's traffic (one pair table per "
                                 "evaluated stem + region position lists); the kernel is INT-issue bound, see alu",
                         "alu": {"algorithmic_int_ops_per_launch": exp_iops_launch,
                                 "achieved_Giops": exp_iops_launch / exp_launch_s / 1e9,
                                 "nominal_peak_Giops": 148 * 128 * 1.965,
                                 "frac_of_nominal": exp_iops_launch / exp_launch_s / 1e9 / (148 * 128 * 1.965)},
                         "whole_step": {"algorithmic_bytes": step_bytes, "achieved": step_bytes / per_step_s / 1e9,
                                        "frac": step_bytes / per_step_s / 1e9 / peak,
                                        "algorithmic_int_ops": step_iops}},
        }
        if not args.no_cpu_baseline:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import pyoracle  # CPU baseline leg only
            cores = os.cpu_count() or 1
            if args.cpu_sample:
                sample, k = cpu_sample(seqs_all, args.cpu_sample)
            else:  # about 15 s of host work
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
