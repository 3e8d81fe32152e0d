"""B1 "reference-python" (BASELINE.md section 3): the UNMODIFIED /root/reference/rafft (rafft.rafft.fold)
with the ViennaRNA calls served by oracle/shim, fanned out with multiprocessing.Pool(P) over the
sequences as benchmark_results/bench_fft.py:17-21 does, on the stratified subsample of the cfg 2 / cfg 3
batch (every 39th sequence -> 256 sequences).  Build container only (/root/reference is not on the
GPU box).  usage: time_reference_python.py cfg2|cfg3 [stride] [P]"""
import json, os, sys, time
from multiprocessing import Pool
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.append(ROOT)                      # rafft_b200.synth only; the repo's own `rafft` alias must NOT win:
sys.path.insert(0, "/root/reference")      # `rafft` = the reference package
sys.path.insert(0, os.path.join(ROOT, "oracle", "shim"))
wl = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
stride = int(sys.argv[2]) if len(sys.argv) > 2 else 39
P = int(sys.argv[3]) if len(sys.argv) > 3 else (os.cpu_count() or 1)
n, ms = (100, 50) if wl == "cfg2" else (200, 200)


def run(seq):
    from rafft.rafft import fold
    assert "/root/reference" in sys.modules["rafft"].__file__
    t = time.perf_counter()
    out = fold(seq, n, ms, 1000, 3, 0.0, False, 37.0, 3.0, 2.0, 1.0)
    return time.perf_counter() - t, len(out)


if __name__ == "__main__":
    from rafft_b200.synth import config_batch
    seqs = config_batch("cfg2")[::stride]
    t0 = time.perf_counter()
    with Pool(P) as pool:
        res = pool.map(run, seqs, chunksize=1)
    dt = time.perf_counter() - t0
    rec = {"workload": wl, "nb_mode": n, "max_stack": ms, "sequences": len(seqs), "nt": sum(map(len, seqs)), "processes": P,
           "wall_s": dt, "seq_per_s": len(seqs) / dt, "nt_per_s": sum(map(len, seqs)) / dt,
           "cpu_s_per_sequence_mean": sum(r[0] for r in res) / len(res), "sample": "every %dth sequence of the 10k batch" % stride,
           "what": "unmodified /root/reference rafft.rafft.fold, RNA module = oracle/shim (C evaluator), multiprocessing.Pool"}
    print(json.dumps(rec))
