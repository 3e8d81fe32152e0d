"""Synthetic inputs of the benchmark configurations."""
import json
import os

import numpy as np

_M = (1 << 64) - 1
HERE = os.path.dirname(os.path.abspath(__file__))


def _splitmix_stream(seed, count):
    "vectorised splitmix64: the `count` outputs following state `seed`"
    with np.errstate(over="ignore"):
        k = np.arange(1, count + 1, dtype=np.uint64)
        z = np.uint64(seed & _M) + k * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def splitmix_seq(seed, length):
    'base k = "ACGU"[splitmix64 >> 62], iid uniform'
    z = _splitmix_stream(seed, length)
    return np.frombuffer(b"ACGU", dtype=np.uint8)[(z >> np.uint64(62)).astype(np.int64)].tobytes().decode()


def benchmark_lengths():
    """sorted multiset of the 2,248 reference sequence lengths in [50, 500]
    (benchmark_results/benchmark_cleaned_all_length.csv; extracted by tools/make_lengths.py)"""
    with open(os.path.join(HERE, "data", "bench_lengths.json")) as fh:
        return json.load(fh)["lengths"]


def config_batch(name, count=None):
    """Sequences of a BASELINE.json configuration: 'cfg1' (one 120-nt), 'cfg2'/'cfg3' (10k,
    50-500 nt benchmark length mix), 'cfg4' (1k x 3000 nt), 'cfg5' (one 200-nt)."""
    if name == "cfg1":
        return [splitmix_seq(1, 120)]
    if name == "cfg5":
        return [splitmix_seq(200, 200)]
    if name in ("cfg2", "cfg3"):
        n = 10000 if count is None else count
        pool = benchmark_lengths()
        idx = (_splitmix_stream(20240001, n) % np.uint64(len(pool))).astype(np.int64)
        return [splitmix_seq(20240002 + k, pool[int(i)]) for k, i in enumerate(idx)]
    if name == "cfg4":
        n = 1000 if count is None else count
        return [splitmix_seq(3000 + k, 3000) for k in range(n)]
    raise ValueError(name)
