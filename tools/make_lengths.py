"""Extract the benchmark length multiset (50 <= L <= 500) from the reference's dataset
(benchmark_results/benchmark_cleaned_all_length.csv) -> rafft_b200/data/bench_lengths.json.
Build container only; the GPU box reads the committed json."""
import csv, json, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rows = list(csv.reader(open("/root/reference/benchmark_results/benchmark_cleaned_all_length.csv")))
lens = sorted(len(r[0]) for r in rows if 50 <= len(r[0]) <= 500)  # columns: seq, struct, name (no header)
json.dump({"source": "benchmark_results/benchmark_cleaned_all_length.csv, 50<=L<=500", "lengths": lens},
          open(os.path.join(ROOT, "rafft_b200", "data", "bench_lengths.json"), "w"))
print(len(lens), sum(lens) / len(lens), lens[len(lens) // 2])
