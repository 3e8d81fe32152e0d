"""ncu per-launch metrics CSV of a 1-step bench run -> per-kernel-family counters of ONE fold, merged
into profiles/r02_kernel_counters.json (what bench.py's roofline reads).

usage: ncu_counters.py <metrics.csv> <workload> [fold-index] [--source text]
The CSV comes from
  RAFFT_LOOP=stream RAFFT_LANES=1 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,\
smsp__thread_inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 900 --csv \
  --log-file <metrics.csv> python bench.py --workload <workload> --steps 1 --warmup 1 --no-cpu-baseline ...
(plain stream launches so that every kernel is a launch ncu lists; same kernels as the graph loop)."""
import csv, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rafft_b200._native import FAMILY_KERNELS

path, workload = sys.argv[1], sys.argv[2]
fold = int(sys.argv[3]) if len(sys.argv) > 3 and not sys.argv[3].startswith("--") else 1
source = sys.argv[sys.argv.index("--source") + 1] if "--source" in sys.argv else os.path.basename(path)
rows = [r for r in csv.reader(open(path)) if len(r) > 10]
H = rows[0]
ki, mi, ui, vi, ii = (H.index(x) for x in ("Kernel Name", "Metric Name", "Metric Unit", "Metric Value", "ID"))
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1.0, "us": 1e3, "ms": 1e6, "inst": 1.0}
launch = {}
for r in rows[1:]:
    name = r[ki].split("(")[0].replace("void ", "").split("<")[0]
    d = launch.setdefault(int(r[ii]), {"k": name})
    d[r[mi]] = float(r[vi].replace(",", "")) * scale.get(r[ui], 1.0)
seq = [launch[k] for k in sorted(launch)]
inits = [i for i, s in enumerate(seq) if s["k"] == "k_init"] + [len(seq)]
part = seq[inits[fold]:inits[fold + 1]]
fam_of = {k: f for f, ks in FAMILY_KERNELS.items() for k in ks}
out = {}
for s in part:
    f = fam_of.get(s["k"])
    if f is None:
        continue
    o = out.setdefault(f, {"launches": 0, "warp_inst_per_step": 0.0, "thread_inst": 0.0, "dram_bytes_per_step": 0.0, "ncu_time_ms": 0.0})
    o["launches"] += 1
    o["warp_inst_per_step"] += s.get("smsp__inst_executed.sum", 0.0)
    o["thread_inst"] += s.get("smsp__thread_inst_executed.sum", 0.0)
    o["dram_bytes_per_step"] += s.get("dram__bytes_read.sum", 0.0) + s.get("dram__bytes_write.sum", 0.0)
    o["ncu_time_ms"] += s.get("gpu__time_duration.sum", 0.0) / 1e6
tot = sum(o["ncu_time_ms"] for o in out.values()) or 1.0
for f, o in out.items():
    o["threads_per_inst"] = o.pop("thread_inst") / max(o["warp_inst_per_step"], 1.0)
    o["ncu_share_of_fold"] = o["ncu_time_ms"] / tot
    o["source"] = source
dst = os.path.join(ROOT, "profiles", "r02_kernel_counters.json")
db = json.load(open(dst)) if os.path.exists(dst) else {}
db[workload] = out
json.dump(db, open(dst, "w"), indent=1, sort_keys=True)
print("%-16s %8s %14s %10s %12s %10s %8s" % ("family", "launches", "warp inst", "lanes", "DRAM MB", "ncu ms", "share"))
for f, o in sorted(out.items(), key=lambda kv: -kv[1]["ncu_time_ms"]):
    print("%-16s %8d %14.3e %10.1f %12.1f %10.3f %7.1f%%" % (f, o["launches"], o["warp_inst_per_step"], o["threads_per_inst"],
                                                          o["dram_bytes_per_step"] / 1e6, o["ncu_time_ms"], 100 * o["ncu_share_of_fold"]))
