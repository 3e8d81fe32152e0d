#!/bin/bash
# round-2 probe 3: full GPU test tier (full-batch parity), bench line, lane split sweep, ncu counters pass
out=gpurun_out; mkdir -p $out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py > $out/r2p3_bench_cfg2.json 2> $out/r2p3_bench_cfg2.err || tail -20 $out/r2p3_bench_cfg2.err
python - <<PY
import json
d = json.load(open("$out/r2p3_bench_cfg2.json"))
print("cfg2: value %.0f ms %.3f e2e %.0f (median %.2f p95 %.2f) py_api %.0f / %.0f lanes %s" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["median_ms"], d["e2e"]["p95_ms"], d["e2e_python_api"]["value"], d["e2e_python_api"]["with_best_structures"]["value"], d["config"]["lanes"]))
print("roofline:", json.dumps(d["roofline"])[:900])
print("parity", d["parity_checked"], d["parity_ok"], "cpu", d.get("cpu_baseline"))
PY
for split in 50 75 85 92; do
  RAFFT_LANES=2 RAFFT_LANE_SPLIT=$split python bench.py --steps 8 --no-cpu-baseline --parity-sample 0 > $out/r2p3_split$split.json 2> $out/r2p3_split$split.err
  python - <<PY
import json
d = json.load(open("$out/r2p3_split$split.json"))
print("lanes 2 split $split: device %.3f ms, e2e mean %.3f median %.3f" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["e2e"]["median_ms"]))
PY
done
RAFFT_LANES=1 python bench.py --steps 8 --no-cpu-baseline --parity-sample 0 > $out/r2p3_lanes1.json 2> $out/r2p3_lanes1.err
python - <<PY
import json
d = json.load(open("$out/r2p3_lanes1.json"))
print("lanes 1: device %.3f ms, e2e mean %.3f median %.3f" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["e2e"]["median_ms"]))
PY
# ncu counters (stream loop, one lane: every kernel is a plain launch)
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum
RAFFT_LOOP=stream RAFFT_LANES=1 ncu --metrics $M --clock-control none -c 900 --csv --log-file $out/r2p3_counters_cfg2.csv \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --parity-sample 0 > $out/r2p3_ncu_cfg2.log 2>&1
python tools/ncu_counters.py $out/r2p3_counters_cfg2.csv cfg2 1 --source "r02 probe 3 (pre-kernel-work baseline)"
cp profiles/r02_kernel_counters.json $out/r2p3_kernel_counters.json
