#!/bin/bash
# round-2 probe 1: conditional-graph feasibility, INT issue peak, per-GPU time of the strong-scaling shards
# (no exchange step: the shard of N GPUs = --count 10000/N on one GPU), e2e step histogram
out=gpurun_out; mkdir -p $out
nvidia-smi --query-gpu=name,driver_version --format=csv,noheader
nproc
./tools/probe/condgraph 2>&1 | tee $out/r2p1_condgraph.txt
./tools/probe/intpeak 2>&1 | tee $out/r2p1_intpeak.txt
for wl in cfg2 cfg3; do
  for c in 1250 2500 5000 10000; do
    python bench.py --workload $wl --count $c --steps 5 --warmup 3 --no-cpu-baseline > $out/r2p1_${wl}_c$c.json 2> $out/r2p1_${wl}_c$c.err
    python - <<PY
import json
d = json.load(open("$out/r2p1_${wl}_c$c.json"))
print("$wl count $c: ms/step %.3f value %.0f e2e %.0f launches %d e2e steps %s" % (d["ms_per_step"], d["value"], d["e2e"]["value"], d["gpu_launches"], d["e2e"]["ms_steps_rank0"]))
PY
  done
done
python bench.py --workload cfg2 --steps 100 --warmup 3 --no-cpu-baseline > $out/r2p1_cfg2_e2e100.json 2> $out/r2p1_cfg2_e2e100.err
python - <<PY
import json, statistics as S
d = json.load(open("$out/r2p1_cfg2_e2e100.json"))
x = sorted(d["e2e"]["ms_steps_rank0"])
print("e2e 100 steps: min %.2f median %.2f p95 %.2f max %.2f; device ms/step %.2f" % (x[0], S.median(x), x[94], x[-1], d["ms_per_step"]))
PY
RAFFT_DEBUG_ALLOC=1 python tools/e2e_probe.py cfg2 2>&1 | grep -v "high water" | tail -60 > $out/r2p1_laps.txt
tail -30 $out/r2p1_laps.txt
