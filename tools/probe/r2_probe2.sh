#!/bin/bash
# round-2 probe 2: the restructured loop (graph WHILE / stream look-ahead / profile), lanes, multi entry
out=gpurun_out; mkdir -p $out
run() { echo "== $*"; "$@" 2>&1 | tail -${TAILN:-4}; }
TAILN=6 run python -m pytest tests -m gpu -x -q
RAFFT_LOOP=stream TAILN=3 run python -m pytest tests/test_gpu_parity.py -m gpu -x -q
RAFFT_LANES=1 TAILN=3 run python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "cfg2 or cfg3 or full or random or edge"
RAFFT_PROFILE=1 TAILN=3 run python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "cfg2 or random or edge"
for wl in cfg2 cfg3; do
  for c in 1250 10000; do
    for mode in "" "RAFFT_LANES=1" "RAFFT_LOOP=stream"; do
      env $mode python bench.py --workload $wl --count $c --steps 5 --warmup 3 --no-cpu-baseline > $out/r2p2_${wl}_c$c.json 2> $out/r2p2_${wl}_c$c.err || tail -5 $out/r2p2_${wl}_c$c.err
      python - <<PY
import json
d = json.load(open("$out/r2p2_${wl}_c$c.json"))
print("$wl count $c [$mode]: ms/step %.3f value %.0f e2e %.0f launches %d e2e steps %s" % (d["ms_per_step"], d["value"], d["e2e"]["value"], d["gpu_launches"], d["e2e"]["ms_steps_rank0"]))
PY
    done
  done
done
