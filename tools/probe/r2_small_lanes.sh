#!/bin/bash
# do two lanes help SMALL shards (the strong-scaling case)?  device ms per step at 1250 / 2500 sequences
out=gpurun_out
for wl in cfg2 cfg3; do for c in 1250 2500; do for mode in "RAFFT_LANES=1" "RAFFT_LANES=2" "RAFFT_LANES=2 RAFFT_LANE_SPLIT=60"; do
  env $mode python bench.py --workload $wl --count $c --steps 8 --no-cpu-baseline --parity-sample 0 > $out/sl.json 2> $out/sl.err || tail -3 $out/sl.err
  python - <<PY
import json
d = json.load(open("$out/sl.json"))
print("$wl $c [$mode]: device %.3f ms, e2e median %.3f" % (d["ms_per_step"], d["e2e"]["median_ms"]))
PY
done; done; done
