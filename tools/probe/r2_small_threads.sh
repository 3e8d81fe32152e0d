#!/bin/bash
# block sizes of the beam kernels on SMALL shards (the strong-scaling case)
out=gpurun_out
for wl in cfg2 cfg3; do for c in 1250 2500 5000; do for mode in "A=1" "RAFFT_SELECT_THREADS=128 RAFFT_MATERIALISE_THREADS=128" "RAFFT_SELECT_THREADS=128 RAFFT_MATERIALISE_THREADS=128 RAFFT_COMBINE_THREADS=256" "RAFFT_SELECT_THREADS=256 RAFFT_MATERIALISE_THREADS=256 RAFFT_COMBINE_THREADS=256"; do
  env $mode python bench.py --workload $wl --count $c --steps 8 --no-cpu-baseline --parity-sample 0 > $out/st.json 2> $out/st.err || tail -3 $out/st.err
  python - <<PY
import json
d = json.load(open("$out/st.json"))
f = d["roofline"]["family_ms_per_step"]
print("$wl $c [$mode]: device %.3f ms; combine %.2f select %.2f materialise %.2f" % (d["ms_per_step"], f["combine"], f["select"], f["materialise"]))
PY
done; done; done
