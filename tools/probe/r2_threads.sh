#!/bin/bash
out=gpurun_out
run() { # label env...
  label=$1; shift
  for c in 1250 10000; do
    env "$@" python bench.py --workload cfg2 --count $c --steps 5 --no-cpu-baseline --parity-sample 0 > $out/thr.json 2> $out/thr.err || tail -3 $out/thr.err
    python - <<PY
import json
d = json.load(open("$out/thr.json"))
print("$label cfg2 $c: device %.3f ms; families %s" % (d["ms_per_step"], {k: round(v, 2) for k, v in d["roofline"]["family_ms_per_step"].items() if v > 0.15}))
PY
  done
}
run base A=1
run comb256 RAFFT_COMBINE_THREADS=256
run chunk512 RAFFT_CHUNK=512
run chunk512_c256 RAFFT_CHUNK=512 RAFFT_COMBINE_THREADS=256
run chunk1024_c256 RAFFT_CHUNK=1024 RAFFT_COMBINE_THREADS=256
run mat128 RAFFT_MATERIALISE_THREADS=128 RAFFT_SELECT_THREADS=128
run gather128 RAFFT_GATHER_THREADS=128
