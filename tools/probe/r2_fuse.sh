#!/bin/bash
out=gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "lanes_and_loop or full_batch_cfg2 or random or edge or reference or cfg4 or parameter" 2>&1 | tail -3
for fuse in 1 0; do for c in 1250 10000; do for wl in cfg2 cfg3; do
RAFFT_FUSE=$fuse python bench.py --workload $wl --count $c --steps 5 --no-cpu-baseline --parity-sample 0 > $out/fuse${fuse}_${wl}_$c.json 2> $out/fuse${fuse}_${wl}_$c.err || tail -3 $out/fuse${fuse}_${wl}_$c.err
python - <<PY
import json
d = json.load(open("$out/fuse${fuse}_${wl}_$c.json"))
print("fuse=$fuse $wl $c: device %.3f ms, e2e median %.3f; launches %d; families %s" % (d["ms_per_step"], d["e2e"]["median_ms"], d["gpu_launches"] / d["steps"], {k: round(v, 2) for k, v in d["roofline"]["family_ms_per_step"].items()}))
PY
done; done; done
