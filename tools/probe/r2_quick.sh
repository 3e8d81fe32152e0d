#!/bin/bash
# quick check of a kernel change: parity subset + cfg2/cfg3 device time + per-family times
out=gpurun_out; tag=${1:-q}
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "${2:-full_batch_cfg2 or cfg3_2500 or random or edge or reference or example or kats or long or cfg4}" 2>&1 | tail -3
for wl in cfg2 cfg3; do
python bench.py --workload $wl --steps 5 --no-cpu-baseline --parity-sample 0 > $out/${tag}_$wl.json 2> $out/${tag}_$wl.err || tail -5 $out/${tag}_$wl.err
python - <<PY
import json
d = json.load(open("$out/${tag}_$wl.json"))
print("$wl: device %.3f ms (%.0f seq/s), e2e median %.3f; families %s" % (d["ms_per_step"], d["value"], d["e2e"]["median_ms"], {k: round(v, 2) for k, v in d["roofline"]["family_ms_per_step"].items()}))
PY
done
