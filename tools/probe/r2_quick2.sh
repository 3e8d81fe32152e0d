#!/bin/bash
# quick check of a beam-kernel change: parity subset + cfg2/cfg3 at 1250 and 10000 sequences, per-family times
out=gpurun_out; tag=${1:-q2}
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "${2:-full_batch_cfg2 or cfg3_2500 or random or edge or reference or example or cfg4}" 2>&1 | tail -2
for wl in cfg2 cfg3; do for c in 1250 10000; do
python bench.py --workload $wl --count $c --steps 5 --no-cpu-baseline --parity-sample 0 > $out/${tag}_${wl}_$c.json 2> $out/${tag}_${wl}_$c.err || tail -5 $out/${tag}_${wl}_$c.err
python - <<PY
import json
d = json.load(open("$out/${tag}_${wl}_$c.json"))
print("$wl $c: device %.3f ms (%.0f seq/s), e2e median %.3f; families %s" % (d["ms_per_step"], d["value"], d["e2e"]["median_ms"], {k: round(v, 2) for k, v in d["roofline"]["family_ms_per_step"].items() if v > 0.15}))
PY
done; done
