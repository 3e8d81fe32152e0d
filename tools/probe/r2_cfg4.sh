#!/bin/bash
out=gpurun_out; tag=${1:-c4}
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "cfg4 or long or edge or full_batch_cfg2" 2>&1 | tail -3
for big in 1 0; do
RAFFT_BIG=$big python bench.py --workload cfg4 --steps 2 --warmup 2 --no-cpu-baseline --parity-sample 0 > $out/${tag}_cfg4_big$big.json 2> $out/${tag}_cfg4_big$big.err || tail -5 $out/${tag}_cfg4_big$big.err
python - <<PY
import json
d = json.load(open("$out/${tag}_cfg4_big$big.json"))
print("cfg4 big=$big: device %.1f ms (%.0f seq/s), e2e median %.1f; iterations %s; families %s" % (d["ms_per_step"], d["value"], d["e2e"]["median_ms"], d["config"]["iterations"], {k: round(v, 1) for k, v in d["roofline"]["family_ms_per_step"].items()}))
PY
done
RAFFT_DEBUG_ALLOC=1 python tools/iter_profile.py cfg4 200 2>&1 | grep "regions per list\|it  [0-9]:\|it 1[0-9]:" | head -24
