#!/bin/bash
# sweep the launch shapes of k_expand_fast (RAFFT_FAST_SHAPE) on the default bench workload
for s in ${SHAPES:-0 1 2 3 4 5 6 7 8 9}; do
  echo -n "shape $s: "
  RAFFT_FAST_SHAPE=$s python bench.py --steps 3 --warmup 2 --no-cpu-baseline ${BENCH_ARGS} 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'])"
done
