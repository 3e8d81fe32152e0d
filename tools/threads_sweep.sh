#!/bin/bash
# sweep the block size of each block-scope phase kernel (one at a time) on a bench workload:
#   KERNELS="COMBINE SELECT" THREADS="96 128" BENCH_ARGS="--workload cfg3" tools/threads_sweep.sh
run() { python bench.py --steps 3 --warmup 2 --no-cpu-baseline ${BENCH_ARGS} 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'])"; }
echo -n "default: "; run
for k in ${KERNELS:-GATHER COMBINE SELECT MATERIALISE FINISH}; do
  for t in ${THREADS:-32 64 128 256}; do
    echo -n "$k $t: "; env RAFFT_${k}_THREADS=$t bash -c "$(declare -f run); run"
  done
done
