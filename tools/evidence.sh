#!/bin/bash
# Round evidence, run on the GPU box under gpurun:  tools/evidence.sh <tag> [kernels...]
#   1. bench.py (cfg2, default flags minus the CPU baseline) -> gpurun_out/<tag>_bench.json
#   2. ncu launch list of a 1-step bench                     -> gpurun_out/<tag>_launches.csv
#   2b. ncu DRAM byte counters of every launch           -> gpurun_out/<tag>_dram.csv
#   3. one `ncu --set full` capture per kernel named (folding iteration 2, the largest, of the first
#      fold): raw and source pages exported as CSV on the box (the .ncu-rep files together exceed
#      the 64 MiB that travel back); only the first kernel's .ncu-rep is kept.
# Every ncu pass runs only after the same command exited 0 without ncu.
set -u
tag=${1:?tag}; shift
kernels=${*:-k_expand_fast k_combine k_materialise}
out=gpurun_out
mkdir -p $out
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > $out/${tag}_bench.json 2> $out/${tag}_bench.err || { echo "bench failed"; tail -5 $out/${tag}_bench.err; exit 1; }
python bench.py --steps 1 --warmup 1 --no-cpu-baseline > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_launches.csv \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline > $out/${tag}_ncu.log 2>&1
# DRAM bytes of every launch (two folds: warm-up + step) -> per-kernel traffic per launch (tools/traffic_table.py)
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 --csv --log-file $out/${tag}_dram.csv \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline > $out/${tag}_dram.log 2>&1
first=1
for k in $kernels; do
  rep=/tmp/${tag}_$k
  [ $first = 1 ] && rep=$out/${tag}_$k
  ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -o $rep \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > $out/${tag}_$k.log 2>&1
  ncu -i $rep.ncu-rep --page raw --csv > $out/${tag}_$k.raw.csv 2>/dev/null
  ncu -i $rep.ncu-rep --page source --csv 2>/dev/null | gzip > $out/${tag}_$k.source.csv.gz
  first=0
done
ls -la $out | grep $tag
