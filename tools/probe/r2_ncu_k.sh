#!/bin/bash
# --set full capture of one kernel at a given batch size: r2_ncu_k.sh <tag> <kernel> <skip> <workload> <count>
out=gpurun_out; tag=$1; k=$2; skip=$3; wl=$4; cnt=$5
export RAFFT_LOOP=stream RAFFT_LANES=1
python bench.py --workload $wl --count $cnt --steps 1 --warmup 1 --no-cpu-baseline --parity-sample 0 > /dev/null 2>&1 || { echo bench failed; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -o /tmp/${tag}_$k \
  python bench.py --workload $wl --count $cnt --steps 1 --warmup 1 --no-cpu-baseline --parity-sample 0 > $out/${tag}_$k.log 2>&1
ncu -i /tmp/${tag}_$k.ncu-rep --page raw --csv > $out/${tag}_$k.raw.csv 2>/dev/null
ncu -i /tmp/${tag}_$k.ncu-rep --page source --csv 2>/dev/null | gzip > $out/${tag}_$k.source.csv.gz
ls -la $out/${tag}_$k.*
