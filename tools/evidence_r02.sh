#!/bin/bash
# Round-2 evidence, run on the GPU box under gpurun:  tools/evidence_r02.sh <tag>
#   1. pytest -m gpu, smoke()
#   2. bench lines of cfg 2 / 3 / 4 and the reference arm
#   3. per workload: ncu metrics pass (time, warp instructions, thread instructions, DRAM bytes per launch; stream
#      loop so that every kernel is a plain launch) -> launch list + tools/ncu_counters.py table
#   4. ncu --set full of the dominant kernels (raw + source pages as CSV)
# Every ncu pass runs only after the same command exited 0 without ncu.
set -u
tag=${1:?tag}
out=gpurun_out
mkdir -p $out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee $out/${tag}_pytest.txt
python __graft_entry__.py smoke 2>&1 | tail -1 | tee $out/${tag}_smoke.txt
python bench.py > $out/${tag}_bench_cfg2.json 2> $out/${tag}_bench_cfg2.err; cut -c1-200 $out/${tag}_bench_cfg2.json
python bench.py --impl reference --steps 2 --warmup 1 > $out/${tag}_bench_cfg2_ref.json 2> $out/${tag}_bench_cfg2_ref.err
python bench.py --workload cfg3 --steps 5 > $out/${tag}_bench_cfg3.json 2> $out/${tag}_bench_cfg3.err; cut -c1-200 $out/${tag}_bench_cfg3.json
python bench.py --workload cfg4 --steps 3 --warmup 3 > $out/${tag}_bench_cfg4.json 2> $out/${tag}_bench_cfg4.err; cut -c1-200 $out/${tag}_bench_cfg4.json
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum
for wl in cfg2 cfg3 cfg4; do
  RAFFT_LOOP=stream RAFFT_LANES=1 python bench.py --workload $wl --steps 1 --warmup 1 --no-cpu-baseline --parity-sample 0 > /dev/null 2>&1 || continue
  RAFFT_LOOP=stream RAFFT_LANES=1 ncu --metrics $M --clock-control none -c 3000 --csv --log-file $out/${tag}_counters_$wl.csv \
    python bench.py --workload $wl --steps 1 --warmup 1 --no-cpu-baseline --parity-sample 0 > $out/${tag}_ncu_$wl.log 2>&1
  python tools/ncu_counters.py $out/${tag}_counters_$wl.csv $wl 1 --source "profiles/${tag}_counters_$wl.txt (ncu metrics pass of bench.py --workload $wl --steps 1 --warmup 1, second fold)" | tee $out/${tag}_counters_$wl.txt
  gzip -f $out/${tag}_counters_$wl.csv
done
cp profiles/r02_kernel_counters.json $out/${tag}_kernel_counters.json
cap() { # kernel skip workload
  RAFFT_LOOP=stream RAFFT_LANES=1 ncu --set full --clock-control none --import-source on -k regex:$1 -s $2 -c 1 -o /tmp/${tag}_$3_$1 \
    python bench.py --workload $3 --steps 1 --warmup 1 --no-cpu-baseline --parity-sample 0 > $out/${tag}_$3_$1.log 2>&1
  ncu -i /tmp/${tag}_$3_$1.ncu-rep --page raw --csv > $out/${tag}_$3_$1.raw.csv 2>/dev/null
  ncu -i /tmp/${tag}_$3_$1.ncu-rep --page source --csv 2>/dev/null | gzip > $out/${tag}_$3_$1.source.csv.gz
}
cap k_expand_fast 1 cfg2
cap k_combine 1 cfg2
cap k_materialise 1 cfg2
cap k_expand_big 1 cfg4
cap k_expand_fast 1 cfg3
ls -la $out | grep ${tag}_ | wc -l
