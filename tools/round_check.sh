#!/bin/bash
# Whole-round check on the GPU box (under gpurun):  tools/round_check.sh <tag>
# GPU tests, smoke, the bench lines of cfg 2 / 3 / 4, then tools/evidence.sh (ncu passes).
tag=${1:?tag}
out=gpurun_out
mkdir -p $out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python __graft_entry__.py smoke 2>&1 | tail -1
python bench.py > $out/${tag}_bench_cfg2.json 2> $out/${tag}_bench_cfg2.err; cut -c1-160 $out/${tag}_bench_cfg2.json
python bench.py --impl reference --steps 2 --warmup 1 > $out/${tag}_bench_cfg2_ref.json 2> $out/${tag}_bench_cfg2_ref.err
python bench.py --workload cfg3 --steps 3 > $out/${tag}_bench_cfg3.json 2> $out/${tag}_bench_cfg3.err; cut -c1-160 $out/${tag}_bench_cfg3.json
python bench.py --workload cfg4 --steps 2 --warmup 3 --no-cpu-baseline > $out/${tag}_bench_cfg4.json 2> $out/${tag}_bench_cfg4.err; cut -c1-160 $out/${tag}_bench_cfg4.json
bash tools/evidence.sh $tag k_expand_fast k_combine k_materialise > $out/${tag}_evidence.log 2>&1; tail -3 $out/${tag}_evidence.log
