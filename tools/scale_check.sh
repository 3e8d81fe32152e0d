#!/bin/bash
# weak-scaling bench lines on one box (under gpurun --gpus 8):  tools/scale_check.sh <tag> [Ns...]
tag=${1:?tag}; shift
mkdir -p gpurun_out
for n in ${*:-2 4 8}; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
    bench.py --gpus $n --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_n$n.json 2> gpurun_out/${tag}_n$n.err
  tail -1 gpurun_out/${tag}_n$n.json | cut -c1-200
done
