#!/bin/bash
# 2-GPU (or N-GPU) lines of the three multi-GPU workloads: bash tools/gpu_n2.sh <tag> [N]
tag=${1:-n2}; n=${2:-2}; port=29610
mkdir -p gpurun_out
for wl in c2r c5 c2; do
  extra=""; [ $wl = c2 ] && extra="--md5-steps 0"
  port=$((port + 1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --workload $wl \
    --steps 5 --warmup 3 --no-cpu-baseline $extra > gpurun_out/${tag}_${wl}_n${n}.json 2> gpurun_out/${tag}_${wl}_n${n}.err
  echo "$wl rc=$?"; tail -1 gpurun_out/${tag}_${wl}_n${n}.json | cut -c1-160
done
