#!/bin/bash
# Multi-GPU bench lines (run under `gpurun --gpus N`): c2 (strong), c3 (weak), c4 (row-tiled single image).
N=${1:-8}
mkdir -p gpurun_out
for wl in ${WLS:-c2 c3 c4}; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --steps 5 --warmup 3 --workload $wl --no-also --no-cpu > gpurun_out/scale_${wl}_n$N.json 2> gpurun_out/scale_${wl}_n$N.err
  echo "$wl rc=$?"; tail -c 1200 gpurun_out/scale_${wl}_n$N.json | cut -c1-1200
done
