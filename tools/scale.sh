#!/bin/bash
# image-space scaling run on one box: N = 1, 2, 4, 8 (default bench settings), then the single-frame launch at 8
# usage: tools/scale.sh tag [max_gpus]     (results in gpurun_out/<tag>_n<N>.json)
tag=$1; maxn=${2:-8}
python bench.py --no-cpu-baseline --no-upload-probe > gpurun_out/${tag}_n1.json 2> gpurun_out/${tag}_n1.err
port=29611
for n in 2 4 8; do
  [ $n -le $maxn ] || continue
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --no-cpu-baseline --no-upload-probe > gpurun_out/${tag}_n$n.json 2> gpurun_out/${tag}_n$n.err
  port=$((port+1))
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node $maxn --master-addr 127.0.0.1 --master-port $port bench.py --gpus $maxn --no-cpu-baseline --no-upload-probe --batch 1 > gpurun_out/${tag}_n${maxn}_b1.json 2> gpurun_out/${tag}_n${maxn}_b1.err
