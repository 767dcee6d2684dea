set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517"
timeout 400 $TR8 bench.py --gpus 8 --mode volume --size 4096 --dtype bf16 --steps 12 --warmup 3 --no-cpu-baseline > gpurun_out/vs8_auto.json 2> gpurun_out/vs8_auto.err; echo rc=$?
tail -c 300 gpurun_out/vs8_auto.err
