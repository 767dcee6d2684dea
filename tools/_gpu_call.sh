set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
TR2="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR2 bench.py --gpus 2 --mode volume --size 2048 --dtype bf16 --steps 12 --warmup 3 --no-cpu-baseline > gpurun_out/vs2_2048_final.json 2> gpurun_out/vs2_2048_final.err; echo rc=$?
timeout 600 $TR2 bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo rc=$?
