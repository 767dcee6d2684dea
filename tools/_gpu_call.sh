set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517"
timeout 300 $TR8 bench.py --gpus 8 --mode volume --size 4096 --dtype bf16 --steps 12 --warmup 3 --no-cpu-baseline > gpurun_out/vs8_final.json 2> gpurun_out/vs8_final.err; echo rc=$?
timeout 300 $TR8 bench.py --gpus 8 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo rc=$?
timeout 300 python -m pytest tests/test_gpu_multirank.py tests/test_gpu_group.py tests/test_gpu_latency_mode.py -m gpu -q 2>&1 | tail -4 | tee gpurun_out/pytest_8gpu.log
timeout 120 ./harness/vrbench --size 512 --image 2048 2048 --frames 6 --grad --lights 2 --gpus 8 2>&1 | tail -8 | tee gpurun_out/vrbench_gpus8.log
timeout 120 ./harness/vrbench --size 512 --image 2048 2048 --frames 6 --grad --lights 2 --gpus 8 --ray-slices 16 2>&1 | tail -8 | tee gpurun_out/vrbench_gpus8_slices.log
