set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python tools/latency_mode.py --gpus 8 --slices 1,4,8,16 2>&1 | tail -4 | tee gpurun_out/latency_8.log
timeout 300 python tools/latency_mode.py --gpus 4 --slices 1,8,16 2>&1 | tail -3 | tee gpurun_out/latency_4.log
timeout 300 python tools/latency_mode.py --gpus 2 --slices 1,8 2>&1 | tail -2 | tee gpurun_out/latency_2.log
