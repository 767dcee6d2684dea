set -x
cd $GRAFT_REPO_ROOT
timeout 60 python -m pytest tests/test_gpu_volume_space.py tests/test_gpu_latency_mode.py -m gpu -q -x 2>&1 | tail -3
timeout 60 python tools/vs_one_rank.py --rank 0 --world 2 --size 1024 --dtype f32 --slices 5 --frames 360 2>&1 | grep -c "^frame"
timeout 60 python tools/vs_one_rank.py --rank 1 --world 2 --size 1024 --dtype f32 --slices 8 --frames 360 2>&1 | grep -c "^frame"
