set -x
cd $GRAFT_REPO_ROOT
timeout 100 python -m pytest tests/test_gpu_volume_space.py tests/test_gpu_latency_mode.py -m gpu -q -x 2>&1 | tail -3
