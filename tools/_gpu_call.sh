set -x
cd $GRAFT_REPO_ROOT
timeout 30 python -m pytest tests/test_gpu_volume_space.py -m gpu -q -x 2>&1 | tail -2
