set -x
cd $GRAFT_REPO_ROOT
timeout 40 python tools/vs_one_rank.py --rank 0 --world 2 --size 1024 --dtype f32 --slices 16 --frames 360 2>&1 | grep -c "^frame"
timeout 40 python -m pytest tests/test_gpu_hostpath.py -m gpu -q -k "harness" 2>&1 | tail -2
