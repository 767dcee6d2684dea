set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_volume_space.py tests/test_gpu_multirank.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r2k_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2k_pytest.log
