set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/pytest_1gpu.log; cat gpurun_out/pytest_1gpu.log
timeout 300 python tools/vs_one_rank.py --rank 1 --frames 3 --slices 0 > gpurun_out/vs_one_rank.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/vs_launches.csv python tools/vs_one_rank.py --rank 1 --frames 3 --slices 0 > gpurun_out/vs_ncu.log 2>&1
tail -2 gpurun_out/vs_one_rank.log | cut -c1-160
grep -c . gpurun_out/vs_launches.csv
