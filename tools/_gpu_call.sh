set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python tools/vs_one_rank.py --rank 0 --world 2 --size 1024 --dtype f32 --slices 0 --frames 3 2>&1 | grep "VrError\|frame" | cut -c1-110
timeout 300 python tools/vs_one_rank.py --rank 1 --world 2 --size 2048 --slices 0 --frames 3 2>&1 | grep "VrError\|frame" | cut -c1-110
timeout 300 python tools/vs_one_rank.py --rank 1 --slices 0 --frames 3 2>&1 | grep "VrError\|frame" | cut -c1-110
timeout 300 python tools/vs_one_rank.py --rank 6 --slices 0 --frames 3 2>&1 | grep "VrError\|frame" | cut -c1-110
timeout 300 python tools/vs_one_rank.py --rank 7 --slices 0 --frames 3 2>&1 | grep "VrError\|frame" | cut -c1-110
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; tail -c 600 gpurun_out/bench_n1.json
