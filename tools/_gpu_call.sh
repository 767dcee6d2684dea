set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/r2i_pytest_2gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2i_pytest_2gpu.log
timeout 400 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2i_bench_n2.json 2> gpurun_out/r2i_bench_n2.err; echo "bench n2 rc=$?"; tail -5 gpurun_out/r2i_bench_n2.err
timeout 400 $TR bench.py --gpus 2 --mode volume --size 1024 --dtype f32 --steps 6 --warmup 2 --no-cpu-baseline > gpurun_out/r2i_vol_n2_f32.json 2> gpurun_out/r2i_vol_n2_f32.err; echo "vol f32 rc=$?"; tail -5 gpurun_out/r2i_vol_n2_f32.err
timeout 600 $TR bench.py --gpus 2 --mode volume --size 2048 --dtype bf16 --steps 6 --warmup 2 --no-cpu-baseline > gpurun_out/r2i_vol_n2_bf16.json 2> gpurun_out/r2i_vol_n2_bf16.err; echo "vol bf16 rc=$?"; tail -5 gpurun_out/r2i_vol_n2_bf16.err
timeout 300 ./harness/vrbench --gpus 2 --size 512 --image 2048 2048 --frames 6 --grad --lights 2 > gpurun_out/r2i_vrbench_g2.log 2>&1; echo "vrbench g2 rc=$?"; grep -E "frame|total|group" gpurun_out/r2i_vrbench_g2.log | tail -9
timeout 300 ./harness/vrbench --gpus 1 --size 512 --image 2048 2048 --frames 6 --grad --lights 2 > gpurun_out/r2i_vrbench_g1.log 2>&1; grep -E "total" gpurun_out/r2i_vrbench_g1.log
timeout 500 $TR bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r2i_ref_n2.json 2> gpurun_out/r2i_ref_n2.err; echo "ref n2 rc=$?"
python - <<'PY'
import json
for f in ("r2i_bench_n2","r2i_vol_n2_f32","r2i_vol_n2_bf16","r2i_ref_n2"):
    try:
        d=json.load(open("gpurun_out/%s.json"%f))
        print(f, "ms/step %.3f value %.2f e2e %.2f" % (d["ms_per_step"], d["value"], d["e2e"]["value"]), d.get("phase_ms_per_frame_max_over_ranks"), d.get("cpu_baseline",{}).get("cores"))
        if d.get("per_rank"): print("   ", d["per_rank"])
    except Exception as ex:
        print(f, "failed", repr(ex))
PY
