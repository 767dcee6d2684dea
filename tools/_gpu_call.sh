set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2o_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2o_pytest.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2o_bench.json 2> gpurun_out/r2o_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2o_bench.err
python bench.py --config example1 --steps 20 --warmup 5 --no-upload-probe --no-load-leg --cpu-seconds 8 > gpurun_out/r2o_cfg_example1.json 2> gpurun_out/r2o_cfg_example1.err; echo "ex1 rc=$?"
python bench.py --config example2 --warmup 5 --no-upload-probe --no-load-leg --no-cpu-baseline > gpurun_out/r2o_cfg_example2.json 2> gpurun_out/r2o_cfg_example2.err; echo "ex2 rc=$?"
python bench.py --config example3 --steps 8 --warmup 2 > gpurun_out/r2o_cfg_example3.json 2> gpurun_out/r2o_cfg_example3.err; echo "ex3 rc=$?"
python bench.py --dtype bf16 --steps 20 --warmup 5 --no-cpu-baseline --no-upload-probe --no-gather-probe > gpurun_out/r2o_bf16.json 2> gpurun_out/r2o_bf16.err; echo "bf16 rc=$?"
P="--steps 2 --warmup 1 --no-cpu-baseline --batch 2 --no-upload-probe --no-gather-probe --no-load-leg"
python bench.py $P > gpurun_out/r2o_ncu_plain.json 2> gpurun_out/r2o_ncu_plain.err; echo "plain rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_raymarch -s 1 -c 1 -f -o gpurun_out/r2o_final python bench.py $P > gpurun_out/r2o_ncu.log 2>&1; echo "ncu rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2o_launches.csv python bench.py $P > gpurun_out/r2o_ncu_l.log 2>&1; echo "ncu list rc=$?"
VR_WARP_TIMES=1 python tools/warp_times.py > gpurun_out/r2o_warp_times.log 2>&1; echo "warp_times rc=$?"; tail -8 gpurun_out/r2o_warp_times.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/r2o_*.json")):
    try:
        d=json.load(open(f))
        print(f, "ms/frame %.3f value %.1f shaded %.2f e2e %.1f" % (d["ms_per_step"], d["value"], d["gsamples_shaded_per_s"], d["e2e"]["value"]), (d.get("parity") or {}).get("max_abs_norm"), (d.get("parity") or {}).get("psnr_db"))
    except Exception as ex:
        print(f, "failed", ex)
PY
