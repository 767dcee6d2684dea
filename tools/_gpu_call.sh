set -x
cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -q > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2c_pytest.log
tail -60 gpurun_out/r2c_pytest.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; echo "bench rc=$?"; tail -5 gpurun_out/r2c_bench.err
Q="--steps 20 --warmup 5 --arith 2 --no-cpu-baseline --no-upload-probe --no-gather-probe --no-load-leg"
for v in mb9 mb10 mb12; do
VR_B200_LIB=$PWD/volume_renderer_b200/csrc/ab/libvr_b200_$v.so python bench.py $Q > gpurun_out/r2c_ab_${v}_l.json 2> gpurun_out/r2c_ab_${v}_l.err
VR_B200_LIB=$PWD/volume_renderer_b200/csrc/ab/libvr_b200_$v.so python bench.py $Q --gradient compute --lights 1 > gpurun_out/r2c_ab_${v}_c.json 2> gpurun_out/r2c_ab_${v}_c.err
done
python bench.py $Q --gradient compute --lights 1 > gpurun_out/r2c_ab_base_c.json 2> gpurun_out/r2c_ab_base_c.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/r2c_ab_*.json"))+["gpurun_out/r2c_bench.json"]:
    try:
        d=json.load(open(f))
        print(f, "ms/frame %.3f value %.1f shaded %.2f e2e %.1f single %.2f" % (d["ms_per_step"], d["value"], d["gsamples_shaded_per_s"], d["e2e"]["value"], d["e2e"]["single_frame"]["ms_per_frame"]))
    except Exception as ex:
        print(f, "failed", ex)
PY
P="--steps 2 --warmup 1 --no-cpu-baseline --batch 2 --no-upload-probe --no-gather-probe --no-load-leg"
ncu --set full --clock-control none --import-source on -k regex:k_raymarch -s 1 -c 1 -f -o gpurun_out/r2c_a2 python bench.py $P > gpurun_out/r2c_ncu.log 2>&1; echo "ncu rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2c_launches.csv python bench.py $P > gpurun_out/r2c_ncu_l.log 2>&1; echo "ncu list rc=$?"
ls -la gpurun_out/
