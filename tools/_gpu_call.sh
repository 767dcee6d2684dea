set -x
cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -q -x > gpurun_out/r2d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2d_pytest.log
tail -15 gpurun_out/r2d_pytest.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; echo "bench rc=$?"; tail -5 gpurun_out/r2d_bench.err
Q="--steps 20 --warmup 5 --no-cpu-baseline --no-upload-probe --no-gather-probe --no-load-leg"
for v in spec pairs specpairs; do
VR_B200_LIB=$PWD/volume_renderer_b200/csrc/ab/libvr_b200_$v.so python bench.py $Q > gpurun_out/r2d_ab_${v}_l.json 2> gpurun_out/r2d_ab_${v}_l.err
done
VR_B200_LIB=$PWD/volume_renderer_b200/csrc/ab/libvr_b200_pairs.so python bench.py $Q --gradient compute --lights 2 > gpurun_out/r2d_ab_pairs_c2.json 2> gpurun_out/r2d_ab_pairs_c2.err
python bench.py $Q --gradient compute --lights 2 > gpurun_out/r2d_ab_base_c2.json 2> gpurun_out/r2d_ab_base_c2.err
for w in 2 4; do python bench.py $Q --wshape $w > gpurun_out/r2d_ab_wshape${w}_l.json 2> gpurun_out/r2d_ab_wshape${w}_l.err; done
python bench.py --config example1 --steps 20 --warmup 5 --no-upload-probe --no-load-leg --cpu-seconds 8 > gpurun_out/r2d_cfg_example1.json 2> gpurun_out/r2d_cfg_example1.err; echo "ex1 rc=$?"
python bench.py --config example2 --warmup 5 --no-upload-probe --no-load-leg --no-cpu-baseline > gpurun_out/r2d_cfg_example2.json 2> gpurun_out/r2d_cfg_example2.err; echo "ex2 rc=$?"
python bench.py --config example3 --steps 8 --warmup 2 > gpurun_out/r2d_cfg_example3.json 2> gpurun_out/r2d_cfg_example3.err; echo "ex3 rc=$?"; tail -3 gpurun_out/r2d_cfg_example3.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/r2d_*.json")):
    try:
        d=json.load(open(f))
        print(f, "ms/frame %.3f value %.1f shaded %.2f e2e %.1f" % (d["ms_per_step"], d["value"], d["gsamples_shaded_per_s"], d["e2e"]["value"]), d.get("parity",{}).get("max_abs_norm"))
    except Exception as ex:
        print(f, "failed", ex)
PY
