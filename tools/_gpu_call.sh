set -x
cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -q --deselect tests/test_gpu_parity.py::test_resident_volume_beyond_2_to_32_elements > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
tail -40 gpurun_out/r2b_pytest.log
python tools/arith_ab.py > gpurun_out/r2b_arith_ab.log 2>&1
cat gpurun_out/r2b_arith_ab.log
for a in 0 1 2; do
python bench.py --steps 20 --warmup 5 --arith $a --no-cpu-baseline --no-upload-probe --no-gather-probe --no-load-leg > gpurun_out/r2b_bench_a$a.json 2> gpurun_out/r2b_bench_a$a.err; echo "bench a$a rc=$?"
python bench.py --steps 20 --warmup 5 --arith $a --gradient compute --lights 1 --no-cpu-baseline --no-upload-probe --no-gather-probe --no-load-leg > gpurun_out/r2b_bench_c_a$a.json 2> gpurun_out/r2b_bench_c_a$a.err; echo "bench compute a$a rc=$?"
done
python - <<'PY'
import json
for a in (0,1,2):
    for t in ("", "_c"):
        try:
            d=json.load(open("gpurun_out/r2b_bench%s_a%d.json" % (t,a)))
            print(t or "_l", a, "ms/frame %.3f value %.1f shaded %.2f e2e %.1f single %.2f" % (d["ms_per_step"], d["value"], d["gsamples_shaded_per_s"], d["e2e"]["value"], d["e2e"]["single_frame"]["ms_per_frame"]))
        except Exception as ex:
            print(t, a, "failed", ex)
PY
