python -m pytest tests -m gpu -x -q > gpurun_out/r02b_gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02b_gputests.log
tail -25 gpurun_out/r02b_gputests.log
python bench.py --steps 40 --warmup 5 --no-cpu > gpurun_out/r02b_bench_base.json 2> gpurun_out/r02b_bench_base.err; echo "bench rc=$?"
for v in occ ilp occilp; do
  PMPS_LIB=$PWD/polymps_b200/libpmps_$v.so python bench.py --steps 20 --warmup 5 --no-cpu --rest-steps 10 --e2e-steps 2 > gpurun_out/r02b_bench_$v.json 2> gpurun_out/r02b_bench_$v.err; echo "bench $v rc=$?"
done
python - <<'PY'
import json
for v in ["base","occ","ilp","occilp"]:
    try:
        d=json.load(open("gpurun_out/r02b_bench_%s.json"%v))
        print(v, "ms/step %.2f"%d["ms_per_step"], "rest %.2f"%d["config"]["at_rest"]["ms_per_step"], "rebuilds",d["config"]["bucket_rebuilds"],"refilters",d["config"]["hit_refilters"], "e2e %.3g"%d["e2e"]["value"])
        print("   ", {k:v2 for k,v2 in list(d["roofline"]["kernels_ms_per_step"].items())[:14]})
    except Exception as e:
        print(v, "failed", e)
PY
tail -3 gpurun_out/r02b_bench_base.err
