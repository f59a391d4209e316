for v in base cur; do
  L=$PWD/polymps_b200/libpmps.so; [ $v = base ] && L=$PWD/polymps_b200/libpmps_base.so
  PMPS_LIB=$L python bench.py --steps 20 --warmup 5 --no-cpu --rest-steps 10 --e2e-steps 2 > gpurun_out/r02g_bench_$v.json 2> gpurun_out/r02g_bench_$v.err; echo "bench $v rc=$?"
done
python - <<'PY'
import json
for v in ["base","cur"]:
    d=json.load(open("gpurun_out/r02g_bench_%s.json"%v))
    print(v, "ms/step %.2f"%d["ms_per_step"], "rest %.2f"%d["config"]["at_rest"]["ms_per_step"], "rebuilds",d["config"]["bucket_rebuilds"],"refilters",d["config"]["hit_refilters"])
    print("   ", {k:v2 for k,v2 in list(d["roofline"]["kernels_ms_per_step"].items())[:10]})
PY
