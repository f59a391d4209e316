# baseline of the restored HEAD: all GPU tests, the flowing bench (default), launch list
python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r02d_gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02d_gputests.log
tail -15 gpurun_out/r02d_gputests.log
python bench.py --steps 40 --warmup 5 > gpurun_out/r02d_bench.json 2> gpurun_out/r02d_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02d_bench.json"))
print("ms/step %.2f"%d["ms_per_step"], "value %.3g"%d["value"], "rest", d["config"].get("at_rest"), "rebuilds",d["config"].get("bucket_rebuilds"),"refilters",d["config"].get("hit_refilters"), "e2e %.3g"%d["e2e"]["value"])
print(d["roofline"].get("kernels_ms_per_step"))
print(d.get("cpu_baseline"))
PY
tail -3 gpurun_out/r02d_bench.err
