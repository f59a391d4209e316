python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r02f_gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02f_gputests.log
tail -8 gpurun_out/r02f_gputests.log
python bench.py --steps 40 --warmup 5 --no-cpu > gpurun_out/r02f_bench.json 2> gpurun_out/r02f_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02f_bench.json"))
print("ms/step %.2f"%d["ms_per_step"], "rest %.2f"%d["config"]["at_rest"]["ms_per_step"], "rebuilds",d["config"]["bucket_rebuilds"],"refilters",d["config"]["hit_refilters"], "e2e %.3g"%d["e2e"]["value"])
print({k:v2 for k,v2 in list(d["roofline"]["kernels_ms_per_step"].items())[:12]})
PY
