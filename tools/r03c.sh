python bench.py > gpurun_out/r03c_bench_n1.json 2> gpurun_out/r03c_bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r03c_bench_n1.json")); r=d["roofline"]
print("N=1 ms/step %.2f value %.4g rest %.2f e2e %.4g cpu %s"%(d["ms_per_step"], d["value"], d["config"]["at_rest"]["ms_per_step"], d["e2e"]["value"], d.get("cpu_baseline",{}).get("value")))
print(r["kernel"], r["kernel_ms"], r["frac"], r["traffic"], r["fp64_active"], r["profile_pass"])
PY
