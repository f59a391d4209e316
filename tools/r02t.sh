
python bench.py --scheme inc --state rest --steps 10 --warmup 4 --no-cpu --rest-steps 0 --e2e-steps 1 --profile-steps 2 > gpurun_out/r02t_inc.json 2> gpurun_out/r02t_inc.err; echo "inc rc=$?"
PMPS_CG_HOST_LOOP=1 python bench.py --scheme inc --state rest --steps 10 --warmup 4 --no-cpu --rest-steps 0 --e2e-steps 1 --profile-steps 2 > gpurun_out/r02t_inc_old.json 2> gpurun_out/r02t_inc_old.err; echo "inc rc=$?"
python - <<'PY'
import json
for f in ["r02t_inc","r02t_inc_old"]:
    d=json.loads([l for l in open("gpurun_out/%s.json"%f).read().splitlines() if l.startswith("{")][-1])
    print(f, "ms/step %.2f value %.3g cg iters %s"%(d["ms_per_step"], d["value"], d["config"]["cg_iterations_last_step"]))
    print("  ", {k:v for k,v in list(d["roofline"]["kernels_ms_per_step"].items())[:14]})
PY
