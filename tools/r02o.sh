python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r02o_gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02o_gputests.log
tail -6 gpurun_out/r02o_gputests.log
python bench.py --steps 20 --warmup 5 --no-cpu --rest-steps 10 --e2e-steps 2 > gpurun_out/r02o_n1.json 2> gpurun_out/r02o_n1.err; echo "bench rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --e2e-steps 4 > gpurun_out/r02o_n2.json 2> gpurun_out/r02o_n2.err; echo "bench2 rc=$?"
python - <<'PY'
import json
def load(p):
    txt=open(p).read(); return json.loads([l for l in txt.splitlines() if l.startswith("{")][-1])
d=load("gpurun_out/r02o_n1.json"); k=d["roofline"]["kernels_ms_per_step"]
print("N=1 ms/step %.2f rest %.2f"%(d["ms_per_step"],d["config"]["at_rest"]["ms_per_step"]), " ".join("%s %.2f"%(n.split('.')[0][:2]+n.split('.')[-1][:5],k[n]) for n in ["F3.pnd+diffusive","F1.viscosity+predict","F5.gradient","shift.gradient","shift.concentration","F2.collision","W5.wall+correct","list.build","list.hits"]))
d=load("gpurun_out/r02o_n2.json")
print("N=2 ms/step %.2f value %.3g e2e %.3g"%(d["ms_per_step"], d["value"], d["e2e"]["value"]))
for r in d["config"]["ranks"]: print(r)
print({k:v for k,v in list(d["roofline"]["kernels_ms_per_step"].items())[:14]})
PY
tail -3 gpurun_out/r02o_n2.err
