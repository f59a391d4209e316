python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --e2e-steps 6 > gpurun_out/r02w_n2.json 2> gpurun_out/r02w_n2.err; echo "bench2 rc=$?"
python - <<'PY'
import json
def load(p):
    txt=open(p).read(); return json.loads([l for l in txt.splitlines() if l.startswith("{")][-1])
d=load("gpurun_out/r02w_n2.json")
print("N=2 ms/step %.2f value %.3g e2e %.3g"%(d["ms_per_step"], d["value"], d["e2e"]["value"]), d["config"].get("host_binding"))
PY
nvidia-smi topo -m | head -20
lscpu | grep -i "numa\|socket\|model name\|^CPU(s)"
