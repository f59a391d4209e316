python -m pytest tests/test_slab_gpu.py -m gpu -q > gpurun_out/r02s_slabtests.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r02s_slabtests.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tests/slab_parity.py --backend nccl --steps 40 > gpurun_out/r02s_parity_nccl2.log 2>&1; echo "parity rc=$?"; tail -3 gpurun_out/r02s_parity_nccl2.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --e2e-steps 6 > gpurun_out/r02s_n2.json 2> gpurun_out/r02s_n2.err; echo "bench2 rc=$?"
python - <<'PY'
import json
def load(p):
    txt=open(p).read(); return json.loads([l for l in txt.splitlines() if l.startswith("{")][-1])
d=load("gpurun_out/r02s_n2.json")
print("N=2 ms/step %.2f value %.3g e2e %.3g"%(d["ms_per_step"], d["value"], d["e2e"]["value"]))
for r in d["config"]["ranks"]: print(r)
PY
tail -5 gpurun_out/r02s_n2.err
