# usage: bash tools/r02_scale.sh N  — weak-scaling flowing bench, the 50 M block (strong) and, at N = 8, the 200 M block; NCCL parity at N
N=$1
P=$((29500 + N))
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P "$@"; }
run bench.py --gpus $N --steps 20 --warmup 5 --e2e-steps 6 > gpurun_out/r02_weak_n$N.json 2> gpurun_out/r02_weak_n$N.err; echo "weak rc=$?"
run bench.py --gpus $N --total-k 292 --steps 20 --warmup 5 --e2e-steps 4 > gpurun_out/r02_50M_n$N.json 2> gpurun_out/r02_50M_n$N.err; echo "50M rc=$?"
if [ "$N" = "8" ]; then run bench.py --gpus $N --total-k 464 --steps 20 --warmup 5 --e2e-steps 3 > gpurun_out/r02_200M_n$N.json 2> gpurun_out/r02_200M_n$N.err; echo "200M rc=$?"; fi
run tests/slab_parity.py --backend nccl --steps 40 > gpurun_out/r02_slab_parity_nccl_${N}gpu.log 2>&1; echo "parity rc=$?"; tail -2 gpurun_out/r02_slab_parity_nccl_${N}gpu.log
python - <<PY
import json
def load(p):
    try:
        txt=open(p).read(); return json.loads([l for l in txt.splitlines() if l.startswith("{")][-1])
    except Exception as e: return None
for tag in ["weak","50M","200M"]:
    d=load("gpurun_out/r02_%s_n$N.json"%tag)
    if d: print(tag, "N=$N ms/step %.2f value %.4g e2e %.4g rebuilds %s refilters %s"%(d["ms_per_step"], d["value"], d["e2e"]["value"], d["config"]["bucket_rebuilds"], d["config"]["hit_refilters"]), [ (r["owned"], r["halo"], r["repartitions"], r["migrants_out"]) for r in d["config"]["ranks"]])
PY
