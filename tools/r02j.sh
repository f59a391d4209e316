for v in base cur fix; do
  L=$PWD/polymps_b200/libpmps_$v.so; [ $v = cur ] && L=$PWD/polymps_b200/libpmps.so
  PMPS_LIB=$L python bench.py --state rest --steps 10 --warmup 5 --no-cpu --rest-steps 0 --e2e-steps 1 --profile-steps 2 > gpurun_out/r02j_$v.json 2> gpurun_out/r02j_$v.err; echo "bench $v rc=$?"
done
python - <<'PY'
import json
for v in ["base","cur","fix"]:
    try:
        d=json.load(open("gpurun_out/r02j_%s.json"%v))
        k=d["roofline"]["kernels_ms_per_step"]
        print("%-6s ms/step %.2f"%(v,d["ms_per_step"]), " ".join("%s %.2f"%(n.split('.')[0][:2]+n.split('.')[-1][:5],k[n]) for n in ["F3.pnd+diffusive","F1.viscosity+predict","F5.gradient","shift.gradient","shift.concentration","F2.collision","W5.wall+correct"]))
    except Exception as e: print(v, "failed", e)
PY
