# A/B of two builds of the library on one box: bench without the CPU legs, per-kernel times from the profile pass
for lib in "" base ""; do
  if [ -n "$lib" ]; then export PMPS_LIB=$PWD/polymps_b200/libpmps_$lib.so; else unset PMPS_LIB; fi
  python bench.py --steps 30 --warmup 6 --no-cpu --rest-steps 10 --e2e-steps 1 --profile-steps 4 > gpurun_out/ab_${lib:-new}.json 2> gpurun_out/ab_${lib:-new}.err
  python - <<PY
import json
d=json.load(open("gpurun_out/ab_${lib:-new}.json")); k=d["roofline"]["kernels_ms_per_step"]
print("${lib:-new}", "flow %.2f rest %.2f"%(d["ms_per_step"], d["config"]["at_rest"]["ms_per_step"]), {n:k[n] for n in ("F3.pnd+diffusive","F1.viscosity+predict","F5.gradient","shift.gradient","shift.concentration","list.build")})
PY
done
