# after the stale-list visibility fix: the three failing tests, then everything, then ncu (full set) of one flowing step at 2.5 M
python -m pytest tests/test_flow_gpu.py tests/test_gpu_parity.py -m gpu -q -x -k "guard or kick or o3 or ood" > gpurun_out/r02e_fix.log 2>&1; echo "fix rc=$?"; tail -5 gpurun_out/r02e_fix.log
python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r02e_gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02e_gputests.log
tail -8 gpurun_out/r02e_gputests.log
python bench.py --steps 40 --warmup 5 --no-cpu > gpurun_out/r02e_bench.json 2> gpurun_out/r02e_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02e_bench.json"))
print("ms/step %.2f"%d["ms_per_step"], "rest %.2f"%d["config"]["at_rest"]["ms_per_step"], "rebuilds",d["config"]["bucket_rebuilds"],"refilters",d["config"]["hit_refilters"], "e2e %.3g"%d["e2e"]["value"])
print({k:v2 for k,v2 in list(d["roofline"]["kernels_ms_per_step"].items())[:12]})
PY
CMD="python bench.py --k 108 --steps 4 --warmup 3 --no-cpu --rest-steps 0 --e2e-steps 1 --profile-steps 1"
$CMD > gpurun_out/r02e_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_tv|k_hl_build" -s 36 -c 16 -o gpurun_out/r02e_prof $CMD > gpurun_out/r02e_ncu.log 2>&1; tail -3 gpurun_out/r02e_ncu.log
