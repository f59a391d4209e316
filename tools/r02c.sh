python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r02c_gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02c_gputests.log
tail -40 gpurun_out/r02c_gputests.log
python bench.py --steps 40 --warmup 5 --no-cpu > gpurun_out/r02c_bench.json 2> gpurun_out/r02c_bench.err; echo "bench rc=$?"
PMPS_DEBUG=1 python bench.py --k 108 --steps 14 --warmup 2 --no-cpu --rest-steps 0 --e2e-steps 1 --profile-steps 1 > gpurun_out/r02c_dbg.json 2> gpurun_out/r02c_dbg.err
grep "mid-step" gpurun_out/r02c_dbg.err | head -20
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02c_bench.json"))
print("ms/step %.2f"%d["ms_per_step"], "rest %.2f"%d["config"]["at_rest"]["ms_per_step"], "rebuilds",d["config"]["bucket_rebuilds"],"refilters",d["config"]["hit_refilters"], "e2e %.3g"%d["e2e"]["value"])
print({k:v2 for k,v2 in list(d["roofline"]["kernels_ms_per_step"].items())[:16]}, d["roofline"]["profile_pass"])
PY
python bench.py --k 108 --steps 6 --warmup 3 --no-cpu --rest-steps 0 --e2e-steps 1 --profile-steps 1 > gpurun_out/r02c_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_hl_build|BPndNeighNpcd|BConcentration" -s 12 -c 8 -o gpurun_out/r02c_prof python bench.py --k 108 --steps 6 --warmup 3 --no-cpu --rest-steps 0 --e2e-steps 1 --profile-steps 1 > gpurun_out/r02c_ncu.log 2>&1; tail -3 gpurun_out/r02c_ncu.log
