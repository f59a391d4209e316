# final one-GPU package of round 2
# (the ncu passes first: the bench line then carries roofline.traffic / fp64_active of THIS build)
CMD="python bench.py --steps 4 --warmup 5 --no-cpu --rest-steps 0 --e2e-steps 1 --profile-steps 1"
$CMD > gpurun_out/r04_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 700 -c 400 --csv --log-file gpurun_out/r04_launches.csv $CMD > gpurun_out/r04_ncu1.log 2>&1; echo "launch list rc=$?"
ncu --set full --clock-control none -k regex:"k_tv|k_hl_build|k_collide|k_wall_closest" -s 120 -c 20 -o gpurun_out/r04_prof $CMD > gpurun_out/r04_ncu2.log 2>&1; echo "full rc=$?"; tail -2 gpurun_out/r04_ncu2.log
python profiles/make_traffic.py gpurun_out/r04_prof.ncu-rep 10000422 gpurun_out/r04_traffic.json > gpurun_out/r04_traffic.txt 2>&1; cat gpurun_out/r04_traffic.txt
python profiles/ncu_summary.py gpurun_out/r04_prof.ncu-rep > gpurun_out/r04_ncu_summaries.txt 2>&1
ls -la gpurun_out/r04_prof.ncu-rep; if [ $(stat -c %s gpurun_out/r04_prof.ncu-rep) -gt 30000000 ]; then rm gpurun_out/r04_prof.ncu-rep; fi; du -sh gpurun_out
cp gpurun_out/r04_traffic.json profiles/r02_traffic.json
python bench.py > gpurun_out/r04_bench_n1.json 2> gpurun_out/r04_bench_n1.err; echo "bench rc=$?"
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r04_ref.json 2> gpurun_out/r04_ref.err; echo "ref rc=$?"
python bench.py --scheme inc --state rest --steps 10 --warmup 4 --no-cpu --rest-steps 0 --e2e-steps 2 > gpurun_out/r04_inc.json 2> gpurun_out/r04_inc.err; echo "inc rc=$?"
for k in 17 26; do python bench.py --k $k --steps 200 --warmup 20 --no-cpu --rest-steps 200 --e2e-steps 20 --profile-steps 4 > gpurun_out/r04_small_k$k.json 2> gpurun_out/r04_small_k$k.err; echo "small $k rc=$?"; done
python - <<'PY'
import json
d=json.load(open("gpurun_out/r04_bench_n1.json"))
print("N=1 ms/step %.2f value %.4g rest %.2f e2e %.4g cpu %s"%(d["ms_per_step"], d["value"], d["config"]["at_rest"]["ms_per_step"], d["e2e"]["value"], d.get("cpu_baseline",{}).get("value")))
PY
python tools/vtu_timing.py 171 > gpurun_out/r04_vtu_timing.json 2> gpurun_out/r04_vtu_timing.err; echo "vtu rc=$?"; cat gpurun_out/r04_vtu_timing.json
python tools/long_run.py 171 400 > gpurun_out/r04_long_run.json 2> gpurun_out/r04_long_run.err; echo "long run rc=$?"; python -c "import json; d=json.load(open('gpurun_out/r04_long_run.json')); print({k:d[k] for k in d if k!='history'}, d['history'][::5])"
