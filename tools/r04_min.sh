# the part of the one-GPU package that is tied to the build id: ncu launch list + full capture -> traffic json, then the bench line
CMD="python bench.py --steps 4 --warmup 5 --no-cpu --rest-steps 0 --e2e-steps 1 --profile-steps 1"
$CMD > gpurun_out/r04_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 700 -c 400 --csv --log-file gpurun_out/r04_launches.csv $CMD > gpurun_out/r04_ncu1.log 2>&1; echo "launch list rc=$?"
ncu --set full --clock-control none -k regex:"k_tv|k_hl_build|k_collide|k_wall_closest" -s 120 -c 20 -o gpurun_out/r04_prof $CMD > gpurun_out/r04_ncu2.log 2>&1; echo "full rc=$?"
python profiles/make_traffic.py gpurun_out/r04_prof.ncu-rep 10000422 gpurun_out/r04_traffic.json > gpurun_out/r04_traffic.txt 2>&1; tail -3 gpurun_out/r04_traffic.txt
python profiles/ncu_summary.py gpurun_out/r04_prof.ncu-rep > gpurun_out/r04_ncu_summaries.txt 2>&1
rm -f gpurun_out/r04_prof.ncu-rep
cp gpurun_out/r04_traffic.json profiles/r02_traffic.json
python bench.py > gpurun_out/r04_bench_n1.json 2> gpurun_out/r04_bench_n1.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r04_bench_n1.json')); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['traffic'], d['roofline']['traffic_source'])"
