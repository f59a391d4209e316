python -m pytest tests/test_flow_gpu.py -m gpu -x -q -s > gpurun_out/r02a_flowtest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02a_flowtest.log
python bench.py --steps 40 --warmup 5 --no-cpu > gpurun_out/r02a_bench_flow.json 2> gpurun_out/r02a_bench_flow.err; echo "bench rc=$?"
python bench.py --k 108 --steps 6 --warmup 3 --no-cpu --rest-steps 0 --e2e-steps 1 --profile-steps 1 > gpurun_out/r02a_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_tv_build|k_tv_hits" -c 3 -o gpurun_out/r02a_build_hits python bench.py --k 108 --steps 6 --warmup 3 --no-cpu --rest-steps 0 --e2e-steps 1 --profile-steps 1 > gpurun_out/r02a_ncu.log 2>&1
tail -30 gpurun_out/r02a_flowtest.log; head -c 6000 gpurun_out/r02a_bench_flow.json; tail -5 gpurun_out/r02a_bench_flow.err
