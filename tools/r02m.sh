CMD="python bench.py --k 108 --steps 6 --warmup 4 --no-cpu --rest-steps 0 --e2e-steps 1 --profile-steps 1"
$CMD > gpurun_out/r02m_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_tv|k_hl_build" -s 60 -c 14 -o gpurun_out/r02m_prof $CMD > gpurun_out/r02m_ncu.log 2>&1; tail -3 gpurun_out/r02m_ncu.log
