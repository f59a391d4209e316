CMD="python bench.py --k 108 --steps 12 --warmup 6 --no-cpu --rest-steps 0 --e2e-steps 1 --profile-steps 1"
$CMD > gpurun_out/r02x_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_tv|k_hl_build|k_collide|k_wall_closest" -s 150 -c 16 -o gpurun_out/r02x_prof $CMD > gpurun_out/r02x_ncu.log 2>&1; tail -2 gpurun_out/r02x_ncu.log
