# capture for the line-level view of the heaviest kernels of a flowing step (after the same command has run without ncu); the report is
# read where nvdisasm is at hand:  python profiles/ncu_summary.py gpurun_out/r04l.ncu-rep --lines <mangled name substring> --top 28
CMD="python bench.py --steps 4 --warmup 5 --no-cpu --rest-steps 0 --e2e-steps 1 --profile-steps 1"
$CMD > gpurun_out/r04l_plain.log 2>&1 || exit 1
ncu --set full --clock-control none -k regex:"k_tv|k_hl_build" -s 120 -c 14 -o gpurun_out/r04l $CMD > gpurun_out/r04l_ncu.log 2>&1; echo "capture rc=$?"
ls -la gpurun_out/r04l.ncu-rep
