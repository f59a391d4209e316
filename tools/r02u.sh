python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r02u_gputests.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02u_gputests.log
bash tools/r02_scale.sh 2
