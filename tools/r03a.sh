python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r03a_gputests.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r03a_gputests.log
