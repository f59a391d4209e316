python -m pytest tests -m gpu -q --maxfail=8 -k "arc or observ or dropin or stage_by_stage" > gpurun_out/r02z_arc.log 2>&1; echo "arc rc=$?"; tail -6 gpurun_out/r02z_arc.log
