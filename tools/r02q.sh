PMPS_DEBUG=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 tests/slab_parity.py --backend gloo --one-gpu --steps 8 --drift 0.5 > gpurun_out/r02q_dbg.log 2>&1; echo rc=$?
grep -c "hit-list check" gpurun_out/r02q_dbg.log; grep "hit-list check\|mid-step" gpurun_out/r02q_dbg.log | head -80
