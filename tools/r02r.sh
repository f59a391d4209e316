PMPS_SLAB_TIMING=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 6 --warmup 5 --e2e-steps 1 --profile-steps 1 > gpurun_out/r02r_n2.log 2> gpurun_out/r02r_n2.err; echo "rc=$?"
grep "\[slab\]" gpurun_out/r02r_n2.log | sed -n 120,260p
