#!/bin/bash
# two GPUs: the finite-difference order parameter across a rank boundary
mkdir -p gpurun_out
timeout 30 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29602 tests/mgpu_worker.py --nvel 19 --halo reduced --fd 2 --gradient 27 --visc 1 --steps 8 > gpurun_out/mgpu2_fd.log 2>&1; echo "rc=$?" >> gpurun_out/mgpu2_fd.log
grep -E "mgpu world|rc=|Error|error" gpurun_out/mgpu2_fd.log | tail -8
