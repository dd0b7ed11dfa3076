#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_host_shim.py -q -m gpu -k "lees or host_mirror" > gpurun_out/t_le.log 2>&1; echo "rc=$?" >> gpurun_out/t_le.log
tail -40 gpurun_out/t_le.log
timeout 120 python bench.py --le 4 --steps 50 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_le.json 2> gpurun_out/bench_le.err; tail -3 gpurun_out/bench_le.err; cat gpurun_out/bench_le.json
