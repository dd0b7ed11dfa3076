#!/bin/bash
# one-off GPU job: new-feature tests + micro benchmarks
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "wall or field_stats or io_records or gradient_27pt" > gpurun_out/t_new.log 2>&1; echo "rc=$?" >> gpurun_out/t_new.log
tail -25 gpurun_out/t_new.log
python -m pytest tests/test_regression_logs.py tests/test_host_shim.py -x -q -m gpu > gpurun_out/t_reg.log 2>&1; echo "rc=$?" >> gpurun_out/t_reg.log
tail -25 gpurun_out/t_reg.log
python tools/time_aux_kernels.py > gpurun_out/aux_kernels.log 2>&1
cat gpurun_out/aux_kernels.log
python bench.py --ndist 2 --local 128 128 128 --steps 100 --no-cpu --no-e2e --gradient 27 > gpurun_out/b_bin27.json 2> gpurun_out/b_bin27.err
python bench.py --steps 100 --no-cpu --no-e2e --walls 0 1 1 > gpurun_out/b_walls.json 2> gpurun_out/b_walls.err
tail -3 gpurun_out/b_walls.err
cat gpurun_out/b_bin27.json gpurun_out/b_walls.json | cut -c1-300
