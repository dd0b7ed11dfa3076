#!/bin/bash
# one-off GPU job: new-feature tests + micro benchmarks
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "gradient_27pt or noise or fluctuat" > gpurun_out/t_new.log 2>&1; echo "rc=$?" >> gpurun_out/t_new.log
tail -5 gpurun_out/t_new.log
python -m pytest tests/test_regression_logs.py tests/test_host_shim.py -x -q -m gpu > gpurun_out/t_reg.log 2>&1; echo "rc=$?" >> gpurun_out/t_reg.log
tail -5 gpurun_out/t_reg.log
python tools/time_aux_kernels.py > gpurun_out/aux_kernels.log 2>&1
cat gpurun_out/aux_kernels.log
python bench.py --ndist 2 --local 128 128 128 --steps 100 --no-cpu --no-e2e > gpurun_out/b_bin7.json 2> gpurun_out/b_bin7.err
python bench.py --ndist 2 --local 128 128 128 --steps 100 --no-cpu --no-e2e --gradient 27 > gpurun_out/b_bin27.json 2> gpurun_out/b_bin27.err
python bench.py --steps 100 --no-cpu --no-e2e --kt 1e-4 > gpurun_out/b_noise.json 2> gpurun_out/b_noise.err
python bench.py --steps 100 --no-cpu --no-e2e > gpurun_out/b_plain.json 2> gpurun_out/b_plain.err
cat gpurun_out/b_bin7.json gpurun_out/b_bin27.json gpurun_out/b_noise.json gpurun_out/b_plain.json | cut -c1-700
